// dn_matrix<r_t>: row-major N x M, leading dimension exactly M, shallow-copy value type over a
// shared buffer (src/matrix.hpp:478-639).  Included by matrix.hpp.
#pragma once

namespace mg_gcn {

template <typename r_t>
class dist_row_dn_matrix;
template <typename r_t>
class repl_dn_matrix;

template <typename r_t>
class dn_matrix {
    std::size_t N_ = 0, M_ = 0;
    cuda_ptr<r_t> buffer_;

    // "u32 N, u32 M, N*M elements" (test/data/prep.py:67-76; reader src/matrix.hpp:486-492)
    void read_file(const std::filesystem::path &path) {
        static_assert(sizeof(r_t) == 4, "the on-disk dense format holds 4-byte elements");
        std::int64_t n = 0, m = 0;
        if (mggcn_dense_file_shape(path.c_str(), &n, &m) != MGGCN_OK) throw matrix_error(mggcn_last_error());
        N_ = (std::size_t)n;
        M_ = (std::size_t)m;
        buffer_ = cuda_malloc_managed<r_t>(N_ * M_);
        if (mggcn_dense_file_read(path.c_str(), buffer_.get(), sizeof(r_t)) != MGGCN_OK)
            throw matrix_error(mggcn_last_error());
    }

   public:
    dn_matrix() = default;
    dn_matrix(const std::filesystem::path &path) {
        if (path.extension() != ".bin") throw matrix_error("File type is not supported.");
        read_file(path);
    }
    dn_matrix(std::size_t N, std::size_t M) : N_(N), M_(M), buffer_(cuda_malloc_managed<r_t>(N * M)) {}
    // non-explicit on purpose: callers rely on pair -> dn_matrix conversion (src/gcn.hpp:255-261)
    dn_matrix(std::pair<std::size_t, std::size_t> shape) : dn_matrix(shape.first, shape.second) {}
    dn_matrix(std::size_t N, std::size_t M, cuda_ptr<r_t> buffer) : N_(N), M_(M), buffer_(std::move(buffer)) {
        if (!buffer_) buffer_ = cuda_malloc_managed<r_t>(N_ * M_);
    }
    dn_matrix(std::pair<std::size_t, std::size_t> shape, cuda_ptr<r_t> buffer)
        : dn_matrix(shape.first, shape.second, std::move(buffer)) {}

    // Weight init of the reference: default_random_engine(99) re-seeded per matrix,
    // uniform(-g, g) with g = gain * sqrt(3 / rows)  (src/matrix.hpp:539-545).
    void init(r_t gain = (r_t)std::sqrt(2 / (1 + 0.01 * 0.01))) {
        static_assert(std::is_same<r_t, float>::value, "init is defined for float matrices");
        CHECK_MGGCN(mggcn_init_weights((std::int64_t)N_, (std::int64_t)M_, gain, buffer_.get()));
    }
    void init(const std::vector<r_t> &f) { std::copy(f.begin(), f.end(), begin()); }

    void copy_to(const context &ctx, dn_matrix<r_t> other, cudaMemcpyKind kind = cudaMemcpyDefault) const {
        ctx.set();
        CHECK_CUDA(cudaMemcpyAsync(other.buffer(), buffer(), size() * sizeof(r_t), kind, ctx.stream(0)));
    }
    dn_matrix copy(const context &ctx, cudaMemcpyKind kind = cudaMemcpyDefault) const {
        dn_matrix clone(N_, M_);
        copy_to(ctx, clone, kind);
        return clone;
    }
    void zero(const context &ctx) {
        ctx.set();
        CHECK_CUDA(cudaMemsetAsync(buffer(), 0, size() * sizeof(r_t), ctx.stream(0)));
    }

    r_t operator[](std::size_t i) const { return buffer_[i]; }
    r_t operator[](std::pair<std::size_t, std::size_t> p) const { return buffer_[p.first * M_ + p.second]; }
    r_t &operator[](std::size_t i) { return buffer_[i]; }
    r_t &operator[](std::pair<std::size_t, std::size_t> p) { return buffer_[p.first * M_ + p.second]; }

    std::size_t n() const { return N_; }
    std::size_t m() const { return M_; }
    std::size_t size() const { return N_ * M_; }
    r_t *buffer() const { return buffer_.get(); }
    cuda_ptr<r_t> shared_buffer() const { return buffer_; }
    r_t *begin() const { return buffer_.get(); }
    r_t *end() const { return buffer_.get() + N_ * M_; }
    std::pair<std::size_t, std::size_t> shape() const { return {N_, M_}; }

    // The reference transposes with cublasSgeam (src/matrix.hpp:620-625); nothing on the hot path
    // transposes a dense matrix (the GEMMs take transpose flags), so this is a host loop after a sync.
    dn_matrix<r_t> transpose(const context &ctx) const {
        ctx.sync();
        dn_matrix<r_t> t(M_, N_);
        for (std::size_t i = 0; i < N_; ++i)
            for (std::size_t j = 0; j < M_; ++j) t.buffer_[j * N_ + i] = buffer_[i * M_ + j];
        return t;
    }

    void print(std::ostream &out = std::cout) const {
        for (std::size_t i = 0; i < N_; ++i) {
            for (std::size_t j = 0; j < M_; ++j) out << buffer_[i * M_ + j] << ' ';
            out << '\n';
        }
        out << std::endl << std::endl;
    }

    friend class dist_row_dn_matrix<r_t>;
    friend class repl_dn_matrix<r_t>;
};

}  // namespace mg_gcn
