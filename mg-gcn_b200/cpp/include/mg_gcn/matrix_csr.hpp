// csr_matrix<x_t, v_t, r_t>: N x M CSR, x_t row pointers (32 or 64 bit), v_t = uint32 column ids,
// r_t = float values (src/matrix.hpp:214-468).  Included by matrix.hpp.
#pragma once

namespace mg_gcn {

template <typename x_t, typename v_t, typename r_t>
class dist_row_csr_matrix;

// The per-matrix SpMM plan (work list of csrc/spmm.cu) and its segment workspace; replaces the
// cuSPARSE descriptor + external buffer pair of the reference (src/matrix.hpp:269-273,
// src/cuda_utils.hpp:94-102).  Built on first use, shared by all copies of the matrix.
struct spmm_plan_cache {
    mggcn_spmm_plan_t *plan = nullptr;
    cuda_ptr<char> ws;
    std::size_t ws_bytes = 0;
    std::mutex mu;
    ~spmm_plan_cache() {
        if (plan) mggcn_spmm_plan_destroy(plan);
    }
};

template <typename x_t, typename v_t, typename r_t>
class csr_matrix {
    static_assert(sizeof(v_t) == 4 && sizeof(r_t) == 4, "column ids are 32-bit, values are float");
    static_assert(sizeof(x_t) == 4 || sizeof(x_t) == 8, "row pointers are 32- or 64-bit");

    v_t N_ = 0, M_ = 0;
    cuda_ptr<x_t> indptr_;
    cuda_ptr<v_t> indices_;
    cuda_ptr<r_t> data_;
    std::shared_ptr<spmm_plan_cache> plan_ = std::make_shared<spmm_plan_cache>();

    std::vector<std::uint64_t> wide_indptr() const {
        std::vector<std::uint64_t> w((std::size_t)N_ + 1);
        for (std::size_t i = 0; i <= N_; ++i) w[i] = (std::uint64_t)indptr_[i];
        return w;
    }

    // "PIGO-CSR-v2" (test/data/prep.py:46-62; consumer src/matrix.hpp:224-234)
    void read_file(const std::filesystem::path &path) {
        mggcn_csr_file_info info{};
        if (mggcn_csr_file_info_read(path.c_str(), &info) != MGGCN_OK) throw matrix_error(mggcn_last_error());
        if (sizeof(x_t) == 4 && info.nnz > 0xFFFFFFFFll)
            throw matrix_error("graph has more than 2^32 entries: instantiate csr_matrix with a 64-bit x_t");
        N_ = (v_t)info.nrows;
        M_ = (v_t)info.ncols;
        std::vector<std::uint64_t> w((std::size_t)info.n + 1);
        indptr_ = cuda_malloc_managed<x_t>((std::size_t)info.n + 1);
        indices_ = cuda_malloc_managed<v_t>((std::size_t)info.nnz);
        data_ = cuda_malloc_managed<r_t>((std::size_t)info.nnz);
        if (mggcn_csr_file_read(path.c_str(), w.data(), reinterpret_cast<std::uint32_t *>(indices_.get()),
                                data_.get()) != MGGCN_OK)
            throw matrix_error(mggcn_last_error());
        for (std::size_t i = 0; i < w.size(); ++i) indptr_[i] = (x_t)w[i];
    }

   public:
    csr_matrix() = default;
    csr_matrix(const std::filesystem::path &path) {
        if (path.extension() != ".bin") throw matrix_error("File type is not supported.");
        read_file(path);
    }
    csr_matrix(const std::vector<x_t> &indptr, const std::vector<v_t> &indices, const std::vector<r_t> &data, v_t M)
        : N_((v_t)(indptr.size() - 1)), M_(M) {
        indptr_ = cuda_malloc_managed<x_t>(indptr.size());
        indices_ = cuda_malloc_managed<v_t>(indices.size());
        data_ = cuda_malloc_managed<r_t>(data.size());
        std::copy(indptr.begin(), indptr.end(), indptr_.get());
        std::copy(indices.begin(), indices.end(), indices_.get());
        std::copy(data.begin(), data.end(), data_.get());
    }
    csr_matrix(cuda_ptr<x_t> indptr, cuda_ptr<v_t> indices, cuda_ptr<r_t> data, v_t N, v_t M)
        : N_(N), M_(M), indptr_(std::move(indptr)), indices_(std::move(indices)), data_(std::move(data)) {}

    auto buffer() const { return std::make_tuple(indptr_, indices_, data_); }
    void init_mat() {}  // kept for source compatibility: there is no descriptor to build

    v_t n() const { return N_; }
    v_t m() const { return M_; }
    x_t nnz() const { return indptr_ ? indptr_[N_] : 0; }
    x_t begin(v_t i) const { return indptr_[i]; }
    x_t end(v_t i) const { return indptr_[i + 1]; }
    std::pair<v_t, r_t> operator[](x_t e) const { return {indices_[e], data_[e]}; }
    std::pair<v_t, v_t> shape() const { return {N_, M_}; }
    const x_t *indptr() const { return indptr_.get(); }
    const v_t *indices() const { return indices_.get(); }
    const r_t *data() const { return data_.get(); }

    dn_matrix<r_t> as_dn() const {  // src/matrix.hpp:328-337
        dn_matrix<r_t> d(N_, M_);
        std::fill(d.begin(), d.end(), (r_t)0);
        for (v_t i = 0; i < N_; ++i)
            for (x_t e = begin(i); e < end(i); ++e) d[{i, indices_[e]}] = data_[e];
        return d;
    }

    // axis = true: divide every entry by its COLUMN sum; false: by its row sum.  In place on the
    // shared buffers, so every copy of the matrix sees it (src/matrix.hpp:340-390).
    void normalize(bool axis = false) {
        auto w = wide_indptr();
        CHECK_MGGCN(mggcn_csr_normalize(N_, M_, w.data(), reinterpret_cast<const std::uint32_t *>(indices_.get()),
                                        data_.get(), axis ? 1 : 0));
    }

    // Counting-sort transpose, ascending source row inside each output row (src/matrix.hpp:426-453).
    csr_matrix transpose() const {
        auto w = wide_indptr();
        const std::size_t nz = (std::size_t)nnz();
        std::vector<std::uint64_t> tw((std::size_t)M_ + 1);
        auto ti = cuda_malloc_managed<v_t>(nz);
        auto td = cuda_malloc_managed<r_t>(nz);
        CHECK_MGGCN(mggcn_csr_transpose(N_, M_, w.data(), reinterpret_cast<const std::uint32_t *>(indices_.get()),
                                        data_.get(), tw.data(), reinterpret_cast<std::uint32_t *>(ti.get()), td.get()));
        auto tp = cuda_malloc_managed<x_t>((std::size_t)M_ + 1);
        for (std::size_t i = 0; i <= M_; ++i) tp[i] = (x_t)tw[i];
        return csr_matrix(tp, ti, td, M_, N_);
    }

    void print(std::ostream &out = std::cout) const {
        for (v_t i = 0; i < N_; ++i) {
            out << i << ':';
            for (x_t e = begin(i); e < end(i); ++e) out << ' ' << indices_[e] << '(' << data_[e] << ')';
            out << '\n';
        }
    }

    // Plan and workspace for a feature width F (lazy; the reference's get_matmul_buffer).
    std::pair<const mggcn_spmm_plan_t *, void *> plan(std::size_t F) const {
        spmm_plan_cache &c = *plan_;
        std::lock_guard<std::mutex> g(c.mu);
        if (!c.plan) {
            if (sizeof(x_t) == 8)
                CHECK_MGGCN(mggcn_spmm_plan_create_i64(N_, reinterpret_cast<const std::uint64_t *>(indptr_.get()), 0, &c.plan));
            else
                CHECK_MGGCN(mggcn_spmm_plan_create_i32(N_, reinterpret_cast<const std::uint32_t *>(indptr_.get()), 0, &c.plan));
        }
        const std::size_t need = mggcn_spmm_plan_workspace_bytes(c.plan, (int)F);
        if (need > c.ws_bytes) {
            c.ws = cuda_malloc<char>(need);
            c.ws_bytes = need;
        }
        return {c.plan, c.ws.get()};
    }

    friend class dist_row_csr_matrix<x_t, v_t, r_t>;
};

}  // namespace mg_gcn
