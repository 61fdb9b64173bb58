// Stand-in for the un-vendored PIGO submodule (/root/reference/.gitmodules:1-4), TEST INFRASTRUCTURE
// ONLY: just the two entry points the reference uses (src/matrix.hpp:225-233 pigo::CSR, :487-491
// pigo::ROFile), reading the PIGO-CSR-v2 layout recovered in SURVEY.md appendix A.  It lets the
// reference's own sources compile, unmodified and where they lie, into oracle/_ref.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace pigo {

class ROFile {
    FILE *f_;

public:
    explicit ROFile(const std::string &path) : f_(std::fopen(path.c_str(), "rb")) {
        if (!f_) throw std::runtime_error("cannot open " + path);
    }
    ~ROFile() { if (f_) std::fclose(f_); }
    template <typename T>
    T read() {
        T v;
        if (std::fread(&v, sizeof(T), 1, f_) != 1) throw std::runtime_error("short read");
        return v;
    }
    void parallel_read(char *dst, std::size_t bytes) {
        while (bytes) {
            std::size_t want = bytes < (std::size_t(1) << 30) ? bytes : (std::size_t(1) << 30);
            if (std::fread(dst, 1, want, f_) != want) throw std::runtime_error("short read");
            dst += want;
            bytes -= want;
        }
    }
};

template <typename Label, typename Ordinal, typename LabelStorage, typename OrdinalStorage, bool weighted,
          typename Weight, typename WeightStorage>
class CSR {
    Label n_ = 0, nrows_ = 0, ncols_ = 0;
    std::size_t m_ = 0;
    std::vector<std::remove_pointer_t<OrdinalStorage>> offsets_;
    std::vector<Label> endpoints_;
    std::vector<Weight> weights_;

public:
    explicit CSR(const std::string &path) {
        ROFile f(path);
        char magic[11];
        f.parallel_read(magic, 11);
        if (std::memcmp(magic, "PIGO-CSR-v2", 11) != 0) throw std::runtime_error("not a PIGO-CSR-v2 file");
        const std::uint8_t lb = f.read<std::uint8_t>(), ob = f.read<std::uint8_t>();
        using O = std::remove_pointer_t<OrdinalStorage>;
        if (lb != sizeof(Label) || ob != sizeof(O)) throw std::runtime_error("PIGO type widths do not match the file");
        n_ = f.read<Label>();
        m_ = f.read<O>();
        nrows_ = f.read<Label>();
        ncols_ = f.read<Label>();
        offsets_.resize(std::size_t(n_) + 1);
        f.parallel_read(reinterpret_cast<char *>(offsets_.data()), offsets_.size() * sizeof(O));
        endpoints_.resize(m_);
        f.parallel_read(reinterpret_cast<char *>(endpoints_.data()), m_ * sizeof(Label));
        weights_.resize(m_);
        if (weighted) f.parallel_read(reinterpret_cast<char *>(weights_.data()), m_ * sizeof(Weight));
    }
    Label nrows() const { return nrows_; }
    Label ncols() const { return ncols_; }
    std::size_t m() const { return m_; }
    auto offsets() { return offsets_.data(); }
    auto endpoints() { return endpoints_.data(); }
    auto weights() { return weights_.data(); }
};

}  // namespace pigo
