// On-disk formats of the reference dataset directory (graph.bin / features.bin / labels.bin /
// sets.bin).  Writer-by-example: /root/reference/test/data/prep.py:46-76; consumers:
// src/matrix.hpp:224-234 (pigo::CSR) and :486-492 (pigo::ROFile).  PIGO itself is an un-vendored
// submodule, so the layout below was recovered from prep.py and the toyA/toyB fixtures
//  and is checked byte-for-byte against them in tests/test_io.py.
#include <cerrno>
#include <cstdlib>
#include <string>
#include <vector>

#include "common.hpp"

namespace mggcn {

thread_local char t_err[512] = {0};
char *last_error_buf() { return t_err; }
std::atomic<uint64_t> g_launches{0};

int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_err, sizeof t_err, fmt, ap);
    va_end(ap);
    return code;
}

namespace {

struct file_closer {
    FILE *f;
    ~file_closer() { if (f) fclose(f); }
};

constexpr char kMagic[] = "PIGO-CSR-v2";  // 11 bytes, no NUL on disk
constexpr size_t kMagicLen = 11;

bool read_exact(FILE *f, void *dst, size_t bytes) {
    // fread in 1 GiB slices: papers100M-shape files are tens of GB
    char *d = static_cast<char *>(dst);
    while (bytes) {
        size_t want = bytes < (size_t(1) << 30) ? bytes : (size_t(1) << 30);
        size_t got = fread(d, 1, want, f);
        if (got != want) return false;
        d += got;
        bytes -= got;
    }
    return true;
}

bool write_exact(FILE *f, const void *src, size_t bytes) {
    const char *s = static_cast<const char *>(src);
    while (bytes) {
        size_t want = bytes < (size_t(1) << 30) ? bytes : (size_t(1) << 30);
        if (fwrite(s, 1, want, f) != want) return false;
        s += want;
        bytes -= want;
    }
    return true;
}

bool read_uint(FILE *f, int width, uint64_t *out) {
    if (width == 4) {
        uint32_t v;
        if (!read_exact(f, &v, 4)) return false;
        *out = v;
        return true;
    }
    return read_exact(f, out, 8);
}

int read_header(FILE *f, const char *path, mggcn_csr_file_info *info) {
    char magic[kMagicLen];
    if (!read_exact(f, magic, kMagicLen) || memcmp(magic, kMagic, kMagicLen) != 0)
        return fail(MGGCN_E_FORMAT, "%s: not a PIGO-CSR-v2 file", path);
    unsigned char widths[2];
    if (!read_exact(f, widths, 2)) return fail(MGGCN_E_IO, "%s: short header", path);
    if ((widths[0] != 4 && widths[0] != 8) || (widths[1] != 4 && widths[1] != 8))
        return fail(MGGCN_E_FORMAT, "%s: label/ordinal width %d/%d unsupported", path, widths[0],
                    widths[1]);
    info->label_bytes = widths[0];
    info->ordinal_bytes = widths[1];
    uint64_t n, m, nr, nc;
    if (!read_uint(f, widths[0], &n) || !read_uint(f, widths[1], &m) ||
        !read_uint(f, widths[0], &nr) || !read_uint(f, widths[0], &nc))
        return fail(MGGCN_E_IO, "%s: short header", path);
    info->n = (int64_t)n;
    info->nnz = (int64_t)m;
    info->nrows = (int64_t)nr;
    info->ncols = (int64_t)nc;
    return MGGCN_OK;
}

}  // namespace
}  // namespace mggcn

using namespace mggcn;

extern "C" {

int mggcn_abi_version(void) { return 1; }
const char *mggcn_last_error(void) { return last_error_buf(); }
uint64_t mggcn_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int mggcn_csr_file_info_read(const char *path, mggcn_csr_file_info *info) {
    if (!path || !info) return fail(MGGCN_E_INVALID, "null argument");
    FILE *f = fopen(path, "rb");
    if (!f) return fail(MGGCN_E_IO, "%s: %s", path, strerror(errno));
    file_closer fc{f};
    int rc = read_header(f, path, info);
    if (rc) return rc;
    // weights are optional in the format; prep.py always writes them
    long hdr = ftell(f);
    fseek(f, 0, SEEK_END);
    int64_t total = ftell(f);
    int64_t body = (info->n + 1) * (int64_t)info->ordinal_bytes + info->nnz * (int64_t)info->label_bytes;
    info->has_weights = (total - hdr - body) >= info->nnz * 4 ? 1 : 0;
    if (total - hdr < body) return fail(MGGCN_E_FORMAT, "%s: truncated (%lld < %lld)", path,
                                        (long long)(total - hdr), (long long)body);
    return MGGCN_OK;
}

int mggcn_csr_file_read(const char *path, uint64_t *rowptr, uint32_t *colidx, float *vals) {
    if (!path || !rowptr || !colidx) return fail(MGGCN_E_INVALID, "null argument");
    mggcn_csr_file_info info;
    int rc = mggcn_csr_file_info_read(path, &info);
    if (rc) return rc;
    if (info.label_bytes != 4)
        return fail(MGGCN_E_UNSUPPORTED, "%s: 64-bit vertex labels are not supported", path);
    FILE *f = fopen(path, "rb");
    if (!f) return fail(MGGCN_E_IO, "%s: %s", path, strerror(errno));
    file_closer fc{f};
    mggcn_csr_file_info again;
    rc = read_header(f, path, &again);
    if (rc) return rc;
    const size_t np1 = (size_t)info.n + 1;
    if (info.ordinal_bytes == 8) {
        if (!read_exact(f, rowptr, np1 * 8)) return fail(MGGCN_E_IO, "%s: short offsets", path);
    } else {
        // read the 32-bit offsets into the top half of the output, then widen front to back
        uint32_t *narrow = reinterpret_cast<uint32_t *>(rowptr) + np1;
        if (!read_exact(f, narrow, np1 * 4)) return fail(MGGCN_E_IO, "%s: short offsets", path);
        for (size_t i = 0; i < np1; ++i) rowptr[i] = narrow[i];
    }
    if (!read_exact(f, colidx, (size_t)info.nnz * 4)) return fail(MGGCN_E_IO, "%s: short endpoints", path);
    if (vals) {
        if (info.has_weights) {
            if (!read_exact(f, vals, (size_t)info.nnz * 4)) return fail(MGGCN_E_IO, "%s: short weights", path);
        } else {
            for (int64_t e = 0; e < info.nnz; ++e) vals[e] = 1.0f;
        }
    }
    if (rowptr[0] != 0 || rowptr[info.n] != (uint64_t)info.nnz)
        return fail(MGGCN_E_FORMAT, "%s: offsets do not span nnz", path);
    return MGGCN_OK;
}

int mggcn_csr_file_write(const char *path, int64_t n, int64_t ncols, const uint64_t *rowptr,
                         const uint32_t *colidx, const float *vals, int ordinal_bytes) {
    if (!path || !rowptr || !colidx || !vals || n < 0) return fail(MGGCN_E_INVALID, "bad argument");
    if (ordinal_bytes != 4 && ordinal_bytes != 8) return fail(MGGCN_E_INVALID, "ordinal_bytes must be 4 or 8");
    const uint64_t nnz = rowptr[n];
    if (ordinal_bytes == 4 && nnz > 0xFFFFFFFFull)
        return fail(MGGCN_E_INVALID, "nnz %llu needs 64-bit ordinals", (unsigned long long)nnz);
    FILE *f = fopen(path, "wb");
    if (!f) return fail(MGGCN_E_IO, "%s: %s", path, strerror(errno));
    file_closer fc{f};
    unsigned char widths[2] = {4, (unsigned char)ordinal_bytes};
    bool ok = write_exact(f, kMagic, kMagicLen) && write_exact(f, widths, 2);
    uint32_t n32 = (uint32_t)n, nc32 = (uint32_t)ncols;
    ok = ok && write_exact(f, &n32, 4);
    if (ordinal_bytes == 4) {
        uint32_t m32 = (uint32_t)nnz;
        ok = ok && write_exact(f, &m32, 4);
    } else {
        ok = ok && write_exact(f, &nnz, 8);
    }
    ok = ok && write_exact(f, &n32, 4) && write_exact(f, &nc32, 4);
    if (ordinal_bytes == 8) {
        ok = ok && write_exact(f, rowptr, ((size_t)n + 1) * 8);
    } else {
        std::vector<uint32_t> narrow((size_t)n + 1);
        for (int64_t i = 0; i <= n; ++i) narrow[i] = (uint32_t)rowptr[i];
        ok = ok && write_exact(f, narrow.data(), narrow.size() * 4);
    }
    ok = ok && write_exact(f, colidx, nnz * 4) && write_exact(f, vals, nnz * 4);
    if (!ok) return fail(MGGCN_E_IO, "%s: write failed", path);
    return MGGCN_OK;
}

int mggcn_dense_file_shape(const char *path, int64_t *n, int64_t *m) {
    if (!path || !n || !m) return fail(MGGCN_E_INVALID, "null argument");
    FILE *f = fopen(path, "rb");
    if (!f) return fail(MGGCN_E_IO, "%s: %s", path, strerror(errno));
    file_closer fc{f};
    uint32_t shape[2];
    if (!read_exact(f, shape, 8)) return fail(MGGCN_E_FORMAT, "%s: short header", path);
    *n = shape[0];
    *m = shape[1];
    return MGGCN_OK;
}

int mggcn_dense_file_read(const char *path, void *data, size_t elem_bytes) {
    if (!path || !data) return fail(MGGCN_E_INVALID, "null argument");
    FILE *f = fopen(path, "rb");
    if (!f) return fail(MGGCN_E_IO, "%s: %s", path, strerror(errno));
    file_closer fc{f};
    uint32_t shape[2];
    if (!read_exact(f, shape, 8)) return fail(MGGCN_E_FORMAT, "%s: short header", path);
    if (!read_exact(f, data, (size_t)shape[0] * shape[1] * elem_bytes))
        return fail(MGGCN_E_IO, "%s: short body", path);
    return MGGCN_OK;
}

int mggcn_dense_file_write(const char *path, int64_t n, int64_t m, const void *data,
                           size_t elem_bytes) {
    if (!path || !data || n < 0 || m < 0 || n > 0xFFFFFFFFll || m > 0xFFFFFFFFll)
        return fail(MGGCN_E_INVALID, "bad argument");
    FILE *f = fopen(path, "wb");
    if (!f) return fail(MGGCN_E_IO, "%s: %s", path, strerror(errno));
    file_closer fc{f};
    uint32_t shape[2] = {(uint32_t)n, (uint32_t)m};
    if (!write_exact(f, shape, 8) || !write_exact(f, data, (size_t)n * m * elem_bytes))
        return fail(MGGCN_E_IO, "%s: write failed", path);
    return MGGCN_OK;
}

}  // extern "C"
