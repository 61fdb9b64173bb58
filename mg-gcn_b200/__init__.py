"""mg-gcn_b200: ctypes view of the C-ABI in include/mggcn.h (lib/libmggcn.so).

This module is plumbing for tests/ and bench.py.  The product is the shared library and the C++
surface (mg-gcn_b200/cpp).  torch is used only to own device memory and streams and for the
process-group bootstrap; every compute call below goes through the C-ABI.  There is no CPU or
PyTorch fallback: if the library is missing or a call fails, an exception is raised.

The directory name contains a hyphen (the repository naming convention), so import it with
`load_package()` from tests/conftest.py / bench.py, which registers it as `mg_gcn_b200`.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libmggcn.so")
INCLUDE = os.path.join(os.path.dirname(_HERE), "include", "mggcn.h")

EPI_NONE, EPI_LEAKY_RELU, EPI_LEAKY_RELU_BWD = 0, 1, 2
COMM_NCCL, COMM_P2P, COMM_P2P_STAGED, COMM_P2P_GROUPED = 0, 1, 2, 3


class MggcnError(RuntimeError):
    pass


def build(verbose=False):
    """make -C mg-gcn_b200 (nvcc, sm_100a only).  Cross-compiles without a GPU."""
    out = subprocess.run(["make", "-C", _HERE, "-j8"], capture_output=True, text=True)
    if out.returncode != 0:
        raise MggcnError("building libmggcn.so failed:\n" + out.stdout[-4000:] + out.stderr[-4000:])
    if verbose:
        print(out.stdout[-2000:])
    return LIB_PATH


class Epilogue(C.Structure):
    _fields_ = [("kind", C.c_int), ("slope", C.c_float), ("bias", C.c_void_p), ("aux", C.c_void_p),
                ("ld_aux", C.c_int64), ("keep_mask", C.c_void_p), ("ld_mask", C.c_int64),
                ("keep_scale", C.c_float)]


class CsrFileInfo(C.Structure):
    _fields_ = [("n", C.c_int64), ("nnz", C.c_int64), ("nrows", C.c_int64), ("ncols", C.c_int64),
                ("label_bytes", C.c_int), ("ordinal_bytes", C.c_int), ("has_weights", C.c_int)]


class EngineConfig(C.Structure):
    _fields_ = [("world", C.c_int), ("rank", C.c_int), ("device", C.c_int), ("num_layers", C.c_int),
                ("sizes", C.POINTER(C.c_int64)), ("n_global", C.c_int64), ("p", C.POINTER(C.c_uint32)),
                ("overlap", C.c_int), ("exact_grads", C.c_int), ("use_graph", C.c_int),
                ("nccl_id", C.c_void_p), ("lrelu_slope", C.c_float), ("comm_backend", C.c_int),
                ("timers", C.c_int)]


_lib = None
_vp, _i64, _i32, _f32, _sz = C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_size_t

_SIGS = {
    "mggcn_abi_version": (C.c_int, []),
    "mggcn_last_error": (C.c_char_p, []),
    "mggcn_launch_count": (C.c_uint64, []),
    "mggcn_spmm_csr_f32_i32": (C.c_int, [_vp, _i64, _i64, _vp, _vp, _vp, _vp, _i64, _vp, _i64, _i32, _f32, _f32,
                                         C.POINTER(Epilogue), _vp, _vp]),
    "mggcn_spmm_csr_f32_i64": (C.c_int, [_vp, _i64, _i64, _vp, _vp, _vp, _vp, _i64, _vp, _i64, _i32, _f32, _f32,
                                         C.POINTER(Epilogue), _vp, _vp]),
    "mggcn_spmm_gathered_f32_i32": (C.c_int, [_vp, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _i64, _i32, _f32, _f32,
                                              _i32, _i32, _vp, _vp, _vp, C.POINTER(Epilogue), _vp, _vp]),
    "mggcn_spmm_plan_create_i32": (C.c_int, [_i64, _vp, _i32, C.POINTER(_vp)]),
    "mggcn_spmm_plan_create_i64": (C.c_int, [_i64, _vp, _i32, C.POINTER(_vp)]),
    "mggcn_spmm_plan_create_blocked_i32": (C.c_int, [_i64, _i64, _vp, _vp, C.c_uint32, _i64, _i32, _i32, C.POINTER(_vp)]),
    "mggcn_spmm_plan_blocks": (C.c_int, [_vp]),
    "mggcn_last_spmm_config": (C.c_char_p, []),
    "mggcn_spmm_plan_workspace_bytes": (_sz, [_vp, _i32]),
    "mggcn_spmm_plan_destroy": (None, [_vp]),
    "mggcn_tune": (C.c_int, [C.c_char_p, _i32]),
    "mggcn_gemm_f32": (C.c_int, [_vp, _i32, _i32, _i64, _i64, _i64, _f32, _vp, _i64, _vp, _i64, _f32, _vp, _i64,
                                 C.POINTER(Epilogue), _vp, _sz]),
    "mggcn_gemm_workspace_bytes": (_sz, [_i32, _i32, _i64, _i64, _i64]),
    "mggcn_colsum_f32": (C.c_int, [_vp, _i64, _i64, _vp, _i64, _vp, _vp, _sz]),
    "mggcn_colsum_workspace_bytes": (_sz, [_i64, _i64]),
    "mggcn_colsum_static_f32": (C.c_int, [_vp, _i64, _i64, _vp, _i64, _vp]),
    "mggcn_leaky_relu_forward_f32": (C.c_int, [_vp, _vp, _vp, _sz, _f32]),
    "mggcn_leaky_relu_backward_f32": (C.c_int, [_vp, _vp, _vp, _vp, _sz, _f32]),
    "mggcn_broadcast_rows_f32": (C.c_int, [_vp, _vp, _vp, _sz, _sz, _i32]),
    "mggcn_axpy_f32": (C.c_int, [_vp, _sz, _f32, _vp, _vp]),
    "mggcn_axpby_f32": (C.c_int, [_vp, _sz, _f32, _vp, _f32, _vp]),
    "mggcn_softmax_xent_f32": (C.c_int, [_vp, _vp, _i64, _i64, _i32, _vp, _f32, _vp, _vp, _sz]),
    "mggcn_softmax_xent_workspace_bytes": (_sz, [_i64]),
    "mggcn_adam_f32": (C.c_int, [_vp, _sz, _vp, _vp, _vp, _vp, _f32, _f32, _f32, _f32, _f32, _f32, _f32, _sz]),
    "mggcn_csr_normalize": (C.c_int, [_i64, _i64, _vp, _vp, _vp, _i32]),
    "mggcn_csr_transpose": (C.c_int, [_i64, _i64, _vp, _vp, _vp, _vp, _vp, _vp]),
    "mggcn_csr_block_counts": (C.c_int, [_i64, _vp, _vp, _i32, _vp, _i32, _vp, _vp]),
    "mggcn_csr_block_fill": (C.c_int, [_i64, _vp, _vp, _vp, _i32, _vp, _i32, _vp, _vp, _vp, _vp]),
    "mggcn_csr_needed_rows": (C.c_int, [_i64, _i64, _vp, _vp, _i32, _vp, _i32, _vp, _vp]),
    "mggcn_csr_file_info_read": (C.c_int, [C.c_char_p, C.POINTER(CsrFileInfo)]),
    "mggcn_csr_file_read": (C.c_int, [C.c_char_p, _vp, _vp, _vp]),
    "mggcn_csr_file_write": (C.c_int, [C.c_char_p, _i64, _i64, _vp, _vp, _vp, _i32]),
    "mggcn_dense_file_shape": (C.c_int, [C.c_char_p, C.POINTER(_i64), C.POINTER(_i64)]),
    "mggcn_dense_file_read": (C.c_int, [C.c_char_p, _vp, _sz]),
    "mggcn_dense_file_write": (C.c_int, [C.c_char_p, _i64, _i64, _vp, _sz]),
    "mggcn_gen_graph_begin": (C.c_int, [_i64, _i64, _i32, C.c_double, _i64, C.c_uint64, _i32, C.POINTER(_vp),
                                        C.POINTER(_i64), C.POINTER(_i64)]),
    "mggcn_gen_graph_finish": (C.c_int, [_vp, _vp, _vp, _vp]),
    "mggcn_gen_rows_begin": (C.c_int, [_i64, _i64, _i32, C.c_double, _i64, C.c_uint64, _i32, _i64, _i64, C.POINTER(_vp),
                                       C.POINTER(_i64), C.POINTER(_i64)]),
    "mggcn_gen_features_rows": (C.c_int, [_i64, _i64, _i64, _i64, C.c_uint64, _i32, _i64, _i64, _vp]),
    "mggcn_gen_features_ids": (C.c_int, [_i64, _i64, _i64, _i64, C.c_uint64, _i32, _vp, _i64, _vp]),
    "mggcn_gen_labels_rows": (C.c_int, [_i64, _i64, _i32, C.c_uint64, _i32, _i64, _i64, _vp]),
    "mggcn_gen_features": (C.c_int, [_i64, _i64, _i64, _i64, C.c_uint64, _i32, _vp]),
    "mggcn_gen_labels": (C.c_int, [_i64, _i64, _i32, C.c_uint64, _i32, _vp]),
    "mggcn_init_weights": (C.c_int, [_i64, _i64, _f32, _vp]),
    "mggcn_nccl_unique_id": (C.c_int, [_vp]),
    "mggcn_engine_create": (C.c_int, [C.POINTER(EngineConfig), C.POINTER(_vp)]),
    "mggcn_engine_comm_handle": (C.c_int, [_vp, _vp, _sz, C.POINTER(_sz)]),
    "mggcn_engine_comm_connect": (C.c_int, [_vp, _vp, _sz]),
    "mggcn_engine_set_adjacency": (C.c_int, [_vp, _i32, _vp, _vp, _vp]),
    "mggcn_engine_set_inputs": (C.c_int, [_vp, _vp, _vp, _i32]),
    "mggcn_engine_train_forward": (C.c_int, [_vp]),
    "mggcn_engine_backward": (C.c_int, [_vp]),
    "mggcn_engine_adam_update": (C.c_int, [_vp, _f32, _f32, _f32, _f32, _f32]),
    "mggcn_engine_epoch": (C.c_int, [_vp, _f32, _f32, _f32, _f32, _f32]),
    "mggcn_engine_prepare": (C.c_int, [_vp, _f32, _f32, _f32, _f32, _f32]),
    "mggcn_engine_loss": (C.c_int, [_vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "mggcn_engine_sync": (C.c_int, [_vp]),
    "mggcn_engine_stream": (_vp, [_vp, _i32]),
    "mggcn_engine_get": (C.c_int, [_vp, C.c_char_p, _vp, _sz, C.POINTER(_sz)]),
    "mggcn_engine_set": (C.c_int, [_vp, C.c_char_p, _vp, _sz]),
    "mggcn_engine_get_rows": (C.c_int, [_vp, C.c_char_p, _i64, _i64, _vp, _sz, C.POINTER(_sz)]),
    "mggcn_engine_spmm_configs": (C.c_int, [_vp, _vp, _sz, C.POINTER(_sz)]),
    "mggcn_engine_dump_timers": (C.c_int, [_vp, C.c_char_p, _vp, _sz, C.POINTER(_sz)]),
    "mggcn_engine_set_mode": (C.c_int, [_vp, _i32, _i32]),
    "mggcn_engine_launches_per_epoch": (_i64, [_vp]),
    "mggcn_engine_destroy": (None, [_vp]),
    "mggcn_bcast_stage_f32": (C.c_int, [_vp, _vp, _vp, _vp, _sz, _i32]),
}


def exported_symbols():
    return sorted(_SIGS)


def lib():
    """The loaded C-ABI.  Fails loudly when the shared library has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MggcnError(f"{LIB_PATH} is missing: run `make -C mg-gcn_b200` (or __graft_entry__.build())")
        L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        for name, (res, args) in _SIGS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise MggcnError(f"mggcn error {rc}: {lib().mggcn_last_error().decode(errors='replace')}")


def _np(a, dtype):
    a = np.ascontiguousarray(a, dtype=dtype)
    return a, a.ctypes.data_as(C.c_void_p)


# ------------------------------------------------------------------------------------------------
# host pre-processing and files (bit-exact parts of the path)
# ------------------------------------------------------------------------------------------------
class Csr:
    """Host CSR: rowptr uint64[n+1], colidx uint32[nnz], vals float32[nnz]; n rows, m columns."""

    def __init__(self, rowptr, colidx, vals, m):
        self.rowptr = np.ascontiguousarray(rowptr, dtype=np.uint64)
        self.colidx = np.ascontiguousarray(colidx, dtype=np.uint32)
        self.vals = np.ascontiguousarray(vals, dtype=np.float32)
        self.n = len(self.rowptr) - 1
        self.m = int(m)

    @property
    def nnz(self):
        return int(self.rowptr[-1])

    def copy(self):
        return Csr(self.rowptr.copy(), self.colidx.copy(), self.vals.copy(), self.m)

    def normalize(self, by_column=True):
        """In place, like csr_matrix::normalize(axis) (src/matrix.hpp:340-390)."""
        check(lib().mggcn_csr_normalize(self.n, self.m, self.rowptr.ctypes.data, self.colidx.ctypes.data,
                                        self.vals.ctypes.data, int(bool(by_column))))
        return self

    def transpose(self):
        """csr_matrix::transpose (src/matrix.hpp:392-453), sorted rows."""
        t_rowptr = np.zeros(self.m + 1, dtype=np.uint64)
        t_colidx = np.empty(self.nnz, dtype=np.uint32)
        t_vals = np.empty(self.nnz, dtype=np.float32)
        check(lib().mggcn_csr_transpose(self.n, self.m, self.rowptr.ctypes.data, self.colidx.ctypes.data,
                                        self.vals.ctypes.data, t_rowptr.ctypes.data, t_colidx.ctypes.data,
                                        t_vals.ctypes.data))
        return Csr(t_rowptr, t_colidx, t_vals, self.n)

    def block_rows(self, p, q, only_row=None):
        """dist_row_csr_matrix (src/dist_matrix.hpp:215-259): blocks[i][j] as Csr with block-local column
        ids.  only_row = i extracts just block row i (what one rank needs)."""
        p = np.ascontiguousarray(p, dtype=np.uint32)
        q = np.ascontiguousarray(q, dtype=np.uint32)
        rows_sel = range(len(p) - 1) if only_row is None else [only_row]
        Q = len(q) - 1
        out = []
        for i in rows_sel:
            pi = np.array([p[i], p[i + 1]], dtype=np.uint32)
            nr = int(pi[1] - pi[0])
            rps = [np.zeros(nr + 1, dtype=np.uint64) for _ in range(Q)]
            ptrs = (C.c_void_p * Q)(*[r.ctypes.data for r in rps])
            check(lib().mggcn_csr_block_counts(self.n, self.rowptr.ctypes.data, self.colidx.ctypes.data, 1,
                                               pi.ctypes.data, Q, q.ctypes.data, ptrs))
            cis = [np.empty(int(r[-1]), dtype=np.uint32) for r in rps]
            vas = [np.empty(int(r[-1]), dtype=np.float32) for r in rps]
            cptrs = (C.c_void_p * Q)(*[c.ctypes.data for c in cis])
            vptrs = (C.c_void_p * Q)(*[v.ctypes.data for v in vas])
            check(lib().mggcn_csr_block_fill(self.n, self.rowptr.ctypes.data, self.colidx.ctypes.data,
                                             self.vals.ctypes.data, 1, pi.ctypes.data, Q, q.ctypes.data, ptrs,
                                             cptrs, vptrs))
            out.append([Csr(rps[j], cis[j], vas[j], int(q[j + 1] - q[j])) for j in range(Q)])
        return out if only_row is None else out[0]

    def needed_rows(self, p, q):
        """prep.py compute_comm (test/data/prep.py:232-272): L[i, j] = distinct columns of block row i inside
        column block j -- the rows of H_j that part i gathers from."""
        p = np.ascontiguousarray(p, dtype=np.uint32)
        q = np.ascontiguousarray(q, dtype=np.uint32)
        L = np.zeros((len(p) - 1, len(q) - 1), dtype=np.uint64)
        check(lib().mggcn_csr_needed_rows(self.n, self.m, self.rowptr.ctypes.data, self.colidx.ctypes.data, len(p) - 1,
                                          p.ctypes.data, len(q) - 1, q.ctypes.data, L.ctypes.data))
        return L


def read_csr(path):
    info = CsrFileInfo()
    check(lib().mggcn_csr_file_info_read(os.fsencode(path), C.byref(info)))
    rowptr = np.empty(info.n + 1, dtype=np.uint64)
    colidx = np.empty(info.nnz, dtype=np.uint32)
    vals = np.empty(info.nnz, dtype=np.float32)
    check(lib().mggcn_csr_file_read(os.fsencode(path), rowptr.ctypes.data, colidx.ctypes.data, vals.ctypes.data))
    return Csr(rowptr, colidx, vals, info.ncols), info


def write_csr(path, csr, ordinal_bytes=4):
    check(lib().mggcn_csr_file_write(os.fsencode(path), csr.n, csr.m, csr.rowptr.ctypes.data,
                                     csr.colidx.ctypes.data, csr.vals.ctypes.data, ordinal_bytes))


def read_dense(path, dtype=np.float32):
    n, m = C.c_int64(), C.c_int64()
    check(lib().mggcn_dense_file_shape(os.fsencode(path), C.byref(n), C.byref(m)))
    out = np.empty((n.value, m.value), dtype=dtype)
    check(lib().mggcn_dense_file_read(os.fsencode(path), out.ctypes.data, out.dtype.itemsize))
    return out


def write_dense(path, a):
    a = np.ascontiguousarray(a)
    a2 = a.reshape(a.shape[0], -1)
    check(lib().mggcn_dense_file_write(os.fsencode(path), a2.shape[0], a2.shape[1], a2.ctypes.data, a2.dtype.itemsize))


def init_weights(rows, cols, gain=None):
    """dn_matrix::init (src/matrix.hpp:539-545)."""
    if gain is None:
        gain = float(np.float32(np.sqrt(2.0 / (1.0 + 0.01 * 0.01))))
    out = np.empty((rows, cols), dtype=np.float32)
    check(lib().mggcn_init_weights(rows, cols, gain, out.ctypes.data))
    return out


# named shapes of BASELINE.json's configs: (n_raw, directed edges, F_raw, classes, degree model, gamma, max_deg)
SHAPES = {
    "oracle4k": (4096, 4096 * 16, 64, 16, 0, 0.0, 0),
    "reddit": (232965, 114615892, 602, 41, 1, 2.6, 21657),
    "products": (2449029, 123718280, 100, 47, 1, 2.9, 17481),
    "proteins": (132534, 79122504, 8, 112, 0, 0.0, 0),
    # degree model 2 (union of k = 14 affine bijections of Z_n and their inverses, csrc/host_gen.cpp): uniform
    # degrees, a row costs O(k) on whichever rank builds it.  Walking a global sample stream (models 0 / 1) costs
    # every rank 1.6 G samples here -- minutes per rank with 4 host threads each.
    "papers100M": (111059956, 3231371744, 128, 172, 2, 0.0, 0),
}
# shapes whose generator output needs no balancing permutation (uniform degrees)
UNPERMUTED = {"papers100M"}
# hidden layers of BASELINE.json's configs
HIDDEN = {"oracle4k": [64], "reddit": [128, 128, 128], "products": [256, 256, 256], "proteins": [128, 128, 128],
          "papers100M": [128, 128, 128]}


def generate(shape, seed=1, permute=True, scale=1.0, features=True):
    """Synthetic stand-in for test/data/prep.py: returns (Csr A, X float32[n, F8], Y int32[n], classes).
    `scale` < 1 shrinks vertices and edges together (same average degree) for tests."""
    n_raw, edges, f_raw, classes, model, gamma, max_deg = shape_params(shape, scale)
    g = C.c_void_p()
    n, nnz = C.c_int64(), C.c_int64()
    check(lib().mggcn_gen_graph_begin(n_raw, edges, model, gamma, max_deg, seed, int(permute), C.byref(g),
                                      C.byref(n), C.byref(nnz)))
    rowptr = np.empty(n.value + 1, dtype=np.uint64)
    colidx = np.empty(nnz.value, dtype=np.uint32)
    vals = np.empty(nnz.value, dtype=np.float32)
    check(lib().mggcn_gen_graph_finish(g, rowptr.ctypes.data, colidx.ctypes.data, vals.ctypes.data))
    A = Csr(rowptr, colidx, vals, n.value)
    X = Y = None
    if features:
        f8 = (f_raw + 7) // 8 * 8
        X = np.empty((n.value, f8), dtype=np.float32)
        check(lib().mggcn_gen_features(n_raw, n.value, f_raw, f8, seed, int(permute), X.ctypes.data))
        Y = np.empty(n.value, dtype=np.int32)
        check(lib().mggcn_gen_labels(n_raw, n.value, classes, seed, int(permute), Y.ctypes.data))
    return A, X, Y, classes


def shape_params(shape, scale=1.0):
    n_raw, edges, f_raw, classes, model, gamma, max_deg = SHAPES[shape] if isinstance(shape, str) else shape
    if scale != 1.0:
        n_raw = max(8, int(n_raw * scale))
        edges = max(n_raw, int(edges * scale))
        max_deg = max(8, int(max_deg * scale)) if max_deg else 0
    return n_raw, edges, f_raw, classes, model, gamma, max_deg


def padded_n(shape, scale=1.0):
    return (shape_params(shape, scale)[0] + 7) // 8 * 8


def generate_rows(shape, r0, r1, seed=1, permute=True, scale=1.0, features=True):
    """Rows [r0, r1) of generate(shape, ...): (Csr slab with r1-r0 rows and GLOBAL column ids, X rows, Y rows,
    classes) -- what one rank of the 1-D row partition needs, built without the rest of the graph."""
    n_raw, edges, f_raw, classes, model, gamma, max_deg = shape_params(shape, scale)
    g = C.c_void_p()
    n, nnz = C.c_int64(), C.c_int64()
    check(lib().mggcn_gen_rows_begin(n_raw, edges, model, gamma, max_deg, seed, int(permute), r0, r1, C.byref(g),
                                     C.byref(n), C.byref(nnz)))
    rowptr = np.empty(r1 - r0 + 1, dtype=np.uint64)
    colidx = np.empty(nnz.value, dtype=np.uint32)
    vals = np.empty(nnz.value, dtype=np.float32)
    check(lib().mggcn_gen_graph_finish(g, rowptr.ctypes.data, colidx.ctypes.data, vals.ctypes.data))
    slab = Csr(rowptr, colidx, vals, n.value)
    X = Y = None
    if features:
        f8 = (f_raw + 7) // 8 * 8
        X = np.empty((r1 - r0, f8), dtype=np.float32)
        check(lib().mggcn_gen_features_rows(n_raw, n.value, f_raw, f8, seed, int(permute), r0, r1, X.ctypes.data))
        Y = np.empty(r1 - r0, dtype=np.int32)
        check(lib().mggcn_gen_labels_rows(n_raw, n.value, classes, seed, int(permute), r0, r1, Y.ctypes.data))
    return slab, X, Y, classes


def generate_feature_rows(shape, ids, seed=1, permute=True, scale=1.0):
    """Feature rows of arbitrary vertex ids of generate(shape, ...)."""
    n_raw, edges, f_raw, classes, model, gamma, max_deg = shape_params(shape, scale)
    n = (n_raw + 7) // 8 * 8
    f8 = (f_raw + 7) // 8 * 8
    ids = np.ascontiguousarray(ids, dtype=np.uint32)
    out = np.empty((len(ids), f8), dtype=np.float32)
    check(lib().mggcn_gen_features_ids(n_raw, n, f_raw, f8, seed, int(permute), ids.ctypes.data, len(ids), out.ctypes.data))
    return out


def write_dataset(dirname, A, X, Y, ordinal_bytes=4):
    """The reference's dataset directory: graph.bin, features.bin, labels.bin, sets.bin."""
    os.makedirs(dirname, exist_ok=True)
    write_csr(os.path.join(dirname, "graph.bin"), A, ordinal_bytes)
    write_dense(os.path.join(dirname, "features.bin"), X)
    write_dense(os.path.join(dirname, "labels.bin"), Y.astype(np.uint32).reshape(-1, 1))
    write_dense(os.path.join(dirname, "sets.bin"), np.zeros((len(Y), 1), dtype=np.uint32))


# ------------------------------------------------------------------------------------------------
# device entry points (torch tensors only as memory owners)
# ------------------------------------------------------------------------------------------------
def _stream(stream=None):
    import torch
    return C.c_void_p((stream or torch.cuda.current_stream()).cuda_stream)


def make_epilogue(kind=EPI_NONE, slope=0.01, bias=None, aux=None, keep_mask=None, keep_scale=1.0):
    e = Epilogue()
    e.kind, e.slope = kind, slope
    e.bias = bias.data_ptr() if bias is not None else None
    e.aux = aux.data_ptr() if aux is not None else None
    e.ld_aux = aux.stride(0) if aux is not None else 0
    e.keep_mask = keep_mask.data_ptr() if keep_mask is not None else None
    e.ld_mask = keep_mask.stride(0) if keep_mask is not None else 0
    e.keep_scale = keep_scale
    return e


class SpmmPlan:
    def __init__(self, rowptr_host, split_nnz=0, colidx_host=None, cols=None, F_hint=0, col_rot=0, col_span=0):
        """Plain plan from the row pointers; with colidx_host (+ cols, F_hint) a column-blocked plan
        (mggcn_spmm_plan_create_blocked_i32) whose block count the library chooses."""
        rp = np.ascontiguousarray(rowptr_host)
        self.rows = len(rp) - 1
        self.handle = C.c_void_p()
        if colidx_host is not None:
            rp = rp.astype(np.uint32)
            ci = np.ascontiguousarray(colidx_host, dtype=np.uint32)
            check(lib().mggcn_spmm_plan_create_blocked_i32(self.rows, cols, rp.ctypes.data, ci.ctypes.data, col_rot, col_span,
                                                           F_hint, split_nnz, C.byref(self.handle)))
        elif rp.dtype == np.uint64:
            check(lib().mggcn_spmm_plan_create_i64(self.rows, rp.ctypes.data, split_nnz, C.byref(self.handle)))
        else:
            rp = rp.astype(np.uint32)
            check(lib().mggcn_spmm_plan_create_i32(self.rows, rp.ctypes.data, split_nnz, C.byref(self.handle)))
        self._ws = None

    @property
    def blocks(self):
        return lib().mggcn_spmm_plan_blocks(self.handle)

    def workspace(self, F):
        import torch
        need = lib().mggcn_spmm_plan_workspace_bytes(self.handle, F)
        if need == 0:
            return None
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device="cuda")
        return self._ws

    def __del__(self):
        if getattr(self, "handle", None) and _lib is not None:
            _lib.mggcn_spmm_plan_destroy(self.handle)
            self.handle = None


def spmm(rowptr, colidx, vals, B, C_out, cols=None, F=None, alpha=1.0, beta=0.0, epi=None, plan=None, stream=None):
    """C = alpha*A*B + beta*C through mggcn_spmm_csr_f32_{i32,i64}; all tensors on the GPU.
    rowptr int32/int64 tensor (bit pattern of uint32/uint64), colidx int32, vals float32."""
    import torch
    rows = rowptr.numel() - 1
    F = B.shape[1] if F is None else F
    cols = B.shape[0] if cols is None else cols
    fn = lib().mggcn_spmm_csr_f32_i64 if rowptr.dtype == torch.int64 else lib().mggcn_spmm_csr_f32_i32
    ws = plan.workspace(F) if plan is not None else None
    check(fn(_stream(stream), rows, cols, rowptr.data_ptr(), colidx.data_ptr(), vals.data_ptr(), B.data_ptr(),
             B.stride(0), C_out.data_ptr(), C_out.stride(0), F, alpha, beta,
             C.byref(epi) if epi is not None else None, plan.handle if plan is not None else None,
             ws.data_ptr() if ws is not None else None))
    return C_out


def gemm(A, B, C_out, ta=False, tb=False, alpha=1.0, beta=0.0, epi=None, stream=None):
    import torch
    M = A.shape[1] if ta else A.shape[0]
    K = A.shape[0] if ta else A.shape[1]
    N = B.shape[0] if tb else B.shape[1]
    need = lib().mggcn_gemm_workspace_bytes(int(ta), int(tb), M, N, K)
    ws = torch.empty(max(need, 1), dtype=torch.uint8, device=A.device)
    check(lib().mggcn_gemm_f32(_stream(stream), int(ta), int(tb), M, N, K, alpha, A.data_ptr(), A.stride(0),
                               B.data_ptr(), B.stride(0), beta, C_out.data_ptr(), C_out.stride(0),
                               C.byref(epi) if epi is not None else None, ws.data_ptr(), need))
    return C_out


def colsum(G, stream=None):
    import torch
    rows, N = G.shape
    out = torch.empty(N, dtype=torch.float32, device=G.device)
    need = lib().mggcn_colsum_workspace_bytes(rows, N)
    ws = torch.empty(max(need, 1), dtype=torch.uint8, device=G.device)
    check(lib().mggcn_colsum_f32(_stream(stream), rows, N, G.data_ptr(), G.stride(0), out.data_ptr(), ws.data_ptr(), need))
    return out


def softmax_xent(logits, labels, n_global=None, stream=None):
    """In place on logits; returns (sum |log p_y|, #correct) as Python floats (synchronises)."""
    import torch
    rows, Cn = logits.shape
    n_global = rows if n_global is None else n_global
    out = torch.zeros(2, dtype=torch.float64, device=logits.device)
    need = lib().mggcn_softmax_xent_workspace_bytes(rows)
    ws = torch.empty(max(need, 1), dtype=torch.uint8, device=logits.device)
    check(lib().mggcn_softmax_xent_f32(_stream(stream), logits.data_ptr(), logits.stride(0), rows, Cn,
                                       labels.data_ptr(), 1.0 / float(n_global), out.data_ptr(), ws.data_ptr(), need))
    o = out.cpu()
    return float(o[0]), float(o[1])


def adam(p, g, m, v, lr, beta1, beta2, wd, eps, step, n_decay=None, stream=None):
    bc1 = float(np.float32(1 - beta1 ** step))
    bc2 = float(np.float32(1 - beta2 ** step))
    n = p.numel()
    check(lib().mggcn_adam_f32(_stream(stream), n, p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), lr, beta1,
                               beta2, wd, eps, bc1, bc2, n if n_decay is None else n_decay))


def leaky_relu_forward(x, out, slope=0.01, stream=None):
    check(lib().mggcn_leaky_relu_forward_f32(_stream(stream), x.data_ptr(), out.data_ptr(), x.numel(), slope))


def leaky_relu_backward(x, g_in, g_out, slope=0.01, stream=None):
    check(lib().mggcn_leaky_relu_backward_f32(_stream(stream), x.data_ptr(), g_in.data_ptr(), g_out.data_ptr(),
                                              x.numel(), slope))


def broadcast_rows(row, mat, discard=True, stream=None):
    check(lib().mggcn_broadcast_rows_f32(_stream(stream), row.data_ptr(), mat.data_ptr(), mat.shape[0], mat.shape[1],
                                         int(discard)))


def row_cuts(N, P):
    """p[i] = i*N/P in 64-bit (src/main.cpp:139-141)."""
    return np.array([i * N // P for i in range(P + 1)], dtype=np.uint32)


def pad_classes(c, P):
    """src/main.cpp:135: class count rounded up to a multiple of P when P > 1."""
    return c if P <= 1 else (c + P - 1) // P * P


class Engine:
    """One rank of gcn / dist_gcn<true,...> (src/gcn.hpp:937-1056) behind mggcn_engine_*."""

    def __init__(self, sizes, n_global, world=1, rank=0, device=0, cuts=None, overlap=True, exact_grads=False,
                 use_graph=False, nccl_id=None, comm=COMM_NCCL, timers=False, slope=0.01):
        self.sizes = np.ascontiguousarray(sizes, dtype=np.int64)
        self.cuts = row_cuts(n_global, world) if cuts is None else np.ascontiguousarray(cuts, dtype=np.uint32)
        self.world, self.rank, self.n_global = world, rank, n_global
        self.n_local = int(self.cuts[rank + 1] - self.cuts[rank])
        cfg = EngineConfig()
        cfg.world, cfg.rank, cfg.device = world, rank, device
        cfg.num_layers = len(self.sizes) - 1
        cfg.sizes = self.sizes.ctypes.data_as(C.POINTER(C.c_int64))
        cfg.n_global = n_global
        cfg.p = self.cuts.ctypes.data_as(C.POINTER(C.c_uint32))
        cfg.overlap, cfg.exact_grads, cfg.use_graph = int(overlap), int(exact_grads), int(use_graph)
        self._id = None
        if nccl_id is not None:
            self._id = (C.c_char * 128).from_buffer_copy(bytes(nccl_id))
            cfg.nccl_id = C.cast(self._id, C.c_void_p)
        cfg.lrelu_slope = slope
        cfg.comm_backend = comm
        cfg.timers = int(timers)
        self.handle = C.c_void_p()
        check(lib().mggcn_engine_create(C.byref(cfg), C.byref(self.handle)))

    # -- wiring
    def comm_handle(self):
        n = C.c_size_t()
        check(lib().mggcn_engine_comm_handle(self.handle, None, 0, C.byref(n)))
        buf = (C.c_char * n.value)()
        check(lib().mggcn_engine_comm_handle(self.handle, buf, n.value, C.byref(n)))
        return bytes(buf)

    def comm_connect(self, handles):
        blob = b"".join(handles)
        check(lib().mggcn_engine_comm_connect(self.handle, blob, len(handles[0])))

    def set_adjacency(self, which, blocks):
        """blocks: this rank's block row as a list of P Csr (block-local column ids)."""
        P = self.world
        assert len(blocks) == P
        rp = (C.c_void_p * P)(*[b.rowptr.ctypes.data for b in blocks])
        ci = (C.c_void_p * P)(*[b.colidx.ctypes.data for b in blocks])
        va = (C.c_void_p * P)(*[b.vals.ctypes.data for b in blocks])
        check(lib().mggcn_engine_set_adjacency(self.handle, which, rp, ci, va))

    def set_inputs(self, X_host=None, Y_host=None, async_=False):
        """X_host: float32 [n_local, F_in] (C-contiguous numpy or pinned torch CPU tensor), Y_host int32."""
        xp = yp = None
        if X_host is not None:
            xp = X_host.data_ptr() if hasattr(X_host, "data_ptr") else X_host.ctypes.data
        if Y_host is not None:
            yp = Y_host.data_ptr() if hasattr(Y_host, "data_ptr") else Y_host.ctypes.data
        check(lib().mggcn_engine_set_inputs(self.handle, xp, yp, int(async_)))

    # -- the reference's gcn interface
    def train_forward(self):
        check(lib().mggcn_engine_train_forward(self.handle))
        return self.loss()

    def backward(self):
        check(lib().mggcn_engine_backward(self.handle))

    def adam_update(self, lr=1e-2, beta1=0.9, beta2=0.999, wd=5e-4, eps=1e-8):
        check(lib().mggcn_engine_adam_update(self.handle, lr, beta1, beta2, wd, eps))

    def epoch(self, lr=1e-2, beta1=0.9, beta2=0.999, wd=5e-4, eps=1e-8):
        check(lib().mggcn_engine_epoch(self.handle, lr, beta1, beta2, wd, eps))

    def loss(self):
        """(sum |log p_y|, #correct) over this rank's rows of the last forward; blocks."""
        a, b = C.c_double(), C.c_double()
        check(lib().mggcn_engine_loss(self.handle, C.byref(a), C.byref(b)))
        return a.value, b.value

    def sync(self):
        check(lib().mggcn_engine_sync(self.handle))

    def stream(self, which=0):
        return lib().mggcn_engine_stream(self.handle, which)

    def get(self, name):
        n = C.c_size_t()
        check(lib().mggcn_engine_get(self.handle, name.encode(), None, 0, C.byref(n)))
        out = np.empty(n.value, dtype=np.float32)
        check(lib().mggcn_engine_get(self.handle, name.encode(), out.ctypes.data, n.value, C.byref(n)))
        return out

    def get_rows(self, name, r0, r1):
        n = C.c_size_t()
        check(lib().mggcn_engine_get_rows(self.handle, name.encode(), r0, r1, None, 0, C.byref(n)))
        out = np.empty(n.value, dtype=np.float32)
        check(lib().mggcn_engine_get_rows(self.handle, name.encode(), r0, r1, out.ctypes.data, n.value, C.byref(n)))
        return out.reshape(r1 - r0, -1)

    def spmm_configs(self):
        """{timer name: kernel instance + schedule} of the SpMM regions of the last eager epoch with timers."""
        n = C.c_size_t()
        check(lib().mggcn_engine_spmm_configs(self.handle, None, 0, C.byref(n)))
        buf = (C.c_char * (n.value + 1))()
        check(lib().mggcn_engine_spmm_configs(self.handle, buf, n.value + 1, C.byref(n)))
        return dict(l.split("\t", 1) for l in bytes(buf).rstrip(b"\0").decode().splitlines() if "\t" in l)

    def set(self, name, value):
        v = np.ascontiguousarray(value, dtype=np.float32).ravel()
        check(lib().mggcn_engine_set(self.handle, name.encode(), v.ctypes.data, v.size))

    def timers(self, prefix=""):
        n = C.c_size_t()
        check(lib().mggcn_engine_dump_timers(self.handle, prefix.encode(), None, 0, C.byref(n)))
        buf = (C.c_char * (n.value + 1))()
        check(lib().mggcn_engine_dump_timers(self.handle, prefix.encode(), buf, n.value + 1, C.byref(n)))
        return bytes(buf).rstrip(b"\0").decode()

    def set_mode(self, use_graph=False, timers=False):
        check(lib().mggcn_engine_set_mode(self.handle, int(use_graph), int(timers)))

    @property
    def launches_per_epoch(self):
        return lib().mggcn_engine_launches_per_epoch(self.handle)

    def close(self):
        if getattr(self, "handle", None) and _lib is not None:
            _lib.mggcn_engine_destroy(self.handle)
            self.handle = None

    def __del__(self):
        self.close()


def build_rank_engine(A, sizes, world=1, rank=0, device=0, cuts=None, **kw):
    """Host pre-processing of src/main.cpp:139-151 for one rank: normalise by column, transpose,
    P x P split, then an Engine holding this rank's block rows of A_T (forward) and A (backward).
    A is consumed (normalised in place), like the reference's csr_matrix::normalize."""
    N = A.n
    cuts = row_cuts(N, world) if cuts is None else np.ascontiguousarray(cuts, dtype=np.uint32)
    A.normalize(by_column=True)
    A_T = A.transpose()
    eng = Engine(sizes, N, world=world, rank=rank, device=device, cuts=cuts, **kw)
    eng.set_adjacency(0, A_T.block_rows(cuts, cuts, only_row=rank))
    eng.set_adjacency(1, A.block_rows(cuts, cuts, only_row=rank))
    return eng, A_T


def slab_operands(slab, deg_all, r0):
    """The two aggregation operands of one rank from its row slab of the UNnormalised graph (unit weights,
    symmetric pattern, global column ids) and every vertex's degree -- without the global normalise / transpose:
      forward  operand = rows of A_T:  A_T[i][j] = A[j][i] = 1 / colsum(i) = 1 / deg(i)     (src/main.cpp:143-151)
      backward operand = rows of A:    A[i][j]   = 1 / colsum(j) = 1 / deg(j)
    The pattern is symmetric, so both have the slab's pattern; the values are the same IEEE quotients
    csr_matrix::normalize(true) computes (1.0f / float column sum; the sum of deg ones is exact), so the blocks are
    bit-identical to normalise -> transpose -> split of the whole graph (tests/test_host.py)."""
    deg_all = np.ascontiguousarray(deg_all, dtype=np.uint32)
    n = slab.n
    inv = (np.float32(1.0) / deg_all.astype(np.float32)).astype(np.float32)
    row_len = np.diff(slab.rowptr.astype(np.int64))
    fwd_vals = np.repeat(inv[r0:r0 + n], row_len)
    bwd_vals = inv[slab.colidx]
    return Csr(slab.rowptr, slab.colidx, fwd_vals, slab.m), Csr(slab.rowptr, slab.colidx, bwd_vals, slab.m)


def build_rank_engine_sharded(slab, deg_all, sizes, world=1, rank=0, device=0, cuts=None, **kw):
    """build_rank_engine for a rank that holds only its own row slab (generate_rows) plus the degree of every
    vertex (all-gathered by the caller): no rank ever builds, normalises or transposes the whole graph."""
    N = slab.m
    cuts = row_cuts(N, world) if cuts is None else np.ascontiguousarray(cuts, dtype=np.uint32)
    r0 = int(cuts[rank])
    assert slab.n == int(cuts[rank + 1]) - r0
    fwd, bwd = slab_operands(slab, deg_all, r0)
    eng = Engine(sizes, N, world=world, rank=rank, device=device, cuts=cuts, **kw)
    local = np.array([0, slab.n], dtype=np.uint32)
    eng.set_adjacency(0, fwd.block_rows(local, cuts, only_row=0))
    eng.set_adjacency(1, bwd.block_rows(local, cuts, only_row=0))
    return eng, fwd, bwd
