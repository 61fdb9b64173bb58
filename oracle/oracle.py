"""ctypes view of oracle/mggcn_oracle.c -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module.  The product package (mg-gcn_b200) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")
_SRC = os.path.join(_HERE, "mggcn_oracle.c")


def build(force=False):
    """gcc the C restatement (portable x86-64-v3 so the .so also runs on the GPU box's host)."""
    if not force and os.path.exists(_SO) and os.path.getmtime(_SO) >= os.path.getmtime(_SRC):
        return _SO
    subprocess.check_call(
        ["gcc", "-O3", "-march=x86-64-v3", "-fopenmp", "-fPIC", "-shared", "-std=c99",
         "-Wall", "-Wextra", _SRC, "-o", _SO, "-lm"])
    return _SO


_lib = None

u64p = np.ctypeslib.ndpointer(np.uint64, flags="C_CONTIGUOUS")
u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        L.orc_threads.restype = C.c_int
        L.orc_normalize.argtypes = [C.c_int64, C.c_int64, u64p, u32p, f32p, C.c_int]
        L.orc_transpose.argtypes = [C.c_int64, C.c_int64, u64p, u32p, f32p, u64p, u32p, f32p]
        L.orc_block.argtypes = [u64p, u32p, f32p, C.c_int64, C.c_int64, C.c_uint32, C.c_uint32,
                                u64p, C.c_void_p, C.c_void_p]
        L.orc_csr_to_dense.argtypes = [C.c_int64, C.c_int64, u64p, u32p, f32p, f32p]
        L.orc_spmm.argtypes = [C.c_int64, u64p, u32p, f32p, C.c_void_p, C.c_int64, C.c_void_p,
                               C.c_int64, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int,
                               C.c_void_p]
        L.orc_gemm.argtypes = [C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_float,
                               C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_float, C.c_void_p,
                               C.c_int64]
        L.orc_leaky_relu_forward.argtypes = [f32p, f32p, C.c_size_t, C.c_float]
        L.orc_leaky_relu_backward.argtypes = [f32p, f32p, f32p, C.c_size_t, C.c_float]
        L.orc_broadcast_rows.argtypes = [f32p, f32p, C.c_size_t, C.c_size_t, C.c_int]
        L.orc_softmax_xent.argtypes = [f32p, C.c_int64, C.c_int, i32p, C.c_int64,
                                       C.POINTER(C.c_float), C.POINTER(C.c_float)]
        L.orc_adam.argtypes = [C.c_size_t, f32p, f32p, f32p, f32p, C.c_float, C.c_float, C.c_float,
                               C.c_float, C.c_float, C.c_int]
        L.orc_init_weights.argtypes = [C.c_int64, C.c_int64, C.c_float, f32p]
        L.orc_gcn_create.restype = C.c_void_p
        L.orc_gcn_create.argtypes = [C.c_int, i64p, C.c_int64, C.c_int, C.c_void_p, u64p, u32p,
                                     f32p, u64p, u32p, f32p, C.c_int]
        L.orc_gcn_destroy.argtypes = [C.c_void_p]
        L.orc_gcn_train_forward.argtypes = [C.c_void_p, f32p, i32p, C.POINTER(C.c_float),
                                            C.POINTER(C.c_float)]
        L.orc_gcn_backward.argtypes = [C.c_void_p]
        L.orc_gcn_adam.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float]
        L.orc_gcn_tensor.restype = C.POINTER(C.c_float)
        L.orc_gcn_tensor.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_int64)]
        _lib = L
    return _lib


def threads():
    return lib().orc_threads()


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def normalize(rowptr, colidx, vals, cols, axis=True):
    vals = np.array(vals, dtype=np.float32, copy=True)
    lib().orc_normalize(len(rowptr) - 1, cols, rowptr, colidx, vals, int(bool(axis)))
    return vals


def transpose(rowptr, colidx, vals, cols):
    rows = len(rowptr) - 1
    t_rowptr = np.zeros(cols + 1, np.uint64)
    t_colidx = np.zeros(len(colidx), np.uint32)
    t_vals = np.zeros(len(colidx), np.float32)
    lib().orc_transpose(rows, cols, rowptr, colidx, vals, t_rowptr, t_colidx, t_vals)
    return t_rowptr, t_colidx, t_vals


def block(rowptr, colidx, vals, r0, r1, c0, c1):
    rp = np.zeros(r1 - r0 + 1, np.uint64)
    lib().orc_block(rowptr, colidx, vals, r0, r1, c0, c1, rp, None, None)
    ci = np.zeros(int(rp[-1]), np.uint32)
    va = np.zeros(int(rp[-1]), np.float32)
    lib().orc_block(rowptr, colidx, vals, r0, r1, c0, c1, rp, _ptr(ci), _ptr(va))
    return rp, ci, va


def csr_to_dense(rowptr, colidx, vals, cols):
    rows = len(rowptr) - 1
    out = np.zeros((rows, cols), np.float32)
    lib().orc_csr_to_dense(rows, cols, rowptr, colidx, vals, out)
    return out


def spmm(rowptr, colidx, vals, B, C_in=None, alpha=1.0, beta=0.0, acc64=False, cuts=None, F=None):
    """C = alpha*A*B + beta*C.  B may be wider than F (ldb = B.shape[1])."""
    rows = len(rowptr) - 1
    B = np.ascontiguousarray(B, np.float32)
    F = B.shape[1] if F is None else F
    Cm = np.zeros((rows, F), np.float32) if C_in is None else np.array(C_in, np.float32, copy=True)
    ldc = Cm.shape[1]
    if cuts is None:
        nblk, cp = 1, None
    else:
        cuts = np.ascontiguousarray(cuts, np.uint32)
        nblk, cp = len(cuts) - 1, _ptr(cuts)
    lib().orc_spmm(rows, rowptr, colidx, vals, _ptr(B), B.shape[1], _ptr(Cm), ldc, F,
                   alpha, beta, int(acc64), nblk, cp)
    return Cm


def gemm(A, B, ta=False, tb=False, alpha=1.0, beta=0.0, C_in=None):
    A = np.ascontiguousarray(A, np.float32)
    B = np.ascontiguousarray(B, np.float32)
    M, K = (A.shape[1], A.shape[0]) if ta else A.shape
    N = B.shape[0] if tb else B.shape[1]
    Cm = np.zeros((M, N), np.float32) if C_in is None else np.array(C_in, np.float32, copy=True)
    lib().orc_gemm(int(ta), int(tb), M, N, K, alpha, _ptr(A), A.shape[1], _ptr(B), B.shape[1],
                   beta, _ptr(Cm), N)
    return Cm


def leaky_relu_forward(x, alpha=0.01):
    x = np.ascontiguousarray(x, np.float32)
    out = np.empty_like(x)
    lib().orc_leaky_relu_forward(x.reshape(-1), out.reshape(-1), x.size, alpha)
    return out


def leaky_relu_backward(x, g, alpha=0.01):
    x = np.ascontiguousarray(x, np.float32)
    g = np.ascontiguousarray(g, np.float32)
    out = np.empty_like(g)
    lib().orc_leaky_relu_backward(x.reshape(-1), g.reshape(-1), out.reshape(-1), x.size, alpha)
    return out


def broadcast_rows(row, rows):
    row = np.ascontiguousarray(row, np.float32).reshape(-1)
    mat = np.zeros((rows, row.size), np.float32)
    lib().orc_broadcast_rows(row, mat.reshape(-1), rows, row.size, 1)
    return mat


def softmax_xent(logits, labels, n_global=None):
    """Returns (loss_sum, correct, G) with G = (softmax - onehot) / n_global."""
    G = np.array(logits, np.float32, copy=True)
    labels = np.ascontiguousarray(labels, np.int32).reshape(-1)
    rows, Cc = G.shape
    ls, cs = C.c_float(), C.c_float()
    lib().orc_softmax_xent(G.reshape(-1), rows, Cc, labels, rows if n_global is None else n_global,
                           C.byref(ls), C.byref(cs))
    return ls.value, cs.value, G


def adam(p, g, m, v, lr, beta1, beta2, wd, eps, step):
    """In place on float32 contiguous arrays (g is modified like the reference does)."""
    lib().orc_adam(p.size, p.reshape(-1), g.reshape(-1), m.reshape(-1), v.reshape(-1), lr, beta1,
                   beta2, wd, eps, step)


def init_weights(rows, cols, gain=None):
    if gain is None:
        gain = np.float32(np.sqrt(2 / (1 + 0.01 * 0.01)))
    out = np.zeros((rows, cols), np.float32)
    lib().orc_init_weights(rows, cols, float(gain), out.reshape(-1))
    return out


class Gcn:
    """orc_gcn: the reference's gcn / dist_gcn<true> as one serial host program."""

    def __init__(self, sizes, N, fwd, bwd, P=1, cuts=None, exact_grads=False):
        self.sizes = np.ascontiguousarray(sizes, np.int64)
        self.N, self.P = N, P
        self._keep = (fwd, bwd)  # the C side borrows these arrays
        cp = None
        if cuts is not None:
            self._cuts = np.ascontiguousarray(cuts, np.uint32)
            cp = _ptr(self._cuts)
        self._h = lib().orc_gcn_create(len(sizes) - 1, self.sizes, N, P, cp, fwd[0], fwd[1], fwd[2],
                                       bwd[0], bwd[1], bwd[2], int(exact_grads))

    def train_forward(self, X, Y):
        self._X = np.ascontiguousarray(X, np.float32)
        self._Y = np.ascontiguousarray(Y, np.int32).reshape(-1)
        loss, acc = C.c_float(), C.c_float()
        lib().orc_gcn_train_forward(self._h, self._X.reshape(-1), self._Y, C.byref(loss), C.byref(acc))
        return loss.value, acc.value

    def backward(self):
        lib().orc_gcn_backward(self._h)

    def adam_update(self, lr=1e-2, beta1=0.9, beta2=0.999, wd=5e-4, eps=1e-8):
        lib().orc_gcn_adam(self._h, lr, beta1, beta2, wd, eps)

    def tensor(self, name):
        n = C.c_int64()
        p = lib().orc_gcn_tensor(self._h, name.encode(), C.byref(n))
        if not p:
            raise KeyError(name)
        return np.ctypeslib.as_array(p, shape=(n.value,)).copy()

    def set_tensor(self, name, value):
        n = C.c_int64()
        p = lib().orc_gcn_tensor(self._h, name.encode(), C.byref(n))
        if not p:
            raise KeyError(name)
        v = np.ascontiguousarray(value, np.float32).reshape(-1)
        assert v.size == n.value
        np.ctypeslib.as_array(p, shape=(n.value,))[:] = v

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_gcn_destroy(self._h)
            self._h = None
