"""L2 -> SM read bandwidth of the chip, measured the way the SpMM uses it: every SM streams float4 loads over a
buffer that stays L2-resident (48 MB, read 40 times by every CTA at a different phase).  The figure is the
denominator of the SECOND roofline bench.py quotes for the reddit-shape aggregation (gather traffic nnz*F*4 bytes
is served by the L2, not by HBM).  torch is the harness only; the kernel is compiled here with nvcc -> cubin and
launched through cuda-python is not needed: torch.utils.cpp_extension is avoided, a tiny .so is built instead.
Usage: python tools/l2_peak.py [--mb 48]"""
import argparse
import ctypes as C
import json
import os
import subprocess
import tempfile

SRC = r'''
#include <cuda_runtime.h>
#include <cstdint>
__global__ void __launch_bounds__(512) l2_read(const float4 *p, size_t n4, int reps, float *sink) {
    float4 acc = make_float4(0, 0, 0, 0);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int r = 0; r < reps; ++r) {
        size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x + (size_t)r * 7919 * 512) % n4;
        for (size_t k = 0; k < n4 / stride; k += 4) {
            float4 a = p[i]; i += stride; if (i >= n4) i -= n4;
            float4 b = p[i]; i += stride; if (i >= n4) i -= n4;
            float4 c = p[i]; i += stride; if (i >= n4) i -= n4;
            float4 d = p[i]; i += stride; if (i >= n4) i -= n4;
            acc.x += a.x + b.x + c.x + d.x; acc.y += a.y + b.y + c.y + d.y;
            acc.z += a.z + b.z + c.z + d.z; acc.w += a.w + b.w + c.w + d.w;
        }
    }
    if (acc.x + acc.y + acc.z + acc.w == 123.456f) *sink = acc.x;
}
extern "C" double l2_peak_gbs(size_t bytes, int reps, int iters) {
    float4 *p; float *sink;
    cudaMalloc(&p, bytes); cudaMalloc(&sink, 4); cudaMemset(p, 0, bytes);
    const size_t n4 = bytes / 16;
    int dev, sms; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = sms * 4;
    l2_read<<<grid, 512>>>(p, n4, 2, sink);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    double best = 0;
    for (int it = 0; it < iters; ++it) {
        cudaEventRecord(a);
        l2_read<<<grid, 512>>>(p, n4, reps, sink);
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        const size_t per_rep = (n4 / ((size_t)grid * 512) / 4 * 4) * (size_t)grid * 512 * 16;
        const double gbs = (double)per_rep * reps / (ms * 1e-3) / 1e9;
        if (gbs > best) best = gbs;
    }
    cudaFree(p); cudaFree(sink);
    return best;
}
'''


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, nargs="+", default=[24, 48, 60])
    a = ap.parse_args()
    d = tempfile.mkdtemp()
    cu, so = os.path.join(d, "l2.cu"), os.path.join(d, "l2.so")
    open(cu, "w").write(SRC)
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-shared", "-Xcompiler", "-fPIC", cu, "-o", so])
    L = C.CDLL(so)
    L.l2_peak_gbs.restype = C.c_double
    L.l2_peak_gbs.argtypes = [C.c_size_t, C.c_int, C.c_int]
    for mb in a.mb:
        print(json.dumps({"l2_resident_read_GBps": round(L.l2_peak_gbs(mb << 20, 40, 5), 1), "buffer_MB": mb}), flush=True)


if __name__ == "__main__":
    main()
