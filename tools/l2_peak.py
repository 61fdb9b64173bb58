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
// every thread keeps 8 independent 16-byte loads in flight; a warp reads 512 contiguous bytes per load (one row of
// B at F = 128), rows are visited in a scrambled order so consecutive warps do not walk DRAM pages in step
__global__ void __launch_bounds__(256) l2_read(const float4 *p, uint32_t rows_mask, int reps, float *sink) {
    float4 acc = make_float4(0, 0, 0, 0);
    const uint32_t lane = threadIdx.x & 31;
    uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nw = (gridDim.x * blockDim.x) >> 5;
    for (int r = 0; r < reps; ++r) {
        for (uint32_t k = w; k <= rows_mask; k += 8 * nw) {
            float4 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const uint32_t row = ((k + u * nw) * 2654435761u) & rows_mask;   // odd multiplier: a permutation of the rows
                v[u] = p[(size_t)row * 32 + lane];
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) { acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w; }
        }
    }
    if (acc.x + acc.y + acc.z + acc.w == 123.456f) *sink = acc.x;
}
extern "C" double l2_peak_gbs(size_t bytes, int reps, int iters) {
    // rows of 512 bytes, a power-of-two count
    uint32_t rows = 1; while ((size_t)rows * 2 * 512 <= bytes) rows *= 2;
    float4 *p; float *sink;
    cudaMalloc(&p, (size_t)rows * 512); cudaMalloc(&sink, 4); cudaMemset(p, 0, (size_t)rows * 512);
    int dev, sms; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = sms * 8;
    l2_read<<<grid, 256>>>(p, rows - 1, 2, sink);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    double best = 0;
    for (int it = 0; it < iters; ++it) {
        cudaEventRecord(a);
        l2_read<<<grid, 256>>>(p, rows - 1, reps, sink);
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        const double gbs = (double)rows * 512.0 * reps / (ms * 1e-3) / 1e9;
        if (gbs > best) best = gbs;
    }
    cudaFree(p); cudaFree(sink);
    return best;
}
'''


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, nargs="+", default=[16, 32, 64, 128, 1024])
    a = ap.parse_args()
    d = tempfile.mkdtemp()
    cu, so = os.path.join(d, "l2.cu"), os.path.join(d, "l2.so")
    open(cu, "w").write(SRC)
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-shared", "-Xcompiler", "-fPIC", cu, "-o", so])
    L = C.CDLL(so)
    L.l2_peak_gbs.restype = C.c_double
    L.l2_peak_gbs.argtypes = [C.c_size_t, C.c_int, C.c_int]
    for mb in a.mb:
        print(json.dumps({"read_GBps": round(L.l2_peak_gbs(mb << 20, max(2, 2048 // mb), 5), 1), "buffer_MB": mb,
                          "what": "512-byte rows in scrambled order, 8 loads in flight per thread, all SMs; <= 64 MB stays L2-resident"}), flush=True)


if __name__ == "__main__":
    main()
