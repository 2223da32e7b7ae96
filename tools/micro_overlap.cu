// Does the FP64 pipe overlap with shared-memory (LSU) traffic on sm_100a?  Each warp runs either a
// DFMA stream (8 independent chains), an LDS.128 stream (conflict-free), or both interleaved.
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

#define LD(i) do { unsigned x0, x1, x2, x3; unsigned ad = sbase + (i) * 16; \
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(x0), "=r"(x1), "=r"(x2), "=r"(x3) : "r"(ad)); \
    acc ^= x0 ^ x1 ^ x2 ^ x3; } while (0)
template <int MODE>   // 0: fp64 only, 1: lds only, 2: warps alternate roles, 3: every warp interleaves
__global__ void __launch_bounds__(256) k(double *out, int iters, int nf, int nl) {
    __shared__ double4 buf[1024];
    const unsigned sbase = (unsigned)__cvta_generic_to_shared(buf);
    const int t = threadIdx.x, w = t >> 5;
    for (int i = t; i < 1024; i += blockDim.x) buf[i] = make_double4(i, 1, 2, 3);
    __syncthreads();
    double a[8];
    for (int j = 0; j < 8; j++) a[j] = 1.0 + j + t;
    double4 s = make_double4(0, 0, 0, 0); unsigned acc = 0;
    const bool do_f = MODE == 0 || MODE == 3 || (MODE == 2 && (w & 1) == 0);
    const bool do_l = MODE == 1 || MODE == 3 || (MODE == 2 && (w & 1) == 1);
    const double c1 = 1.0000001, c2 = 1e-9;
    int idx = t;
    for (int it = 0; it < iters; it++) {
        if (MODE == 3) {
#pragma unroll
            for (int r = 0; r < 8; r++) {
#pragma unroll
                for (int j = 0; j < 8; j++) a[j] = fma(a[j], c1, c2);
                LD(((idx + r * 32) & 1023));
            }
        } else {
            if (do_f) {
#pragma unroll
                for (int r = 0; r < 8; r++)
#pragma unroll
                    for (int j = 0; j < 8; j++) a[j] = fma(a[j], c1, c2);
            }
            if (do_l) {
#pragma unroll
                for (int r = 0; r < 8; r++) {
                    LD(((idx + r * 32) & 1023));
                }
            }
        }
        idx = (idx + 256) & 1023;
    }
    double r = s.x + (double)acc;
    for (int j = 0; j < 8; j++) r += a[j];
    if (r == 123.456) out[0] = r;
}

int main() {
    double *d; CK(cudaMalloc(&d, 8));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000;
    auto run = [&](int mode, int warps) {
        float best = 1e9;
        for (int rep = 0; rep < 3; rep++) {
            cudaEventRecord(e0);
            if (mode == 0) k<0><<<148, warps * 32>>>(d, iters, 0, 0);
            if (mode == 1) k<1><<<148, warps * 32>>>(d, iters, 0, 0);
            if (mode == 2) k<2><<<148, warps * 32>>>(d, iters, 0, 0);
            if (mode == 3) k<3><<<148, warps * 32>>>(d, iters, 0, 0);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
        }
        return best;
    };
    for (int warps : {4, 8}) {
        float f = run(0, warps), l = run(1, warps), alt = run(2, warps), il = run(3, warps);
        // per iteration per warp: 64 DFMA warp-instr, 8 LDS.128 (4 wavefronts each)
        printf("warps/SM %d: fp64-only %.3f ms, lds-only %.3f ms, alternate-warps(half each) %.3f ms, interleaved(all both) %.3f ms\n", warps, f, l, alt, il);
        double clk = 1.965e6;  // per ms
        printf("   fp64-only: %.2f clk per DFMA warp-instr per SMSP;  lds-only: %.2f clk per wavefront per SM\n",
               f * clk / (iters * 64.0 * warps / 4), l * clk / (iters * 32.0 * warps));
    }
    CK(cudaDeviceSynchronize());
    return 0;
}
