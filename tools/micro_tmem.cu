// TMEM read path on sm_100a as this repo uses it (lane-private scratch, tcgen05.ld.32x32b): bytes per clock
// per SM and the round-trip time of ld -> wait::ld, for 1..8 warps per SM, alone and next to a DFMA stream.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o micro_tmem micro_tmem.cu && ./micro_tmem
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ void ld16(unsigned ta, unsigned (&r)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                   "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]) : "r"(ta));
}
__device__ __forceinline__ void ld32(unsigned ta, unsigned (&r)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                   "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]),
                   "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]),
                   "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]) : "r"(ta));
}
__device__ __forceinline__ void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// MODE 0: x16, wait after every load (round trip).  1: x32 round trip.  2: x32, 4 loads in flight per wait.
// 3: x16 round trip with 16 independent DFMA between load and wait.  4: DFMA only (16 per step).
template <int MODE>
__global__ void __launch_bounds__(256) k(unsigned *out, int iters, long long *clk) {
    __shared__ unsigned tbase_s;
    const int w = threadIdx.x >> 5;
    if (w == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"((unsigned)__cvta_generic_to_shared(&tbase_s)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tb = tbase_s + (((unsigned)(w & 3) * 32u) << 16);
    unsigned acc = 0;
    double a[16];
    for (int j = 0; j < 16; j++) a[j] = 1.0 + j + threadIdx.x;
    const double c1 = 1.0000001, c2 = 1e-9;
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
        if (MODE == 0 || MODE == 3) {
            unsigned r[16];
            ld16(tb + 16 * (it & 15), r);
            if (MODE == 3) {
#pragma unroll
                for (int j = 0; j < 16; j++) a[j] = fma(a[j], c1, c2);
            }
            wait_ld();
#pragma unroll
            for (int j = 0; j < 16; j++) acc ^= r[j];
        } else if (MODE == 1) {
            unsigned r[32];
            ld32(tb + 32 * (it & 7), r);
            wait_ld();
#pragma unroll
            for (int j = 0; j < 32; j++) acc ^= r[j];
        } else if (MODE == 2) {
            unsigned r0[32], r1[32], r2[32], r3[32];
            ld32(tb, r0); ld32(tb + 32, r1); ld32(tb + 64, r2); ld32(tb + 96, r3);
            wait_ld();
#pragma unroll
            for (int j = 0; j < 32; j++) acc ^= r0[j] ^ r1[j] ^ r2[j] ^ r3[j];
        } else {
#pragma unroll
            for (int j = 0; j < 16; j++) a[j] = fma(a[j], c1, c2);
        }
    }
    const long long t1 = clock64();
    double s = 0;
    for (int j = 0; j < 16; j++) s += a[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc + (unsigned)s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (w == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tbase_s) : "memory");
}

template <int MODE>
int run(const char *name, int warps, int bytes_per_step, unsigned *d, long long *dclk) {
    const int iters = 20000;
    k<MODE><<<148, 32 * warps>>>(d, 10, dclk);
    k<MODE><<<148, 32 * warps>>>(d, iters, dclk);
    CK(cudaDeviceSynchronize());
    long long c;
    CK(cudaMemcpy(&c, dclk, 8, cudaMemcpyDeviceToHost));
    const double per = (double)c / iters;
    printf("  %-44s warps/SM %d: %7.1f clk per step per warp", name, warps, per);
    if (bytes_per_step) printf("   %6.1f B/clk/SM", (double)bytes_per_step * warps / per);
    printf("\n");
    return 0;
}

int main() {
    unsigned *d; long long *dclk;
    CK(cudaMalloc(&d, 148 * 256 * 4)); CK(cudaMalloc(&dclk, 8));
    for (int warps : {1, 4, 8}) {
        run<0>("ld.x16 -> wait (2 KB per step)", warps, 2048, d, dclk);
        run<1>("ld.x32 -> wait (4 KB per step)", warps, 4096, d, dclk);
        run<2>("4 x ld.x32 -> wait (16 KB per step)", warps, 16384, d, dclk);
        run<3>("ld.x16, 16 DFMA, wait", warps, 2048, d, dclk);
        run<4>("16 DFMA", warps, 0, d, dclk);
    }
    return 0;
}
