// Machine-peak microbenchmarks used as roofline denominators that MEASURED_PEAKS.json does not
// carry: the FP64 FMA pipe (the bound of the blind-rotation kernel).
#include "fhs_internal.cuh"

namespace fhs {
namespace {

__global__ void __launch_bounds__(256) fp64_fma_kernel(double *out, int iters, double a, double b) {
    double x[8];
#pragma unroll
    for (int j = 0; j < 8; j++) x[j] = (double)(threadIdx.x + j) * 1e-3;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 4; u++) {
#pragma unroll
            for (int j = 0; j < 8; j++) x[j] = fma(x[j], a, b);
        }
    }
    double s = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) s += x[j];
    if (s == 123456.789) out[blockIdx.x * blockDim.x + threadIdx.x] = s;  // keep the chain alive
}

}  // namespace
}  // namespace fhs

extern "C" int fhs_measure_fp64_tflops(fhs_ctx *h, double *tflops_out) {
    fhs::Ctx *c = reinterpret_cast<fhs::Ctx *>(h);
    if (!c || !tflops_out) return FHS_ERR_INVALID;
    FHS_CUDA(cudaSetDevice(c->device));
    double *d = nullptr;
    const int blocks = c->sm_count * 8, threads = 256, iters = 8192;
    FHS_CUDA(cudaMalloc(&d, (size_t)blocks * threads * sizeof(double)));
    cudaEvent_t e0, e1;
    FHS_CUDA(cudaEventCreate(&e0));
    FHS_CUDA(cudaEventCreate(&e1));
    double best = 0.0;
    for (int rep = 0; rep < 5; rep++) {
        FHS_CUDA(cudaEventRecord(e0, c->stream));
        fhs::fp64_fma_kernel<<<blocks, threads, 0, c->stream>>>(d, iters, 1.0000001, 1e-9);
        FHS_CUDA(cudaEventRecord(e1, c->stream));
        FHS_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f;
        FHS_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double flops = 2.0 * 32.0 * (double)iters * (double)blocks * threads;
        const double tf = flops / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    *tflops_out = best;
    return FHS_OK;
}
