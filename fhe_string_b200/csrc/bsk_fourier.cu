// Upload path, SURVEY.md 8(a1) row a1.6 (Fourier-domain bootstrapping key): standard-domain BSK
// polynomial -> the two Fourier layouts the blind-rotation kernels read, computed in DOUBLE-DOUBLE
// arithmetic and rounded to f64 once at the end.
//
// Why: with q = 2^64 and 23-bit digits the f64 FFT itself is a noise source of the PBS output
// (its rounding errors are ~2^-53 of values that reach 2^91 torus units; measured: about a third of
// the output variance).  The error of a product spectrum digit^ * bsk^ is the sum of the relative
// errors of both factors plus that of the inverse transform.  The key's spectrum is computed once
// per upload, so its ~12 rounding steps can be removed for free: every value stored here is the
// correctly rounded f64 of the exact transform (error 2^-53 instead of ~sqrt(12) * 2^-53).  The
// per-ciphertext transforms in the hot loop are unchanged.
//
// Transform (same convention as bsk_transform_kernel in blind_rotate.cu, which stays for A/B runs:
// FHS_BSK_XF=0):  z_j = p_j + i p_{j+1024},  X[K] = sum_j z_j e^{2 pi i j (K + 1/4) / 1024},
// stored value = X[K] * 2^-10 (inverse FFT normalisation) * 2^-34 (accumulator unit).
#include <cmath>
#include <cstdlib>
#include <vector>

#include "fhs_internal.cuh"

namespace fhs {
namespace {

struct dd {
    double hi, lo;
};
struct cdd {
    dd re, im;
};

__device__ __forceinline__ dd dd_renorm(double s, double e) {
    const double hi = __dadd_rn(s, e);
    return dd{hi, __dadd_rn(e, -__dadd_rn(hi, -s))};
}
__device__ __forceinline__ dd dd_add(dd a, dd b) {
    const double s = __dadd_rn(a.hi, b.hi);
    const double bb = __dadd_rn(s, -a.hi);
    double e = __dadd_rn(__dadd_rn(a.hi, -__dadd_rn(s, -bb)), __dadd_rn(b.hi, -bb));
    e = __dadd_rn(e, __dadd_rn(a.lo, b.lo));
    return dd_renorm(s, e);
}
__device__ __forceinline__ dd dd_neg(dd a) { return dd{-a.hi, -a.lo}; }
__device__ __forceinline__ dd dd_mul(dd a, dd b) {
    const double p = __dmul_rn(a.hi, b.hi);
    double e = __fma_rn(a.hi, b.hi, -p);
    e = __fma_rn(a.hi, b.lo, e);
    e = __fma_rn(a.lo, b.hi, e);
    return dd_renorm(p, e);
}
__device__ __forceinline__ cdd cdd_mul(cdd a, cdd b) {
    return cdd{dd_add(dd_mul(a.re, b.re), dd_neg(dd_mul(a.im, b.im))),
               dd_add(dd_mul(a.re, b.im), dd_mul(a.im, b.re))};
}
// signed 64-bit integer -> double-double, exact
__device__ __forceinline__ dd dd_from_i64(uint64_t x) {
    const double h = (double)(int32_t)(x >> 32) * 4294967296.0;
    const double l = (double)(uint32_t)x;
    const double s = __dadd_rn(h, l);
    const double bb = __dadd_rn(s, -h);
    const double e = __dadd_rn(__dadd_rn(h, -__dadd_rn(s, -bb)), __dadd_rn(l, -bb));
    return dd{s, e};
}

constexpr int kDdThreads = 256;
// 2^-10 (inverse FFT normalisation) * 2^-34 (accumulator unit): a power of two, exact
constexpr double kStoreScale = 1.0 / 1024.0 / 17179869184.0;

// X[K] of polynomial (GGSW i, row r, column c) -> its place in the two kernel layouts
__device__ __forceinline__ void store_bsk_value(double2 *__restrict__ bsk_f, double2 *__restrict__ bsk_w,
                                                int i, int r, int c, int K, double2 val) {
    const int k1 = K >> 5, k2 = K & 31;
    const int ww = k2 >> 4, lp = (r << 4) | (k2 & 15), sl = (c == r) ? 1 : 0;
    bsk_f[((((size_t)i * 32 + k1) * 2 + ww) * 32 + lp) * 2 + sl] = val;
    const int cc = K & 127, f = 8 * (cc >> 6) + (K >> 7), qq = cc & 63;
    bsk_w[((((size_t)i * 16 + f) * 4 + (qq >> 4)) * 32 + ((r << 4) | (qq & 15))) * 2 + sl] = val;
}

// one CTA per polynomial q = ((i * 2 + r) * 2 + c): GGSW i, row r, column c
__global__ void __launch_bounds__(kDdThreads)
bsk_transform_dd_kernel(const uint64_t *__restrict__ bsk, double2 *__restrict__ bsk_f,
                        double2 *__restrict__ bsk_w, const cdd *__restrict__ twist,
                        const cdd *__restrict__ tw) {
    __shared__ cdd x[kH];
    const int q = blockIdx.x;
    const int i = q >> 2, r = (q >> 1) & 1, c = q & 1;
    const uint64_t *__restrict__ P = bsk + (size_t)q * kN;
    for (int j = threadIdx.x; j < kH; j += kDdThreads) {
        const cdd z{dd_from_i64(P[j]), dd_from_i64(P[j + kH])};
        x[__brev((unsigned)j) >> 22] = cdd_mul(z, twist[j]);       // e^{2 pi i j / 4096}
    }
    __syncthreads();
    for (int len = 2; len <= kH; len <<= 1) {
        const int half = len >> 1, step = kH / len;
        for (int b = threadIdx.x; b < kH / 2; b += kDdThreads) {
            const int j = b & (half - 1);
            const int lo = ((b - j) << 1) + j, hi = lo + half;
            const cdd a = x[lo];
            const cdd t = cdd_mul(x[hi], tw[j * step]);           // e^{+2 pi i j / len}
            x[lo] = cdd{dd_add(a.re, t.re), dd_add(a.im, t.im)};
            x[hi] = cdd{dd_add(a.re, dd_neg(t.re)), dd_add(a.im, dd_neg(t.im))};
        }
        __syncthreads();
    }
    for (int K = threadIdx.x; K < kH; K += kDdThreads) {
        // after renormalisation hi is the f64 nearest to hi + lo
        store_bsk_value(bsk_f, bsk_w, i, r, c, K, make_double2(x[K].re.hi * kStoreScale, x[K].im.hi * kStoreScale));
    }
}

// Fourier-domain import (fhs_upload_keys_fourier): the caller already holds the spectra (tfhe-rs keeps
// only the Fourier BSK inside its ServerKey), natural frequency order, un-normalised integer units.
//   conj_index = 0:  in[m] = sum_j z_j e^{i pi j / N} e^{+2 pi i j m / (N/2)}   (this library's own K)
//   conj_index = 1:  in[m] = sum_j z_j e^{i pi j / N} e^{-2 pi i j m / (N/2)}   (forward DFT with the
//                    usual negative exponent after the same twist: the same roots, visited as K = -m)
__global__ void __launch_bounds__(kDdThreads)
bsk_import_fourier_kernel(const double2 *__restrict__ in, double2 *__restrict__ bsk_f,
                          double2 *__restrict__ bsk_w, int conj_index) {
    const int q = blockIdx.x;
    const int i = q >> 2, r = (q >> 1) & 1, c = q & 1;
    for (int K = threadIdx.x; K < kH; K += kDdThreads) {
        const int m = conj_index ? ((kH - K) & (kH - 1)) : K;
        const double2 v = in[(size_t)q * kH + m];
        store_bsk_value(bsk_f, bsk_w, i, r, c, K, make_double2(v.x * kStoreScale, v.y * kStoreScale));
    }
}

void split_ld(long double v, double &hi, double &lo) {
    hi = (double)v;
    lo = (double)(v - (long double)hi);
}

}  // namespace

int launch_bsk_transform_dd(Ctx *c, const uint64_t *d_bsk_std) {
    // twiddles to 64 mantissa bits (x87 long double), split into double-double
    const long double PI = 3.14159265358979323846264338327950288L;
    std::vector<double> h(4 * (kH + kH / 2));
    for (int j = 0; j < kH; j++) {
        const long double a = 2.0L * PI * (long double)j / 4096.0L;
        split_ld(cosl(a), h[4 * j + 0], h[4 * j + 1]);
        split_ld(sinl(a), h[4 * j + 2], h[4 * j + 3]);
    }
    for (int j = 0; j < kH / 2; j++) {
        const long double a = 2.0L * PI * (long double)j / (long double)kH;
        double *o = h.data() + 4 * (kH + j);
        split_ld(cosl(a), o[0], o[1]);
        split_ld(sinl(a), o[2], o[3]);
    }
    double *d_tab = nullptr;
    FHS_CUDA(cudaMalloc(&d_tab, h.size() * sizeof(double)));
    cudaError_t e = cudaMemcpyAsync(d_tab, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) {
        const cdd *twist = reinterpret_cast<const cdd *>(d_tab);
        bsk_transform_dd_kernel<<<(unsigned)c->n * 4, kDdThreads, 0, c->stream>>>(d_bsk_std, c->d_bsk_f, c->d_bsk_w,
                                                                                   twist, twist + kH);
        c->launches++;
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);   // `h` and `d_tab` go out of scope
    cudaFree(d_tab);
    if (e != cudaSuccess) return cuda_fail(e, "bsk transform (double-double)");
    return FHS_OK;
}

int launch_bsk_import_fourier(Ctx *c, const double2 *d_bsk_fourier, int conj_index) {
    bsk_import_fourier_kernel<<<(unsigned)c->n * 4, kDdThreads, 0, c->stream>>>(d_bsk_fourier, c->d_bsk_f,
                                                                                 c->d_bsk_w, conj_index);
    c->launches++;
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "bsk import (fourier)");
    return FHS_OK;
}

}  // namespace fhs
