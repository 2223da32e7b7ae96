// LWE keyswitch (tfhe-rs keyswitch_lwe_ciphertext + SignedDecomposer{base_log 3, level 5};
// SURVEY.md 8(a1) row a1.4, Appendix A.3/A.4) as a batched mod-2^64 integer matrix product:
//     out[b] = (0, ..., 0, body_b) - sum_{i < kN} sum_{t < L} digit_{b,i,t} * KSK[i][t][.]
// Digits are produced in registers from the top base_log*L bits of each mask element and staged in
// shared memory once per CTA tile; every thread keeps a (ciphertexts x columns) tile of u64
// accumulators in registers and streams the KSK rows straight from L2/HBM with coalesced loads, so
// one KSK pass serves kKsTb ciphertexts.  Digits are offset to [0, 2^base_log] so each MAC is
// IMAD.WIDE.U32 + IMAD; the offset is removed with a per-column constant computed at upload.
// Integer arithmetic only: bit-exact against the oracle.
#include <cstdlib>

#include "fhs_internal.cuh"

namespace fhs {
namespace {

constexpr int kKsTb = 28;        // largest ciphertext tile per CTA (4096 / 28 -> 147 CTAs = one wave of 148 SMs)
constexpr int kKsThreads = 256;
constexpr int kKsCols = 3;       // columns per thread: t, t + 256, t + 512  (n + 1 <= 768)
constexpr int kKsIc = 16;        // mask elements per shared-memory digit chunk
constexpr int kKsLevels = 5;
constexpr int kKsBaseLog = 3;

// Signed decomposition, least significant digit first, returned offset by +2^(base_log-1).
__device__ __forceinline__ void ks_digits(uint64_t x, uint32_t (&d)[kKsLevels]) {
    constexpr int bits = kKsBaseLog * kKsLevels;
    constexpr int r = 64 - bits;
    uint64_t st = ((x >> r) + ((x >> (r - 1)) & 1ull)) & ((1ull << bits) - 1ull);
    uint32_t state = (uint32_t)st;
#pragma unroll
    for (int t = 0; t < kKsLevels; t++) {
        uint32_t res = state & ((1u << kKsBaseLog) - 1u);
        state >>= kKsBaseLog;
        uint32_t carry = (((res - 1u) | state) & res) >> (kKsBaseLog - 1);
        state += carry;
        // digit = res - (carry << base_log) in [-2^(b-1), 2^(b-1)]; store digit + 2^(b-1)
        d[t] = res + (1u << (kKsBaseLog - 1)) - (carry << kKsBaseLog);
    }
}

template <int TB>
__global__ void __launch_bounds__(kKsThreads, 1)
keyswitch_kernel(const uint64_t *__restrict__ in, const uint64_t *__restrict__ ksk,
                 const uint64_t *__restrict__ corr, uint64_t *__restrict__ out, unsigned int batch,
                 int n_small, int big, int cts_per_cta) {
    __shared__ __align__(16) uint32_t dig[kKsIc][kKsLevels][TB];
    // blockIdx.y = slice of the mask range (split-K for small batches: more CTAs stream the KSK)
    const int t = threadIdx.x;
    const int ksplit = gridDim.y;
    const int ic_begin = (int)((long long)big * blockIdx.y / ksplit);
    const int ic_end = (int)((long long)big * (blockIdx.y + 1) / ksplit);
    const unsigned int c0 = blockIdx.x * cts_per_cta;
    const int nb = (int)min((unsigned int)cts_per_cta, batch - c0);
    const int out_w = n_small + 1;
    const size_t in_w = (size_t)big + 1;

    uint64_t acc[TB][kKsCols];
#pragma unroll
    for (int b = 0; b < TB; b++)
#pragma unroll
        for (int c = 0; c < kKsCols; c++) acc[b][c] = 0ull;

    bool col_ok[kKsCols];
#pragma unroll
    for (int c = 0; c < kKsCols; c++) col_ok[c] = (t + c * kKsThreads) < out_w;

    for (int ic = ic_begin; ic < ic_end; ic += kKsIc) {
        // ---- stage digits of mask elements [ic, ic+16) for the tile's ciphertexts ----
        for (int p = t; p < kKsIc * TB; p += kKsThreads) {
            const int b = p / kKsIc, ii = p % kKsIc;
            uint32_t d[kKsLevels];
            if (b < nb) {
                ks_digits(in[(size_t)(c0 + b) * in_w + ic + ii], d);
            } else {
#pragma unroll
                for (int l = 0; l < kKsLevels; l++) d[l] = 0u;
            }
#pragma unroll
            for (int l = 0; l < kKsLevels; l++) dig[ii][l][b] = d[l];
        }
        __syncthreads();
#pragma unroll 1
        for (int ii = 0; ii < kKsIc; ii++) {
            const uint64_t *__restrict__ rowp = ksk + (size_t)(ic + ii) * kKsLevels * out_w + t;
            uint64_t kv[kKsLevels][kKsCols];
#pragma unroll
            for (int l = 0; l < kKsLevels; l++)
#pragma unroll
                for (int c = 0; c < kKsCols; c++)
                    kv[l][c] = col_ok[c] ? __ldg(rowp + (size_t)l * out_w + c * kKsThreads) : 0ull;
#pragma unroll
            for (int l = 0; l < kKsLevels; l++) {
#pragma unroll
                for (int b4 = 0; b4 < TB / 4; b4++) {
                    const uint4 d = *reinterpret_cast<const uint4 *>(&dig[ii][l][4 * b4]);
#pragma unroll
                    for (int c = 0; c < kKsCols; c++) {
                        acc[4 * b4 + 0][c] += (uint64_t)d.x * kv[l][c];
                        acc[4 * b4 + 1][c] += (uint64_t)d.y * kv[l][c];
                        acc[4 * b4 + 2][c] += (uint64_t)d.z * kv[l][c];
                        acc[4 * b4 + 3][c] += (uint64_t)d.w * kv[l][c];
                    }
                }
            }
        }
        __syncthreads();
    }

#pragma unroll
    for (int c = 0; c < kKsCols; c++) {
        const int j = t + c * kKsThreads;
        if (j < out_w) {
            const uint64_t cr = corr[j];
#pragma unroll
            for (int b = 0; b < TB; b++) {
                if (b < nb) {
                    const uint64_t body = (j == n_small) ? in[(size_t)(c0 + b) * in_w + big] : 0ull;
                    if (ksplit == 1) {
                        out[(size_t)(c0 + b) * out_w + j] = body - (acc[b][c] - cr);
                    } else {  // out was zeroed by the launcher; slice 0 also carries body and offset
                        const uint64_t v = (blockIdx.y == 0 ? body + cr : 0ull) - acc[b][c];
                        atomicAdd(reinterpret_cast<unsigned long long *>(out + (size_t)(c0 + b) * out_w + j),
                                  (unsigned long long)v);
                    }
                }
            }
        }
    }
}

// raw KSK [i][level rows][n+1] -> internal [i][digit index t (least significant first)][n+1]
__global__ void ksk_permute_kernel(const uint64_t *__restrict__ raw, uint64_t *__restrict__ dst,
                                   size_t rows, int out_w, int levels, int level_order) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= rows * (size_t)out_w) return;
    const size_t row = idx / out_w;
    const int j = (int)(idx % out_w);
    const size_t i = row / levels;
    const int t = (int)(row % levels);
    // digit t (t = 0 least significant) multiplies level (levels - t); MSB-first storage keeps
    // level l at row l-1, LSB-first storage keeps it at row levels-l.
    const int src_row = (level_order == FHS_KSK_LEVEL_MSB_FIRST) ? (levels - 1 - t) : t;
    dst[idx] = raw[(i * levels + src_row) * (size_t)out_w + j];
}

// corr[j] = 2^(base_log-1) * sum over all rows of ksk[row][j]   (mod 2^64)
__global__ void ksk_corr_kernel(const uint64_t *__restrict__ ksk, uint64_t *__restrict__ corr,
                                size_t rows, int out_w) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= out_w) return;
    uint64_t s = 0;
    for (size_t r = 0; r < rows; r++) s += ksk[r * out_w + j];
    corr[j] = s << (kKsBaseLog - 1);
}

}  // namespace

int launch_ksk_prepare(Ctx *c, const uint64_t *d_ksk_raw, int level_order) {
    const size_t rows = (size_t)c->big * c->p.ks_level;
    const int out_w = (int)c->n + 1;
    const size_t total = rows * out_w;
    ksk_permute_kernel<<<(unsigned)((total + 255) / 256), 256, 0, c->stream>>>(
        d_ksk_raw, c->d_ksk, rows, out_w, (int)c->p.ks_level, level_order);
    c->launches++;
    FHS_CUDA(cudaGetLastError());
    return FHS_OK;
}

int launch_ksk_corr(Ctx *c) {
    const size_t rows = (size_t)c->big * c->p.ks_level;
    const int out_w = (int)c->n + 1;
    ksk_corr_kernel<<<(out_w + 127) / 128, 128, 0, c->stream>>>(c->d_ksk, c->d_ksk_corr, rows, out_w);
    c->launches++;
    FHS_CUDA(cudaGetLastError());
    return FHS_OK;
}

int launch_keyswitch(Ctx *c, const uint64_t *d_in, size_t batch, uint64_t *d_small) {
    if (batch == 0) return FHS_OK;
    if ((int)c->n + 1 > kKsCols * kKsThreads || c->p.ks_level != kKsLevels ||
        c->p.ks_base_log != kKsBaseLog || (c->big % kKsIc) != 0) {
        set_error("keyswitch kernel built for ks_base_log=3, ks_level=5, n+1<=768");
        return FHS_ERR_INVALID;
    }
    // FHS_KS_VARIANT: 3 (default) = tcgen05 kernel (keyswitch_umma.cu; faster at every batch size:
    // 0.04 ms for <= 128 ciphertexts, 0.23 ms for 4096) from FHS_KS_UMMA_MIN (default 1) ciphertexts up and
    // the mma.sync kernel (keyswitch_mma.cu, split-K) below; 2 = mma.sync only; 1 = IMAD kernel
    // (knobs are per context, read in fhs_ctx_create; the key forms of the two baseline kernels are only
    // resident when the context was created with FHS_KS_VARIANT != 3 or FHS_KS_UMMA_MIN > 1)
    const int variant = c->knobs.ks_variant;
    if (variant == 3 && batch >= (size_t)c->knobs.ks_umma_min) return launch_keyswitch_umma(c, d_in, batch, d_small);
    if (!c->knobs.ks_baseline_keys) { set_error("baseline keyswitch kernels need FHS_KS_VARIANT set at context creation"); return FHS_ERR_INVALID; }
    const bool use_mma = variant == 2 || variant == 3;
    if (use_mma && batch >= (size_t)c->knobs.ks_mma_min) return launch_keyswitch_mma(c, d_in, batch, d_small);
    // tile size: fill the SMs first, then grow tiles up to kKsTb (KSK traffic ~ 1/tile)
    int per = (int)((batch + c->sm_count - 1) / c->sm_count);
    if (per < 1) per = 1;
    if (per > kKsTb) per = kKsTb;
    const unsigned tiles = (unsigned)((batch + per - 1) / per);
    unsigned ksplit = 1;
    while (ksplit < 32 && tiles * ksplit * 2 <= (unsigned)c->sm_count) ksplit *= 2;
    if (ksplit > 1)
        FHS_CUDA(cudaMemsetAsync(d_small, 0, batch * ((size_t)c->n + 1) * sizeof(uint64_t), c->stream));
    const dim3 grid(tiles, ksplit);
    if (per <= 4)
        keyswitch_kernel<4><<<grid, kKsThreads, 0, c->stream>>>(
            d_in, c->d_ksk, c->d_ksk_corr, d_small, (unsigned int)batch, (int)c->n, (int)c->big, per);
    else if (per <= 12)
        keyswitch_kernel<12><<<grid, kKsThreads, 0, c->stream>>>(
            d_in, c->d_ksk, c->d_ksk_corr, d_small, (unsigned int)batch, (int)c->n, (int)c->big, per);
    else
        keyswitch_kernel<kKsTb><<<grid, kKsThreads, 0, c->stream>>>(
            d_in, c->d_ksk, c->d_ksk_corr, d_small, (unsigned int)batch, (int)c->n, (int)c->big, per);
    c->launches++;
    FHS_CUDA(cudaGetLastError());
    return FHS_OK;
}

}  // namespace fhs
