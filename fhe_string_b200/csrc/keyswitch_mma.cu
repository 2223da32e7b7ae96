// LWE keyswitch on the integer tensor cores (large batches).
//   out[b][j] = body_b*[j==n] - sum_{i,t} digit_t(a_{b,i}) * KSK[i][t][j]          (mod 2^64)
// is a [B x 10240] x [10240 x 743] product whose left factor holds signed digits in [-4, 4] (s8) and
// whose right factor is split into its 8 little-endian BYTE limbs (u8): 8 independent s8 x u8 -> s32
// products C_limb, recombined exactly as sum_limb C_limb << 8*limb (|C_limb| <= 10240*4*255 < 2^31).
// The byte matrix of the KSK is its own memory image, re-tiled once at upload:
//   Kt[n_tile][kb][t][n (128 byte columns)][32 mask elements]   (20 KB per (n_tile, kb) block)
// so one K-block is ONE bulk async copy (cp.async.bulk + mbarrier, the TMA 1-D path).  Digits are
// produced in registers from the ciphertext masks and staged in shared memory.  MMA: legacy
// mma.sync m16n8k32 s8*u8 (IMMA); CTA tile 128 ciphertexts x 128 byte columns (16 u64 columns),
// 8 warps of 64 x 32.  Rows within a tile are XOR-swizzled by 16 B to keep fragment loads
// conflict-free.  Integer arithmetic only: bit-exact against the oracle.
#include "fhs_internal.cuh"

namespace fhs {
namespace {

constexpr int kMT = 128, kNT = 128;     // CTA tile: ciphertexts x byte columns
constexpr int kKI = 32;                 // mask elements per K-block
constexpr int kLv = 5, kBl = 3;         // ks_level, ks_base_log
constexpr int kStageBytes = kLv * kNT * kKI;  // 20480
constexpr int kThreadsM = 256;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ int tile_off(int t, int row, int kk) {
    return t * (kNT * kKI) + row * kKI + ((((kk >> 4) ^ ((row >> 2) & 1)) & 1) << 4) + (kk & 15);
}

// signed digits (least significant first) of the closest 15-bit representable of x
__device__ __forceinline__ void signed_digits(uint64_t x, int (&d)[kLv]) {
    uint32_t state = (uint32_t)(((x >> 49) + ((x >> 48) & 1ull)) & 0x7FFFull);
#pragma unroll
    for (int t = 0; t < kLv; t++) {
        uint32_t res = state & 7u;
        state >>= kBl;
        uint32_t carry = (((res - 1u) | state) & res) >> (kBl - 1);
        state += carry;
        d[t] = (int)res - (int)(carry << kBl);
    }
}

__device__ __forceinline__ void mma_s8u8(int (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

__global__ void __launch_bounds__(kThreadsM, 1)
keyswitch_mma_kernel(const uint64_t *__restrict__ in, const uint8_t *__restrict__ kt,
                     uint64_t *__restrict__ out, unsigned int batch, int n_small, int big, int n_kb_total) {
    // split-K over blockIdx.z (small batches: fill the SMs; partial sums are combined with u64
    // atomics into a zeroed output, the body term is added by the first slice)
    const int kb_base = (int)(((long long)blockIdx.z * n_kb_total) / gridDim.z);
    const int n_kb = (int)(((long long)(blockIdx.z + 1) * n_kb_total) / gridDim.z) - kb_base;
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char *Bs = smem;                       // 2 stages
    unsigned char *As = smem + 2 * kStageBytes;     // 2 stages
    __shared__ __align__(8) uint64_t mbar[2];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, tg = lane & 3;
    const int wm = warp >> 2, wn = warp & 3;        // warp tile: rows wm*64.., byte cols wn*32..
    const unsigned int m0 = blockIdx.y * kMT;
    const int n_tile = blockIdx.x;
    const size_t in_w = (size_t)big + 1;
    const unsigned char *kt_tile = kt + ((size_t)n_tile * n_kb_total + kb_base) * kStageBytes;

    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar[0])));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar[1])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto issue_b = [&](int kb) {
        const int st = kb & 1;
        const uint32_t bar = smem_u32(&mbar[st]);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(kStageBytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(Bs + st * kStageBytes)), "l"(kt_tile + (size_t)kb * kStageBytes),
                       "r"(kStageBytes), "r"(bar) : "memory");
    };
    // A production: item q = tid + 256*r (r < 4): row b = q >> 3, mask elements 4*(q & 7) .. +3
    uint64_t pre[4][4];
    auto load_a = [&](int kb) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const int q = tid + kThreadsM * r, b = q >> 3, i4 = (q & 7) * 4;
            const unsigned int ct = m0 + b;
            if (ct < batch) {
                const uint64_t *p = in + (size_t)ct * in_w + (size_t)(kb_base + kb) * kKI + i4;
#pragma unroll
                for (int e = 0; e < 4; e++) pre[r][e] = __ldg(p + e);
            } else {
#pragma unroll
                for (int e = 0; e < 4; e++) pre[r][e] = 0ull;
            }
        }
    };
    auto store_a = [&](int kb) {
        unsigned char *A = As + (kb & 1) * kStageBytes;
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const int q = tid + kThreadsM * r, b = q >> 3, i4 = (q & 7) * 4;
            uint32_t packed[kLv] = {0, 0, 0, 0, 0};
#pragma unroll
            for (int e = 0; e < 4; e++) {
                int d[kLv];
                signed_digits(pre[r][e], d);
#pragma unroll
                for (int t = 0; t < kLv; t++) packed[t] |= ((uint32_t)d[t] & 0xFFu) << (8 * e);
            }
#pragma unroll
            for (int t = 0; t < kLv; t++) *reinterpret_cast<uint32_t *>(A + tile_off(t, b, i4)) = packed[t];
        }
    };

    int acc[4][4][4];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 4; b++)
#pragma unroll
            for (int c = 0; c < 4; c++) acc[a][b][c] = 0;

    if (tid == 0) issue_b(0);
    load_a(0);
    store_a(0);
    if (n_kb > 1) load_a(1);
    __syncthreads();

    for (int kb = 0; kb < n_kb; kb++) {
        if (kb + 1 < n_kb) {
            if (tid == 0) issue_b(kb + 1);
            store_a(kb + 1);
            if (kb + 2 < n_kb) load_a(kb + 2);
        }
        {   // wait for the bulk copy of this K-block
            const uint32_t bar = smem_u32(&mbar[kb & 1]);
            const uint32_t parity = (uint32_t)((kb >> 1) & 1);
            uint32_t done = 0;
            while (!done) {
                asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                             : "=r"(done) : "r"(bar), "r"(parity) : "memory");
            }
        }
        const unsigned char *A = As + (kb & 1) * kStageBytes;
        const unsigned char *Bt = Bs + (kb & 1) * kStageBytes;
#pragma unroll
        for (int t = 0; t < kLv; t++) {
            uint32_t af[4][4], bf[4][2];
#pragma unroll
            for (int mi = 0; mi < 4; mi++) {
                const int r0 = wm * 64 + mi * 16 + g;
                af[mi][0] = *reinterpret_cast<const uint32_t *>(A + tile_off(t, r0, 4 * tg));
                af[mi][1] = *reinterpret_cast<const uint32_t *>(A + tile_off(t, r0 + 8, 4 * tg));
                af[mi][2] = *reinterpret_cast<const uint32_t *>(A + tile_off(t, r0, 16 + 4 * tg));
                af[mi][3] = *reinterpret_cast<const uint32_t *>(A + tile_off(t, r0 + 8, 16 + 4 * tg));
            }
#pragma unroll
            for (int ni = 0; ni < 4; ni++) {
                const int c0 = wn * 32 + ni * 8 + g;
                bf[ni][0] = *reinterpret_cast<const uint32_t *>(Bt + tile_off(t, c0, 4 * tg));
                bf[ni][1] = *reinterpret_cast<const uint32_t *>(Bt + tile_off(t, c0, 16 + 4 * tg));
            }
#pragma unroll
            for (int mi = 0; mi < 4; mi++)
#pragma unroll
                for (int ni = 0; ni < 4; ni++) mma_s8u8(acc[mi][ni], af[mi], bf[ni]);
        }
        __syncthreads();
    }

    // epilogue: one n8 tile = the 8 limbs of one u64 column; lanes tg hold limbs 2tg, 2tg+1
    const int out_w = n_small + 1;
#pragma unroll
    for (int mi = 0; mi < 4; mi++) {
#pragma unroll
        for (int ni = 0; ni < 4; ni++) {
            const int j = (n_tile * kNT + wn * 32 + ni * 8) >> 3;
#pragma unroll
            for (int h = 0; h < 2; h++) {
                uint64_t v = ((uint64_t)(int64_t)acc[mi][ni][2 * h] << (16 * tg)) +
                             ((uint64_t)(int64_t)acc[mi][ni][2 * h + 1] << (16 * tg + 8));
                v += __shfl_xor_sync(0xffffffffu, v, 1);
                v += __shfl_xor_sync(0xffffffffu, v, 2);
                const unsigned int ct = m0 + wm * 64 + mi * 16 + g + 8 * h;
                if (tg == 0 && ct < batch && j < out_w) {
                    const uint64_t body = (j == n_small && blockIdx.z == 0) ? in[(size_t)ct * in_w + big] : 0ull;
                    if (gridDim.z == 1) out[(size_t)ct * out_w + j] = body - v;
                    else atomicAdd(reinterpret_cast<unsigned long long *>(out + (size_t)ct * out_w + j),
                                   (unsigned long long)(body - v));
                }
            }
        }
    }
}

// internal KSK [i][t][j] u64 -> Kt tiles (see header comment); bytes of columns >= n+1 are zero
__global__ void ksk_tile_kernel(const uint64_t *__restrict__ ksk, uint8_t *__restrict__ kt, int out_w,
                                int n_kb, size_t total) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    // decode destination: (((n_tile*n_kb + kb)*5 + t)*128 + n)*32 + pos
    const int pos = (int)(idx & 31);
    size_t r = idx >> 5;
    const int n = (int)(r % kNT); r /= kNT;
    const int t = (int)(r % kLv); r /= kLv;
    const int kb = (int)(r % n_kb);
    const int n_tile = (int)(r / n_kb);
    // un-swizzle pos -> kk
    const int half = ((pos >> 4) ^ ((n >> 2) & 1)) & 1;
    const int kk = (half << 4) | (pos & 15);
    const int i = kb * kKI + kk;
    const int ng = n_tile * kNT + n;
    const int j = ng >> 3, limb = ng & 7;
    uint8_t v = 0;
    if (j < out_w) v = (uint8_t)(ksk[((size_t)i * kLv + t) * out_w + j] >> (8 * limb));
    kt[idx] = v;
}

}  // namespace

size_t ksk_tiles_bytes(const Ctx *c) {
    const int n_tiles = (int)(((size_t)(c->n + 1) * 8 + kNT - 1) / kNT);
    return (size_t)n_tiles * (c->big / kKI) * kStageBytes;
}

int launch_ksk_tile(Ctx *c) {
    const size_t total = ksk_tiles_bytes(c);
    ksk_tile_kernel<<<(unsigned)((total + 255) / 256), 256, 0, c->stream>>>(c->d_ksk, c->d_ksk_t, (int)c->n + 1,
                                                                            (int)(c->big / kKI), total);
    c->launches++;
    FHS_CUDA(cudaGetLastError());
    return FHS_OK;
}

int launch_keyswitch_mma(Ctx *c, const uint64_t *d_in, size_t batch, uint64_t *d_small) {
    const int smem_bytes = 4 * kStageBytes;
    if (!c->attr_ks) {
        FHS_CUDA(cudaFuncSetAttribute(keyswitch_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
        c->attr_ks = true;
    }
    const int n_tiles = (int)(((size_t)(c->n + 1) * 8 + kNT - 1) / kNT);
    const unsigned m_tiles = (unsigned)((batch + kMT - 1) / kMT);
    unsigned ksplit = 1;
    while (ksplit < 8 && n_tiles * m_tiles * (ksplit + 1) <= (unsigned)c->sm_count) ksplit++;
    if (ksplit > 1)
        FHS_CUDA(cudaMemsetAsync(d_small, 0, batch * ((size_t)c->n + 1) * sizeof(uint64_t), c->stream));
    dim3 grid(n_tiles, m_tiles, ksplit);
    keyswitch_mma_kernel<<<grid, kThreadsM, smem_bytes, c->stream>>>(d_in, c->d_ksk_t, d_small, (unsigned int)batch,
                                                                      (int)c->n, (int)c->big, (int)(c->big / kKI));
    c->launches++;
    FHS_CUDA(cudaGetLastError());
    return FHS_OK;
}

}  // namespace fhs
