// LWE keyswitch on the 5th-generation tensor cores (tcgen05.mma kind::i8, accumulators in TMEM).
//   out[b][j] = body_b*[j==n] - sum_{i,t} digit_t(a_{b,i}) * KSK[i][t][j]          (mod 2^64)
// Same arithmetic as keyswitch_mma.cu (signed digits s8 x byte limbs u8 -> s32, recombined exactly:
// |C_limb| <= 10240*4*255 < 2^31), restructured for the UMMA programming model:
//   * a prologue kernel decomposes the masks ONCE into digit tiles At[m_tile][kb][t][128 x 32] s8
//     (keyswitch_mma.cu recomputes them per column tile);
//   * the KSK byte limbs are tiled at upload as Ku[n_tile][kb][t][256 x 32] u8;
//   * both tile kinds are stored in global memory as the exact shared-memory image of a K-major,
//     non-swizzled UMMA operand (8-row x 16-byte core matrices; the two 16-byte K halves 128 B apart,
//     8-row groups 256 B apart), so a pipeline stage is two bulk async copies (cp.async.bulk + mbarrier);
//   * one CTA = 128 ciphertexts x 256 byte columns (32 u64 columns); warp 0 feeds a 3-stage ring,
//     one thread of warp 1 issues 5 MMAs (M128 N256 K32) per stage and commits them to the stage's
//     "empty" barrier, warps 2-5 read the 256 s32 accumulator columns back with tcgen05.ld and
//     recombine the 8 limbs of every output coefficient.
// Integer arithmetic only: bit-exact against the oracle (tests/test_gpu_parity.py).
#include <mutex>

#include "fhs_internal.cuh"
#include "tmem.cuh"

namespace fhs {
namespace {

constexpr int kUM = 128, kUN = 256, kUK = 32;   // UMMA shape (kind::i8: K = 32)
constexpr int kLv = 5, kBl = 3;                 // ks_level, ks_base_log
constexpr int kAStage = kLv * kUM * kUK;        // 20480 B of digits per K-block
constexpr int kBStage = kLv * kUN * kUK;        // 40960 B of key bytes per K-block
constexpr int kStages = 3;
constexpr int kThreadsU = 192;
constexpr int kTmemCols = 256;

__host__ __device__ __forceinline__ int canon_off(int row, int k) {
    return ((row >> 3) << 8) + ((k >> 4) << 7) + ((row & 7) << 4) + (k & 15);
}
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

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

// masks -> digit tiles.  grid (n_kb, m_tiles), 256 threads: item q = tid + 256 r: row q >> 3, mask
// elements 4 (q & 7) .. +3 of K-block kb.  Rows past the batch are zero.  zero_out / zero_words: see below.
__global__ void __launch_bounds__(256)
ks_digits_kernel(const uint64_t *__restrict__ in, uint8_t *__restrict__ at, unsigned int batch, int big, int n_kb,
                 uint64_t *__restrict__ zero_out, size_t zero_words) {
    const int kb = blockIdx.x, m_tile = blockIdx.y, tid = threadIdx.x;
    // split-K keyswitch (small batches): the partial sums are added into `out` atomically, so it starts at zero
    for (size_t i = ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * 256 + tid; i < zero_words;
         i += (size_t)gridDim.x * gridDim.y * 256)
        zero_out[i] = 0ull;
    uint8_t *dst = at + ((size_t)m_tile * n_kb + kb) * kAStage;
    const size_t in_w = (size_t)big + 1;
#pragma unroll
    for (int r = 0; r < 4; r++) {
        const int q = tid + 256 * r, b = q >> 3, i4 = (q & 7) * 4;
        const unsigned int ct = (unsigned int)m_tile * kUM + b;
        uint32_t packed[kLv] = {0, 0, 0, 0, 0};
        if (ct < batch) {
            const uint64_t *p = in + (size_t)ct * in_w + (size_t)kb * kUK + i4;
#pragma unroll
            for (int e = 0; e < 4; e++) {
                int d[kLv];
                signed_digits(__ldg(p + e), d);
#pragma unroll
                for (int t = 0; t < kLv; t++) packed[t] |= ((uint32_t)d[t] & 0xFFu) << (8 * e);
            }
        }
#pragma unroll
        for (int t = 0; t < kLv; t++)
            *reinterpret_cast<uint32_t *>(dst + t * (kUM * kUK) + canon_off(b, i4)) = packed[t];
    }
}

// internal KSK [i][t][j] u64 -> Ku tiles; bytes of columns >= n+1 are zero
__global__ void ksk_tile_umma_kernel(const uint64_t *__restrict__ ksk, uint8_t *__restrict__ ku, int out_w,
                                     int n_kb, size_t total) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const int off = (int)(idx % (kUN * kUK));
    size_t r = idx / (kUN * kUK);
    const int t = (int)(r % kLv); r /= kLv;
    const int kb = (int)(r % n_kb);
    const int n_tile = (int)(r / n_kb);
    const int n = ((off >> 8) << 3) | ((off >> 4) & 7);
    const int kk = (((off >> 7) & 1) << 4) | (off & 15);
    const int i = kb * kUK + kk;
    const int ng = n_tile * kUN + n;
    const int j = ng >> 3, limb = ng & 7;
    uint8_t v = 0;
    if (j < out_w) v = (uint8_t)(ksk[((size_t)i * kLv + t) * out_w + j] >> (8 * limb));
    ku[idx] = v;
}

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    while (!done) {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    }
}

// K-major, no swizzle: start address, leading (K-half) byte offset 128, stride (8-row group) byte
// offset 256, descriptor version 1 (sm_100)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
    return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)(128u >> 4) << 16) | ((uint64_t)(256u >> 4) << 32) |
           (1ull << 46);
}
// instruction descriptor: D = s32, A = s8 (K-major), B = u8 (K-major), N = 256, M = 128
constexpr uint32_t kIdesc = (2u << 4) | (1u << 7) | (0u << 10) | ((uint32_t)(kUN >> 3) << 17) | ((uint32_t)(kUM >> 4) << 24);

__global__ void __launch_bounds__(kThreadsU, 1)
keyswitch_umma_kernel(const uint8_t *__restrict__ at, const uint8_t *__restrict__ ku,
                      const uint64_t *__restrict__ in, uint64_t *__restrict__ out, unsigned int batch,
                      int n_small, int big, int n_kb, int ksplit) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) uint64_t full_bar[kStages], empty_bar[kStages], accum_bar;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n_tile = blockIdx.x, m_tile = blockIdx.y;
    // split-K (levels of a few ciphertexts: 24 column tiles would leave 124 SMs idle, each streaming 3.8 MB of key
    // at one SM's L2 bandwidth): CTA z of ksplit takes K-blocks [kb0, kb1) and adds its partial sums atomically
    // (wrapping u64 adds commute, so the result is the same bits whatever the order)
    const int kb0 = (int)(((long)blockIdx.z * n_kb) / ksplit), kb1 = (int)(((long)(blockIdx.z + 1) * n_kb) / ksplit);
    const int my_kb = kb1 - kb0;
    unsigned char *As = smem;                          // kStages x 20 KB
    unsigned char *Bs = smem + kStages * kAStage;      // kStages x 40 KB

    if (tid == 0) {
        for (int s = 0; s < kStages; s++) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&full_bar[s])));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&empty_bar[s])));
        }
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&accum_bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) tmem_alloc(&tmem_base_s, kTmemCols);
    tmem_fence_before_sync();
    __syncthreads();
    tmem_fence_after_sync();
    const uint32_t tmem_base = tmem_base_s;

    if (warp == 0) {
        // producer: one bulk copy per operand per K-block
        const uint8_t *a_src = at + (size_t)m_tile * n_kb * kAStage;
        const uint8_t *b_src = ku + (size_t)n_tile * n_kb * kBStage;
        for (int kk = 0; kk < my_kb; kk++) {
            const int kb = kb0 + kk;
            const int s = kk % kStages, it = kk / kStages;
            if (it > 0) mbar_wait(smem_u32(&empty_bar[s]), (uint32_t)((it - 1) & 1));
            if (lane == 0) {
                const uint32_t bar = smem_u32(&full_bar[s]);
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(kAStage + kBStage) : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(smem_u32(As + s * kAStage)), "l"(a_src + (size_t)kb * kAStage), "r"(kAStage), "r"(bar) : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(smem_u32(Bs + s * kBStage)), "l"(b_src + (size_t)kb * kBStage), "r"(kBStage), "r"(bar) : "memory");
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // MMA issuer: 5 levels per K-block, each M128 N256 K32, accumulating into TMEM columns 0..255
        for (int kk = 0; kk < my_kb; kk++) {
            const int s = kk % kStages, it = kk / kStages;
            mbar_wait(smem_u32(&full_bar[s]), (uint32_t)(it & 1));
            tmem_fence_after_sync();
            if (lane == 0) {
                const uint32_t a0 = smem_u32(As + s * kAStage), b0 = smem_u32(Bs + s * kBStage);
#pragma unroll
                for (int t = 0; t < kLv; t++) {
                    const uint64_t da = umma_desc(a0 + t * (kUM * kUK)), db = umma_desc(b0 + t * (kUN * kUK));
                    const uint32_t accumulate = (kk > 0 || t > 0) ? 1u : 0u;
                    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
                                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}"
                                 ::"r"(tmem_base), "l"(da), "l"(db), "r"(kIdesc), "r"(accumulate) : "memory");
                }
                // frees the smem slot once these MMAs have read it (implies fence::before_thread_sync)
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                             ::"r"(smem_u32(&empty_bar[s])) : "memory");
                if (kk == my_kb - 1)
                    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                                 ::"r"(smem_u32(&accum_bar)) : "memory");
            }
            __syncwarp();
        }
    } else {
        // epilogue: thread <-> TMEM lane <-> ciphertext row; 8 consecutive columns = the limbs of one u64
        const int quarter = warp & 3;
        const int row = quarter * 32 + lane;
        const unsigned int ct = (unsigned int)m_tile * kUM + row;
        const int out_w = n_small + 1;
        const size_t in_w = (size_t)big + 1;
        mbar_wait(smem_u32(&accum_bar), 0u);
        tmem_fence_after_sync();
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16);
#pragma unroll 1
        for (int c4 = 0; c4 < kUN / 32; c4++) {
            uint32_t r[32];
            tmem_ld32(taddr + c4 * 32, r);
            tmem_wait_ld();
#pragma unroll
            for (int e = 0; e < 4; e++) {
                uint64_t v = 0;
#pragma unroll
                for (int l = 0; l < 8; l++) v += (uint64_t)(int64_t)(int32_t)r[e * 8 + l] << (8 * l);
                const int j = n_tile * (kUN / 8) + c4 * 4 + e;
                if (ct < batch && j < out_w) {
                    const uint64_t body = (j == n_small && blockIdx.z == 0) ? in[(size_t)ct * in_w + big] : 0ull;
                    if (ksplit == 1) out[(size_t)ct * out_w + j] = body - v;
                    else atomicAdd(reinterpret_cast<unsigned long long *>(out + (size_t)ct * out_w + j), (unsigned long long)(body - v));
                }
            }
        }
    }
    tmem_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

}  // namespace

size_t ksk_umma_tiles_bytes(const Ctx *c) {
    const int n_tiles = (int)(((size_t)(c->n + 1) * 8 + kUN - 1) / kUN);
    return (size_t)n_tiles * (c->big / kUK) * kBStage;
}

int launch_ksk_tile_umma(Ctx *c) {
    const size_t total = ksk_umma_tiles_bytes(c);
    ksk_tile_umma_kernel<<<(unsigned)((total + 255) / 256), 256, 0, c->stream>>>(c->d_ksk, c->d_ksk_u, (int)c->n + 1,
                                                                                 (int)(c->big / kUK), total);
    c->launches++;
    FHS_CUDA(cudaGetLastError());
    return FHS_OK;
}

int launch_keyswitch_umma(Ctx *c, const uint64_t *d_in, size_t batch, uint64_t *d_small) {
    const int smem_bytes = kStages * (kAStage + kBStage);
    if (!c->attr_ksu) {
        FHS_CUDA(cudaFuncSetAttribute(keyswitch_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
        c->attr_ksu = true;
    }
    const int n_kb = (int)(c->big / kUK);
    const unsigned m_tiles = (unsigned)((batch + kUM - 1) / kUM);
    const size_t need = (size_t)m_tiles * n_kb * kAStage;
    if (need > c->cap_digits) {
        // grown outside the steady state only (first call at a larger batch)
        FHS_CUDA(cudaStreamSynchronize(c->stream));
        cudaFree(c->d_ks_digits);
        c->d_ks_digits = nullptr;
        c->cap_digits = 0;
        FHS_CUDA(cudaMalloc(&c->d_ks_digits, need));
        c->cap_digits = need;
    }
    const int n_tiles = (int)(((size_t)(c->n + 1) * 8 + kUN - 1) / kUN);
    // split-K while the column x row tiles alone would leave more than half of the SMs idle
    int ksplit = 1;
    if (c->knobs.ks_splitk) {
        ksplit = c->sm_count / (n_tiles * (int)m_tiles);
        if (ksplit > 8) ksplit = 8;
        if (ksplit < 2) ksplit = 1;
    }
    ks_digits_kernel<<<dim3(n_kb, m_tiles), 256, 0, c->stream>>>(d_in, c->d_ks_digits, (unsigned int)batch, (int)c->big, n_kb,
                                                                 d_small, ksplit > 1 ? batch * ((size_t)c->n + 1) : 0);
    c->launches++;
    FHS_CUDA(cudaGetLastError());
    keyswitch_umma_kernel<<<dim3(n_tiles, m_tiles, ksplit), kThreadsU, smem_bytes, c->stream>>>(
        c->d_ks_digits, c->d_ksk_u, d_in, d_small, (unsigned int)batch, (int)c->n, (int)c->big, n_kb, ksplit);
    c->launches++;
    FHS_CUDA(cudaGetLastError());
    return FHS_OK;
}

}  // namespace fhs

// ---- allocation-free KS-PBS (include/fhs_cuda.h fhs_ks_pbs_batch_dev_noalloc) ----------------------------
// For callers that may not allocate (a stream under graph capture, a serving loop, a harness that forbids
// cudaMalloc): the two scratch arrays of the path -- the digit tiles of the tensor-core keyswitch and the
// keyswitched small LWE ciphertexts handed to the blind rotation -- live in file-scope device arrays sized
// for FHS_NOALLOC_MAX_BATCH ciphertexts instead of the context's grow-on-demand buffers.  Every GPU that
// loads the library holds them (6.3 MB) until exit.  Both arrays are overwritten in full before they are
// read (ks_digits_kernel zero-fills the rows past the batch; the keyswitch writes every word of every row
// below it), so nothing carries over from the previous call; but all calls on a GPU share them: calls that
// overlap on the device (two streams, two contexts) race.
namespace fhs {
namespace {
constexpr size_t kNoAllocMTiles = (FHS_NOALLOC_MAX_BATCH + kUM - 1) / kUM;
constexpr size_t kNoAllocKb = 2048 / kUK;                       // K-blocks of the big LWE mask (k N = 2048)
constexpr size_t kNoAllocSmallW = 768;                          // >= n + 1 (keyswitch.cu: n + 1 <= 768)
static __device__ __align__(256) uint8_t g_noalloc_digits[kNoAllocMTiles * kNoAllocKb * kAStage];
static __device__ __align__(256) uint64_t g_noalloc_small[(size_t)FHS_NOALLOC_MAX_BATCH * kNoAllocSmallW];

// The arrays' names are host-side stand-ins here: the device addresses come from cudaGetSymbolAddress,
// once per device, with that device current.
constexpr int kNoAllocMaxDev = 64;
struct NoAllocAddr { uint8_t *digits; uint64_t *small; bool ok; };
NoAllocAddr g_noalloc_addr[kNoAllocMaxDev];
std::mutex g_noalloc_mu;

int noalloc_addresses(int device, uint8_t **digits, uint64_t **small) {
    if (device < 0 || device >= kNoAllocMaxDev) { set_error("device %d out of range", device); return FHS_ERR_INVALID; }
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    NoAllocAddr &a = g_noalloc_addr[device];
    if (!a.ok) {
        FHS_CUDA(cudaSetDevice(device));
        FHS_CUDA(cudaGetSymbolAddress(reinterpret_cast<void **>(&a.digits), g_noalloc_digits));
        FHS_CUDA(cudaGetSymbolAddress(reinterpret_cast<void **>(&a.small), g_noalloc_small));
        a.ok = true;
    }
    *digits = a.digits;
    *small = a.small;
    return FHS_OK;
}
}  // namespace
}  // namespace fhs

extern "C" int fhs_ks_pbs_batch_dev_noalloc(fhs_ctx *h, const uint64_t *d_in, const uint32_t *d_lut_id,
                                            const uint64_t *d_tables, size_t n_luts, size_t batch, uint64_t *d_out) {
    using namespace fhs;
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return FHS_ERR_INVALID;
    if (!c->has_keys) { set_error("keys not uploaded"); return FHS_ERR_NO_KEYS; }
    if (batch == 0) return FHS_OK;
    if (!d_in || !d_lut_id || !d_tables || !d_out || n_luts == 0) { set_error("null buffer or empty LUT table"); return FHS_ERR_INVALID; }
    if (batch > (size_t)FHS_NOALLOC_MAX_BATCH) {
        set_error("fhs_ks_pbs_batch_dev_noalloc: batch %zu exceeds the static scratch (FHS_NOALLOC_MAX_BATCH = %d); "
                  "split the batch or use fhs_ks_pbs_batch_dev", batch, FHS_NOALLOC_MAX_BATCH);
        return FHS_ERR_INVALID;
    }
    if (c->big != kNoAllocKb * kUK || (size_t)c->n + 1 > kNoAllocSmallW || c->p.ks_level != kLv || c->p.ks_base_log != kBl) {
        set_error("fhs_ks_pbs_batch_dev_noalloc: static scratch is sized for k N = 2048, n + 1 <= 768, KS 2^3 x 5");
        return FHS_ERR_INVALID;
    }
    FHS_CUDA(cudaSetDevice(c->device));
    uint8_t *digits = nullptr;
    uint64_t *small = nullptr;
    int rc = noalloc_addresses(c->device, &digits, &small);
    if (rc != FHS_OK) return rc;
    const int smem_bytes = kStages * (kAStage + kBStage);
    if (!c->attr_ksu) {
        FHS_CUDA(cudaFuncSetAttribute(keyswitch_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
        c->attr_ksu = true;
    }
    const int n_kb = (int)kNoAllocKb;
    const unsigned m_tiles = (unsigned)((batch + kUM - 1) / kUM);
    ks_digits_kernel<<<dim3(n_kb, m_tiles), 256, 0, c->stream>>>(d_in, digits, (unsigned int)batch, (int)c->big, n_kb, nullptr, 0);
    c->launches++;
    FHS_CUDA(cudaGetLastError());
    const int n_tiles = (int)(((size_t)(c->n + 1) * 8 + kUN - 1) / kUN);
    keyswitch_umma_kernel<<<dim3(n_tiles, m_tiles), kThreadsU, smem_bytes, c->stream>>>(
        digits, c->d_ksk_u, d_in, small, (unsigned int)batch, (int)c->n, (int)c->big, n_kb, 1);
    c->launches++;
    FHS_CUDA(cudaGetLastError());
    return launch_blind_rotate(c, small, d_lut_id, d_tables, n_luts, batch, d_out);
}
