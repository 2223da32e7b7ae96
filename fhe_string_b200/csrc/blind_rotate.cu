// Blind rotation (tfhe-rs blind_rotate_assign / cmux / add_external_product_assign; SURVEY.md
// 8(a1) rows a1.1, a1.5, a1.6, a1.7) for k = 1, N = 2048, one PBS level, as one sm_100a kernel:
//   modulus switch -> LUT accumulator (closed form, already rotated by X^{-b~}) -> n CMUX
//   iterations on an f64 negacyclic FFT -> sample extract.
//
// FFT design ("32x32 four-step, one warp per polynomial"):
//   A real polynomial p of N = 2048 coefficients is folded to 1024 complex points
//   z_j = p_j + i p_{j+1024} and evaluated at the roots x_k = e^{i pi (4k+1)/N} of X^N + 1:
//       X[k] = sum_j z_j e^{2 pi i j (k + 1/4)/1024},   j = n1 + 32 n2,  k = 32 k1 + k2.
//   Column pass (lane = n1, 32 values n2 in registers): shifted 32-point DFT = constant pre-twist
//   c_{n2} = e^{2 pi i n2/128} + plain DFT-32 codelet.  Twiddle T[k2][n1] = e^{2 pi i n1 (k2+1/4)/1024}.
//   Transpose through shared memory.  Row pass (lane = k2, 32 values n1 in registers): DFT-32.
//   The negacyclic twist is therefore folded into compile-time constants and the twiddle table.
//   The inverse runs the same passes backwards with conjugate factors; 1/1024 lives in the BSK.
//
// One ciphertext = 2 warps (64 threads): warp w owns accumulator polynomial w (0 = mask, 1 = body)
// for the coefficient-domain work and the column passes; for the row passes warp w owns
// frequencies k2 in [16w, 16w+16) of BOTH polynomials (lanes 0-15: poly 0, lanes 16-31: poly 1), so
// the 2x2 GGSW contraction needs only one shfl_xor(16) exchange per frequency.
//
// A CTA holds kSlots2 ciphertext slots that share the twiddle table; the two warps of a slot
// synchronise with a named barrier (64 threads); CTAs are persistent and stride over slot groups.
#include <cstdlib>

#include "fhs_internal.cuh"
#include "fft32_codelets.cuh"
#include "tmem.cuh"

namespace fhs {
namespace {

constexpr int kTile = 32 * kTStride;  // complex elements in one padded 32x32 tile

__constant__ double kPreRe[32] = FHS_PRETWIST_RE;
__constant__ double kPreIm[32] = FHS_PRETWIST_IM;

__device__ __forceinline__ void slot_bar(int slot) {
    asm volatile("bar.sync %0, 64;" ::"r"(slot + 1) : "memory");
}

// a1.5 pbs_modulus_switch for N = 2048: ((x >> 51) + 1) >> 1, reduced mod 2N.
__device__ __forceinline__ uint32_t modswitch(uint64_t x) {
    return (uint32_t)((((x >> 51) + 1ull) >> 1) & (uint64_t)(2 * kN - 1));
}

// a1.1 + initial rotation (SURVEY.md A.8 closed form): coefficient j of X^{-bt} * LUT_f.
__device__ __forceinline__ uint64_t lut_coeff(const uint64_t *__restrict__ tab, uint32_t j,
                                              uint32_t bt, int log_box, uint64_t delta) {
    const uint32_t v = (j + bt + (1u << (log_box - 1))) & (2 * kN - 1);
    const uint64_t f = tab[(v & (kN - 1)) >> log_box] * delta;
    return (v & kN) ? (0ull - f) : f;
}

// Column pass, forward: inputs in BR5 slots -> slot k2 holds Y'[n1 = lane][k2].
__device__ __forceinline__ void fwd_cols(double (&xr)[32], double (&xi)[32],
                                         const double2 *__restrict__ T, int lane) {
    fft32_pos(xr, xi);
#pragma unroll
    for (int k2 = 0; k2 < 32; k2++) {
        const double2 t = T[k2 * kTStride + lane];
        const double a = xr[k2], b = xi[k2];
        xr[k2] = a * t.x - b * t.y;
        xi[k2] = a * t.y + b * t.x;
    }
}

// =================================================================================================
// v2: "lock-step CTA with TMEM": 4 ciphertext slots (8 compute warps) + 4 loader warps per CTA.
//   * The accumulator lives in Tensor Memory: every compute thread owns 64 coefficients of its
//     warp's polynomial as 128 lane-private 32-bit columns (64 high words, 64 low words).  Shared
//     memory keeps only the HIGH 32 bits of each coefficient (the "hi plane", 16 KB per ciphertext)
//     for the rotated reads of (X^a - 1) * acc: the signed digit needs the top 23 bits, and the
//     dropped borrow perturbs the value being rounded by < 2^-32 (torus), far below the 2^-24
//     rounding step of the decomposition itself (noise-bounded tier).
//   * The Fourier BSK row of iteration i is shared by the 4 ciphertexts of the CTA: loader warp q
//     (one per TMEM lane quarter) streams the 32 KB slice of role w = q & 1 from L2 into registers
//     and parks it in TMEM columns [256, 512) of its quarter; compute warps pull 4 frequency rows
//     (32 columns) at a time with tcgen05.ld.  Named barriers kBarFull / kBarFree hand the single
//     buffer back and forth; the loads of slice i+1 are already in flight while slice i is consumed.
//   * All compute warps advance through i in lock step (they meet at kBarFull every iteration), so
//     one L2 read of the BSK serves 4 ciphertexts and the instruction stream is shared.
constexpr int kSlots2 = 4;
constexpr int kComputeThreads2 = 64 * kSlots2;   // 256
constexpr int kLoaderThreads2 = 128;             // one loader warp per TMEM lane quarter
constexpr int kThreads2 = kComputeThreads2 + kLoaderThreads2;
constexpr int kBarFull = 5, kBarFree = 6;     // group 0 (or the whole CTA when not staggered)
constexpr int kBarFull1 = 7, kBarFree1 = 8;   // group 1 (slots 2, 3) in staggered mode
constexpr int kBarTokA = 9, kBarTokB = 10;    // transpose token: A = group 0 -> 1, B = group 1 -> 0
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kTmemBskCol = 256;

struct SlotSmem2 {
    double2 S[2][kTile];     // transpose tiles
    uint16_t amod[768];
};
struct Smem2 {
    // high words of the accumulator polynomials, first so that every 8 KB plane is 8 KB aligned in
    // the shared window (the kernel aligns the struct itself): a rotated read address is then
    // plane | (offset & 0x1FFC) -- one LOP3 -- and its partner (+N/2) is address ^ 0x1000.
    uint32_t hi[kSlots2][2][kN];
    double2 T[kTile];
    SlotSmem2 slot[kSlots2];
    uint32_t tmem_base;
    uint32_t pad[3];
};
constexpr size_t kSmem2Bytes = sizeof(Smem2) + 8192 - 16;   // worst-case alignment padding
static_assert(kSmem2Bytes <= 227 * 1024, "shared memory budget (v2)");

__device__ __forceinline__ void named_bar_sync(int id, int n) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory");
}
__device__ __forceinline__ void named_bar_arrive(int id, int n) {
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory");
}

// Signed digit (base 2^23, one level) from the high word of the value: round(x / 2^41) as a
// signed 23-bit integer.  Equals the tfhe rule except at the single tie x/2^41 = 2^22 - 1/2, where
// it returns -2^22 instead of +2^22 (the same value mod 2^64).
__device__ __forceinline__ int pbs_digit_hi(uint32_t hi) {
    return (int)(hi + 0x100u) >> 9;
}

__device__ __forceinline__ void ldg_b128(const void *p, uint32_t &a, uint32_t &b, uint32_t &c,
                                         uint32_t &d) {
    asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(a), "=r"(b), "=r"(c), "=r"(d)
                 : "l"(p));
}

// tcgen05.wait::ld carrying the loaded registers as read-write operands (uses cannot move above it)
__device__ __forceinline__ void wait_ld_dep16(uint32_t (&r)[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
                 :: "memory");
}
__device__ __forceinline__ void wait_ld_dep32(uint32_t (&r)[32]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]),
                   "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]),
                   "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
                 :: "memory");
}
// one 32-column chunk of the BSK slice = 4 frequencies: 2x2 GGSW contraction with one shfl_xor(16)
__device__ __forceinline__ void contract4(double (&xr)[32], double (&xi)[32], const uint32_t (&br)[32], int k0) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int k1 = k0 + k;
        const double csx = __hiloint2double((int)br[8 * k + 1], (int)br[8 * k + 0]);
        const double csy = __hiloint2double((int)br[8 * k + 3], (int)br[8 * k + 2]);
        const double ckx = __hiloint2double((int)br[8 * k + 5], (int)br[8 * k + 4]);
        const double cky = __hiloint2double((int)br[8 * k + 7], (int)br[8 * k + 6]);
        const double dr = xr[k1], di = xi[k1];
        const double sr = dr * csx - di * csy;
        const double si = dr * csy + di * csx;
        const double gr = __shfl_xor_sync(0xffffffffu, sr, 16);
        const double gi = __shfl_xor_sync(0xffffffffu, si, 16);
        xr[k1] = fma(dr, ckx, fma(-di, cky, gr));
        xi[k1] = fma(dr, cky, fma(di, ckx, gi));
    }
}
__device__ __forceinline__ void inv_twiddle4(double (&xr)[32], double (&xi)[32], const uint32_t (&tw)[16], int n0) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int n1 = n0 + k;
        const double tx = __hiloint2double((int)tw[4 * k + 1], (int)tw[4 * k + 0]);
        const double ty = __hiloint2double((int)tw[4 * k + 3], (int)tw[4 * k + 2]);
        const double a0 = xr[n1], b0 = xi[n1];
        xr[n1] = a0 * tx + b0 * ty;
        xi[n1] = b0 * tx - a0 * ty;
    }
}

// kOpt bit 2: rotated gather of the hi plane
// kOpt bit 3: TMEM loads (BSK slice, inverse twiddles) are requested one chunk ahead of their use. with one LOP3 per address (8 KB aligned planes) and the
// negacyclic sign taken from a pre-shifted offset (-160 integer instructions per CMUX).
// kOpt bit 0: inverse twiddles (the row T[rk2][0..31] every thread multiplies by after the inverse
// row pass) live in the 128 TMEM columns the accumulator and the BSK slice leave free, instead of
// being re-read from shared memory every iteration (shared-memory wavefronts are as scarce as FP64
// issue slots in this kernel: ncu shows elapsed = FP64 cycles + LSU wavefronts).
template <int kOpt, int kStag>
__global__ void __launch_bounds__(kThreads2, 1)
blind_rotate_v2_kernel(const uint64_t *__restrict__ lwe_small, const uint32_t *__restrict__ lut_id,
                       const uint64_t *__restrict__ lut_tables, const double2 *__restrict__ bsk_f,
                       const double2 *__restrict__ twiddle_g, uint64_t *__restrict__ lwe_out,
                       unsigned int batch, int n_lwe, int log_box, uint64_t delta, int active_slots,
                       int stagger_rt, unsigned int n_luts) {
    // active_slots in {1, 2, 4}: small levels use fewer ciphertexts per CTA (more CTAs, shorter
    // iterations); idle compute / loader warps skip the main loops entirely.
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const uint32_t smem_pad = (0u - (uint32_t)__cvta_generic_to_shared(smem_raw)) & 8191u;
    Smem2 &sm = *reinterpret_cast<Smem2 *>(smem_raw + smem_pad);
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    for (int t = threadIdx.x; t < kTile; t += kThreads2) sm.T[t] = twiddle_g[t];
    if (warp == 0) tmem_alloc(&sm.tmem_base, kTmemCols);
    tmem_fence_before_sync();
    __syncthreads();
    tmem_fence_after_sync();
    const uint32_t tbase = sm.tmem_base + (((uint32_t)(warp & 3) * 32u) << 16);
    const unsigned int n_groups = (batch + active_slots - 1) / active_slots;
    const int active_loaders = active_slots == 1 ? 2 : 4;
    // Staggered mode (4 slots): slots {0,1} and {2,3} form two groups with their own full / free
    // barriers.  Every SM sub-partition hosts one warp of each group; a token passed between the
    // groups serialises their shared-memory transposes (G0, G1, G0, G1, ...), which keeps the groups
    // out of phase: while one group is in an LSU-only transpose the other owns the FP64 pipe.
    // kStag 0: no stagger code (baseline for A/B runs); 3: stagger behind a run-time flag.  The flag
    // is deliberately NOT a template parameter: with predicated token barriers ptxas keeps the
    // hand-over next to the transposed reads; the compile-time variant sank it below the FFT and
    // was 6 ms slower.
    const bool stag = kStag != 0 && stagger_rt && active_slots == kSlots2;
    const int bar_threads = stag ? 128 + 32 * active_loaders : 64 * active_slots + 32 * active_loaders;

    if (warp >= 2 * kSlots2) {
        const bool loader_active = (warp & 3) < active_loaders;
        // =========================== BSK loader warps ===========================
        // Register budget: the pool that setmaxnreg.inc draws from is what the loaders release, so
        // (R_compute - 168) * 256 <= (168 - R_loader) * 128  ->  56 / 224.
        asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
        const int role = warp & 1;  // lane quarter q = warp & 3 serves compute warps with w = q & 1
        bool first = true;
        for (unsigned int g = blockIdx.x; loader_active && g < n_groups; g += gridDim.x) {
            for (int i = 0; i < n_lwe; i++) {
                const uint4 *__restrict__ src = reinterpret_cast<const uint4 *>(
                    bsk_f + ((((size_t)i * 32) * 2 + role) * 32 + lane) * 2);
#pragma unroll 1
                for (int b = 0; b < 8; b++) {   // 4 frequency rows (32 columns) per batch
                    uint32_t r[32];
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const uint4 *p = src + (size_t)(4 * b + k) * 128;  // row stride 128 double2
                        ldg_b128(p, r[8 * k + 0], r[8 * k + 1], r[8 * k + 2], r[8 * k + 3]);
                        ldg_b128(p + 1, r[8 * k + 4], r[8 * k + 5], r[8 * k + 6], r[8 * k + 7]);
                    }
                    if (b == 0 && !first) {     // previous slice fully consumed?
                        named_bar_sync(kBarFree, bar_threads);
                        if (stag) named_bar_sync(kBarFree1, bar_threads);
                        tmem_fence_after_sync();
                    }
                    tmem_st32(tbase + kTmemBskCol + 32 * b, r);
                }
                first = false;
                tmem_wait_st();
                tmem_fence_before_sync();
                named_bar_arrive(kBarFull, bar_threads);
                if (stag) named_bar_arrive(kBarFull1, bar_threads);
            }
        }
    } else {
        // =========================== compute warps ===========================
        asm volatile("setmaxnreg.inc.sync.aligned.u32 224;");
        const int slot = warp >> 1;
        const int w = warp & 1;
        SlotSmem2 &ss = sm.slot[slot];
        const int table_len = kN >> log_box;
        const int rp = lane >> 4;
        const int rk2 = 16 * w + (lane & 15);
        uint32_t *__restrict__ H = sm.hi[slot][w];
        const uint32_t h_plane = (uint32_t)__cvta_generic_to_shared(H);   // 8 KB aligned
        const uint32_t tacc = tbase + (uint32_t)(warp >> 2) * 128u;  // 64 columns: high words of 64 coefficients
        const uint32_t tbsk = tbase + kTmemBskCol;
        // free TMEM columns of this lane quarter: [64,128) and [192,256); warps q and q+4 share the
        // quarter and have the same (w, lane) -> the same twiddle row, so both write identical data.
        const uint32_t ttw0 = tbase + 64, ttw1 = tbase + 192;
        if ((kOpt & 1) && slot < active_slots) {
#pragma unroll
            for (int h = 0; h < 2; h++) {
                uint32_t tw[64];
#pragma unroll
                for (int q = 0; q < 16; q++) {
                    const double2 t = sm.T[rk2 * kTStride + 16 * h + q];
                    tw[4 * q + 0] = (uint32_t)__double2loint(t.x);
                    tw[4 * q + 1] = (uint32_t)__double2hiint(t.x);
                    tw[4 * q + 2] = (uint32_t)__double2loint(t.y);
                    tw[4 * q + 3] = (uint32_t)__double2hiint(t.y);
                }
                tmem_st64(h ? ttw1 : ttw0, tw);
            }
            tmem_wait_st();
        }
        const unsigned int last_group = ((n_groups - 1 - blockIdx.x) / gridDim.x) * gridDim.x + blockIdx.x;
        const int grp = stag ? (warp >> 2) : 0;
        const int bar_full = grp ? kBarFull1 : kBarFull;
        const int bar_free = grp ? kBarFree1 : kBarFree;
        const int tok_wait = grp ? kBarTokA : kBarTokB;   // token this group waits for
        const int tok_pass = grp ? kBarTokB : kBarTokA;   // token this group hands over
        if (stag && grp == 1) named_bar_arrive(kBarTokB, 256);  // group 0 owns the first transpose

        for (unsigned int g = blockIdx.x; slot < active_slots && g < n_groups; g += gridDim.x) {
            unsigned int ct = g * active_slots + slot;
            if (ct >= batch) ct = batch - 1;  // idle slot of the last group: recompute the last one
            const uint64_t *__restrict__ lwe = lwe_small + (size_t)ct * (n_lwe + 1);
            for (int t = threadIdx.x & 63; t < n_lwe; t += 64) ss.amod[t] = (uint16_t)modswitch(lwe[t]);
            {
                // a1.1 + rotation by X^{-b~}: this thread's 64 coefficients -> TMEM + hi plane
                const uint32_t bt = modswitch(lwe[n_lwe]);
                const uint64_t *__restrict__ tab = lut_tables + (size_t)min(lut_id[ct], n_luts - 1u) * table_len;
                uint32_t vhi[64];
#pragma unroll
                for (int q = 0; q < 64; q++) {
                    const int j = lane + 32 * (q >> 1) + kH * (q & 1);
                    const uint64_t v = (w == 0) ? 0ull : lut_coeff(tab, (uint32_t)j, bt, log_box, delta);
                    vhi[q] = (uint32_t)(v >> 32);   // low words of f * Delta are zero
                    H[j] = vhi[q];
                }
                tmem_st64(tacc, vhi);
                tmem_wait_st();
            }
            slot_bar(slot);  // amod + hi plane visible

            for (int i = 0; i < n_lwe; i++) {
                const uint32_t a = ss.amod[i];
                double xr[32], xi[32];
                // ---- (X^a - 1) * acc_w on high words, signed digit, pre-twist ----
                {
                    uint32_t own_l[64];
                    tmem_ld64(tacc, own_l);
                    const uint32_t base = (uint32_t)(lane - (int)a) & (2 * kN - 1);
                    // byte offset of coefficient (lane - a) mod 2N, pre-shifted so that the
                    // negacyclic sign (bit 13) is the top bit: + (m << 25) walks the 32 rows mod 2^32
                    const uint32_t off_s = base << 20;
                    tmem_wait_ld();
                    auto digits8 = [&](const uint32_t *own, int m0) {
#pragma unroll
                        for (int mm = 0; mm < 8; mm++) {
                            const int m = m0 + mm;
                            uint32_t r0, r1;
                            if (kOpt & 4) {
                                const uint32_t t0 = off_s + ((uint32_t)m << 25);     // idx0 << 20
                                const uint32_t t1 = t0 + (1u << 30);                // (idx0 + N/2) << 20
                                const uint32_t a0 = h_plane | ((t0 >> 18) & 0x1FFCu);
                                const uint32_t a1 = a0 ^ 0x1000u;
                                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r0) : "r"(a0) : "memory");
                                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r1) : "r"(a1) : "memory");
                                r0 ^= (uint32_t)((int)t0 >> 31);
                                r1 ^= (uint32_t)((int)t1 >> 31);
                            } else {
                                const uint32_t idx0 = (base + 32 * m) & (2 * kN - 1);
                                const uint32_t idx1 = (idx0 + kH) & (2 * kN - 1);
                                r0 = H[idx0 & (kN - 1)];
                                r1 = H[idx1 & (kN - 1)];
                                // negacyclic wrap: -x on the high word is ~x up to the dropped borrow
                                r0 ^= (uint32_t)((int)(idx0 << 20) >> 31);
                                r1 ^= (uint32_t)((int)(idx1 << 20) >> 31);
                            }
                            const double d0 = __int2double_rn(pbs_digit_hi(r0 - own[2 * mm]));
                            const double d1 = __int2double_rn(pbs_digit_hi(r1 - own[2 * mm + 1]));
                            if (m == 0) {
                                xr[FHS_BR5(m)] = d0;
                                xi[FHS_BR5(m)] = d1;
                            } else {
                                xr[FHS_BR5(m)] = d0 * fhs_pre_re(m) - d1 * fhs_pre_im(m);
                                xi[FHS_BR5(m)] = d0 * fhs_pre_im(m) + d1 * fhs_pre_re(m);
                            }
                        }
                    };
                    digits8(own_l, 0);
                    digits8(own_l + 16, 8);
                    digits8(own_l + 32, 16);
                    digits8(own_l + 48, 24);
                }
                fwd_cols(xr, xi, sm.T, lane);
                if (stag) named_bar_sync(tok_wait, 256);
                {
                    double2 *__restrict__ Sw = ss.S[w];
#pragma unroll
                    for (int k2 = 0; k2 < 32; k2++)
                        Sw[k2 * kTStride + lane] = make_double2(xr[k2], xi[k2]);
                }
                slot_bar(slot);
                {
                    const double2 *__restrict__ Sr = ss.S[rp] + rk2 * kTStride;
#pragma unroll
                    for (int n1 = 0; n1 < 32; n1++) {
                        const double2 v = Sr[n1];
                        xr[FHS_BR5(n1)] = v.x;
                        xi[FHS_BR5(n1)] = v.y;
                    }
                }
                if (stag) named_bar_arrive(tok_pass, 256);
                fft32_pos(xr, xi);
                // ---- GGSW contraction against the BSK slice parked in TMEM ----
                named_bar_sync(bar_full, bar_threads);
                tmem_fence_after_sync();
                if (kOpt & 8) {
                    uint32_t br0[32], br1[32];
                    tmem_ld32(tbsk, br0);
#pragma unroll
                    for (int b = 0; b < 8; b += 2) {
                        wait_ld_dep32(br0);
                        tmem_ld32(tbsk + 32 * (b + 1), br1);
                        contract4(xr, xi, br0, 4 * b);
                        wait_ld_dep32(br1);
                        if (b + 2 < 8) tmem_ld32(tbsk + 32 * (b + 2), br0);
                        contract4(xr, xi, br1, 4 * b + 4);
                    }
                } else {
#pragma unroll
                    for (int b = 0; b < 8; b++) {
                        uint32_t br[32];
                        tmem_ld32(tbsk + 32 * b, br);
                        tmem_wait_ld();
    #pragma unroll
                        for (int k = 0; k < 4; k++) {
                            const int k1 = 4 * b + k;
                            const double csx = __hiloint2double((int)br[8 * k + 1], (int)br[8 * k + 0]);
                            const double csy = __hiloint2double((int)br[8 * k + 3], (int)br[8 * k + 2]);
                            const double ckx = __hiloint2double((int)br[8 * k + 5], (int)br[8 * k + 4]);
                            const double cky = __hiloint2double((int)br[8 * k + 7], (int)br[8 * k + 6]);
                            const double dr = xr[k1], di = xi[k1];
                            const double sr = dr * csx - di * csy;
                            const double si = dr * csy + di * csx;
                            const double gr = __shfl_xor_sync(0xffffffffu, sr, 16);
                            const double gi = __shfl_xor_sync(0xffffffffu, si, 16);
                            xr[k1] = fma(dr, ckx, fma(-di, cky, gr));
                            xi[k1] = fma(dr, cky, fma(di, ckx, gi));
                        }
                    }
                }
                if (!(g == last_group && i == n_lwe - 1)) {
                    tmem_fence_before_sync();
                    named_bar_arrive(bar_free, bar_threads);
                }
#pragma unroll
                for (int q = 0; q < 32; q++) {
                    if (FHS_BR5(q) > q) {
                        double t0 = xr[q]; xr[q] = xr[FHS_BR5(q)]; xr[FHS_BR5(q)] = t0;
                        double t1 = xi[q]; xi[q] = xi[FHS_BR5(q)]; xi[FHS_BR5(q)] = t1;
                    }
                }
                // ---- inverse row pass, conjugate twiddle, transpose back ----
                if ((kOpt & 9) == 9) {
                    uint32_t tw0[16], tw1[16];
                    tmem_ld16(ttw0, tw0);          // in flight during the inverse row pass
                    fft32_neg(xr, xi);
#pragma unroll
                    for (int c = 0; c < 8; c += 2) {
                        wait_ld_dep16(tw0);
                        tmem_ld16((((c + 1) & 4) ? ttw1 : ttw0) + 16 * ((c + 1) & 3), tw1);
                        inv_twiddle4(xr, xi, tw0, 4 * c);
                        wait_ld_dep16(tw1);
                        if (c + 2 < 8) tmem_ld16((((c + 2) & 4) ? ttw1 : ttw0) + 16 * ((c + 2) & 3), tw0);
                        inv_twiddle4(xr, xi, tw1, 4 * c + 4);
                    }
                } else if (kOpt & 1) {
                    fft32_neg(xr, xi);
#pragma unroll
                    for (int c = 0; c < 8; c++) {   // 4 twiddles (16 columns) per TMEM load
                        uint32_t tw[16];
                        tmem_ld16(((c & 4) ? ttw1 : ttw0) + 16 * (c & 3), tw);
                        tmem_wait_ld();
#pragma unroll
                        for (int k = 0; k < 4; k++) {
                            const int n1 = 4 * c + k;
                            const double tx = __hiloint2double((int)tw[4 * k + 1], (int)tw[4 * k + 0]);
                            const double ty = __hiloint2double((int)tw[4 * k + 3], (int)tw[4 * k + 2]);
                            const double a0 = xr[n1], b0 = xi[n1];
                            xr[n1] = a0 * tx + b0 * ty;
                            xi[n1] = b0 * tx - a0 * ty;
                        }
                    }
                } else {
                    fft32_neg(xr, xi);
                    const double2 *__restrict__ Tr = sm.T + rk2 * kTStride;
#pragma unroll
                    for (int n1 = 0; n1 < 32; n1++) {
                        const double2 t = Tr[n1];
                        const double a0 = xr[n1], b0 = xi[n1];
                        xr[n1] = a0 * t.x + b0 * t.y;
                        xi[n1] = b0 * t.x - a0 * t.y;
                    }
                }
                if (stag) named_bar_sync(tok_wait, 256);
                {
                    double2 *__restrict__ Sq = ss.S[rp] + rk2 * kTStride;
#pragma unroll
                    for (int n1 = 0; n1 < 32; n1++) Sq[n1] = make_double2(xr[n1], xi[n1]);
                }
                slot_bar(slot);
                // the last token of the kernel is not handed over (nobody would collect it)
                const bool pass_tok = stag && !(grp == 1 && g == last_group && i == n_lwe - 1);
                {
                    const double2 *__restrict__ Sw = ss.S[w];
#pragma unroll
                    for (int k2 = 0; k2 < 32; k2++) {
                        const double2 v = Sw[k2 * kTStride + lane];
                        xr[FHS_BR5(k2)] = v.x;
                        xi[FHS_BR5(k2)] = v.y;
                    }
                }
                if (pass_tok) named_bar_arrive(tok_pass, 256);
                // ---- inverse column pass; conjugate pre-twist, add as torus into the TMEM accumulator ----
                auto accumulate8 = [&](uint32_t (&ohi)[16], int c) {
#pragma unroll
                    for (int mm = 0; mm < 8; mm++) {
                        const int m = 8 * c + mm;
                        double ur, ui;
                        if (m == 0) {
                            ur = xr[m];
                            ui = xi[m];
                        } else {
                            ur = xr[m] * fhs_pre_re(m) + xi[m] * fhs_pre_im(m);
                            ui = xi[m] * fhs_pre_re(m) - xr[m] * fhs_pre_im(m);
                        }
                        // values arrive in units of 2^34 (scale folded into the Fourier BSK):
                        // hi word += (round(v / 2^34) << 2) mod 2^32
                        ohi[2 * mm] += (uint32_t)__double2ll_rn(ur) << 2;
                        ohi[2 * mm + 1] += (uint32_t)__double2ll_rn(ui) << 2;
                        H[lane + 32 * m] = ohi[2 * mm];
                        H[lane + 32 * m + kH] = ohi[2 * mm + 1];
                    }
                    tmem_st16(tacc + 16 * c, ohi);
                };
                if (kOpt & 8) {
                    uint32_t oa[16], ob[16];
                    tmem_ld16(tacc, oa);           // in flight during the inverse column pass
                    fft32_neg(xr, xi);
                    wait_ld_dep16(oa);
                    tmem_ld16(tacc + 16, ob);
                    accumulate8(oa, 0);
                    wait_ld_dep16(ob);
                    tmem_ld16(tacc + 32, oa);      // issued after st16(chunk 0): same registers, in order
                    accumulate8(ob, 1);
                    wait_ld_dep16(oa);
                    tmem_ld16(tacc + 48, ob);
                    accumulate8(oa, 2);
                    wait_ld_dep16(ob);
                    accumulate8(ob, 3);
                } else {
                    fft32_neg(xr, xi);
#pragma unroll
                    for (int c = 0; c < 4; c++) {   // 8 values of m (16 coefficients) per chunk
                        uint32_t ohi[16];
                        tmem_ld16(tacc + 16 * c, ohi);
                        tmem_wait_ld();
                        accumulate8(ohi, c);
                    }
                }
                tmem_wait_st();
                __syncwarp();  // hi plane of polynomial w complete before this warp's rotated reads
            }

            // ---- a1.7 sample extract at coefficient 0, from the TMEM accumulator ----
            {
                uint32_t vhi[64];
                tmem_ld64(tacc, vhi);
                tmem_wait_ld();
                uint64_t *__restrict__ out = lwe_out + (size_t)ct * (kN + 1);
                if (w == 0) {
                    // out[0] = A[0]; out[j] = -A[N - j]  <=>  coefficient c goes to out[(N - c) % N]
#pragma unroll
                    for (int q = 0; q < 64; q++) {
                        const int cidx = lane + 32 * (q >> 1) + kH * (q & 1);
                        const uint64_t v = (uint64_t)vhi[q] << 32;
                        if (cidx == 0) out[0] = v;
                        else out[kN - cidx] = 0ull - v;
                    }
                } else if (lane == 0) {
                    out[kN] = (uint64_t)vhi[0] << 32;
                }
            }
            slot_bar(slot);  // amod / hi plane reuse by the next group
        }
    }
    tmem_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(sm.tmem_base, kTmemCols);
}

// Forward transform of the standard-domain BSK into the layout the kernel above reads:
//   bsk_f[(((i*32 + k1)*2 + w)*32 + lane)*2 + slot],  rp = lane >> 4, k2 = 16 w + (lane & 15),
//   slot 0 = GGSW row rp, column 1-rp ("send"), slot 1 = row rp, column rp ("keep"),
//   value = (1/1024) * FFT(signed torus polynomial)[32 k1 + k2].
constexpr int kXfWarps = 4;
struct XfSmem {
    double2 T[kTile];
    double2 S[kXfWarps][kTile];
};

__global__ void __launch_bounds__(32 * kXfWarps)
bsk_transform_kernel(const uint64_t *__restrict__ bsk, double2 *__restrict__ bsk_f,
                     double2 *__restrict__ bsk_w, const double2 *__restrict__ twiddle_g, int n_polys) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    XfSmem &sm = *reinterpret_cast<XfSmem *>(smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int t = threadIdx.x; t < kTile; t += 32 * kXfWarps) sm.T[t] = twiddle_g[t];
    __syncthreads();
    const int q = blockIdx.x * kXfWarps + warp;
    if (q >= n_polys) return;
    const int i = q >> 2, r = (q >> 1) & 1, c = q & 1;
    const uint64_t *__restrict__ P = bsk + (size_t)q * kN;
    double xr[32], xi[32];
#pragma unroll
    for (int m = 0; m < 32; m++) {
        const double d0 = __ll2double_rn((long long)P[lane + 32 * m]);
        const double d1 = __ll2double_rn((long long)P[lane + 32 * m + kH]);
        xr[FHS_BR5(m)] = d0 * kPreRe[m] - d1 * kPreIm[m];
        xi[FHS_BR5(m)] = d0 * kPreIm[m] + d1 * kPreRe[m];
    }
    fwd_cols(xr, xi, sm.T, lane);
    double2 *__restrict__ S = sm.S[warp];
#pragma unroll
    for (int k2 = 0; k2 < 32; k2++) S[k2 * kTStride + lane] = make_double2(xr[k2], xi[k2]);
    __syncwarp();
#pragma unroll
    for (int n1 = 0; n1 < 32; n1++) {
        const double2 v = S[lane * kTStride + n1];
        xr[FHS_BR5(n1)] = v.x;
        xi[FHS_BR5(n1)] = v.y;
    }
    fft32_pos(xr, xi);
    const int k2 = lane;
    const int ww = k2 >> 4, lp = (r << 4) | (k2 & 15), sl = (c == r) ? 1 : 0;
    const double sc = 1.0 / 1024.0 / 17179869184.0;  // 1/1024 (inverse FFT) and 2^-34 (accumulator unit)
#pragma unroll
    for (int k1 = 0; k1 < 32; k1++) {
        const double2 val = make_double2(xr[k1] * sc, xi[k1] * sc);
        bsk_f[((((size_t)i * 32 + k1) * 2 + ww) * 32 + lp) * 2 + sl] = val;
        // wide layout (blind_rotate_wide.cu): frequency K = c + 128 k3, c = q + 64 e; thread
        // (quarter q >> 4, lane (r << 4) | (q & 15)) holds it in slot f = 8 e + k3
        const int K = 32 * k1 + k2, cc = K & 127, f = 8 * (cc >> 6) + (K >> 7), qq = cc & 63;
        bsk_w[((((size_t)i * 16 + f) * 4 + (qq >> 4)) * 32 + ((r << 4) | (qq & 15))) * 2 + sl] = val;
    }
}

__global__ void build_lut_kernel(const uint64_t *__restrict__ table, uint64_t *__restrict__ acc,
                                 int log_box, uint64_t delta) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < kN) {
        acc[j] = 0ull;
        acc[kN + j] = lut_coeff(table, (uint32_t)j, 0u, log_box, delta);
    }
}

int ilog2(uint32_t x) {
    int l = 0;
    while ((1u << l) < x) l++;
    return l;
}

}  // namespace

size_t bsk_fourier_elems(const Ctx *c) { return (size_t)c->n * 32 * 2 * 32 * 2; }

int launch_bsk_transform(Ctx *c, const uint64_t *d_bsk_std) {
    const int n_polys = (int)c->n * 4;
    FHS_CUDA(cudaFuncSetAttribute(bsk_transform_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)sizeof(XfSmem)));
    bsk_transform_kernel<<<(n_polys + kXfWarps - 1) / kXfWarps, 32 * kXfWarps, sizeof(XfSmem),
                           c->stream>>>(d_bsk_std, c->d_bsk_f, c->d_bsk_w, c->d_twiddle, n_polys);
    c->launches++;
    FHS_CUDA(cudaGetLastError());
    return FHS_OK;
}

int launch_build_lut(Ctx *c, const uint64_t *d_table, uint64_t *d_acc) {
    const int log_box = ilog2(kN / c->space);
    const uint64_t delta = (1ull << 63) / c->space;
    build_lut_kernel<<<kN / 256, 256, 0, c->stream>>>(d_table, d_acc, log_box, delta);
    c->launches++;
    FHS_CUDA(cudaGetLastError());
    return FHS_OK;
}

// Dispatch of one level of `batch` blind rotations over the three kernels (all limits per context):
//   <= pair_limit (one ciphertext per two SMs, every cluster resident)  -> pair kernel, 1 wave
//   <= wide_limit (2 per SM)                                            -> wide kernel, 1 wave
//   otherwise whole 4-slot rounds of the throughput kernel, and the remainder as one pair wave, one wide
//   wave, or wide + pair when that is cheaper than one more 4-slot round.
int launch_blind_rotate(Ctx *c, const uint64_t *d_small, const uint32_t *d_lut_id,
                        const uint64_t *d_tables, size_t n_luts, size_t batch, uint64_t *d_out) {
    if (batch == 0) return FHS_OK;
    if (n_luts == 0) { set_error("blind rotation without a LUT"); return FHS_ERR_INVALID; }
    const long wide_limit = c->knobs.wide_limit;
    const long pair_cap = c->knobs.pair_on ? blind_rotate_pair_limit(c) : 0;
    const size_t in_w = (size_t)c->n + 1, out_w = (size_t)c->big + 1;
    auto small_wave = [&](size_t lo, size_t nb) -> int {   // nb <= max(pair_cap, wide_limit)
        if ((long)nb <= pair_cap)
            return launch_blind_rotate_pair(c, d_small + lo * in_w, d_lut_id + lo, d_tables, n_luts, nb, d_out + lo * out_w);
        return launch_blind_rotate_wide(c, d_small + lo * in_w, d_lut_id + lo, d_tables, n_luts, nb, d_out + lo * out_w);
    };
    if ((long)batch <= pair_cap || (long)batch <= wide_limit) return small_wave(0, batch);
    {
        const size_t round4 = (size_t)kSlots2 * c->sm_count, rem = batch % round4, head = batch - rem;
        // a remainder of at most 2 ciphertexts per SM after whole 4-slot rounds is cheaper as a separate
        // small wave (1.7 / 2.3 / 3.6 ms) than as one more 4-slot round (6.2 ms)
        if (rem > 0 && ((long)rem <= wide_limit || (long)rem <= pair_cap)) {
            int rc = launch_blind_rotate(c, d_small, d_lut_id, d_tables, n_luts, head, d_out);
            if (rc != FHS_OK) return rc;
            return small_wave(head, rem);
        }
        // up to one pair wave above two ciphertexts per SM: wide (3.6 ms) + pair (1.7 ms) instead of one
        // more 4-slot round (5.9 ms)
        if (wide_limit > 0 && pair_cap > 0 && (long)rem > wide_limit && (long)rem <= wide_limit + pair_cap) {
            int rc = head ? launch_blind_rotate(c, d_small, d_lut_id, d_tables, n_luts, head, d_out) : FHS_OK;
            if (rc != FHS_OK) return rc;
            if ((rc = small_wave(head, (size_t)wide_limit)) != FHS_OK) return rc;
            return small_wave(head + (size_t)wide_limit, rem - (size_t)wide_limit);
        }
    }
    const int stagger = c->knobs.stagger, opt = c->knobs.opt;
    if (!c->attr_br) {
#define FHS_BR_ATTR(o, st) FHS_CUDA(cudaFuncSetAttribute(blind_rotate_v2_kernel<o, st>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem2Bytes))
        FHS_BR_ATTR(0, 0); FHS_BR_ATTR(1, 3); FHS_BR_ATTR(5, 3); FHS_BR_ATTR(13, 3);
#undef FHS_BR_ATTR
        c->attr_br = true;
    }
    const int log_box = ilog2(kN / c->space);
    const uint64_t delta = (1ull << 63) / c->space;
    {
        const int active = batch <= (size_t)c->sm_count ? 1 : (batch <= 2 * (size_t)c->sm_count ? 2 : kSlots2);
        size_t ctas = (batch + active - 1) / active;
        if (ctas > (size_t)c->sm_count) ctas = c->sm_count;
        // stagger: 0 = plain kernel, 3/4 = kernel with the run-time stagger flag off/on (default 4).
        // The flagged kernel is the fast one even with the flag off: ptxas keeps the predicated
        // token barriers next to the transposes, which it does not for the compile-time variants.
        const int st = stagger;
        using kern_t = void (*)(const uint64_t *, const uint32_t *, const uint64_t *, const double2 *,
                                const double2 *, uint64_t *, unsigned int, int, int, uint64_t, int, int, unsigned int);
        // FHS_BR_OPT=0 FHS_BR_STAGGER=0 selects the round-1 baseline kernel (A/B runs); default is
        // the TMEM-twiddle kernel with the stagger flag on (4); 3 = same kernel, flag off.
        kern_t kern = (opt == 0 && st == 0) ? blind_rotate_v2_kernel<0, 0>
                      : (opt & 8)            ? blind_rotate_v2_kernel<13, 3>
                      : (opt & 4)            ? blind_rotate_v2_kernel<5, 3> : blind_rotate_v2_kernel<1, 3>;
        const int stagger_rt = st == 4;
        kern<<<(unsigned)ctas, kThreads2, kSmem2Bytes, c->stream>>>(
            d_small, d_lut_id, d_tables, c->d_bsk_f, c->d_twiddle, d_out, (unsigned int)batch,
            (int)c->n, log_box, delta, active, stagger_rt, (unsigned int)n_luts);
    }
    c->launches++;
    FHS_CUDA(cudaGetLastError());
    return FHS_OK;
}

}  // namespace fhs
