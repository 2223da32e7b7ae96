// Blind rotation, "wide" variant for SMALL levels (<= 2 ciphertexts per SM): the same algorithm as
// blind_rotate.cu (tfhe-rs blind_rotate_assign / cmux / add_external_product_assign, SURVEY.md 8(a1)
// rows a1.1, a1.5, a1.6, a1.7), but ONE ciphertext is spread over FOUR warps instead of two, so the
// dependent chain of a CMUX iteration (what bounds a level of <= 148 LUT evaluations: 742 strictly
// sequential iterations) is half as long per thread.
//
// FFT: 1024 complex points per polynomial as 16 x 8 x 8 (three passes, two shared-memory exchanges):
//   j = 64 j1 + t,  t = 8 j2 + j3      (thread t of the polynomial's 64 threads owns 16 values j1)
//   k = k1 + 16 k2 + 128 k3
//   X[k] = sum_{j3} w^{128 j3 k3} . w^{j3 (k1 + 16 k2)} . sum_{j2} w^{128 j2 k2} . w^{8 j2 k1 + t/4}
//          . sum_{j1} w^{64 j1 k1} . w^{16 j1} z_j                     w = e^{2 pi i / 1024}
//   pass 1: thread (j2, j3): constant pre-twist w^{16 j1}, DFT-16 over j1, twiddle tw1[k1]
//   pass 2: thread (k1, j3) x 2: DFT-8 over j2, twiddle tw2[k2]
//   pass 3: thread (k1, k2) x 2: DFT-8 over j3 -> 16 frequencies per thread
// In pass 3 lanes 0-15 of every warp hold polynomial 0 and lanes 16-31 polynomial 1 of the same
// frequencies, so the 2x2 GGSW contraction is one shfl_xor(16) per frequency as in the 2-warp kernel.
// The inverse runs the passes backwards with conjugate factors.  All twiddles (32 per thread) and
// the accumulator's own coefficients live in TMEM; the BSK slice of an iteration is 128 columns per
// lane, loaded by the loader warps from a second, wide-layout copy of the Fourier BSK.
#include <cstdlib>

#include "fhs_internal.cuh"
#include "fft_small_codelets.cuh"
#include "tmem.cuh"

namespace fhs {
namespace {

constexpr int kWCts = 2;                        // ciphertext slots per CTA
constexpr int kWCompute = 128 * kWCts;          // 4 warps per ciphertext
constexpr int kWThreads = kWCompute + 128;      // + one loader warp per TMEM lane quarter
constexpr int kE1Stride = 65;                   // E1[16 k1][64 t], padded rows
constexpr int kE2Stride = 9;                    // E2[128 c][8 j3], padded rows
constexpr int kWBarCt = 1;                      // + ct        (128 threads)
constexpr int kWBarPoly = 3;                    // + 2 ct + p  (64 threads)
constexpr int kWBarFull = 8, kWBarFree = 9;
constexpr uint32_t kWColOwn = 0;                // 32 columns per ciphertext slot
constexpr uint32_t kWColTw1 = 64, kWColTw2 = 128, kWColBsk = 256;

struct WSmem {
    uint32_t hi[kWCts][2][kN];                  // 8 KB planes, 8 KB aligned (kernel aligns the struct)
    double2 E1[kWCts][2][16 * kE1Stride];
    double2 E2[kWCts][2][128 * kE2Stride];
    uint16_t amod[kWCts][768];
    uint32_t tmem_base;
    uint32_t pad[3];
};
constexpr size_t kWSmemBytes = sizeof(WSmem) + 8192 - 16;
static_assert(kWSmemBytes <= 227 * 1024, "shared memory budget (wide)");

__device__ __forceinline__ void wbar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void wbar_arrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }

__device__ __forceinline__ uint32_t w_modswitch(uint64_t x) {
    return (uint32_t)((((x >> 51) + 1ull) >> 1) & (uint64_t)(2 * kN - 1));
}
__device__ __forceinline__ uint64_t w_lut_coeff(const uint64_t *__restrict__ tab, uint32_t j, uint32_t bt,
                                                int log_box, uint64_t delta) {
    const uint32_t v = (j + bt + (1u << (log_box - 1))) & (2 * kN - 1);
    const uint64_t f = tab[(v & (kN - 1)) >> log_box] * delta;
    return (v & kN) ? (0ull - f) : f;
}
__device__ __forceinline__ int w_digit_hi(uint32_t hi) { return (int)(hi + 0x100u) >> 9; }
__device__ __forceinline__ void w_ldg_b128(const void *p, uint32_t &a, uint32_t &b, uint32_t &c, uint32_t &d) {
    asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(p));
}
__device__ __forceinline__ double w_dbl(uint32_t lo, uint32_t hi) { return __hiloint2double((int)hi, (int)lo); }

// tcgen05.wait::ld with the loaded registers as read-write operands: later uses cannot be scheduled
// above the wait, and the registers stay reserved while the asynchronous load is in flight.
__device__ __forceinline__ void w_wait_ld16(uint32_t (&r)[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
                 :: "memory");
}

__device__ __forceinline__ void w_wait_ld32(uint32_t (&r)[32]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]),
                   "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]),
                   "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
                 :: "memory");
}

template <bool kConj, int kPerm, int kChunk>
__device__ __forceinline__ void w_twiddle4(double (&xr)[16], double (&xi)[16], const uint32_t (&tw)[16]) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int n = 4 * kChunk + k;
        // kPerm 0: natural slot; 4: slot BR4(n); 3: slot 8*(n>>3) + BR3(n&7)
        const int s = kPerm == 4 ? FHS_BR4(n) : (kPerm == 3 ? ((n & 8) | FHS_BR3(n & 7)) : n);
        const double tx = w_dbl(tw[4 * k + 0], tw[4 * k + 1]);
        const double ty = kConj ? -w_dbl(tw[4 * k + 2], tw[4 * k + 3]) : w_dbl(tw[4 * k + 2], tw[4 * k + 3]);
        const double a = xr[s], b = xi[s];
        xr[s] = a * tx - b * ty;
        xi[s] = a * ty + b * tx;
    }
}
// x[slot] *= (tx + i ty) or its conjugate, twiddles 4 at a time from 16 TMEM columns.  The caller has
// already requested chunk 0 into `tw0` (before the codelet that precedes the multiplication), and each
// further chunk is requested before the previous one is consumed, so the TMEM round trips overlap the
// arithmetic instead of serialising with it.
template <bool kConj, int kPerm>
__device__ __forceinline__ void w_twiddle16(double (&xr)[16], double (&xi)[16], uint32_t tcol, uint32_t (&tw0)[16]) {
    uint32_t tw1[16];
    w_wait_ld16(tw0);
    tmem_ld16(tcol + 16, tw1);
    w_twiddle4<kConj, kPerm, 0>(xr, xi, tw0);
    w_wait_ld16(tw1);
    tmem_ld16(tcol + 32, tw0);
    w_twiddle4<kConj, kPerm, 1>(xr, xi, tw1);
    w_wait_ld16(tw0);
    tmem_ld16(tcol + 48, tw1);
    w_twiddle4<kConj, kPerm, 2>(xr, xi, tw0);
    w_wait_ld16(tw1);
    w_twiddle4<kConj, kPerm, 3>(xr, xi, tw1);
}

__global__ void __launch_bounds__(kWThreads, 1)
blind_rotate_wide_kernel(const uint64_t *__restrict__ lwe_small, const uint32_t *__restrict__ lut_id,
                         const uint64_t *__restrict__ lut_tables, const double2 *__restrict__ bsk_w,
                         uint64_t *__restrict__ lwe_out, unsigned int batch, int n_lwe, int log_box,
                         uint64_t delta, int active_cts, unsigned int n_luts) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const uint32_t smem_pad = (0u - (uint32_t)__cvta_generic_to_shared(smem_raw)) & 8191u;
    WSmem &sm = *reinterpret_cast<WSmem *>(smem_raw + smem_pad);
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    if (warp == 0) tmem_alloc(&sm.tmem_base, 512);
    tmem_fence_before_sync();
    __syncthreads();
    tmem_fence_after_sync();
    const int v = warp & 3;                                        // TMEM lane quarter
    const uint32_t tbase = sm.tmem_base + (((uint32_t)v * 32u) << 16);
    const unsigned int n_groups = (batch + active_cts - 1) / active_cts;
    const int bar_threads = 128 * active_cts + 128;

    if (warp >= 4 * kWCts) {
        // =========================== BSK loader warps ===========================
        bool first = true;
        for (unsigned int g = blockIdx.x; g < n_groups; g += gridDim.x) {
            for (int i = 0; i < n_lwe; i++) {
                const uint4 *__restrict__ src = reinterpret_cast<const uint4 *>(
                    bsk_w + ((((size_t)i * 16) * 4 + v) * 32 + lane) * 2);
#pragma unroll 1
                for (int b = 0; b < 4; b++) {       // 4 frequency slots (32 columns) per batch
                    uint32_t r[32];
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const uint4 *p = src + (size_t)(4 * b + k) * (4 * 32 * 2);   // slot stride: 256 double2
                        w_ldg_b128(p, r[8 * k + 0], r[8 * k + 1], r[8 * k + 2], r[8 * k + 3]);
                        w_ldg_b128(p + 1, r[8 * k + 4], r[8 * k + 5], r[8 * k + 6], r[8 * k + 7]);
                    }
                    if (b == 0 && !first) {
                        wbar_sync(kWBarFree, bar_threads);
                        tmem_fence_after_sync();
                    }
                    tmem_st32(tbase + kWColBsk + 32 * b, r);
                }
                first = false;
                tmem_wait_st();
                tmem_fence_before_sync();
                wbar_arrive(kWBarFull, bar_threads);
            }
        }
    } else if ((warp >> 2) < active_cts) {
        // =========================== compute warps ===========================
        const int ct = warp >> 2;
        const int p = v >> 1;                       // polynomial in the coefficient domain / passes 1, 2
        const int t = 32 * (v & 1) + lane;          // thread of the polynomial: j2 = t >> 3, j3 = t & 7
        const int k1u = t & 15, j3u = t >> 4;       // pass 2: (k1u, j3u) and (k1u, j3u + 4)
        const int rp = lane >> 4;                   // pass 3: polynomial (= GGSW row)
        const int q = 16 * v + (lane & 15);         // pass 3: frequency classes c = q, q + 64
        uint32_t *__restrict__ H = sm.hi[ct][p];
        const uint32_t h_plane = (uint32_t)__cvta_generic_to_shared(H);
        double2 *__restrict__ E1 = sm.E1[ct][p];
        double2 *__restrict__ E2p = sm.E2[ct][p];
        double2 *__restrict__ E2r = sm.E2[ct][rp];
        const uint32_t town = tbase + kWColOwn + 32u * (uint32_t)ct;
        const uint32_t ttw1 = tbase + kWColTw1, ttw2 = tbase + kWColTw2, tbsk = tbase + kWColBsk;
        const int bar_ct = kWBarCt + ct, bar_poly = kWBarPoly + 2 * ct + p;
        const int table_len = kN >> log_box;
        const unsigned int last_group = ((n_groups - 1 - blockIdx.x) / gridDim.x) * gridDim.x + blockIdx.x;

        {   // twiddles of this thread -> TMEM (both ciphertext slots write identical values)
            uint32_t tw[64];
#pragma unroll
            for (int k1 = 0; k1 < 16; k1++) {       // tw1[k1] = w^{8 j2 k1 + t/4}
                double s, c;
                sincospi((double)(8 * (t >> 3) * k1) / 512.0 + (double)t / 2048.0, &s, &c);
                tw[4 * k1 + 0] = (uint32_t)__double2loint(c); tw[4 * k1 + 1] = (uint32_t)__double2hiint(c);
                tw[4 * k1 + 2] = (uint32_t)__double2loint(s); tw[4 * k1 + 3] = (uint32_t)__double2hiint(s);
            }
            tmem_st64(ttw1, tw);
#pragma unroll
            for (int n = 0; n < 16; n++) {          // tw2[8 e + k2] = w^{(j3u + 4 e)(k1u + 16 k2)}
                const int e = n >> 3, k2 = n & 7;
                double s, c;
                sincospi((double)(((j3u + 4 * e) * (k1u + 16 * k2)) & 1023) / 512.0, &s, &c);
                tw[4 * n + 0] = (uint32_t)__double2loint(c); tw[4 * n + 1] = (uint32_t)__double2hiint(c);
                tw[4 * n + 2] = (uint32_t)__double2loint(s); tw[4 * n + 3] = (uint32_t)__double2hiint(s);
            }
            tmem_st64(ttw2, tw);
            tmem_wait_st();
        }

        for (unsigned int g = blockIdx.x; g < n_groups; g += gridDim.x) {
            unsigned int cti = g * active_cts + ct;
            if (cti >= batch) cti = batch - 1;      // idle slot of the last group: recompute the last one
            const uint64_t *__restrict__ lwe = lwe_small + (size_t)cti * (n_lwe + 1);
            for (int x = threadIdx.x & 127; x < n_lwe; x += 128) sm.amod[ct][x] = (uint16_t)w_modswitch(lwe[x]);
            {   // a1.1 + rotation by X^{-b~}: this thread's 32 coefficients -> TMEM + hi plane
                const uint32_t bt = w_modswitch(lwe[n_lwe]);
                const uint64_t *__restrict__ tab = lut_tables + (size_t)min(lut_id[cti], n_luts - 1u) * table_len;
                uint32_t vhi[32];
#pragma unroll
                for (int x = 0; x < 32; x++) {
                    const int j = t + 64 * (x >> 1) + kH * (x & 1);
                    const uint64_t val = (p == 0) ? 0ull : w_lut_coeff(tab, (uint32_t)j, bt, log_box, delta);
                    vhi[x] = (uint32_t)(val >> 32);
                    H[j] = vhi[x];
                }
                tmem_st32(town, vhi);
                tmem_wait_st();
            }
            wbar_sync(bar_ct, 128);                 // amod + hi planes visible

            for (int i = 0; i < n_lwe; i++) {
                const uint32_t a = sm.amod[ct][i];
                double xr[16], xi[16];
                // ---- (X^a - 1) * acc on high words, signed digit, pre-twist w^{16 j1} ----
                {
                    uint32_t own[32];
                    tmem_ld32(town, own);
                    const uint32_t off_s = ((uint32_t)(t - (int)a) & (2 * kN - 1)) << 20;
                    tmem_wait_ld();
#pragma unroll
                    for (int j1 = 0; j1 < 16; j1++) {
                        const uint32_t t0 = off_s + ((uint32_t)j1 << 26);   // idx0 << 20, idx0 = t - a + 64 j1
                        const uint32_t t1 = t0 + (1u << 30);
                        const uint32_t a0 = h_plane | ((t0 >> 18) & 0x1FFCu);
                        const uint32_t a1 = a0 ^ 0x1000u;
                        uint32_t r0, r1;
                        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r0) : "r"(a0) : "memory");
                        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r1) : "r"(a1) : "memory");
                        r0 ^= (uint32_t)((int)t0 >> 31);
                        r1 ^= (uint32_t)((int)t1 >> 31);
                        const double d0 = __int2double_rn(w_digit_hi(r0 - own[2 * j1]));
                        const double d1 = __int2double_rn(w_digit_hi(r1 - own[2 * j1 + 1]));
                        if (j1 == 0) {
                            xr[FHS_BR4(j1)] = d0;
                            xi[FHS_BR4(j1)] = d1;
                        } else {
                            xr[FHS_BR4(j1)] = d0 * fhs_pre16_re(j1) - d1 * fhs_pre16_im(j1);
                            xi[FHS_BR4(j1)] = d0 * fhs_pre16_im(j1) + d1 * fhs_pre16_re(j1);
                        }
                    }
                }
                // ---- forward pass 1, twiddle 1, exchange 1 ----
                uint32_t twp[16];
                tmem_ld16(ttw1, twp);
                fft16_pos(xr, xi);
                w_twiddle16<false, 0>(xr, xi, ttw1, twp);
#pragma unroll
                for (int k1 = 0; k1 < 16; k1++) E1[k1 * kE1Stride + t] = make_double2(xr[k1], xi[k1]);
                wbar_sync(bar_poly, 64);
#pragma unroll
                for (int n = 0; n < 16; n++) {
                    const int e = n >> 3, j2 = n & 7;
                    const double2 val = E1[k1u * kE1Stride + 8 * j2 + j3u + 4 * e];
                    xr[8 * e + FHS_BR3(j2)] = val.x;
                    xi[8 * e + FHS_BR3(j2)] = val.y;
                }
                // ---- forward pass 2, twiddle 2, exchange 2 ----
                tmem_ld16(ttw2, twp);
                fft8_pos_lo(xr, xi);
                fft8_pos_hi(xr, xi);
                w_twiddle16<false, 0>(xr, xi, ttw2, twp);
#pragma unroll
                for (int n = 0; n < 16; n++) {
                    const int e = n >> 3, k2 = n & 7;
                    E2p[(k1u + 16 * k2) * kE2Stride + j3u + 4 * e] = make_double2(xr[n], xi[n]);
                }
                wbar_sync(bar_ct, 128);
#pragma unroll
                for (int n = 0; n < 16; n++) {
                    const int e = n >> 3, j3 = n & 7;
                    const double2 val = E2r[(q + 64 * e) * kE2Stride + j3];
                    xr[8 * e + FHS_BR3(j3)] = val.x;
                    xi[8 * e + FHS_BR3(j3)] = val.y;
                }
                // ---- forward pass 3: slot 8 e + k3 = frequency (q + 64 e) + 128 k3 of polynomial rp ----
                fft8_pos_lo(xr, xi);
                fft8_pos_hi(xr, xi);
                // ---- GGSW contraction against the BSK slice parked in TMEM ----
                wbar_sync(kWBarFull, bar_threads);
                tmem_fence_after_sync();
                {
                    uint32_t br0[16], br1[16];
                    auto contract2 = [&](const uint32_t (&br)[16], int f0) {
#pragma unroll
                        for (int k = 0; k < 2; k++) {
                            const int f = f0 + k;
                            const double csx = w_dbl(br[8 * k + 0], br[8 * k + 1]);
                            const double csy = w_dbl(br[8 * k + 2], br[8 * k + 3]);
                            const double ckx = w_dbl(br[8 * k + 4], br[8 * k + 5]);
                            const double cky = w_dbl(br[8 * k + 6], br[8 * k + 7]);
                            const double dr = xr[f], di = xi[f];
                            const double sr = dr * csx - di * csy;
                            const double si = dr * csy + di * csx;
                            const double gr = __shfl_xor_sync(0xffffffffu, sr, 16);
                            const double gi = __shfl_xor_sync(0xffffffffu, si, 16);
                            xr[f] = fma(dr, ckx, fma(-di, cky, gr));
                            xi[f] = fma(dr, cky, fma(di, ckx, gi));
                        }
                    };
                    tmem_ld16(tbsk, br0);
#pragma unroll
                    for (int b = 0; b < 8; b += 2) {
                        w_wait_ld16(br0);
                        tmem_ld16(tbsk + 16 * (b + 1), br1);
                        contract2(br0, 2 * b);
                        w_wait_ld16(br1);
                        if (b + 2 < 8) tmem_ld16(tbsk + 16 * (b + 2), br0);
                        contract2(br1, 2 * b + 2);
                    }
                }
                if (!(g == last_group && i == n_lwe - 1)) {
                    tmem_fence_before_sync();
                    wbar_arrive(kWBarFree, bar_threads);
                }
                // natural -> bit-reversed order inside each 8-block (BR3 swaps 1<->4, 3<->6)
#pragma unroll
                for (int e = 0; e < 2; e++) {
                    double s0;
                    s0 = xr[8 * e + 1]; xr[8 * e + 1] = xr[8 * e + 4]; xr[8 * e + 4] = s0;
                    s0 = xi[8 * e + 1]; xi[8 * e + 1] = xi[8 * e + 4]; xi[8 * e + 4] = s0;
                    s0 = xr[8 * e + 3]; xr[8 * e + 3] = xr[8 * e + 6]; xr[8 * e + 6] = s0;
                    s0 = xi[8 * e + 3]; xi[8 * e + 3] = xi[8 * e + 6]; xi[8 * e + 6] = s0;
                }
                // ---- inverse pass 3, exchange 2 back ----
                fft8_neg_lo(xr, xi);
                fft8_neg_hi(xr, xi);
#pragma unroll
                for (int n = 0; n < 16; n++) {
                    const int e = n >> 3, j3 = n & 7;
                    E2r[(q + 64 * e) * kE2Stride + j3] = make_double2(xr[n], xi[n]);
                }
                wbar_sync(bar_ct, 128);
                tmem_ld16(ttw2, twp);
#pragma unroll
                for (int n = 0; n < 16; n++) {
                    const int e = n >> 3, k2 = n & 7;
                    const double2 val = E2p[(k1u + 16 * k2) * kE2Stride + j3u + 4 * e];
                    xr[8 * e + FHS_BR3(k2)] = val.x;
                    xi[8 * e + FHS_BR3(k2)] = val.y;
                }
                // ---- conjugate twiddle 2, inverse pass 2, exchange 1 back ----
                w_twiddle16<true, 3>(xr, xi, ttw2, twp);
                fft8_neg_lo(xr, xi);
                fft8_neg_hi(xr, xi);
#pragma unroll
                for (int n = 0; n < 16; n++) {
                    const int e = n >> 3, j2 = n & 7;
                    E1[k1u * kE1Stride + 8 * j2 + j3u + 4 * e] = make_double2(xr[n], xi[n]);
                }
                wbar_sync(bar_poly, 64);
                tmem_ld16(ttw1, twp);
#pragma unroll
                for (int k1 = 0; k1 < 16; k1++) {
                    const double2 val = E1[k1 * kE1Stride + t];
                    xr[FHS_BR4(k1)] = val.x;
                    xi[FHS_BR4(k1)] = val.y;
                }
                // ---- conjugate twiddle 1, inverse pass 1 ----
                w_twiddle16<true, 4>(xr, xi, ttw1, twp);
                uint32_t ohi[32];
                tmem_ld32(town, ohi);               // in flight during the inverse pass
                fft16_neg(xr, xi);
                // ---- conjugate pre-twist, add as torus into the TMEM accumulator + hi plane ----
                {
                    w_wait_ld32(ohi);
#pragma unroll
                    for (int j1 = 0; j1 < 16; j1++) {
                        double ur, ui;
                        if (j1 == 0) {
                            ur = xr[j1];
                            ui = xi[j1];
                        } else {
                            ur = xr[j1] * fhs_pre16_re(j1) + xi[j1] * fhs_pre16_im(j1);
                            ui = xi[j1] * fhs_pre16_re(j1) - xr[j1] * fhs_pre16_im(j1);
                        }
                        ohi[2 * j1] += (uint32_t)__double2ll_rn(ur) << 2;
                        ohi[2 * j1 + 1] += (uint32_t)__double2ll_rn(ui) << 2;
                        H[t + 64 * j1] = ohi[2 * j1];
                        H[t + 64 * j1 + kH] = ohi[2 * j1 + 1];
                    }
                    tmem_st32(town, ohi);
                    tmem_wait_st();
                }
                wbar_sync(bar_poly, 64);            // hi plane of this polynomial complete (2 warps write it)
            }

            // ---- a1.7 sample extract at coefficient 0, from the TMEM accumulator ----
            {
                uint32_t vhi[32];
                tmem_ld32(town, vhi);
                tmem_wait_ld();
                uint64_t *__restrict__ out = lwe_out + (size_t)cti * (kN + 1);
                if (p == 0) {
#pragma unroll
                    for (int x = 0; x < 32; x++) {
                        const int cidx = t + 64 * (x >> 1) + kH * (x & 1);
                        const uint64_t val = (uint64_t)vhi[x] << 32;
                        if (cidx == 0) out[0] = val;
                        else out[kN - cidx] = 0ull - val;
                    }
                } else if (t == 0) {
                    out[kN] = (uint64_t)vhi[0] << 32;
                }
            }
            wbar_sync(bar_ct, 128);                 // amod / hi plane reuse by the next group
        }
    }
    tmem_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(sm.tmem_base, 512);
}

int w_ilog2(uint32_t x) {
    int l = 0;
    while ((1u << l) < x) l++;
    return l;
}

}  // namespace

// One wave of the wide kernel (the caller, launch_blind_rotate, picks the kernel per wave).
int launch_blind_rotate_wide(Ctx *c, const uint64_t *d_small, const uint32_t *d_lut_id,
                             const uint64_t *d_tables, size_t n_luts, size_t batch, uint64_t *d_out) {
    if (batch == 0) return FHS_OK;
    if (!c->attr_brw) {
        FHS_CUDA(cudaFuncSetAttribute(blind_rotate_wide_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)kWSmemBytes));
        c->attr_brw = true;
    }
    const int log_box = w_ilog2(kN / c->space);
    const uint64_t delta = (1ull << 63) / c->space;
    const int active = batch <= (size_t)c->sm_count ? 1 : kWCts;
    size_t ctas = (batch + active - 1) / active;
    if (ctas > (size_t)c->sm_count) ctas = c->sm_count;
    blind_rotate_wide_kernel<<<(unsigned)ctas, kWThreads, kWSmemBytes, c->stream>>>(
        d_small, d_lut_id, d_tables, c->d_bsk_w, d_out, (unsigned int)batch, (int)c->n, log_box, delta, active, (unsigned int)n_luts);
    c->launches++;
    FHS_CUDA(cudaGetLastError());
    return FHS_OK;
}

}  // namespace fhs
