// Blind rotation, "pair" variant for the SMALLEST levels (at most one ciphertext per two SMs): the
// same algorithm as blind_rotate.cu / blind_rotate_wide.cu (tfhe-rs blind_rotate_assign / cmux /
// add_external_product_assign), but ONE ciphertext is
// spread over a thread-block CLUSTER of two CTAs (two SMs).  A level of a few dozen LUT evaluations
// is bound by the 742 strictly sequential CMUX iterations; with one ciphertext per SM the FP64 pipe
// of that SM is the floor of an iteration, so the only way below it is a second SM per ciphertext.
//
// CTA p of the pair owns accumulator polynomial p (0 = mask, 1 = body) end to end:
//   decomposition of (X^a - 1) acc_p, forward FFT, both products of the GGSW row p
//       keep_p[k] = D_p[k] B[p][p][k]        send_p[k] = D_p[k] B[p][1-p][k],
//   send_p goes straight into the PEER's shared memory with st.async, whose bytes complete a
//   transaction on the peer's mbarrier (no fence, no remote arrive), out_p = keep_p + recv, inverse FFT,
//   torus accumulation.  One 16 KB exchange per direction and iteration, no cluster-wide barrier in
//   the loop; receive buffers and mbarriers are double-buffered so the peer may run one iteration ahead.
//
// FFT: 1024 complex points over 128 threads x 8 values as 8 x 8 x (8 x 2):
//   j = 128 j1 + t                       thread t owns 8 values j1 (16 coefficients)
//   k = k1 + 8 k0' + 64 m + 512 h
//   pass 1: constant pre-twist e^{2 pi i j1/32}, DFT-8 over j1, twiddle w^{t k1 + t/4}   (w = e^{2 pi i/1024})
//   pass 2: thread (k1, t0), t = 16 t1 + t0: DFT-8 over t1, twiddle e^{2 pi i t0 k0'/128}
//   pass 3: thread (k1, k0', s), t0 = 2 u + s: DFT-8 over u, twiddle e^{2 pi i s m/16}, then a radix-2
//           butterfly with the neighbouring lane (shfl_xor 1): thread s ends up with h = s.
// The inverse runs the passes backwards with conjugate factors; 1/1024 and the accumulator unit are
// folded into the Fourier BSK as in the other kernels.  Twiddles, the thread's 16 accumulator words
// and the BSK slice (64 columns per lane) live in TMEM.
#include <cstdlib>

#include "fhs_internal.cuh"
#include "fft_small_codelets.cuh"
#include "tmem.cuh"

namespace fhs {
namespace {

constexpr int kPThreads = 256;                  // 4 compute warps + one loader warp per TMEM lane quarter
constexpr int kPE1Stride = 128;                 // E1[8 k1][128 t]
constexpr int kPE2Stride = 18;                  // E2[64 (k1, k0')][16 t0], padded rows
constexpr int kPBarAll = 1, kPBarFull = 2, kPBarFree = 3;
constexpr uint32_t kPColOwn = 0, kPColTw1 = 32, kPColTw2 = 64, kPColTw3 = 96, kPColBsk = 128, kPTmemCols = 256;

struct PSmem {
    uint32_t hi[kN];                            // 8 KB plane of this CTA's polynomial, 8 KB aligned
    double2 recv[2][8 * 128];                   // [iteration parity][m][thread]: written by the peer CTA
    double2 E1[8 * kPE1Stride];
    double2 E2[64 * kPE2Stride];
    unsigned long long mbar[2];
    uint16_t amod[768];
    uint32_t tmem_base;
    uint32_t pad[3];
};
constexpr size_t kPSmemBytes = sizeof(PSmem) + 8192 - 16;

__device__ __forceinline__ void pbar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void pbar_arrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
// 16-byte store into the peer's shared memory that reports its bytes to the peer's mbarrier when it
// has landed: no release fence on the sender (an mbarrier.arrive.release.cluster after plain remote
// stores compiles to MEMBAR.ALL.GPU + ERRBAR and cost 40 % of an iteration)
__device__ __forceinline__ void st_peer_f64x2(uint32_t caddr, double x, double y, uint32_t cmbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f64 [%0], {%1, %2}, [%3];"
                 ::"r"(caddr), "d"(x), "d"(y), "r"(cmbar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t saddr, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(saddr), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_init(uint32_t saddr, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(saddr), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t saddr, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"(saddr), "r"(parity) : "memory");
    } while (!ok);
}

__device__ __forceinline__ uint32_t p_modswitch(uint64_t x) {
    return (uint32_t)((((x >> 51) + 1ull) >> 1) & (uint64_t)(2 * kN - 1));
}
__device__ __forceinline__ uint64_t p_lut_coeff(const uint64_t *__restrict__ tab, uint32_t j, uint32_t bt,
                                                int log_box, uint64_t delta) {
    const uint32_t v = (j + bt + (1u << (log_box - 1))) & (2 * kN - 1);
    const uint64_t f = tab[(v & (kN - 1)) >> log_box] * delta;
    return (v & kN) ? (0ull - f) : f;
}
__device__ __forceinline__ int p_digit_hi(uint32_t hi) { return (int)(hi + 0x100u) >> 9; }
__device__ __forceinline__ void p_ldg_b128(const void *p, uint32_t &a, uint32_t &b, uint32_t &c, uint32_t &d) {
    asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(p));
}
__device__ __forceinline__ double p_dbl(uint32_t lo, uint32_t hi) { return __hiloint2double((int)hi, (int)lo); }
__device__ __forceinline__ void p_wait_ld32(uint32_t (&r)[32]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]),
                   "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]),
                   "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
                 :: "memory");
}
__device__ __forceinline__ void p_wait_ld16(uint32_t (&r)[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
                 :: "memory");
}

// x[slot(n)] *= tw[n] (or its conjugate), n < 8; kBitRev: value n sits in slot BR3(n)
template <bool kConj, bool kBitRev>
__device__ __forceinline__ void p_twiddle8(double (&xr)[16], double (&xi)[16], const uint32_t (&tw)[32]) {
#pragma unroll
    for (int n = 0; n < 8; n++) {
        const int s = kBitRev ? FHS_BR3(n) : n;
        const double tx = p_dbl(tw[4 * n + 0], tw[4 * n + 1]);
        const double ty = kConj ? -p_dbl(tw[4 * n + 2], tw[4 * n + 3]) : p_dbl(tw[4 * n + 2], tw[4 * n + 3]);
        const double a = xr[s], b = xi[s];
        xr[s] = a * tx - b * ty;
        xi[s] = a * ty + b * tx;
    }
}
__device__ __forceinline__ void p_store_tw(uint32_t (&tw)[32], int n, double c, double s) {
    tw[4 * n + 0] = (uint32_t)__double2loint(c); tw[4 * n + 1] = (uint32_t)__double2hiint(c);
    tw[4 * n + 2] = (uint32_t)__double2loint(s); tw[4 * n + 3] = (uint32_t)__double2hiint(s);
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kPThreads, 1)
blind_rotate_pair_kernel(const uint64_t *__restrict__ lwe_small, const uint32_t *__restrict__ lut_id,
                         const uint64_t *__restrict__ lut_tables, const double2 *__restrict__ bsk_x,
                         uint64_t *__restrict__ lwe_out, int n_lwe, int log_box, uint64_t delta,
                         unsigned int n_luts) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    // the pad is the same in both CTAs of the pair (same kernel, same dynamic shared memory window), so a
    // local shared address maps to the peer's copy of the same field
    const uint32_t smem_pad = (0u - (uint32_t)__cvta_generic_to_shared(smem_raw)) & 8191u;
    PSmem &sm = *reinterpret_cast<PSmem *>(smem_raw + smem_pad);
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int v = warp & 3;                                        // TMEM lane quarter
    const uint32_t p = cluster_ctarank();                          // polynomial / GGSW row of this CTA
    const unsigned int cti = blockIdx.x >> 1;                      // ciphertext of this pair
    const uint32_t mbar0 = (uint32_t)__cvta_generic_to_shared(&sm.mbar[0]);
    if (warp == 0) tmem_alloc(&sm.tmem_base, kPTmemCols);
    if (threadIdx.x == 32) {
        mbar_init(mbar0, 1);                                       // one local arrival (expect_tx) + 16 KB of peer stores
        mbar_init(mbar0 + 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    tmem_fence_before_sync();
    __syncthreads();
    tmem_fence_after_sync();
    cluster_sync_all();                                            // the peer's mbarriers exist
    const uint32_t tbase = sm.tmem_base + (((uint32_t)v * 32u) << 16);

    if (warp >= 4) {
        // =========================== BSK loader warps ===========================
        for (int i = 0; i < n_lwe; i++) {
            const uint4 *__restrict__ src = reinterpret_cast<const uint4 *>(
                bsk_x + (((((size_t)i * 2 + p) * 4 + v) * 8) * 32 + lane) * 2);
#pragma unroll 1
            for (int b = 0; b < 2; b++) {           // 4 frequency slots m (32 columns) per batch
                uint32_t r[32];
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const uint4 *q = src + (size_t)(4 * b + k) * 64;           // slot stride: 32 lanes x 2 double2
                    p_ldg_b128(q, r[8 * k + 0], r[8 * k + 1], r[8 * k + 2], r[8 * k + 3]);
                    p_ldg_b128(q + 1, r[8 * k + 4], r[8 * k + 5], r[8 * k + 6], r[8 * k + 7]);
                }
                if (b == 0 && i > 0) {
                    pbar_sync(kPBarFree, kPThreads);
                    tmem_fence_after_sync();
                }
                tmem_st32(tbase + kPColBsk + 32 * b, r);
            }
            tmem_wait_st();
            tmem_fence_before_sync();
            pbar_arrive(kPBarFull, kPThreads);
        }
    } else {
        // =========================== compute warps ===========================
        const int t = 32 * v + lane;                // pass 1: coefficients 128 j1 + t (+ N/2)
        const int t0 = t & 15, k1p = t >> 4;        // pass 2: thread (k1p, t0)
        const int s = t & 1, kq = (t >> 1) & 7, k1q = t >> 4;   // pass 3: thread (k1q, k0' = kq, s)
        const double sgn = s ? -1.0 : 1.0;
        uint32_t *__restrict__ H = sm.hi;
        const uint32_t h_plane = (uint32_t)__cvta_generic_to_shared(H);
        double2 *__restrict__ E1 = sm.E1;
        double2 *__restrict__ E2 = sm.E2;
        const uint32_t town = tbase + kPColOwn, ttw1 = tbase + kPColTw1, ttw2 = tbase + kPColTw2,
                       ttw3 = tbase + kPColTw3, tbsk = tbase + kPColBsk;
        const uint32_t peer = p ^ 1u;
        const uint32_t peer_recv = map_to_cta((uint32_t)__cvta_generic_to_shared(&sm.recv[0][t]), peer);
        const uint32_t peer_mbar = map_to_cta(mbar0, peer);
        const int table_len = kN >> log_box;

        {   // twiddles of this thread -> TMEM
            uint32_t tw[32];
#pragma unroll
            for (int k1 = 0; k1 < 8; k1++) {        // tw1[k1] = w^{t k1 + t/4}
                double sn, cs;
                sincospi((double)(t * k1) / 512.0 + (double)t / 2048.0, &sn, &cs);
                p_store_tw(tw, k1, cs, sn);
            }
            tmem_st32(ttw1, tw);
#pragma unroll
            for (int k0 = 0; k0 < 8; k0++) {        // tw2[k0'] = e^{2 pi i t0 k0'/128}
                double sn, cs;
                sincospi((double)(t0 * k0) / 64.0, &sn, &cs);
                p_store_tw(tw, k0, cs, sn);
            }
            tmem_st32(ttw2, tw);
#pragma unroll
            for (int m = 0; m < 8; m++) {           // tw3[m] = e^{2 pi i s m/16}
                double sn, cs;
                sincospi((double)(s * m) / 8.0, &sn, &cs);
                p_store_tw(tw, m, cs, sn);
            }
            tmem_st32(ttw3, tw);
            tmem_wait_st();
        }

        const uint64_t *__restrict__ lwe = lwe_small + (size_t)cti * (n_lwe + 1);
        for (int x = t; x < n_lwe; x += 128) sm.amod[x] = (uint16_t)p_modswitch(lwe[x]);
        {   // a1.1 + rotation by X^{-b~}: this thread's 16 coefficients -> TMEM + hi plane
            const uint32_t bt = p_modswitch(lwe[n_lwe]);
            const uint64_t *__restrict__ tab = lut_tables + (size_t)min(lut_id[cti], n_luts - 1u) * table_len;
            uint32_t vhi[16];
#pragma unroll
            for (int x = 0; x < 16; x++) {
                const int j = t + 128 * (x >> 1) + kH * (x & 1);
                const uint64_t val = (p == 0) ? 0ull : p_lut_coeff(tab, (uint32_t)j, bt, log_box, delta);
                vhi[x] = (uint32_t)(val >> 32);
                H[j] = vhi[x];
            }
            tmem_st16(town, vhi);
            tmem_wait_st();
        }
        pbar_sync(kPBarAll, 128);                   // amod + hi plane visible

        for (int i = 0; i < n_lwe; i++) {
            const uint32_t a = sm.amod[i];
            const int buf = i & 1;
            double xr[16], xi[16];
            uint32_t tw[32];
            // ---- (X^a - 1) * acc_p on high words, signed digit, pre-twist e^{2 pi i j1/32} ----
            {
                uint32_t own[16];
                tmem_ld16(town, own);
                const uint32_t off_s = ((uint32_t)(t - (int)a) & (2 * kN - 1)) << 20;
                tmem_wait_ld();
#pragma unroll
                for (int j1 = 0; j1 < 8; j1++) {
                    const uint32_t u0 = off_s + ((uint32_t)j1 << 27);       // idx0 << 20, idx0 = t - a + 128 j1
                    const uint32_t u1 = u0 + (1u << 30);
                    const uint32_t a0 = h_plane | ((u0 >> 18) & 0x1FFCu);
                    const uint32_t a1 = a0 ^ 0x1000u;
                    uint32_t r0, r1;
                    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r0) : "r"(a0) : "memory");
                    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r1) : "r"(a1) : "memory");
                    r0 ^= (uint32_t)((int)u0 >> 31);
                    r1 ^= (uint32_t)((int)u1 >> 31);
                    const double d0 = __int2double_rn(p_digit_hi(r0 - own[2 * j1]));
                    const double d1 = __int2double_rn(p_digit_hi(r1 - own[2 * j1 + 1]));
                    if (j1 == 0) {
                        xr[FHS_BR3(j1)] = d0;
                        xi[FHS_BR3(j1)] = d1;
                    } else {
                        xr[FHS_BR3(j1)] = d0 * fhs_pre16_re(2 * j1) - d1 * fhs_pre16_im(2 * j1);
                        xi[FHS_BR3(j1)] = d0 * fhs_pre16_im(2 * j1) + d1 * fhs_pre16_re(2 * j1);
                    }
                }
            }
            // ---- forward pass 1, twiddle 1, exchange 1 ----
            tmem_ld32(ttw1, tw);
            fft8_pos_lo(xr, xi);
            p_wait_ld32(tw);
            p_twiddle8<false, false>(xr, xi, tw);
#pragma unroll
            for (int k1 = 0; k1 < 8; k1++) E1[k1 * kPE1Stride + t] = make_double2(xr[k1], xi[k1]);
            pbar_sync(kPBarAll, 128);
            tmem_ld32(ttw2, tw);
#pragma unroll
            for (int t1 = 0; t1 < 8; t1++) {
                const double2 val = E1[k1p * kPE1Stride + 16 * t1 + t0];
                xr[FHS_BR3(t1)] = val.x;
                xi[FHS_BR3(t1)] = val.y;
            }
            // ---- forward pass 2, twiddle 2, exchange 2 ----
            fft8_pos_lo(xr, xi);
            p_wait_ld32(tw);
            p_twiddle8<false, false>(xr, xi, tw);
#pragma unroll
            for (int k0 = 0; k0 < 8; k0++) E2[(k1p * 8 + k0) * kPE2Stride + t0] = make_double2(xr[k0], xi[k0]);
            // rows 16 v .. 16 v + 15 of E2 are written and read by warp v alone (k1p = k1q = t >> 4): a
            // warp-level sync replaces the 128-thread barrier of round 1
            __syncwarp();
            tmem_ld32(ttw3, tw);
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const double2 val = E2[(k1q * 8 + kq) * kPE2Stride + 2 * u + s];
                xr[FHS_BR3(u)] = val.x;
                xi[FHS_BR3(u)] = val.y;
            }
            // ---- forward pass 3: DFT-8 over u, twiddle 3, radix-2 butterfly with lane ^ 1 ----
            fft8_pos_lo(xr, xi);
            p_wait_ld32(tw);
            p_twiddle8<false, false>(xr, xi, tw);
#pragma unroll
            for (int m = 0; m < 8; m++) {
                const double gr = __shfl_xor_sync(0xffffffffu, xr[m], 1);
                const double gi = __shfl_xor_sync(0xffffffffu, xi[m], 1);
                xr[m] = fma(sgn, xr[m], gr);        // s = 0: E0 + E1'   s = 1: E0 - E1'
                xi[m] = fma(sgn, xi[m], gi);
            }
            // slot m = frequency k1q + 8 kq + 64 m + 512 s of polynomial p
            // ---- GGSW row p against the BSK slice parked in TMEM; "send" products go to the peer ----
            pbar_sync(kPBarFull, kPThreads);
            tmem_fence_after_sync();
            if (t == 0) mbar_expect_tx(mbar0 + 8u * (uint32_t)buf, 8u * 128u * (uint32_t)sizeof(double2));
            {
                uint32_t br0[32], br1[32];
                tmem_ld32(tbsk, br0);
                auto contract4 = [&](const uint32_t (&br)[32], int m0) {
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const int m = m0 + k;
                        const double csx = p_dbl(br[8 * k + 0], br[8 * k + 1]);
                        const double csy = p_dbl(br[8 * k + 2], br[8 * k + 3]);
                        const double ckx = p_dbl(br[8 * k + 4], br[8 * k + 5]);
                        const double cky = p_dbl(br[8 * k + 6], br[8 * k + 7]);
                        const double dr = xr[m], di = xi[m];
                        st_peer_f64x2(peer_recv + (uint32_t)((buf * 8 + m) * 128 * sizeof(double2)),
                                      dr * csx - di * csy, dr * csy + di * csx, peer_mbar + 8u * (uint32_t)buf);
                        xr[m] = dr * ckx - di * cky;
                        xi[m] = dr * cky + di * ckx;
                    }
                };
                p_wait_ld32(br0);
                tmem_ld32(tbsk + 32, br1);
                contract4(br0, 0);
                p_wait_ld32(br1);
                contract4(br1, 4);
            }
            if (i != n_lwe - 1) {
                tmem_fence_before_sync();
                pbar_arrive(kPBarFree, kPThreads);
            }
            mbar_wait(mbar0 + 8u * (uint32_t)buf, (uint32_t)(i >> 1) & 1u);
#pragma unroll
            for (int m = 0; m < 8; m++) {
                const double2 val = sm.recv[buf][m * 128 + t];
                xr[m] += val.x;
                xi[m] += val.y;
            }
            // ---- inverse pass 3: butterfly, conjugate twiddle 3, DFT-8 over m; exchange 2 back ----
            tmem_ld32(ttw3, tw);
#pragma unroll
            for (int m = 0; m < 8; m++) {
                const double gr = __shfl_xor_sync(0xffffffffu, xr[m], 1);
                const double gi = __shfl_xor_sync(0xffffffffu, xi[m], 1);
                xr[m] = fma(sgn, xr[m], gr);
                xi[m] = fma(sgn, xi[m], gi);
            }
            p_wait_ld32(tw);
            p_twiddle8<true, false>(xr, xi, tw);
            {   // natural -> bit-reversed order (BR3 swaps 1<->4, 3<->6)
                double s0;
                s0 = xr[1]; xr[1] = xr[4]; xr[4] = s0;
                s0 = xi[1]; xi[1] = xi[4]; xi[4] = s0;
                s0 = xr[3]; xr[3] = xr[6]; xr[6] = s0;
                s0 = xi[3]; xi[3] = xi[6]; xi[6] = s0;
            }
            fft8_neg_lo(xr, xi);
#pragma unroll
            for (int u = 0; u < 8; u++) E2[(k1q * 8 + kq) * kPE2Stride + 2 * u + s] = make_double2(xr[u], xi[u]);
            __syncwarp();                           // warp-private rows, as in the forward exchange
            tmem_ld32(ttw2, tw);
#pragma unroll
            for (int k0 = 0; k0 < 8; k0++) {
                const double2 val = E2[(k1p * 8 + k0) * kPE2Stride + t0];
                xr[FHS_BR3(k0)] = val.x;
                xi[FHS_BR3(k0)] = val.y;
            }
            // ---- conjugate twiddle 2, inverse pass 2, exchange 1 back ----
            p_wait_ld32(tw);
            p_twiddle8<true, true>(xr, xi, tw);
            fft8_neg_lo(xr, xi);
#pragma unroll
            for (int t1 = 0; t1 < 8; t1++) E1[k1p * kPE1Stride + 16 * t1 + t0] = make_double2(xr[t1], xi[t1]);
            pbar_sync(kPBarAll, 128);
            tmem_ld32(ttw1, tw);
#pragma unroll
            for (int k1 = 0; k1 < 8; k1++) {
                const double2 val = E1[k1 * kPE1Stride + t];
                xr[FHS_BR3(k1)] = val.x;
                xi[FHS_BR3(k1)] = val.y;
            }
            // ---- conjugate twiddle 1, inverse pass 1, conjugate pre-twist, torus accumulation ----
            p_wait_ld32(tw);
            p_twiddle8<true, true>(xr, xi, tw);
            uint32_t ohi[16];
            tmem_ld16(town, ohi);                   // in flight during the inverse pass
            fft8_neg_lo(xr, xi);
            p_wait_ld16(ohi);
#pragma unroll
            for (int j1 = 0; j1 < 8; j1++) {
                double ur, ui;
                if (j1 == 0) {
                    ur = xr[j1];
                    ui = xi[j1];
                } else {
                    ur = xr[j1] * fhs_pre16_re(2 * j1) + xi[j1] * fhs_pre16_im(2 * j1);
                    ui = xi[j1] * fhs_pre16_re(2 * j1) - xr[j1] * fhs_pre16_im(2 * j1);
                }
                ohi[2 * j1] += (uint32_t)__double2ll_rn(ur) << 2;
                ohi[2 * j1 + 1] += (uint32_t)__double2ll_rn(ui) << 2;
                H[t + 128 * j1] = ohi[2 * j1];
                H[t + 128 * j1 + kH] = ohi[2 * j1 + 1];
            }
            tmem_st16(town, ohi);
            tmem_wait_st();
            pbar_sync(kPBarAll, 128);               // hi plane complete (4 warps write it)
        }

        // ---- a1.7 sample extract at coefficient 0, from the TMEM accumulator ----
        {
            uint32_t vhi[16];
            tmem_ld16(town, vhi);
            tmem_wait_ld();
            uint64_t *__restrict__ out = lwe_out + (size_t)cti * (kN + 1);
            if (p == 0) {
#pragma unroll
                for (int x = 0; x < 16; x++) {
                    const int cidx = t + 128 * (x >> 1) + kH * (x & 1);
                    const uint64_t val = (uint64_t)vhi[x] << 32;
                    if (cidx == 0) out[0] = val;
                    else out[kN - cidx] = 0ull - val;
                }
            } else if (t == 0) {
                out[kN] = (uint64_t)vhi[0] << 32;
            }
        }
    }
    tmem_fence_before_sync();
    __syncthreads();
    cluster_sync_all();                             // no CTA of the pair exits while its peer may still write to it
    if (warp == 0) tmem_dealloc(sm.tmem_base, kPTmemCols);
}

// wide layout (blind_rotate_wide.cu) -> pair layout:
//   bsk_x[(((((i*2 + r)*4 + v)*8 + m)*32 + lane)*2 + sl],  thread T = 32 v + lane = 2 (k0' + 8 k1) + h,
//   frequency K = k1 + 8 k0' + 64 m + 512 h, sl 0 = GGSW row r column 1-r ("send"), 1 = column r ("keep")
__global__ void bsk_pair_repack_kernel(const double2 *__restrict__ bsk_w, double2 *__restrict__ bsk_x, int n_lwe) {
    const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;      // (i, r, K, sl)
    if (g >= (size_t)n_lwe * 2 * kH * 2) return;
    const int sl = (int)(g & 1), K = (int)((g >> 1) & (kH - 1)), r = (int)((g >> 11) & 1);
    const size_t i = g >> 12;
    const int cc = K & 127, f = 8 * (cc >> 6) + (K >> 7), qq = cc & 63;
    const double2 val = bsk_w[((((size_t)i * 16 + f) * 4 + (qq >> 4)) * 32 + ((r << 4) | (qq & 15))) * 2 + sl];
    const int k1 = K & 7, k0 = (K >> 3) & 7, m = (K >> 6) & 7, h = K >> 9;
    const int T = 2 * (k0 + 8 * k1) + h;
    bsk_x[(((((size_t)i * 2 + r) * 4 + (T >> 5)) * 8 + m) * 32 + (T & 31)) * 2 + sl] = val;
}

int p_ilog2(uint32_t x) {
    int l = 0;
    while ((1u << l) < x) l++;
    return l;
}

}  // namespace

int launch_bsk_pair_repack(Ctx *c) {
    const size_t n = (size_t)c->n * 2 * kH * 2;
    bsk_pair_repack_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(c->d_bsk_w, c->d_bsk_x, (int)c->n);
    c->launches++;
    FHS_CUDA(cudaGetLastError());
    return FHS_OK;
}

// Largest batch the pair kernel takes on this device: every pair must be resident at once (a second
// wave would double the latency this kernel exists to cut), at most one ciphertext per two SMs.
int blind_rotate_pair_limit(Ctx *c) {
    if (c->pair_limit >= 0) return c->pair_limit;
    c->pair_limit = 0;
    if (cudaFuncSetAttribute(blind_rotate_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)kPSmemBytes) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * (unsigned)(c->sm_count / 2));
    cfg.blockDim = dim3(kPThreads);
    cfg.dynamicSmemBytes = kPSmemBytes;
    int clusters = 0;
    if (cudaOccupancyMaxActiveClusters(&clusters, blind_rotate_pair_kernel, &cfg) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    const int half = c->sm_count / 2;
    c->pair_limit = clusters < half ? clusters : half;
    return c->pair_limit;
}

int launch_blind_rotate_pair(Ctx *c, const uint64_t *d_small, const uint32_t *d_lut_id,
                             const uint64_t *d_tables, size_t n_luts, size_t batch, uint64_t *d_out) {
    if (batch == 0) return FHS_OK;
    const int log_box = p_ilog2(kN / c->space);
    const uint64_t delta = (1ull << 63) / c->space;
    blind_rotate_pair_kernel<<<(unsigned)(2 * batch), kPThreads, kPSmemBytes, c->stream>>>(
        d_small, d_lut_id, d_tables, c->d_bsk_x, d_out, (int)c->n, log_box, delta, (unsigned int)n_luts);
    c->launches++;
    FHS_CUDA(cudaGetLastError());
    return FHS_OK;
}

}  // namespace fhs
