/*
 * fhs_oracle.c -- CPU ORACLE (test infrastructure) restating the tfhe-rs 0.5 KS-PBS path that
 * sits under every fhe_string operation.  PARITY UNPINNED at the tfhe-rs boundary -- see the
 * header comment in fhs_oracle.h.  Plain C, no dependencies (OpenMP optional).
 *
 * Reference call sites that reach this arithmetic: every `k.k.<op>` in
  * /root/reference/src/ciphertext/ (all .rs files), e.g. compare.rs:32, search.rs:17-41.
 * Algorithms follow SURVEY.md Appendix A (tfhe-rs 0.5, recalled):
 *   A.2 encoding / encrypt / decrypt          A.3 signed decomposer
 *   A.4 keyswitch                             A.5 modulus switch
 *   A.6 blind rotation (f64 negacyclic FFT)   A.7 sample extract      A.8 LUT accumulator
 */
#include "fhs_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ----------------------------------------------------------------------------------------- */
/* RNG: xoshiro256** seeded with splitmix64; Box-Muller gaussians. Deterministic per stream.  */
typedef struct {
    uint64_t s[4];
    int has_spare;
    double spare;
} orc_rng;

static uint64_t splitmix64(uint64_t *x) {
    uint64_t z = (*x += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
static void rng_seed(orc_rng *r, uint64_t seed, uint64_t stream) {
    uint64_t x = seed ^ (stream * 0xD6E8FEB86659FD93ULL + 0x2545F4914F6CDD1DULL);
    for (int i = 0; i < 4; i++) r->s[i] = splitmix64(&x);
    r->has_spare = 0;
    r->spare = 0.0;
}
static inline uint64_t rotl64(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
static inline uint64_t rng_u64(orc_rng *r) {
    uint64_t *s = r->s;
    uint64_t result = rotl64(s[1] * 5, 7) * 9;
    uint64_t t = s[1] << 17;
    s[2] ^= s[0];
    s[3] ^= s[1];
    s[1] ^= s[2];
    s[0] ^= s[3];
    s[2] ^= t;
    s[3] = rotl64(s[3], 45);
    return result;
}
static inline double rng_unit(orc_rng *r) { /* (0,1] */
    return ((double)(rng_u64(r) >> 11) + 1.0) * (1.0 / 9007199254740992.0);
}
static double rng_gauss(orc_rng *r) {
    if (r->has_spare) {
        r->has_spare = 0;
        return r->spare;
    }
    double u1 = rng_unit(r), u2 = rng_unit(r);
    double mag = sqrt(-2.0 * log(u1));
    double a = 6.283185307179586476925286766559 * u2;
    r->spare = mag * sin(a);
    r->has_spare = 1;
    return mag * cos(a);
}
/* torus noise: round(N(0, (std * 2^64)^2)) as a wrapping u64 */
static inline uint64_t rng_noise(orc_rng *r, double std) {
    double v = rng_gauss(r) * std * 18446744073709551616.0;
    return (uint64_t)(int64_t)llrint(v);
}

/* ----------------------------------------------------------------------------------------- */
typedef struct {
    double re, im;
} cplx;

struct orc_keyset {
    orc_params p;
    uint64_t *lwe_sk;  /* n */
    uint64_t *glwe_sk; /* k*N, also the big LWE key */
    uint64_t *ksk;     /* [kN][ks_level][n+1] */
    uint64_t *bsk;     /* [n][lvl][k+1][k+1][N] */
    cplx *bsk_f;       /* [n][lvl][k+1][k+1][N/2] */
    cplx *twist;       /* N/2: e^{i pi j / N} */
    cplx *tw;          /* N/4 twiddles of the N/2-point FFT */
    uint32_t *brev;    /* bit reversal for N/2 */
#ifdef ORC_FAST
    double *ftw;       /* per-stage radix-4 twiddles, SoA */
    double *twr, *twi; /* twist, SoA */
    double *bsk_fr, *bsk_fi; /* Fourier BSK in the fast transform's own order, SoA [n][lvl][k+1][k+1][N/2] */
#endif
};

void orc_default_params(orc_params *p) {
    memset(p, 0, sizeof(*p));
    p->lwe_dim = 742;
    p->glwe_dim = 1;
    p->poly_size = 2048;
    p->pbs_base_log = 23;
    p->pbs_level = 1;
    p->ks_base_log = 3;
    p->ks_level = 5;
    p->message_modulus = 4;
    p->carry_modulus = 4;
    p->lwe_std = 7.069849454709433e-6;
    p->glwe_std = 0.00000000000000029403601535432533;
}

const orc_params *orc_keyset_params(const orc_keyset *ks) { return &ks->p; }
const uint64_t *orc_ksk(const orc_keyset *ks) { return ks->ksk; }
const uint64_t *orc_bsk(const orc_keyset *ks) { return ks->bsk; }
const uint64_t *orc_lwe_sk(const orc_keyset *ks) { return ks->lwe_sk; }
const uint64_t *orc_glwe_sk(const orc_keyset *ks) { return ks->glwe_sk; }
size_t orc_ksk_len(const orc_keyset *ks) {
    return (size_t)ks->p.glwe_dim * ks->p.poly_size * ks->p.ks_level * (ks->p.lwe_dim + 1);
}
size_t orc_bsk_len(const orc_keyset *ks) {
    size_t k1 = ks->p.glwe_dim + 1;
    return (size_t)ks->p.lwe_dim * ks->p.pbs_level * k1 * k1 * ks->p.poly_size;
}
int orc_is_fast_build(void) {
#ifdef ORC_FAST
    return 1;
#else
    return 0;
#endif
}
int orc_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ----------------------------------------------------------------------------------------- */
/* A.6 FFT convention: N reals -> N/2 complex z_j = (p_j + i p_{j+N/2}) * e^{i pi j/N}, then a   */
/* N/2-point complex DFT.  Radix-2 DIT with a bit-reversal permutation.                         */
static void fft_tables(orc_keyset *ks) {
    const uint32_t N = ks->p.poly_size, H = N / 2;
    ks->twist = (cplx *)malloc(sizeof(cplx) * H);
    ks->tw = (cplx *)malloc(sizeof(cplx) * (H / 2));
    ks->brev = (uint32_t *)malloc(sizeof(uint32_t) * H);
    const long double PI = 3.14159265358979323846264338327950288L;
    for (uint32_t j = 0; j < H; j++) {
        long double a = PI * (long double)j / (long double)N;
        ks->twist[j].re = (double)cosl(a);
        ks->twist[j].im = (double)sinl(a);
    }
    for (uint32_t j = 0; j < H / 2; j++) {
        long double a = -2.0L * PI * (long double)j / (long double)H;
        ks->tw[j].re = (double)cosl(a);
        ks->tw[j].im = (double)sinl(a);
    }
    int lg = 0;
    while ((1u << lg) < H) lg++;
    for (uint32_t j = 0; j < H; j++) {
        uint32_t r = 0;
        for (int b = 0; b < lg; b++)
            if (j & (1u << b)) r |= 1u << (lg - 1 - b);
        ks->brev[j] = r;
    }
}

/* in-place forward DFT (sign -1) or inverse (sign +1, unnormalised) of H points */
static void fft_core(const orc_keyset *ks, cplx *x, int inverse) {
    const uint32_t H = ks->p.poly_size / 2;
    for (uint32_t j = 0; j < H; j++) {
        uint32_t r = ks->brev[j];
        if (r > j) {
            cplx t = x[j];
            x[j] = x[r];
            x[r] = t;
        }
    }
    for (uint32_t len = 2; len <= H; len <<= 1) {
        uint32_t half = len >> 1, step = H / len;
        for (uint32_t s = 0; s < H; s += len) {
            for (uint32_t j = 0; j < half; j++) {
                cplx w = ks->tw[j * step];
                if (inverse) w.im = -w.im;
                cplx a = x[s + j], b = x[s + j + half];
                cplx t = {b.re * w.re - b.im * w.im, b.re * w.im + b.im * w.re};
                x[s + j].re = a.re + t.re;
                x[s + j].im = a.im + t.im;
                x[s + j + half].re = a.re - t.re;
                x[s + j + half].im = a.im - t.im;
            }
        }
    }
}

/* forward_as_integer: signed integer polynomial -> Fourier */
static void fft_fwd_int(const orc_keyset *ks, const int64_t *poly, cplx *out) {
    const uint32_t H = ks->p.poly_size / 2;
    for (uint32_t j = 0; j < H; j++) {
        double a = (double)poly[j], b = (double)poly[j + H];
        cplx w = ks->twist[j];
        out[j].re = a * w.re - b * w.im;
        out[j].im = a * w.im + b * w.re;
    }
    fft_core(ks, out, 0);
}
/* forward_as_torus: u64 torus polynomial (as signed, scaled by 2^-64) -> Fourier */
static void fft_fwd_torus(const orc_keyset *ks, const uint64_t *poly, cplx *out) {
    const uint32_t H = ks->p.poly_size / 2;
    const double sc = 1.0 / 18446744073709551616.0;
    for (uint32_t j = 0; j < H; j++) {
        double a = (double)(int64_t)poly[j] * sc, b = (double)(int64_t)poly[j + H] * sc;
        cplx w = ks->twist[j];
        out[j].re = a * w.re - b * w.im;
        out[j].im = a * w.im + b * w.re;
    }
    fft_core(ks, out, 0);
}
static inline uint64_t torus_from_f64(double x) {
    double fr = x - nearbyint(x); /* [-0.5, 0.5] */
    double y = fr * 18446744073709551616.0;
    if (y >= 9223372036854775807.0) return 0x8000000000000000ULL;
    if (y <= -9223372036854775808.0) return 0x8000000000000000ULL;
    return (uint64_t)(int64_t)llrint(y);
}
/* add_backward_as_torus: Fourier -> torus polynomial, wrapping-added into `poly` */
static void fft_bwd_add_torus(const orc_keyset *ks, cplx *f, uint64_t *poly) {
    const uint32_t H = ks->p.poly_size / 2;
    fft_core(ks, f, 1);
    const double inv = 1.0 / (double)H;
    for (uint32_t j = 0; j < H; j++) {
        cplx w = ks->twist[j];
        double re = (f[j].re * w.re + f[j].im * w.im) * inv;
        double im = (f[j].im * w.re - f[j].re * w.im) * inv;
        poly[j] += torus_from_f64(re);
        poly[j + H] += torus_from_f64(im);
    }
}


#ifdef ORC_FAST
static void fast_bsk(orc_keyset *ks);
/* ------------------------------------------------------------------------------------------ */
/* ORC_FAST build (oracle/Makefile target `fast`: -O3 -march=native -ffp-contract=fast): the SAME    */
/* algorithm with the engineering a production CPU library has -- structure-of-arrays complex data,   */
/* radix-4 decimation-in-frequency forward / decimation-in-time inverse transforms that skip the    */
/* digit-reversal permutation (the spectra are only multiplied pointwise, so any order that the two  */
/* directions share is fine: concrete-fft does the same), per-stage contiguous twiddles, scratch     */
/* buffers allocated once per thread, monomial rotation without per-element modulo.  Results are    */
/* NOT bit-identical to the portable build (different summation order, FMA contraction); the        */
/* golden vectors pin the portable build only.  This build is what bench.py times as the CPU arm.   */
typedef struct {
    double *re, *im;
} soa;

/* per-stage twiddles for spans m = H, H/4, ..., 4: w^j, w^2j, w^3j with w = e^{-2 pi i / m} */
static void fast_tables(orc_keyset *ks) {
    const uint32_t H = ks->p.poly_size / 2;
    size_t total = 0;
    for (uint32_t m = H; m >= 4; m >>= 2) total += (size_t)(m / 4) * 6;
    ks->ftw = (double *)aligned_alloc(64, ((total * sizeof(double) + 63) / 64) * 64);
    const long double PI = 3.14159265358979323846264338327950288L;
    double *t = ks->ftw;
    for (uint32_t m = H; m >= 4; m >>= 2) {
        const uint32_t q = m / 4;
        for (int k = 1; k <= 3; k++) {
            for (uint32_t j = 0; j < q; j++) {
                long double a = -2.0L * PI * (long double)(k * j) / (long double)m;
                t[(size_t)(2 * (k - 1)) * q + j] = (double)cosl(a);
                t[(size_t)(2 * (k - 1) + 1) * q + j] = (double)sinl(a);
            }
        }
        t += (size_t)q * 6;
    }
    ks->twr = (double *)aligned_alloc(64, sizeof(double) * H);
    ks->twi = (double *)aligned_alloc(64, sizeof(double) * H);
    for (uint32_t j = 0; j < H; j++) { ks->twr[j] = ks->twist[j].re; ks->twi[j] = ks->twist[j].im; }
}

/* one radix-4 decimation-in-frequency stage on one block: q butterflies, unit stride */
static inline void r4_fwd_block(double *restrict r0, double *restrict r1, double *restrict r2, double *restrict r3,
                                double *restrict i0, double *restrict i1, double *restrict i2, double *restrict i3,
                                const double *restrict t, long q) {
    const double *restrict w1r = t, *restrict w1i = t + q, *restrict w2r = t + 2 * q, *restrict w2i = t + 3 * q;
    const double *restrict w3r = t + 4 * q, *restrict w3i = t + 5 * q;
#pragma omp simd
    for (long j = 0; j < q; j++) {
        const double b0r = r0[j] + r2[j], b0i = i0[j] + i2[j], b1r = r0[j] - r2[j], b1i = i0[j] - i2[j];
        const double b2r = r1[j] + r3[j], b2i = i1[j] + i3[j];
        const double b3r = i1[j] - i3[j], b3i = r3[j] - r1[j];          /* (a1 - a3) * (-i) */
        const double y1r = b1r + b3r, y1i = b1i + b3i, y2r = b0r - b2r, y2i = b0i - b2i;
        const double y3r = b1r - b3r, y3i = b1i - b3i;
        r0[j] = b0r + b2r; i0[j] = b0i + b2i;
        r1[j] = y1r * w1r[j] - y1i * w1i[j]; i1[j] = y1r * w1i[j] + y1i * w1r[j];
        r2[j] = y2r * w2r[j] - y2i * w2i[j]; i2[j] = y2r * w2i[j] + y2i * w2r[j];
        r3[j] = y3r * w3r[j] - y3i * w3i[j]; i3[j] = y3r * w3i[j] + y3i * w3r[j];
    }
}
/* its exact mirror (decimation in time, conjugate twiddles), unnormalised (x 4) */
static inline void r4_inv_block(double *restrict r0, double *restrict r1, double *restrict r2, double *restrict r3,
                                double *restrict i0, double *restrict i1, double *restrict i2, double *restrict i3,
                                const double *restrict t, long q) {
    const double *restrict w1r = t, *restrict w1i = t + q, *restrict w2r = t + 2 * q, *restrict w2i = t + 3 * q;
    const double *restrict w3r = t + 4 * q, *restrict w3i = t + 5 * q;
#pragma omp simd
    for (long j = 0; j < q; j++) {
        const double u1r = r1[j] * w1r[j] + i1[j] * w1i[j], u1i = i1[j] * w1r[j] - r1[j] * w1i[j];
        const double u2r = r2[j] * w2r[j] + i2[j] * w2i[j], u2i = i2[j] * w2r[j] - r2[j] * w2i[j];
        const double u3r = r3[j] * w3r[j] + i3[j] * w3i[j], u3i = i3[j] * w3r[j] - r3[j] * w3i[j];
        const double b0r = r0[j] + u2r, b0i = i0[j] + u2i, b2r = r0[j] - u2r, b2i = i0[j] - u2i;
        const double b1r = u1r + u3r, b1i = u1i + u3i;
        const double b3r = u3i - u1i, b3i = u1r - u3r;                  /* (u1 - u3) * (+i) */
        r0[j] = b0r + b1r; i0[j] = b0i + b1i;
        r2[j] = b0r - b1r; i2[j] = b0i - b1i;
        r1[j] = b2r + b3r; i1[j] = b2i + b3i;
        r3[j] = b2r - b3r; i3[j] = b2i - b3i;
    }
}
/* the last two forward stages (spans 16 and 4) on one 16-point block, twiddles as constants in registers:
 * unit-stride loops of 4 and 1 butterflies do not vectorise, a fully unrolled block does */
static inline void r4_fwd_tail16(double *restrict xr, double *restrict xi, const double *restrict t16) {
    double ar[16], ai[16];
    for (int j = 0; j < 4; j++) {     /* span 16: q = 4 */
        const double b0r = xr[j] + xr[j + 8], b0i = xi[j] + xi[j + 8], b1r = xr[j] - xr[j + 8], b1i = xi[j] - xi[j + 8];
        const double b2r = xr[j + 4] + xr[j + 12], b2i = xi[j + 4] + xi[j + 12];
        const double b3r = xi[j + 4] - xi[j + 12], b3i = xr[j + 12] - xr[j + 4];
        const double y1r = b1r + b3r, y1i = b1i + b3i, y2r = b0r - b2r, y2i = b0i - b2i, y3r = b1r - b3r, y3i = b1i - b3i;
        ar[j] = b0r + b2r; ai[j] = b0i + b2i;
        ar[j + 4] = y1r * t16[j] - y1i * t16[4 + j]; ai[j + 4] = y1r * t16[4 + j] + y1i * t16[j];
        ar[j + 8] = y2r * t16[8 + j] - y2i * t16[12 + j]; ai[j + 8] = y2r * t16[12 + j] + y2i * t16[8 + j];
        ar[j + 12] = y3r * t16[16 + j] - y3i * t16[20 + j]; ai[j + 12] = y3r * t16[20 + j] + y3i * t16[16 + j];
    }
    for (int s = 0; s < 16; s += 4) { /* span 4: q = 1, all twiddles are 1 */
        const double b0r = ar[s] + ar[s + 2], b0i = ai[s] + ai[s + 2], b1r = ar[s] - ar[s + 2], b1i = ai[s] - ai[s + 2];
        const double b2r = ar[s + 1] + ar[s + 3], b2i = ai[s + 1] + ai[s + 3];
        const double b3r = ai[s + 1] - ai[s + 3], b3i = ar[s + 3] - ar[s + 1];
        xr[s] = b0r + b2r; xi[s] = b0i + b2i;
        xr[s + 1] = b1r + b3r; xi[s + 1] = b1i + b3i;
        xr[s + 2] = b0r - b2r; xi[s + 2] = b0i - b2i;
        xr[s + 3] = b1r - b3r; xi[s + 3] = b1i - b3i;
    }
}
static inline void r4_inv_head16(double *restrict xr, double *restrict xi, const double *restrict t16) {
    double ar[16], ai[16];
    for (int s = 0; s < 16; s += 4) { /* span 4 */
        const double b0r = xr[s] + xr[s + 2], b0i = xi[s] + xi[s + 2], b2r = xr[s] - xr[s + 2], b2i = xi[s] - xi[s + 2];
        const double b1r = xr[s + 1] + xr[s + 3], b1i = xi[s + 1] + xi[s + 3];
        const double b3r = xi[s + 3] - xi[s + 1], b3i = xr[s + 1] - xr[s + 3];
        ar[s] = b0r + b1r; ai[s] = b0i + b1i;
        ar[s + 2] = b0r - b1r; ai[s + 2] = b0i - b1i;
        ar[s + 1] = b2r + b3r; ai[s + 1] = b2i + b3i;
        ar[s + 3] = b2r - b3r; ai[s + 3] = b2i - b3i;
    }
    for (int j = 0; j < 4; j++) {     /* span 16 */
        const double u1r = ar[j + 4] * t16[j] + ai[j + 4] * t16[4 + j], u1i = ai[j + 4] * t16[j] - ar[j + 4] * t16[4 + j];
        const double u2r = ar[j + 8] * t16[8 + j] + ai[j + 8] * t16[12 + j], u2i = ai[j + 8] * t16[8 + j] - ar[j + 8] * t16[12 + j];
        const double u3r = ar[j + 12] * t16[16 + j] + ai[j + 12] * t16[20 + j], u3i = ai[j + 12] * t16[16 + j] - ar[j + 12] * t16[20 + j];
        const double b0r = ar[j] + u2r, b0i = ai[j] + u2i, b2r = ar[j] - u2r, b2i = ai[j] - u2i;
        const double b1r = u1r + u3r, b1i = u1i + u3i, b3r = u3i - u1i, b3i = u1r - u3r;
        xr[j] = b0r + b1r; xi[j] = b0i + b1i;
        xr[j + 8] = b0r - b1r; xi[j + 8] = b0i - b1i;
        xr[j + 4] = b2r + b3r; xi[j + 4] = b2i + b3i;
        xr[j + 12] = b2r - b3r; xi[j + 12] = b2i - b3i;
    }
}

/* forward DFT (sign -1), natural order in, digit-reversed order out; in place on SoA data.
 * H must be a power of 4 with H >= 16 (N = 2048: H = 1024 = 4^5). */
static void fast_fwd(const orc_keyset *ks, double *restrict xr, double *restrict xi) {
    const long H = ks->p.poly_size / 2;
    const double *t = ks->ftw;
    long m = H;
    for (; m > 16; m >>= 2) {
        const long q = m / 4;
        for (long s = 0; s < H; s += m)
            r4_fwd_block(xr + s, xr + s + q, xr + s + 2 * q, xr + s + 3 * q, xi + s, xi + s + q, xi + s + 2 * q, xi + s + 3 * q, t, q);
        t += q * 6;
    }
    for (long s = 0; s < H; s += 16) r4_fwd_tail16(xr + s, xi + s, t);
}

/* exact mirror of fast_fwd: digit-reversed in, natural out, unnormalised (x H) */
static void fast_inv(const orc_keyset *ks, double *restrict xr, double *restrict xi) {
    const long H = ks->p.poly_size / 2;
    long total = 0;
    for (long m = H; m >= 4; m >>= 2) total += (m / 4) * 6;
    const double *t = ks->ftw + total - 6 - 24;   /* tables of spans 16 (24 doubles) and 4 (6) come last */
    for (long s = 0; s < H; s += 16) r4_inv_head16(xr + s, xi + s, t);
    for (long m = 64; m <= H; m <<= 2) {
        const long q = m / 4;
        t -= q * 6;
        for (long s = 0; s < H; s += m)
            r4_inv_block(xr + s, xr + s + q, xr + s + 2 * q, xr + s + 3 * q, xi + s, xi + s + q, xi + s + 2 * q, xi + s + 3 * q, t, q);
    }
}
#endif /* ORC_FAST */

/* ----------------------------------------------------------------------------------------- */
/* A.3 signed decomposer                                                                      */
static inline uint64_t closest_state(uint64_t x, int base_log, int level) {
    int r = 64 - base_log * level;
    uint64_t msb = (x >> (r - 1)) & 1ULL;
    uint64_t res = (x >> r) + msb;
    /* the wrap of `res << r` in tfhe-rs is a reduction mod 2^(base_log*level) of the state */
    int bits = base_log * level;
    if (bits < 64) res &= ((1ULL << bits) - 1ULL);
    return res;
}
static inline int64_t decomp_next(uint64_t *state, int base_log) {
    uint64_t mask = (1ULL << base_log) - 1ULL;
    uint64_t res = *state & mask;
    *state >>= base_log;
    uint64_t carry = ((res - 1ULL) | *state) & res;
    carry >>= (base_log - 1);
    *state += carry;
    return (int64_t)(res - (carry << base_log));
}
void orc_decompose(uint64_t x, int base_log, int level, int64_t *digits_out) {
    uint64_t st = closest_state(x, base_log, level);
    for (int l = 0; l < level; l++) digits_out[l] = decomp_next(&st, base_log);
}

/* ----------------------------------------------------------------------------------------- */
/* negacyclic helpers                                                                         */
/* out = in * X^a, a in [0, 2N) */
static void poly_mul_monomial(const uint64_t *in, uint32_t a, uint32_t N, uint64_t *out) {
    for (uint32_t j = 0; j < N; j++) {
        uint32_t idx = (j + 2 * N - a) % (2 * N);
        out[j] = idx < N ? in[idx] : (uint64_t)0 - in[idx - N];
    }
}
/* body += mask * S for a binary secret polynomial S (sum of rotations) */
static void poly_add_mul_binary(uint64_t *body, const uint64_t *mask, const uint64_t *sbits,
                                uint32_t N) {
    for (uint32_t j = 0; j < N; j++) {
        if (!sbits[j]) continue;
        /* mask * X^j */
        for (uint32_t t = 0; t < N - j; t++) body[t + j] += mask[t];
        for (uint32_t t = N - j; t < N; t++) body[t + j - N] -= mask[t];
    }
}

/* ----------------------------------------------------------------------------------------- */
orc_keyset *orc_keygen(const orc_params *p, uint64_t seed) {
    orc_keyset *ks = (orc_keyset *)calloc(1, sizeof(orc_keyset));
    ks->p = *p;
    const uint32_t n = p->lwe_dim, k = p->glwe_dim, N = p->poly_size;
    const uint32_t kN = k * N, k1 = k + 1;
    ks->lwe_sk = (uint64_t *)malloc(sizeof(uint64_t) * n);
    ks->glwe_sk = (uint64_t *)malloc(sizeof(uint64_t) * kN);
    orc_rng r;
    rng_seed(&r, seed, 1);
    for (uint32_t i = 0; i < kN; i++) ks->glwe_sk[i] = rng_u64(&r) >> 63;
    rng_seed(&r, seed, 2);
    for (uint32_t i = 0; i < n; i++) ks->lwe_sk[i] = rng_u64(&r) >> 63;
    fft_tables(ks);

    /* KSK big -> small (A.4): block i, level l (1-based) encrypts s_big[i] * 2^(64 - b*l). */
    const uint32_t kl = p->ks_level, kb = p->ks_base_log;
    ks->ksk = (uint64_t *)malloc(sizeof(uint64_t) * orc_ksk_len(ks));
#pragma omp parallel for schedule(static)
    for (uint32_t i = 0; i < kN; i++) {
        orc_rng rr;
        rng_seed(&rr, seed, 0x1000000ULL + i);
        for (uint32_t l = 1; l <= kl; l++) {
            uint64_t *row = ks->ksk + ((size_t)i * kl + (l - 1)) * (n + 1);
            uint64_t acc = 0;
            for (uint32_t j = 0; j < n; j++) {
                row[j] = rng_u64(&rr);
                if (ks->lwe_sk[j]) acc += row[j];
            }
            uint64_t pt = ks->glwe_sk[i] << (64 - kb * l);
            row[n] = acc + pt + rng_noise(&rr, p->lwe_std);
        }
    }

    /* BSK (A.6): GGSW_i encrypts bit s_i of the small key under the GLWE key. */
    const uint32_t bl = p->pbs_level, bb = p->pbs_base_log;
    ks->bsk = (uint64_t *)malloc(sizeof(uint64_t) * orc_bsk_len(ks));
    ks->bsk_f = (cplx *)malloc(sizeof(cplx) * orc_bsk_len(ks) / 2);
#pragma omp parallel for schedule(dynamic, 4)
    for (uint32_t i = 0; i < n; i++) {
        orc_rng rr;
        rng_seed(&rr, seed, 0x2000000ULL + i);
        const uint64_t m = ks->lwe_sk[i];
        for (uint32_t l = 1; l <= bl; l++) {
            const uint64_t factor = ((uint64_t)0 - m) << (64 - bb * l); /* -m * q/B^l */
            for (uint32_t row = 0; row < k1; row++) {
                uint64_t *glwe = ks->bsk + ((((size_t)i * bl + (l - 1)) * k1 + row) * k1) * N;
                uint64_t *body = glwe + (size_t)k * N;
                /* plaintext polynomial */
                if (row < k) {
                    for (uint32_t j = 0; j < N; j++) body[j] = factor * ks->glwe_sk[row * N + j];
                } else {
                    memset(body, 0, sizeof(uint64_t) * N);
                    body[0] = (uint64_t)0 - factor;
                }
                for (uint32_t c = 0; c < k; c++) {
                    uint64_t *mask = glwe + (size_t)c * N;
                    for (uint32_t j = 0; j < N; j++) mask[j] = rng_u64(&rr);
                    poly_add_mul_binary(body, mask, ks->glwe_sk + (size_t)c * N, N);
                }
                for (uint32_t j = 0; j < N; j++) body[j] += rng_noise(&rr, p->glwe_std);
                for (uint32_t c = 0; c < k1; c++)
                    fft_fwd_torus(ks, glwe + (size_t)c * N,
                                  ks->bsk_f +
                                      (((((size_t)i * bl + (l - 1)) * k1 + row) * k1) + c) * (N / 2));
            }
        }
    }
#ifdef ORC_FAST
    fast_tables(ks);
    fast_bsk(ks);
#endif
    return ks;
}

void orc_keyset_free(orc_keyset *ks) {
    if (!ks) return;
    free(ks->lwe_sk);
    free(ks->glwe_sk);
    free(ks->ksk);
    free(ks->bsk);
    free(ks->bsk_f);
    free(ks->twist);
    free(ks->tw);
    free(ks->brev);
#ifdef ORC_FAST
    free(ks->ftw); free(ks->twr); free(ks->twi); free(ks->bsk_fr); free(ks->bsk_fi);
#endif
    free(ks);
}

/* ----------------------------------------------------------------------------------------- */
/* A.2 encoding                                                                               */
static inline uint64_t delta_of(const orc_params *p) {
    return (1ULL << 63) / ((uint64_t)p->message_modulus * p->carry_modulus);
}
void orc_encrypt_block(const orc_keyset *ks, uint64_t m, uint64_t rng_seed_, uint64_t *ct) {
    const uint32_t kN = ks->p.glwe_dim * ks->p.poly_size;
    orc_rng r;
    rng_seed(&r, rng_seed_, 0x3000000ULL);
    uint64_t acc = 0;
    for (uint32_t j = 0; j < kN; j++) {
        ct[j] = rng_u64(&r);
        if (ks->glwe_sk[j]) acc += ct[j];
    }
    ct[kN] = acc + m * delta_of(&ks->p) + rng_noise(&r, ks->p.glwe_std);
}
void orc_trivial_block(const orc_keyset *ks, uint64_t m, uint64_t *ct) {
    const uint32_t kN = ks->p.glwe_dim * ks->p.poly_size;
    memset(ct, 0, sizeof(uint64_t) * kN);
    ct[kN] = m * delta_of(&ks->p);
}
uint64_t orc_phase_big(const orc_keyset *ks, const uint64_t *ct) {
    const uint32_t kN = ks->p.glwe_dim * ks->p.poly_size;
    uint64_t acc = 0;
    for (uint32_t j = 0; j < kN; j++)
        if (ks->glwe_sk[j]) acc += ct[j];
    return ct[kN] - acc;
}
uint64_t orc_phase_small(const orc_keyset *ks, const uint64_t *ct) {
    const uint32_t n = ks->p.lwe_dim;
    uint64_t acc = 0;
    for (uint32_t j = 0; j < n; j++)
        if (ks->lwe_sk[j]) acc += ct[j];
    return ct[n] - acc;
}
uint64_t orc_decode(const orc_keyset *ks, uint64_t phase) {
    const uint64_t d = delta_of(&ks->p);
    const uint64_t space = 2ULL * ks->p.message_modulus * ks->p.carry_modulus;
    return ((phase + d / 2) / d) % space;
}

/* ----------------------------------------------------------------------------------------- */
/* A.8 / a1.1 LUT accumulator: boxes of N/(msg*carry), first half-box negated, rotate_left.   */
void orc_build_lut(const orc_params *p, const uint64_t *table, uint64_t *acc) {
    const uint32_t N = p->poly_size, k = p->glwe_dim;
    const uint32_t space = p->message_modulus * p->carry_modulus;
    const uint32_t box = N / space, half = box / 2;
    const uint64_t d = delta_of(p);
    memset(acc, 0, sizeof(uint64_t) * (size_t)(k + 1) * N);
    uint64_t *body = acc + (size_t)k * N;
    uint64_t *tmp = (uint64_t *)malloc(sizeof(uint64_t) * N);
    for (uint32_t i = 0; i < space; i++)
        for (uint32_t j = 0; j < box; j++) tmp[i * box + j] = table[i] * d;
    for (uint32_t j = 0; j < half; j++) tmp[j] = (uint64_t)0 - tmp[j];
    for (uint32_t j = 0; j < N; j++) body[j] = tmp[(j + half) % N]; /* rotate_left(half) */
    free(tmp);
}

/* A.4 / a1.4 keyswitch: out = (0,..,0,b) - sum_i sum_l digit_{i,l} * KSK[i][l]               */
void orc_keyswitch(const orc_keyset *ks, const uint64_t *in, uint64_t *out) {
    const orc_params *p = &ks->p;
    const uint32_t n = p->lwe_dim, kN = p->glwe_dim * p->poly_size;
    const uint32_t kl = p->ks_level, kb = p->ks_base_log;
    memset(out, 0, sizeof(uint64_t) * (n + 1));
    out[n] = in[kN];
    for (uint32_t i = 0; i < kN; i++) {
        uint64_t st = closest_state(in[i], kb, kl);
        /* digits come least-significant (level kl) first; rows are stored level 1 first and
         * visited in reverse (block.iter().rev() zipped with the decomposition iterator). */
        for (uint32_t l = kl; l >= 1; l--) {
            int64_t dg = decomp_next(&st, kb);
            if (dg == 0) continue;
            const uint64_t *row = ks->ksk + ((size_t)i * kl + (l - 1)) * (n + 1);
            uint64_t d = (uint64_t)dg;
            for (uint32_t j = 0; j <= n; j++) out[j] -= d * row[j];
        }
    }
}

/* A.5 / a1.5 */
uint64_t orc_modswitch(const orc_params *p, uint64_t x) {
    int lg = 0;
    while ((1u << lg) < p->poly_size) lg++;
    /* ((x >> (64 - log2(N) - 2)) + 1) >> 1  in [0, 2N] */
    return ((x >> (64 - lg - 2)) + 1ULL) >> 1;
}


#ifdef ORC_FAST
/* torus polynomial (u64, read as signed, scaled 2^-64) -> fast-order spectrum, SoA */
static void fast_fwd_torus(const orc_keyset *ks, const uint64_t *poly, double *restrict fr, double *restrict fi) {
    const uint32_t H = ks->p.poly_size / 2;
    const double sc = 1.0 / 18446744073709551616.0;
    for (uint32_t j = 0; j < H; j++) {
        const double a = (double)(int64_t)poly[j] * sc, b = (double)(int64_t)poly[j + H] * sc;
        fr[j] = a * ks->twr[j] - b * ks->twi[j];
        fi[j] = a * ks->twi[j] + b * ks->twr[j];
    }
    fast_fwd(ks, fr, fi);
}

static void fast_bsk(orc_keyset *ks) {
    const uint32_t n = ks->p.lwe_dim, N = ks->p.poly_size, H = N / 2, k1 = ks->p.glwe_dim + 1, bl = ks->p.pbs_level;
    const size_t polys = (size_t)n * bl * k1 * k1;
    ks->bsk_fr = (double *)aligned_alloc(64, polys * H * sizeof(double));
    ks->bsk_fi = (double *)aligned_alloc(64, polys * H * sizeof(double));
#pragma omp parallel for schedule(static)
    for (long q = 0; q < (long)polys; q++)
        fast_fwd_torus(ks, ks->bsk + (size_t)q * N, ks->bsk_fr + (size_t)q * H, ks->bsk_fi + (size_t)q * H);
}

/* per-thread scratch of the blind rotation, allocated once */
typedef struct {
    uint32_t N;
    uint64_t *ct1;
    double *dr, *di;     /* digit spectrum */
    double *or_, *oi;    /* output spectra, k1 polynomials */
    uint64_t *small, *acc;
} fast_scratch;
static __thread fast_scratch tls_fs;

static fast_scratch *fast_get_scratch(const orc_params *p) {
    const uint32_t N = p->poly_size, k1 = p->glwe_dim + 1, H = N / 2;
    if (tls_fs.N != N) {
        free(tls_fs.ct1); free(tls_fs.dr); free(tls_fs.di); free(tls_fs.or_); free(tls_fs.oi); free(tls_fs.small); free(tls_fs.acc);
        tls_fs.ct1 = (uint64_t *)aligned_alloc(64, sizeof(uint64_t) * (size_t)k1 * N);
        tls_fs.dr = (double *)aligned_alloc(64, sizeof(double) * H);
        tls_fs.di = (double *)aligned_alloc(64, sizeof(double) * H);
        tls_fs.or_ = (double *)aligned_alloc(64, sizeof(double) * (size_t)k1 * H);
        tls_fs.oi = (double *)aligned_alloc(64, sizeof(double) * (size_t)k1 * H);
        tls_fs.small = (uint64_t *)aligned_alloc(64, ((sizeof(uint64_t) * (p->lwe_dim + 1) + 63) / 64) * 64);
        tls_fs.acc = (uint64_t *)aligned_alloc(64, sizeof(uint64_t) * (size_t)k1 * N);
        tls_fs.N = N;
    }
    return &tls_fs;
}

/* acc += GGSW_i (x) ct1, one PBS level (the parameter family of this path has pbs_level = 1; other
 * level counts go through the generic code) */
static void fast_external_product(const orc_keyset *ks, uint32_t i, const uint64_t *ct1, uint64_t *acc, fast_scratch *fs) {
    const orc_params *p = &ks->p;
    const uint32_t N = p->poly_size, H = N / 2, k1 = p->glwe_dim + 1;
    const int bb = (int)p->pbs_base_log, r = 64 - bb;
    const uint64_t half = 1ULL << (bb - 1), mask = (1ULL << bb) - 1ULL;
    memset(fs->or_, 0, sizeof(double) * (size_t)k1 * H);
    memset(fs->oi, 0, sizeof(double) * (size_t)k1 * H);
    for (uint32_t row = 0; row < k1; row++) {
        const uint64_t *c = ct1 + (size_t)row * N;
        double *restrict dr = fs->dr, *restrict di = fs->di;
        for (uint32_t j = 0; j < H; j++) {
            /* one-level signed digit: closest multiple of 2^r, balanced representative (A.3 with l = 1) */
            uint64_t sa = ((c[j] >> r) + ((c[j] >> (r - 1)) & 1ULL)) & mask;
            uint64_t sb = ((c[j + H] >> r) + ((c[j + H] >> (r - 1)) & 1ULL)) & mask;
            const double a = (double)((int64_t)sa - (int64_t)((sa > half) ? (mask + 1ULL) : 0ULL));
            const double b = (double)((int64_t)sb - (int64_t)((sb > half) ? (mask + 1ULL) : 0ULL));
            dr[j] = a * ks->twr[j] - b * ks->twi[j];
            di[j] = a * ks->twi[j] + b * ks->twr[j];
        }
        fast_fwd(ks, dr, di);
        for (uint32_t cc = 0; cc < k1; cc++) {
            const size_t q = (((size_t)i * k1 + row) * k1 + cc) * H;
            const double *restrict br = ks->bsk_fr + q, *restrict bi = ks->bsk_fi + q;
            double *restrict o_r = fs->or_ + (size_t)cc * H, *restrict o_i = fs->oi + (size_t)cc * H;
            for (uint32_t t = 0; t < H; t++) {
                o_r[t] += dr[t] * br[t] - di[t] * bi[t];
                o_i[t] += dr[t] * bi[t] + di[t] * br[t];
            }
        }
    }
    const double inv = 1.0 / (double)H, two64 = 18446744073709551616.0;
    for (uint32_t cc = 0; cc < k1; cc++) {
        double *restrict o_r = fs->or_ + (size_t)cc * H, *restrict o_i = fs->oi + (size_t)cc * H;
        fast_inv(ks, o_r, o_i);
        uint64_t *restrict pa = acc + (size_t)cc * N;
        for (uint32_t j = 0; j < H; j++) {
            double re = (o_r[j] * ks->twr[j] + o_i[j] * ks->twi[j]) * inv;
            double im = (o_i[j] * ks->twr[j] - o_r[j] * ks->twi[j]) * inv;
            re -= __builtin_rint(re);
            im -= __builtin_rint(im);
            pa[j] += (uint64_t)(int64_t)(re * two64);       /* |re| <= 0.5: the product fits except at +0.5 */
            pa[j + H] += (uint64_t)(int64_t)(im * two64);
        }
    }
}

/* ct1 = acc * X^a - acc without a modulo per element */
static void fast_rotate_sub(const uint64_t *restrict acc, uint32_t a, uint32_t N, uint64_t *restrict out) {
    /* (acc * X^a)[j] = acc[j - a] for j >= a, -acc[N + j - a] for j < a  (a < N); a >= N: negate and use a - N */
    const int neg = a >= N;
    if (neg) a -= N;
    if (!neg) {
        for (uint32_t j = 0; j < a; j++) out[j] = (uint64_t)0 - acc[N + j - a] - acc[j];
        for (uint32_t j = a; j < N; j++) out[j] = acc[j - a] - acc[j];
    } else {
        for (uint32_t j = 0; j < a; j++) out[j] = acc[N + j - a] - acc[j];
        for (uint32_t j = a; j < N; j++) out[j] = (uint64_t)0 - acc[j - a] - acc[j];
    }
}
#endif /* ORC_FAST */

/* external product accumulate: acc += GGSW_i (x) ct1                                          */
static void add_external_product(const orc_keyset *ks, uint32_t i, const uint64_t *ct1,
                                 uint64_t *acc, int64_t *dig, uint64_t *states, cplx *fdig,
                                 cplx *fout) {
    const orc_params *p = &ks->p;
    const uint32_t N = p->poly_size, H = N / 2, k1 = p->glwe_dim + 1;
    const uint32_t bl = p->pbs_level, bb = p->pbs_base_log;
    for (size_t j = 0; j < (size_t)k1 * N; j++) states[j] = closest_state(ct1[j], bb, bl);
    memset(fout, 0, sizeof(cplx) * (size_t)k1 * H);
    for (uint32_t l = bl; l >= 1; l--) { /* least significant level first */
        for (size_t j = 0; j < (size_t)k1 * N; j++) dig[j] = decomp_next(&states[j], bb);
        for (uint32_t r = 0; r < k1; r++) {
            fft_fwd_int(ks, dig + (size_t)r * N, fdig);
            for (uint32_t c = 0; c < k1; c++) {
                const cplx *b =
                    ks->bsk_f + (((((size_t)i * bl + (l - 1)) * k1 + r) * k1) + c) * H;
                cplx *o = fout + (size_t)c * H;
                for (uint32_t t = 0; t < H; t++) {
                    o[t].re += fdig[t].re * b[t].re - fdig[t].im * b[t].im;
                    o[t].im += fdig[t].re * b[t].im + fdig[t].im * b[t].re;
                }
            }
        }
    }
    for (uint32_t c = 0; c < k1; c++) fft_bwd_add_torus(ks, fout + (size_t)c * H, acc + (size_t)c * N);
}

/* Exposed for tests/test_cpu_oracle_first_principles.py: one acc += GGSW_i (x) ct1 step, so that the
 * FFT algebra above can be compared with an exact integer negacyclic convolution. */
void orc_external_product(const orc_keyset *ks, uint32_t i, const uint64_t *ct1, uint64_t *acc) {
    const uint32_t N = ks->p.poly_size, k1 = ks->p.glwe_dim + 1;
    uint64_t *states = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)k1 * N);
    int64_t *dig = (int64_t *)malloc(sizeof(int64_t) * (size_t)k1 * N);
    cplx *fdig = (cplx *)malloc(sizeof(cplx) * (N / 2));
    cplx *fout = (cplx *)malloc(sizeof(cplx) * (size_t)k1 * (N / 2));
    add_external_product(ks, i, ct1, acc, dig, states, fdig, fout);
    free(states);
    free(dig);
    free(fdig);
    free(fout);
}

/* A.6 / a1.6 blind rotation */
void orc_blind_rotate(const orc_keyset *ks, const uint64_t *lwe, uint64_t *acc) {
    const orc_params *p = &ks->p;
    const uint32_t n = p->lwe_dim, N = p->poly_size, k1 = p->glwe_dim + 1;
#ifdef ORC_FAST
    if (p->pbs_level == 1) {
        fast_scratch *fs = fast_get_scratch(p);
        uint32_t bt0 = (uint32_t)(orc_modswitch(p, lwe[n]) % (2 * N));
        for (uint32_t c = 0; c < k1; c++)
            poly_mul_monomial(acc + (size_t)c * N, (2 * N - bt0) % (2 * N), N, fs->ct1 + (size_t)c * N);
        memcpy(acc, fs->ct1, sizeof(uint64_t) * (size_t)k1 * N);
        for (uint32_t i = 0; i < n; i++) {
            uint32_t a = (uint32_t)(orc_modswitch(p, lwe[i]) % (2 * N));
            if (a == 0) continue;
            for (uint32_t c = 0; c < k1; c++) fast_rotate_sub(acc + (size_t)c * N, a, N, fs->ct1 + (size_t)c * N);
            fast_external_product(ks, i, fs->ct1, acc, fs);
        }
        return;
    }
#endif
    uint64_t *ct1 = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)k1 * N);
    uint64_t *states = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)k1 * N);
    int64_t *dig = (int64_t *)malloc(sizeof(int64_t) * (size_t)k1 * N);
    cplx *fdig = (cplx *)malloc(sizeof(cplx) * (N / 2));
    cplx *fout = (cplx *)malloc(sizeof(cplx) * (size_t)k1 * (N / 2));

    uint32_t bt = (uint32_t)(orc_modswitch(p, lwe[n]) % (2 * N));
    /* acc <- acc * X^{-b~}  (monic monomial division) */
    for (uint32_t c = 0; c < k1; c++)
        poly_mul_monomial(acc + (size_t)c * N, (2 * N - bt) % (2 * N), N, ct1 + (size_t)c * N);
    memcpy(acc, ct1, sizeof(uint64_t) * (size_t)k1 * N);

    for (uint32_t i = 0; i < n; i++) {
        uint32_t a = (uint32_t)(orc_modswitch(p, lwe[i]) % (2 * N));
        if (a == 0) continue;
        for (uint32_t c = 0; c < k1; c++) {
            poly_mul_monomial(acc + (size_t)c * N, a, N, ct1 + (size_t)c * N);
            for (uint32_t j = 0; j < N; j++) ct1[(size_t)c * N + j] -= acc[(size_t)c * N + j];
        }
        add_external_product(ks, i, ct1, acc, dig, states, fdig, fout);
    }
    free(ct1);
    free(states);
    free(dig);
    free(fdig);
    free(fout);
}

/* A.7 / a1.7 */
void orc_sample_extract(const orc_params *p, const uint64_t *acc, uint64_t *out) {
    const uint32_t N = p->poly_size, k = p->glwe_dim;
    for (uint32_t c = 0; c < k; c++) {
        const uint64_t *a = acc + (size_t)c * N;
        uint64_t *o = out + (size_t)c * N;
        o[0] = a[0];
        for (uint32_t j = 1; j < N; j++) o[j] = (uint64_t)0 - a[N - j];
    }
    out[(size_t)k * N] = acc[(size_t)k * N];
}

/* a1.3 */
void orc_ks_pbs(const orc_keyset *ks, const uint64_t *lwe_in, const uint64_t *table,
                uint64_t *lwe_out) {
    const orc_params *p = &ks->p;
    const uint32_t N = p->poly_size, k1 = p->glwe_dim + 1;
#ifdef ORC_FAST
    fast_scratch *fs = fast_get_scratch(p);
    orc_keyswitch(ks, lwe_in, fs->small);
    orc_build_lut(p, table, fs->acc);
    orc_blind_rotate(ks, fs->small, fs->acc);
    orc_sample_extract(p, fs->acc, lwe_out);
    return;
#endif
    uint64_t *small = (uint64_t *)malloc(sizeof(uint64_t) * (p->lwe_dim + 1));
    uint64_t *acc = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)k1 * N);
    orc_keyswitch(ks, lwe_in, small);
    orc_build_lut(p, table, acc);
    orc_blind_rotate(ks, small, acc);
    orc_sample_extract(p, acc, lwe_out);
    free(small);
    free(acc);
}

int orc_ks_pbs_batch(const orc_keyset *ks, const uint64_t *lwe_in, const uint32_t *lut_id,
                     const uint64_t *lut_tables, size_t batch, uint64_t *lwe_out,
                     int num_threads) {
    const orc_params *p = &ks->p;
    const size_t big = (size_t)p->glwe_dim * p->poly_size + 1;
    const uint32_t space = p->message_modulus * p->carry_modulus;
    int used = 1;
#ifdef _OPENMP
    if (num_threads <= 0) num_threads = omp_get_max_threads();
    used = num_threads;
#pragma omp parallel for schedule(dynamic, 1) num_threads(num_threads)
#endif
    for (long b = 0; b < (long)batch; b++) {
        orc_ks_pbs(ks, lwe_in + (size_t)b * big, lut_tables + (size_t)lut_id[b] * space,
                   lwe_out + (size_t)b * big);
    }
    (void)num_threads;
    return used;
}
