/*
 * fhs_oracle.c -- CPU ORACLE (test infrastructure) restating the tfhe-rs 0.5 KS-PBS path that
 * sits under every fhe_string operation.  PARITY UNPINNED at the tfhe-rs boundary -- see the
 * header comment in fhs_oracle.h.  Plain C, no dependencies (OpenMP optional).
 *
 * Reference call sites that reach this arithmetic: every `k.k.<op>` in
 * /root/reference/src/ciphertext/*.rs, e.g. compare.rs:32, search.rs:17-41.
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
