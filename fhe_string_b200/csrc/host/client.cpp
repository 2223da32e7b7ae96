// Client side of the fhe_string API on the host CPU (reference: src/lib.rs:44-77 generate_keys*,
// src/client_key.rs:16-62, src/ciphertext/mod.rs:35-37,58-92,121-129): key generation for the
// KS-PBS parameter sets, block / string encryption under the big key, decryption.  This is one-off
// CPU work in the reference as well (tfhe-rs gen_keys_radix, RadixClientKey::encrypt) and is not on
// the accelerated path; it exists so that the package is usable without any test infrastructure.
// Randomness: ChaCha20 (RFC 8439 block function, 64-bit block counter + 64-bit stream id) keyed with
// TWO independent 256-bit keys drawn from the OS (getrandom): one for everything secret (key bits, noise),
// one for the public masks (KSK / BSK rows, ciphertext a-vectors) -- the split tfhe-rs makes between its
// secret and encryption generators (concrete-csprng, AES-CTR there).  Public masks are raw generator
// output, so the generator must be one whose output does not reveal its key; ADVICE r1 showed that the
// earlier xoshiro256** streams did.  A caller seed != 0 derives both keys from that 64-bit value: the
// REPRODUCIBLE TEST / BENCH MODE, at most 64 bits of security, never for real data (README).
#include <sys/random.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../../include/fhs_string.h"

namespace {

struct ChaKey { uint32_t w[8]; };

inline uint32_t rotl32(uint32_t x, int k) { return (x << k) | (x >> (32 - k)); }
#define FHS_QR(a, b, c, d) \
    a += b; d ^= a; d = rotl32(d, 16); c += d; b ^= c; b = rotl32(b, 12); \
    a += b; d ^= a; d = rotl32(d, 8);  c += d; b ^= c; b = rotl32(b, 7);

// One ChaCha20 keystream per (key, stream id); 64-byte blocks, 64-bit block counter.
struct Rng {
    uint32_t st[16];
    uint64_t buf[8];
    int pos = 8;
    Rng(const ChaKey &key, uint64_t stream) {
        st[0] = 0x61707865u; st[1] = 0x3320646eu; st[2] = 0x79622d32u; st[3] = 0x6b206574u;
        for (int i = 0; i < 8; i++) st[4 + i] = key.w[i];
        st[12] = 0; st[13] = 0;
        st[14] = (uint32_t)stream; st[15] = (uint32_t)(stream >> 32);
    }
    void refill() {
        uint32_t x[16];
        for (int i = 0; i < 16; i++) x[i] = st[i];
        for (int r = 0; r < 10; r++) {
            FHS_QR(x[0], x[4], x[8], x[12]) FHS_QR(x[1], x[5], x[9], x[13])
            FHS_QR(x[2], x[6], x[10], x[14]) FHS_QR(x[3], x[7], x[11], x[15])
            FHS_QR(x[0], x[5], x[10], x[15]) FHS_QR(x[1], x[6], x[11], x[12])
            FHS_QR(x[2], x[7], x[8], x[13]) FHS_QR(x[3], x[4], x[9], x[14])
        }
        for (int i = 0; i < 8; i++)
            buf[i] = (uint64_t)(x[2 * i] + st[2 * i]) | ((uint64_t)(x[2 * i + 1] + st[2 * i + 1]) << 32);
        if (++st[12] == 0) ++st[13];
        pos = 0;
    }
    uint64_t next() { if (pos == 8) refill(); return buf[pos++]; }
    double unit() { return ((double)(next() >> 11) + 0.5) * (1.0 / 9007199254740992.0); }
    uint64_t torus_noise(double std) {  // round(N(0, (std 2^64)^2)) mod 2^64
        const double u1 = unit(), u2 = unit();
        const double g = std::sqrt(-2.0 * std::log(u1)) * std::cos(6.283185307179586 * u2);
        return (uint64_t)(int64_t)std::llrint(g * std * 18446744073709551616.0);
    }
};
#undef FHS_QR

bool os_random(void *dst, size_t n) {
    uint8_t *p = (uint8_t *)dst;
    while (n) {
        ssize_t got = getrandom(p, n, 0);
        if (got <= 0) return false;
        p += got; n -= (size_t)got;
    }
    return true;
}

}  // namespace

struct fhs_client_key {
    fhs_params p;
    double lwe_std, glwe_std;
    uint32_t nb_char, nb_usize;
    size_t max_len;
    ChaKey secret_key, public_key;         // secret: key bits + noise; public: masks
    uint64_t enc_counter = 0;
    bool reproducible = false;             // keys derived from a caller seed (test / bench mode)
    std::vector<uint8_t> lwe_sk, glwe_sk;  // bits
    std::vector<uint64_t> ksk, bsk;        // server key material (fhs_upload_keys layouts)
};

namespace {

thread_local std::string c_err;

uint32_t ceil_log(uint64_t base, uint64_t v) {  // smallest e with base^e >= v  (lib.rs:51-62)
    uint32_t e = 0;
    uint64_t x = 1;
    while (x < v) { x *= base; e++; }
    return e;
}

void gen_ksk(fhs_client_key &k, int nthreads) {
    const uint32_t n = k.p.lwe_dim, big = k.p.glwe_dim * k.p.poly_size, lv = k.p.ks_level, bl = k.p.ks_base_log;
    k.ksk.assign((size_t)big * lv * (n + 1), 0);
    auto work = [&](uint32_t i0, uint32_t i1) {
        for (uint32_t i = i0; i < i1; i++) {
            Rng r(k.public_key, 0x10000000ull + i), e(k.secret_key, 0x10000000ull + i);
            for (uint32_t l = 1; l <= lv; l++) {  // rows level 1 (most significant) .. lv: FHS_KSK_LEVEL_MSB_FIRST
                uint64_t *row = &k.ksk[((size_t)i * lv + (l - 1)) * (n + 1)];
                uint64_t dot = 0;
                for (uint32_t j = 0; j < n; j++) { row[j] = r.next(); if (k.lwe_sk[j]) dot += row[j]; }
                row[n] = dot + ((uint64_t)k.glwe_sk[i] << (64 - bl * l)) + e.torus_noise(k.lwe_std);
            }
        }
    };
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; t++) th.emplace_back(work, (uint32_t)((uint64_t)big * t / nthreads), (uint32_t)((uint64_t)big * (t + 1) / nthreads));
    for (auto &t : th) t.join();
}

// GGSW(s_i) rows: for level l, row r < k: GLWE encryption of -s_i * S_r(X) * q/B^l; row k: of s_i * q/B^l
void gen_bsk(fhs_client_key &k, int nthreads) {
    const uint32_t n = k.p.lwe_dim, K = k.p.glwe_dim, N = k.p.poly_size, lv = k.p.pbs_level, bl = k.p.pbs_base_log;
    const uint32_t k1 = K + 1;
    k.bsk.assign((size_t)n * lv * k1 * k1 * N, 0);
    std::vector<std::vector<uint32_t>> ones(K);  // positions of the set key bits per key polynomial
    for (uint32_t c = 0; c < K; c++)
        for (uint32_t j = 0; j < N; j++) if (k.glwe_sk[c * N + j]) ones[c].push_back(j);
    auto work = [&](uint32_t i0, uint32_t i1) {
        for (uint32_t i = i0; i < i1; i++) {
            Rng r(k.public_key, 0x20000000ull + i), e(k.secret_key, 0x20000000ull + i);
            const uint64_t bit = k.lwe_sk[i];
            for (uint32_t l = 1; l <= lv; l++) {
                const uint64_t scale = bit << (64 - bl * l);
                for (uint32_t row = 0; row < k1; row++) {
                    uint64_t *glwe = &k.bsk[((((size_t)i * lv + (l - 1)) * k1 + row) * k1) * N];
                    uint64_t *body = glwe + (size_t)K * N;
                    if (row < K) { for (uint32_t j = 0; j < N; j++) body[j] = (0 - scale) * k.glwe_sk[row * N + j]; }
                    else body[0] = scale;
                    for (uint32_t c = 0; c < K; c++) {
                        uint64_t *mask = glwe + (size_t)c * N;
                        for (uint32_t j = 0; j < N; j++) mask[j] = r.next();
                        for (uint32_t pos : ones[c]) {  // body += mask * X^pos (negacyclic)
                            for (uint32_t t = 0; t < N - pos; t++) body[t + pos] += mask[t];
                            for (uint32_t t = N - pos; t < N; t++) body[t + pos - N] -= mask[t];
                        }
                    }
                    for (uint32_t j = 0; j < N; j++) body[j] += e.torus_noise(k.glwe_std);
                }
            }
        }
    };
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; t++) th.emplace_back(work, (uint32_t)((uint64_t)n * t / nthreads), (uint32_t)((uint64_t)n * (t + 1) / nthreads));
    for (auto &t : th) t.join();
}

}  // namespace

extern "C" {

const char *fhs_client_last_error(void) { return c_err.c_str(); }

// Self-test hook: one ChaCha20 block for key[8] and state words 12..15 (RFC 8439 2.3.2 known answer).
void fhs_chacha20_block(const uint32_t key[8], const uint32_t ctr_nonce[4], uint8_t out[64]) {
    ChaKey k;
    memcpy(k.w, key, sizeof(k.w));
    Rng r(k, 0);
    for (int i = 0; i < 4; i++) r.st[12 + i] = ctr_nonce[i];
    r.refill();
    memcpy(out, r.buf, 64);
}

int fhs_generate_keys(const fhs_params *params, const double *stds, size_t l, uint64_t seed,
                      fhs_client_key **out) {
    if (!out) return FHS_ERR_INVALID;
    *out = nullptr;
    fhs_client_key *k = new fhs_client_key();
    if (params) k->p = *params; else fhs_default_params(&k->p);
    k->lwe_std = stds ? stds[0] : 7.069849454709433e-6;
    k->glwe_std = stds ? stds[1] : 0.00000000000000029403601535432533;
    if (k->p.message_modulus < 2 || k->p.lwe_dim == 0 || k->p.poly_size == 0) { delete k; c_err = "bad parameters"; return FHS_ERR_INVALID; }
    // lib.rs:60-62: blocks per char = ceil(log_msg(128)), per usize = ceil(log_msg(l+1))
    k->nb_char = ceil_log(k->p.message_modulus, 128);
    k->nb_usize = std::max<uint32_t>(1, ceil_log(k->p.message_modulus, (uint64_t)l + 1));
    k->max_len = l;
    if (seed == 0) {
        if (!os_random(&k->secret_key, sizeof(ChaKey)) || !os_random(&k->public_key, sizeof(ChaKey))) {
            delete k; c_err = "getrandom failed"; return FHS_ERR_INVALID;
        }
    } else {   // reproducible test / bench mode: both keys are functions of the 64-bit seed -- NOT secure
        ChaKey boot{};
        boot.w[0] = (uint32_t)seed; boot.w[1] = (uint32_t)(seed >> 32); boot.w[2] = 0x46485321u;
        Rng d(boot, 0);
        for (auto &w : k->secret_key.w) w = (uint32_t)d.next();
        for (auto &w : k->public_key.w) w = (uint32_t)d.next();
        k->reproducible = true;
    }
    const uint32_t big = k->p.glwe_dim * k->p.poly_size;
    Rng r1(k->secret_key, 1), r2(k->secret_key, 2);
    k->glwe_sk.resize(big); for (auto &b : k->glwe_sk) b = (uint8_t)(r1.next() >> 63);
    k->lwe_sk.resize(k->p.lwe_dim); for (auto &b : k->lwe_sk) b = (uint8_t)(r2.next() >> 63);
    int nt = (int)std::thread::hardware_concurrency();
    if (nt < 1) nt = 1;
    if (nt > 32) nt = 32;
    gen_ksk(*k, nt);
    gen_bsk(*k, nt);
    *out = k;
    return FHS_OK;
}

void fhs_client_key_destroy(fhs_client_key *k) { delete k; }

int fhs_client_key_info(const fhs_client_key *k, fhs_params *p, uint32_t *nb_char, uint32_t *nb_usize, size_t *max_len) {
    if (!k) return FHS_ERR_INVALID;
    if (p) *p = k->p;
    if (nb_char) *nb_char = k->nb_char;
    if (nb_usize) *nb_usize = k->nb_usize;
    if (max_len) *max_len = (size_t)((k->nb_usize >= 32) ? ~0ull : ((1ull << (2 * k->nb_usize)) - 1));  // FheUsize::max, mod.rs:435
    return FHS_OK;
}

const uint64_t *fhs_client_ksk(const fhs_client_key *k, size_t *len) { if (len) *len = k->ksk.size(); return k->ksk.data(); }
const uint64_t *fhs_client_bsk(const fhs_client_key *k, size_t *len) { if (len) *len = k->bsk.size(); return k->bsk.data(); }

// Encrypt n plaintext block values (message + carry space, < 2*msg*carry incl. padding bit) under the big key.
int fhs_encrypt_blocks(fhs_client_key *k, const uint64_t *msgs, size_t n, uint64_t *out) {
    if (!k || (!msgs && n) || (!out && n)) return FHS_ERR_INVALID;
    const uint32_t big = k->p.glwe_dim * k->p.poly_size;
    const uint64_t delta = (1ull << 63) / ((uint64_t)k->p.message_modulus * k->p.carry_modulus);
    for (size_t b = 0; b < n; b++) {
        const uint64_t id = 0x30000000ull + k->enc_counter++;
        Rng r(k->public_key, id), e(k->secret_key, id);
        uint64_t *ct = out + b * (big + 1), dot = 0;
        for (uint32_t j = 0; j < big; j++) { ct[j] = r.next(); if (k->glwe_sk[j]) dot += ct[j]; }
        ct[big] = dot + msgs[b] * delta + e.torus_noise(k->glwe_std);
    }
    return FHS_OK;
}

// decrypt_message_and_carry: value mod 2*msg*carry (bit log2(msg*carry) is the padding bit)
int fhs_decrypt_blocks(const fhs_client_key *k, const uint64_t *cts, size_t n, uint64_t *msgs_out) {
    if (!k || (!cts && n) || (!msgs_out && n)) return FHS_ERR_INVALID;
    const uint32_t big = k->p.glwe_dim * k->p.poly_size;
    const uint64_t space = (uint64_t)k->p.message_modulus * k->p.carry_modulus;
    const uint64_t delta = (1ull << 63) / space;
    for (size_t b = 0; b < n; b++) {
        const uint64_t *ct = cts + b * (big + 1);
        uint64_t dot = 0;
        for (uint32_t j = 0; j < big; j++) if (k->glwe_sk[j]) dot += ct[j];
        const uint64_t phase = ct[big] - dot;
        msgs_out[b] = ((phase + delta / 2) / delta) % (2 * space);
    }
    return FHS_OK;
}

// FheString::new (mod.rs:58-92) validation; writes the plaintext block values of the string
// (chars + terminator + padding, nb_char digits each) and returns the number of chars.
int fhs_string_layout(const fhs_client_key *k, const char *s, int64_t pad, uint64_t *block_values, size_t cap_blocks, size_t *n_chars) {
    if (!k || !s || !n_chars) return FHS_ERR_INVALID;
    const size_t len = strlen(s);
    size_t maxl = 0;
    fhs_client_key_info(k, nullptr, nullptr, nullptr, &maxl);
    for (size_t i = 0; i < len; i++) if ((unsigned char)s[i] >= 128) { c_err = "string is not ascii"; return FHS_ERR_INVALID; }
    // a C string cannot contain the terminator; kept for parity: "string contains terminator character"
    if (len > maxl) { c_err = "string length exceeds maximum length"; return FHS_ERR_INVALID; }
    if (pad >= 0) {
        if ((size_t)pad > maxl) { c_err = "pad length exceeds maximum length"; return FHS_ERR_INVALID; }
        if ((size_t)pad < len) { c_err = "string length exceeds pad length"; return FHS_ERR_INVALID; }
    }
    const size_t chars = (pad >= 0 ? (size_t)pad : len) + 1;
    *n_chars = chars;
    if (!block_values) return FHS_OK;
    if (cap_blocks < chars * k->nb_char) { c_err = "buffer too small"; return FHS_ERR_INVALID; }
    const uint64_t mm = k->p.message_modulus;
    for (size_t i = 0; i < chars; i++) {
        uint64_t c = i < len ? (unsigned char)s[i] : 0;
        for (uint32_t b = 0; b < k->nb_char; b++) { block_values[i * k->nb_char + b] = c % mm; c /= mm; }
    }
    return FHS_OK;
}

}  // extern "C"
