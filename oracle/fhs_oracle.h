/*
 * fhs_oracle.h -- CPU ORACLE for the KS-PBS hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * PARITY UNPINNED at the tfhe-rs boundary: the arithmetic of this path lives in
 * the un-vendored crate `tfhe = "0.5.0"` (/root/reference/Cargo.toml:7, no
 * Cargo.lock), which is absent from this machine, and the reference's tests hold
 * no golden ciphertexts or known-answer vectors at the PBS boundary
 * (/root/reference/src/ciphertext/tests/ (all .rs files) only compare decrypted strings with
 * Rust std::str).  This file restates the published tfhe-rs 0.5 algorithms
 * (SURVEY.md Appendix A.1-A.8); end-to-end parity is pinned instead on the
 * reference's own test tables (decrypted string results == std::str).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library.  Nothing under fhe_string_b200/ does.
 */
#ifndef FHS_ORACLE_H
#define FHS_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* PARAM_MESSAGE_2_CARRY_2_KS_PBS (reference: src/lib.rs:31,45). */
typedef struct {
    uint32_t lwe_dim;         /* n = 742   */
    uint32_t glwe_dim;        /* k = 1     */
    uint32_t poly_size;       /* N = 2048  */
    uint32_t pbs_base_log;    /* 23        */
    uint32_t pbs_level;       /* 1         */
    uint32_t ks_base_log;     /* 3         */
    uint32_t ks_level;        /* 5         */
    uint32_t message_modulus; /* 4         */
    uint32_t carry_modulus;   /* 4         */
    uint32_t reserved;
    double lwe_std;           /* 7.069849454709433e-6 */
    double glwe_std;          /* 2.94036e-16          */
} orc_params;

typedef struct orc_keyset orc_keyset;

void orc_default_params(orc_params *p);

/* Key generation (client side; reference: src/lib.rs:50-77 -> tfhe gen_keys_radix). */
orc_keyset *orc_keygen(const orc_params *p, uint64_t seed);
void orc_keyset_free(orc_keyset *ks);
const orc_params *orc_keyset_params(const orc_keyset *ks);

/* Raw key material, in the layouts of SURVEY A.4 / A.6:
 *   ksk  u64 [k*N][ks_level][n+1]            level 1 (most significant) first
 *   bsk  u64 [n][pbs_level][k+1][k+1][N]     standard (coefficient) domain     */
const uint64_t *orc_ksk(const orc_keyset *ks);
const uint64_t *orc_bsk(const orc_keyset *ks);
const uint64_t *orc_lwe_sk(const orc_keyset *ks);  /* n   bits as u64 0/1 */
const uint64_t *orc_glwe_sk(const orc_keyset *ks); /* k*N bits as u64 0/1 */
size_t orc_ksk_len(const orc_keyset *ks);
size_t orc_bsk_len(const orc_keyset *ks);

/* Client-side block encryption under the big key.
 * `m` is the raw plaintext value placed at Delta = 2^63/(msg*carry): it may use
 * the padding bit (m in [0, 2*msg*carry)).  `rng_seed` makes noise reproducible. */
void orc_encrypt_block(const orc_keyset *ks, uint64_t m, uint64_t rng_seed, uint64_t *ct_out);
void orc_trivial_block(const orc_keyset *ks, uint64_t m, uint64_t *ct_out);
/* phase = b - <a,s> under the big key. */
uint64_t orc_phase_big(const orc_keyset *ks, const uint64_t *ct);
/* phase under the small key (after keyswitch). */
uint64_t orc_phase_small(const orc_keyset *ks, const uint64_t *ct);
/* decrypt_message_and_carry: ((phase + Delta/2) / Delta) mod (2*msg*carry) --
 * the padding bit is returned as bit log2(msg*carry). */
uint64_t orc_decode(const orc_keyset *ks, uint64_t phase);

/* --- the hot path, one function per SURVEY 8(a1) row ----------------------- */
/* a1.1 generate_lookup_table: 16-entry table -> trivial GLWE accumulator
 * ((k+1)*N u64, mask polynomials zero). */
void orc_build_lut(const orc_params *p, const uint64_t *table, uint64_t *acc_out);
/* a1.4 keyswitch_lwe_ciphertext: big LWE (k*N+1) -> small LWE (n+1). */
void orc_keyswitch(const orc_keyset *ks, const uint64_t *lwe_big, uint64_t *lwe_small);
/* a1.5 pbs_modulus_switch of one torus element to [0, 2N]. */
uint64_t orc_modswitch(const orc_params *p, uint64_t x);
/* a1.6 blind_rotate_assign on an accumulator (in place). */
void orc_blind_rotate(const orc_keyset *ks, const uint64_t *lwe_small, uint64_t *acc);
/* one step of a1.6: acc += GGSW_i (x) ct1 (decompose, forward FFT, Fourier BSK contraction, inverse
 * FFT, torus add); ct1 and acc are GLWE ciphertexts of (k+1)*N u64.  For first-principles tests. */
void orc_external_product(const orc_keyset *ks, uint32_t i, const uint64_t *ct1, uint64_t *acc);
/* a1.7 extract_lwe_sample_from_glwe_ciphertext at coefficient 0. */
void orc_sample_extract(const orc_params *p, const uint64_t *acc, uint64_t *lwe_big_out);
/* a1.3 keyswitch_programmable_bootstrap: full KS-PBS of one block. */
void orc_ks_pbs(const orc_keyset *ks, const uint64_t *lwe_in, const uint64_t *table,
                uint64_t *lwe_out);
/* Batched KS-PBS with OpenMP over ciphertexts (the "rayon" stand-in).
 * lut_id[b] selects a row of lut_tables[n_luts][msg*carry]. Returns threads used. */
int orc_ks_pbs_batch(const orc_keyset *ks, const uint64_t *lwe_in, const uint32_t *lut_id,
                     const uint64_t *lut_tables, size_t batch, uint64_t *lwe_out,
                     int num_threads);

/* Signed decomposition, exposed for unit tests.
 * Writes `level` digits, least-significant (level = `level`) first. */
void orc_decompose(uint64_t x, int base_log, int level, int64_t *digits_out);

int orc_max_threads(void);
/* 1 for the ORC_FAST build (oracle/Makefile target `fast`): same algorithm, -march=native, SoA radix-4
 * FFT without permutations, per-thread scratch -- the CPU arm bench.py times; NOT bit-identical to the
 * portable build the golden vectors pin. */
int orc_is_fast_build(void);

#ifdef __cplusplus
}
#endif
#endif
