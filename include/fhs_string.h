/*
 * fhs_string.h -- C ABI of the host-side level scheduler: fhe_string operations compiled to level
 * programs and executed on the GPU through include/fhs_cuda.h.
 *
 * Replaces the bodies of the reference's FheString methods (src/ciphertext/compare.rs:15-151,
 * convert.rs:14-61, search.rs:17-244, trim.rs:15-132, split.rs:238-515, replace.rs:17-104, insert.rs:14-124, mod.rs:132-237):
 * instead of issuing one tfhe-rs PBS at a time under rayon, an operation is built ONCE per operand
 * shape as a DAG of LUT evaluations over linear combinations of blocks; all nodes of equal depth
 * form one level = one fhs_level call.  Ciphertexts are plain tfhe shortint LWE containers
 * (u64[2049]); operands are passed as arrays of blocks in the reference's own layout:
 * FheString = chars in order, each FheAsciiChar = 4 blocks little-endian base 4; FheUsize =
 * num_blocks_usize blocks little-endian; BooleanBlock = 1 block.
 */
#ifndef FHS_STRING_H
#define FHS_STRING_H
#include <stddef.h>
#include <stdint.h>

#include "fhs_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct fhs_prog fhs_prog;
/* exchange callback of fhs_prog_run_sharded / fhs_prog_run_ex (see there) */
typedef int (*fhs_allgather_fn)(void *user, uint64_t *device_base, size_t bytes_per_rank);

/* Build the program of operation `op` (the reference method name: "eq", "ne", "lt", "le", "gt",
 * "ge", "eq_ignore_ascii_case", "is_empty", "len", "to_uppercase", "to_lowercase", "contains",
 * "find", "rfind", "starts_with", "ends_with", "trim", "trim_start", "trim_end", "strip_prefix",
 * "strip_suffix", "substr_to", "substr_from", "split*", "replace", "replacen", "add", "insert",
 * "repeat", ...) for operand shapes:
 *   n_chars[i] = number of FheAsciiChar of string operand i (terminator and padding included),
 *   n_usize_args = number of FheUsize operands that follow the strings,
 *   scalars = clear arguments of the method (e.g. l of replace, clear n), may be NULL.
 * Returns FHS_ERR_INVALID for unknown ops or shapes the reference would panic on. */
int fhs_prog_build(const char *op, const uint32_t *n_chars, size_t n_strings, size_t n_usize_args,
                   const uint64_t *scalars, size_t n_scalars, uint32_t num_blocks_char,
                   uint32_t num_blocks_usize, fhs_prog **out);
void fhs_prog_destroy(fhs_prog *p);
/* Message of the last FHS_ERR_* returned by fhs_prog_build on this thread. */
const char *fhs_host_last_error(void);

/* Shape of the program. info[0..7] = nodes, input blocks, output blocks, terms, luts, depth (levels),
 * LUT evaluations (PBS count), max variance multiplier fed to a PBS (rounded up). */
int fhs_prog_info(const fhs_prog *p, uint64_t info[8]);
/* Output layout, e.g. "bool", "usize", "str:65", "opt_usize", "opt_str:65", "slices:67:str:65". */
const char *fhs_prog_output_desc(const fhs_prog *p);
/* Export for inspection / clear-text evaluation by test code.  Arrays sized from fhs_prog_info:
 * kind (0 input, 1 pbs, 2 linear), lut, t0, tn, cst (message units), level per node; term_src /
 * term_coef per term; luts[n_luts][16]; outputs[n_outputs] node ids (-1-c encodes the clear
 * constant c for outputs that do not depend on any ciphertext). */
int fhs_prog_export(const fhs_prog *p, uint8_t *kind, int32_t *lut, uint32_t *t0, uint32_t *tn,
                    int64_t *cst, int32_t *level, int32_t *term_src, int32_t *term_coef,
                    uint64_t *luts, int64_t *outputs);

/* Largest variance multiplier (in units of one PBS output) any output block of the program carries.  The next
 * operation treats its operands as fresh (1), so the builders keep this small (outputs above 4 are refreshed). */
double fhs_prog_output_noise(const fhs_prog *p);

/* Per node: bounds of the linear combination feeding it (in_lo, in_hi), of its output (out_lo, out_hi),
 * and whether a PBS input may be negative (negacyclic use of the padding bit).  For evaluators that
 * assert the builder's range claims. */
int fhs_prog_export_ranges(const fhs_prog *p, int32_t *in_lo, int32_t *in_hi, int32_t *out_lo, int32_t *out_hi,
                           uint8_t *signed_in);

/* Schedule the executor uses on `world` GPUs (shard_min 0 = 74): level[node] after the level-balancing pass
 * (a level costs what its blind-rotation WAVES cost -- 74 / 148 / 296 / 370 / 592 ciphertexts on a B200 -- so
 * evaluations with slack move to the next level where that saves a wave; depth and results are unchanged), and
 * model_ms = {as-soon-as-possible, balanced} modelled device time.  FHS_LEVEL_BALANCE=0 switches the pass off.
 * Replaces nothing in the reference (rayon runs each op's PBS as they come); for tests and tools. */
int fhs_prog_export_schedule(const fhs_prog *p, int world, size_t shard_min, int32_t *level, double model_ms[2]);

/* Execute on the GPU: in_blocks [n_inputs][k*N+1] host -> out_blocks [n_outputs][k*N+1] host.
 * Ciphertexts stay in a device pool between levels.  stats[0..3] (may be NULL) = levels run,
 * LUT evaluations, kernels launched, device milliseconds (events around the whole program). */
int fhs_prog_run(const fhs_prog *p, fhs_ctx *ctx, const uint64_t *in_blocks, uint64_t *out_blocks,
                 double stats[4]);

/* General form: operands given piecewise, each piece a HOST array or a DEVICE-resident array
 * (fhs_dev_alloc, fhs_cuda.h) of n_blocks blocks, concatenated in order to the program's inputs; the
 * result goes to out_host and / or out_dev (either may be NULL).  Results left on the device feed the
 * next operation without crossing PCIe (to_uppercase -> contains).  Page-locked host memory
 * (fhs_host_alloc) is transferred by direct DMA, pageable memory through a page-locked staging buffer.
 * The planned pool with its resident level descriptors is cached on the context per (program, rank,
 * world, shard_min): a repeated call is transfers + kernel launches only.  Synchronous: operands and
 * results are settled when the call returns.
 * stats[0..7] = levels, LUT evaluations, kernels launched, wall ms of the call, DEVICE ms (events on the
 * context's stream around operand copies + levels, -1 if unavailable), host planning ms (0.0x when
 * cached), levels that were split by rank, 1 if the execution state came from the cache. */
typedef struct { const uint64_t *host; const uint64_t *dev; size_t n_blocks; } fhs_operand;
int fhs_prog_run_ex(const fhs_prog *p, fhs_ctx *ctx, const fhs_operand *ops, size_t n_ops, uint64_t *out_host,
                    uint64_t *out_dev, double stats[8], int rank, int world, size_t shard_min,
                    fhs_allgather_fn allgather, void *user);

/* M independent instances of the same program in one pass (e.g. the 1024 string comparisons of a
 * throughput sweep): in_blocks [M][n_inputs][k*N+1], out_blocks [M][n_outputs][k*N+1].  Every level
 * is evaluated for all instances at once (chunks of <= 32768 LUT evaluations per fhs_level call), so
 * levels that are small per instance still fill the GPU.  Independent instances need no exchange:
 * multi-GPU callers give each rank its own slice of the instances. */
int fhs_prog_run_many(const fhs_prog *p, fhs_ctx *ctx, size_t n_instances, const uint64_t *in_blocks,
                      uint64_t *out_blocks, double stats[4]);

/* Multi-GPU (one process per GPU, every rank holds a key replica and runs the same program on the
 * same inputs): levels with more than shard_min evaluations (0 = default 74, one wave of the pair blind-rotation kernel) are split
 * contiguously by rank; after such a level `allgather(user, device_base, bytes_per_rank)` must make
 * the world * bytes_per_rank bytes at device_base identical on all ranks (rank q contributes chunk q
 * in place: ncclAllGather / torch.distributed.all_gather_into_tensor).  Smaller levels are
 * replicated.  fhs_prog_plan reports, per level, {B, blocks per rank (0 = replicated), lo, hi}.
 * allgather == NULL selects the library's own exchange: ncclAllGather in place on the context's
 * stream through the communicator set up with fhs_comm_init (fhs_cuda.h) -- no host
 * synchronisation between levels, no host framework in the loop; FHS_ERR_COMM if the context has no
 * communicator of this rank / world. */
int fhs_prog_run_sharded(const fhs_prog *p, fhs_ctx *ctx, const uint64_t *in_blocks, uint64_t *out_blocks,
                         double stats[4], int rank, int world, size_t shard_min,
                         fhs_allgather_fn allgather, void *user);
int fhs_prog_plan(const fhs_prog *p, int rank, int world, size_t shard_min, uint64_t *levels_out,
                  size_t cap_levels, size_t *n_levels);

/* ---- client side (CPU): generate_keys / ClientKey of the reference ---------------------------
 * src/lib.rs:44-77 generate_keys_with_params(p, l); src/client_key.rs:16-62; mod.rs:58-92,121-129.
 * Not accelerated (one-off CPU work in the reference too).  seed == 0 draws from the OS. */
typedef struct fhs_client_key fhs_client_key;
int fhs_generate_keys(const fhs_params *params /* NULL: PARAM_MESSAGE_2_CARRY_2_KS_PBS */,
                      const double *stds /* {lwe_std, glwe_std} or NULL */, size_t l, uint64_t seed,
                      fhs_client_key **out);
void fhs_client_key_destroy(fhs_client_key *k);
const char *fhs_client_last_error(void);
/* Self-test hook for the client's generator (ChaCha20 block function, RFC 8439 2.3.2): state words
 * 12..15 = ctr_nonce.  Keys, masks and noise come from this generator keyed from the OS (getrandom,
 * 2 x 256 bit: one key for secrets and noise, one for public masks) when seed == 0; a non-zero seed is
 * the reproducible test / bench mode and is NOT secure (64-bit seed). */
void fhs_chacha20_block(const uint32_t key[8], const uint32_t ctr_nonce[4], uint8_t out[64]);
int fhs_client_key_info(const fhs_client_key *k, fhs_params *p, uint32_t *num_blocks_char,
                        uint32_t *num_blocks_usize, size_t *max_len);
/* Server key material in the layouts fhs_upload_keys takes (KSK level order MSB first). */
const uint64_t *fhs_client_ksk(const fhs_client_key *k, size_t *len);
const uint64_t *fhs_client_bsk(const fhs_client_key *k, size_t *len);
int fhs_encrypt_blocks(fhs_client_key *k, const uint64_t *msgs, size_t n, uint64_t *out);
int fhs_decrypt_blocks(const fhs_client_key *k, const uint64_t *cts, size_t n, uint64_t *msgs_out);
/* FheString::new validation + layout: same five error messages as mod.rs:59-71. */
int fhs_string_layout(const fhs_client_key *k, const char *s, int64_t pad, uint64_t *block_values,
                      size_t cap_blocks, size_t *n_chars);

#ifdef __cplusplus
}
#endif
#endif
