/*
 * fhs_cuda.h -- C ABI of the B200-native batched KS-PBS engine under fhe_string.
 *
 * The reference has no FFI for this path: fhe_string calls tfhe::integer::ServerKey methods
 * (`k.k.<op>`, /root/reference/src/server_key.rs:12-33 and every call site listed in SURVEY.md
 * Appendix C, e.g. src/ciphertext/compare.rs:32, src/ciphertext/search.rs:33-41), and each of those
 * ends in tfhe-rs `keyswitch_programmable_bootstrap_assign` on one block at a time.  This header is
 * the boundary a `fhe_string_cuda` crate would bind instead (INTEGRATION.md shows the Rust
 * `extern "C"` block): one call evaluates a whole LEVEL of independent LUT evaluations.
 *
 * Conventions: every function returns 0 on success and a negative fhs_status on failure; the
 * message is available from fhs_last_error().  Nothing unwinds across this boundary.  The caller
 * owns all host buffers; the library owns all device memory.  A context is bound to one GPU and
 * is NOT thread-safe; use one context per GPU (one process per GPU in the multi-GPU layout).
 * There is no CPU fallback: without a CUDA device fhs_ctx_create fails with FHS_ERR_CUDA.
 */
#ifndef FHS_CUDA_H
#define FHS_CUDA_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    FHS_OK = 0,
    FHS_ERR_INVALID = -1,     /* bad argument / unsupported parameter set */
    FHS_ERR_CUDA = -2,        /* CUDA runtime error, see fhs_last_error */
    FHS_ERR_NO_KEYS = -3,     /* keys not uploaded yet */
    FHS_ERR_OOM = -4,
    FHS_ERR_COMM = -5         /* NCCL error */
} fhs_status;

/* tfhe-rs ClassicPBSParameters as used by the reference (src/lib.rs:31,45,50:
 * PARAM_MESSAGE_2_CARRY_2_KS_PBS): n=742, k=1, N=2048, pbs 23/1, ks 3/5, msg 4, carry 4. */
typedef struct {
    uint32_t lwe_dim;
    uint32_t glwe_dim;
    uint32_t poly_size;
    uint32_t pbs_base_log;
    uint32_t pbs_level;
    uint32_t ks_base_log;
    uint32_t ks_level;
    uint32_t message_modulus;
    uint32_t carry_modulus;
    uint32_t reserved;
} fhs_params;

typedef struct fhs_ctx fhs_ctx;

/* KSK row order inside one input-coefficient block (version-sensitive in tfhe-rs). */
#define FHS_KSK_LEVEL_MSB_FIRST 0 /* rows stored level 1 (most significant) .. level l  (tfhe-rs 0.5) */
#define FHS_KSK_LEVEL_LSB_FIRST 1 /* rows stored level l .. level 1 */

/* Frequency convention of a caller-supplied Fourier BSK (fhs_upload_keys_fourier).  With
 * z_j = p_j + i p_{j+N/2} (coefficients as signed 64-bit integers) and the twist w^j, w = e^{i pi/N}:
 *   POS: X[m] = sum_j z_j w^j e^{+2 pi i j m/(N/2)}  = P(e^{i pi (4m+1)/N})
 *   NEG: X[m] = sum_j z_j w^j e^{-2 pi i j m/(N/2)}  = P(e^{i pi (1-4m)/N})   (twist, then a forward DFT
 *        with the usual negative exponent: what tfhe-rs' canonical/serialized order is recalled to be)
 * Both are un-normalised and in natural order m = 0..N/2-1. */
#define FHS_FOURIER_NATURAL_POS 0
#define FHS_FOURIER_NATURAL_NEG 1

void fhs_default_params(fhs_params *p);
const char *fhs_last_error(void);
int fhs_version(void);

/* Context on CUDA device `device` (ordinal).  Replaces: construction of tfhe ServerKey state
 * (src/server_key.rs:12-17). */
int fhs_ctx_create(int device, fhs_ctx **out);
void fhs_ctx_destroy(fhs_ctx *ctx);
/* Run all work of this context on an existing CUDA stream (cudaStream_t as void*); NULL = the
 * context's own stream.  Lets a host framework time launches with its own events. */
int fhs_ctx_set_stream(fhs_ctx *ctx, void *cuda_stream);
int fhs_ctx_synchronize(fhs_ctx *ctx);
/* Parameters of the uploaded keys (FHS_ERR_NO_KEYS before fhs_upload_keys). */
int fhs_ctx_get_params(fhs_ctx *ctx, fhs_params *out);

/* Upload server key material once per GPU (north_star item 3).  Replaces the key held by
 * tfhe::integer::ServerKey (src/lib.rs:64-72).
 *   ksk : u64 [k*N][ks_level][n+1]                 (level order per `ksk_level_order`)
 *   bsk : u64 [n][pbs_level][k+1][k+1][N]          standard (coefficient) domain GGSW rows
 * The BSK is transformed to the Fourier domain on the GPU in the layout the blind-rotation kernel
 * consumes; the KSK is re-tiled to digit order. */
int fhs_upload_keys(fhs_ctx *ctx, const fhs_params *p, const uint64_t *ksk, int ksk_level_order,
                    const uint64_t *bsk);
/* Same, for a caller that only holds the Fourier-domain key: tfhe::integer::ServerKey keeps a
 * FourierLweBootstrapKey and drops the standard one at key generation (src/lib.rs:64-72 returns only
 * that ServerKey), so a reference-made server key can be uploaded without the client key.
 *   bsk_fourier : f64 pairs (re, im) [n][pbs_level][k+1][k+1][N/2], natural frequency order,
 *                 un-normalised (integer units of the torus), `fourier_convention` = FHS_FOURIER_*.
 * The spectra are used as they are (scaled by a power of two and re-tiled); PBS output noise is then
 * that of the caller's f64 transform instead of this library's double-double one. */
int fhs_upload_keys_fourier(fhs_ctx *ctx, const fhs_params *p, const uint64_t *ksk, int ksk_level_order,
                            const double *bsk_fourier, int fourier_convention);

/* The two calls above behind ONE entry point with a format flag, as SURVEY.md 8(b) words the boundary:
 * bsk_format FHS_BSK_STANDARD = u64 [n][lvl][k+1][k+1][N] coefficient-domain rows, FHS_BSK_FOURIER_POS /
 * _NEG = c64 natural-order spectra with that exponent sign. */
#define FHS_BSK_STANDARD (-1)
#define FHS_BSK_FOURIER_POS FHS_FOURIER_NATURAL_POS
#define FHS_BSK_FOURIER_NEG FHS_FOURIER_NATURAL_NEG
int fhs_upload_keys_fmt(fhs_ctx *ctx, const fhs_params *p, const uint64_t *ksk, int ksk_level_order,
                        const void *bsk, int bsk_format);

/* One level of independent LUT evaluations, host buffers (copies inside the call).
 * Replaces B calls of shortint apply_lookup_table (tfhe keyswitch_programmable_bootstrap_assign),
 * reached from e.g. src/ciphertext/compare.rs:32,74,130-132, src/ciphertext/logic.rs:13-65.
 *   lwe_in  : u64 [B][k*N+1]   big-key LWE blocks (mask || body)
 *   lut_id  : u32 [B]          row of lut_tables for block b
 *   lut_tables : u64 [n_luts][message_modulus*carry_modulus]  f(x) values (not scaled by Delta)
 *   lwe_out : u64 [B][k*N+1] */
int fhs_ks_pbs_batch(fhs_ctx *ctx, const uint64_t *lwe_in, const uint32_t *lut_id,
                     const uint64_t *lut_tables, size_t n_luts, size_t batch, uint64_t *lwe_out);
/* Same, all pointers are DEVICE pointers on the context's GPU; asynchronous on its stream. */
int fhs_ks_pbs_batch_dev(fhs_ctx *ctx, const uint64_t *d_lwe_in, const uint32_t *d_lut_id,
                         const uint64_t *d_lut_tables, size_t n_luts, size_t batch,
                         uint64_t *d_lwe_out);
/* Same as fhs_ks_pbs_batch_dev for callers that may not allocate (a stream under CUDA graph capture, a
 * tight serving loop, a harness that forbids cudaMalloc).  No counterpart in the reference.
 * The path's scratch (keyswitch digit tiles + the keyswitched small LWE
 * ciphertexts) is a pair of file-scope device arrays in the library, 6.3 MB per GPU held until process exit,
 * instead of the context's grow-on-demand buffers; the call makes no cudaMalloc / cudaFree and calls no
 * allocating routine.  batch > FHS_NOALLOC_MAX_BATCH -> FHS_ERR_INVALID (fhs_last_error names the limit);
 * parameter sets other than k N = 2048, n + 1 <= 768, KS 2^3 x 5 likewise.  Results are bit-identical to
 * fhs_ks_pbs_batch_dev.  All calls on one GPU share the arrays: each call overwrites what it reads, so
 * calls in sequence on one stream are independent, but calls that overlap on the GPU (two streams, two
 * contexts on the same device) race.  Asynchronous on the context's stream, like fhs_ks_pbs_batch_dev. */
#define FHS_NOALLOC_MAX_BATCH 384
int fhs_ks_pbs_batch_dev_noalloc(fhs_ctx *ctx, const uint64_t *d_lwe_in, const uint32_t *d_lut_id,
                                 const uint64_t *d_lut_tables, size_t n_luts, size_t batch,
                                 uint64_t *d_lwe_out);

/* Verification entry points for the bit-exact tiers. */
/* tfhe keyswitch_lwe_ciphertext on a batch: [B][k*N+1] -> [B][n+1]; host buffers. */
int fhs_keyswitch_batch(fhs_ctx *ctx, const uint64_t *lwe_in, size_t batch, uint64_t *lwe_small_out);
int fhs_keyswitch_batch_dev(fhs_ctx *ctx, const uint64_t *d_lwe_in, size_t batch,
                            uint64_t *d_lwe_small_out);
/* tfhe generate_lookup_table: 16-entry table -> trivial GLWE accumulator [(k+1)][N] (mask 0),
 * produced by the same device code the blind rotation uses (closed form, rotation 0). */
int fhs_build_lut(fhs_ctx *ctx, const uint64_t *table, uint64_t *acc_out);
/* Blind rotation + sample extract only (no keyswitch): small LWE [B][n+1] -> big LWE [B][k*N+1]. */
int fhs_pbs_batch(fhs_ctx *ctx, const uint64_t *lwe_small_in, const uint32_t *lut_id,
                  const uint64_t *lut_tables, size_t n_luts, size_t batch, uint64_t *lwe_out);
int fhs_pbs_batch_dev(fhs_ctx *ctx, const uint64_t *d_lwe_small_in, const uint32_t *d_lut_id,
                      const uint64_t *d_lut_tables, size_t n_luts, size_t batch,
                      uint64_t *d_lwe_out);

/* Device-time of the most recent *_batch call's kernels in milliseconds (CUDA events on the
 * context's stream): [0]=keyswitch, [1]=blind rotation (+LUT build, sample extract).  Both are 0
 * after a host-buffer call that was pipelined (fhs_ks_pbs_batch with >= 12 ciphertexts per SM runs
 * in three chunks whose transfers overlap the other chunks' kernels). */
int fhs_last_kernel_ms(fhs_ctx *ctx, float *ms2);
/* Number of kernels this context has launched so far. */
uint64_t fhs_launch_count(fhs_ctx *ctx);
/* Measured FP64 FMA throughput of the context's GPU in TFLOP/s (dependent-free DFMA chains, best
 * of 4 after a warm-up): the roofline denominator of the blind-rotation kernel. */
int fhs_measure_fp64_tflops(fhs_ctx *ctx, double *tflops_out);

/* ---- device-resident ciphertext pool + level execution ------------------------
 * A pool is a device array of big-key LWE blocks [capacity][k*N+1].  fhs_level evaluates one level
 * of a radix/string operation: for b < B
 *     in_b = sum_{t in [term_off[b], term_off[b+1])} term_coef[t] * pool[term_idx[t]]  + (0,..,0,const_body[b])
 *     pool[out_slot + b] = lut_id ? KS-PBS(in_b, lut_tables[lut_id[b]]) : in_b
 * (const_body is already scaled by Delta).  Replaces the linear shortint ops + apply_lookup_table
 * sequences tfhe-rs runs per block (unchecked_scalar_mul/add/sub + keyswitch_programmable_bootstrap).
 * Descriptor arrays are host pointers; ciphertexts never leave the GPU.  Asynchronous on the
 * context's stream; fhs_pool_read synchronises. */
typedef struct fhs_pool fhs_pool;
int fhs_pool_create(fhs_ctx *ctx, size_t capacity_blocks, fhs_pool **out);
void fhs_pool_destroy(fhs_pool *pool);
size_t fhs_pool_capacity(fhs_pool *pool);
uint64_t *fhs_pool_device_ptr(fhs_pool *pool);
int fhs_pool_write(fhs_pool *pool, size_t slot, size_t n, const uint64_t *host_blocks);
int fhs_pool_read(fhs_pool *pool, size_t slot, size_t n, uint64_t *host_blocks);
int fhs_level(fhs_pool *pool, size_t B, const uint32_t *term_off, const uint32_t *term_idx,
              const int32_t *term_coef, const uint64_t *const_body, const uint32_t *lut_id,
              const uint64_t *lut_tables, size_t n_luts, size_t out_slot);

/* fhs_level with its arguments as one descriptor (the form SURVEY.md 8(b) gives):
 *   B evaluations; gather idx / coef as CSR (term_off[B+1]); const_body[B] scaled by Delta; lut_id[B] or
 *   NULL for a purely linear level; results in pool slots [out_slot, out_slot + B). */
typedef struct {
    size_t B;
    const uint32_t *term_off, *term_idx;
    const int32_t *term_coef;
    const uint64_t *const_body;
    const uint32_t *lut_id;
    const uint64_t *lut_tables;
    size_t n_luts;
    size_t out_slot;
} fhs_level_desc;
int fhs_level_d(fhs_pool *pool, const fhs_level_desc *level);

/* Side-stream upload into a pool (overlaps with levels already queued on the context's stream) and
 * the matching wait of the context's stream.  Used by fhs_prog_run_many to hide the H2D of the next
 * chunk of instances behind the levels of the current one. */
int fhs_pool_write_async(fhs_pool *pool, size_t slot, size_t n, const uint64_t *host_blocks);
int fhs_pool_wait_writes(fhs_pool *pool);

/* Resident level program: upload the descriptors of EVERY level of a program once (rows in
 * execution order; row_off[rows+1] are absolute offsets into term_idx / term_coef; lut_id is ignored
 * for rows that are run with is_pbs = 0), then run levels with kernel launches only.  This is what
 * the level scheduler (include/fhs_string.h) uses: a 71-level split costs 71 x 3 launches and no
 * host<->device traffic between its first input block and its last output block. */
int fhs_pool_load_program(fhs_pool *pool, size_t rows, size_t terms, const uint32_t *row_off,
                          const uint32_t *term_idx, const int32_t *term_coef, const uint64_t *const_body,
                          const uint32_t *lut_id, const uint64_t *lut_tables, size_t n_luts,
                          size_t max_level_rows);
int fhs_level_run(fhs_pool *pool, size_t row_lo, size_t B, int is_pbs, size_t out_slot);

/* ---- resident operands and cached executions (SURVEY.md 8b "ciphertext-pool handles") ---------
 * Un-synchronised pool transfers on the context's stream (the caller keeps the host buffer alive and
 * unchanged until its next fhs_ctx_synchronize) and device-to-device copies between a pool and a
 * resident ciphertext array. */
int fhs_pool_enqueue_write(fhs_pool *pool, size_t slot, size_t n, const uint64_t *host_blocks);
int fhs_pool_enqueue_read(fhs_pool *pool, size_t slot, size_t n, uint64_t *host_blocks);
int fhs_pool_copy_in(fhs_pool *pool, size_t slot, size_t n, const uint64_t *d_src);
int fhs_pool_copy_out(fhs_pool *pool, size_t slot, size_t n, uint64_t *d_dst);
/* Device-resident ciphertext arrays [n_blocks][k*N+1] (stream-ordered allocation on the context's
 * stream): the result of one operation is the operand of the next without crossing PCIe.  Replaces the
 * host-side Vec<Ciphertext> the reference passes between its ops (src/ciphertext/mod.rs:43-47). */
int fhs_dev_alloc(fhs_ctx *ctx, size_t n_blocks, uint64_t **d_out);
void fhs_dev_free(fhs_ctx *ctx, uint64_t *d);
int fhs_dev_write(fhs_ctx *ctx, uint64_t *d, size_t n_blocks, const uint64_t *host_blocks);   /* synchronous */
int fhs_dev_read(fhs_ctx *ctx, const uint64_t *d, size_t n_blocks, uint64_t *host_blocks);    /* synchronous */
/* Page-locked host memory for operand / result staging (direct DMA instead of a staged copy). */
int fhs_host_alloc(fhs_ctx *ctx, size_t bytes, void **out);
void fhs_host_free(fhs_ctx *ctx, void *p);
int fhs_host_is_pinned(const void *p);
/* Attachment cache of a context: a caller (the level scheduler, fhs_string.h) parks per-program state
 * that lives on this GPU -- the planned pool with its resident level descriptors -- under a 4-word key
 * (key[0] = owner id).  Entries are destroyed (dtor) with the context, when their owner goes away
 * (fhs_cache_drop_owner, every live context) or least-recently-used first when the sum of device_bytes
 * exceeds the budget (default 16 GiB, FHS_CACHE_BYTES / fhs_ctx_cache_limit). */
typedef void (*fhs_cache_dtor)(void *obj);
int fhs_ctx_cache_get(fhs_ctx *ctx, const uint64_t key[4], void **obj);   /* *obj = NULL when absent */
int fhs_ctx_cache_put(fhs_ctx *ctx, const uint64_t key[4], void *obj, fhs_cache_dtor dtor, size_t device_bytes);
void fhs_cache_drop_owner(uint64_t owner);
int fhs_ctx_cache_stats(fhs_ctx *ctx, size_t *entries, size_t *device_bytes);
int fhs_ctx_cache_limit(fhs_ctx *ctx, size_t device_bytes);
/* Device timer: CUDA events on the context's stream around whatever is enqueued between start and stop;
 * fhs_ctx_timer_ms waits for the stop event. */
int fhs_ctx_timer_start(fhs_ctx *ctx);
int fhs_ctx_timer_stop(fhs_ctx *ctx);
int fhs_ctx_timer_ms(fhs_ctx *ctx, float *ms);

/* ---- multi-GPU: one process per GPU, a key replica per GPU ---------------------
 * The path shards over independent LUT evaluations; the only traffic is the server keys once
 * (broadcast) and the result blocks of a level that was split by rank (all-gather).  Both are NCCL
 * collectives over NVLink enqueued on the context's stream.  NCCL is resolved with dlopen at the
 * first call (the copy already loaded in the process, e.g. torch's, else the system libnccl.so.2;
 * FHS_NCCL_LIB overrides); single-GPU callers never need it.
 * Bootstrap: rank 0 calls fhs_comm_unique_id and ships the 128 bytes to the other ranks by any
 * means (file, socket, MPI, torch.distributed); every rank then calls fhs_comm_init. */
#define FHS_COMM_ID_BYTES 128
int fhs_comm_unique_id(uint8_t id[FHS_COMM_ID_BYTES]);
int fhs_comm_init(fhs_ctx *ctx, const uint8_t id[FHS_COMM_ID_BYTES], int rank, int world);
int fhs_comm_info(fhs_ctx *ctx, int *rank, int *world);   /* FHS_ERR_COMM when there is none */
void fhs_comm_destroy(fhs_ctx *ctx);
/* In-place all-gather on device memory, asynchronous on the context's stream: rank q has written
 * [d_base + q*bytes_per_rank, +bytes_per_rank); afterwards all world*bytes_per_rank bytes are equal
 * on every rank.  Used by fhs_prog_run_sharded after every level that was split by rank. */
int fhs_comm_allgather(fhs_ctx *ctx, uint64_t *d_base, size_t bytes_per_rank);
/* Peer windows: the library's own collective for the levels a sharded program splits by rank.
 * fhs_pool_create_shared is fhs_pool_create for a pool that every other rank of the context's communicator maps
 * into its address space (CUDA IPC over NVLink / NVSwitch peer access; the handles travel through the NCCL
 * communicator once).  COLLECTIVE: every rank calls it at the same point with the same capacity.  Without a
 * communicator (or with FHS_COMM_P2P=0, or when any rank cannot export / map the memory) the result is an
 * ordinary pool: ask fhs_pool_is_shared.
 * fhs_pool_peer_allgather(pool, slot_base, per_rank, my_blocks): rank q has computed slots
 * [slot_base + q*per_rank, + n_q) of its own pool (this rank's n is my_blocks <= per_rank); ONE kernel stores them
 * into the same slots of every peer's pool, publishes a sequence number there and waits for the peers' own, so
 * that kernels queued behind it see the blocks of all ranks -- the all-gather of fhs_comm_allgather without
 * NCCL's launch and protocol latency.  Asynchronous on the context's stream; every rank calls it for the same
 * levels in the same order.  fhs_pool_peer_run_done closes a run (the peers may then overwrite this run's blocks);
 * fhs_pool_peer_status, after a stream synchronisation, is FHS_ERR_COMM when a wait for another rank timed out
 * (FHS_PEER_TIMEOUT_MS, default 30 s: a dead peer costs an error, not a hung GPU).  Reference: none -- the
 * reference is single-process; this replaces rayon's fan-out of a level (src/ciphertext/search.rs:33-41). */
int fhs_pool_create_shared(fhs_ctx *ctx, size_t capacity_blocks, fhs_pool **out);
int fhs_pool_is_shared(fhs_pool *pool);
int fhs_pool_peer_allgather(fhs_pool *pool, size_t slot_base, size_t per_rank, size_t my_blocks);
int fhs_pool_peer_run_done(fhs_pool *pool);
int fhs_pool_peer_status(fhs_pool *pool);
unsigned long long fhs_comm_peer_pushes(fhs_ctx *ctx);   /* peer-window all-gathers launched on this context */
/* Broadcast of a HOST buffer from `root` (staged through device memory, synchronous): how the
 * server keys made on rank 0 reach the other ranks before each calls fhs_upload_keys. */
int fhs_comm_broadcast_host(fhs_ctx *ctx, void *host_buf, size_t bytes, int root);

#ifdef __cplusplus
}
#endif
#endif /* FHS_CUDA_H */
