// Device-resident ciphertext pool and level execution (SURVEY.md 8b "device-resident variant"):
//   K1 gather + linear-combination prologue:  in[b] = sum_t coef[b][t] * pool[idx[b][t]] + const[b]
//   (bivariate pack 4x+y, compare x-y, block pack 4hi+lo, <=5-term sums, scalar add; replaces the
//   linear shortint ops tfhe-rs runs before each PBS: unchecked_scalar_mul / unchecked_add / sub),
//   then keyswitch + blind rotation of the whole level, results written to consecutive pool slots.
#include <cstring>
#include <vector>

#include "fhs_internal.cuh"

namespace fhs {

// Control words of a peer window; every word that another GPU writes sits in its own 128-byte line.
constexpr int kMaxPeers = 8;
struct PeerCtrl {
    uint32_t arrive[kMaxPeers][32];   // [q][0]: last step whose blocks from rank q have all landed here
    uint32_t done[32];                // [0]: runs THIS rank has completed (it no longer reads that run's blocks);
                                      // the peers poll it over NVLink before the first stores of their next run, so
                                      // nobody writes into a rank's memory after that rank's last run has ended
    uint32_t cta_count[32];           // local: CTAs of the current push kernel that have finished their stores
    uint32_t error[32];               // local: a wait timed out
};
struct PeerPtrs {
    uint64_t *pool[kMaxPeers];
    PeerCtrl *ctrl[kMaxPeers];
};

struct Pool {
    Ctx *ctx;
    uint64_t *d = nullptr;
    size_t capacity = 0;
    // level descriptor staging (device)
    uint32_t *d_off = nullptr, *d_idx = nullptr, *d_lut = nullptr;
    int32_t *d_coef = nullptr;
    uint64_t *d_cst = nullptr, *d_tables = nullptr;
    size_t cap_b = 0, cap_t = 0, cap_luts = 0;
    uint64_t *d_stage = nullptr;  // [cap_b][big+1] linear-combination results
    // resident program (fhs_pool_load_program): descriptors of every level, uploaded once
    uint32_t *r_off = nullptr, *r_idx = nullptr, *r_lut = nullptr;
    int32_t *r_coef = nullptr;
    uint64_t *r_cst = nullptr, *r_tables = nullptr, *r_stage = nullptr;
    size_t r_rows = 0, r_luts = 0, r_max_rows = 0;
    // side-stream uploads (fhs_pool_write_async): overlap the next batch's H2D with this batch's levels
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t write_done = nullptr;
    bool write_pending = false;
    // peer window (fhs_pool_create_shared): the pool is a cudaMalloc block followed by a control area, mapped
    // into every other rank's address space; result blocks of a level that was split by rank are pushed
    // into the peers' pools by this library's own kernel instead of an NCCL all-gather
    bool shared = false;            // mapped on every rank
    bool own_malloc = false;        // p->d came from cudaMalloc (IPC needs that), not from the stream-ordered pool
    int rank = 0, world = 1;
    uint64_t *peer_pool[kMaxPeers] = {};
    PeerCtrl *peer_ctrl[kMaxPeers] = {};
    uint32_t step = 0, runs = 0;    // sharded levels pushed so far / runs completed (same on every rank)
    bool first_of_run = true;
    uint32_t *h_err = nullptr;      // page-locked copy of the control area's error word (valid after a stream sync)
    unsigned long long pushes = 0;
};

namespace {

__global__ void __launch_bounds__(256)
gather_linear_kernel(const uint64_t *__restrict__ pool, const uint32_t *__restrict__ off,
                     const uint32_t *__restrict__ idx, const int32_t *__restrict__ coef,
                     const uint64_t *__restrict__ cst, uint64_t *__restrict__ out, int width) {
    const unsigned b = blockIdx.x;
    const uint32_t t0 = off[b], t1 = off[b + 1];
    uint64_t *__restrict__ o = out + (size_t)b * width;
    for (int j = threadIdx.x; j < width; j += blockDim.x) {
        uint64_t acc = (j == width - 1) ? cst[b] : 0ull;
        for (uint32_t t = t0; t < t1; t++)
            acc += (uint64_t)(int64_t)coef[t] * pool[(size_t)idx[t] * width + j];
        o[j] = acc;
    }
}

// ---- peer windows: all-gather of a split level's result blocks as ONE kernel over NVLink peer memory ----
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// spin until *p has reached `want` (sequence numbers, wrap-safe); gives up after `timeout` clocks so that a
// dead peer costs an error code, not a hung GPU
__device__ __forceinline__ bool wait_reached(const uint32_t *p, uint32_t want, long long timeout, const uint32_t *broken) {
    if (*reinterpret_cast<const volatile uint32_t *>(broken)) return false;   // an earlier wait timed out: do not wait again
    const long long t0 = clock64();
    while ((int32_t)(ld_acquire_sys(p) - want) < 0) {
        if (clock64() - t0 > timeout) return false;
        __nanosleep(64);
    }
    return true;
}

// Rank `rank` owns n_words u64 at word_off of its pool (the blocks it has just computed).  Every CTA copies a
// stride of them into the same place of every peer's pool (plain stores over NVLink), fences, and counts
// itself; the last CTA publishes `step` in arrive[rank] of every peer and then waits until every peer's
// `step` has arrived here -- when the kernel ends, the level's blocks of all ranks are in the local pool.
// need_done != 0 (first split level of a run on a reused pool): wait first until every peer has finished run
// need_done (polling ITS done word), because it may still be reading the slots these stores overwrite.
__global__ void __launch_bounds__(256)
peer_allgather_kernel(PeerPtrs pp, int rank, int world, size_t word_off, size_t n_words, uint32_t step,
                      uint32_t need_done, long long timeout) {
    PeerCtrl *mine = pp.ctrl[rank];
    __shared__ int s_flag;
    if (need_done) {
        if (threadIdx.x < world && threadIdx.x != rank && !wait_reached(&pp.ctrl[threadIdx.x]->done[0], need_done, timeout, &mine->error[0]))
            atomicExch(&mine->error[0], 1u);
        __syncthreads();
    }
    const uint64_t *__restrict__ src = pp.pool[rank] + word_off;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_words; i += (size_t)gridDim.x * blockDim.x) {
        const uint64_t v = src[i];
        for (int d = 1; d < world; d++) {
            int q = rank + d;
            if (q >= world) q -= world;
            pp.pool[q][word_off + i] = v;
        }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) s_flag = atomicAdd(&mine->cta_count[0], 1u) == gridDim.x - 1 ? 1 : 0;
    __syncthreads();
    if (!s_flag) return;
    if (threadIdx.x == 0) mine->cta_count[0] = 0;       // the next push kernel is stream-ordered after this one
    __threadfence_system();
    if (threadIdx.x < world && threadIdx.x != rank) {
        st_release_sys(&pp.ctrl[threadIdx.x]->arrive[rank][0], step);
        if (!wait_reached(&mine->arrive[threadIdx.x][0], step, timeout, &mine->error[0])) atomicExch(&mine->error[0], 1u);
    }
}

__global__ void peer_done_kernel(PeerCtrl *mine, uint32_t runs) {
    if (threadIdx.x == 0) st_release_sys(&mine->done[0], runs);
}

// Pool memory comes from the device's stream-ordered allocator on the context's stream (the default
// mempool keeps freed blocks cached: fhs_ctx_create lifts its release threshold), so creating and
// destroying a pool per operation costs microseconds instead of a cudaMalloc / cudaFree round trip each.
template <class T> cudaError_t pool_alloc(Ctx *c, T **p, size_t bytes) {
    return cudaMallocAsync(reinterpret_cast<void **>(p), bytes, c->stream);
}
inline void pool_free(Ctx *c, void *p) {
    if (p) cudaFreeAsync(p, c->stream);
}

int ensure_desc(Pool *p, size_t B, size_t T, size_t n_luts) {
    Ctx *c = p->ctx;
    if (B > p->cap_b) {
        size_t cap = p->cap_b ? p->cap_b : 1024;
        while (cap < B) cap *= 2;
        pool_free(c, p->d_off); pool_free(c, p->d_lut); pool_free(c, p->d_cst); pool_free(c, p->d_stage);
        p->d_off = p->d_lut = nullptr; p->d_cst = nullptr; p->d_stage = nullptr; p->cap_b = 0;
        FHS_CUDA(fhs::pool_alloc(c, &p->d_off, (cap + 1) * sizeof(uint32_t)));
        FHS_CUDA(fhs::pool_alloc(c, &p->d_lut, cap * sizeof(uint32_t)));
        FHS_CUDA(fhs::pool_alloc(c, &p->d_cst, cap * sizeof(uint64_t)));
        FHS_CUDA(fhs::pool_alloc(c, &p->d_stage, cap * ((size_t)c->big + 1) * sizeof(uint64_t)));
        p->cap_b = cap;
    }
    if (T > p->cap_t) {
        size_t cap = p->cap_t ? p->cap_t : 4096;
        while (cap < T) cap *= 2;
        pool_free(c, p->d_idx); pool_free(c, p->d_coef);
        p->d_idx = nullptr; p->d_coef = nullptr; p->cap_t = 0;
        FHS_CUDA(fhs::pool_alloc(c, &p->d_idx, cap * sizeof(uint32_t)));
        FHS_CUDA(fhs::pool_alloc(c, &p->d_coef, cap * sizeof(int32_t)));
        p->cap_t = cap;
    }
    if (n_luts > p->cap_luts) {
        size_t cap = p->cap_luts ? p->cap_luts : 64;
        while (cap < n_luts) cap *= 2;
        pool_free(c, p->d_tables);
        p->d_tables = nullptr; p->cap_luts = 0;
        FHS_CUDA(fhs::pool_alloc(c, &p->d_tables, cap * c->space * sizeof(uint64_t)));
        p->cap_luts = cap;
    }
    return FHS_OK;
}

}  // namespace
}  // namespace fhs

using fhs::Ctx;
using fhs::Pool;

extern "C" {

int fhs_pool_create(fhs_ctx *h, size_t capacity_blocks, fhs_pool **out) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !out || capacity_blocks == 0) return FHS_ERR_INVALID;
    if (!c->has_keys) { fhs::set_error("keys not uploaded"); return FHS_ERR_NO_KEYS; }
    FHS_CUDA(cudaSetDevice(c->device));
    Pool *p = new Pool();
    p->ctx = c;
    cudaError_t e = fhs::pool_alloc(c, &p->d, capacity_blocks * ((size_t)c->big + 1) * sizeof(uint64_t));
    if (e != cudaSuccess) { delete p; return fhs::cuda_fail(e, "cudaMallocAsync(pool)"); }
    // other streams (the pool's copy stream, a caller's NCCL stream) may touch the block right away
    if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess) { delete p; return fhs::cuda_fail(e, "pool alloc sync"); }
    p->capacity = capacity_blocks;
    *out = reinterpret_cast<fhs_pool *>(p);
    return FHS_OK;
}

void fhs_pool_destroy(fhs_pool *hp) {
    Pool *p = reinterpret_cast<Pool *>(hp);
    if (!p) return;
    cudaSetDevice(p->ctx->device);
    Ctx *c = p->ctx;
    cudaStreamSynchronize(c->stream);
    if (p->copy_stream) { cudaStreamSynchronize(p->copy_stream); cudaStreamDestroy(p->copy_stream); }
    for (int q = 0; q < fhs::kMaxPeers; q++)
        if (p->shared && q != p->rank && p->peer_pool[q]) cudaIpcCloseMemHandle(p->peer_pool[q]);
    if (p->own_malloc) { cudaFree(p->d); p->d = nullptr; }
    if (p->h_err) cudaFreeHost(p->h_err);
    void *all[] = {p->d, p->d_off, p->d_idx, p->d_lut, p->d_coef, p->d_cst, p->d_tables, p->d_stage,
                   p->r_off, p->r_idx, p->r_lut, p->r_coef, p->r_cst, p->r_tables, p->r_stage};
    for (void *q : all) fhs::pool_free(c, q);
    if (p->write_done) cudaEventDestroy(p->write_done);
    delete p;
}

size_t fhs_pool_capacity(fhs_pool *hp) { return hp ? reinterpret_cast<Pool *>(hp)->capacity : 0; }
uint64_t *fhs_pool_device_ptr(fhs_pool *hp) { return hp ? reinterpret_cast<Pool *>(hp)->d : nullptr; }

int fhs_pool_write(fhs_pool *hp, size_t slot, size_t n, const uint64_t *host_blocks) {
    Pool *p = reinterpret_cast<Pool *>(hp);
    if (!p || (!host_blocks && n) || slot + n > p->capacity) { fhs::set_error("pool write out of range"); return FHS_ERR_INVALID; }
    Ctx *c = p->ctx;
    FHS_CUDA(cudaSetDevice(c->device));
    const size_t w = (size_t)c->big + 1;
    FHS_CUDA(cudaMemcpyAsync(p->d + slot * w, host_blocks, n * w * sizeof(uint64_t), cudaMemcpyHostToDevice, c->stream));
    FHS_CUDA(cudaStreamSynchronize(c->stream));
    return FHS_OK;
}

int fhs_pool_read(fhs_pool *hp, size_t slot, size_t n, uint64_t *host_blocks) {
    Pool *p = reinterpret_cast<Pool *>(hp);
    if (!p || (!host_blocks && n) || slot + n > p->capacity) { fhs::set_error("pool read out of range"); return FHS_ERR_INVALID; }
    Ctx *c = p->ctx;
    FHS_CUDA(cudaSetDevice(c->device));
    const size_t w = (size_t)c->big + 1;
    FHS_CUDA(cudaMemcpyAsync(host_blocks, p->d + slot * w, n * w * sizeof(uint64_t), cudaMemcpyDeviceToHost, c->stream));
    FHS_CUDA(cudaStreamSynchronize(c->stream));
    return FHS_OK;
}

// Un-synchronised transfers on the context's stream, for callers that own the host buffer until their
// next fhs_ctx_synchronize (the cached executions of libfhs_string: page-locked staging), and device to
// device copies between a pool and a resident ciphertext array (fhs_dev_alloc).
int fhs_pool_enqueue_write(fhs_pool *hp, size_t slot, size_t n, const uint64_t *host_blocks) {
    Pool *p = reinterpret_cast<Pool *>(hp);
    if (!p || (!host_blocks && n) || slot + n > p->capacity) { fhs::set_error("pool write out of range"); return FHS_ERR_INVALID; }
    Ctx *c = p->ctx;
    FHS_CUDA(cudaSetDevice(c->device));
    const size_t w = (size_t)c->big + 1;
    FHS_CUDA(cudaMemcpyAsync(p->d + slot * w, host_blocks, n * w * sizeof(uint64_t), cudaMemcpyHostToDevice, c->stream));
    return FHS_OK;
}

int fhs_pool_enqueue_read(fhs_pool *hp, size_t slot, size_t n, uint64_t *host_blocks) {
    Pool *p = reinterpret_cast<Pool *>(hp);
    if (!p || (!host_blocks && n) || slot + n > p->capacity) { fhs::set_error("pool read out of range"); return FHS_ERR_INVALID; }
    Ctx *c = p->ctx;
    FHS_CUDA(cudaSetDevice(c->device));
    const size_t w = (size_t)c->big + 1;
    FHS_CUDA(cudaMemcpyAsync(host_blocks, p->d + slot * w, n * w * sizeof(uint64_t), cudaMemcpyDeviceToHost, c->stream));
    return FHS_OK;
}

int fhs_pool_copy_in(fhs_pool *hp, size_t slot, size_t n, const uint64_t *d_src) {
    Pool *p = reinterpret_cast<Pool *>(hp);
    if (!p || (!d_src && n) || slot + n > p->capacity) { fhs::set_error("pool copy out of range"); return FHS_ERR_INVALID; }
    Ctx *c = p->ctx;
    FHS_CUDA(cudaSetDevice(c->device));
    const size_t w = (size_t)c->big + 1;
    FHS_CUDA(cudaMemcpyAsync(p->d + slot * w, d_src, n * w * sizeof(uint64_t), cudaMemcpyDeviceToDevice, c->stream));
    return FHS_OK;
}

int fhs_pool_copy_out(fhs_pool *hp, size_t slot, size_t n, uint64_t *d_dst) {
    Pool *p = reinterpret_cast<Pool *>(hp);
    if (!p || (!d_dst && n) || slot + n > p->capacity) { fhs::set_error("pool copy out of range"); return FHS_ERR_INVALID; }
    Ctx *c = p->ctx;
    FHS_CUDA(cudaSetDevice(c->device));
    const size_t w = (size_t)c->big + 1;
    FHS_CUDA(cudaMemcpyAsync(d_dst, p->d + slot * w, n * w * sizeof(uint64_t), cudaMemcpyDeviceToDevice, c->stream));
    return FHS_OK;
}

// One level: B linear combinations of pool blocks; if lut_id != NULL each is keyswitched and
// bootstrapped through lut_tables[lut_id[b]]; results land in pool slots [out_slot, out_slot + B).
int fhs_level(fhs_pool *hp, size_t B, const uint32_t *term_off, const uint32_t *term_idx,
              const int32_t *term_coef, const uint64_t *const_body, const uint32_t *lut_id,
              const uint64_t *lut_tables, size_t n_luts, size_t out_slot) {
    Pool *p = reinterpret_cast<Pool *>(hp);
    if (!p) return FHS_ERR_INVALID;
    if (B == 0) return FHS_OK;
    Ctx *c = p->ctx;
    if (!term_off || !const_body || out_slot + B > p->capacity || (lut_id && (!lut_tables || !n_luts))) {
        fhs::set_error("fhs_level: bad arguments (out_slot %zu + B %zu vs capacity %zu)", out_slot, B, p->capacity);
        return FHS_ERR_INVALID;
    }
    const size_t T = term_off[B];
    if (T && (!term_idx || !term_coef)) return FHS_ERR_INVALID;
    for (size_t t = 0; t < T; t++)
        if (term_idx[t] >= p->capacity) { fhs::set_error("fhs_level: term_idx[%zu] = %u outside the pool (%zu)", t, term_idx[t], p->capacity); return FHS_ERR_INVALID; }
    if (lut_id)
        for (size_t b = 0; b < B; b++)
            if (lut_id[b] >= n_luts) { fhs::set_error("fhs_level: lut_id[%zu] = %u but there are %zu tables", b, lut_id[b], n_luts); return FHS_ERR_INVALID; }
    FHS_CUDA(cudaSetDevice(c->device));
    int rc = fhs::ensure_desc(p, B, T, lut_id ? n_luts : 0);
    if (rc != FHS_OK) return rc;
    cudaStream_t s = c->stream;
    FHS_CUDA(cudaMemcpyAsync(p->d_off, term_off, (B + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, s));
    if (T) {
        FHS_CUDA(cudaMemcpyAsync(p->d_idx, term_idx, T * sizeof(uint32_t), cudaMemcpyHostToDevice, s));
        FHS_CUDA(cudaMemcpyAsync(p->d_coef, term_coef, T * sizeof(int32_t), cudaMemcpyHostToDevice, s));
    }
    FHS_CUDA(cudaMemcpyAsync(p->d_cst, const_body, B * sizeof(uint64_t), cudaMemcpyHostToDevice, s));
    const int width = (int)c->big + 1;
    uint64_t *dst = p->d + out_slot * (size_t)width;
    if (!lut_id) {
        fhs::gather_linear_kernel<<<(unsigned)B, 256, 0, s>>>(p->d, p->d_off, p->d_idx, p->d_coef, p->d_cst, dst, width);
        c->launches++;
        FHS_CUDA(cudaGetLastError());
        return FHS_OK;
    }
    FHS_CUDA(cudaMemcpyAsync(p->d_lut, lut_id, B * sizeof(uint32_t), cudaMemcpyHostToDevice, s));
    FHS_CUDA(cudaMemcpyAsync(p->d_tables, lut_tables, n_luts * c->space * sizeof(uint64_t), cudaMemcpyHostToDevice, s));
    fhs::gather_linear_kernel<<<(unsigned)B, 256, 0, s>>>(p->d, p->d_off, p->d_idx, p->d_coef, p->d_cst, p->d_stage, width);
    c->launches++;
    FHS_CUDA(cudaGetLastError());
    return fhs_ks_pbs_batch_dev(reinterpret_cast<fhs_ctx *>(c), p->d_stage, p->d_lut, p->d_tables, n_luts, B, dst);
}

int fhs_level_d(fhs_pool *hp, const fhs_level_desc *d) {
    if (!d) return FHS_ERR_INVALID;
    return fhs_level(hp, d->B, d->term_off, d->term_idx, d->term_coef, d->const_body, d->lut_id, d->lut_tables,
                     d->n_luts, d->out_slot);
}

// Upload on a side stream: returns once the host buffer has been consumed (pageable memory is staged
// by the runtime), the DMA itself overlaps with kernels already queued on the context's stream.
// fhs_pool_wait_writes makes the context's stream wait for it.  The caller guarantees that no
// kernel still reads or writes the target slots (e.g. it has read back the previous results).
int fhs_pool_write_async(fhs_pool *hp, size_t slot, size_t n, const uint64_t *host_blocks) {
    Pool *p = reinterpret_cast<Pool *>(hp);
    if (!p || (!host_blocks && n) || slot + n > p->capacity) { fhs::set_error("pool write out of range"); return FHS_ERR_INVALID; }
    Ctx *c = p->ctx;
    FHS_CUDA(cudaSetDevice(c->device));
    if (!p->copy_stream) {
        FHS_CUDA(cudaStreamCreateWithFlags(&p->copy_stream, cudaStreamNonBlocking));
        FHS_CUDA(cudaEventCreateWithFlags(&p->write_done, cudaEventDisableTiming));
    }
    const size_t w = (size_t)c->big + 1;
    FHS_CUDA(cudaMemcpyAsync(p->d + slot * w, host_blocks, n * w * sizeof(uint64_t), cudaMemcpyHostToDevice, p->copy_stream));
    FHS_CUDA(cudaEventRecord(p->write_done, p->copy_stream));
    p->write_pending = true;
    return FHS_OK;
}

int fhs_pool_wait_writes(fhs_pool *hp) {
    Pool *p = reinterpret_cast<Pool *>(hp);
    if (!p) return FHS_ERR_INVALID;
    if (!p->write_pending) return FHS_OK;
    FHS_CUDA(cudaSetDevice(p->ctx->device));
    FHS_CUDA(cudaStreamWaitEvent(p->ctx->stream, p->write_done, 0));
    p->write_pending = false;
    return FHS_OK;
}

// Resident program: all level descriptors of a level program are uploaded ONCE (rows in execution
// order, row_off absolute term offsets), so that running a level is kernel launches only -- no
// host->device descriptor copies and no host synchronisation between the levels of a deep program.
int fhs_pool_load_program(fhs_pool *hp, size_t rows, size_t terms, const uint32_t *row_off,
                          const uint32_t *term_idx, const int32_t *term_coef, const uint64_t *const_body,
                          const uint32_t *lut_id, const uint64_t *lut_tables, size_t n_luts,
                          size_t max_level_rows) {
    Pool *p = reinterpret_cast<Pool *>(hp);
    if (!p || !row_off || (terms && (!term_idx || !term_coef)) || (rows && (!const_body || !lut_id)) ||
        (n_luts && !lut_tables)) return FHS_ERR_INVALID;
    for (size_t t = 0; t < terms; t++)
        if (term_idx[t] >= p->capacity) { fhs::set_error("fhs_pool_load_program: term_idx[%zu] = %u outside the pool (%zu)", t, term_idx[t], p->capacity); return FHS_ERR_INVALID; }
    for (size_t r = 0; r < rows; r++)   // rows run with is_pbs = 0 carry lut_id 0
        if (lut_id[r] >= (n_luts ? n_luts : 1)) { fhs::set_error("fhs_pool_load_program: lut_id[%zu] = %u but there are %zu tables", r, lut_id[r], n_luts); return FHS_ERR_INVALID; }
    Ctx *c = p->ctx;
    FHS_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = c->stream;
    FHS_CUDA(cudaStreamSynchronize(s));
    {
        void *old[] = {p->r_off, p->r_idx, p->r_lut, p->r_coef, p->r_cst, p->r_tables, p->r_stage};
        for (void *q : old) fhs::pool_free(c, q);
    }
    p->r_off = p->r_idx = p->r_lut = nullptr; p->r_coef = nullptr;
    p->r_cst = p->r_tables = p->r_stage = nullptr;
    p->r_rows = 0;
    const size_t w = (size_t)c->big + 1;
    FHS_CUDA(fhs::pool_alloc(c, &p->r_off, (rows + 1) * sizeof(uint32_t)));
    FHS_CUDA(fhs::pool_alloc(c, &p->r_idx, (terms + 1) * sizeof(uint32_t)));
    FHS_CUDA(fhs::pool_alloc(c, &p->r_coef, (terms + 1) * sizeof(int32_t)));
    FHS_CUDA(fhs::pool_alloc(c, &p->r_cst, (rows + 1) * sizeof(uint64_t)));
    FHS_CUDA(fhs::pool_alloc(c, &p->r_lut, (rows + 1) * sizeof(uint32_t)));
    FHS_CUDA(fhs::pool_alloc(c, &p->r_tables, (n_luts + 1) * c->space * sizeof(uint64_t)));
    FHS_CUDA(fhs::pool_alloc(c, &p->r_stage, (max_level_rows + 1) * w * sizeof(uint64_t)));
    FHS_CUDA(cudaMemcpyAsync(p->r_off, row_off, (rows + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, s));
    if (terms) {
        FHS_CUDA(cudaMemcpyAsync(p->r_idx, term_idx, terms * sizeof(uint32_t), cudaMemcpyHostToDevice, s));
        FHS_CUDA(cudaMemcpyAsync(p->r_coef, term_coef, terms * sizeof(int32_t), cudaMemcpyHostToDevice, s));
    }
    if (rows) {
        FHS_CUDA(cudaMemcpyAsync(p->r_cst, const_body, rows * sizeof(uint64_t), cudaMemcpyHostToDevice, s));
        FHS_CUDA(cudaMemcpyAsync(p->r_lut, lut_id, rows * sizeof(uint32_t), cudaMemcpyHostToDevice, s));
    }
    if (n_luts)
        FHS_CUDA(cudaMemcpyAsync(p->r_tables, lut_tables, n_luts * c->space * sizeof(uint64_t), cudaMemcpyHostToDevice, s));
    FHS_CUDA(cudaStreamSynchronize(s));   // host arrays may be freed by the caller
    p->r_rows = rows; p->r_luts = n_luts; p->r_max_rows = max_level_rows;
    return FHS_OK;
}

// Rows [row_lo, row_lo + B) of the resident program -> pool slots [out_slot, out_slot + B).
// Asynchronous on the context's stream.
int fhs_level_run(fhs_pool *hp, size_t row_lo, size_t B, int is_pbs, size_t out_slot) {
    Pool *p = reinterpret_cast<Pool *>(hp);
    if (!p) return FHS_ERR_INVALID;
    if (B == 0) return FHS_OK;
    Ctx *c = p->ctx;
    if (row_lo + B > p->r_rows || out_slot + B > p->capacity || (is_pbs && (B > p->r_max_rows || !p->r_luts))) {
        fhs::set_error("fhs_level_run: rows [%zu, %zu) of %zu, out_slot %zu, capacity %zu", row_lo, row_lo + B,
                       p->r_rows, out_slot, p->capacity);
        return FHS_ERR_INVALID;
    }
    FHS_CUDA(cudaSetDevice(c->device));
    const int width = (int)c->big + 1;
    uint64_t *dst = p->d + out_slot * (size_t)width;
    fhs::gather_linear_kernel<<<(unsigned)B, 256, 0, c->stream>>>(p->d, p->r_off + row_lo, p->r_idx, p->r_coef,
                                                                   p->r_cst + row_lo, is_pbs ? p->r_stage : dst, width);
    c->launches++;
    FHS_CUDA(cudaGetLastError());
    if (!is_pbs) return FHS_OK;
    return fhs_ks_pbs_batch_dev(reinterpret_cast<fhs_ctx *>(c), p->r_stage, p->r_lut + row_lo, p->r_tables,
                                p->r_luts, B, dst);
}

// ---- peer windows (include/fhs_cuda.h "own collective") ------------------------------------------------
// Collective over the context's communicator: every rank calls it at the same point with the same
// capacity.  The pool is a cudaMalloc block (legacy CUDA IPC cannot export stream-ordered allocations) whose
// handle travels through the NCCL communicator once; from then on the ranks exchange result blocks with
// peer_allgather_kernel.  If any rank cannot export / map (no P2P, IPC closed in the container), all ranks
// agree to leave the pool unshared and the caller keeps using fhs_comm_allgather.
int fhs_pool_create_shared(fhs_ctx *h, size_t capacity_blocks, fhs_pool **out) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !out || capacity_blocks == 0) return FHS_ERR_INVALID;
    if (!c->has_keys) { fhs::set_error("keys not uploaded"); return FHS_ERR_NO_KEYS; }
    if (!c->nccl_comm || c->comm_world < 2 || c->comm_world > fhs::kMaxPeers || !c->knobs.peer_p2p)
        return fhs_pool_create(h, capacity_blocks, out);
    FHS_CUDA(cudaSetDevice(c->device));
    const int rank = c->comm_rank, world = c->comm_world;
    const size_t pool_bytes = (capacity_blocks * ((size_t)c->big + 1) * sizeof(uint64_t) + 255) & ~(size_t)255;
    Pool *p = new Pool();
    p->ctx = c; p->rank = rank; p->world = world; p->capacity = capacity_blocks;
    cudaError_t e = cudaMalloc(reinterpret_cast<void **>(&p->d), pool_bytes + sizeof(fhs::PeerCtrl));
    if (e != cudaSuccess) { delete p; return fhs::cuda_fail(e, "cudaMalloc(shared pool)"); }
    p->own_malloc = true;
    fhs::PeerCtrl *ctrl = reinterpret_cast<fhs::PeerCtrl *>(reinterpret_cast<uint8_t *>(p->d) + pool_bytes);
    struct Slot { cudaIpcMemHandle_t handle; uint32_t ok; uint32_t pad[15]; };
    static_assert(sizeof(Slot) == 128, "handle slot");
    std::vector<Slot> hs(world);
    memset(hs.data(), 0, world * sizeof(Slot));
    Slot *d_slots = nullptr;
    int rc = FHS_OK;
    bool agreed = false;
    do {
        if ((e = cudaHostAlloc(reinterpret_cast<void **>(&p->h_err), sizeof(uint32_t), cudaHostAllocDefault)) != cudaSuccess) break;
        *p->h_err = 0;
        if ((e = cudaMalloc(reinterpret_cast<void **>(&d_slots), world * sizeof(Slot))) != cudaSuccess) break;
        if ((e = cudaMemsetAsync(ctrl, 0, sizeof(fhs::PeerCtrl), c->stream)) != cudaSuccess) break;
        // round 1: handles.  round 2: did every rank map every peer?
        bool mine_ok = cudaIpcGetMemHandle(&hs[rank].handle, p->d) == cudaSuccess;
        if (!mine_ok) cudaGetLastError();
        for (int round = 0; round < 2 && rc == FHS_OK; round++) {
            hs[rank].ok = mine_ok ? 1u : 0u;
            if ((e = cudaMemcpyAsync(d_slots + rank, &hs[rank], sizeof(Slot), cudaMemcpyHostToDevice, c->stream)) != cudaSuccess) break;
            if ((rc = fhs_comm_allgather(h, reinterpret_cast<uint64_t *>(d_slots), sizeof(Slot))) != FHS_OK) break;
            if ((e = cudaMemcpyAsync(hs.data(), d_slots, world * sizeof(Slot), cudaMemcpyDeviceToHost, c->stream)) != cudaSuccess) break;
            if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess) break;
            bool all_ok = true;
            for (int q = 0; q < world; q++) all_ok = all_ok && hs[q].ok;
            if (!all_ok) { mine_ok = false; if (round == 1) break; continue; }
            if (round == 0) {
                for (int q = 0; q < world && mine_ok; q++) {
                    if (q == rank) continue;
                    void *ptr = nullptr;
                    if (cudaIpcOpenMemHandle(&ptr, hs[q].handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                        cudaGetLastError();
                        mine_ok = false;
                    } else {
                        p->peer_pool[q] = static_cast<uint64_t *>(ptr);
                    }
                }
            } else {
                agreed = true;
            }
        }
    } while (0);
    if (d_slots) cudaFree(d_slots);
    if (rc == FHS_OK && e != cudaSuccess) rc = fhs::cuda_fail(e, "fhs_pool_create_shared");
    if (rc != FHS_OK || !agreed) {
        for (int q = 0; q < world; q++)
            if (q != rank && p->peer_pool[q]) { cudaIpcCloseMemHandle(p->peer_pool[q]); p->peer_pool[q] = nullptr; }
    }
    if (rc != FHS_OK) { cudaFree(p->d); if (p->h_err) cudaFreeHost(p->h_err); delete p; return rc; }
    if (agreed) {
        p->shared = true;
        p->peer_pool[rank] = p->d;
        for (int q = 0; q < world; q++)
            p->peer_ctrl[q] = reinterpret_cast<fhs::PeerCtrl *>(reinterpret_cast<uint8_t *>(p->peer_pool[q]) + pool_bytes);
    }
    *out = reinterpret_cast<fhs_pool *>(p);
    return FHS_OK;
}

int fhs_pool_is_shared(fhs_pool *hp) { return hp && reinterpret_cast<Pool *>(hp)->shared ? 1 : 0; }

// All-gather of the blocks of one split level: rank q has written slots [slot_base + q*per_rank, +n_q) of its
// own pool (n_q <= per_rank; this rank's n is my_blocks); afterwards every rank's pool holds all of them.
// Asynchronous on the context's stream; every rank calls it for the same levels in the same order.
int fhs_pool_peer_allgather(fhs_pool *hp, size_t slot_base, size_t per_rank, size_t my_blocks) {
    Pool *p = reinterpret_cast<Pool *>(hp);
    if (!p || !p->shared) { fhs::set_error("pool has no peer window"); return FHS_ERR_COMM; }
    if (my_blocks > per_rank || slot_base + per_rank * (size_t)p->world > p->capacity) {
        fhs::set_error("fhs_pool_peer_allgather: slots out of range");
        return FHS_ERR_INVALID;
    }
    Ctx *c = p->ctx;
    FHS_CUDA(cudaSetDevice(c->device));
    const size_t w = (size_t)c->big + 1;
    fhs::PeerPtrs pp;
    for (int q = 0; q < fhs::kMaxPeers; q++) { pp.pool[q] = p->peer_pool[q]; pp.ctrl[q] = p->peer_ctrl[q]; }
    const size_t n_words = my_blocks * w;
    unsigned grid = (unsigned)((n_words + 4095) / 4096);
    if (grid < 1) grid = 1;
    if (grid > 2u * (unsigned)c->sm_count) grid = 2u * (unsigned)c->sm_count;
    const uint32_t need_done = p->first_of_run ? p->runs : 0u;
    p->first_of_run = false;
    p->step++;
    const long long timeout = (long long)c->knobs.peer_timeout_ms * 2000000ll;   // ~2 GHz
    fhs::peer_allgather_kernel<<<grid, 256, 0, c->stream>>>(pp, p->rank, p->world, (slot_base + per_rank * (size_t)p->rank) * w,
                                                             n_words, p->step, need_done, timeout);
    c->launches++;
    p->pushes++;
    c->peer_pushes++;
    FHS_CUDA(cudaGetLastError());
    return FHS_OK;
}

// End of a run on a shared pool: tells the peers that this rank no longer reads the blocks of this run (they may
// overwrite them with the next run's) and fetches the error word.  Call after the last kernel that reads the pool.
int fhs_pool_peer_run_done(fhs_pool *hp) {
    Pool *p = reinterpret_cast<Pool *>(hp);
    if (!p || !p->shared) return FHS_OK;
    if (p->first_of_run) return FHS_OK;          // no split level ran
    Ctx *c = p->ctx;
    FHS_CUDA(cudaSetDevice(c->device));
    p->runs++;
    p->first_of_run = true;
    fhs::peer_done_kernel<<<1, 32, 0, c->stream>>>(p->peer_ctrl[p->rank], p->runs);
    c->launches++;
    FHS_CUDA(cudaGetLastError());
    FHS_CUDA(cudaMemcpyAsync(p->h_err, &p->peer_ctrl[p->rank]->error[0], sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    return FHS_OK;
}

// After a stream synchronisation: FHS_ERR_COMM if a wait of the last run(s) timed out (a peer died or fell
// more than FHS_PEER_TIMEOUT_MS behind); the blocks of that run are then not to be trusted.
int fhs_pool_peer_status(fhs_pool *hp) {
    Pool *p = reinterpret_cast<Pool *>(hp);
    if (!p || !p->shared || !p->h_err || *p->h_err == 0) return FHS_OK;
    fhs::set_error("peer window: a wait for another rank timed out");
    return FHS_ERR_COMM;
}

unsigned long long fhs_comm_peer_pushes(fhs_ctx *h) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    return c ? c->peer_pushes : 0ull;
}


}  // extern "C"
