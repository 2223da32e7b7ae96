// Device-resident ciphertext pool and level execution (SURVEY.md 8b "device-resident variant"):
//   K1 gather + linear-combination prologue:  in[b] = sum_t coef[b][t] * pool[idx[b][t]] + const[b]
//   (bivariate pack 4x+y, compare x-y, block pack 4hi+lo, <=5-term sums, scalar add; replaces the
//   linear shortint ops tfhe-rs runs before each PBS: unchecked_scalar_mul / unchecked_add / sub),
//   then keyswitch + blind rotation of the whole level, results written to consecutive pool slots.
#include "fhs_internal.cuh"

namespace fhs {

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

}  // extern "C"
