// Residency services of the context (include/fhs_cuda.h "resident operands and cached executions"):
//   * device-resident ciphertext arrays: results of one operation feed the next without a PCIe round trip;
//   * page-locked host buffers for operand / result staging;
//   * an attachment cache on the context: the level scheduler (libfhs_string) parks the planned +
//     descriptor-loaded pool of a program here, so that the second call of an operation is kernel
//     launches only.  Entries die with the context, with their owner (fhs_cache_drop_owner) or by LRU when
//     the cached device bytes exceed the budget;
//   * a device timer (events on the context's stream) for "device ms" next to wall-clock ms.
#include <cstdlib>
#include <cstring>
#include <list>
#include <mutex>
#include <set>

#include "fhs_internal.cuh"

namespace fhs {

struct CacheEntry {
    uint64_t key[4];
    void *obj;
    fhs_cache_dtor dtor;
    size_t bytes;
};

struct Resident {
    std::list<CacheEntry> lru;      // front = most recently used
    size_t bytes = 0, limit = 0;
    cudaEvent_t t0 = nullptr, t1 = nullptr, in_done = nullptr;
    bool timing = false;
};

namespace {
std::mutex g_reg_mu;
std::set<Ctx *> g_live;            // contexts whose caches fhs_cache_drop_owner must visit

Resident *res_of(Ctx *c) {
    if (!c->resident) {
        Resident *r = new Resident();
        const char *e = getenv("FHS_CACHE_BYTES");
        r->limit = e ? (size_t)atoll(e) : ((size_t)16 << 30);
        c->resident = r;
        std::lock_guard<std::mutex> lk(g_reg_mu);
        g_live.insert(c);
    }
    return static_cast<Resident *>(c->resident);
}

void evict(Resident *r, const CacheEntry *keep) {
    while (r->bytes > r->limit && !r->lru.empty()) {
        CacheEntry &e = r->lru.back();
        if (&e == keep) break;      // never the entry that was just inserted / used
        r->bytes -= e.bytes;
        if (e.dtor) e.dtor(e.obj);
        r->lru.pop_back();
    }
}
}  // namespace

void resident_release(Ctx *c) {
    if (!c->resident) return;
    {
        std::lock_guard<std::mutex> lk(g_reg_mu);
        g_live.erase(c);
    }
    Resident *r = static_cast<Resident *>(c->resident);
    for (auto &e : r->lru)
        if (e.dtor) e.dtor(e.obj);  // pools are destroyed while the context is still whole
    r->lru.clear();
    if (r->t0) cudaEventDestroy(r->t0);
    if (r->t1) cudaEventDestroy(r->t1);
    if (r->in_done) cudaEventDestroy(r->in_done);
    delete r;
    c->resident = nullptr;
}

}  // namespace fhs

using fhs::Ctx;
using fhs::Resident;

extern "C" {

int fhs_ctx_cache_get(fhs_ctx *h, const uint64_t key[4], void **obj) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !key || !obj) return FHS_ERR_INVALID;
    *obj = nullptr;
    Resident *r = fhs::res_of(c);
    for (auto it = r->lru.begin(); it != r->lru.end(); ++it)
        if (!memcmp(it->key, key, sizeof(it->key))) {
            r->lru.splice(r->lru.begin(), r->lru, it);
            *obj = it->obj;
            break;
        }
    return FHS_OK;
}

int fhs_ctx_cache_put(fhs_ctx *h, const uint64_t key[4], void *obj, fhs_cache_dtor dtor, size_t device_bytes) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !key || !obj) return FHS_ERR_INVALID;
    Resident *r = fhs::res_of(c);
    for (auto it = r->lru.begin(); it != r->lru.end(); ++it)
        if (!memcmp(it->key, key, sizeof(it->key))) {   // replace
            r->bytes -= it->bytes;
            if (it->dtor && it->obj != obj) it->dtor(it->obj);
            r->lru.erase(it);
            break;
        }
    fhs::CacheEntry e;
    memcpy(e.key, key, sizeof(e.key));
    e.obj = obj; e.dtor = dtor; e.bytes = device_bytes;
    r->lru.push_front(e);
    r->bytes += device_bytes;
    fhs::evict(r, &r->lru.front());
    return FHS_OK;
}

void fhs_cache_drop_owner(uint64_t owner) {
    std::lock_guard<std::mutex> lk(fhs::g_reg_mu);
    for (Ctx *c : fhs::g_live) {
        Resident *r = static_cast<Resident *>(c->resident);
        if (!r) continue;
        for (auto it = r->lru.begin(); it != r->lru.end();) {
            if (it->key[0] == owner) {
                r->bytes -= it->bytes;
                if (it->dtor) it->dtor(it->obj);
                it = r->lru.erase(it);
            } else {
                ++it;
            }
        }
    }
}

int fhs_ctx_cache_stats(fhs_ctx *h, size_t *entries, size_t *device_bytes) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return FHS_ERR_INVALID;
    Resident *r = fhs::res_of(c);
    if (entries) *entries = r->lru.size();
    if (device_bytes) *device_bytes = r->bytes;
    return FHS_OK;
}

int fhs_ctx_cache_limit(fhs_ctx *h, size_t device_bytes) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return FHS_ERR_INVALID;
    Resident *r = fhs::res_of(c);
    r->limit = device_bytes;
    fhs::evict(r, nullptr);
    return FHS_OK;
}

// ---- page-locked host memory ------------------------------------------------------------------
int fhs_host_alloc(fhs_ctx *h, size_t bytes, void **out) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !out) return FHS_ERR_INVALID;
    *out = nullptr;
    FHS_CUDA(cudaSetDevice(c->device));
    FHS_CUDA(cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault));
    return FHS_OK;
}

void fhs_host_free(fhs_ctx *h, void *p) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!p) return;
    if (c) cudaSetDevice(c->device);
    cudaFreeHost(p);
}

int fhs_host_is_pinned(const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return 0; }
    return a.type == cudaMemoryTypeHost;
}

// ---- device-resident ciphertext arrays ----------------------------------------------------------
int fhs_dev_alloc(fhs_ctx *h, size_t n_blocks, uint64_t **d_out) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !d_out) return FHS_ERR_INVALID;
    if (!c->has_keys) { fhs::set_error("keys not uploaded (block width unknown)"); return FHS_ERR_NO_KEYS; }
    FHS_CUDA(cudaSetDevice(c->device));
    const size_t bytes = (n_blocks ? n_blocks : 1) * ((size_t)c->big + 1) * sizeof(uint64_t);
    FHS_CUDA(cudaMallocAsync(reinterpret_cast<void **>(d_out), bytes, c->stream));
    return FHS_OK;
}

void fhs_dev_free(fhs_ctx *h, uint64_t *d) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !d) return;
    cudaSetDevice(c->device);
    cudaFreeAsync(d, c->stream);
}

int fhs_dev_write(fhs_ctx *h, uint64_t *d, size_t n_blocks, const uint64_t *host) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || (!d && n_blocks) || (!host && n_blocks)) return FHS_ERR_INVALID;
    FHS_CUDA(cudaSetDevice(c->device));
    FHS_CUDA(cudaMemcpyAsync(d, host, n_blocks * ((size_t)c->big + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, c->stream));
    FHS_CUDA(cudaStreamSynchronize(c->stream));
    return FHS_OK;
}

int fhs_dev_read(fhs_ctx *h, const uint64_t *d, size_t n_blocks, uint64_t *host) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || (!d && n_blocks) || (!host && n_blocks)) return FHS_ERR_INVALID;
    FHS_CUDA(cudaSetDevice(c->device));
    FHS_CUDA(cudaMemcpyAsync(host, d, n_blocks * ((size_t)c->big + 1) * sizeof(uint64_t), cudaMemcpyDeviceToHost, c->stream));
    FHS_CUDA(cudaStreamSynchronize(c->stream));
    return FHS_OK;
}

// ---- device timer ---------------------------------------------------------------------------------
int fhs_ctx_timer_start(fhs_ctx *h) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return FHS_ERR_INVALID;
    Resident *r = fhs::res_of(c);
    FHS_CUDA(cudaSetDevice(c->device));
    if (!r->t0) { FHS_CUDA(cudaEventCreate(&r->t0)); FHS_CUDA(cudaEventCreate(&r->t1)); }
    FHS_CUDA(cudaEventRecord(r->t0, c->stream));
    r->timing = true;
    return FHS_OK;
}

int fhs_ctx_timer_stop(fhs_ctx *h) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return FHS_ERR_INVALID;
    Resident *r = fhs::res_of(c);
    if (!r->timing) return FHS_ERR_INVALID;
    FHS_CUDA(cudaEventRecord(r->t1, c->stream));
    return FHS_OK;
}

int fhs_ctx_timer_ms(fhs_ctx *h, float *ms) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !ms) return FHS_ERR_INVALID;
    Resident *r = fhs::res_of(c);
    if (!r->timing) return FHS_ERR_INVALID;
    FHS_CUDA(cudaEventSynchronize(r->t1));
    FHS_CUDA(cudaEventElapsedTime(ms, r->t0, r->t1));
    return FHS_OK;
}

}  // extern "C"
