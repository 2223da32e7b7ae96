// Multi-GPU plumbing of the path: one process per GPU, a key replica per GPU, and
// only two kinds of traffic -- the server keys once (broadcast) and the result blocks of a level
// that was split by rank (all-gather, in place on the ciphertext pool).  Both are NCCL collectives
// enqueued on the context's own stream, so a sharded level program runs without a host
// synchronisation between its levels and without any host framework in the loop (a Rust host binds
// these entry points directly).
//
// NCCL is resolved at run time with dlopen: inside a Python/torch process this picks up the
// libnccl.so.2 torch has already loaded (one NCCL per process), elsewhere the system library.  The
// library itself has no link-time dependency on NCCL; single-GPU users never touch it.
#include <dlfcn.h>

#include <cstring>
#include <mutex>

#include "fhs_internal.cuh"

namespace fhs {
namespace {

// the few NCCL 2.x declarations used here (stable ABI since 2.0; nccl.h is not needed to build)
struct NcclUniqueId { char internal[128]; };
typedef void *NcclComm;
constexpr int kNcclSuccess = 0;
constexpr int kNcclUint8 = 1;

struct NcclApi {
    void *lib = nullptr;
    int (*GetUniqueId)(NcclUniqueId *) = nullptr;
    int (*CommInitRank)(NcclComm *, int, NcclUniqueId, int) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, NcclComm, cudaStream_t) = nullptr;
    int (*Broadcast)(const void *, void *, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    bool tried = false, ok = false;
};
NcclApi g_nccl;

std::mutex g_nccl_mu;   // contexts of different threads may initialise their communicators concurrently

bool load_nccl() {
    std::lock_guard<std::mutex> lock(g_nccl_mu);
    if (g_nccl.tried) {
        if (!g_nccl.ok) set_error("NCCL unavailable (earlier dlopen / dlsym failure); set FHS_NCCL_LIB");
        return g_nccl.ok;
    }
    g_nccl.tried = true;
    const char *names[] = {getenv("FHS_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    for (const char *n : names) {
        if (!n || !*n) continue;
        g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (g_nccl.lib) break;
    }
    if (!g_nccl.lib) {
        set_error("NCCL not found (dlopen libnccl.so.2: %s); set FHS_NCCL_LIB", dlerror());
        return false;
    }
#define FHS_SYM(field, name)                                                      \
    *reinterpret_cast<void **>(&g_nccl.field) = dlsym(g_nccl.lib, name);          \
    if (!g_nccl.field) {                                                          \
        set_error("NCCL symbol %s missing", name);                                \
        return false;                                                             \
    }
    FHS_SYM(GetUniqueId, "ncclGetUniqueId");
    FHS_SYM(CommInitRank, "ncclCommInitRank");
    FHS_SYM(CommDestroy, "ncclCommDestroy");
    FHS_SYM(AllGather, "ncclAllGather");
    FHS_SYM(Broadcast, "ncclBroadcast");
    FHS_SYM(GetErrorString, "ncclGetErrorString");
#undef FHS_SYM
    g_nccl.ok = true;
    return true;
}

int nccl_fail(int r, const char *what) {
    set_error("%s: NCCL error %d (%s)", what, r, g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?");
    return FHS_ERR_COMM;
}

}  // namespace

void comm_release(Ctx *c) {
    if (c->nccl_comm && g_nccl.ok) g_nccl.CommDestroy(reinterpret_cast<NcclComm>(c->nccl_comm));
    c->nccl_comm = nullptr;
    c->comm_rank = 0;
    c->comm_world = 1;
}

}  // namespace fhs

using fhs::Ctx;

extern "C" {

int fhs_comm_unique_id(uint8_t id[FHS_COMM_ID_BYTES]) {
    if (!id) return FHS_ERR_INVALID;
    if (!fhs::load_nccl()) return FHS_ERR_COMM;
    fhs::NcclUniqueId u;
    const int r = fhs::g_nccl.GetUniqueId(&u);
    if (r != fhs::kNcclSuccess) return fhs::nccl_fail(r, "ncclGetUniqueId");
    static_assert(sizeof(u) == FHS_COMM_ID_BYTES, "unique id size");
    memcpy(id, &u, sizeof(u));
    return FHS_OK;
}

int fhs_comm_init(fhs_ctx *h, const uint8_t id[FHS_COMM_ID_BYTES], int rank, int world) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !id || world < 1 || rank < 0 || rank >= world) {
        fhs::set_error("fhs_comm_init: bad argument");
        return FHS_ERR_INVALID;
    }
    if (!fhs::load_nccl()) return FHS_ERR_COMM;
    FHS_CUDA(cudaSetDevice(c->device));
    fhs::comm_release(c);
    fhs::NcclUniqueId u;
    memcpy(&u, id, sizeof(u));
    fhs::NcclComm comm = nullptr;
    const int r = fhs::g_nccl.CommInitRank(&comm, world, u, rank);
    if (r != fhs::kNcclSuccess) return fhs::nccl_fail(r, "ncclCommInitRank");
    c->nccl_comm = comm;
    c->comm_rank = rank;
    c->comm_world = world;
    return FHS_OK;
}

int fhs_comm_info(fhs_ctx *h, int *rank, int *world) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return FHS_ERR_INVALID;
    if (!c->nccl_comm) {
        fhs::set_error("no communicator (fhs_comm_init)");
        return FHS_ERR_COMM;
    }
    if (rank) *rank = c->comm_rank;
    if (world) *world = c->comm_world;
    return FHS_OK;
}

void fhs_comm_destroy(fhs_ctx *h) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    fhs::comm_release(c);
}

int fhs_comm_allgather(fhs_ctx *h, uint64_t *d_base, size_t bytes_per_rank) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !d_base) return FHS_ERR_INVALID;
    if (!c->nccl_comm) {
        fhs::set_error("no communicator (fhs_comm_init)");
        return FHS_ERR_COMM;
    }
    if (bytes_per_rank == 0) return FHS_OK;
    FHS_CUDA(cudaSetDevice(c->device));
    // in place: rank q's contribution already sits at d_base + q * bytes_per_rank
    uint8_t *base = reinterpret_cast<uint8_t *>(d_base);
    const int r = fhs::g_nccl.AllGather(base + (size_t)c->comm_rank * bytes_per_rank, base, bytes_per_rank,
                                        fhs::kNcclUint8, reinterpret_cast<fhs::NcclComm>(c->nccl_comm), c->stream);
    if (r != fhs::kNcclSuccess) return fhs::nccl_fail(r, "ncclAllGather");
    return FHS_OK;
}

int fhs_comm_broadcast_host(fhs_ctx *h, void *host_buf, size_t bytes, int root) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || (!host_buf && bytes)) return FHS_ERR_INVALID;
    if (!c->nccl_comm) {
        fhs::set_error("no communicator (fhs_comm_init)");
        return FHS_ERR_COMM;
    }
    if (root < 0 || root >= c->comm_world) {
        fhs::set_error("fhs_comm_broadcast_host: bad root");
        return FHS_ERR_INVALID;
    }
    if (bytes == 0) return FHS_OK;
    FHS_CUDA(cudaSetDevice(c->device));
    void *d = nullptr;
    FHS_CUDA(cudaMalloc(&d, bytes));
    int rc = FHS_OK;
    cudaError_t e = cudaSuccess;
    if (c->comm_rank == root) e = cudaMemcpyAsync(d, host_buf, bytes, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) {
        const int r = fhs::g_nccl.Broadcast(d, d, bytes, fhs::kNcclUint8, root,
                                            reinterpret_cast<fhs::NcclComm>(c->nccl_comm), c->stream);
        if (r != fhs::kNcclSuccess) rc = fhs::nccl_fail(r, "ncclBroadcast");
    }
    if (rc == FHS_OK && e == cudaSuccess && c->comm_rank != root)
        e = cudaMemcpyAsync(host_buf, d, bytes, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d);
    if (rc == FHS_OK && e != cudaSuccess) rc = fhs::cuda_fail(e, "fhs_comm_broadcast_host");
    return rc;
}

}  // extern "C"
