// C ABI of libfhs_cuda.so (include/fhs_cuda.h): context, key upload, batched KS-PBS.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "fhs_internal.cuh"

namespace fhs {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char *what) {
    set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
    return e == cudaErrorMemoryAllocation ? FHS_ERR_OOM : FHS_ERR_CUDA;
}

static int ensure_capacity(Ctx *c, size_t batch, size_t n_luts) {
    if (batch > c->cap) {
        size_t cap = c->cap ? c->cap : 256;
        while (cap < batch) cap *= 2;
        cudaFree(c->d_small);
        cudaFree(c->d_in);
        cudaFree(c->d_out);
        cudaFree(c->d_lut_id);
        c->d_small = c->d_in = c->d_out = nullptr;
        c->d_lut_id = nullptr;
        c->cap = 0;
        const size_t in_w = (size_t)c->big + 1;
        FHS_CUDA(cudaMalloc(&c->d_small, cap * in_w * sizeof(uint64_t)));  // also holds small-LWE input
        FHS_CUDA(cudaMalloc(&c->d_in, cap * in_w * sizeof(uint64_t)));
        FHS_CUDA(cudaMalloc(&c->d_out, cap * in_w * sizeof(uint64_t)));
        FHS_CUDA(cudaMalloc(&c->d_lut_id, cap * sizeof(uint32_t)));
        c->cap = cap;
    }
    if (n_luts > c->cap_tables) {
        size_t cap = c->cap_tables ? c->cap_tables : 64;
        while (cap < n_luts) cap *= 2;
        cudaFree(c->d_tables);
        c->d_tables = nullptr;
        c->cap_tables = 0;
        FHS_CUDA(cudaMalloc(&c->d_tables, cap * c->space * sizeof(uint64_t)));
        c->cap_tables = cap;
    }
    return FHS_OK;
}

static void free_keys(Ctx *c) {
    cudaFree(c->d_ksk);
    cudaFree(c->d_ksk_corr);
    cudaFree(c->d_ksk_t);
    cudaFree(c->d_ksk_u);
    c->d_ksk_t = c->d_ksk_u = nullptr;
    cudaFree(c->d_bsk_f);
    cudaFree(c->d_bsk_w);
    cudaFree(c->d_bsk_x);
    c->d_ksk = c->d_ksk_corr = nullptr;
    c->d_bsk_f = c->d_bsk_w = c->d_bsk_x = nullptr;
    c->has_keys = false;
}

}  // namespace fhs

using fhs::Ctx;

extern "C" {

void fhs_default_params(fhs_params *p) {
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
}

const char *fhs_last_error(void) { return fhs::g_err; }
int fhs_version(void) { return 100; }

static int ctx_init(Ctx *c, int device);

static void read_knobs(Ctx *c) {
    auto env_int = [](const char *name, long dflt) { const char *e = getenv(name); return e ? atol(e) : dflt; };
    c->knobs.wide_limit = env_int("FHS_BR_WIDE", 2L * c->sm_count);
    c->knobs.pair_on = (int)env_int("FHS_BR_PAIR", 1);
    c->knobs.stagger = (int)env_int("FHS_BR_STAGGER", 4);
    c->knobs.opt = (int)env_int("FHS_BR_OPT", 13) & 13;
    c->knobs.ks_variant = (int)env_int("FHS_KS_VARIANT", 3);
    c->knobs.ks_umma_min = (int)env_int("FHS_KS_UMMA_MIN", 1);
    c->knobs.ks_mma_min = (int)env_int("FHS_KS_MMA_MIN", 1);
    c->knobs.ks_splitk = (int)env_int("FHS_KS_SPLITK", 1);
    c->knobs.bsk_f64 = env_int("FHS_BSK_XF", 1) == 0;
    c->knobs.ks_baseline_keys = c->knobs.ks_variant != 3 || c->knobs.ks_umma_min > 1;
    c->knobs.peer_p2p = (int)env_int("FHS_COMM_P2P", 1);
    c->knobs.peer_timeout_ms = env_int("FHS_PEER_TIMEOUT_MS", 30000);
    if (c->knobs.peer_timeout_ms < 1) c->knobs.peer_timeout_ms = 1;
}

int fhs_ctx_create(int device, fhs_ctx **out) {
    if (!out) return FHS_ERR_INVALID;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        fhs::set_error("no CUDA device available (%s): this library has no CPU fallback",
                       cudaGetErrorString(e));
        return FHS_ERR_CUDA;
    }
    if (device < 0 || device >= count) {
        fhs::set_error("device %d out of range (have %d)", device, count);
        return FHS_ERR_INVALID;
    }
    FHS_CUDA(cudaSetDevice(device));
    Ctx *c = new (std::nothrow) Ctx();
    if (!c) return FHS_ERR_OOM;
    c->device = device;
    const int rc = ctx_init(c, device);
    if (rc != FHS_OK) {          // release whatever was created before the failing call
        fhs_ctx_destroy(reinterpret_cast<fhs_ctx *>(c));
        return rc;
    }
    *out = reinterpret_cast<fhs_ctx *>(c);
    return FHS_OK;
}

static int ctx_init(Ctx *c, int device) {
    cudaDeviceProp prop;
    FHS_CUDA(cudaGetDeviceProperties(&prop, device));
    c->sm_count = prop.multiProcessorCount;
    if (prop.major < 10) {
        fhs::set_error("device %d is sm_%d%d; this library is built for sm_100a only", device,
                       prop.major, prop.minor);
        return FHS_ERR_CUDA;
    }
    read_knobs(c);
    FHS_CUDA(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
    c->stream = c->own_stream;
    {   // ciphertext pools use the stream-ordered allocator (level.cu): keep freed blocks cached
        cudaMemPool_t mp;
        FHS_CUDA(cudaDeviceGetDefaultMemPool(&mp, device));
        uint64_t keep = ~0ull;
        FHS_CUDA(cudaMemPoolSetAttribute(mp, cudaMemPoolAttrReleaseThreshold, &keep));
    }
    for (int i = 0; i < 4; i++) FHS_CUDA(cudaEventCreate(&c->ev[i]));
    FHS_CUDA(cudaMalloc(&c->d_counter, 64));
    // inter-pass twiddles T[k2][n1] = e^{2 pi i n1 (k2 + 1/4)/1024}, padded rows of 33
    std::vector<double2> T(32 * fhs::kTStride, make_double2(0.0, 0.0));
    const long double PI = 3.14159265358979323846264338327950288L;
    for (int k2 = 0; k2 < 32; k2++)
        for (int n1 = 0; n1 < 32; n1++) {
            // angle = 2 pi n1 (4 k2 + 1) / 4096, reduced exactly in integers
            const int num = (n1 * (4 * k2 + 1)) % 4096;
            const long double a = 2.0L * PI * (long double)num / 4096.0L;
            T[k2 * fhs::kTStride + n1] = make_double2((double)cosl(a), (double)sinl(a));
        }
    FHS_CUDA(cudaMalloc(&c->d_twiddle, T.size() * sizeof(double2)));
    FHS_CUDA(cudaMemcpy(c->d_twiddle, T.data(), T.size() * sizeof(double2), cudaMemcpyHostToDevice));
    return FHS_OK;
}

void fhs_ctx_destroy(fhs_ctx *h) {
    if (!h) return;
    Ctx *c = reinterpret_cast<Ctx *>(h);
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    fhs::resident_release(c);
    fhs::comm_release(c);
    fhs::free_keys(c);
    cudaFree(c->d_twiddle);
    cudaFree(c->d_small);
    cudaFree(c->d_in);
    cudaFree(c->d_out);
    cudaFree(c->d_lut_id);
    cudaFree(c->d_tables);
    cudaFree(c->d_counter);
    cudaFree(c->d_ks_digits);
    for (int i = 0; i < 4; i++)
        if (c->ev[i]) cudaEventDestroy(c->ev[i]);
    if (c->copy_stream) {
        cudaStreamSynchronize(c->copy_stream);
        cudaStreamDestroy(c->copy_stream);
        for (int i = 0; i < 7; i++)
            if (c->ev_pipe[i]) cudaEventDestroy(c->ev_pipe[i]);
    }
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
}

int fhs_ctx_set_stream(fhs_ctx *h, void *cuda_stream) {
    if (!h) return FHS_ERR_INVALID;
    Ctx *c = reinterpret_cast<Ctx *>(h);
    c->stream = cuda_stream ? reinterpret_cast<cudaStream_t>(cuda_stream) : c->own_stream;
    return FHS_OK;
}

int fhs_ctx_synchronize(fhs_ctx *h) {
    if (!h) return FHS_ERR_INVALID;
    Ctx *c = reinterpret_cast<Ctx *>(h);
    FHS_CUDA(cudaSetDevice(c->device));
    FHS_CUDA(cudaStreamSynchronize(c->stream));
    return FHS_OK;
}

int fhs_ctx_get_params(fhs_ctx *h, fhs_params *out) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !out) return FHS_ERR_INVALID;
    if (!c->has_keys) {
        fhs::set_error("keys not uploaded");
        return FHS_ERR_NO_KEYS;
    }
    *out = c->p;
    return FHS_OK;
}

// bsk_format: -1 = standard domain (u64 coefficients), else an FHS_FOURIER_* frequency convention
static int upload_keys_impl(fhs_ctx *h, const fhs_params *p, const uint64_t *ksk, int ksk_level_order,
                            const void *bsk, int bsk_format) {
    if (!h || !p || !ksk || !bsk) {
        fhs::set_error("fhs_upload_keys: null argument");
        return FHS_ERR_INVALID;
    }
    if (bsk_format != -1 && bsk_format != FHS_FOURIER_NATURAL_POS && bsk_format != FHS_FOURIER_NATURAL_NEG) {
        fhs::set_error("bad fourier_convention");
        return FHS_ERR_INVALID;
    }
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (p->glwe_dim != 1 || p->poly_size != (uint32_t)fhs::kN || p->pbs_level != 1 ||
        p->pbs_base_log != 23) {
        fhs::set_error(
            "unsupported PBS parameters (kernels are built for k=1, N=2048, pbs_base_log=23, "
            "pbs_level=1: the PARAM_MESSAGE_2_CARRY_2_KS_PBS family)");
        return FHS_ERR_INVALID;
    }
    if (p->ks_base_log != 3 || p->ks_level != 5 || p->lwe_dim + 1 > 768 || p->lwe_dim == 0) {
        fhs::set_error("unsupported keyswitch parameters (built for ks_base_log=3, ks_level=5, n<768)");
        return FHS_ERR_INVALID;
    }
    const uint32_t space = p->message_modulus * p->carry_modulus;
    if (space == 0 || (space & (space - 1)) != 0 || space > (uint32_t)fhs::kN / 2) {
        fhs::set_error("message_modulus*carry_modulus must be a power of two <= N/2");
        return FHS_ERR_INVALID;
    }
    if (ksk_level_order != FHS_KSK_LEVEL_MSB_FIRST && ksk_level_order != FHS_KSK_LEVEL_LSB_FIRST) {
        fhs::set_error("bad ksk_level_order");
        return FHS_ERR_INVALID;
    }
    FHS_CUDA(cudaSetDevice(c->device));
    fhs::free_keys(c);
    c->p = *p;
    c->n = p->lwe_dim;
    c->big = p->glwe_dim * p->poly_size;
    c->space = space;
    const size_t ksk_len = (size_t)c->big * p->ks_level * (c->n + 1);
    const size_t bsk_len = (size_t)c->n * 4 * fhs::kN;
    uint64_t *d_raw = nullptr;
    const size_t raw_len = ksk_len > bsk_len ? ksk_len : bsk_len;
    FHS_CUDA(cudaMalloc(&d_raw, raw_len * sizeof(uint64_t)));
    int rc = FHS_OK;
    do {
        cudaError_t e;
        // the tcgen05 keyswitch needs only its own operand tiles; the digit-order key is a staging
        // buffer for the tilers, and the key forms of the two baseline kernels (offset-correction row,
        // mma.sync tiles; ~120 MB and ~1.3 ms of upload kernels) exist only for A/B contexts
        const bool base = c->knobs.ks_baseline_keys;
        if ((e = cudaMalloc(&c->d_ksk, ksk_len * sizeof(uint64_t))) != cudaSuccess ||
            (base && (e = cudaMalloc(&c->d_ksk_corr, (c->n + 1) * sizeof(uint64_t))) != cudaSuccess) ||
            (base && (e = cudaMalloc(&c->d_ksk_t, fhs::ksk_tiles_bytes(c))) != cudaSuccess) ||
            (e = cudaMalloc(&c->d_ksk_u, fhs::ksk_umma_tiles_bytes(c))) != cudaSuccess ||
            (e = cudaMalloc(&c->d_bsk_f, fhs::bsk_fourier_elems(c) * sizeof(double2))) != cudaSuccess ||
            (e = cudaMalloc(&c->d_bsk_w, fhs::bsk_fourier_elems(c) * sizeof(double2))) != cudaSuccess ||
            (e = cudaMalloc(&c->d_bsk_x, fhs::bsk_fourier_elems(c) * sizeof(double2))) != cudaSuccess) {
            rc = fhs::cuda_fail(e, "cudaMalloc(keys)");
            break;
        }
        if ((e = cudaMemcpyAsync(d_raw, ksk, ksk_len * sizeof(uint64_t), cudaMemcpyHostToDevice,
                                 c->stream)) != cudaSuccess) {
            rc = fhs::cuda_fail(e, "cudaMemcpyAsync(ksk)");
            break;
        }
        if ((rc = fhs::launch_ksk_prepare(c, d_raw, ksk_level_order)) != FHS_OK) break;
        if (base && (rc = fhs::launch_ksk_corr(c)) != FHS_OK) break;
        if (base && (rc = fhs::launch_ksk_tile(c)) != FHS_OK) break;
        if ((rc = fhs::launch_ksk_tile_umma(c)) != FHS_OK) break;
        if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess) {
            rc = fhs::cuda_fail(e, "ksk prepare");
            break;
        }
        if (!base) { cudaFree(c->d_ksk); c->d_ksk = nullptr; }
        if ((e = cudaMemcpyAsync(d_raw, bsk, bsk_len * sizeof(uint64_t), cudaMemcpyHostToDevice,
                                 c->stream)) != cudaSuccess) {
            rc = fhs::cuda_fail(e, "cudaMemcpyAsync(bsk)");
            break;
        }
        if (bsk_format != -1) {
            // spectra computed by the caller (f64, 8 bytes per real: same size as the standard key)
            if ((rc = fhs::launch_bsk_import_fourier(c, reinterpret_cast<const double2 *>(d_raw),
                                                     bsk_format == FHS_FOURIER_NATURAL_NEG)) != FHS_OK)
                break;
        } else {
            // Fourier BSK rounded once from a double-double transform (bsk_fourier.cu); the f64
            // transform that shares the hot loop's FFT code stays selectable for A/B noise runs
            const bool f64_only = c->knobs.bsk_f64 != 0;
            if ((rc = f64_only ? fhs::launch_bsk_transform(c, d_raw) : fhs::launch_bsk_transform_dd(c, d_raw)) != FHS_OK)
                break;
        }
        if ((rc = fhs::launch_bsk_pair_repack(c)) != FHS_OK) break;
        if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess) {
            rc = fhs::cuda_fail(e, "bsk transform");
            break;
        }
    } while (0);
    cudaFree(d_raw);
    if (rc != FHS_OK) {
        fhs::free_keys(c);
        return rc;
    }
    c->has_keys = true;
    return FHS_OK;
}

int fhs_upload_keys(fhs_ctx *h, const fhs_params *p, const uint64_t *ksk, int ksk_level_order,
                    const uint64_t *bsk) {
    return upload_keys_impl(h, p, ksk, ksk_level_order, bsk, -1);
}

int fhs_upload_keys_fourier(fhs_ctx *h, const fhs_params *p, const uint64_t *ksk, int ksk_level_order,
                            const double *bsk_fourier, int fourier_convention) {
    return upload_keys_impl(h, p, ksk, ksk_level_order, bsk_fourier, fourier_convention);
}

int fhs_upload_keys_fmt(fhs_ctx *h, const fhs_params *p, const uint64_t *ksk, int ksk_level_order,
                        const void *bsk, int bsk_format) {
    return upload_keys_impl(h, p, ksk, ksk_level_order, bsk, bsk_format);
}

// host-pointer entry points: every lut_id must name a row of lut_tables (device-pointer calls cannot be
// checked without a synchronisation; there the kernels clamp the id)
static int check_lut_ids(const uint32_t *lut_id, size_t batch, size_t n_luts) {
    for (size_t b = 0; b < batch; b++)
        if (lut_id[b] >= n_luts) {
            fhs::set_error("lut_id[%zu] = %u but there are %zu tables", b, lut_id[b], n_luts);
            return FHS_ERR_INVALID;
        }
    return FHS_OK;
}

static int check_batch_args(Ctx *c, const void *a, const void *b, const void *t, const void *o,
                            size_t n_luts, size_t batch) {
    if (!c) return FHS_ERR_INVALID;
    if (!c->has_keys) {
        fhs::set_error("keys not uploaded");
        return FHS_ERR_NO_KEYS;
    }
    if (batch == 0) return FHS_OK;
    if (!a || !b || !t || !o || n_luts == 0) {
        fhs::set_error("null buffer or empty LUT table");
        return FHS_ERR_INVALID;
    }
    if (batch > 0xFFFFFFF0ull) {
        fhs::set_error("batch too large");
        return FHS_ERR_INVALID;
    }
    return FHS_OK;
}

int fhs_ks_pbs_batch_dev(fhs_ctx *h, const uint64_t *d_in, const uint32_t *d_lut_id,
                         const uint64_t *d_tables, size_t n_luts, size_t batch, uint64_t *d_out) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    int rc = check_batch_args(c, d_in, d_lut_id, d_tables, d_out, n_luts, batch);
    if (rc != FHS_OK || batch == 0) return rc;
    FHS_CUDA(cudaSetDevice(c->device));
    if ((rc = fhs::ensure_capacity(c, batch, 0)) != FHS_OK) return rc;
    FHS_CUDA(cudaEventRecord(c->ev[0], c->stream));
    if ((rc = fhs::launch_keyswitch(c, d_in, batch, c->d_small)) != FHS_OK) return rc;
    FHS_CUDA(cudaEventRecord(c->ev[1], c->stream));
    if ((rc = fhs::launch_blind_rotate(c, c->d_small, d_lut_id, d_tables, n_luts, batch, d_out)) != FHS_OK)
        return rc;
    FHS_CUDA(cudaEventRecord(c->ev[2], c->stream));
    c->ev_valid[0] = c->ev_valid[1] = true;
    return FHS_OK;
}

int fhs_ks_pbs_batch(fhs_ctx *h, const uint64_t *lwe_in, const uint32_t *lut_id,
                     const uint64_t *lut_tables, size_t n_luts, size_t batch, uint64_t *lwe_out) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    int rc = check_batch_args(c, lwe_in, lut_id, lut_tables, lwe_out, n_luts, batch);
    if (rc != FHS_OK || batch == 0) return rc;
    if ((rc = check_lut_ids(lut_id, batch, n_luts)) != FHS_OK) return rc;
    FHS_CUDA(cudaSetDevice(c->device));
    if ((rc = fhs::ensure_capacity(c, batch, n_luts)) != FHS_OK) return rc;
    const size_t w = (size_t)c->big + 1;
    // Ciphertexts are independent, so the batch is cut at blind-rotation wave boundaries
    // (4 ciphertexts per SM) into a short head, a body and the last partial wave: the body's
    // upload and the head's download hide behind compute, only the head's upload and the tail's
    // download stay exposed.  Batches of less than three waves go through in one piece.
    const size_t wave = (size_t)4 * c->sm_count;
    size_t cut[4] = {0, batch, batch, batch};
    int n_chunks = 1;
    if (batch >= 3 * wave) {
        const size_t tail = batch % wave ? batch % wave : wave;
        cut[1] = wave;
        cut[2] = batch - tail;
        cut[3] = batch;
        n_chunks = 3;
    }
    FHS_CUDA(cudaMemcpyAsync(c->d_lut_id, lut_id, batch * sizeof(uint32_t), cudaMemcpyHostToDevice,
                             c->stream));
    FHS_CUDA(cudaMemcpyAsync(c->d_tables, lut_tables, n_luts * c->space * sizeof(uint64_t),
                             cudaMemcpyHostToDevice, c->stream));
    if (n_chunks == 1) {
        FHS_CUDA(cudaMemcpyAsync(c->d_in, lwe_in, batch * w * sizeof(uint64_t), cudaMemcpyHostToDevice,
                                 c->stream));
        if ((rc = fhs_ks_pbs_batch_dev(h, c->d_in, c->d_lut_id, c->d_tables, n_luts, batch, c->d_out)) !=
            FHS_OK)
            return rc;
        FHS_CUDA(cudaMemcpyAsync(lwe_out, c->d_out, batch * w * sizeof(uint64_t), cudaMemcpyDeviceToHost,
                                 c->stream));
        FHS_CUDA(cudaStreamSynchronize(c->stream));
        return FHS_OK;
    }
    if (!c->copy_stream) {
        FHS_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
        for (int i = 0; i < 7; i++) FHS_CUDA(cudaEventCreateWithFlags(&c->ev_pipe[i], cudaEventDisableTiming));
    }
    // the copy stream starts after whatever the compute stream still has queued on the staging buffers
    FHS_CUDA(cudaEventRecord(c->ev_pipe[6], c->stream));
    FHS_CUDA(cudaStreamWaitEvent(c->copy_stream, c->ev_pipe[6], 0));
    for (int i = 0; i < n_chunks; i++) {
        const size_t lo = cut[i], nb = cut[i + 1] - cut[i];
        FHS_CUDA(cudaMemcpyAsync(c->d_in + lo * w, lwe_in + lo * w, nb * w * sizeof(uint64_t),
                                 cudaMemcpyHostToDevice, c->copy_stream));
        FHS_CUDA(cudaEventRecord(c->ev_pipe[i], c->copy_stream));
    }
    FHS_CUDA(cudaEventRecord(c->ev[0], c->stream));
    for (int i = 0; i < n_chunks; i++) {
        const size_t lo = cut[i], nb = cut[i + 1] - cut[i];
        FHS_CUDA(cudaStreamWaitEvent(c->stream, c->ev_pipe[i], 0));
        uint64_t *small = c->d_small + lo * ((size_t)c->n + 1);
        if ((rc = fhs::launch_keyswitch(c, c->d_in + lo * w, nb, small)) != FHS_OK) return rc;
        if ((rc = fhs::launch_blind_rotate(c, small, c->d_lut_id + lo, c->d_tables, n_luts, nb, c->d_out + lo * w)) !=
            FHS_OK)
            return rc;
        FHS_CUDA(cudaEventRecord(c->ev_pipe[3 + i], c->stream));
        FHS_CUDA(cudaStreamWaitEvent(c->copy_stream, c->ev_pipe[3 + i], 0));
        FHS_CUDA(cudaMemcpyAsync(lwe_out + lo * w, c->d_out + lo * w, nb * w * sizeof(uint64_t),
                                 cudaMemcpyDeviceToHost, c->copy_stream));
    }
    FHS_CUDA(cudaEventRecord(c->ev[2], c->stream));
    c->ev_valid[0] = c->ev_valid[1] = false;   // per-kernel split is not defined for interleaved chunks
    FHS_CUDA(cudaStreamSynchronize(c->copy_stream));
    FHS_CUDA(cudaStreamSynchronize(c->stream));
    return FHS_OK;
}

int fhs_keyswitch_batch_dev(fhs_ctx *h, const uint64_t *d_in, size_t batch, uint64_t *d_small_out) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return FHS_ERR_INVALID;
    if (!c->has_keys) {
        fhs::set_error("keys not uploaded");
        return FHS_ERR_NO_KEYS;
    }
    if (batch == 0) return FHS_OK;
    if (!d_in || !d_small_out) return FHS_ERR_INVALID;
    FHS_CUDA(cudaSetDevice(c->device));
    FHS_CUDA(cudaEventRecord(c->ev[0], c->stream));
    int rc = fhs::launch_keyswitch(c, d_in, batch, d_small_out);
    if (rc != FHS_OK) return rc;
    FHS_CUDA(cudaEventRecord(c->ev[1], c->stream));
    c->ev_valid[0] = true;
    c->ev_valid[1] = false;
    return FHS_OK;
}

int fhs_keyswitch_batch(fhs_ctx *h, const uint64_t *lwe_in, size_t batch, uint64_t *small_out) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return FHS_ERR_INVALID;
    if (!c->has_keys) {
        fhs::set_error("keys not uploaded");
        return FHS_ERR_NO_KEYS;
    }
    if (batch == 0) return FHS_OK;
    if (!lwe_in || !small_out) return FHS_ERR_INVALID;
    FHS_CUDA(cudaSetDevice(c->device));
    int rc = fhs::ensure_capacity(c, batch, 0);
    if (rc != FHS_OK) return rc;
    const size_t w = (size_t)c->big + 1;
    FHS_CUDA(cudaMemcpyAsync(c->d_in, lwe_in, batch * w * sizeof(uint64_t), cudaMemcpyHostToDevice,
                             c->stream));
    if ((rc = fhs_keyswitch_batch_dev(h, c->d_in, batch, c->d_small)) != FHS_OK) return rc;
    FHS_CUDA(cudaMemcpyAsync(small_out, c->d_small, batch * (c->n + 1) * sizeof(uint64_t),
                             cudaMemcpyDeviceToHost, c->stream));
    FHS_CUDA(cudaStreamSynchronize(c->stream));
    return FHS_OK;
}

int fhs_build_lut(fhs_ctx *h, const uint64_t *table, uint64_t *acc_out) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !table || !acc_out) return FHS_ERR_INVALID;
    if (!c->has_keys) {
        fhs::set_error("keys not uploaded (parameters unknown)");
        return FHS_ERR_NO_KEYS;
    }
    FHS_CUDA(cudaSetDevice(c->device));
    int rc = fhs::ensure_capacity(c, 4, 1);
    if (rc != FHS_OK) return rc;
    FHS_CUDA(cudaMemcpyAsync(c->d_tables, table, c->space * sizeof(uint64_t), cudaMemcpyHostToDevice,
                             c->stream));
    if ((rc = fhs::launch_build_lut(c, c->d_tables, c->d_out)) != FHS_OK) return rc;
    FHS_CUDA(cudaMemcpyAsync(acc_out, c->d_out, 2 * (size_t)fhs::kN * sizeof(uint64_t),
                             cudaMemcpyDeviceToHost, c->stream));
    FHS_CUDA(cudaStreamSynchronize(c->stream));
    return FHS_OK;
}

int fhs_pbs_batch_dev(fhs_ctx *h, const uint64_t *d_small_in, const uint32_t *d_lut_id,
                      const uint64_t *d_tables, size_t n_luts, size_t batch, uint64_t *d_out) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    int rc = check_batch_args(c, d_small_in, d_lut_id, d_tables, d_out, n_luts, batch);
    if (rc != FHS_OK || batch == 0) return rc;
    FHS_CUDA(cudaSetDevice(c->device));
    FHS_CUDA(cudaEventRecord(c->ev[1], c->stream));
    if ((rc = fhs::launch_blind_rotate(c, d_small_in, d_lut_id, d_tables, n_luts, batch, d_out)) != FHS_OK)
        return rc;
    FHS_CUDA(cudaEventRecord(c->ev[2], c->stream));
    c->ev_valid[0] = false;
    c->ev_valid[1] = true;
    return FHS_OK;
}

int fhs_pbs_batch(fhs_ctx *h, const uint64_t *small_in, const uint32_t *lut_id,
                  const uint64_t *lut_tables, size_t n_luts, size_t batch, uint64_t *lwe_out) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    int rc = check_batch_args(c, small_in, lut_id, lut_tables, lwe_out, n_luts, batch);
    if (rc != FHS_OK || batch == 0) return rc;
    if ((rc = check_lut_ids(lut_id, batch, n_luts)) != FHS_OK) return rc;
    FHS_CUDA(cudaSetDevice(c->device));
    if ((rc = fhs::ensure_capacity(c, batch, n_luts)) != FHS_OK) return rc;
    const size_t w = (size_t)c->big + 1;
    FHS_CUDA(cudaMemcpyAsync(c->d_small, small_in, batch * (c->n + 1) * sizeof(uint64_t),
                             cudaMemcpyHostToDevice, c->stream));
    FHS_CUDA(cudaMemcpyAsync(c->d_lut_id, lut_id, batch * sizeof(uint32_t), cudaMemcpyHostToDevice,
                             c->stream));
    FHS_CUDA(cudaMemcpyAsync(c->d_tables, lut_tables, n_luts * c->space * sizeof(uint64_t),
                             cudaMemcpyHostToDevice, c->stream));
    if ((rc = fhs_pbs_batch_dev(h, c->d_small, c->d_lut_id, c->d_tables, n_luts, batch, c->d_out)) !=
        FHS_OK)
        return rc;
    FHS_CUDA(cudaMemcpyAsync(lwe_out, c->d_out, batch * w * sizeof(uint64_t), cudaMemcpyDeviceToHost,
                             c->stream));
    FHS_CUDA(cudaStreamSynchronize(c->stream));
    return FHS_OK;
}

int fhs_last_kernel_ms(fhs_ctx *h, float *ms2) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !ms2) return FHS_ERR_INVALID;
    FHS_CUDA(cudaSetDevice(c->device));
    ms2[0] = ms2[1] = 0.f;
    if (c->ev_valid[0]) {
        FHS_CUDA(cudaEventSynchronize(c->ev[1]));
        FHS_CUDA(cudaEventElapsedTime(&ms2[0], c->ev[0], c->ev[1]));
    }
    if (c->ev_valid[1]) {
        FHS_CUDA(cudaEventSynchronize(c->ev[2]));
        FHS_CUDA(cudaEventElapsedTime(&ms2[1], c->ev[1], c->ev[2]));
    }
    return FHS_OK;
}

uint64_t fhs_launch_count(fhs_ctx *h) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    return c ? c->launches : 0;
}

}  // extern "C"
