// Internal declarations shared by the CUDA translation units of libfhs_cuda.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

#include "../../include/fhs_cuda.h"

namespace fhs {

// Compile-time geometry of the blind-rotation kernels (PARAM_MESSAGE_2_CARRY_2_KS_PBS family:
// k = 1, N = 2048, one PBS level).  Other sets are rejected at fhs_upload_keys.
constexpr int kN = 2048;         // polynomial size
constexpr int kH = kN / 2;       // complex points per polynomial (folded real FFT)
constexpr int kTStride = 33;     // padded row stride (complex elements) of 32x32 smem tiles

struct Ctx {
    int device = 0;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    fhs_params p{};
    bool has_keys = false;
    uint32_t n = 0, big = 0;       // lwe_dim, k*N
    uint32_t space = 0;            // message_modulus * carry_modulus
    // device key material
    uint64_t *d_ksk = nullptr;     // [k*N][ks_level (digit order, least significant first)][n+1]
    uint8_t *d_ksk_t = nullptr;    // byte-limb tiles of the KSK for the tensor-core keyswitch
    uint8_t *d_ksk_u = nullptr;    // the same bytes as UMMA operand tiles (keyswitch_umma.cu)
    uint8_t *d_ks_digits = nullptr;// scratch: signed digits of a batch as UMMA operand tiles
    size_t cap_digits = 0;
    uint64_t *d_ksk_corr = nullptr;// [n+1]  2^(base_log-1) * sum of all KSK rows (unsigned-digit offset)
    double2 *d_bsk_f = nullptr;    // Fourier BSK, layout in blind_rotate.cu
    double2 *d_bsk_w = nullptr;    // the same values in the layout of blind_rotate_wide.cu
    double2 *d_bsk_x = nullptr;    // ... and in the layout of blind_rotate_pair.cu (one GGSW row per CTA of a pair)
    int pair_limit = -1;           // largest batch of the pair kernel on this device (-1: not queried yet)
    // Dispatch knobs, read from the environment ONCE per context in fhs_ctx_create (they used to be
    // process-wide statics initialised from the first context's device).
    struct Knobs {
        long wide_limit = 0;       // FHS_BR_WIDE: largest batch of the wide kernel (default 2 per SM; 0 disables)
        int pair_on = 1;           // FHS_BR_PAIR=0 disables the 2-CTA-cluster kernel
        int stagger = 4;           // FHS_BR_STAGGER (0 | 3 | 4)
        int opt = 13;              // FHS_BR_OPT (0 | 1 | 5 | 13)
        int ks_variant = 3;        // FHS_KS_VARIANT: 3 tcgen05 (default), 2 mma.sync baseline, 1 IMAD baseline
        int ks_umma_min = 1;       // FHS_KS_UMMA_MIN
        int ks_mma_min = 1;        // FHS_KS_MMA_MIN
        int ks_splitk = 1;         // FHS_KS_SPLITK=0: no split-K in the tcgen05 keyswitch at small batches
        int bsk_f64 = 0;           // FHS_BSK_XF=0: plain f64 key transform instead of double-double
        int ks_baseline_keys = 0;  // key forms of the baseline keyswitch kernels are resident (variant != 3)
        int peer_p2p = 1;          // FHS_COMM_P2P=0: sharded levels exchange with ncclAllGather instead of the peer-window kernel
        long peer_timeout_ms = 30000;   // FHS_PEER_TIMEOUT_MS: a wait for another rank gives up after this long
    } knobs;
    double2 *d_twiddle = nullptr;  // [32][33] inter-pass twiddles T[k2][n1]
    // scratch
    uint64_t *d_small = nullptr;   // [cap][n+1] keyswitched ciphertexts
    uint64_t *d_in = nullptr;      // [cap][big+1] staging for host-buffer calls
    uint64_t *d_out = nullptr;
    uint32_t *d_lut_id = nullptr;
    uint64_t *d_tables = nullptr;
    size_t cap = 0, cap_tables = 0;
    unsigned int *d_counter = nullptr;
    cudaStream_t copy_stream = nullptr;   // host-buffer calls: transfers overlapped with compute
    cudaEvent_t ev_pipe[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    bool ev_valid[2] = {false, false};
    uint64_t launches = 0;
    unsigned long long peer_pushes = 0;   // level.cu: peer-window all-gathers launched
    int sm_count = 0;
    // cudaFuncSetAttribute is per device: done once per context, not once per process
    bool attr_br = false, attr_brw = false, attr_ks = false, attr_ksu = false;
    // comm.cu: NCCL communicator of this rank (one process per GPU); null on single-GPU contexts
    void *nccl_comm = nullptr;
    void *resident = nullptr;      // resident.cu: attachment cache, device timer (created on first use)
    int comm_rank = 0, comm_world = 1;
};

void set_error(const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what);

#define FHS_CUDA(expr)                                            \
    do {                                                          \
        cudaError_t _e = (expr);                                  \
        if (_e != cudaSuccess) return ::fhs::cuda_fail(_e, #expr); \
    } while (0)

// keyswitch.cu
int launch_keyswitch(Ctx *c, const uint64_t *d_in, size_t batch, uint64_t *d_small);
int launch_ksk_prepare(Ctx *c, const uint64_t *d_ksk_raw, int level_order);
int launch_ksk_corr(Ctx *c);     // offset-correction row of the IMAD baseline kernel only
// keyswitch_mma.cu
size_t ksk_tiles_bytes(const Ctx *c);
int launch_ksk_tile(Ctx *c);
int launch_keyswitch_mma(Ctx *c, const uint64_t *d_in, size_t batch, uint64_t *d_small);
// keyswitch_umma.cu: tcgen05.mma kind::i8, accumulators in TMEM (large batches)
size_t ksk_umma_tiles_bytes(const Ctx *c);
int launch_ksk_tile_umma(Ctx *c);
int launch_keyswitch_umma(Ctx *c, const uint64_t *d_in, size_t batch, uint64_t *d_small);
// blind_rotate.cu
// every blind-rotation launcher takes n_luts: the kernels clamp lut_id[b] to n_luts - 1, so a bad id in a
// device-pointer call reads a wrong table instead of memory outside lut_tables
int launch_blind_rotate(Ctx *c, const uint64_t *d_small, const uint32_t *d_lut_id,
                        const uint64_t *d_tables, size_t n_luts, size_t batch, uint64_t *d_out);
int launch_bsk_transform(Ctx *c, const uint64_t *d_bsk_std);
// bsk_fourier.cu: the same transform in double-double arithmetic (default; FHS_BSK_XF=0 selects the f64 one)
int launch_bsk_transform_dd(Ctx *c, const uint64_t *d_bsk_std);
// bsk_fourier.cu: spectra supplied by the caller in natural frequency order (fhs_upload_keys_fourier)
int launch_bsk_import_fourier(Ctx *c, const double2 *d_bsk_fourier, int conj_index);
// blind_rotate_wide.cu: 4 warps per ciphertext, for levels of at most 2 ciphertexts per SM
int launch_blind_rotate_wide(Ctx *c, const uint64_t *d_small, const uint32_t *d_lut_id,
                             const uint64_t *d_tables, size_t n_luts, size_t batch, uint64_t *d_out);
int launch_build_lut(Ctx *c, const uint64_t *d_table, uint64_t *d_acc);
size_t bsk_fourier_elems(const Ctx *c);
// blind_rotate_pair.cu: one ciphertext per 2-CTA cluster, for levels of at most one ciphertext per two SMs
int launch_bsk_pair_repack(Ctx *c);
int blind_rotate_pair_limit(Ctx *c);
int launch_blind_rotate_pair(Ctx *c, const uint64_t *d_small, const uint32_t *d_lut_id,
                             const uint64_t *d_tables, size_t n_luts, size_t batch, uint64_t *d_out);
// comm.cu
void comm_release(Ctx *c);
// resident.cu: runs the destructors of everything cached on the context (before its memory goes away)
void resident_release(Ctx *c);

}  // namespace fhs
