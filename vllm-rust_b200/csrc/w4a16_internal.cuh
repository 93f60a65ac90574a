// w4a16_internal.cuh -- handle layout, error plumbing and kernel launch entry points
// shared by the translation units of libw4a16_b200.so.  Not part of the ABI.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include <atomic>
#include <string>
#include <vector>

#include "../../include/w4a16_b200.h"
#include "w4a16_epilogue.cuh"
#include "w4a16_layout.cuh"

namespace w4 {

// ---- errors ----------------------------------------------------------------
void set_error(const std::string &msg);
int fail(const std::string &msg);  // sets the error, returns 1
#define W4_CUDA_OK(expr)                                                                          \
    do {                                                                                          \
        cudaError_t _e = (expr);                                                                  \
        if (_e != cudaSuccess)                                                                    \
            return ::w4::fail(std::string(#expr) + ": " + cudaGetErrorString(_e));                \
    } while (0)

extern std::atomic<uint64_t> g_kernel_launches;
inline void count_launch() { g_kernel_launches.fetch_add(1, std::memory_order_relaxed); }

// ---- environment switches -----------------------------------------------------
// The library reads a handful of DOCUMENTED variables at run time: W4A16_GV_DEBUG (per-CTA timing stamps), W4A16_NO_PDL
// (no programmatic launches), W4A16_NO_FUSED_AR / W4A16_NO_CHAIN_KERNEL (fall back to NCCL / to per-linear launches),
// W4A16_IMMA8 (opt-in two-token-tile GEMV, parity-tested).  Every tuning / experiment knob (kernel selection, stage
// counts, split-K model constants, prefetch distance ...) is compiled out of the shipped library and only read when it
// is built with -DW4A16_EXPERIMENT_ENV (make EXTRA_NVFLAGS=-DW4A16_EXPERIMENT_ENV): the tools under tools/ say when
// they need that.
inline int exp_env_int(const char *name, int dflt) {
#ifdef W4A16_EXPERIMENT_ENV
    const char *v = getenv(name);
    return v && *v ? atoi(v) : dflt;
#else
    (void)name;
    return dflt;
#endif
}
inline double exp_env_double(const char *name, double dflt) {
#ifdef W4A16_EXPERIMENT_ENV
    const char *v = getenv(name);
    return v && *v ? atof(v) : dflt;
#else
    (void)name;
    return dflt;
#endif
}

// ---- per-device facts ----------------------------------------------------------
struct DeviceScratch {
    int sm_count = 0;
    int cc_major = 0, cc_minor = 0;
};
int get_device_scratch(int device, DeviceScratch **out);

// ---- the handle ---------------------------------------------------------------
struct Linear {
    // full (checkpoint) geometry
    int64_t K_full = 0, N_full = 0;
    int group_size_cfg = 0;  // as given: -1 or 32/64/128
    int format = 0;
    bool has_bias = false;
    int device = 0;
    int tp_mode = 0, tp_rank = 0, tp_size = 1;
    // local shard geometry
    int64_t k_off = 0, n_off = 0;
    Dims d{};
    // resident tensors
    uint32_t *qw = nullptr;        // [NR][KB][128]
    uint8_t *meta = nullptr;       // [NR][G][48]
    __nv_bfloat16 *bias = nullptr; // [N] or null
    int32_t *perm = nullptr;       // act-order: tiled row i holds checkpoint row perm[i]; null otherwise
    bool perm_external = false;    // act-order ROW shard: the rows were sorted over the FULL K before sharding; x arrives
                                   // already in that order (no gather at forward), hooks answer in tiled row order
    int32_t *tickets = nullptr;    // [NR] split-K arrival counters (self-resetting)
    float *splitk_scratch = nullptr;  // [SC][32][N] fp32 partials when the decode kernel splits K over CTAs
    int64_t splitk_floats = 0;
    bool loaded = false;
    // fused sibling projections (w4a16_load_*_part): column intervals filled so far; epilogue 1 = SwiGLU pair (the handle
    // holds gate rows in the g half and up rows in the g+8 half of every row block and writes silu(gate) * up, N / 2 wide)
    int epilogue = 0;
    int64_t part_cols[8][2] = {};
    int n_parts_loaded = 0;
    std::vector<int32_t> part_g_idx;   // act-order: every part must bring the same g_idx
    // scratch that scales with M
    __nv_bfloat16 *x_gather = nullptr;  // act-order gathered activations [cap, K]
    int64_t x_gather_cap = 0;
    __nv_bfloat16 *stage_x = nullptr, *stage_y = nullptr;  // forward_host staging
    int64_t stage_cap = 0;
    std::vector<void *> retired;   // outgrown scratch buffers: freed with the handle, never earlier (a captured CUDA graph
                                   // may still hold their addresses)
    float *rowpar_f32 = nullptr;  // unused for now (fp32 all-reduce staging)
    int *gv_tab = nullptr;        // decode stream kernel: per-CTA unit range + row-block sharing table (device)
    int gv_grid = 0, gv_max_slots = 0;
    bool gv_aligned = false;      // every CTA range is a whole number of row blocks
    DeviceScratch *scratch = nullptr;
};

// ---- kernels (defined in the .cu files) ----------------------------------------
struct RepackSrc {
    const uint32_t *qweight;  // AWQ [K_full, N_full/8]  or GPTQ [K_full/8, N_full]
    const uint32_t *qzeros;   // [G_full, N_full/8] or null (sym)
    const void *scales;       // [G_full, N_full]
    int scales_dtype;
    int format;
    int64_t K_full, N_full, G_full;
    int64_t k_off, n_off;     // shard offsets (rows / columns of the checkpoint)
    int gs;                   // effective group size
    const int32_t *perm;      // device, local row i -> checkpoint row (already includes k_off) or null
    const int32_t *row_group; // device, local row i -> checkpoint group index (act-order) or null
    int64_t dst_R0 = 0, NR_part = 0;   // column part: destination row blocks [dst_R0, dst_R0 + NR_part); 0 = the whole handle
    int half = -1;                     // SwiGLU pair: 0 = gate rows (g), 1 = up rows (g + 8); -1 = plain
};
int launch_repack(const Linear &L, const RepackSrc &src, cudaStream_t stream);
int launch_unpack(const Linear &L, uint8_t *q, uint8_t *z, cudaStream_t stream);
int launch_dequant(const Linear &L, __nv_bfloat16 *w, cudaStream_t stream);
int launch_scales_out(const Linear &L, __nv_bfloat16 *s, cudaStream_t stream);
int launch_dequant_d1(const Linear &L, const void *scales, int scales_dtype, __half *w, cudaStream_t stream);
int launch_convert_bias(const void *src, int dtype, __nv_bfloat16 *dst, int64_t n, cudaStream_t stream, int half = -1);
int launch_gather_x(const __nv_bfloat16 *x, const int32_t *perm, __nv_bfloat16 *xg, int64_t M, int64_t K,
                    cudaStream_t stream);
int launch_add_bias(__nv_bfloat16 *y, const __nv_bfloat16 *bias, int64_t M, int64_t N, cudaStream_t stream);

// decode: M <= 32 tokens per launch
int launch_decode(const Linear &L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias,
                  cudaStream_t stream);
const char *decode_kernel_name(const Linear &L, int64_t M);
int decode_scratch_need(const Linear &L, int sm_count, int64_t *scratch_floats, int64_t *tickets);
// decode, M <= 16: the single-wave stream kernel (w4a16_gemv.cu); the register-ring kernel above is its fallback
bool gemv_stream_usable(const Linear &L, int64_t M);
int gemv_build_table(Linear &L, int sm_count);
int64_t gemv_scratch_floats(int sm_count);
int launch_gemv_stream(const Linear &L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias,
                       cudaStream_t stream);
// decode, M <= 4: int8 tensor-core stream kernel (w4a16_gemv_imma.cu)
bool gemv_imma_usable(const Linear &L, int64_t M);
int launch_gemv_imma(const Linear &L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias,
                     cudaStream_t stream);
// decode, M <= 4, K % 4096 == 0: second-generation int8 tensor-core stream kernel (w4a16_gemv_imma2.cu)
bool gemv_imma2_usable(const Linear &L, int64_t M);
int launch_gemv_imma2(const Linear &L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias,
                      cudaStream_t stream);
struct PeerView;   // w4a16_gemv_common.cuh
bool gemv_imma_peer_usable(const Linear &L, int64_t M, const PeerView &peer);
int launch_gemv_imma_peer(const Linear &L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias,
                          cudaStream_t stream, const PeerView *peer);
int gemv_debug_read(unsigned long long *out, int64_t words, unsigned long long *launches);
// prefill (tcgen05): any M
int launch_prefill(const Linear &L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias,
                   cudaStream_t stream);
bool prefill_available();
// 11 <= M <= 48: mma.sync skinny GEMM, every warp a row block x k slice (w4a16_midm.cu).  launch_midm returns 2 when it
// cannot run this call (split-K scratch would have to grow under stream capture): the caller uses the prefill kernel.
bool midm_usable(const Linear &L, int64_t M);
int launch_midm(const Linear &L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias, cudaStream_t stream);
int splitk_scratch(int device, size_t need, cudaStream_t stream, float **part, int **tickets);   // per (device, stream)
constexpr int SPLITK_MAX_TILES = 1024;   // tickets per device

}  // namespace w4
