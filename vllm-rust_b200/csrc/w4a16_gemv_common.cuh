// w4a16_gemv_common.cuh -- parameters and PTX helpers shared by the decode stream kernels
// (w4a16_gemv.cu: dp4a / mma.sync f16 path; w4a16_gemv_imma.cu: int8 tensor-core path for M <= 4).
#pragma once
#include "w4a16_epilogue.cuh"
#include <stdlib.h>

#include "w4a16_internal.cuh"

namespace w4 {

constexpr int GV_THREADS = 512;
constexpr int GV_WARPS = 16;
constexpr int GV_UNIT_K = 128;        // 2 chunks = 1 KB of weights
constexpr int GV_UNIT_BYTES = 1024;
constexpr int GV_STAGE_UNITS = GV_WARPS;
constexpr int GV_STAGE_W_BYTES = GV_STAGE_UNITS * GV_UNIT_BYTES;
constexpr int GV_MAX_STAGES = 12;
constexpr int GV_SMEM_HALF_SM = 112 * 1024;   // two CTAs (this linear + the prefetching next one) per SM
constexpr int GV_SMEM_FULL_SM = 220 * 1024;
constexpr int GV_SMEM_2CTA = (228 * 1024 - 2 * 1024) / 2;   // dynamic bytes per CTA that still let two CTAs share an SM
constexpr uint32_t GV_SPIN_LIMIT = 1u << 26;

// One rank's view of the symmetric buffers of a tensor-parallel group (w4a16_peer_*, w4a16_api.cu).
// Every rank owns one allocation  data[3 rotating buffers][world sources][cap fp32 words], every word PEER_EMPTY
// until a source stores into it, and holds the (IPC-opened) base pointer of every peer's allocation.
constexpr int PEER_MAX_WORLD = 8;
constexpr int PEER_TOK = 4;                    // token slots per output row in a peer buffer (the fused path takes M <= 4)
constexpr uint32_t PEER_EMPTY = 0x80000000u;   // -0.0f: never stored as data (canonicalised to +0.0f)
struct PeerView {
    float *base[PEER_MAX_WORLD];   // base[r] = rank r's allocation as mapped into this process (base[rank] = own)
    int world, rank;
    uint32_t *epoch_ptr;           // device-resident launch counter of this rank (same value on every rank: SPMD call
                                   // sequence); a launch uses *epoch_ptr + 1 and its last CTA stores it back, so the
                                   // kernel is CUDA-graph capturable (nothing launch-specific in the parameters)
    int *ticket;                   // CTAs of the current launch that have finished (self-resetting)
    uint32_t cap;                  // fp32 words per (buffer, source)
    uint32_t flags;                // bit 0: release programmatic dependents early (experiment, W4A16_FUSED_EARLY_PDL)
};

struct GemvParams {
    const uint8_t *qw;
    const uint8_t *meta;
    const __nv_bfloat16 *bias;  // null: no bias add in this launch
    const __nv_bfloat16 *x;
    __nv_bfloat16 *y;
    float4 *scratch;            // [grid][2][P][8] boundary partials
    int32_t *tickets;           // [NR]
    const int4 *tab;            // [grid][2]: {a, b, first_lo, first_hi}, {last_lo, last_hi, -, -}
    int M, K, N;
    int swiglu;                 // SwiGLU handle: rows g / g+8 are (gate, up); y is [M, N/2] (w4a16_epilogue.cuh)
    int UPR;                    // units per row block = K / 128
    int meta_stride;            // bytes of meta per unit (MPU*48), 0 for per-channel scales
    int P;                      // ceil(M/2): fragment column pairs that hold real tokens
    int fast;                   // whole row blocks per CTA, full ring stages, grouped meta: the streamlined loop applies
    int n_stages_ring;          // stages in the shared-memory ring
    int stage_m_bytes;          // meta bytes per ring stage (16-byte multiple)
    // shared-memory carve-up (bytes from the start of dynamic smem)
    int off_xs, off_gtab, off_flags, off_part, off_wring, off_mring;
    unsigned long long *dbg;    // W4A16_GV_DEBUG: per-CTA globaltimer stamps {start, after wait, staged, streamed, combined, end}
    // fused row-parallel all-reduce over NVLink peer memory (w4a16_gemv_imma.cu epilogue); world <= 1: off
    PeerView peer;
};

__device__ __forceinline__ uint32_t gv_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void gv_mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void gv_mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ uint32_t gv_mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok;
}
// A wait that cannot hang the GPU: a lost completion traps (the launch fails loudly).
__device__ __forceinline__ void gv_mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t spins = 0;
    while (!gv_mbar_try_wait(bar, parity))
        if (++spins > GV_SPIN_LIMIT) __trap();
}
__device__ __forceinline__ void gv_bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ uint32_t gv_lop3_and_or(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
template <bool ZERO_C>
__device__ __forceinline__ void gv_mma_f16(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                           uint32_t b0, uint32_t b1) {
    if (ZERO_C) {
        asm volatile(
            "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
            "{%10,%10,%10,%10};"
            : "=f"(c[0]), "=f"(c[1]), "=f"(c[2]), "=f"(c[3])
            : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "f"(0.f));
    } else {
        asm volatile(
            "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
            "{%0,%1,%2,%3};"
            : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
            : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
}

__device__ __forceinline__ int gv_dp4a_u8s8(uint32_t a_u8x4, uint32_t b_s8x4, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a_u8x4), "r"(b_s8x4), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t gv_pack4(int b0, int b1, int b2, int b3) {
    return (uint32_t)(b0 & 0xFF) | ((uint32_t)(b1 & 0xFF) << 8) | ((uint32_t)(b2 & 0xFF) << 16) | ((uint32_t)(b3 & 0xFF) << 24);
}

// Reductions over an aligned group of GL lanes (activation staging: the lanes that hold one quantisation group).
// 16-lane groups use two FULL-warp redux.sync over inputs masked to one half each (a partial member mask sends
// redux.sync down a slow divergent path; the shuffle tree costs 4 dependent SHFL + op per reduction).
template <int GL>
__device__ __forceinline__ uint32_t gv_group_max(uint32_t v, int lane) {
    if constexpr (GL == 16) {
        uint32_t lo, hi;
        asm volatile("redux.sync.max.u32 %0, %1, 0xffffffff;" : "=r"(lo) : "r"(lane < 16 ? v : 0u));
        asm volatile("redux.sync.max.u32 %0, %1, 0xffffffff;" : "=r"(hi) : "r"(lane < 16 ? 0u : v));
        return lane < 16 ? lo : hi;
    } else {
#pragma unroll
        for (int o = GL / 2; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
        return v;
    }
}
template <int GL>
__device__ __forceinline__ int gv_group_add(int v, int lane) {
    if constexpr (GL == 16) {
        int lo, hi;
        asm volatile("redux.sync.add.s32 %0, %1, 0xffffffff;" : "=r"(lo) : "r"(lane < 16 ? v : 0));
        asm volatile("redux.sync.add.s32 %0, %1, 0xffffffff;" : "=r"(hi) : "r"(lane < 16 ? 0 : v));
        return lane < 16 ? lo : hi;
    } else {
#pragma unroll
        for (int o = GL / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        return v;
    }
}

inline int gv_env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}
unsigned long long *gemv_debug_slot();   // W4A16_GV_DEBUG stamps (w4a16_gemv.cu)

}  // namespace w4
