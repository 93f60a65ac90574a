// w4a16_decode.cu -- the memory-bound skinny GEMM for M <= 32 tokens (decode).
//
// Replaces awq_gemv_int4_kt_bf16 / kt_mloop_m{4,8,12,16} (awq_gemv.cu:377-665) and
// gptq_gemm_int4_bf16_tiled (gptq_dequant.cu:243-318) as dispatched by
// marlin_gemm_pooled / marlin_gemm (marlin_cuda.rs:843-953, 1655-1753).
//
// Arithmetic (the reference's GEMV form, awq_gemv.cu:597-664, regrouped):
//     y[m,n] = sum_g  s[g,n] * ( sum_{k in g} (q[k,n] - z[g,n]) * x[m,k] )      fp32
// (q - z) is an exact small integer in bf16, its product with a bf16 x is exact in fp32,
// the inner sum runs on mma.sync.m16n8k16 (weights = A: 16 output rows x 16 k, tokens = B)
// with fp32 accumulators, and the bf16 scale is applied once per 64-k chunk in fp32.
// W is read ONCE for all M tokens.
//
// Data movement: a warp streams a K-run of one 16-row block: each 512-byte chunk is one
// fully coalesced 128-bit load per lane (L1::no_allocate, evict-first) that IS the set of
// mma A fragments after a LOP3 -- no shared-memory staging, no transposition.  A register
// ring keeps 2*DEPTH chunks (4 KB) per warp in flight.  K is split over the warps of a CTA
// (shared-memory reduction in a fixed order) and, for short-N shapes, over CTAs
// (per-handle fp32 scratch + a ticket; the last CTA sums in split order: deterministic).
// Activations come through L1 (they are the only reused data).
// Programmatic dependent launch: the weight prefetch is issued BEFORE griddepcontrol.wait,
// so in a stream of linears the next layer's weights are already in flight while the
// previous kernel drains; x / y / scratch are only touched after the wait.
#include <algorithm>
#include <mutex>

#include "w4a16_internal.cuh"
#include "w4a16_dequant.cuh"

namespace w4 {

constexpr int DEC_THREADS = 256;
constexpr int DEC_WARPS = 8;
constexpr int DEC_DEPTH = 2;   // 128-k units per register-ring half (ring = 2 halves = 8 chunks = 4 KB per warp)
constexpr int UNIT_K = 128;    // the K granule of the decode kernel: 2 chunks; K % 128 == 0 by the shape gate

struct DecodeParams {
    const uint4 *qw;
    const uint8_t *meta;
    const __nv_bfloat16 *bias;  // null: no bias add in this launch
    const __nv_bfloat16 *x;
    __nv_bfloat16 *y;
    float4 *scratch;   // [SC][NR][NT][32] fragment-layout fp32 partials, SC > 1 only
    int32_t *tickets;  // [gridDim.x]
    int M, K, N, KB, G;
    int meta_unit_stride;  // bytes of meta per 128-k unit: MPU * 48, or 0 for per-channel scales
    int wk_shift;          // log2(warps along K inside a CTA)
    int sc;                // CTAs along K
    int ups;               // units per slice
    int swiglu;            // SwiGLU handle: rows g / g+8 are (gate, up), y is [M, N/2] (w4a16_epilogue.cuh)
};

__device__ __forceinline__ uint4 ldg_stream(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint32_t ldg_stream_u32(const void *p) {
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
__device__ __forceinline__ uint32_t ldg_stream_u8(const void *p) {
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u8 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
// D = A*B (+ C).  ZERO_C feeds literal zeros so ptxas uses RZ instead of clearing registers.
template <bool ZERO_C>
__device__ __forceinline__ void mma_bf16(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                         uint32_t b0, uint32_t b1) {
    if (ZERO_C) {
        asm volatile(
            "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
            "{%10,%10,%10,%10};"
            : "=f"(c[0]), "=f"(c[1]), "=f"(c[2]), "=f"(c[3])
            : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "f"(0.f));
    } else {
        asm volatile(
            "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
            "{%0,%1,%2,%3};"
            : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
            : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
}

// One prefetched 128-k unit of one 16-row block: 2 chunks (8 words per lane) + the scale pair /
// zero byte of each group inside it.  MPU = groups per unit: 1 (g=128 or per-channel), 2 (g=64), 4 (g=32).
template <int MPU>
struct Unit {
    uint4 w[2];
    uint32_t s2[MPU];
    uint32_t zb[MPU];
};

// wp: this lane's 16 bytes of the unit's first chunk.  ms / mz: this lane's scale pair / zero byte
// of the unit's first group.  All offsets below are compile-time immediates.
template <int MPU>
__device__ __forceinline__ void load_unit(Unit<MPU> &u, const uint4 *wp, const uint8_t *ms, const uint8_t *mz) {
    u.w[0] = ldg_stream(wp);
    u.w[1] = ldg_stream(wp + 32);
#pragma unroll
    for (int h = 0; h < MPU; ++h) {
        u.s2[h] = ldg_stream_u32(ms + h * META_BYTES);
        u.zb[h] = ldg_stream_u8(mz + h * META_BYTES);
    }
}

template <int NT, int MPU>
__device__ __forceinline__ void compute_unit(const Unit<MPU> &u, const __nv_bfloat16 *const (&xp)[NT],
                                             float (&yacc)[NT][4]) {
    // B fragments: token 8nt+g, k = 16*jm + 4t .. +3 of this unit (8-byte loads through L1)
    uint2 xb[NT][8];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int jm = 0; jm < 8; ++jm) xb[nt][jm] = __ldg(reinterpret_cast<const uint2 *>(xp[nt] + 16 * jm));
    const uint32_t wj[8] = {u.w[0].x, u.w[0].y, u.w[0].z, u.w[0].w, u.w[1].x, u.w[1].y, u.w[1].z, u.w[1].w};
    constexpr int MMAS_PER_GROUP = 8 / MPU;
#pragma unroll
    for (int h = 0; h < MPU; ++h) {
        const uint32_t zlo = zero_magic(u.zb[h] & 0xFu), zhi = zero_magic(u.zb[h] >> 4);
        float acc[NT][4];
#pragma unroll
        for (int jj = 0; jj < MMAS_PER_GROUP; ++jj) {
            const int jm = h * MMAS_PER_GROUP + jj;
            const uint32_t w = wj[jm];
            const uint32_t a0 = centered_pair(w, zlo);        // row g   , k-slots 4t, 4t+1
            const uint32_t a1 = centered_pair(w >> 4, zhi);   // row g+8 , k-slots 4t, 4t+1
            const uint32_t a2 = centered_pair(w >> 8, zlo);   // row g   , k-slots 4t+2, 4t+3
            const uint32_t a3 = centered_pair(w >> 12, zhi);  // row g+8 , k-slots 4t+2, 4t+3
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                if (jj == 0) mma_bf16<true>(acc[nt], a0, a1, a2, a3, xb[nt][jm].x, xb[nt][jm].y);
                else mma_bf16<false>(acc[nt], a0, a1, a2, a3, xb[nt][jm].x, xb[nt][jm].y);
            }
        }
        const float slo = __uint_as_float(u.s2[h] << 16), shi = __uint_as_float(u.s2[h] & 0xFFFF0000u);
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            yacc[nt][0] = fmaf(acc[nt][0], slo, yacc[nt][0]);
            yacc[nt][1] = fmaf(acc[nt][1], slo, yacc[nt][1]);
            yacc[nt][2] = fmaf(acc[nt][2], shi, yacc[nt][2]);
            yacc[nt][3] = fmaf(acc[nt][3], shi, yacc[nt][3]);
        }
    }
}

// Fragment -> y: lane (g,t) holds rows 16R+g and 16R+g+8 for tokens 8nt+2t and 8nt+2t+1.
template <int NT>
__device__ __forceinline__ void store_fragment(const DecodeParams &p, int R, int lane, const float4 (&v)[NT]) {
    const int g = lane >> 2, t = lane & 3;
    const int n0 = R * 16 + g;
    const float b0 = p.bias ? __bfloat162float(p.bias[n0]) : 0.f;
    const float b1 = p.bias ? __bfloat162float(p.bias[n0 + 8]) : 0.f;
    if (p.swiglu) {
        const size_t I = (size_t)(p.N >> 1), c = (size_t)R * 8 + g;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int tok = nt * 8 + 2 * t;
            if (tok < p.M) p.y[(size_t)tok * I + c] = swiglu_bf16(v[nt].x + b0, v[nt].z + b1);
            if (tok + 1 < p.M) p.y[(size_t)(tok + 1) * I + c] = swiglu_bf16(v[nt].y + b0, v[nt].w + b1);
        }
        return;
    }
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
        const int tok = nt * 8 + 2 * t;
        if (tok < p.M) {
            p.y[(size_t)tok * p.N + n0] = __float2bfloat16_rn(v[nt].x + b0);
            p.y[(size_t)tok * p.N + n0 + 8] = __float2bfloat16_rn(v[nt].z + b1);
        }
        if (tok + 1 < p.M) {
            p.y[(size_t)(tok + 1) * p.N + n0] = __float2bfloat16_rn(v[nt].y + b0);
            p.y[(size_t)(tok + 1) * p.N + n0 + 8] = __float2bfloat16_rn(v[nt].w + b1);
        }
    }
}

// NT = token tiles of 8 (M <= 8*NT).  grid = (NR / WR, SC), WR = 8 >> wk_shift.
template <int NT, int MPU>
__global__ void __launch_bounds__(DEC_THREADS, NT >= 4 ? 1 : 2) decode_kernel(const DecodeParams p) {
    __shared__ float4 red[DEC_WARPS][NT][32];
    __shared__ int s_last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int WK = 1 << p.wk_shift;
    const int wr = warp >> p.wk_shift, wkk = warp & (WK - 1);
    const int R = blockIdx.x * (DEC_WARPS >> p.wk_shift) + wr;
    const int slice = blockIdx.y * WK + wkk;
    const int u0 = slice * p.ups;
    const int n_units = min(p.K / UNIT_K, u0 + p.ups) - u0;  // may be <= 0 for a trailing slice

    const uint4 *wp = p.qw + ((size_t)R * p.KB + 2 * (size_t)u0) * 32 + lane;
    const uint8_t *ms = p.meta + (size_t)R * p.G * META_BYTES + (size_t)u0 * p.meta_unit_stride + 4 * g;
    const uint8_t *mz = ms + (META_ZERO_OFF - 3 * g);
    const __nv_bfloat16 *xp[NT];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)  // token rows beyond M alias row M-1: their columns are never stored
        xp[nt] = p.x + (size_t)min(nt * 8 + g, p.M - 1) * p.K + (size_t)u0 * UNIT_K + 4 * t;
    const size_t mstride = (size_t)p.meta_unit_stride;

    float yacc[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int i = 0; i < 4; ++i) yacc[nt][i] = 0.f;

    Unit<MPU> ring[2 * DEC_DEPTH];
#pragma unroll
    for (int d = 0; d < 2 * DEC_DEPTH; ++d)
        if (d < n_units) load_unit<MPU>(ring[d], wp + d * 64, ms + d * mstride, mz + d * mstride);

    // Weights are immutable: everything above is independent of the previous kernel in the
    // stream.  Let the next kernel start its own prefetch, then wait for our producer.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");

    for (int base = 0; base < n_units; base += 2 * DEC_DEPTH) {
        const int left = n_units - base;  // units not yet computed, > 0
#pragma unroll
        for (int hlf = 0; hlf < 2; ++hlf) {
#pragma unroll
            for (int d = 0; d < DEC_DEPTH; ++d) {
                const int i = hlf * DEC_DEPTH + d;
                if (i < left) {
                    const __nv_bfloat16 *xq[NT];
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt) xq[nt] = xp[nt] + i * UNIT_K;
                    compute_unit<NT, MPU>(ring[i], xq, yacc);
                }
            }
#pragma unroll
            for (int d = 0; d < DEC_DEPTH; ++d) {  // refill the half just consumed, one ring turn ahead
                const int i = hlf * DEC_DEPTH + d + 2 * DEC_DEPTH;
                if (i < left) load_unit<MPU>(ring[i - 2 * DEC_DEPTH], wp + i * 64, ms + i * mstride, mz + i * mstride);
            }
        }
        wp += 2 * DEC_DEPTH * 64;
        ms += 2 * DEC_DEPTH * mstride;
        mz += 2 * DEC_DEPTH * mstride;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) xp[nt] += 2 * DEC_DEPTH * UNIT_K;
    }

    // ---- reduce the K-split warps of this CTA: warp (wr, 0) sums its row block in slice order ----
    if (WK > 1) {
        if (wkk != 0) {
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
                red[warp][nt][lane] = make_float4(yacc[nt][0], yacc[nt][1], yacc[nt][2], yacc[nt][3]);
        }
        __syncthreads();
        if (wkk != 0 && p.sc == 1) return;
    }
    float4 v[NT];
    if (wkk == 0) {
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            v[nt] = make_float4(yacc[nt][0], yacc[nt][1], yacc[nt][2], yacc[nt][3]);
            for (int k = 1; k < WK; ++k) {
                const float4 o = red[warp + k][nt][lane];
                v[nt].x += o.x; v[nt].y += o.y; v[nt].z += o.z; v[nt].w += o.w;
            }
        }
        if (p.sc == 1) {
            store_fragment<NT>(p, R, lane, v);
            return;
        }
        // cross-CTA split-K: publish this CTA's partial fragment
        float4 *part = p.scratch + (((size_t)blockIdx.y * (p.N >> 4) + R) * NT) * 32 + lane;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) __stcg(part + nt * 32, v[nt]);
    }
    // ---- the last CTA of this column (ticket) sums the SC partials in slice order: deterministic ----
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const int prev = atomicAdd(&p.tickets[blockIdx.x], 1);
        s_last = prev == p.sc - 1;
        if (s_last) p.tickets[blockIdx.x] = 0;  // self-reset for the next launch
    }
    __syncthreads();
    if (!s_last || wkk != 0) return;
    __threadfence();
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
        v[nt] = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int s = 0; s < p.sc; ++s) {
            const float4 o = __ldcg(p.scratch + (((size_t)s * (p.N >> 4) + R) * NT + nt) * 32 + lane);
            v[nt].x += o.x; v[nt].y += o.y; v[nt].z += o.z; v[nt].w += o.w;
        }
    }
    store_fragment<NT>(p, R, lane, v);
}

// ---- host side ---------------------------------------------------------------------
struct DecodeCfg { int wk_shift, sc, ups; };

static int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

// Heuristic: enough CTAs to cover the chip about twice, at least 4 chunks per warp,
// and as little K-splitting across CTAs as that allows.
static DecodeCfg pick_cfg(const Linear &L, int sm_count) {
    const int NR = (int)L.d.NR, U = (int)(L.d.K / UNIT_K);
    int wk_shift = 3;  // 8 warps along K, one row block per CTA
    // Long-N shapes: put 2 or 4 row blocks in a CTA (NR is a multiple of 4 by the N % 64 gate).
    while (wk_shift > 1 && (NR >> (3 - wk_shift + 1)) >= 4 * sm_count && (U >> (wk_shift - 1)) >= 4) --wk_shift;
    int WR = DEC_WARPS >> wk_shift, WK = 1 << wk_shift;
    int ctas_x = NR / WR;
    int sc = 1;
    while (ctas_x * sc < 2 * sm_count && (U / (WK * sc * 2)) >= 2 && sc < 16) sc *= 2;
    int ov_wk = exp_env_int("W4A16_DECODE_WK_SHIFT", -1), ov_sc = exp_env_int("W4A16_DECODE_SC", -1);
    if (ov_wk >= 1 && ov_wk <= 3) wk_shift = ov_wk;
    if (ov_sc >= 1 && ov_sc <= 64) sc = ov_sc;
    WK = 1 << wk_shift;
    int slices = WK * sc;
    int ups = (U + slices - 1) / slices;
    return {wk_shift, sc, ups};
}

int decode_scratch_need(const Linear &L, int sm_count, int64_t *scratch_floats, int64_t *tickets) {
    // worst case over the env overrides is bounded by sc <= 64
    DecodeCfg c = pick_cfg(L, sm_count);
    *scratch_floats = c.sc > 1 ? (int64_t)c.sc * L.d.NR * 4 * 32 * 4 : 0;  // [SC][NR][NT<=4][32] float4
    *scratch_floats = std::max<int64_t>(*scratch_floats, gemv_scratch_floats(sm_count));
    *tickets = L.d.NR;  // >= gridDim.x for any WR
    return 0;
}

template <int NT, int MPU>
static int launch_t(const DecodeParams &p, dim3 grid, cudaStream_t stream, bool pdl) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(DEC_THREADS);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    W4_CUDA_OK(cudaLaunchKernelEx(&cfg, decode_kernel<NT, MPU>, p));
    count_launch();
    return 0;
}

const char *decode_kernel_name(const Linear &L, int64_t M) {
    if (gemv_imma2_usable(L, M)) return "gemv_stream_imma2";
    if (gemv_imma_usable(L, M)) return "gemv_stream_imma";
    if (M <= 4 && gemv_stream_usable(L, M)) return M <= 2 ? "gemv_stream_dp4a" : "gemv_stream_nt1";
    return M <= 8 ? "decode_mma_nt1" : M <= 16 ? "decode_mma_nt2" : "decode_mma_nt4";
}

int launch_decode(const Linear &L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias,
                  cudaStream_t stream) {
    if (M < 1 || M > 32) return fail("launch_decode: M must be in 1..32");
    if (gemv_imma2_usable(L, M)) return launch_gemv_imma2(L, x, y, M, add_bias, stream);
    if (gemv_imma_usable(L, M)) return launch_gemv_imma(L, x, y, M, add_bias, stream);
    // measured: above 4 tokens the register-ring kernel below beats the stream kernel (its per-token staging dominates)
    if (M <= 4 && gemv_stream_usable(L, M)) return launch_gemv_stream(L, x, y, M, add_bias, stream);
    const int gs = L.d.gs;
    DecodeCfg c = pick_cfg(L, L.scratch->sm_count);
    DecodeParams p;
    p.qw = reinterpret_cast<const uint4 *>(L.qw);
    p.meta = L.meta;
    p.bias = add_bias && L.has_bias ? L.bias : nullptr;
    p.x = x; p.y = y;
    p.scratch = reinterpret_cast<float4 *>(L.splitk_scratch); p.tickets = L.tickets;
    p.M = (int)M; p.K = (int)L.d.K; p.N = (int)L.d.N; p.KB = (int)L.d.KB; p.G = (int)L.d.G;
    p.swiglu = L.epilogue == W4A16_EPILOGUE_SWIGLU;
    if (gs != 32 && gs != 64 && gs != 128 && gs != L.d.K) return fail("launch_decode: unsupported group size");
    const int mpu = gs == 32 ? 4 : gs == 64 ? 2 : 1;
    p.meta_unit_stride = gs == L.d.K && gs != 128 ? 0 : mpu * META_BYTES;  // per-channel: one group for all of K
    p.wk_shift = c.wk_shift; p.sc = c.sc; p.ups = c.ups;
    const int nt = M <= 8 ? 1 : M <= 16 ? 2 : 4;
    if (c.sc > 1 && (!L.splitk_scratch || (int64_t)c.sc * L.d.NR * nt * 32 * 4 > L.splitk_floats))
        return fail("launch_decode: split-K scratch too small for this configuration");
    const int WR = DEC_WARPS >> c.wk_shift;
    dim3 grid((unsigned)(L.d.NR / WR), (unsigned)c.sc);
    const bool pdl = env_int("W4A16_NO_PDL", 0) == 0;
#define W4_DISPATCH(MPU_)                                               \
    do {                                                                \
        if (nt == 1) return launch_t<1, MPU_>(p, grid, stream, pdl);    \
        if (nt == 2) return launch_t<2, MPU_>(p, grid, stream, pdl);    \
        return launch_t<4, MPU_>(p, grid, stream, pdl);                 \
    } while (0)
    if (mpu == 1) W4_DISPATCH(1);
    if (mpu == 2) W4_DISPATCH(2);
    W4_DISPATCH(4);
#undef W4_DISPATCH
}

// ---- w4a16_forward_decode_static: the same kernel with its split-K scratch in a static device array ----------------
// (no counterpart in the reference, see include/w4a16_b200.h)
// One array per GPU that loads the library, held until exit: [0, 16 KiB) the per-column arrival tickets, then 8 MiB of
// fp32 partial fragments.  Every call on a device shares it.
constexpr size_t NA_TICKET_BYTES = 16 * 1024;          // 4096 tickets: N <= 65536
constexpr size_t NA_PART_BYTES = 8u * 1024 * 1024;     // sc * (N / 16) * token tiles * 512 B of partials
static __device__ __align__(256) unsigned char g_static_scratch[NA_TICKET_BYTES + NA_PART_BYTES];

// In host code g_static_scratch is only a stand-in for the symbol: the device address comes from the runtime, once per
// device, with that device current.
static unsigned char *g_static_ptr[64] = {};
static std::mutex g_static_mu;
static int static_scratch_ptr(int device, unsigned char **out) {
    if (device < 0 || device >= 64) return fail("forward_decode_static: device index out of range");
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_ptr[device]) {
        int cur = -1;
        W4_CUDA_OK(cudaGetDevice(&cur));
        if (cur != device) return fail("forward_decode_static: the handle's device is not current");
        void *ptr = nullptr;
        W4_CUDA_OK(cudaGetSymbolAddress(&ptr, g_static_scratch));
        if (!ptr) return fail("forward_decode_static: cudaGetSymbolAddress returned a null address");
        g_static_ptr[device] = static_cast<unsigned char *>(ptr);
    }
    *out = g_static_ptr[device];
    return 0;
}

int launch_decode_static(const Linear &L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias,
                         cudaStream_t stream) {
    if (M < 1 || M > 32) return fail("forward_decode_static: M must be in 1..32, got " + std::to_string(M));
    const int gs = L.d.gs;
    if (gs != 32 && gs != 64 && gs != 128 && gs != L.d.K) return fail("forward_decode_static: unsupported group size");
    const DecodeCfg c = pick_cfg(L, L.scratch->sm_count);
    const int nt = M <= 8 ? 1 : M <= 16 ? 2 : 4;
    const int WR = DEC_WARPS >> c.wk_shift;
    const dim3 grid((unsigned)(L.d.NR / WR), (unsigned)c.sc);
    const size_t part_need = c.sc > 1 ? (size_t)c.sc * L.d.NR * nt * 32 * sizeof(float4) : 0;
    if (part_need > NA_PART_BYTES)
        return fail("forward_decode_static: this shape needs " + std::to_string(part_need) + " bytes of split-K scratch, the static array holds " +
                    std::to_string(NA_PART_BYTES) + " (split " + std::to_string(c.sc) + " x " + std::to_string(L.d.NR) +
                    " row blocks x " + std::to_string(nt) + " token tiles x 512 B)");
    if ((size_t)grid.x * sizeof(int32_t) > NA_TICKET_BYTES)
        return fail("forward_decode_static: out_features " + std::to_string(L.d.N) + " exceeds the static ticket table (65536 columns)");
    unsigned char *base = nullptr;
    if (static_scratch_ptr(L.device, &base)) return 1;
    DecodeParams p;
    p.qw = reinterpret_cast<const uint4 *>(L.qw);
    p.meta = L.meta;
    p.bias = add_bias && L.has_bias ? L.bias : nullptr;
    p.x = x; p.y = y;
    p.tickets = reinterpret_cast<int32_t *>(base);
    p.scratch = reinterpret_cast<float4 *>(base + NA_TICKET_BYTES);
    p.M = (int)M; p.K = (int)L.d.K; p.N = (int)L.d.N; p.KB = (int)L.d.KB; p.G = (int)L.d.G;
    p.swiglu = L.epilogue == W4A16_EPILOGUE_SWIGLU;
    const int mpu = gs == 32 ? 4 : gs == 64 ? 2 : 1;
    p.meta_unit_stride = gs == L.d.K && gs != 128 ? 0 : mpu * META_BYTES;
    p.wk_shift = c.wk_shift; p.sc = c.sc; p.ups = c.ups;
    // The array keeps the last call's contents.  Partials are fully rewritten before they are read; the tickets must
    // start at zero (they reset themselves, but a launch that died half-way would leave them counting).
    if (c.sc > 1) W4_CUDA_OK(cudaMemsetAsync(p.tickets, 0, (size_t)grid.x * sizeof(int32_t), stream));
#define W4_DISPATCH(MPU_)                                                 \
    do {                                                                  \
        if (nt == 1) return launch_t<1, MPU_>(p, grid, stream, false);    \
        if (nt == 2) return launch_t<2, MPU_>(p, grid, stream, false);    \
        return launch_t<4, MPU_>(p, grid, stream, false);                 \
    } while (0)
    if (mpu == 1) W4_DISPATCH(1);
    if (mpu == 2) W4_DISPATCH(2);
    W4_DISPATCH(4);
#undef W4_DISPATCH
}

}  // namespace w4

extern "C" int w4a16_forward_decode_static(w4a16_handle_t h, const void *x_bf16, void *y_bf16, int64_t M, w4a16_stream_t stream) {
    using namespace w4;
    const Linear *L = reinterpret_cast<const Linear *>(h);
    if (!L) return fail("forward_decode_static: null handle");
    if (!L->loaded) return fail("forward_decode_static: weights not loaded (call load_weights first)");
    if (!x_bf16 || !y_bf16) return fail("forward_decode_static: null x or y");
    if ((reinterpret_cast<uintptr_t>(x_bf16) & 15) || (reinterpret_cast<uintptr_t>(y_bf16) & 15))
        return fail("forward_decode_static: x and y must be 16-byte aligned");
    if (L->perm && !L->perm_external)
        return fail("forward_decode_static: act-order handles gather x into an allocated buffer; use w4a16_forward_into");
    int prev = -1;
    W4_CUDA_OK(cudaGetDevice(&prev));
    if (prev != L->device) W4_CUDA_OK(cudaSetDevice(L->device));
    const bool add_bias = L->tp_mode != W4A16_TP_ROW || L->tp_size == 1;
    const int rc = launch_decode_static(*L, reinterpret_cast<const __nv_bfloat16 *>(x_bf16), reinterpret_cast<__nv_bfloat16 *>(y_bf16), M,
                                        add_bias, reinterpret_cast<cudaStream_t>(stream));
    if (prev != L->device) cudaSetDevice(prev);
    return rc;
}
