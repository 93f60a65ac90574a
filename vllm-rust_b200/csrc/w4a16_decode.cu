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
#include "w4a16_internal.cuh"
#include "w4a16_dequant.cuh"

namespace w4 {

constexpr int DEC_THREADS = 256;
constexpr int DEC_WARPS = 8;
constexpr int DEC_DEPTH = 4;  // chunks per register-ring half

struct DecodeParams {
    const uint4 *qw;
    const uint8_t *meta;
    const __nv_bfloat16 *bias;  // null: no bias add in this launch
    const __nv_bfloat16 *x;
    __nv_bfloat16 *y;
    float *scratch;    // [SC][M][N] fp32 partials, SC > 1 only
    int32_t *tickets;  // [gridDim.x]
    int M, K, N, KB, G;
    int chunks_per_group_shift;  // gs >= 64: log2(gs/64); per-channel: 31 (every chunk -> group 0)
    int wk_shift;                // log2(warps along K inside a CTA)
    int sc;                      // CTAs along K
    int cps;                     // chunks per slice
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
__device__ __forceinline__ void mma_bf16(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                         uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// One prefetched chunk: 4 words of packed nibbles + the scale pair / zero byte of the
// group(s) it belongs to.  GPC = groups per chunk (2 only for group_size 32).
template <int GPC>
struct Slot {
    uint4 w;
    uint32_t s2[GPC];
    uint32_t zb[GPC];
};

template <int GPC>
__device__ __forceinline__ void load_slot(Slot<GPC> &s, const DecodeParams &p, const uint4 *wrow,
                                          const uint8_t *mrow, int kb, int kb_end, int lane) {
    if (kb < kb_end) {
        s.w = ldg_stream(wrow + (size_t)kb * 32 + lane);
        const int g = lane >> 2;
#pragma unroll
        for (int h = 0; h < GPC; ++h) {
            int grp = GPC == 2 ? kb * 2 + h : (kb >> p.chunks_per_group_shift);
            const uint8_t *m = mrow + (size_t)grp * META_BYTES;
            s.s2[h] = ldg_stream_u32(m + 4 * g);
            s.zb[h] = ldg_stream_u8(m + META_ZERO_OFF + g);
        }
    } else {
        s.w = make_uint4(0, 0, 0, 0);
#pragma unroll
        for (int h = 0; h < GPC; ++h) { s.s2[h] = 0; s.zb[h] = 0; }
    }
}

template <int NT, int GPC>
__device__ __forceinline__ void compute_slot(const Slot<GPC> &s, const DecodeParams &p, int kb, int lane,
                                             float (&yacc)[NT][4]) {
    const int g = lane >> 2, t = lane & 3;
    // B fragments: token 8nt+g, k = 64kb + 16j + 4t .. +3  (one 8-byte load through L1)
    uint2 xb[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
        int tok = nt * 8 + g;
        const __nv_bfloat16 *xr = p.x + (size_t)tok * p.K + (size_t)kb * CHUNK_K + 4 * t;
#pragma unroll
        for (int j = 0; j < 4; ++j)
            xb[nt][j] = tok < p.M ? __ldg(reinterpret_cast<const uint2 *>(xr + 16 * j)) : make_uint2(0u, 0u);
    }
    const uint32_t wj[4] = {s.w.x, s.w.y, s.w.z, s.w.w};
#pragma unroll
    for (int h = 0; h < GPC; ++h) {
        const uint32_t zlo = zero_magic(s.zb[h] & 0xFu), zhi = zero_magic(s.zb[h] >> 4);
        float acc[NT][4];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[nt][i] = 0.f;
#pragma unroll
        for (int jj = 0; jj < 4 / GPC; ++jj) {
            const int j = h * (4 / GPC) + jj;
            const uint32_t w = wj[j];
            const uint32_t a0 = centered_pair(w, zlo);        // row g   , k-slots 4t, 4t+1
            const uint32_t a1 = centered_pair(w >> 4, zhi);   // row g+8 , k-slots 4t, 4t+1
            const uint32_t a2 = centered_pair(w >> 8, zlo);   // row g   , k-slots 4t+2, 4t+3
            const uint32_t a3 = centered_pair(w >> 12, zhi);  // row g+8 , k-slots 4t+2, 4t+3
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) mma_bf16(acc[nt], a0, a1, a2, a3, xb[nt][j].x, xb[nt][j].y);
        }
        const float slo = __uint_as_float(s.s2[h] << 16), shi = __uint_as_float(s.s2[h] & 0xFFFF0000u);
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            yacc[nt][0] = fmaf(acc[nt][0], slo, yacc[nt][0]);
            yacc[nt][1] = fmaf(acc[nt][1], slo, yacc[nt][1]);
            yacc[nt][2] = fmaf(acc[nt][2], shi, yacc[nt][2]);
            yacc[nt][3] = fmaf(acc[nt][3], shi, yacc[nt][3]);
        }
    }
}

// NT = token tiles of 8 (M <= 8*NT).  grid = (ceil(NR / WR), SC), WR = 8 >> wk_shift.
template <int NT, int GPC>
__global__ void __launch_bounds__(DEC_THREADS, NT >= 4 ? 1 : 2) decode_kernel(const DecodeParams p) {
    __shared__ float red[DEC_WARPS][NT][32][4];
    __shared__ int s_last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int WK = 1 << p.wk_shift, WR = DEC_WARPS >> p.wk_shift;
    const int wr = warp >> p.wk_shift, wkk = warp & (WK - 1);
    const int NR = p.N >> 4;
    const int R = blockIdx.x * WR + wr;
    const int slice = blockIdx.y * WK + wkk;
    const int kb0 = slice * p.cps;
    const int kb_end = R < NR ? min(p.KB, kb0 + p.cps) : kb0;  // out-of-range row block: empty slice

    const uint4 *wrow = p.qw + (size_t)(R < NR ? R : 0) * p.KB * 32;
    const uint8_t *mrow = p.meta + (size_t)(R < NR ? R : 0) * p.G * META_BYTES;

    float yacc[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int i = 0; i < 4; ++i) yacc[nt][i] = 0.f;

    Slot<GPC> cur[DEC_DEPTH], nxt[DEC_DEPTH];
#pragma unroll
    for (int d = 0; d < DEC_DEPTH; ++d) load_slot<GPC>(cur[d], p, wrow, mrow, kb0 + d, kb_end, lane);
#pragma unroll
    for (int d = 0; d < DEC_DEPTH; ++d) load_slot<GPC>(nxt[d], p, wrow, mrow, kb0 + DEC_DEPTH + d, kb_end, lane);

    // Weights are immutable: everything above is independent of the previous kernel in the
    // stream.  Let the next kernel start its own prefetch, then wait for our producer.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");

    for (int base = kb0; base < kb_end; base += 2 * DEC_DEPTH) {
#pragma unroll
        for (int d = 0; d < DEC_DEPTH; ++d)
            if (base + d < kb_end) compute_slot<NT, GPC>(cur[d], p, base + d, lane, yacc);
#pragma unroll
        for (int d = 0; d < DEC_DEPTH; ++d) load_slot<GPC>(cur[d], p, wrow, mrow, base + 2 * DEC_DEPTH + d, kb_end, lane);
#pragma unroll
        for (int d = 0; d < DEC_DEPTH; ++d)
            if (base + DEC_DEPTH + d < kb_end) compute_slot<NT, GPC>(nxt[d], p, base + DEC_DEPTH + d, lane, yacc);
#pragma unroll
        for (int d = 0; d < DEC_DEPTH; ++d) load_slot<GPC>(nxt[d], p, wrow, mrow, base + 3 * DEC_DEPTH + d, kb_end, lane);
    }

    // ---- reduce the K-split warps of this CTA (fixed order) -------------------------
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
        *reinterpret_cast<float4 *>(&red[warp][nt][lane][0]) = make_float4(yacc[nt][0], yacc[nt][1], yacc[nt][2], yacc[nt][3]);
    __syncthreads();

    // Output element e of the CTA tile: token-major, n fastest, so stores are contiguous.
    const int tile_n = WR * 16;
    const int n_out = tile_n * p.M;
    const int n_base = blockIdx.x * tile_n;
    auto tile_value = [&](int e, int &tok, int &n) -> float {
        tok = e / tile_n;
        int nl = e - tok * tile_n;
        n = n_base + nl;
        int wr_e = nl >> 4, r = nl & 15;
        int nt = tok >> 3, c = tok & 7;
        int ln = (r & 7) * 4 + (c >> 1), i = ((r >> 3) << 1) | (c & 1);
        float v = 0.f;
        for (int k = 0; k < WK; ++k) v += red[(wr_e << p.wk_shift) + k][nt][ln][i];
        return v;
    };

    if (p.sc == 1) {
        for (int e = threadIdx.x; e < n_out; e += DEC_THREADS) {
            int tok, n;
            float v = tile_value(e, tok, n);
            if (n < p.N) {
                if (p.bias) v += __bfloat162float(p.bias[n]);
                p.y[(size_t)tok * p.N + n] = __float2bfloat16_rn(v);
            }
        }
        return;
    }

    // ---- cross-CTA split-K: publish the partial, the last CTA of the column sums -----
    float *part = p.scratch + (size_t)blockIdx.y * p.M * p.N;
    for (int e = threadIdx.x; e < n_out; e += DEC_THREADS) {
        int tok, n;
        float v = tile_value(e, tok, n);
        if (n < p.N) part[(size_t)tok * p.N + n] = v;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        int prev = atomicAdd(&p.tickets[blockIdx.x], 1);
        s_last = prev == p.sc - 1;
        if (s_last) p.tickets[blockIdx.x] = 0;  // self-reset for the next launch
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    for (int e = threadIdx.x; e < n_out; e += DEC_THREADS) {
        int tok = e / tile_n, n = n_base + (e - tok * tile_n);
        if (n >= p.N) continue;
        float v = 0.f;
        for (int s = 0; s < p.sc; ++s) v += __ldcg(p.scratch + ((size_t)s * p.M + tok) * p.N + n);
        if (p.bias) v += __bfloat162float(p.bias[n]);
        p.y[(size_t)tok * p.N + n] = __float2bfloat16_rn(v);
    }
}

// ---- host side ---------------------------------------------------------------------
struct DecodeCfg { int wk_shift, sc, cps; };

static int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

// Heuristic: enough CTAs to cover the chip about twice, at least 4 chunks per warp,
// and as little K-splitting across CTAs as that allows.
static DecodeCfg pick_cfg(const Linear &L, int sm_count) {
    const int NR = (int)L.d.NR, KB = (int)L.d.KB;
    int wk_shift = 3;  // 8 warps along K, one row block per CTA
    // Long-N shapes: put 2 or 4 row blocks in a CTA so x is fetched less often.
    while (wk_shift > 1 && (NR >> (3 - wk_shift + 1)) >= 4 * sm_count && (KB >> (wk_shift - 1)) >= 8) --wk_shift;
    int WR = DEC_WARPS >> wk_shift, WK = 1 << wk_shift;
    int ctas_x = (NR + WR - 1) / WR;
    int sc = 1;
    while (ctas_x * sc < 2 * sm_count && (KB / (WK * sc * 2)) >= 4 && sc < 16) sc *= 2;
    int ov_wk = env_int("W4A16_DECODE_WK_SHIFT", -1), ov_sc = env_int("W4A16_DECODE_SC", -1);
    if (ov_wk >= 0 && ov_wk <= 3) wk_shift = ov_wk;
    if (ov_sc >= 1 && ov_sc <= 64) sc = ov_sc;
    WK = 1 << wk_shift;
    int slices = WK * sc;
    int cps = (KB + slices - 1) / slices;
    return {wk_shift, sc, cps};
}

int decode_scratch_need(const Linear &L, int sm_count, int64_t *scratch_floats, int64_t *tickets) {
    // worst case over the env overrides is bounded by sc <= 64
    DecodeCfg c = pick_cfg(L, sm_count);
    *scratch_floats = c.sc > 1 ? (int64_t)c.sc * 32 * L.d.N : 0;
    *tickets = L.d.NR;  // >= gridDim.x for any WR
    return 0;
}

template <int NT, int GPC>
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
    W4_CUDA_OK(cudaLaunchKernelEx(&cfg, decode_kernel<NT, GPC>, p));
    count_launch();
    return 0;
}

const char *decode_kernel_name(int64_t M) {
    return M <= 8 ? "decode_mma_nt1" : M <= 16 ? "decode_mma_nt2" : "decode_mma_nt4";
}

int launch_decode(const Linear &L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias,
                  cudaStream_t stream) {
    if (M < 1 || M > 32) return fail("launch_decode: M must be in 1..32");
    const int gs = L.d.gs;
    DecodeCfg c = pick_cfg(L, L.scratch->sm_count);
    DecodeParams p;
    p.qw = reinterpret_cast<const uint4 *>(L.qw);
    p.meta = L.meta;
    p.bias = add_bias && L.has_bias ? L.bias : nullptr;
    p.x = x; p.y = y;
    p.scratch = L.splitk_scratch; p.tickets = L.tickets;
    p.M = (int)M; p.K = (int)L.d.K; p.N = (int)L.d.N; p.KB = (int)L.d.KB; p.G = (int)L.d.G;
    p.chunks_per_group_shift = gs == 64 ? 0 : gs == 128 ? 1 : 31;  // gs == K (per-channel): one group
    if (gs != 32 && gs != 64 && gs != 128 && gs != L.d.K) return fail("launch_decode: unsupported group size");
    p.wk_shift = c.wk_shift; p.sc = c.sc; p.cps = c.cps;
    if (c.sc > 1 && (!L.splitk_scratch || (int64_t)c.sc * M * L.d.N > L.splitk_floats))
        return fail("launch_decode: split-K scratch too small for this configuration");
    const int WR = DEC_WARPS >> c.wk_shift;
    dim3 grid((unsigned)((L.d.NR + WR - 1) / WR), (unsigned)c.sc);
    const bool pdl = env_int("W4A16_NO_PDL", 0) == 0;
    const int gpc = gs == 32 ? 2 : 1;
    const int nt = M <= 8 ? 1 : M <= 16 ? 2 : 4;
    if (gpc == 1) {
        if (nt == 1) return launch_t<1, 1>(p, grid, stream, pdl);
        if (nt == 2) return launch_t<2, 1>(p, grid, stream, pdl);
        return launch_t<4, 1>(p, grid, stream, pdl);
    }
    if (nt == 1) return launch_t<1, 2>(p, grid, stream, pdl);
    if (nt == 2) return launch_t<2, 2>(p, grid, stream, pdl);
    return launch_t<4, 2>(p, grid, stream, pdl);
}

}  // namespace w4
