// w4a16_gemv.cu -- the decode GEMV / skinny GEMM for M <= 16 tokens: an evenly partitioned,
// single-wave stream over the packed weights.
//
// Replaces awq_gemv_int4_kt_bf16 / kt_mloop_m{4,8,12,16} (awq_gemv.cu:377-665) and
// gptq_gemm_int4_bf16_tiled (gptq_dequant.cu:243-318) as dispatched by marlin_gemm_pooled /
// marlin_gemm (marlin_cuda.rs:843-953, 1655-1753).
//
// Arithmetic (the reference's GEMV form, awq_gemv.cu:597-664, regrouped so that the zero point
// leaves the inner loop):
//     y[m,n] = sum_g s[g,n] * 2^e[g,m] * ( sum_{k in g} (c + q[k,n]) * xt[m,k]  -  (c + z[g,n]) * X[g,m] )
//   xt[m,k] = x[m,k] * 2^-e[g,m]  held as fp16 (exact: a bf16 mantissa fits, and the per-group, per-token
//             power of two puts max|x| of the group in [2^13, 2^14), so nothing overflows; values more than
//             2^-27 below the group's maximum start to lose low bits, far below the GEMM tolerance)
//   X[g,m]  = sum_{k in g} xt[m,k]   (fp32, computed once per CTA while the first weights are in flight)
//   c       = 1024 for rows g of a 16-row block, 64 for rows g+8: the fp16 "magic" exponents that turn a
//             nibble sitting at mantissa bits [0:3] / [4:7] into the exact number c + q with ONE LOP3.
// The inner sum runs on mma.sync.m16n8k16 (f16 x f16 -> f32): weights are the A operand straight from the
// W4T chunk (one 16-byte load per lane = four MMAs' fragments after 1 SHF + 4 LOP3 per word), tokens the
// B operand from shared memory.  5 ALU + 1 LDS + 1 HMMA issue slots per 8 weights x 8 tokens.
//
// Work distribution: the W4T weight array is ONE linear stream of 1 KB units (16 rows x 128 k).  It is
// cut into gridDim.x equal contiguous ranges (+-1 unit), one CTA each, grid = 2 x SM count = one wave
// with half of every SM's registers left free, so that under programmatic dependent launch the NEXT
// linear's CTAs are already resident and have their first weights in flight while this one drains.
// Inside a CTA the 8 warps take units round-robin; each keeps GV_DEPTH units (4 KB) in flight in
// registers.  Partial sums of a 16-row block are combined in a fixed order: across the warps of a CTA in
// shared memory, across the (at most two, for ranges >= one row block) CTAs that share a row block through
// a per-handle scratch and a self-resetting ticket.  No atomics on data: results are deterministic.
#include "w4a16_internal.cuh"
#include "w4a16_dequant.cuh"

namespace w4 {

constexpr int GV_THREADS = 256;
constexpr int GV_WARPS = 8;
constexpr int GV_DEPTH = 4;      // units a warp keeps in flight
constexpr int GV_UNIT_K = 128;   // 2 chunks = 1 KB of weights
constexpr int GV_UNIT_V4 = 64;   // uint4 per unit

struct GemvParams {
    const uint4 *qw;
    const uint8_t *meta;
    const __nv_bfloat16 *bias;  // null: no bias add in this launch
    const __nv_bfloat16 *x;
    __nv_bfloat16 *y;
    float4 *scratch;            // [grid][2][P][8] boundary partials
    int32_t *tickets;           // [NR]
    int M, K, N;
    int UPR;                    // units per row block = K / 128
    int UT;                     // total units
    int q, r;                   // CTA i owns units [i*q + min(i,r), +q + (i<r))
    int meta_stride;            // bytes of meta per unit (MPU*48), 0 for per-channel scales
    int G;                      // groups per row block (meta row stride = G*48)
    int P;                      // ceil(M/2): fragment column pairs that hold real tokens
    int n_slots;                // row blocks a CTA range can touch
    // shared-memory carve-up (bytes from the start of dynamic smem)
    int off_xsum, off_tok, off_part;
    unsigned long long *dbg;    // W4A16_GV_DEBUG: per-CTA globaltimer stamps {start, after wait, staged, streamed, combined, end}
};

__device__ __forceinline__ uint4 gv_ldg_stream(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint32_t gv_ldg_u32(const void *p) {
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
__device__ __forceinline__ uint32_t gv_ldg_u8(const void *p) {
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u8 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
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

// One prefetched unit: 2 chunks (8 words per lane) + scale pair / zero byte of each group inside it.
template <int MPU>
struct GvUnit {
    uint4 w[2];
    uint32_t s2[MPU];
    uint32_t zb[MPU];
};

// ms: this lane's scale pair of the unit's first group (meta block + 4g); zoff = META_ZERO_OFF - 3g reaches its zero byte.
template <int MPU>
__device__ __forceinline__ void gv_load_unit(GvUnit<MPU> &u, const uint4 *wp, const uint8_t *ms, int zoff) {
    u.w[0] = gv_ldg_stream(wp);
    u.w[1] = gv_ldg_stream(wp + 32);
#pragma unroll
    for (int h = 0; h < MPU; ++h) {
        u.s2[h] = gv_ldg_u32(ms + h * META_BYTES);
        u.zb[h] = gv_ldg_u8(ms + h * META_BYTES + zoff);
    }
}

// MR = token rows held in shared memory (M rounded up to 1, 2, 4, 8 or 16).  NT = MMA token tiles.
template <int MR, int MPU>
__global__ void __launch_bounds__(GV_THREADS, MR <= 8 ? 3 : 2) gemv_stream_kernel(const GemvParams p) {
    constexpr int NT = MR <= 8 ? 1 : MR / 8;
    constexpr int PAIRS = MR < 2 ? 1 : MR / 2;     // token pairs per row of the group table
    constexpr int MMAS_PER_GROUP = 8 / MPU;
    constexpr int SUB_K = GV_UNIT_K / MPU;         // k per (sub)group inside a unit
    extern __shared__ __align__(16) uint8_t gv_smem[];
    uint2 *xs = reinterpret_cast<uint2 *>(gv_smem);                          // [K/4][MR] 4 x fp16
    float *gtab = reinterpret_cast<float *>(gv_smem + p.off_xsum);           // [K/SUB_K][PAIRS] {X0, X1, 2^e0, 2^e1}
    int *flags = reinterpret_cast<int *>(gv_smem + p.off_tok);               // [2]
    float4 *part = reinterpret_cast<float4 *>(gv_smem + p.off_part);         // [n_slots][GV_WARPS][P][8]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int cta = blockIdx.x;
    const int a = cta * p.q + min(cta, p.r);
    const int n_cta = p.q + (cta < p.r ? 1 : 0);
    const int b = a + n_cta;
    const int n_w = warp < n_cta ? (n_cta - warp + GV_WARPS - 1) / GV_WARPS : 0;  // units of this warp
    const int R_first = a / p.UPR;

    // ---- prefetch (weights and meta are immutable: independent of the previous kernel) ----
    const int u0 = a + warp;
    int R_l = u0 / p.UPR, ku_l = u0 - R_l * p.UPR;   // load cursor (per-channel meta needs R)
    const uint4 *wp = p.qw + (size_t)u0 * GV_UNIT_V4 + lane;
    const uint8_t *meta_g = p.meta + 4 * g;
    const int zoff = META_ZERO_OFF - 3 * g;
    auto meta_ptr = [&](int R, int ku) -> const uint8_t * {
        return meta_g + ((size_t)R * p.G) * META_BYTES + (size_t)ku * p.meta_stride;
    };
    GvUnit<MPU> ring[GV_DEPTH];
#pragma unroll
    for (int d = 0; d < GV_DEPTH; ++d) {
        if (d < n_w) {
            gv_load_unit<MPU>(ring[d], wp + (size_t)d * GV_WARPS * GV_UNIT_V4, meta_ptr(R_l, ku_l), zoff);
            ku_l += GV_WARPS;
            while (ku_l >= p.UPR) { ku_l -= p.UPR; ++R_l; }
        }
    }
    // zero the partial-sum slots (a warp may never visit some row block of the range)
    {
        const int n4 = p.n_slots * GV_WARPS * p.P * 8;
        for (int i = tid; i < n4; i += GV_THREADS) part[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    auto stamp = [&](int i) {
        if (p.dbg && tid == 0) {
            unsigned long long tt;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tt));
            p.dbg[(size_t)cta * 8 + i] = tt;
        }
    };
    stamp(0);
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    stamp(1);

    // ---- stage the activations: per-(group, token) power-of-two scale, fp16, fragment order, group sums ----
    // One pass: the SUB_K/8 consecutive lanes that hold a (sub)group agree on its max by shuffles.
    const int ppr = p.K >> 3;  // 16-byte pieces per token row
    for (int m = 0; m < p.M; ++m) {
        for (int c0 = tid - lane; c0 < ppr; c0 += GV_THREADS) {   // warp-uniform trip count: the shuffles need all lanes
            const int c = c0 + lane;
            const bool live = c < ppr;
            uint4 v = make_uint4(0, 0, 0, 0);
            if (live) v = __ldg(reinterpret_cast<const uint4 *>(p.x + (size_t)m * p.K) + c);
            const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
            uint32_t mx = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) mx = max(mx, max(w4[i] & 0x7FFFu, (w4[i] >> 16) & 0x7FFFu));
#pragma unroll
            for (int o = SUB_K / 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            int E = (int)(mx >> 7);                                  // biased exponent of the group's max |x| (bf16 bits)
            E = min(max(E, 14), 254);
            const float scale = __int_as_float((267 - E) << 23);     // 2^(13 - (E-127)): max|x| lands in [2^13, 2^14)
            uint32_t h2[4];
            float sum = 0.f;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float lo = __uint_as_float(w4[i] << 16) * scale, hi = __uint_as_float(w4[i] & 0xFFFF0000u) * scale;
                const __half2 hh = __floats2half2_rn(lo, hi);
                h2[i] = *reinterpret_cast<const uint32_t *>(&hh);
                const float2 back = __half22float2(hh);
                sum += back.x + back.y;
            }
#pragma unroll
            for (int o = SUB_K / 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
            if (live) {
                xs[(size_t)(2 * c) * MR + m] = make_uint2(h2[0], h2[1]);
                xs[(size_t)(2 * c + 1) * MR + m] = make_uint2(h2[2], h2[3]);
                if ((c & (SUB_K / 8 - 1)) == 0) {
                    float *e = gtab + ((size_t)(c / (SUB_K / 8)) * PAIRS + (m >> 1)) * 4 + (m & 1);
                    e[0] = sum;
                    e[2] = __int_as_float((E - 13) << 23);           // 2^e: undoes the scale
                }
            }
        }
    }
    __syncthreads();
    stamp(2);

    // ---- main loop ----
    float yacc[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int i = 0; i < 4; ++i) yacc[nt][i] = 0.f;
    int R_c = u0 / p.UPR, ku_c = u0 - R_c * p.UPR;   // compute cursor
    // per-lane shared-memory bases: x fragment of MMA jm in unit ku = xs[((32*ku + 4*jm + t) * MR + row)]
    const uint2 *xb[NT];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) xb[nt] = xs + t * MR + min(nt * 8 + g, p.M - 1);
    auto flush = [&](int R) {
        float4 *dst = part + ((size_t)(R - R_first) * GV_WARPS + warp) * p.P * 8 + g;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int pp = nt * 4 + t;
            if (pp < p.P) dst[pp * 8] = make_float4(yacc[nt][0], yacc[nt][1], yacc[nt][2], yacc[nt][3]);
#pragma unroll
            for (int i = 0; i < 4; ++i) yacc[nt][i] = 0.f;
        }
    };

    for (int base = 0; base < n_w; base += GV_DEPTH) {
#pragma unroll
        for (int d = 0; d < GV_DEPTH; ++d) {
            if (base + d < n_w) {
                const GvUnit<MPU> &u = ring[d];
                const uint32_t wj[8] = {u.w[0].x, u.w[0].y, u.w[0].z, u.w[0].w, u.w[1].x, u.w[1].y, u.w[1].z, u.w[1].w};
#pragma unroll
                for (int h = 0; h < MPU; ++h) {
                    float acc[NT][4];
#pragma unroll
                    for (int jj = 0; jj < MMAS_PER_GROUP; ++jj) {
                        const int jm = h * MMAS_PER_GROUP + jj;
                        const uint32_t w = wj[jm], v = w >> 8;
                        const uint32_t a0 = gv_lop3_and_or(w, 0x000F000Fu, 0x64006400u);  // row g  : 1024 + q, k-slots 4t, 4t+1
                        const uint32_t a1 = gv_lop3_and_or(w, 0x00F000F0u, 0x54005400u);  // row g+8:   64 + q, k-slots 4t, 4t+1
                        const uint32_t a2 = gv_lop3_and_or(v, 0x000F000Fu, 0x64006400u);  // row g  : k-slots 4t+2, 4t+3
                        const uint32_t a3 = gv_lop3_and_or(v, 0x00F000F0u, 0x54005400u);  // row g+8: k-slots 4t+2, 4t+3
#pragma unroll
                        for (int nt = 0; nt < NT; ++nt) {
                            const uint2 xf = xb[nt][(size_t)(32 * ku_c + 4 * jm) * MR];
                            if (jj == 0) gv_mma_f16<true>(acc[nt], a0, a1, a2, a3, xf.x, xf.y);
                            else gv_mma_f16<false>(acc[nt], a0, a1, a2, a3, xf.x, xf.y);
                        }
                    }
                    const float slo = __uint_as_float(u.s2[h] << 16), shi = __uint_as_float(u.s2[h] & 0xFFFF0000u);
                    // exact floats 1024 + z (rows g) and 64 + z (rows g+8)
                    const float zlo = __uint_as_float(0x44800000u + ((u.zb[h] & 0x0Fu) << 13));
                    const float zhi = __uint_as_float(0x42800000u + ((u.zb[h] & 0xF0u) << 13));
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt) {
                        const float4 X = *reinterpret_cast<const float4 *>(
                            gtab + ((size_t)(ku_c * MPU + h) * PAIRS + (PAIRS == 1 ? 0 : nt * 4 + t)) * 4);
                        const float sl0 = slo * X.z, sl1 = slo * X.w, sh0 = shi * X.z, sh1 = shi * X.w;
                        yacc[nt][0] = fmaf(sl0, fmaf(-zlo, X.x, acc[nt][0]), yacc[nt][0]);
                        yacc[nt][1] = fmaf(sl1, fmaf(-zlo, X.y, acc[nt][1]), yacc[nt][1]);
                        yacc[nt][2] = fmaf(sh0, fmaf(-zhi, X.x, acc[nt][2]), yacc[nt][2]);
                        yacc[nt][3] = fmaf(sh1, fmaf(-zhi, X.y, acc[nt][3]), yacc[nt][3]);
                    }
                }
                // refill this ring slot GV_DEPTH units ahead
                if (base + d + GV_DEPTH < n_w) {
                    gv_load_unit<MPU>(ring[d], wp + (size_t)(base + d + GV_DEPTH) * GV_WARPS * GV_UNIT_V4, meta_ptr(R_l, ku_l), zoff);
                    ku_l += GV_WARPS;
                    while (ku_l >= p.UPR) { ku_l -= p.UPR; ++R_l; }
                }
                // advance the compute cursor; leaving a row block publishes its partial sums
                const int R_prev = R_c;
                ku_c += GV_WARPS;
                while (ku_c >= p.UPR) { ku_c -= p.UPR; ++R_c; }
                if (R_c != R_prev || base + d + 1 == n_w) flush(R_prev);
            }
        }
    }
    __syncthreads();
    stamp(3);

    // ---- combine the warps of this CTA (fixed order), write complete row blocks, publish boundary ones ----
    const int R_last = n_cta > 0 ? (b - 1) / p.UPR : R_first - 1;
    const int items = (R_last - R_first + 1) * p.P * 8;
    const bool first_partial = n_cta > 0 && (a != R_first * p.UPR || b < (R_first + 1) * p.UPR);
    const bool last_partial = n_cta > 0 && R_last != R_first && b != (R_last + 1) * p.UPR;
    auto store_y = [&](int R, int pp, int gg, float4 v) {
        const int n0 = R * 16 + gg;
        const int tok = 2 * pp;
        const float b0 = p.bias ? __bfloat162float(p.bias[n0]) : 0.f;
        const float b1 = p.bias ? __bfloat162float(p.bias[n0 + 8]) : 0.f;
        if (tok < p.M) {
            p.y[(size_t)tok * p.N + n0] = __float2bfloat16_rn(v.x + b0);
            p.y[(size_t)tok * p.N + n0 + 8] = __float2bfloat16_rn(v.z + b1);
        }
        if (tok + 1 < p.M) {
            p.y[(size_t)(tok + 1) * p.N + n0] = __float2bfloat16_rn(v.y + b0);
            p.y[(size_t)(tok + 1) * p.N + n0 + 8] = __float2bfloat16_rn(v.w + b1);
        }
    };
    for (int it = tid; it < items; it += GV_THREADS) {
        const int gg = it & 7, rest = it >> 3;
        const int pp = rest % p.P, slot = rest / p.P;
        const float4 *src = part + ((size_t)slot * GV_WARPS * p.P + pp) * 8 + gg;
        float4 v = src[0];
#pragma unroll
        for (int w = 1; w < GV_WARPS; ++w) {
            const float4 o = src[(size_t)w * p.P * 8];
            v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
        }
        const int R = R_first + slot;
        const bool is_first = slot == 0 && first_partial, is_last = R == R_last && last_partial;
        if (is_first || is_last)
            __stcg(p.scratch + (((size_t)cta * 2 + (is_first ? 0 : 1)) * p.P + pp) * 8 + gg, v);
        else
            store_y(R, pp, gg, v);
    }
    stamp(4);
    if (!first_partial && !last_partial) { stamp(5); return; }

    // ---- row blocks shared with neighbouring CTAs: the last arriver sums the contributors in CTA order ----
    __threadfence();
    __syncthreads();
    auto cta_of_unit = [&](int u) -> int {
        const int big = p.r * (p.q + 1);
        return u < big ? u / (p.q + 1) : p.r + (u - big) / p.q;   // q > 0 whenever a unit lies beyond the big ranges
    };
    if (tid < 2) {
        const bool mine = tid == 0 ? first_partial : last_partial;
        int last = 0;
        if (mine) {
            const int R = tid == 0 ? R_first : R_last;
            const int n_contrib = cta_of_unit((R + 1) * p.UPR - 1) - cta_of_unit(R * p.UPR) + 1;
            const int prev = atomicAdd(&p.tickets[R], 1);
            last = prev == n_contrib - 1;
            if (last) p.tickets[R] = 0;   // self-reset for the next launch
        }
        flags[tid] = last;
    }
    __syncthreads();
#pragma unroll 1
    for (int which = 0; which < 2; ++which) {
        if (!flags[which]) continue;
        __threadfence();
        const int R = which == 0 ? R_first : R_last;
        const int i_lo = cta_of_unit(R * p.UPR), i_hi = cta_of_unit((R + 1) * p.UPR - 1);
        for (int it = tid; it < p.P * 8; it += GV_THREADS) {
            const int gg = it & 7, pp = it >> 3;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int i = i_lo; i <= i_hi; ++i) {
                const int a_i = i * p.q + min(i, p.r);
                const int wh = (a_i / p.UPR == R) ? 0 : 1;   // contributor i holds R as its first or its last row block
                const float4 o = __ldcg(p.scratch + (((size_t)i * 2 + wh) * p.P + pp) * 8 + gg);
                v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
            }
            store_y(R, pp, gg, v);
        }
    }
    stamp(5);
}

// ---- host side ---------------------------------------------------------------------------
constexpr int GV_DBG_SLOTS = 4, GV_DBG_WORDS = 1024 * 8;
static unsigned long long *g_gv_dbg = nullptr;
static unsigned long long g_gv_dbg_launch = 0;

int gemv_debug_read(unsigned long long *out, int64_t words, unsigned long long *launches) {
    if (!g_gv_dbg) return fail("debug_stamps: no stamps recorded (set W4A16_GV_DEBUG=1 before the first launch)");
    if (words > (int64_t)GV_DBG_SLOTS * GV_DBG_WORDS) words = (int64_t)GV_DBG_SLOTS * GV_DBG_WORDS;
    W4_CUDA_OK(cudaDeviceSynchronize());
    W4_CUDA_OK(cudaMemcpy(out, g_gv_dbg, (size_t)words * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    if (launches) *launches = g_gv_dbg_launch;
    return 0;
}

struct GemvPlan {
    bool ok;
    int grid, q, r, n_slots, P, MR;
    int off_xsum, off_tok, off_part, smem_bytes;
};

static int gv_env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

static GemvPlan gemv_plan(const Linear &L, int64_t M, int sm_count) {
    GemvPlan pl{};
    pl.ok = false;
    if (M < 1 || M > 16) return pl;
    const int gs = L.d.gs;
    if (gs != 32 && gs != 64 && gs != 128 && gs != L.d.K) return pl;
    const int64_t UPR = L.d.K / GV_UNIT_K, UT64 = L.d.NR * UPR;
    if (UT64 >= (int64_t)1 << 30) return pl;
    const int UT = (int)UT64;
    const int mpu = gs == 32 ? 4 : gs == 64 ? 2 : 1;
    pl.MR = M <= 1 ? 1 : M <= 2 ? 2 : M <= 4 ? 4 : M <= 8 ? 8 : 16;
    pl.P = (int)((M + 1) / 2);
    int grid = gv_env_int("W4A16_GEMV_CTAS_PER_SM", 2) * sm_count;
    if (grid > UT) grid = UT;
    pl.grid = grid;
    pl.q = UT / grid; pl.r = UT % grid;
    const int range = pl.q + (pl.r ? 1 : 0);
    pl.n_slots = (int)((range + UPR - 1) / UPR) + 1;
    const int pairs = pl.MR < 2 ? 1 : pl.MR / 2;
    int64_t off = (int64_t)pl.MR * L.d.K * 2;                       // xs
    pl.off_xsum = (int)off; off += (int64_t)UPR * mpu * pairs * 16; // group table {X0, X1, 2^e0, 2^e1}
    off = (off + 15) & ~(int64_t)15;
    pl.off_tok = (int)off; off += 16;                               // flags[2] (+pad)
    off = (off + 15) & ~(int64_t)15;
    pl.off_part = (int)off; off += (int64_t)pl.n_slots * GV_WARPS * pl.P * 8 * 16;
    if (off > 200 * 1024) return pl;                                 // x does not fit: the register-ring kernel takes it
    pl.smem_bytes = (int)off;
    pl.ok = true;
    return pl;
}

bool gemv_stream_usable(const Linear &L, int64_t M, int sm_count) {
    if (gv_env_int("W4A16_NO_GEMV_STREAM", 0)) return false;
    return gemv_plan(L, M, sm_count).ok;
}

int64_t gemv_scratch_floats(int sm_count) {
    // [grid <= 4*sm][2][P <= 8][8] float4
    return (int64_t)4 * sm_count * 2 * 8 * 8 * 4;
}

template <int MR, int MPU>
static int gv_launch_t(const GemvParams &p, const GemvPlan &pl, cudaStream_t stream, bool pdl) {
    static int max_smem_set = 0;   // grows monotonically; racing threads only repeat the call
    if (pl.smem_bytes > 48 * 1024 && pl.smem_bytes > max_smem_set) {
        W4_CUDA_OK(cudaFuncSetAttribute(gemv_stream_kernel<MR, MPU>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        200 * 1024));
        max_smem_set = 200 * 1024;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)pl.grid);
    cfg.blockDim = dim3(GV_THREADS);
    cfg.dynamicSmemBytes = (size_t)pl.smem_bytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    W4_CUDA_OK(cudaLaunchKernelEx(&cfg, gemv_stream_kernel<MR, MPU>, p));
    count_launch();
    return 0;
}

int launch_gemv_stream(const Linear &L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias,
                       cudaStream_t stream) {
    const GemvPlan pl = gemv_plan(L, M, L.scratch->sm_count);
    if (!pl.ok) return fail("launch_gemv_stream: shape not supported by the stream kernel");
    const int gs = L.d.gs;
    const int mpu = gs == 32 ? 4 : gs == 64 ? 2 : 1;
    GemvParams p;
    p.qw = reinterpret_cast<const uint4 *>(L.qw);
    p.meta = L.meta;
    p.bias = add_bias && L.has_bias ? L.bias : nullptr;
    p.x = x; p.y = y;
    p.scratch = reinterpret_cast<float4 *>(L.splitk_scratch);
    p.tickets = L.tickets;
    p.M = (int)M; p.K = (int)L.d.K; p.N = (int)L.d.N;
    p.UPR = (int)(L.d.K / GV_UNIT_K);
    p.UT = (int)(L.d.NR * p.UPR);
    p.q = pl.q; p.r = pl.r;
    const bool per_channel = gs == L.d.K && gs != 32 && gs != 64 && gs != 128;
    p.meta_stride = per_channel ? 0 : mpu * META_BYTES;
    p.G = (int)L.d.G;
    p.P = pl.P; p.n_slots = pl.n_slots;
    p.off_xsum = pl.off_xsum; p.off_tok = pl.off_tok; p.off_part = pl.off_part;
    p.dbg = nullptr;
    if (gv_env_int("W4A16_GV_DEBUG", 0)) {   // timing experiments: launch i stamps ring slot i % 4
        if (!g_gv_dbg) W4_CUDA_OK(cudaMalloc(&g_gv_dbg, GV_DBG_SLOTS * GV_DBG_WORDS * sizeof(unsigned long long)));
        p.dbg = g_gv_dbg + (size_t)(g_gv_dbg_launch++ % GV_DBG_SLOTS) * GV_DBG_WORDS;
    }
    if ((int64_t)pl.grid * 2 * pl.P * 8 * 4 > L.splitk_floats)
        return fail("launch_gemv_stream: boundary scratch too small for this configuration");
    const bool pdl = gv_env_int("W4A16_NO_PDL", 0) == 0;
#define W4_GV(MR_)                                                               \
    do {                                                                         \
        if (mpu == 1) return gv_launch_t<MR_, 1>(p, pl, stream, pdl);            \
        if (mpu == 2) return gv_launch_t<MR_, 2>(p, pl, stream, pdl);            \
        return gv_launch_t<MR_, 4>(p, pl, stream, pdl);                          \
    } while (0)
    switch (pl.MR) {
        case 1: W4_GV(1);
        case 2: W4_GV(2);
        case 4: W4_GV(4);
        case 8: W4_GV(8);
        default: W4_GV(16);
    }
#undef W4_GV
}

}  // namespace w4
