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
// cut into gridDim.x contiguous ranges (a per-handle table, equal to +-1 unit), one CTA of 16 warps per SM.
// A CTA pulls its range through a shared-memory ring of 16 KB stages with cp.async.bulk (TMA bulk copy,
// mbarrier complete_tx); the copies are issued BEFORE griddepcontrol.wait and a CTA uses at most half of an
// SM (registers and shared memory), so under programmatic dependent launch the NEXT linear's CTAs are
// already resident and have up to ~100 KB of weights per SM landed while this one still computes: the HBM
// stream keeps running across the per-kernel dependency / staging / combine latencies.
// Inside a stage warp w owns unit w.  Partial sums of a 16-row block are combined in a fixed order: across
// the warps of a CTA in shared memory, across the CTAs that share a row block through a per-handle scratch
// and a self-resetting ticket.  No atomics on data: results are deterministic.
#include <algorithm>
#include <vector>

#include "w4a16_internal.cuh"
#include "w4a16_dequant.cuh"
#include "w4a16_gemv_common.cuh"

namespace w4 {


// MR = token rows held in shared memory (M rounded up to 1, 2, 4, 8 or 16).  NT = MMA token tiles.
// MR <= 2 runs on the integer dot-product path (see the header); MR >= 4 on mma.sync.
template <int MR, int MPU>
__global__ void __launch_bounds__(GV_THREADS, MR <= 8 ? 2 : 1) gemv_stream_kernel(const GemvParams p) {
    constexpr bool DP = MR <= 2;
    constexpr int NT = MR <= 8 ? 1 : MR / 8;
    constexpr int PAIRS = MR < 2 ? 1 : MR / 2;     // token pairs per row of the group table
    constexpr int MMAS_PER_GROUP = 8 / MPU;
    constexpr int SUB_K = GV_UNIT_K / MPU;         // k per (sub)group inside a unit
    extern __shared__ __align__(128) uint8_t gv_smem[];
    uint64_t *bars = reinterpret_cast<uint64_t *>(gv_smem);                  // full[GV_MAX_STAGES]
    uint32_t *consumed = reinterpret_cast<uint32_t *>(gv_smem + 8 * GV_MAX_STAGES);   // [GV_MAX_STAGES] arrival counters
    uint2 *xs = reinterpret_cast<uint2 *>(gv_smem + p.off_xs);               // [K/4][MR] 4 x fp16
    float *gtab = reinterpret_cast<float *>(gv_smem + p.off_gtab);           // [K/SUB_K][PAIRS] {X0, X1, 2^e0, 2^e1}
    int *flags = reinterpret_cast<int *>(gv_smem + p.off_flags);             // [2]
    float4 *part = reinterpret_cast<float4 *>(gv_smem + p.off_part);         // [n_slots][GV_WARPS][P][8]
    const uint8_t *wring = gv_smem + p.off_wring;
    const uint8_t *mring = gv_smem + p.off_mring;
    const uint32_t bar0 = gv_smem_u32(bars);
    const uint32_t wring_s = gv_smem_u32(wring), mring_s = gv_smem_u32(mring);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int cta = blockIdx.x;
    const int4 rec0 = __ldg(p.tab + 2 * cta), rec1 = __ldg(p.tab + 2 * cta + 1);
    const int a = rec0.x, b = rec0.y;
    const int n_cta = b - a;
    const int n_stages = (n_cta + GV_STAGE_UNITS - 1) / GV_STAGE_UNITS;
    const int S = p.n_stages_ring;
    const int R_first = a / p.UPR;

    // stage j of this CTA's range -> ring slot j % S
    auto issue_stage = [&](int j) {
        const int s = j % S;
        const int u_lo = a + j * GV_STAGE_UNITS;
        const int cnt = min(GV_STAGE_UNITS, b - u_lo);
        const uint32_t wb = (uint32_t)cnt * GV_UNIT_BYTES, mb = (uint32_t)cnt * (uint32_t)p.meta_stride;
        const uint32_t bar = bar0 + 8 * s;
        gv_mbar_expect_tx(bar, wb + mb);
        gv_bulk_g2s(wring_s + s * GV_STAGE_W_BYTES, p.qw + (size_t)u_lo * GV_UNIT_BYTES, wb, bar);
        if (mb) gv_bulk_g2s(mring_s + s * p.stage_m_bytes, p.meta + (size_t)u_lo * p.meta_stride, mb, bar);
    };

    // ---- prefetch (weights and meta are immutable: independent of the previous kernel) ----
    if (tid == 0) {
        for (int s = 0; s < S; ++s) { gv_mbar_init(bar0 + 8 * s, 1); consumed[s] = 0; }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        const int first = min(S, n_stages);
        for (int j = 0; j < first; ++j) issue_stage(j);
    }
    {   // zero the partial-sum slots (a warp may never visit some row block of the range)
        const int R_last0 = n_cta > 0 ? (b - 1) / p.UPR : R_first - 1;
        const int n4 = (R_last0 - R_first + 1) * GV_WARPS * p.P * 8;
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
    // x is [M][K] contiguous: piece i = m * ppr + c.  BATCH loads are in flight per thread before the first
    // one is consumed (a token-by-token loop pays one L2 round trip per token: 1.45 us each, measured).
    auto stage_piece = [&](const int i, const uint4 v, const bool live) {
        const int m = i / ppr, c = i - m * ppr;
        const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
        uint32_t mx = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) mx = max(mx, max(w4[q] & 0x7FFFu, (w4[q] >> 16) & 0x7FFFu));
#pragma unroll
        for (int o = SUB_K / 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        int E = (int)(mx >> 7);                                  // biased exponent of the group's max |x| (bf16 bits)
        E = min(max(E, 14), 254);
        const float scale = __int_as_float((267 - E) << 23);     // 2^(13 - (E-127)): max|x| lands in [2^13, 2^14)
        if constexpr (DP) {
            // Integer path: xt rounded to an integer in [-16383, 16383] (exact for every element within 2^6 of the
            // group's maximum, |error| <= 2^-14 of that maximum below), split into two signed-byte planes
            // xt = 128*hi + lo, packed per 4-k quad in the byte order (k0,k2,k1,k3) of the weight nibbles.
            int *gti = reinterpret_cast<int *>(gtab);                    // [K/SUB_K][MR]{sum_t0..3, 2^e, -, -, -}
            int xi[8];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                xi[2 * q] = __float2int_rn(__uint_as_float(w4[q] << 16) * scale);
                xi[2 * q + 1] = __float2int_rn(__uint_as_float(w4[q] & 0xFFFF0000u) * scale);
            }
            int sumA = xi[0] + xi[1] + xi[2] + xi[3], sumB = xi[4] + xi[5] + xi[6] + xi[7];
#pragma unroll
            for (int o = SUB_K / 16; o > 1; o >>= 1) {   // pieces of equal parity share the quad positions t
                sumA += __shfl_xor_sync(0xffffffffu, sumA, o);
                sumB += __shfl_xor_sync(0xffffffffu, sumB, o);
            }
            if (live) {
                xs[(size_t)(2 * c) * MR + m] = make_uint2(gv_pack4(xi[0] >> 7, xi[2] >> 7, xi[1] >> 7, xi[3] >> 7),
                                                          gv_pack4(xi[0] & 127, xi[2] & 127, xi[1] & 127, xi[3] & 127));
                xs[(size_t)(2 * c + 1) * MR + m] = make_uint2(gv_pack4(xi[4] >> 7, xi[6] >> 7, xi[5] >> 7, xi[7] >> 7),
                                                              gv_pack4(xi[4] & 127, xi[6] & 127, xi[5] & 127, xi[7] & 127));
                const int in_grp = c & (SUB_K / 8 - 1);
                if (in_grp < 2) {
                    int *e = gti + ((size_t)(c / (SUB_K / 8)) * MR + m) * 8;
                    e[2 * in_grp] = sumA;
                    e[2 * in_grp + 1] = sumB;
                    // bits of 2^e (undoes the scale); NaN when the (group, token) holds Inf / NaN: non-finite in -> NaN out
                    if (in_grp == 0) e[4] = mx >= 0x7F80u ? 0x7FC00000 : (E - 13) << 23;
                }
            }
        } else {
            uint32_t h2[4];
            float sum = 0.f;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float lo = __uint_as_float(w4[q] << 16) * scale, hi = __uint_as_float(w4[q] & 0xFFFF0000u) * scale;
                const __half2 hh = __floats2half2_rn(lo, hi);
                h2[q] = *reinterpret_cast<const uint32_t *>(&hh);
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
    };
    {
        const int total = p.M * ppr;
        const uint4 *x16 = reinterpret_cast<const uint4 *>(p.x);
        constexpr int BATCH = 4;
        for (int base = 0; base < total; base += BATCH * GV_THREADS) {   // warp-uniform trip count: the shuffles need all lanes
            uint4 v[BATCH];
#pragma unroll
            for (int r = 0; r < BATCH; ++r) {
                const int i = base + r * GV_THREADS + tid;
                v[r] = i < total ? __ldg(x16 + i) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int r = 0; r < BATCH; ++r) {
                if (base + r * GV_THREADS >= total) break;               // warp-uniform
                const int i = base + r * GV_THREADS + tid;
                stage_piece(min(i, total - 1), v[r], i < total);
            }
        }
    }
    __syncthreads();
    stamp(2);

    // ---- main loop: stage j, unit a + 16j + warp ----
    float yacc[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int i = 0; i < 4; ++i) yacc[nt][i] = 0.f;
    int u = a + warp;
    int R_c = u / p.UPR, ku_c = u - R_c * p.UPR;   // cursor of this warp's unit
    int R_acc = -1;                                // row block whose partial sums yacc holds (-1: none)
    // per-lane shared-memory bases: x fragment of MMA jm in unit ku = xs[((32*ku + 4*jm + t) * MR + row)]
    const uint2 *xb[NT];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) xb[nt] = xs + t * MR + min(nt * 8 + g, p.M - 1);
    auto flush = [&](int R) {
        float4 *dst = part + ((size_t)(R - R_first) * GV_WARPS + warp) * p.P * 8 + g;
        if constexpr (DP) {   // yacc[0] = {row g tok0, row g tok1, row g+8 tok0, row g+8 tok1} over this lane's k-slots
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float v = yacc[0][i];
                v += __shfl_xor_sync(0xffffffffu, v, 1);
                v += __shfl_xor_sync(0xffffffffu, v, 2);
                yacc[0][i] = v;
            }
            if (t == 0) dst[0] = make_float4(yacc[0][0], yacc[0][1], yacc[0][2], yacc[0][3]);
#pragma unroll
            for (int i = 0; i < 4; ++i) yacc[0][i] = 0.f;
            return;
        }
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int pp = nt * 4 + t;
            if (pp < p.P) dst[pp * 8] = make_float4(yacc[nt][0], yacc[nt][1], yacc[nt][2], yacc[nt][3]);
#pragma unroll
            for (int i = 0; i < 4; ++i) yacc[nt][i] = 0.f;
        }
    };

    bool fast_done = false;
    if constexpr (DP) {
        if (p.fast) {
            // Fast path (whole row blocks per CTA, K % 2048 == 0 so every ring stage is full, grouped meta):
            // the same arithmetic as the general loop below with every address a running 32-bit shared-memory
            // offset and every branch warp-uniform.
            const uint32_t my_w = wring_s + (uint32_t)(warp * GV_UNIT_BYTES + lane * 16);
            const uint32_t my_m = mring_s + (uint32_t)(warp * (MPU * META_BYTES) + 4 * g);
            const uint32_t my_z = mring_s + (uint32_t)(warp * (MPU * META_BYTES) + META_ZERO_OFF + g);
            const uint32_t xk0 = gv_smem_u32(xs) + (uint32_t)((32 * warp + t) * MR * 8);
            const uint32_t gk0 = gv_smem_u32(gtab) + (uint32_t)(warp * MPU * MR * 32 + 4 * t);
            uint32_t xk = xk0, gk = gk0, woff = 0, moff = 0;
            int sk = 0, R = R_first, s = 0;
            uint32_t phase = 0;
            for (int j = 0; j < n_stages; ++j) {
                gv_mbar_wait(bar0 + 8 * s, phase);
                uint32_t wj[8], s2[MPU], zb[MPU];
                asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(wj[0]), "=r"(wj[1]), "=r"(wj[2]), "=r"(wj[3]) : "r"(my_w + woff));
                asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4+512];" : "=r"(wj[4]), "=r"(wj[5]), "=r"(wj[6]), "=r"(wj[7]) : "r"(my_w + woff));
#pragma unroll
                for (int h = 0; h < MPU; ++h) {
                    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(s2[h]) : "r"(my_m + moff + h * META_BYTES));
                    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(zb[h]) : "r"(my_z + moff + h * META_BYTES));
                }
#pragma unroll
                for (int h = 0; h < MPU; ++h) {
                    int acc[MR][4];
#pragma unroll
                    for (int m = 0; m < MR; ++m)
#pragma unroll
                        for (int i = 0; i < 4; ++i) acc[m][i] = 0;
#pragma unroll
                    for (int jj = 0; jj < MMAS_PER_GROUP; ++jj) {
                        const int jm = h * MMAS_PER_GROUP + jj;
                        const uint32_t w = wj[jm];
                        const uint32_t ev = w & 0x0F0F0F0Fu;          // row g  , k-slots (4t, 4t+2, 4t+1, 4t+3)
                        const uint32_t od = (w >> 4) & 0x0F0F0F0Fu;   // row g+8, same k-slots
#pragma unroll
                        for (int m = 0; m < MR; ++m) {
                            uint32_t xh, xl;
                            asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(xh), "=r"(xl) : "r"(xk + (uint32_t)((4 * jm * MR + m) * 8)));
                            acc[m][0] = gv_dp4a_u8s8(ev, xh, acc[m][0]);
                            acc[m][1] = gv_dp4a_u8s8(ev, xl, acc[m][1]);
                            acc[m][2] = gv_dp4a_u8s8(od, xh, acc[m][2]);
                            acc[m][3] = gv_dp4a_u8s8(od, xl, acc[m][3]);
                        }
                    }
                    const float slo = __uint_as_float(s2[h] << 16), shi = __uint_as_float(s2[h] & 0xFFFF0000u);
                    const int zlo = (int)(zb[h] & 0x0Fu), zhi = (int)(zb[h] >> 4);
#pragma unroll
                    for (int m = 0; m < MR; ++m) {
                        int xsum; uint32_t e2b;
                        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(xsum) : "r"(gk + (uint32_t)((h * MR + m) * 32)));
                        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(e2b) : "r"(gk + (uint32_t)((h * MR + m) * 32 + 16) - 4u * (uint32_t)t));
                        const float e2 = __uint_as_float(e2b);
                        const int tg = acc[m][0] * 128 + acc[m][1] - zlo * xsum;     // exact: |.| < 2^24
                        const int tg8 = acc[m][2] * 128 + acc[m][3] - zhi * xsum;
                        yacc[0][m] = fmaf(slo * e2, (float)tg, yacc[0][m]);
                        yacc[0][2 + m] = fmaf(shi * e2, (float)tg8, yacc[0][2 + m]);
                    }
                }
                if (j + S < n_stages) {   // warp-uniform: the 16th arrival refills this ring slot
                    __syncwarp();
                    if (lane == 0) {
                        __threadfence_block();
                        const uint32_t old = atomicAdd(&consumed[s], 1u);
                        if ((old & (GV_WARPS - 1)) == GV_WARPS - 1) {
                            __threadfence_block();
                            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                            issue_stage(j + S);
                        }
                    }
                }
                woff += GV_STAGE_W_BYTES; moff += (uint32_t)p.stage_m_bytes;
                if (++s == S) { s = 0; phase ^= 1u; woff = 0; moff = 0; }
                sk += GV_STAGE_UNITS; xk += GV_STAGE_UNITS * 32 * MR * 8; gk += GV_STAGE_UNITS * MPU * MR * 32;
                if (sk == p.UPR) { flush(R); ++R; sk = 0; xk = xk0; gk = gk0; }
            }
            fast_done = true;
        }
    }
    int s = 0;
    uint32_t phase = 0;
    for (int j = 0; j < (fast_done ? 0 : n_stages); ++j) {
        gv_mbar_wait(bar0 + 8 * s, phase);
        if (u < b) {
            if (R_acc >= 0 && R_acc != R_c) flush(R_acc);
            R_acc = R_c;
            const uint4 *wsrc = reinterpret_cast<const uint4 *>(wring + s * GV_STAGE_W_BYTES + warp * GV_UNIT_BYTES) + lane;
            const uint4 w0 = wsrc[0], w1 = wsrc[32];
            uint32_t s2[MPU], zb[MPU];
            if (p.meta_stride) {
                const uint8_t *m = mring + s * p.stage_m_bytes + warp * (MPU * META_BYTES);
#pragma unroll
                for (int h = 0; h < MPU; ++h) {
                    s2[h] = *reinterpret_cast<const uint32_t *>(m + h * META_BYTES + 4 * g);
                    zb[h] = m[h * META_BYTES + META_ZERO_OFF + g];
                }
            } else {   // per-channel scales: one meta block per row block, read through L1
                const uint8_t *m = p.meta + (size_t)R_c * META_BYTES;
                s2[0] = __ldg(reinterpret_cast<const uint32_t *>(m + 4 * g));
                zb[0] = __ldg(m + META_ZERO_OFF + g);
            }
            const uint32_t wj[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
            if constexpr (DP) {
                const int *gti = reinterpret_cast<const int *>(gtab);
#pragma unroll
                for (int h = 0; h < MPU; ++h) {
                    int acc[MR][4];
#pragma unroll
                    for (int m = 0; m < MR; ++m)
#pragma unroll
                        for (int i = 0; i < 4; ++i) acc[m][i] = 0;
#pragma unroll
                    for (int jj = 0; jj < MMAS_PER_GROUP; ++jj) {
                        const int jm = h * MMAS_PER_GROUP + jj;
                        const uint32_t w = wj[jm];
                        const uint32_t ev = w & 0x0F0F0F0Fu;          // row g  , k-slots (4t, 4t+2, 4t+1, 4t+3)
                        const uint32_t od = (w >> 4) & 0x0F0F0F0Fu;   // row g+8, same k-slots
#pragma unroll
                        for (int m = 0; m < MR; ++m) {
                            const uint2 xv = xs[(size_t)(32 * ku_c + 4 * jm + t) * MR + min(m, p.M - 1)];
                            acc[m][0] = gv_dp4a_u8s8(ev, xv.x, acc[m][0]);
                            acc[m][1] = gv_dp4a_u8s8(ev, xv.y, acc[m][1]);
                            acc[m][2] = gv_dp4a_u8s8(od, xv.x, acc[m][2]);
                            acc[m][3] = gv_dp4a_u8s8(od, xv.y, acc[m][3]);
                        }
                    }
                    const float slo = __uint_as_float(s2[h] << 16), shi = __uint_as_float(s2[h] & 0xFFFF0000u);
                    const int zlo = (int)(zb[h] & 0x0Fu), zhi = (int)(zb[h] >> 4);
#pragma unroll
                    for (int m = 0; m < MR; ++m) {
                        const int *e = gti + ((size_t)(ku_c * MPU + h) * MR + min(m, p.M - 1)) * 8;
                        const int xsum = e[t];
                        const float e2 = __int_as_float(e[4]);
                        const int tg = acc[m][0] * 128 + acc[m][1] - zlo * xsum;     // exact: |.| < 2^24
                        const int tg8 = acc[m][2] * 128 + acc[m][3] - zhi * xsum;
                        yacc[0][m] = fmaf(slo * e2, (float)tg, yacc[0][m]);
                        yacc[0][2 + m] = fmaf(shi * e2, (float)tg8, yacc[0][2 + m]);
                    }
                }
            } else {
#pragma unroll
            for (int h = 0; h < MPU; ++h) {
                // CH independent accumulator chains: a dependent HMMA only issues after the previous one retires
                constexpr int CH = MMAS_PER_GROUP >= 4 ? 4 : MMAS_PER_GROUP;
                float accs[CH][NT][4];
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
                        if (jj < CH) gv_mma_f16<true>(accs[jj % CH][nt], a0, a1, a2, a3, xf.x, xf.y);
                        else gv_mma_f16<false>(accs[jj % CH][nt], a0, a1, a2, a3, xf.x, xf.y);
                    }
                }
                float acc[NT][4];
#pragma unroll
                for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        float v = accs[0][nt][i];
#pragma unroll
                        for (int c = 1; c < CH; ++c) v += accs[c][nt][i];
                        acc[nt][i] = v;
                    }
                const float slo = __uint_as_float(s2[h] << 16), shi = __uint_as_float(s2[h] & 0xFFFF0000u);
                // exact floats 1024 + z (rows g) and 64 + z (rows g+8)
                const float zlo = __uint_as_float(0x44800000u + ((zb[h] & 0x0Fu) << 13));
                const float zhi = __uint_as_float(0x42800000u + ((zb[h] & 0xF0u) << 13));
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
            }
        }
        // this warp is done with ring slot s; the 16th arrival refills it (generic reads -> async-proxy write)
        if (j + S < n_stages) {
            __syncwarp();
            if (lane == 0) {
                __threadfence_block();
                const uint32_t old = atomicAdd(&consumed[s], 1u);
                if ((old & (GV_WARPS - 1)) == GV_WARPS - 1) {
                    __threadfence_block();
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    issue_stage(j + S);
                }
            }
        }
        u += GV_STAGE_UNITS;
        ku_c += GV_STAGE_UNITS;
        while (ku_c >= p.UPR) { ku_c -= p.UPR; ++R_c; }
        if (++s == S) { s = 0; phase ^= 1u; }
    }
    if (!fast_done && R_acc >= 0) flush(R_acc);
    __syncthreads();
    stamp(3);

    // ---- combine the warps of this CTA (fixed order), write complete row blocks, publish boundary ones ----
    const int R_last = n_cta > 0 ? (b - 1) / p.UPR : R_first - 1;
    const int items = (R_last - R_first + 1) * p.P * 8;
    const bool first_partial = rec0.z >= 0, last_partial = rec1.x >= 0;
    auto store_y = [&](int R, int pp, int gg, float4 v) {
        const int n0 = R * 16 + gg;
        const int tok = 2 * pp;
        const float b0 = p.bias ? __bfloat162float(p.bias[n0]) : 0.f;
        const float b1 = p.bias ? __bfloat162float(p.bias[n0 + 8]) : 0.f;
        if (p.swiglu) {   // rows g / g+8 are (gate, up) of output column 8 R + gg
            const size_t I = (size_t)(p.N >> 1), c = (size_t)R * 8 + gg;
            if (tok < p.M) p.y[(size_t)tok * I + c] = swiglu_bf16(v.x + b0, v.z + b1);
            if (tok + 1 < p.M) p.y[(size_t)(tok + 1) * I + c] = swiglu_bf16(v.y + b0, v.w + b1);
            return;
        }
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
        const bool is_first = slot == 0 && first_partial, is_last = slot != 0 && R == R_last && last_partial;
        if (is_first || is_last)
            __stcg(p.scratch + (((size_t)cta * 2 + (is_first ? 0 : 1)) * p.P + pp) * 8 + gg, v);
        else
            store_y(R, pp, gg, v);
    }
    stamp(4);
    if (!first_partial && !last_partial) { stamp(5); return; }

    // ---- row blocks shared with neighbouring CTAs: the last arriver sums the contributors in CTA order ----
    // release: every thread's scratch stores, then the CTA barrier, then one acq_rel ticket per shared row block
    asm volatile("fence.acq_rel.gpu;" ::: "memory");
    __syncthreads();
    if (tid < 2) {
        const bool mine = tid == 0 ? first_partial : last_partial;
        int last = 0;
        if (mine) {
            const int R = tid == 0 ? R_first : R_last;
            const int n_contrib = tid == 0 ? rec0.w - rec0.z + 1 : rec1.y - rec1.x + 1;
            int prev;
            asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], 1;" : "=r"(prev) : "l"(p.tickets + R) : "memory");
            last = prev == n_contrib - 1;
            if (last) p.tickets[R] = 0;   // self-reset for the next launch
        }
        flags[tid] = last;
    }
    __syncthreads();
#pragma unroll 1
    for (int which = 0; which < 2; ++which) {
        if (!flags[which]) continue;
        asm volatile("fence.acq_rel.gpu;" ::: "memory");
        const int R = which == 0 ? R_first : R_last;
        const int i_lo = which == 0 ? rec0.z : rec1.x, i_hi = which == 0 ? rec0.w : rec1.y;
        for (int it = tid; it < p.P * 8; it += GV_THREADS) {
            const int gg = it & 7, pp = it >> 3;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int i = i_lo; i <= i_hi; ++i) {
                const int a_i = __ldg(p.tab + 2 * i).x;
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


// timing experiments (W4A16_GV_DEBUG=1): launch i stamps ring slot i % 4
unsigned long long *gemv_debug_slot() {
    if (!gv_env_int("W4A16_GV_DEBUG", 0)) return nullptr;
    if (!g_gv_dbg && cudaMalloc(&g_gv_dbg, GV_DBG_SLOTS * GV_DBG_WORDS * sizeof(unsigned long long)) != cudaSuccess) return nullptr;
    return g_gv_dbg + (size_t)(g_gv_dbg_launch++ % GV_DBG_SLOTS) * GV_DBG_WORDS;
}

// The work partition of one handle: contiguous unit ranges, one per CTA, and for every range end that cuts
// a row block the CTAs that share it.  Built once per handle (create), lives in device memory.
int gemv_build_table(Linear &L, int sm_count) {
    const int64_t UPR = L.d.K / GV_UNIT_K, UT = L.d.NR * UPR;
    L.gv_grid = 0;
    if (UT >= (int64_t)1 << 30) return 0;   // the register-ring kernel takes it
    int grid = (int)std::min<int64_t>(UT, (int64_t)sm_count * exp_env_int("W4A16_GEMV_CTAS_PER_SM", 1));
    const int align = exp_env_int("W4A16_GEMV_ALIGN", 1);   // 1: ranges of whole row blocks (no shared row blocks, coarser balance)
    // fewer row blocks than SMs but at least half as many: one row block per CTA beats sharing row blocks across CTAs
    if (align && L.d.NR < grid && 2 * L.d.NR >= grid) grid = (int)L.d.NR;
    const int64_t q = UT / grid, r = UT % grid;
    std::vector<int> start(grid + 1);
    for (int i = 0; i <= grid; ++i)
        start[i] = align && L.d.NR >= grid ? (int)((i * L.d.NR / grid) * UPR) : (int)(i * q + std::min<int64_t>(i, r));
    std::vector<int> rec((size_t)grid * 8, -1);
    int max_slots = 1;
    auto contributors = [&](int64_t R, int &lo, int &hi) {
        const int64_t u0 = R * UPR, u1 = (R + 1) * UPR - 1;
        lo = (int)(std::upper_bound(start.begin(), start.end(), (int)u0) - start.begin()) - 1;
        hi = (int)(std::upper_bound(start.begin(), start.end(), (int)u1) - start.begin()) - 1;
    };
    for (int i = 0; i < grid; ++i) {
        const int a = start[i], b = start[i + 1];
        const int64_t Rf = a / UPR, Rl = (b - 1) / UPR;
        int *rc = &rec[(size_t)i * 8];
        rc[0] = a; rc[1] = b;
        max_slots = std::max<int>(max_slots, (int)(Rl - Rf + 1));
        if (a != Rf * UPR || b < (Rf + 1) * UPR) contributors(Rf, rc[2], rc[3]);
        if (Rl != Rf && b != (Rl + 1) * UPR) contributors(Rl, rc[4], rc[5]);
    }
    cudaFree(L.gv_tab); L.gv_tab = nullptr;
    W4_CUDA_OK(cudaMalloc(&L.gv_tab, rec.size() * sizeof(int)));
    W4_CUDA_OK(cudaMemcpy(L.gv_tab, rec.data(), rec.size() * sizeof(int), cudaMemcpyHostToDevice));
    L.gv_grid = grid;
    L.gv_max_slots = max_slots;
    L.gv_aligned = align && L.d.NR >= grid;
    return 0;
}

struct GemvPlan {
    bool ok;
    int P, MR, S, stage_m_bytes;
    int off_xs, off_gtab, off_flags, off_part, off_wring, off_mring, smem_bytes;
};

static GemvPlan gemv_plan(const Linear &L, int64_t M) {
    GemvPlan pl{};
    pl.ok = false;
    if (M < 1 || M > 16 || L.gv_grid <= 0 || !L.gv_tab) return pl;
    const int gs = L.d.gs;
    if (gs != 32 && gs != 64 && gs != 128 && gs != L.d.K) return pl;
    const int64_t UPR = L.d.K / GV_UNIT_K;
    const bool per_channel = gs == L.d.K && gs != 32 && gs != 64 && gs != 128;
    const int mpu = gs == 32 ? 4 : gs == 64 ? 2 : 1;
    pl.MR = M <= 1 ? 1 : M <= 2 ? 2 : M <= 4 ? 4 : M <= 8 ? 8 : 16;
    pl.P = (int)((M + 1) / 2);
    const int pairs = pl.MR < 2 ? 1 : pl.MR / 2;
    int64_t off = 8 * GV_MAX_STAGES + 4 * GV_MAX_STAGES;            // mbarriers + arrival counters
    off = (off + 127) & ~(int64_t)127;
    pl.off_xs = (int)off; off += (int64_t)pl.MR * L.d.K * 2;         // fp16 activations, fragment order
    pl.off_gtab = (int)off;                                          // group table: {X0, X1, 2^e0, 2^e1} per token pair,
    off += pl.MR <= 2 ? (int64_t)UPR * mpu * pl.MR * 32 : (int64_t)UPR * mpu * pairs * 16;   // or 8 ints per token (integer path)
    pl.off_flags = (int)off; off += 16;
    pl.off_part = (int)off; off += (int64_t)L.gv_max_slots * GV_WARPS * pl.P * 8 * 16;
    off = (off + 127) & ~(int64_t)127;
    pl.stage_m_bytes = per_channel ? 0 : GV_STAGE_UNITS * mpu * META_BYTES;
    const int64_t stage = GV_STAGE_W_BYTES + pl.stage_m_bytes;
    const int64_t fixed = off;
    int64_t budget = pl.MR <= 8 ? GV_SMEM_HALF_SM : GV_SMEM_FULL_SM;
    int64_t S = (budget - fixed) / stage;
    if (S < 3) { budget = GV_SMEM_FULL_SM; S = (budget - fixed) / stage; }
    if (S < 2) return pl;                                            // x does not fit: the register-ring kernel takes it
    S = std::min<int64_t>(S, std::min<int64_t>(GV_MAX_STAGES, exp_env_int("W4A16_GEMV_STAGES", GV_MAX_STAGES)));
    pl.S = (int)S;
    pl.off_wring = (int)off; off += S * GV_STAGE_W_BYTES;
    pl.off_mring = (int)off; off += S * pl.stage_m_bytes;
    pl.smem_bytes = (int)off;
    pl.ok = true;
    return pl;
}

bool gemv_stream_usable(const Linear &L, int64_t M) {
    if (exp_env_int("W4A16_NO_GEMV_STREAM", 0)) return false;
    return gemv_plan(L, M).ok;
}

int64_t gemv_scratch_floats(int sm_count) {
    // [grid <= 4*sm][2][P <= 8][8] float4
    return (int64_t)4 * sm_count * 2 * 8 * 8 * 4;
}

template <int MR, int MPU>
static int gv_launch_t(const GemvParams &p, const GemvPlan &pl, int grid, cudaStream_t stream, bool pdl) {
    static bool attr_set = false;   // racing threads only repeat the call
    if (!attr_set) {
        W4_CUDA_OK(cudaFuncSetAttribute(gemv_stream_kernel<MR, MPU>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        GV_SMEM_FULL_SM));
        attr_set = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)grid);
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
    const GemvPlan pl = gemv_plan(L, M);
    if (!pl.ok) return fail("launch_gemv_stream: shape not supported by the stream kernel");
    const int gs = L.d.gs;
    const int mpu = gs == 32 ? 4 : gs == 64 ? 2 : 1;
    GemvParams p{};
    p.qw = reinterpret_cast<const uint8_t *>(L.qw);
    p.meta = L.meta;
    p.bias = add_bias && L.has_bias ? L.bias : nullptr;
    p.x = x; p.y = y;
    p.scratch = reinterpret_cast<float4 *>(L.splitk_scratch);
    p.tickets = L.tickets;
    p.tab = reinterpret_cast<const int4 *>(L.gv_tab);
    p.M = (int)M; p.K = (int)L.d.K; p.N = (int)L.d.N;
    p.swiglu = L.epilogue == W4A16_EPILOGUE_SWIGLU;
    p.UPR = (int)(L.d.K / GV_UNIT_K);
    p.meta_stride = pl.stage_m_bytes ? mpu * META_BYTES : 0;
    p.P = pl.P;
    p.n_stages_ring = pl.S;
    p.fast = L.gv_aligned && (L.d.K % (GV_UNIT_K * GV_STAGE_UNITS) == 0) && pl.stage_m_bytes != 0 &&
             exp_env_int("W4A16_GEMV_FAST", 1);
    p.stage_m_bytes = pl.stage_m_bytes;
    p.off_xs = pl.off_xs; p.off_gtab = pl.off_gtab; p.off_flags = pl.off_flags; p.off_part = pl.off_part;
    p.off_wring = pl.off_wring; p.off_mring = pl.off_mring;
    p.dbg = gemv_debug_slot();
    if ((int64_t)L.gv_grid * 2 * pl.P * 8 * 4 > L.splitk_floats)
        return fail("launch_gemv_stream: boundary scratch too small for this configuration");
    const bool pdl = gv_env_int("W4A16_NO_PDL", 0) == 0;
#define W4_GV(MR_)                                                                     \
    do {                                                                               \
        if (mpu == 1) return gv_launch_t<MR_, 1>(p, pl, L.gv_grid, stream, pdl);       \
        if (mpu == 2) return gv_launch_t<MR_, 2>(p, pl, L.gv_grid, stream, pdl);       \
        return gv_launch_t<MR_, 4>(p, pl, L.gv_grid, stream, pdl);                     \
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
