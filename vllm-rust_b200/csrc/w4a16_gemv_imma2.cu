// w4a16_gemv_imma2.cu -- decode GEMV for M <= 4 tokens, int8 tensor cores, second generation.
//
// Same arithmetic, same W4T stream, same shared-memory ring and the same programmatic-launch protocol as
// gemv_imma_kernel (w4a16_gemv_imma.cu, which documents them and stays the kernel for every other shape); what changes
// is the work per instruction.  The first kernel runs 16 warps x 2 units per iteration under a 64-register cap (two
// 512-thread CTAs per SM) and issues 172 SASS instructions per 2 KB of weights per warp: measured (per-CTA
// %globaltimer stamps, profiles/r2_decode_timeline.md) its main loop is issue-bound at ~7.9 TB/s-equivalent even on
// data that is already in shared memory, which is what the small linears (qkv, o: wholly prefetched before they start)
// and the first 96 KB per SM of the large ones pay.  Here a CTA has 8 warps of 4 units per iteration (still two CTAs
// per SM: this linear + the prefetching next one, now with 128 registers per thread), and for K = 4096 a warp meets
// the SAME four k-units in every row block, so
//   * its B fragments (the staged activations) and the per-group table {X, 2^e} live in registers for the whole kernel
//     (HOIST): no shared-memory reads and no address arithmetic for them inside the loop;
//   * the 16 IMMAs of an iteration are issued as four interleaved chains (no second accumulator set, no merge adds);
//   * ring stages are always whole (K % 4096 == 0), so the refill carries one predicate instead of six.
// Replaces awq_gemv_int4_kt_bf16 / kt_mloop (awq_gemv.cu:377-665) as dispatched by marlin_gemm_pooled
// (marlin_cuda.rs:843-953, 1655-1753) for K % 4096 == 0, group size 32 / 64 / 128, M <= 4.
#include <algorithm>

#include "w4a16_gemv_common.cuh"

namespace w4 {

constexpr int G2_WARPS = 8;
constexpr int G2_THREADS = 256;
constexpr int G2_UPW = 4;            // units per warp per iteration (a stage = 32 units)

__device__ __forceinline__ void g2_imma(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// MR = token slots (1, 2, 4).  MPU = groups per 128-k unit.  DI = ring stages.  HOIST = K == 4096 and MPU == 1: the
// warp's B fragments and group-table entries are loop invariants held in registers.
template <int MR, int MPU, int DI, bool HOIST>
__global__ void __launch_bounds__(G2_THREADS, 2) gemv_imma2_kernel(const GemvParams p) {
    constexpr int COLS = 2 * MR;
    constexpr int IPG = 4 / MPU;
    constexpr int SUB_K = GV_UNIT_K / MPU;
    constexpr int GL = SUB_K / 8;
    extern __shared__ __align__(128) uint8_t gv_smem[];
    uint8_t *xs = gv_smem + p.off_xs;                                   // [K/32][4 t][COLS][8 B]: B fragments
    int2 *gtab = reinterpret_cast<int2 *>(gv_smem + p.off_gtab);        // [K/SUB_K][MR] {X, bits of 2^e}
    float2 *part = reinterpret_cast<float2 *>(gv_smem + p.off_part);    // [slots][G2_WARPS][8 g][MR] {row g, row g+8}

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int cta = blockIdx.x;
    const int4 rec0 = __ldg(p.tab + 2 * cta);
    const int a = rec0.x, b = rec0.y;          // unit range: whole row blocks, each a whole number of 32-unit stages
    const int R_first = a / p.UPR, n_slots = (b - a) / p.UPR;
    const int n_st = (b - a) >> 5;
    constexpr uint32_t STAGE_W = 32 * GV_UNIT_BYTES, STAGE_M = 32 * MPU * META_BYTES, STAGE = STAGE_W + STAGE_M;
    const uint32_t ring = gv_smem_u32(gv_smem + p.off_wring);
    const uint32_t bar_full = gv_smem_u32(gv_smem);

    // ---- prefetch: the first DI stages by bulk copy, before the dependency wait (weights are immutable) ----
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < DI; ++s) gv_mbar_init(bar_full + 8 * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#pragma unroll
        for (int s = 0; s < DI; ++s) {
            if (s < n_st) {
                const int u_lo = a + 32 * s;
                const uint32_t bar = bar_full + 8 * s, dst = ring + (uint32_t)s * STAGE;
                gv_mbar_expect_tx(bar, STAGE);
                gv_bulk_g2s(dst, p.qw + (size_t)u_lo * GV_UNIT_BYTES, STAGE_W, bar);
                gv_bulk_g2s(dst + STAGE_W, p.meta + (size_t)u_lo * (MPU * META_BYTES), STAGE_M, bar);
            }
        }
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

    // ---- stage the activations (identical to gemv_imma_kernel: per-(group, token) power-of-two scale, two signed-byte
    //      planes in B-fragment order, integer group sums; Inf / NaN poison the group's scale) ----
    {
        const int ppr = p.K >> 3, total = p.M * ppr;
        const uint4 *x16 = reinterpret_cast<const uint4 *>(p.x);
        constexpr int BATCH = 4;
        for (int base = 0; base < total; base += BATCH * G2_THREADS) {
            uint4 v[BATCH];
#pragma unroll
            for (int r = 0; r < BATCH; ++r) {
                const int i = base + r * G2_THREADS + tid;
                v[r] = i < total ? __ldg(x16 + i) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int r = 0; r < BATCH; ++r) {
                const int i = base + r * G2_THREADS + tid;
                if (base + r * G2_THREADS >= total) break;               // warp-uniform
                const uint32_t w4[4] = {v[r].x, v[r].y, v[r].z, v[r].w};
                uint32_t mx = 0;
#pragma unroll
                for (int q = 0; q < 4; ++q) mx = max(mx, max(w4[q] & 0x7FFFu, (w4[q] >> 16) & 0x7FFFu));
                mx = gv_group_max<GL>(mx, lane);
                if (p.dbg && tid == 0 && base == 0 && r == 0) {   // stamp 6: the first activation load has arrived
                    unsigned long long tt;
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tt));
                    p.dbg[(size_t)cta * 8 + 6] = tt;
                }
                int E = (int)(mx >> 7);
                E = min(max(E, 14), 254);
                const float scale = __int_as_float((267 - E) << 23);     // max|x| lands in [2^13, 2^14)
                uint32_t pk[4];
                int sum = 0;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float lo = fmaf(__uint_as_float(w4[q] << 16), scale, 12582912.f);
                    const float hi = fmaf(__uint_as_float(w4[q] & 0xFFFF0000u), scale, 12582912.f);
                    pk[q] = __byte_perm(__float_as_uint(lo), __float_as_uint(hi), 0x5410);
                    sum = __dp2a_lo((int)pk[q], 0x0101, sum);
                }
                sum = gv_group_add<GL>(sum, lane);
                if (i < total) {
                    const int m = MR == 1 ? 0 : (i >= ppr) + (MR > 2 ? (i >= 2 * ppr) + (i >= 3 * ppr) : 0);
                    const int c = i - m * ppr;
                    const int j = (c & 15) >> 1;
                    uint8_t *dst = xs + ((size_t)(((c >> 4) * 4 + (j >> 1)) * 4 + 2 * (c & 1)) * COLS + 2 * m) * 8 + (j & 1) * 4;
                    uint32_t hiP[4], loP[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) { hiP[q] = (pk[q] >> 7) & 0x00FF00FFu; loP[q] = pk[q] & 0x007F007Fu; }
                    uint32_t v0 = hiP[0] | (hiP[1] << 8), v1 = loP[0] | (loP[1] << 8);
                    uint32_t v2 = hiP[2] | (hiP[3] << 8), v3 = loP[2] | (loP[3] << 8);
                    if constexpr (MR >= 2) {   // rotate the order of the four stores per lane group: 16-way -> 4-way conflicts
                        const int rot = (lane >> 2) & 3;
                        if (rot & 1) { const uint32_t tv = v0; v0 = v1; v1 = v2; v2 = v3; v3 = tv; }
                        if (rot & 2) { uint32_t tv = v0; v0 = v2; v2 = tv; tv = v1; v1 = v3; v3 = tv; }
                        const uint32_t vv[4] = {v0, v1, v2, v3};
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const int idx = (k + rot) & 3;
                            *reinterpret_cast<uint32_t *>(dst + (idx & 1) * 8 + (idx >> 1) * (COLS * 8)) = vv[k];
                        }
                    } else {
                        *reinterpret_cast<uint32_t *>(dst) = v0;
                        *reinterpret_cast<uint32_t *>(dst + 8) = v1;
                        *reinterpret_cast<uint32_t *>(dst + COLS * 8) = v2;
                        *reinterpret_cast<uint32_t *>(dst + COLS * 8 + 8) = v3;
                    }
                    if ((c & (GL - 1)) == 0)
                        gtab[(size_t)(c / GL) * MR + m] = make_int2(sum, mx >= 0x7F80u ? 0x7FC00000 : (E - 13) << 23);
                }
            }
        }
    }
    __syncthreads();
    stamp(2);

    // ---- main loop: stage `it` = 32 units; this warp owns units warp + 8 j (j = 0..3) of every stage ----
    constexpr uint32_t XU = 4 * 4 * COLS * 8;  // bytes of B fragments per unit
    constexpr uint32_t GU = MPU * MR * 8;      // bytes of group table per unit
    const uint32_t xs_l = gv_smem_u32(xs) + (uint32_t)((t * COLS + (g & (COLS - 1))) * 8);
    const uint32_t gt_l = gv_smem_u32(gtab) + (uint32_t)((t & (MR - 1)) * 8);
    const bool b_lane = COLS >= 8 || g < COLS;   // only the lanes of the data columns need real B values

    uint32_t bq[HOIST ? G2_UPW : 1][4][2];      // HOIST: B fragments of the warp's four k-units
    int gX[HOIST ? G2_UPW : 1];
    float gE[HOIST ? G2_UPW : 1];
    if constexpr (HOIST) {
#pragma unroll
        for (int j = 0; j < G2_UPW; ++j) {
            const uint32_t xk = xs_l + (uint32_t)(warp + 8 * j) * XU;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                bq[j][i][0] = 0; bq[j][i][1] = 0;
                if (b_lane) asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(bq[j][i][0]), "=r"(bq[j][i][1]) : "r"(xk + (uint32_t)(i * 4 * COLS * 8)));
            }
            int e2b;
            asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(gX[j]), "=r"(e2b) : "r"(gt_l + (uint32_t)(warp + 8 * j) * GU));
            gE[j] = __int_as_float(e2b);
        }
    }

    // refill of this warp's four units of the slot with stage it + DI (whole stages only: one predicate)
    const uint32_t offW = (uint32_t)(warp * GV_UNIT_BYTES + lane * 16);
    const uint32_t offM = STAGE_W + (uint32_t)(warp * MPU * META_BYTES);
    const uint8_t *nxt_w = p.qw + ((size_t)(a + DI * 32 + warp) * GV_UNIT_BYTES + lane * 16);
    const uint8_t *nxt_m = p.meta + ((size_t)(a + DI * 32 + warp) * (MPU * META_BYTES) + lane * 16);
    int nxt_st = DI;
    auto refill = [&](uint32_t sb) {
        const uint32_t ok = nxt_st < n_st, okM = ok && lane < MPU * 3;
        asm volatile(
            "{\n\t.reg .pred pa, pm;\n\t"
            "setp.ne.u32 pa, %4, 0;\n\t"
            "setp.ne.u32 pm, %5, 0;\n\t"
            "@pa cp.async.cg.shared.global [%0], [%2], 16;\n\t"
            "@pa cp.async.cg.shared.global [%0+512], [%2+512], 16;\n\t"
            "@pa cp.async.cg.shared.global [%0+8192], [%2+8192], 16;\n\t"
            "@pa cp.async.cg.shared.global [%0+8704], [%2+8704], 16;\n\t"
            "@pa cp.async.cg.shared.global [%0+16384], [%2+16384], 16;\n\t"
            "@pa cp.async.cg.shared.global [%0+16896], [%2+16896], 16;\n\t"
            "@pa cp.async.cg.shared.global [%0+24576], [%2+24576], 16;\n\t"
            "@pa cp.async.cg.shared.global [%0+25088], [%2+25088], 16;\n\t"
            "@pm cp.async.cg.shared.global [%1], [%3], 16;\n\t"
            "@pm cp.async.cg.shared.global [%1+%6], [%3+%6], 16;\n\t"
            "@pm cp.async.cg.shared.global [%1+%7], [%3+%7], 16;\n\t"
            "@pm cp.async.cg.shared.global [%1+%8], [%3+%8], 16;\n\t"
            "cp.async.commit_group;\n\t}"
            ::"r"(sb + offW), "r"(sb + offM + lane * 16), "l"(nxt_w), "l"(nxt_m), "r"(ok), "r"(okM),
              "n"(8 * MPU * META_BYTES), "n"(16 * MPU * META_BYTES), "n"(24 * MPU * META_BYTES) : "memory");
        nxt_w += 32 * GV_UNIT_BYTES;
        nxt_m += 32 * MPU * META_BYTES;
        ++nxt_st;
    };

    const int ipr = p.UPR >> 5;                // stages per row block
    float y0 = 0.f, y1 = 0.f;                  // rows g / g+8 of the current row block, token slot t
    uint32_t sb = ring;
    int slot = 0, it = 0;
    for (int s = 0; s < n_slots; ++s) {
        for (int m = 0; m < ipr; ++m, ++it) {
            if (it < DI) {
                gv_mbar_wait(bar_full + 8 * it, 0);
            } else {
                asm volatile("cp.async.wait_group %0;" ::"n"(DI - 1) : "memory");
                __syncwarp();                   // meta is read across lanes
            }
            uint32_t w[G2_UPW][8], s2[G2_UPW][MPU], zb[G2_UPW][MPU];
#pragma unroll
            for (int j = 0; j < G2_UPW; ++j) {
                const uint32_t sw = sb + offW + (uint32_t)(8 * j * GV_UNIT_BYTES), sm = sb + offM + (uint32_t)(8 * j * MPU * META_BYTES);
                asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(w[j][0]), "=r"(w[j][1]), "=r"(w[j][2]), "=r"(w[j][3]) : "r"(sw));
                asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4+512];" : "=r"(w[j][4]), "=r"(w[j][5]), "=r"(w[j][6]), "=r"(w[j][7]) : "r"(sw));
#pragma unroll
                for (int h = 0; h < MPU; ++h) {
                    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(s2[j][h]) : "r"(sm + h * META_BYTES + 4 * g));
                    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(zb[j][h]) : "r"(sm + h * META_BYTES + META_ZERO_OFF + g));
                }
            }
            __syncwarp();                       // every lane has read this warp's units: refill them with stage it + DI
            refill(sb);
            // 16 IMMAs as four interleaved chains (one per unit); a group's IMMAs accumulate in place
            int acc[G2_UPW][MPU][4];
#pragma unroll
            for (int j = 0; j < G2_UPW; ++j)
#pragma unroll
                for (int h = 0; h < MPU; ++h)
#pragma unroll
                    for (int q = 0; q < 4; ++q) acc[j][h][q] = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
#pragma unroll
                for (int j = 0; j < G2_UPW; ++j) {
                    const uint32_t wa = w[j][2 * i], wb = w[j][2 * i + 1];
                    uint32_t b0, b1;
                    if constexpr (HOIST) {
                        b0 = bq[j][i][0]; b1 = bq[j][i][1];
                    } else {
                        b0 = 0; b1 = 0;
                        if (b_lane)
                            asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(b0), "=r"(b1)
                                         : "r"(xs_l + (uint32_t)(32 * m + warp + 8 * j) * XU + (uint32_t)(i * 4 * COLS * 8)));
                    }
                    // rows g: the low nibbles; rows g+8: the RAW bytes 16 q[g+8] + q[g] (rows g are subtracted afterwards)
                    g2_imma(acc[j][i / IPG], wa & 0x0F0F0F0Fu, wa, wb & 0x0F0F0F0Fu, wb, b0, b1);
                }
            }
#pragma unroll
            for (int j = 0; j < G2_UPW; ++j) {
#pragma unroll
                for (int h = 0; h < MPU; ++h) {
                    int X;
                    float e2;
                    if constexpr (HOIST) {
                        X = gX[j]; e2 = gE[j];
                    } else {
                        int e2b;
                        asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(X), "=r"(e2b)
                                     : "r"(gt_l + (uint32_t)(32 * m + warp + 8 * j) * GU + (uint32_t)(h * MR * 8)));
                        e2 = __int_as_float(e2b);
                    }
                    // acc = {row g hi, row g lo, raw hi, raw lo} of token slot t; raw = 16 * row g+8 + row g
                    const int tg = acc[j][h][0] * 128 + acc[j][h][1];
                    const int tr = acc[j][h][2] * 128 + acc[j][h][3];
                    const int ug = tg - (int)(zb[j][h] & 0x0Fu) * X;
                    const int ug8 = tr - tg - (int)(zb[j][h] & 0xF0u) * X;          // 16 * (row g+8 - z8 * X)
                    const float slo = __uint_as_float(s2[j][h] << 16) * e2;
                    const float shi = __uint_as_float(s2[j][h] & 0xFFFF0000u) * (e2 * 0.0625f);
                    y0 = fmaf(slo, (float)ug, y0);
                    y1 = fmaf(shi, (float)ug8, y1);
                }
            }
            sb += STAGE;
            if (++slot == DI) { slot = 0; sb = ring; }
        }
        if (t < MR) part[(((size_t)s * G2_WARPS + warp) * 8 + g) * MR + t] = make_float2(y0, y1);
        y0 = 0.f; y1 = 0.f;
    }
    __syncthreads();
    stamp(3);

    // ---- combine the 8 warps (fixed tree: deterministic) and write y ----
    for (int i = tid; i < n_slots * 32 * MR; i += G2_THREADS) {         // multiples of 32: whole warps
        const int q = i & 3, e = (i >> 2) & (8 * MR - 1), sl = i / (32 * MR);
        const float2 *src = part + ((size_t)sl * G2_WARPS + 2 * q) * 8 * MR + e;
        float2 v = src[0];
        const float2 o = src[8 * MR];
        v.x += o.x; v.y += o.y;
        v.x += __shfl_xor_sync(0xffffffffu, v.x, 1); v.y += __shfl_xor_sync(0xffffffffu, v.y, 1);
        v.x += __shfl_xor_sync(0xffffffffu, v.x, 2); v.y += __shfl_xor_sync(0xffffffffu, v.y, 2);
        const int tok = e % MR, n0 = (R_first + sl) * 16 + e / MR;
        if (q == 0 && tok < p.M) {
            const float b0 = p.bias ? __bfloat162float(p.bias[n0]) : 0.f;
            const float b1 = p.bias ? __bfloat162float(p.bias[n0 + 8]) : 0.f;
            if (p.swiglu) {   // rows g / g+8 are (gate, up) of output column 8 R + g
                p.y[(size_t)tok * (p.N >> 1) + (n0 >> 4) * 8 + (n0 & 7)] = swiglu_bf16(v.x + b0, v.y + b1);
            } else {
                p.y[(size_t)tok * p.N + n0] = __float2bfloat16_rn(v.x + b0);
                p.y[(size_t)tok * p.N + n0 + 8] = __float2bfloat16_rn(v.y + b1);
            }
        }
    }
    stamp(4);
    stamp(5);
}

// ---- host side ---------------------------------------------------------------------------
struct Imma2Plan {
    bool ok;
    int MR, mpu, S;
    int off_xs, off_gtab, off_part, off_wring, smem_bytes;
    bool hoist;
};

static Imma2Plan imma2_plan(const Linear &L, int64_t M) {
    Imma2Plan pl{};
    pl.ok = false;
    if (M < 1 || M > 4 || L.gv_grid <= 0 || !L.gv_tab || !L.gv_aligned) return pl;
    // K = 4096 only: one 32-unit stage per row block.  Larger multiples of 4096 run correctly (the loop is general)
    // but measured slower than the first kernel: with 256 threads the activation staging of K = 12288 takes 2.0 us
    // against 1.35 us with 512 (Qwen3-8B down_proj: 7.2 us against 6.9 us per launch).
    if (L.d.K != 32 * GV_UNIT_K) return pl;
    const int gs = L.d.gs;
    if (gs != 32 && gs != 64 && gs != 128) return pl;
    pl.mpu = 128 / gs;
    pl.MR = M <= 1 ? 1 : M <= 2 ? 2 : 4;
    pl.hoist = pl.mpu == 1;
    int64_t off = 128;                                               // mbarriers
    pl.off_xs = (int)off; off += L.d.K * 2 * pl.MR;
    pl.off_gtab = (int)off; off += (L.d.K / gs) * pl.MR * 8;
    off = (off + 15) & ~(int64_t)15;
    pl.off_part = (int)off; off += (int64_t)L.gv_max_slots * G2_WARPS * 8 * pl.MR * 8;
    off = (off + 127) & ~(int64_t)127;
    const int64_t stage = 32 * (GV_UNIT_BYTES + pl.mpu * META_BYTES);
    int64_t D = (GV_SMEM_2CTA - off) / stage;                        // two CTAs per SM: this linear + the next one's prefetch
    if (D < 2) return pl;                                            // the first kernel (one CTA per SM) takes it
    D = std::min<int64_t>(D, 3);
    pl.S = (int)D;
    pl.off_wring = (int)off; off += D * stage;
    pl.smem_bytes = (int)off;
    pl.ok = true;
    return pl;
}

bool gemv_imma2_usable(const Linear &L, int64_t M) {
    if (exp_env_int("W4A16_NO_GEMV_IMMA2", 0)) return false;
    return imma2_plan(L, M).ok;
}

template <int MR, int MPU, int DI, bool HOIST>
static int g2_launch_t(const GemvParams &p, const Imma2Plan &pl, int grid, int device, cudaStream_t stream, bool pdl) {
    static bool attr_set[64] = {};   // per device; racing threads only repeat the call
    if (device >= 0 && device < 64 && !attr_set[device]) {
        W4_CUDA_OK(cudaFuncSetAttribute(gemv_imma2_kernel<MR, MPU, DI, HOIST>, cudaFuncAttributeMaxDynamicSharedMemorySize, GV_SMEM_2CTA));
        attr_set[device] = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(G2_THREADS);
    cfg.dynamicSmemBytes = (size_t)pl.smem_bytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    W4_CUDA_OK(cudaLaunchKernelEx(&cfg, gemv_imma2_kernel<MR, MPU, DI, HOIST>, p));
    count_launch();
    return 0;
}

int launch_gemv_imma2(const Linear &L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias, cudaStream_t stream) {
    const Imma2Plan pl = imma2_plan(L, M);
    if (!pl.ok) return fail("launch_gemv_imma2: shape not supported");
    GemvParams p{};
    p.qw = reinterpret_cast<const uint8_t *>(L.qw);
    p.meta = L.meta;
    p.bias = add_bias && L.has_bias ? L.bias : nullptr;
    p.x = x; p.y = y;
    p.tab = reinterpret_cast<const int4 *>(L.gv_tab);
    p.M = (int)M; p.K = (int)L.d.K; p.N = (int)L.d.N;
    p.swiglu = L.epilogue == W4A16_EPILOGUE_SWIGLU;
    p.UPR = (int)(L.d.K / GV_UNIT_K);
    p.meta_stride = pl.mpu * META_BYTES;
    p.n_stages_ring = pl.S;
    p.off_xs = pl.off_xs; p.off_gtab = pl.off_gtab; p.off_part = pl.off_part; p.off_wring = pl.off_wring;
    p.dbg = gemv_debug_slot();
    const bool pdl = gv_env_int("W4A16_NO_PDL", 0) == 0;
#define W4_G2D(MR_, MPU_, HOIST_)                                                                       \
    do {                                                                                               \
        if (pl.S == 3) return g2_launch_t<MR_, MPU_, 3, HOIST_>(p, pl, L.gv_grid, L.device, stream, pdl); \
        return g2_launch_t<MR_, MPU_, 2, HOIST_>(p, pl, L.gv_grid, L.device, stream, pdl);             \
    } while (0)
#define W4_G2(MR_)                                                                                     \
    do {                                                                                               \
        if (pl.mpu == 1) { if (pl.hoist) W4_G2D(MR_, 1, true); W4_G2D(MR_, 1, false); }                \
        if (pl.mpu == 2) W4_G2D(MR_, 2, false);                                                        \
        W4_G2D(MR_, 4, false);                                                                         \
    } while (0)
    switch (pl.MR) {
        case 1: W4_G2(1);
        case 2: W4_G2(2);
        default: W4_G2(4);
    }
#undef W4_G2
#undef W4_G2D
}

}  // namespace w4
