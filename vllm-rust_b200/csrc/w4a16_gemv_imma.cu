// w4a16_gemv_imma.cu -- decode GEMV for M <= 4 tokens on the int8 tensor-core path (IMMA.16832.U8.S8).
//
// Replaces awq_gemv_int4_kt_bf16 / kt_mloop_m4 (awq_gemv.cu:377-665) and gptq_gemm_int4_bf16_tiled
// (gptq_dequant.cu:243-318) as dispatched by marlin_gemm_pooled / marlin_gemm (marlin_cuda.rs:843-953,
// 1655-1753).  One CTA of 16 warps per SM owns a contiguous range of whole row blocks of the W4T stream and pulls it
// through a shared-memory ring of DI stages of 32 units (32 KB of weights + their meta blocks) with cp.async.bulk,
// the first DI stages issued BEFORE griddepcontrol.wait; in a stage warp w owns units w and w + 16.
// Measured (tools/ubench/pdlstream.cu, tmalat.cu): cp.async.bulk executes ONE request at a time per SM (~0.4 us
// each), so 8 KB requests cap the chip at 3.4 TB/s, 16 KB at 5.6, 32 KB reach 6.2-6.5; a per-warp cp.async (LDGSTS)
// ring sustains 6.4-6.8 TB/s but costs 8 LSU cycles per 512 B, which the co-resident CTA of the NEXT kernel spends
// exactly while this one stages its activations.  Hence 32 KB bulk requests.
//
// Arithmetic (the reference's GEMV form, awq_gemv.cu:597-664, regrouped so the zero point leaves the loop):
//     y[m,n] = sum_g s[g,n] * 2^e[g,m] * ( sum_{k in g} q[k,n] * xt[m,k]  -  z[g,n] * X[g,m] )
//   xt[m,k] = round(x[m,k] * 2^-e[g,m]) in [-16383, 16383]  (per group and token the power of two puts max|x| in
//             [2^13, 2^14): exact for every element within 2^6 of the group's maximum, |error| <= 2^-14 of that
//             maximum below), carried as two signed-byte planes xt = 128*hi + lo;
//   X[g,m]  = sum_{k in g} xt[m,k]  (int32, once per CTA while the first weights are in flight).
// One W4T word holds 8 nibbles = rows {g, g+8} x 4 k:  `w & 0x0F0F0F0F` IS the u8x4 A-fragment register of row g,
// `w & 0xF0F0F0F0` the one of row g+8 scaled by 16 (undone exactly in the scale): 2 LOP3 per 8 weights, no
// shift.  Two words make the four A registers of mma.m16n8k32.u8.s8; the 8 B columns are (hi, lo) planes of up to
// 4 tokens; accumulation is exact int32.  Per 1 KB unit (2048 weights) a warp issues 2 LDS.128 + 16 LOP3 +
// 4 LDS.64 + 4 IMMA + ~20 instructions of group epilogue: about half of the dp4a loop it replaces.
#include <algorithm>

#include "w4a16_gemv_common.cuh"

namespace w4 {

__device__ __forceinline__ void gi_imma(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                        uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// MR = token slots (1, 2 or 4; M <= MR).  MPU = groups per 128-k unit (1, 2, 4 for group size 128, 64, 32).
// DI = loop iterations in flight in the warp's ring; one iteration = 2 units (two independent IMMA chains per warp
// hide the tensor-pipe and shared-memory latencies that 4 warps per scheduler cannot).
// REG = every row block is a whole number of 32-unit strides (K % 4096 == 0): the two units of an iteration share
// their row block and every warp visits every row block, so the loop carries no row-block bookkeeping.
// FUSED = the epilogue all-reduces row-parallel partial sums over the peers' symmetric buffers (p.peer); a separate
// instantiation, so the single-GPU kernel carries none of it.
template <int MR, int MPU, int DI, bool REG, bool FUSED>
__global__ void __launch_bounds__(GV_THREADS, MR <= 4 ? 2 : 1) gemv_imma_kernel(const GemvParams p) {
    constexpr int COLS = 2 * MR;               // B columns holding data: (hi, lo) plane of each token slot
    constexpr int NT = MR <= 4 ? 1 : MR / 4;   // IMMA token tiles (8 columns = 4 tokens x 2 planes each); lane t of tile nt = token 4 nt + t
    constexpr int IPG = 4 / MPU;               // IMMAs per group
    constexpr int SUB_K = GV_UNIT_K / MPU;     // k per group inside a unit
    constexpr int GL = SUB_K / 8;              // staging lanes (8-element pieces) per group
    extern __shared__ __align__(128) uint8_t gv_smem[];
    uint8_t *xs = gv_smem + p.off_xs;                                   // [K/32][4 t][COLS][8 B]: B fragments
    int2 *gtab = reinterpret_cast<int2 *>(gv_smem + p.off_gtab);        // [K/SUB_K][MR] {X, bits of 2^e}
    float2 *part = reinterpret_cast<float2 *>(gv_smem + p.off_part);    // [slots][GV_WARPS][8 g][MR] {row g, row g+8}

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int cta = blockIdx.x;
    const int4 rec0 = __ldg(p.tab + 2 * cta);
    const int a = rec0.x, b = rec0.y;          // unit range: whole row blocks
    const int R_first = a / p.UPR, n_slots = (b - a) / p.UPR;
    const int n_st = (b - a + 2 * GV_WARPS - 1) / (2 * GV_WARPS);      // stages of 32 units
    constexpr uint32_t STAGE_W = 2 * GV_WARPS * GV_UNIT_BYTES, STAGE_M = 2 * GV_WARPS * MPU * META_BYTES;
    constexpr uint32_t STAGE = STAGE_W + STAGE_M;
    const uint32_t ring = gv_smem_u32(gv_smem + p.off_wring);
    const uint32_t bar_full = gv_smem_u32(gv_smem);                    // DI mbarriers at the start of smem

    // The first DI stages: one bulk copy of the weights + one of the meta blocks per stage, on full[stage] (no LSU
    // issue slots: the CTA of the previous kernel on this SM is staging its activations right now).
    auto issue_stage = [&](int j, int slot) {
        const int u_lo = a + j * (2 * GV_WARPS);
        const int cnt = min(2 * GV_WARPS, b - u_lo);
        const uint32_t wb = (uint32_t)cnt * GV_UNIT_BYTES, mb = (uint32_t)cnt * (MPU * META_BYTES);
        const uint32_t bar = bar_full + 8 * slot, dst = ring + (uint32_t)slot * STAGE;
        gv_mbar_expect_tx(bar, wb + mb);
        gv_bulk_g2s(dst, p.qw + (size_t)u_lo * GV_UNIT_BYTES, wb, bar);
        gv_bulk_g2s(dst + STAGE_W, p.meta + (size_t)u_lo * (MPU * META_BYTES), mb, bar);
    };
    // Every later stage: each warp refills ITS OWN two units of the slot with cp.async (16 B per lane: a lane copies
    // exactly the pieces it later reads; only this warp ever read the bytes it overwrites, so a __syncwarp orders the
    // reuse).  One commit group per call, empty past the end; running pointers, predicated copies, no branches.
    const uint32_t offW = (uint32_t)(warp * GV_UNIT_BYTES + lane * 16);
    const uint32_t offM = STAGE_W + (uint32_t)(warp * MPU * META_BYTES);
    const uint8_t *nxt_w = p.qw + ((size_t)(a + DI * 2 * GV_WARPS + warp) * GV_UNIT_BYTES + lane * 16);
    const uint8_t *nxt_m = p.meta + ((size_t)(a + DI * 2 * GV_WARPS + warp) * (MPU * META_BYTES) + lane * 16);
    int nxt_u = a + DI * 2 * GV_WARPS + warp;  // unit A of the next stage to refill
    auto refill = [&](uint32_t sb) {
        const uint32_t okA = nxt_u < b, okB = nxt_u + GV_WARPS < b, okM = lane < MPU * 3;
        asm volatile(
            "{\n\t.reg .pred pa, pb, pma, pmb;\n\t"
            "setp.ne.u32 pa, %4, 0;\n\t"
            "setp.ne.u32 pb, %5, 0;\n\t"
            "setp.ne.u32 pma, %6, 0;\n\t"
            "and.pred pmb, pma, pb;\n\t"
            "and.pred pma, pma, pa;\n\t"
            "@pa cp.async.cg.shared.global [%0], [%2], 16;\n\t"
            "@pa cp.async.cg.shared.global [%0+512], [%2+512], 16;\n\t"
            "@pb cp.async.cg.shared.global [%0+16384], [%2+16384], 16;\n\t"
            "@pb cp.async.cg.shared.global [%0+16896], [%2+16896], 16;\n\t"
            "@pma cp.async.cg.shared.global [%1], [%3], 16;\n\t"
            "@pmb cp.async.cg.shared.global [%1+%7], [%3+%7], 16;\n\t"
            "cp.async.commit_group;\n\t}"
            ::"r"(sb + offW), "r"(sb + offM + lane * 16), "l"(nxt_w), "l"(nxt_m), "r"(okA), "r"(okB), "r"(okM),
              "n"(GV_WARPS * MPU * META_BYTES) : "memory");
        nxt_w += 2 * GV_WARPS * GV_UNIT_BYTES;
        nxt_m += 2 * GV_WARPS * MPU * META_BYTES;
        nxt_u += 2 * GV_WARPS;
    };

    // ---- prefetch: weights and meta are immutable, independent of the previous kernel ----
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < DI; ++s) gv_mbar_init(bar_full + 8 * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#pragma unroll
        for (int s = 0; s < DI; ++s) if (s < n_st) issue_stage(s, s);
    }
    if (!REG)   // a warp may never visit some row block of the range (REG: every warp writes every slot)
        for (int i = tid; i < n_slots * GV_WARPS * 8 * MR; i += GV_THREADS) part[i] = make_float2(0.f, 0.f);
    auto stamp = [&](int i) {
        if (p.dbg && tid == 0) {
            unsigned long long tt;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tt));
            p.dbg[(size_t)cta * 8 + i] = tt;
        }
    };
    stamp(0);
    // A launch that ends with the cross-GPU exchange releases its dependents only at exit: with the next linear's
    // CTAs co-resident during the exchange the TP=2 step measured 1.21 ms against 1.02 ms without (bench.py --gpus 2).
    if (!FUSED || (p.peer.flags & 1)) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    stamp(1);
    // fused all-reduce: this launch's epoch (the previous launch's last CTA has stored the counter; read it now so
    // that the load is off the epilogue's critical path)
    constexpr bool fused = FUSED;
    uint32_t epoch = 0u;
    if constexpr (FUSED) epoch = __ldcg(p.peer.epoch_ptr) + 1u;

    // ---- stage the activations: per-(group, token) power-of-two scale, two signed-byte planes in B-fragment
    //      order, integer group sums.  x is [M][K] contiguous, so 8-element piece i sits at x + 8 i. ----
    {
        const int ppr = p.K >> 3, total = p.M * ppr;
        const uint4 *x16 = reinterpret_cast<const uint4 *>(p.x);
        constexpr int BATCH = 4;               // loads in flight per thread
        for (int base = 0; base < total; base += BATCH * GV_THREADS) {   // warp-uniform trip count (shuffles)
            uint4 v[BATCH];
#pragma unroll
            for (int r = 0; r < BATCH; ++r) {
                const int i = base + r * GV_THREADS + tid;
                v[r] = i < total ? __ldg(x16 + i) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int r = 0; r < BATCH; ++r) {
                const int i = base + r * GV_THREADS + tid;
                if (base + r * GV_THREADS >= total) break;               // warp-uniform
                const uint32_t w4[4] = {v[r].x, v[r].y, v[r].z, v[r].w};
                uint32_t mx = 0;
#pragma unroll
                for (int q = 0; q < 4; ++q) mx = max(mx, max(w4[q] & 0x7FFFu, (w4[q] >> 16) & 0x7FFFu));
                mx = gv_group_max<GL>(mx, lane);
                int E = (int)(mx >> 7);                                  // biased exponent of the group's max |x|
                E = min(max(E, 14), 254);
                const float scale = __int_as_float((267 - E) << 23);     // max|x| lands in [2^13, 2^14)
                // round to integers with the 1.5 * 2^23 trick: the low 16 bits of the sum's pattern are xt (two's
                // complement), so one PRMT packs a pair; planes and the group sum then work on s16x2 words
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
                    const int m = MR == 1 ? 0 : MR > 4 ? i / ppr : (i >= ppr) + (MR > 2 ? (i >= 2 * ppr) + (i >= 3 * ppr) : 0);
                    const int c = i - m * ppr;                           // token, piece within the row
                    // k = 8c: unit ku = c >> 4, word j = (c & 15) >> 1, IMMA j >> 1 (half j & 1), quads t = 2(c&1), +1
                    const int j = (c & 15) >> 1;
                    uint8_t *dst = xs + ((size_t)(((c >> 4) * 4 + (j >> 1)) * 4 + 2 * (c & 1)) * COLS + 2 * m) * 8 +
                                   (j & 1) * 4;
                    // xt = 128 hi + lo per 16-bit half; pair (k0,k1) | pair (k2,k3) << 8 gives the byte order
                    // (k0, k2, k1, k3) of the bytes of `w & 0x0F0F0F0F`
                    uint32_t hiP[4], loP[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) { hiP[q] = (pk[q] >> 7) & 0x00FF00FFu; loP[q] = pk[q] & 0x007F007Fu; }
                    uint32_t v0 = hiP[0] | (hiP[1] << 8), v1 = loP[0] | (loP[1] << 8);      // -> dst, dst + 8
                    uint32_t v2 = hiP[2] | (hiP[3] << 8), v3 = loP[2] | (loP[3] << 8);      // -> dst + COLS*8, dst + COLS*8 + 8
                    if constexpr (MR >= 2) {
                        // All 32 pieces of a warp land in rows that are multiples of 128 B apart: issued in the same
                        // order, each of the four stores is a 16-way bank conflict (~2 us per CTA at 4 tokens).  Lanes
                        // rotate the order of their four stores by (lane >> 2) & 3, which spreads them over 4x the banks.
                        const int rot = (lane >> 2) & 3;
                        if (rot & 1) { const uint32_t tv = v0; v0 = v1; v1 = v2; v2 = v3; v3 = tv; }
                        if (rot & 2) { uint32_t tv = v0; v0 = v2; v2 = tv; tv = v1; v1 = v3; v3 = tv; }
                        const uint32_t vv[4] = {v0, v1, v2, v3};
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const int idx = (k + rot) & 3;                                  // which of the four slots vv[k] is
                            *reinterpret_cast<uint32_t *>(dst + (idx & 1) * 8 + (idx >> 1) * (COLS * 8)) = vv[k];
                        }
                    } else {
                        *reinterpret_cast<uint32_t *>(dst) = v0;
                        *reinterpret_cast<uint32_t *>(dst + 8) = v1;
                        *reinterpret_cast<uint32_t *>(dst + COLS * 8) = v2;
                        *reinterpret_cast<uint32_t *>(dst + COLS * 8 + 8) = v3;
                    }
                    // Inf / NaN anywhere in the (group, token): the scale becomes NaN, so every output of that token is
                    // NaN (non-finite in -> non-finite out, like the reference's fp32 FMA chain, awq_gemv.cu:597-664)
                    if ((c & (GL - 1)) == 0)
                        gtab[(size_t)(c / GL) * MR + m] = make_int2(sum, mx >= 0x7F80u ? 0x7FC00000 : (E - 13) << 23);
                }
            }
        }
    }
    __syncthreads();
    stamp(2);

    // ---- main loop: iteration `it` handles this warp's units A = 2 it and B = 2 it + 1 (stream units a + warp + 32 it, + 16) ----
    constexpr uint32_t XU = 4 * 4 * COLS * 8;  // bytes of B fragments per unit
    constexpr uint32_t GU = MPU * MR * 8;      // bytes of group table per unit
    float y0[NT], y1[NT];                      // rows g / g+8 of the current row block, token slot 4 nt + t
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) { y0[nt] = 0.f; y1[nt] = 0.f; }
    const uint32_t xs_l = gv_smem_u32(xs) + (uint32_t)((t * COLS + (g & (COLS - 1))) * 8);   // + 64 nt: columns 8 nt + g
    const uint32_t gt_l = gv_smem_u32(gtab) + (uint32_t)((t & (MR - 1)) * 8);                // + 32 nt: token 4 nt + t
    // integer part of one unit: 4 IMMAs into MPU group accumulators (xk = this lane's B fragments of the unit)
    auto unit_mma = [&](const uint32_t (&wj)[8], uint32_t xk, int (&acc)[NT][MPU][4]) {
#pragma unroll
        for (int h = 0; h < MPU; ++h) {
            int c1[NT][4];                     // second accumulator chain when a group spans several IMMAs
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int q = 0; q < 4; ++q) { acc[nt][h][q] = 0; c1[nt][q] = 0; }
#pragma unroll
            for (int ii = 0; ii < IPG; ++ii) {
                const int i = h * IPG + ii;
                const uint32_t wa = wj[2 * i], wb = wj[2 * i + 1];
                // rows g: the low nibbles; rows g+8: the RAW bytes 16 q[g+8] + q[g] (rows g are subtracted afterwards)
                const uint32_t la = wa & 0x0F0F0F0Fu, lb = wb & 0x0F0F0F0Fu;
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    uint32_t b0 = 0, b1 = 0;
                    // only the lanes of the COLS data columns need real B values: fewer shared-memory wavefronts
                    if (COLS >= 8 || g < COLS)
                        asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(b0), "=r"(b1) : "r"(xk + (uint32_t)(i * 4 * COLS * 8 + nt * 64)));
                    if (ii & 1) gi_imma(c1[nt], la, wa, lb, wb, b0, b1);
                    else gi_imma(acc[nt][h], la, wa, lb, wb, b0, b1);
                }
            }
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                if (IPG > 1) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) acc[nt][h][q] += c1[nt][q];
                }
                acc[nt][h][2] -= acc[nt][h][0];    // 16 * sum(q[g+8] x) = raw - sum(q[g] x)
                acc[nt][h][3] -= acc[nt][h][1];
            }
        }
    };
    // float part: acc = {row g hi, row g lo, row g+8 hi, row g+8 lo} of token slot t; rows g+8 carry a factor 16
    auto unit_fin = [&](const int (&acc)[NT][MPU][4], const uint32_t (&s2)[MPU], const uint32_t (&zb)[MPU], uint32_t gk) {
#pragma unroll
        for (int h = 0; h < MPU; ++h) {
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                int X, e2b;
                asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(X), "=r"(e2b) : "r"(gk + (uint32_t)(h * MR * 8 + nt * 32)));
                const float e2 = __int_as_float(e2b);
                const int tg = acc[nt][h][0] * 128 + acc[nt][h][1] - (int)(zb[h] & 0x0Fu) * X;
                const int tg8 = acc[nt][h][2] * 128 + acc[nt][h][3] - (int)(zb[h] & 0xF0u) * X;
                const float slo = __uint_as_float(s2[h] << 16) * e2;
                const float shi = __uint_as_float(s2[h] & 0xFFFF0000u) * (e2 * 0.0625f);
                y0[nt] = fmaf(slo, (float)tg, y0[nt]);
                y1[nt] = fmaf(shi, (float)tg8, y1[nt]);
            }
        }
    };
    auto load_unit = [&](uint32_t sw, uint32_t sm, uint32_t (&w)[8], uint32_t (&s2)[MPU], uint32_t (&zb)[MPU]) {
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "r"(sw));
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4+512];" : "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "r"(sw));
#pragma unroll
        for (int h = 0; h < MPU; ++h) {
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(s2[h]) : "r"(sm + h * META_BYTES + 4 * g));
            asm volatile("ld.shared.u8 %0, [%1];" : "=r"(zb[h]) : "r"(sm + h * META_BYTES + META_ZERO_OFF + g));
        }
    };
    // stage `it` is complete in its slot: the first DI came by bulk copy (mbarrier), the rest by this warp's cp.async
    auto wait_stage = [&](int it) {
        if (it < DI) {
            gv_mbar_wait(bar_full + 8 * it, 0);
        } else {
            asm volatile("cp.async.wait_group %0;" ::"n"(DI - 1) : "memory");
            __syncwarp();                       // meta is read across lanes
        }
    };
    if constexpr (REG) {
        // every iteration: units (R, ku = warp + 32 m) and (R, ku + 16); UPR / 32 iterations per row block
        const int ipr = p.UPR >> 5;
        uint32_t sb = ring;                    // base of the current ring slot
        int slot = 0, it = 0;
        for (int s = 0; s < n_slots; ++s) {
            uint32_t xk = xs_l + (uint32_t)warp * XU, gk = gt_l + (uint32_t)warp * GU;
            for (int m = 0; m < ipr; ++m, ++it) {
                wait_stage(it);
                if (it == 0 && warp == 0) stamp(6);
                uint32_t wA[8], wB[8], s2A[MPU], zbA[MPU], s2B[MPU], zbB[MPU];
                load_unit(sb + offW, sb + offM, wA, s2A, zbA);
                load_unit(sb + offW + GV_WARPS * GV_UNIT_BYTES, sb + offM + GV_WARPS * MPU * META_BYTES, wB, s2B, zbB);
                __syncwarp();                   // every lane has read this warp's units: refill them with stage it + DI
                refill(sb);
                int accA[NT][MPU][4], accB[NT][MPU][4];
                unit_mma(wA, xk, accA);
                unit_mma(wB, xk + GV_WARPS * XU, accB);
                unit_fin(accA, s2A, zbA, gk);
                unit_fin(accB, s2B, zbB, gk + GV_WARPS * GU);
                xk += 2 * GV_WARPS * XU; gk += 2 * GV_WARPS * GU;
                sb += STAGE;
                if (++slot == DI) { slot = 0; sb = ring; }
            }
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                if (t < MR) part[(((size_t)s * GV_WARPS + warp) * 8 + g) * MR + 4 * nt + t] = make_float2(y0[nt], y1[nt]);
                y0[nt] = 0.f; y1[nt] = 0.f;
            }
        }
    } else {
    int R_A = (a + warp) / p.UPR, ku_A = (a + warp) - R_A * p.UPR;
    int R_acc = -1;
    auto flush = [&](int R) {
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            if (t < MR) part[(((size_t)(R - R_first) * GV_WARPS + warp) * 8 + g) * MR + 4 * nt + t] = make_float2(y0[nt], y1[nt]);
            y0[nt] = 0.f; y1[nt] = 0.f;
        }
    };
    int slot = 0;
    for (int it = 0; it < n_st; ++it) {
        wait_stage(it);
        if (it == 0 && warp == 0) stamp(6);
        const int uA = a + it * 2 * GV_WARPS + warp;
        const bool has_A = uA < b, has_B = uA + GV_WARPS < b;    // warp-uniform
        int R_B = R_A, ku_B = ku_A;
        if (has_B) { ku_B += GV_WARPS; while (ku_B >= p.UPR) { ku_B -= p.UPR; ++R_B; } }
        const uint32_t sb = ring + (uint32_t)slot * STAGE;
        uint32_t wA[8], wB[8], s2A[MPU], zbA[MPU], s2B[MPU], zbB[MPU];
        if (has_A) {
            load_unit(sb + offW, sb + offM, wA, s2A, zbA);
        } else {                                // past the range: zero weights with zero scales contribute nothing
#pragma unroll
            for (int q = 0; q < 8; ++q) wA[q] = 0;
#pragma unroll
            for (int h = 0; h < MPU; ++h) { s2A[h] = 0; zbA[h] = 0; }
        }
        if (has_B) {
            load_unit(sb + offW + GV_WARPS * GV_UNIT_BYTES, sb + offM + GV_WARPS * MPU * META_BYTES, wB, s2B, zbB);
        } else {
#pragma unroll
            for (int q = 0; q < 8; ++q) wB[q] = 0;
#pragma unroll
            for (int h = 0; h < MPU; ++h) { s2B[h] = 0; zbB[h] = 0; }
        }
        __syncwarp();
        refill(sb);
        if (!has_A) {                           // nothing of this stage belongs to this warp (only in the last stage)
            if (++slot == DI) slot = 0;
            continue;
        }
        int accA[NT][MPU][4], accB[NT][MPU][4];
        unit_mma(wA, xs_l + (uint32_t)ku_A * XU, accA);
        unit_mma(wB, xs_l + (uint32_t)ku_B * XU, accB);
        if (R_acc != R_A) { if (R_acc >= 0) flush(R_acc); R_acc = R_A; }
        unit_fin(accA, s2A, zbA, gt_l + (uint32_t)ku_A * GU);
        if (R_acc != R_B) { flush(R_acc); R_acc = R_B; }
        unit_fin(accB, s2B, zbB, gt_l + (uint32_t)ku_B * GU);
        ku_A += 2 * GV_WARPS;
        while (ku_A >= p.UPR) { ku_A -= p.UPR; ++R_A; }
        if (++slot == DI) slot = 0;
    }
    if (R_acc >= 0) flush(R_acc);
    }
    if (warp == 0) stamp(7);
    __syncthreads();
    stamp(3);

    // ---- combine the 16 warps (fixed tree: deterministic) and write y ----
    // thread (item, q): item = (slot, row g, token), q = quarter of the warps; quarters are adjacent lanes
    //
    // Fused row-parallel all-reduce (p.peer.world > 1), "Lamport" style: every rank owns three rotating buffers
    // [3][world sources][cap] of fp32 words that hold the sentinel -0.0f (PEER_EMPTY) when empty.  Launch e stores
    // this rank's partial sums (with -0.0f canonicalised to +0.0f) into slot [e % 3][rank] of EVERY rank over NVLink,
    // polls its own slots [e % 3][0..world) until no word is the sentinel, sums the sources in rank order (the same
    // order on every rank: bit-identical results everywhere), adds the bias and writes bf16.  The data words are
    // their own arrival flags: no fence, no flag round trip, no barrier between ranks.  Only the words of real tokens
    // (tok < M) are stored, every stored word is polled by exactly one thread, and that thread puts the sentinel back
    // as soon as it has read the word: a launch leaves its buffer clean whatever N and M the neighbouring launches
    // use (one PeerGroup serves row-parallel linears of different N).  The next store into such a word belongs to
    // launch e + 3, which a peer issues only after it has seen this rank's launch e + 2 partials -- two kernel
    // boundaries after the reset.  All ranks run the same partition, so CTA c owns the same rows everywhere and the
    // exchange is purely per CTA.
    float *own = nullptr;
    size_t cur_off = 0;
    if constexpr (FUSED) {
        own = p.peer.base[p.peer.rank & (PEER_MAX_WORLD - 1)];
        cur_off = (size_t)(epoch % 3u) * ((size_t)p.peer.world * p.peer.cap);
    }
    for (int it = tid; it < n_slots * 32 * MR; it += GV_THREADS) {      // multiples of 32: whole warps
        const int q = it & 3, e = (it >> 2) & (8 * MR - 1), slot = it / (32 * MR);
        const float2 *src = part + ((size_t)slot * GV_WARPS + 4 * q) * 8 * MR + e;
        float2 v = src[0];
#pragma unroll
        for (int w = 1; w < 4; ++w) { const float2 o = src[w * 8 * MR]; v.x += o.x; v.y += o.y; }
        v.x += __shfl_xor_sync(0xffffffffu, v.x, 1); v.y += __shfl_xor_sync(0xffffffffu, v.y, 1);
        v.x += __shfl_xor_sync(0xffffffffu, v.x, 2); v.y += __shfl_xor_sync(0xffffffffu, v.y, 2);
        const int tok = e % MR, n0 = (R_first + slot) * 16 + e / MR;
        if constexpr (FUSED) {
            uint32_t bx = __float_as_uint(v.x), by = __float_as_uint(v.y);
            if (bx == PEER_EMPTY) bx = 0u;
            if (by == PEER_EMPTY) by = 0u;
            for (int r = q; r < p.peer.world && tok < p.M; r += 4) {    // lane q serves peers q, q + 4; real tokens only
                uint32_t *dst = reinterpret_cast<uint32_t *>(p.peer.base[r] + cur_off + (size_t)p.peer.rank * p.peer.cap);
                asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(dst + (size_t)n0 * PEER_TOK + tok), "r"(bx) : "memory");
                asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(dst + (size_t)(n0 + 8) * PEER_TOK + tok), "r"(by) : "memory");
            }
        } else if (q == 0 && tok < p.M) {
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
    if constexpr (FUSED) {
        unsigned long long t0;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        for (int it = tid; it < n_slots * 16 * MR; it += GV_THREADS) {
            const int tok = it % MR;
            const int n = R_first * 16 + it / MR;
            if (tok >= p.M) continue;                                   // padding token slots are never read
            const uint32_t *srcw = reinterpret_cast<const uint32_t *>(own + cur_off) + (size_t)n * PEER_TOK + tok;
            // all sources' words are requested together and re-read until none is the sentinel (one NVLink latency,
            // not `world` of them in sequence)
            uint32_t w[PEER_MAX_WORLD];
            uint32_t spins = 0;
            for (;;) {
                bool all = true;
#pragma unroll
                for (int r = 0; r < PEER_MAX_WORLD; ++r) {
                    w[r] = 0u;
                    if (r < p.peer.world)
                        asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(w[r]) : "l"(srcw + (size_t)r * p.peer.cap) : "memory");
                }
#pragma unroll
                for (int r = 0; r < PEER_MAX_WORLD; ++r) all = all && w[r] != PEER_EMPTY;
                if (all) break;
                if ((++spins & 1023u) == 0) {
                    unsigned long long t1;
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                    if (t1 - t0 > 20000000000ull) __trap();   // 20 s: a peer never arrived -- fail loudly instead of hanging
                }
            }
            float acc = 0.f;
#pragma unroll
            for (int r = 0; r < PEER_MAX_WORLD; ++r) {
                if (r < p.peer.world) {
                    acc += __uint_as_float(w[r]);               // rank order: bit-identical on every rank
                    const_cast<uint32_t *>(srcw)[(size_t)r * p.peer.cap] = PEER_EMPTY;   // consumed: empty again for launch e + 3
                }
            }
            const float bsv = p.bias ? __bfloat162float(p.bias[n]) : 0.f;
            p.y[(size_t)tok * p.N + n] = __float2bfloat16_rn(acc + bsv);
        }
        // the last CTA of this launch to finish advances the rank's epoch (every CTA read it right after its
        // griddepcontrol.wait, before its own ticket; the next launch reads it after ITS griddepcontrol.wait)
        __syncthreads();
        if (tid == 0) {
            __threadfence();
            const int prev = atomicAdd(p.peer.ticket, 1);
            if (prev == (int)gridDim.x - 1) {
                *p.peer.ticket = 0;
                *p.peer.epoch_ptr = epoch;
                __threadfence();
            }
        }
    }
    stamp(4);
    stamp(5);
}

// ---- host side ---------------------------------------------------------------------------
struct ImmaPlan {
    bool ok;
    int MR, mpu, S, stage_m_bytes;
    int off_xs, off_gtab, off_part, off_wring, off_mring, smem_bytes;
};

static ImmaPlan imma_plan(const Linear &L, int64_t M) {
    ImmaPlan pl{};
    pl.ok = false;
    if (M < 1 || M > 8 || L.gv_grid <= 0 || !L.gv_tab || !L.gv_aligned) return pl;
    // 5..8 tokens (two IMMA token tiles) is parity-tested but opt-in: every CTA quantises all tokens itself (~1 us per
    // token on the Qwen3-8B shapes: qkv M=8 14.5 us against 8.3 us for the register-ring kernel).  It becomes the
    // route once the activation quantisation is hoisted into a pre-kernel.
    if (M > 4 && !gv_env_int("W4A16_IMMA8", 0)) return pl;
    const int gs = L.d.gs;
    if (gs != 32 && gs != 64 && gs != 128) return pl;               // per-channel scales: the f16 stream kernel
    pl.mpu = 128 / gs;
    pl.MR = M <= 1 ? 1 : M <= 2 ? 2 : M <= 4 ? 4 : 8;   // 8: two IMMA token tiles, one CTA per SM (activations take 2 K x 8 bytes)
    int64_t off = 128;                                               // full / empty mbarriers
    pl.off_xs = (int)off; off += L.d.K * 2 * pl.MR;                  // two byte planes per token slot
    pl.off_gtab = (int)off; off += (L.d.K / gs) * pl.MR * 8;
    off = (off + 15) & ~(int64_t)15;
    pl.off_part = (int)off; off += (int64_t)L.gv_max_slots * GV_WARPS * 8 * pl.MR * 8;
    off = (off + 127) & ~(int64_t)127;
    const int64_t depth = 2 * GV_WARPS * (GV_UNIT_BYTES + pl.mpu * META_BYTES);   // one more stage (32 units) in flight
    // two CTAs per SM (this linear + the next one's prefetch): 228 KB minus 1 KB reserved per CTA, halved
    int64_t D = (GV_SMEM_2CTA - off) / depth;
    if (D < 2) D = (GV_SMEM_FULL_SM - off) / depth;                  // large K x M: one CTA per SM
    if (D < 2) return pl;
    D = std::min<int64_t>(D, std::min<int64_t>(pl.MR > 4 ? 2 : 3, exp_env_int("W4A16_GEMV_STAGES", 3)));
    if (D < 2) D = 2;
    pl.S = (int)D;
    pl.stage_m_bytes = 0;
    pl.off_wring = (int)off; off += D * depth;
    pl.off_mring = pl.off_wring;
    pl.smem_bytes = (int)off;
    pl.ok = true;
    return pl;
}

bool gemv_imma_usable(const Linear &L, int64_t M) {
    if (exp_env_int("W4A16_NO_GEMV_IMMA", 0)) return false;
    return imma_plan(L, M).ok;
}

template <int MR, int MPU, int DI, bool REG, bool FUSED>
static int gi_launch_t(const GemvParams &p, const ImmaPlan &pl, int grid, cudaStream_t stream, bool pdl) {
    static bool attr_set = false;   // racing threads only repeat the call
    if (!attr_set) {
        W4_CUDA_OK(cudaFuncSetAttribute(gemv_imma_kernel<MR, MPU, DI, REG, FUSED>, cudaFuncAttributeMaxDynamicSharedMemorySize,
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
    W4_CUDA_OK(cudaLaunchKernelEx(&cfg, gemv_imma_kernel<MR, MPU, DI, REG, FUSED>, p));
    count_launch();
    return 0;
}

int launch_gemv_imma(const Linear &L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias,
                     cudaStream_t stream) {
    return launch_gemv_imma_peer(L, x, y, M, add_bias, stream, nullptr);
}

// peer != null: the epilogue all-reduces the row-parallel partial sums over the peers' symmetric buffers.
bool gemv_imma_peer_usable(const Linear &L, int64_t M, const PeerView &peer) {
    const ImmaPlan pl = imma_plan(L, M);
    return pl.ok && pl.MR <= 4 && peer.world > 1 && peer.world <= PEER_MAX_WORLD && (uint64_t)L.d.N * PEER_TOK <= peer.cap;
}

int launch_gemv_imma_peer(const Linear &L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias,
                          cudaStream_t stream, const PeerView *peer) {
    const ImmaPlan pl = imma_plan(L, M);
    if (!pl.ok) return fail("launch_gemv_imma: shape not supported by the IMMA stream kernel");
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
    p.stage_m_bytes = pl.stage_m_bytes;
    p.off_xs = pl.off_xs; p.off_gtab = pl.off_gtab; p.off_part = pl.off_part;
    p.off_wring = pl.off_wring; p.off_mring = pl.off_mring;
    p.dbg = gemv_debug_slot();
    if (peer) {
        if (!gemv_imma_peer_usable(L, M, *peer)) return fail("launch_gemv_imma: shape not supported by the fused all-reduce epilogue");
        p.peer = *peer;
    }
    // A launch with the fused exchange is launched programmatically early (its weight prefetch runs under the previous
    // kernel) from 4 ranks on, where the shards are a few microseconds long: Llama-3-70B step at TP = 2 / 4 / 8 without ->
    // with the attribute 0.972 -> 1.078 / 0.723 -> 0.713 / 0.632 -> 0.598 ms.  It never releases ITS dependents early.
    const bool pdl = gv_env_int("W4A16_NO_PDL", 0) == 0 && (!peer || exp_env_int("W4A16_FUSED_PDL_ATTR", peer->world >= 4 ? 1 : 0));
    const bool reg = (L.d.K % (32 * GV_UNIT_K)) == 0 && exp_env_int("W4A16_GV_REG", 1);
#define W4_GI3(MR_, MPU_, FUSED_)                                                      \
    do {                                                                               \
        if (reg) {                                                                     \
            if (pl.S == 3) return gi_launch_t<MR_, MPU_, 3, true, FUSED_>(p, pl, L.gv_grid, stream, pdl);  \
            return gi_launch_t<MR_, MPU_, 2, true, FUSED_>(p, pl, L.gv_grid, stream, pdl);     \
        }                                                                              \
        if (pl.S == 3) return gi_launch_t<MR_, MPU_, 3, false, FUSED_>(p, pl, L.gv_grid, stream, pdl); \
        return gi_launch_t<MR_, MPU_, 2, false, FUSED_>(p, pl, L.gv_grid, stream, pdl);        \
    } while (0)
#define W4_GI2(MR_, MPU_)                                                              \
    do {                                                                               \
        if (peer) W4_GI3(MR_, MPU_, true);                                             \
        W4_GI3(MR_, MPU_, false);                                                      \
    } while (0)
#define W4_GI(MR_)                                                                     \
    do {                                                                               \
        if (pl.mpu == 1) W4_GI2(MR_, 1);                                               \
        if (pl.mpu == 2) W4_GI2(MR_, 2);                                               \
        W4_GI2(MR_, 4);                                                                \
    } while (0)
    if (pl.MR == 8) {   // no fused instantiation above 4 tokens (gemv_imma_peer_usable), ring of 2
        if (peer) return fail("launch_gemv_imma: the fused all-reduce epilogue takes M <= 4");
#define W4_GI8(MPU_)                                                                   \
    do {                                                                               \
        if (reg) return gi_launch_t<8, MPU_, 2, true, false>(p, pl, L.gv_grid, stream, pdl);  \
        return gi_launch_t<8, MPU_, 2, false, false>(p, pl, L.gv_grid, stream, pdl);   \
    } while (0)
        if (pl.mpu == 1) W4_GI8(1);
        if (pl.mpu == 2) W4_GI8(2);
        W4_GI8(4);
#undef W4_GI8
    }
    switch (pl.MR) {
        case 1: W4_GI(1);
        case 2: W4_GI(2);
        default: W4_GI(4);
    }
#undef W4_GI2
#undef W4_GI3
#undef W4_GI
}

}  // namespace w4
