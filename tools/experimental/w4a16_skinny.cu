// w4a16_skinny.cu -- the skinny GEMM for 5 <= M <= 128 tokens: a weight-streaming tcgen05 kernel, SS mode.
//
// Replaces kt_mloop_m{8,12,16} (awq_gemv.cu:377-665), the dequant + cuBLAS path at small M
// (awq_marlin_dequant.cu:35-69, marlin_cuda.rs:793-818) and gptq_gemm_int4_bf16_tiled (gptq_dequant.cu:243-318)
// for the batched-decode range of BASELINE configs 3 and 4.
//
// Why not the prefill kernel: with the weights as the TMEM A operand every tcgen05.mma of 128 rows x 16 k costs
// ~170 cycles whatever the token count (measured: 680 cycles per 64-k block at N = 64 and at N = 256), i.e. 6 B of
// packed weights per cycle per SM = 1.7 TB/s on the whole chip.  Here the roles are swapped:
//     D[tok, n] = sum_k X[tok, k] * W[n, k]     UMMA M = 128 token rows (rows >= M zero-filled by TMA),
//                                               UMMA N = 256 output features, UMMA K = 16
//   A operand (tokens)  : shared memory, K-major SWIZZLE_128B tile [128 tok][64 k] loaded by TMA from the caller's x
//   B operand (weights) : shared memory, K-major SWIZZLE_128B tile [256 feat][64 k] of bf16 written by the dequant
//                         warps (a W4T chunk = 16 features x 64 k, four chunks per warp per stage; a lane's four k-consecutive
//                         values of one feature are one 8-byte store, the 128B swizzle makes them conflict-free)
//   accumulator         : TMEM, fp32, two buffers of 256 columns (epilogue of tile i under the main loop of i+1)
// An MMA of 128 x 256 x 16 has a floor of 128 cycles, so a 64-k stage (8 KB of packed weights) costs 512 cycles
// of tensor pipe and ~80 KB of shared-memory traffic (A 16 + B 32 read, B 32 written): 13-16 B of packed weights
// per cycle per SM, 2.2-2.7x the prefill kernel's rate at these token counts.
//
// Work: tile = (256-feature block, K slice).  K is split so that tiles ~ a multiple of the SM count (slices may
// differ by one k block); fp32 partials meet in the per-device split-K scratch and the last slice to arrive sums
// them in slice order (self-resetting tickets, deterministic).  Arithmetic is the prefill kernel's: every weight
// is the reference's D2 value bf16_rn(f32(q - z) * f32(bf16 scale)), fp32 accumulation, one rounding to bf16.
#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "w4a16_dequant.cuh"
#include "w4a16_internal.cuh"
#include "w4a16_tc.cuh"

namespace w4 {

int encode_x_map(CUtensorMap *map, const __nv_bfloat16 *x, int64_t K, int64_t M, int box_rows);

constexpr int SK_THREADS = 640;      // 20 warps: TMA, MMA, TMEM-alloc, spare, 12 dequant (3 groups of 4), 4 epilogue
constexpr int SK_DQ_WARPS = 12;
constexpr int SK_TM = 128;           // token rows (UMMA M)
constexpr int SK_TN = 256;           // output features per tile (UMMA N)
constexpr int SK_BK = 64;            // k per stage
constexpr int SK_STAGES = 3;
constexpr int SK_A_BYTES = SK_TM * 128;
constexpr int SK_B_BYTES = SK_TN * 128;
constexpr int SK_STAGE_BYTES = SK_A_BYTES + SK_B_BYTES;
constexpr int SK_WRING_BYTES = SK_DQ_WARPS * 2 * 4 * CHUNK_BYTES;   // packed weights: 12 warps x 2 iterations x 4 chunks
constexpr int SK_SMEM_BYTES = SK_STAGES * SK_STAGE_BYTES + SK_WRING_BYTES + 1024 /*barriers*/ + 1024 /*alignment*/;

struct SkinnyParams {
    const uint4 *qw;
    const uint8_t *meta;
    const __nv_bfloat16 *bias;
    __nv_bfloat16 *y;
    int M, K, N, KB, G;
    int gs_shift;             // log2(group size), or 31 for per-channel
    int n_tiles, ksplit;      // tile = slice * n_tiles + n_tile
    float *part;              // [ksplit][M][N] fp32 partial sums (ksplit > 1)
    int *tickets;             // [n_tiles]
    int test;                 // W4A16_SK_TEST: debugging switches
    unsigned long long *dbg;  // W4A16_SK_DEBUG: per-CTA cycle counters (timing experiments only)
};

// D[tmem] (+)= A[smem descriptor] * B[smem descriptor]
__device__ __forceinline__ void tc_mma_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void st_shared_v2(uint32_t addr, uint32_t a, uint32_t b) {
    asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(addr), "r"(a), "r"(b) : "memory");
}

// k blocks [kb0, kb1) of slice `sl` out of `ks` (slices differ by at most one block)
__device__ __forceinline__ void sk_slice(int sl, int ks, int kblks, int &kb0, int &kb1) {
    kb0 = (int)(((long long)sl * kblks) / ks);
    kb1 = (int)(((long long)(sl + 1) * kblks) / ks);
}

template <int GPS>
__global__ void __launch_bounds__(SK_THREADS, 1)
skinny_kernel(const __grid_constant__ CUtensorMap x_map, const SkinnyParams p) {
    constexpr int S = SK_STAGES;
    extern __shared__ uint8_t sk_smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(sk_smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + S * SK_STAGE_BYTES + SK_WRING_BYTES);
    const uint32_t smem_a = smem_u32(smem);                       // A stages: [S][128][128 B]
    const uint32_t smem_b = smem_a + S * SK_A_BYTES;              // B stages: [S][256][128 B]
    const uint32_t smem_w = smem_b + S * SK_B_BYTES;              // packed-weight ring: [16 warps][2][4 chunks][512 B]
    const uint32_t full = smem_u32(bars), empty = full + 8 * S;
    const uint32_t tmem_full = empty + 8 * S, tmem_empty = tmem_full + 16;
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(bars + 2 * S + 4);
    int *flag = reinterpret_cast<int *>(bars + 2 * S + 5);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KBLKS = p.K / SK_BK;
    const int ntiles = p.n_tiles * p.ksplit;

    // ---- dequant-warp state (declared here so that the first weight copies go out before the CTA set-up) ----
    const int dq = warp - 4, gq = (dq >> 2) % SK_STAGES, wq = dq & 3;
    const int g = lane >> 2, t = lane & 3;
    const int nr = p.N >> 4;
    const size_t rb_stride = (size_t)p.KB * 32;      // uint4 per row block
    struct Cursor { int tile, kb, kb1; uint32_t vmask; const uint4 *w; const uint8_t *m; };
    auto cur_set = [&](Cursor &c) {     // c.tile < ntiles
        sk_slice(c.tile / p.n_tiles, p.ksplit, KBLKS, c.kb, c.kb1);
        const int R0 = (c.tile % p.n_tiles) * (SK_TN / 16) + 4 * wq;
        c.vmask = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) c.vmask |= (R0 + i < nr ? 1u : 0u) << i;   // row blocks beyond N dequantise to 0
        c.w = p.qw + (size_t)R0 * rb_stride + lane;
        c.m = p.meta + (size_t)R0 * p.G * META_BYTES + 4 * g;
    };
    auto cur_norm = [&](Cursor &c) {    // carry a k index that ran past its slice into the next tile(s)
        while (c.tile < ntiles && c.kb >= c.kb1) {
            const int rem = c.kb - c.kb1;
            c.tile += gridDim.x;
            if (c.tile < ntiles) { cur_set(c); c.kb += rem; }
        }
    };
    auto cur_init = [&](Cursor &c) {
        c.tile = blockIdx.x; c.kb = 0; c.kb1 = 0; c.vmask = 0; c.w = nullptr; c.m = nullptr;
        if (c.tile < ntiles) { cur_set(c); c.kb += gq; cur_norm(c); }
    };
    // four chunks (row blocks) of one k block -> ring slot; always commits a group so that wait_group counts iterations
    auto fetch_w = [&](Cursor &c, uint32_t slot) {
        if (c.tile < ntiles) {
            const uint4 *q = c.w + (size_t)c.kb * 32;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                if ((c.vmask >> i) & 1u)
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(slot + i * CHUNK_BYTES), "l"(q + i * rb_stride) : "memory");
                else
                    asm volatile("st.shared.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(slot + i * CHUNK_BYTES), "r"(0u) : "memory");
            }
            c.kb += SK_STAGES;
            if (c.kb >= c.kb1) cur_norm(c);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    auto fetch_m = [&](Cursor &c, uint32_t (&s2)[4][GPS], uint32_t (&zb)[4][GPS]) {
        if (c.tile < ntiles) {
#pragma unroll
            for (int h = 0; h < GPS; ++h) {
                const int grp = (c.kb * SK_BK + h * 32) >> p.gs_shift;
                const uint8_t *m = c.m + (size_t)grp * META_BYTES;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const bool v = (c.vmask >> i) & 1u;
                    const uint8_t *mi = m + (size_t)i * p.G * META_BYTES;
                    s2[i][h] = v ? __ldg(reinterpret_cast<const uint32_t *>(mi)) : 0u;
                    zb[i][h] = v ? (uint32_t)__ldg(mi + (META_ZERO_OFF - 3 * g)) : 0u;
                }
            }
            c.kb += SK_STAGES;
            if (c.kb >= c.kb1) cur_norm(c);
        }
    };
    const uint32_t wring_l = smem_w + (uint32_t)((unsigned)dq % SK_DQ_WARPS) * (2 * 4 * CHUNK_BYTES) + lane * 16;   // this lane's 16 bytes of slot 0
    Cursor cw;
    cw.tile = ntiles;
    if (warp >= 4 && warp < 4 + SK_DQ_WARPS) {       // weights are immutable: the first two iterations' chunks are requested right away
        cur_init(cw);
        fetch_w(cw, wring_l);
        fetch_w(cw, wring_l + 4 * CHUNK_BYTES);
    }

    if (warp == 1 && lane == 0) {
        for (int i = 0; i < S; ++i) {
            mbar_init(full + 8 * i, 5);     // the producer's arrive.expect_tx (TMA completes the bytes) + 4 dequant warps
            mbar_init(empty + 8 * i, 1);    // tcgen05.commit
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(tmem_full + 8 * i, 1);    // tcgen05.commit after the last k block of a tile
            mbar_init(tmem_empty + 8 * i, 4);   // one arrival per epilogue warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_u32(const_cast<uint32_t *>(tmem_slot))), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (warp == 0 && lane == 0)
        asm volatile("prefetch.tensormap [%0];" ::"l"(&x_map) : "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    if (warp == 0) {
        // ===== TMA producer: x tiles [128 tokens][64 k] =====
        if (lane == 0) {
            asm volatile("griddepcontrol.wait;" ::: "memory");   // x may be the previous kernel's output
            int s = 0; uint32_t ph = 0;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                int kb0, kb1;
                sk_slice(tile / p.n_tiles, p.ksplit, KBLKS, kb0, kb1);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(empty + 8 * s, ph ^ 1);
                    mbar_arrive_expect_tx(full + 8 * s, SK_A_BYTES);
                    tma_load_2d(smem_a + s * SK_A_BYTES, &x_map, kb * SK_BK, 0, full + 8 * s);
                    if (++s == S) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        constexpr uint32_t idesc = make_idesc(SK_TM, SK_TN);
        int s = 0; uint32_t ph = 0;
        uint32_t ti = 0;
        const bool prof = p.dbg != nullptr;
        long long c_start = prof ? clock64() : 0, w_f = 0, w_i = 0, c0 = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++ti) {
            int kb0, kb1;
            sk_slice(tile / p.n_tiles, p.ksplit, KBLKS, kb0, kb1);
            const uint32_t buf = ti & 1, use = ti >> 1;
            mbar_wait(tmem_empty + 8 * buf, (use & 1) ^ 1);   // the epilogue has drained this accumulator buffer
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + buf * SK_TN;
            for (int kb = kb0; kb < kb1; ++kb) {
                if (prof) c0 = clock64();
                mbar_wait(full + 8 * s, ph);       // x tile landed and all 16 row blocks of B are in shared memory
                if (prof) { const long long c1 = clock64(); w_f += c1 - c0; c0 = c1; }
                tc_fence_after();
                if (elect_one()) {
                    const uint64_t adesc = make_b_desc(smem_a + s * SK_A_BYTES);
                    const uint64_t bdesc = make_b_desc(smem_b + s * SK_B_BYTES);
#pragma unroll
                    for (int k = 0; k < SK_BK / 16; ++k)   // +32 bytes of k inside the swizzle atom = +2 in the address field
                        tc_mma_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (uint32_t)((kb != kb0) | (k != 0)));
                    tc_commit(empty + 8 * s);                  // frees both operand tiles of this stage
                    if (kb == kb1 - 1) tc_commit(tmem_full + 8 * buf);
                }
                __syncwarp();
                if (prof) w_i += clock64() - c0;
                if (++s == S) { s = 0; ph ^= 1; }
            }
        }
        if (prof && lane == 0) {
            unsigned long long *d = p.dbg + blockIdx.x * 8;
            d[0] = clock64() - c_start; d[1] = w_f; d[2] = w_i;
        }
    } else if (warp >= 4 && warp < 4 + SK_DQ_WARPS) {
        // ===== dequant warps: packed W4T chunks -> bf16 -> B stage (shared memory, K-major, 128B swizzle) =====
        // Three groups of four warps.  Group gq fills the k blocks with (CTA-global block index) % 3 == gq (ring stage gq); inside a
        // group warp wq owns the four row blocks [4 wq, 4 wq + 4) of the 16 in the tile.  A warp therefore touches
        // every 3rd k block only: its per-block latencies (barrier wait, proxy fence, arrive) are amortised over four
        // chunks and overlap with the other groups' blocks.
        // Packed weights travel through a warp-private shared-memory ring filled with cp.async two iterations
        // (= 6 k blocks, 48 KB per SM) ahead: every lane copies exactly the 16 bytes it later reads, so no barrier
        // is needed, and no registers are spent on the lookahead.  The first copies are issued before the TMEM
        // allocation / barrier set-up of the CTA completes (see the call of `prologue_fetch` above).
        const int my_tiles = blockIdx.x < ntiles ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
        int total = 0;
        for (int i = 0, tile = blockIdx.x; i < my_tiles; ++i, tile += gridDim.x) {
            int kb0, kb1;
            sk_slice(tile / p.n_tiles, p.ksplit, KBLKS, kb0, kb1);
            total += kb1 - kb0;
        }
        Cursor cm;
        cur_init(cm);
        uint32_t s2[4][GPS], zb[4][GPS];        // scales / zeros: one iteration ahead, in registers
        fetch_m(cm, s2, zb);

        // this lane's 8-byte slots inside a B stage: row 64 wq + 16 i + g (and + 8), 16-byte chunk (2 j + (t >> 1)) ^ g
        const uint32_t lane_b = (uint32_t)((64 * wq + g) * 128 + 8 * (t & 1));
        uint32_t xo[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) xo[j] = (uint32_t)((2 * j + (t >> 1)) ^ g) << 4;
        const bool prof = p.dbg != nullptr && dq == 0;
        long long w_e = 0, w_d = 0, w_x = 0, c0 = 0, c_start = prof ? clock64() : 0;
        int n = 0;                               // iteration of this warp: CTA-global k block it = gq + 3 n
        // group gq owns ring stage gq: its consecutive iterations are consecutive uses of that stage, so the
        // parity wait below can never be more than one phase ahead of the barrier
        static_assert(SK_DQ_WARPS == 4 * SK_STAGES, "one dequant group per ring stage");
        for (int it = gq; it < total; it += S, ++n) {
            const int s = gq;
            const uint32_t ph = (uint32_t)n & 1u;
            if (p.test & 1) asm volatile("cp.async.wait_group 0;" ::: "memory");
            else asm volatile("cp.async.wait_group 1;" ::: "memory");   // this iteration's chunks have landed (own bytes only)
            uint4 w[4];
            const uint32_t slot = wring_l + (uint32_t)(n & 1) * (4 * CHUNK_BYTES);
#pragma unroll
            for (int i = 0; i < 4; ++i)
                asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                             : "=r"(w[i].x), "=r"(w[i].y), "=r"(w[i].z), "=r"(w[i].w) : "r"(slot + i * CHUNK_BYTES));
            if (prof) c0 = clock64();
            mbar_wait(empty + 8 * s, ph ^ 1);   // the MMAs that read this stage have retired
            if (prof) { const long long c1 = clock64(); w_e += c1 - c0; c0 = c1; }
            const uint32_t stage_b = smem_b + s * SK_B_BYTES + lane_b;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const uint32_t wj[4] = {w[i].x, w[i].y, w[i].z, w[i].w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int h = GPS == 2 ? (j >> 1) : 0;
                    const uint32_t zlo = zero_magic(zb[i][h] & 0xFu), zhi = zero_magic(zb[i][h] >> 4);
                    const uint32_t slo2 = prmt(s2[i][h], 0, 0x1010), shi2 = prmt(s2[i][h], 0, 0x3232);
                    const uint32_t a = stage_b + i * (16 * 128) + xo[j];
                    // row g: k 16j + 4t .. + 3, then row g + 8
                    st_shared_v2(a, dequant_pair(wj[j], zlo, slo2), dequant_pair(wj[j] >> 8, zlo, slo2));
                    st_shared_v2(a + 8 * 128, dequant_pair(wj[j] >> 4, zhi, shi2), dequant_pair(wj[j] >> 12, zhi, shi2));
                }
            }
            fetch_w(cw, slot);                    // refill the slot just consumed: iteration n + 2
            fetch_m(cm, s2, zb);
            if (prof) { const long long c1 = clock64(); w_d += c1 - c0; c0 = c1; }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic stores -> tensor-core reads
            __syncwarp();
            if (lane == 0) mbar_arrive(full + 8 * s);
            if (prof) w_x += clock64() - c0;
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        if (prof && lane == 0) {
            unsigned long long *d = p.dbg + blockIdx.x * 8;
            d[3] = clock64() - c_start; d[4] = w_e; d[5] = w_d; d[6] = w_x;
        }
    } else if (warp >= 4 + SK_DQ_WARPS) {
        // ===== epilogue warps: TMEM lane = token, columns = features =====
        const int q = warp & 3;
        const int et = threadIdx.x - (4 + SK_DQ_WARPS) * 32;   // 0..127
        const int tok = q * 32 + lane;
        const bool warp_live = q * 32 < p.M;    // warp-uniform: this quarter of the TMEM lanes holds real tokens
        uint32_t ti = 0;
        asm volatile("griddepcontrol.wait;" ::: "memory");       // y / the split-K scratch may still be in use upstream
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++ti) {
            const int n_tile = tile % p.n_tiles, sl = tile / p.n_tiles;
            const uint32_t buf = ti & 1, use = ti >> 1;
            const int n0 = n_tile * SK_TN;
            mbar_wait(tmem_full + 8 * buf, use & 1);
            tc_fence_after();
            const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + buf * SK_TN;
            if (warp_live) {
#pragma unroll 1
                for (int c = 0; c < SK_TN / 32; ++c) {
                    const int n = n0 + c * 32;
                    if (n >= p.N) break;               // N % 64 == 0: a 32-column block is inside or outside as a whole
                    uint32_t v[32];
                    tc_ld_32x32b_x32(t_addr + c * 32, v);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    if (tok < p.M) {
                        if (p.ksplit > 1) {
                            float4 *dst = reinterpret_cast<float4 *>(p.part + ((size_t)sl * p.M + tok) * p.N + n);
#pragma unroll
                            for (int i = 0; i < 8; ++i)
                                __stcg(dst + i, make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]),
                                                            __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3])));
                        } else {
                            uint32_t o[16];
#pragma unroll
                            for (int i = 0; i < 16; ++i) {
                                float b0 = 0.f, b1 = 0.f;
                                if (p.bias) { b0 = __bfloat162float(p.bias[n + 2 * i]); b1 = __bfloat162float(p.bias[n + 2 * i + 1]); }
                                __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(v[2 * i]) + b0, __uint_as_float(v[2 * i + 1]) + b1);
                                o[i] = *reinterpret_cast<uint32_t *>(&h);
                            }
                            uint4 *dst = reinterpret_cast<uint4 *>(p.y + (size_t)tok * p.N + n);
#pragma unroll
                            for (int i = 0; i < 4; ++i) dst[i] = make_uint4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]);
                        }
                    }
                }
            }
            tc_fence_before();
            if (lane == 0) mbar_arrive(tmem_empty + 8 * buf);   // the MMAs of tile ti + 2 may overwrite this buffer
            if (p.ksplit == 1) continue;

            // split-K: the last slice of this feature block to arrive sums all slices in slice order
            asm volatile("fence.acq_rel.gpu;" ::: "memory");
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (et == 0) {
                int prev;
                asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], 1;" : "=r"(prev) : "l"(p.tickets + n_tile) : "memory");
                const int last = prev == p.ksplit - 1;
                if (last) p.tickets[n_tile] = 0;          // self-reset for the next launch
                *flag = last;
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            const int last = *flag;
            asm volatile("bar.sync 1, 128;" ::: "memory");   // the flag may be rewritten for the next tile
            if (!last) continue;
            asm volatile("fence.acq_rel.gpu;" ::: "memory");
            // 128 threads: 64 float4 columns x 2 rows per pass, 4 passes in flight
            const int c4 = et & 63, r0 = et >> 6;
            const int nn = n0 + c4 * 4;
            if (nn >= p.N) continue;
            float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
            if (p.bias) {
                b4.x = __bfloat162float(p.bias[nn]); b4.y = __bfloat162float(p.bias[nn + 1]);
                b4.z = __bfloat162float(p.bias[nn + 2]); b4.w = __bfloat162float(p.bias[nn + 3]);
            }
            const size_t slice = (size_t)p.M * p.N;
#pragma unroll 1
            for (int row0 = r0; row0 < p.M; row0 += 8) {
                float4 acc[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int s2 = 0; s2 < p.ksplit; ++s2) {
                    float4 v[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int row = row0 + 2 * i;
                        v[i] = row < p.M ? __ldcg(reinterpret_cast<const float4 *>(p.part + (size_t)s2 * slice + (size_t)row * p.N + nn))
                                         : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
#pragma unroll
                    for (int i = 0; i < 4; ++i) { acc[i].x += v[i].x; acc[i].y += v[i].y; acc[i].z += v[i].z; acc[i].w += v[i].w; }
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int row = row0 + 2 * i;
                    if (row < p.M) {
                        __nv_bfloat162 lo = __floats2bfloat162_rn(acc[i].x + b4.x, acc[i].y + b4.y);
                        __nv_bfloat162 hi = __floats2bfloat162_rn(acc[i].z + b4.z, acc[i].w + b4.w);
                        uint2 out;
                        out.x = *reinterpret_cast<uint32_t *>(&lo); out.y = *reinterpret_cast<uint32_t *>(&hi);
                        *reinterpret_cast<uint2 *>(p.y + (size_t)row * p.N + nn) = out;
                    }
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// ---- host side -----------------------------------------------------------------------
static int sk_env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return v ? atoi(v) : dflt;
}

bool skinny_usable(const Linear &L, int64_t M) {
    if (L.epilogue != W4A16_EPILOGUE_NONE) return false;   // no SwiGLU epilogue in the opt-in kernel
    if (sk_env_int("W4A16_NO_SKINNY", 0)) return false;
    if (M < 1 || M > SK_TM) return false;
    const int gs = L.d.gs;
    if (gs != 32 && gs != 64 && gs != 128 && gs != L.d.K) return false;
    return L.d.K % SK_BK == 0 && L.d.N % 64 == 0 && L.d.K / SK_BK >= 4;
}

// K slices: tiles ~ a whole number of waves of the persistent grid, slices of at least 4 k blocks.
// Cost in k-block units of one CTA: waves x (blocks per slice + a fixed per-tile cost: pipeline fill, epilogue,
// fp32 partial write, ticket).
static int skinny_pick_ksplit(int n_tiles, int kblks, int sms) {
    const int forced = sk_env_int("W4A16_SK_KSPLIT", 0);
    if (forced > 0 && kblks / forced >= 1) return forced;
    const int c_fix = sk_env_int("W4A16_SK_FIX", 6);
    int best_ks = 1;
    long best = -1;
    for (int ks = 1; ks <= 32 && kblks / ks >= 4; ++ks) {
        if (n_tiles > SPLITK_MAX_TILES && ks > 1) break;
        const long waves = ((long)n_tiles * ks + sms - 1) / sms;
        const long per = (kblks + ks - 1) / ks + c_fix + (ks > 1 ? 2 : 0);
        const long cost = waves * per;
        if (best < 0 || cost < best) { best = cost; best_ks = ks; }
    }
    return best_ks;
}

template <int GPS>
static int launch_sk(const CUtensorMap &map, const SkinnyParams &p, int grid, cudaStream_t stream) {
    static bool attr_set = false;   // racing threads only repeat the call
    if (!attr_set) {
        W4_CUDA_OK(cudaFuncSetAttribute(skinny_kernel<GPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, SK_SMEM_BYTES));
        attr_set = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(SK_THREADS);
    cfg.dynamicSmemBytes = SK_SMEM_BYTES;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = sk_env_int("W4A16_NO_PDL", 0) ? 0 : 1;
    W4_CUDA_OK(cudaLaunchKernelEx(&cfg, skinny_kernel<GPS>, map, p));
    count_launch();
    return 0;
}

int launch_skinny(const Linear &L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias,
                  cudaStream_t stream) {
    if (!skinny_usable(L, M)) return fail("launch_skinny: shape not supported by the skinny kernel");
    const int gs = L.d.gs;
    SkinnyParams p;
    p.qw = reinterpret_cast<const uint4 *>(L.qw);
    p.meta = L.meta;
    p.bias = add_bias && L.has_bias ? L.bias : nullptr;
    p.y = y;
    p.M = (int)M; p.K = (int)L.d.K; p.N = (int)L.d.N; p.KB = (int)L.d.KB; p.G = (int)L.d.G;
    p.gs_shift = gs == L.d.K && gs != 32 && gs != 64 && gs != 128 ? 31 : (gs == 32 ? 5 : gs == 64 ? 6 : 7);
    p.n_tiles = (int)((L.d.N + SK_TN - 1) / SK_TN);
    const int kblks = (int)(L.d.K / SK_BK), sms = L.scratch->sm_count;
    p.ksplit = 1; p.part = nullptr; p.tickets = nullptr;
    const int ks = skinny_pick_ksplit(p.n_tiles, kblks, sms);
    if (ks > 1) {
        float *part = nullptr; int *tickets = nullptr;
        if (splitk_scratch(L.device, (size_t)ks * M * L.d.N * sizeof(float), stream, &part, &tickets) == 0 && part) {
            p.ksplit = ks; p.part = part; p.tickets = tickets;
        }
    }
    CUtensorMap map;
    if (encode_x_map(&map, x, L.d.K, M, SK_TM)) return 1;
    const int grid = std::min(p.n_tiles * p.ksplit, sms);
    static unsigned long long *dbg_buf = nullptr;
    p.dbg = nullptr;
    p.test = sk_env_int("W4A16_SK_TEST", 0);
    if (sk_env_int("W4A16_SK_DEBUG", 0)) {
        if (!dbg_buf) W4_CUDA_OK(cudaMalloc(&dbg_buf, 256 * 8 * sizeof(unsigned long long)));
        W4_CUDA_OK(cudaMemsetAsync(dbg_buf, 0, 256 * 8 * sizeof(unsigned long long), stream));
        p.dbg = dbg_buf;
    }
    const int rc = gs == 32 ? launch_sk<2>(map, p, grid, stream) : launch_sk<1>(map, p, grid, stream);
    if (rc == 0 && p.dbg) {
        static unsigned long long h[256 * 8];
        W4_CUDA_OK(cudaStreamSynchronize(stream));
        W4_CUDA_OK(cudaMemcpy(h, dbg_buf, sizeof(h), cudaMemcpyDeviceToHost));
        double v[7] = {0, 0, 0, 0, 0, 0, 0};
        const int n = std::min(grid, 256);
        for (int i = 0; i < n; ++i) for (int j = 0; j < 7; ++j) v[j] += (double)h[i * 8 + j] / n;
        fprintf(stderr, "[sk dbg] grid=%d ks=%d kblks/tile~%d | mma warp: cycles=%.0f wait_full=%.0f issue+commit=%.0f | dequant warp 0: cycles=%.0f wait_empty=%.0f dequant+st=%.0f fence+arrive=%.0f\n",
                grid, p.ksplit, kblks / p.ksplit, v[0], v[1], v[2], v[3], v[4], v[5], v[6]);
    }
    return rc;
}

}  // namespace w4
