// w4a16_prefill.cu -- the batched / prefill GEMM (M > 32 tokens): tcgen05.mma with TMEM accumulators.
//
// Replaces awq_marlin_dequant_int4_bf16 + cuBLAS (awq_marlin_dequant.cu:35-69, marlin_cuda.rs:793-818),
// gptq_dequant_int4_bf16 + matmul (gptq_dequant.cu:48-87) and marlin_tile_mma (marlin_tile_mma.cu:348-475):
// the dense bf16 [K,N] weight matrix the reference materialises in HBM on every call never exists here.
//
// Arithmetic = the reference's "D2" grid: every weight is bf16_rn(f32(q - z) * f32(bf16 scale))
// (w4a16_dequant.cuh proves the packed bf16x2 sub+mul is that exact rounding), then a bf16 x bf16
// contraction with fp32 accumulation, + bias, one rounding to bf16.
//
// Mapping onto the SM (one persistent CTA per SM, 24 warps, warp-specialised):
//   D^T[n, tok] = sum_k W^T[n, k] * X[tok, k]      UMMA M = 128 output features (TMEM lanes),
//                                                   UMMA N = TM tokens (TMEM columns), UMMA K = 16
//   A operand (weights)  : TMEM.  16 dequant warps read the W4T chunks straight from global (one coalesced
//                          16-byte load per lane per 16x64 chunk, 4 stages ahead in registers), turn every
//                          word into its four bf16x2 fragment registers and write them with ONE
//                          tcgen05.st.16x256b.x4 per chunk -- the W4T nibble order (w4a16_layout.cuh) IS
//                          that instruction's register order, so there is no shuffle and no shared-memory
//                          round trip for weights.
//   B operand (tokens)   : shared memory, K-major, 128B-swizzled [TM tokens][64 k] tiles loaded by TMA
//                          (cp.async.bulk.tensor.2d) from the caller's x; rows beyond M are zero-filled.
//   accumulator          : TMEM columns [0, TM), fp32.
//   epilogue             : 4 warps, tcgen05.ld.32x32b -> +bias -> bf16 -> shared staging -> 16-byte
//                          coalesced stores of y rows; TMEM is released as soon as it has been read.
//   pipeline             : a ring of STAGES {A columns, B tile}; mbarriers full (TMA bytes + 8 dequant warps) / empty
//                          (tcgen05.commit), tmem_full / tmem_empty between MMA and epilogue.
#include <cuda.h>  // CUtensorMap + enums only; the encode entry point is fetched from the runtime, libcuda is not linked

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <vector>

#include "w4a16_dequant.cuh"
#include "w4a16_epilogue.cuh"
#include "w4a16_internal.cuh"
#include "w4a16_tc.cuh"

namespace w4 {

constexpr int PF_THREADS = 768;      // 24 warps: TMA, MMA, TMEM-alloc, spare, 16 dequant, 4 epilogue
constexpr int PF_TILE_N = 128;     // output features per CTA tile (UMMA M)
constexpr int PF_BLOCK_K = 64;     // k per pipeline stage: one W4T chunk column, one 128-byte swizzle row of x
constexpr int PF_A_COL0 = 256;     // TMEM column where the A ring starts (accumulator owns [0, 256))
constexpr int PF_A_STAGE_COLS = 32;  // 64 k of bf16 = 32 32-bit columns
constexpr int PF_WPREF = 4;        // weight chunks each dequant warp keeps in flight

struct PrefillParams {
    const uint4 *qw;
    const uint8_t *meta;
    const __nv_bfloat16 *bias;
    __nv_bfloat16 *y;
    int M, K, N, KB, G;
    int gs_shift;      // log2(group size), or 31 for per-channel (every k maps to group 0)
    int n_tiles, m_tiles;
    int ksplit, kblks_tile;   // split-K for few tokens x few output tiles: a tile = (n_tile, m_tile, k slice of kblks_tile blocks)
    float *part;              // [ksplit][M][N] fp32 partial sums (ksplit > 1)
    int *tickets;             // [n_tiles * m_tiles] arrival counters, self-resetting
    unsigned long long *dbg;   // W4A16_PF_DEBUG & 4: per-CTA {cycles, ns, wait_full_b, wait_full_a, wait_tmem_empty, tiles}
    int swiglu;        // SwiGLU handle: TMEM lanes l / l^8 are (gate, up); y is [M, N/2], 64 output columns per n-tile
    int debug;         // timing experiments only (W4A16_PF_DEBUG): 1 = skip x loads after the first k block, 2 = skip dequant
};

template <int TM> struct PfCfg {
    static constexpr int STAGES = TM == 256 ? 5 : 8;
    static constexpr int B_STAGE_BYTES = TM * 128;
    static constexpr int STAGE_BYTES_TOTAL = STAGES * B_STAGE_BYTES;
    static constexpr int OUT_BYTES = TM * PF_TILE_N * 2;  // bf16 staging tile [TM][128]
    static constexpr int BAR_BYTES = 1024;
    static constexpr int SMEM_USED = STAGE_BYTES_TOTAL + OUT_BYTES + BAR_BYTES;
    // at least ~116 KB so that two CTAs never share an SM: each CTA allocates all 512 TMEM columns
    static constexpr int SMEM_BYTES = (SMEM_USED < 118 * 1024 ? 118 * 1024 : SMEM_USED) + 1024;
};

// One prefetched 16x64 chunk of one row block + the scale pair / zero byte of its group(s).
template <int GPS> struct WEntry {
    uint4 w;
    uint32_t s2[GPS];
    uint32_t zb[GPS];
};

// TM = tokens per tile (UMMA N): 256, 128 or 64.  GPS = groups per 64-k stage: 1 (g >= 64), 2 (g = 32).
// CL2 = launched as clusters of two CTAs that work on the two n-tiles of a pair in lockstep (same m-tile, same k blocks):
// every x tile is fetched from L2 ONCE per pair -- each CTA loads half of the token rows and multicasts them into both
// shared memories.  Measured with the kernel's own switches (W4A16_PF_DEBUG), skipping the x loads after the first
// k block gained 6-12 % at M = 4096: the activation tiles (32 KB per 680 cycles per SM, 10 TB/s chip-wide) sit at the
// L2's bandwidth limit.
template <int TM, int GPS, bool CL2>
__global__ void __launch_bounds__(PF_THREADS, 1)
prefill_kernel(const __grid_constant__ CUtensorMap x_map, const PrefillParams p) {
    using C = PfCfg<TM>;
    constexpr int S = C::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __nv_bfloat16 *smem_out = reinterpret_cast<__nv_bfloat16 *>(smem + C::STAGE_BYTES_TOTAL);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + C::STAGE_BYTES_TOTAL + C::OUT_BYTES);
    const uint32_t smem_b = smem_u32(smem);
    const uint32_t full = smem_u32(bars), empty = full + 8 * S;
    const uint32_t tmem_full = empty + 8 * S, tmem_empty = tmem_full + 8;
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(bars + 2 * S + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KBLKS = p.kblks_tile;              // k blocks per tile (all of K unless split)
    const int nm_tiles = p.n_tiles * p.m_tiles;
    const int ntiles = nm_tiles * p.ksplit;      // tile = ks * nm_tiles + m_tile * n_tiles + n_tile

    if (warp == 1 && lane == 0) {
        for (int i = 0; i < S; ++i) {
            mbar_init(full + 8 * i, 9);     // the producer's arrive.expect_tx (TMA completes the bytes) + 8 dequant warps
            mbar_init(empty + 8 * i, CL2 ? 2 : 1);    // tcgen05.commit (CL2: of both CTAs of the pair)
        }
        mbar_init(tmem_full, 1);            // tcgen05.commit after the last k block of a tile
        mbar_init(tmem_empty, 4);           // one arrival per epilogue warp
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
    if constexpr (CL2) cluster_sync_all();   // the peer's barriers exist before anything is multicast into this CTA
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t crank = CL2 ? cluster_ctarank() : 0u;
    // Programmatic dependent launch: the next kernel of the
    // stream may take this SM as soon as it is free; of this kernel, only the warps that touch what a previous kernel may
    // still be using -- x (TMA producer), y / split-K scratch / tickets (epilogue) -- wait for it; the dequant warps start
    // on the immutable weights at once.  Without the attribute both instructions are no-ops.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    if (warp == 0) {
        // ===== TMA producer: x tiles [TM tokens][64 k] =====
        asm volatile("griddepcontrol.wait;" ::: "memory");
        int s = 0; uint32_t ph = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            const int m_tile = (tile % nm_tiles) / p.n_tiles, kb0 = (tile / nm_tiles) * KBLKS;
            for (int kb = 0; kb < KBLKS; ++kb) {
                mbar_wait(empty + 8 * s, ph ^ 1);
                if (elect_one()) {
                    if constexpr (CL2) {
                        // both halves of the tile land in THIS CTA's stage (own load + the peer's multicast): 32 KB expected
                        mbar_arrive_expect_tx(full + 8 * s, C::B_STAGE_BYTES);
                        tma_load_2d_mc(smem_b + s * C::B_STAGE_BYTES + crank * (C::B_STAGE_BYTES / 2), &x_map, (kb0 + kb) * PF_BLOCK_K,
                                       m_tile * TM + (int)crank * (TM / 2), full + 8 * s, (uint16_t)3);
                    } else if ((p.debug & 1) && kb >= S) {
                        mbar_arrive(full + 8 * s);
                    } else {
                        mbar_arrive_expect_tx(full + 8 * s, C::B_STAGE_BYTES);
                        tma_load_2d(smem_b + s * C::B_STAGE_BYTES, &x_map, (kb0 + kb) * PF_BLOCK_K, m_tile * TM, full + 8 * s);
                    }
                }
                __syncwarp();
                if (++s == S) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: the warp stays converged, one elected lane drives the tensor core for the CTA =====
        constexpr uint32_t idesc = make_idesc(PF_TILE_N, TM);
        int s = 0; uint32_t ph = 0, ti = 0;
        const bool prof = p.debug & 4;
        long long c_start = clock64(), w_f = 0, w_e = 0, w_m = 0, w_c = 0;
        unsigned long long t_start = 0;
        if (prof) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++ti) {
            long long c0 = prof ? clock64() : 0;
            mbar_wait(tmem_empty, (ti & 1) ^ 1);   // the epilogue has drained the previous tile
            if (prof) w_e += clock64() - c0;
            tc_fence_after();
            for (int kb = 0; kb < KBLKS; ++kb) {
                if (prof) c0 = clock64();
                mbar_wait(full + 8 * s, ph);       // x tile landed and all 8 row blocks of A are in TMEM
                if (prof) { long long c1 = clock64(); w_f += c1 - c0; c0 = c1; }
                tc_fence_after();
                if (elect_one()) {
                    const uint64_t bdesc = make_b_desc(smem_b + s * C::B_STAGE_BYTES);
                    const uint32_t a_tmem = tmem_base + PF_A_COL0 + s * PF_A_STAGE_COLS;
#pragma unroll
                    for (int k = 0; k < PF_BLOCK_K / 16; ++k)   // +32 bytes of k inside the swizzle atom = +2 in the address field
                        tc_mma_ts(tmem_base, a_tmem + k * 8, bdesc + 2 * k, idesc, (uint32_t)((kb | k) != 0));
                    if (prof) { long long c1 = clock64(); w_m += c1 - c0; c0 = c1; }
                    if constexpr (CL2) tc_commit_mc(empty + 8 * s, (uint16_t)3);   // ... in both CTAs: the peer multicasts into this stage too
                    else tc_commit(empty + 8 * s);              // frees the A columns and the B tile of this stage
                    if (kb == KBLKS - 1) tc_commit(tmem_full);  // accumulator complete
                    if (prof) w_c += clock64() - c0;
                }
                __syncwarp();
                if (++s == S) { s = 0; ph ^= 1; }
            }
        }
        if (prof && lane == 0) {
            unsigned long long t_end;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
            unsigned long long *d = p.dbg + blockIdx.x * 8;
            d[0] = clock64() - c_start; d[1] = t_end - t_start; d[2] = w_f; d[3] = 0; d[4] = w_e; d[5] = ti;
        }
        if (prof && elect_one()) {
            unsigned long long *d = p.dbg + blockIdx.x * 8;
            d[6] = w_m; d[7] = w_c;
        }
    } else if (warp >= 4 && warp < 20) {
        // ===== dequant warps: W4T chunk -> bf16 fragments -> TMEM (A operand) =====
        // Two sets of 8 warps; set 0 fills the even k blocks, set 1 the odd ones (K % 128 == 0, so the
        // number of k blocks per tile is even and a set keeps its parity across tiles).  Within a set,
        // warp (quarter, half) owns the 16 rows at TMEM lanes 32*quarter + 16*half.
        const int dq = warp - 4;
        const int quarter = warp & 3, half = (dq >> 2) & 1, set = dq >> 3;
        const int rloc = quarter * 2 + half;
        const int g = lane >> 2;
        const uint32_t a_taddr = tmem_base + ((uint32_t)(quarter * 32 + half * 16) << 16) + PF_A_COL0;
        const int my_tiles = blockIdx.x < ntiles ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
        const int total = my_tiles * KBLKS;
        const int nr = p.N >> 4;

        int pf_tile = blockIdx.x, pf_kb = set, pf_k0 = 0;
        const uint4 *pf_w = nullptr;
        const uint8_t *pf_m = nullptr;
        bool pf_valid = false;   // false for the row blocks of a 64-row tail tile that lie beyond N: they dequantise to 0
        auto set_tile = [&](int tile) {
            const int R = (tile % p.n_tiles) * 8 + rloc;
            pf_valid = R < nr;
            pf_k0 = (tile / nm_tiles) * KBLKS;
            pf_w = p.qw + (size_t)R * p.KB * 32 + lane;
            pf_m = p.meta + (size_t)R * p.G * META_BYTES + 4 * g;
        };
        if (pf_tile < ntiles) set_tile(pf_tile);
        auto fetch = [&](WEntry<GPS> &e) {
            if (pf_tile >= ntiles) return;
            if (pf_valid) {
                e.w = ldg_nc_v4(pf_w + (pf_k0 + pf_kb) * 32);
#pragma unroll
                for (int h = 0; h < GPS; ++h) {
                    const int grp = ((pf_k0 + pf_kb) * PF_BLOCK_K + h * 32) >> p.gs_shift;
                    const uint8_t *m = pf_m + grp * META_BYTES;
                    e.s2[h] = __ldg(reinterpret_cast<const uint32_t *>(m));
                    e.zb[h] = __ldg(m + (META_ZERO_OFF - 3 * g));
                }
            } else {
                e.w = make_uint4(0, 0, 0, 0);
#pragma unroll
                for (int h = 0; h < GPS; ++h) { e.s2[h] = 0; e.zb[h] = 0; }
            }
            pf_kb += 2;
            if (pf_kb >= KBLKS) {
                pf_kb -= KBLKS;
                pf_tile += gridDim.x;
                if (pf_tile < ntiles) set_tile(pf_tile);
            }
        };
        WEntry<GPS> ring[PF_WPREF];
#pragma unroll
        for (int u = 0; u < PF_WPREF; ++u) fetch(ring[u]);

        int s = set; uint32_t ph = 0;   // stage / phase of global k-block iteration `it` (it = set, set+2, ...)
        for (int it0 = set; it0 < total; it0 += 2 * PF_WPREF) {
#pragma unroll
            for (int u = 0; u < PF_WPREF; ++u) {
                if (it0 + 2 * u < total) {
                    const WEntry<GPS> &e = ring[u];
                    const uint32_t wj[4] = {e.w.x, e.w.y, e.w.z, e.w.w};
                    uint32_t r[16];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int h = GPS == 2 ? (j >> 1) : 0;
                        const uint32_t zlo = zero_magic(e.zb[h] & 0xFu), zhi = zero_magic(e.zb[h] >> 4);
                        const uint32_t slo2 = prmt(e.s2[h], 0, 0x1010), shi2 = prmt(e.s2[h], 0, 0x3232);
                        r[4 * j + 0] = dequant_pair(wj[j], zlo, slo2);        // row g   , k 4t, 4t+1
                        r[4 * j + 1] = dequant_pair(wj[j] >> 8, zlo, slo2);   // row g   , k 4t+2, 4t+3
                        r[4 * j + 2] = dequant_pair(wj[j] >> 4, zhi, shi2);   // row g+8 , k 4t, 4t+1
                        r[4 * j + 3] = dequant_pair(wj[j] >> 12, zhi, shi2);  // row g+8 , k 4t+2, 4t+3
                    }
                    fetch(ring[u]);
                    mbar_wait(empty + 8 * s, ph ^ 1);   // the MMAs that read this stage's columns have retired
                    tc_fence_after();
                    if (!(p.debug & 2)) {
                    tc_st_16x256b_x4(a_taddr + s * PF_A_STAGE_COLS, r);
                    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                    }
                    tc_fence_before();
                    if (lane == 0) mbar_arrive(full + 8 * s);
                    s += 2;
                    if (s >= S) { s -= S; ph ^= 1; }
                }
            }
        }
    } else if (warp >= 20) {
        // ===== epilogue warps: TMEM -> +bias -> bf16 -> staging -> y =====
        asm volatile("griddepcontrol.wait;" ::: "memory");
        const int q = warp & 3;
        const int et = threadIdx.x - 20 * 32;   // 0..127
        uint32_t ti = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++ti) {
            const int rem = tile % nm_tiles, ks = tile / nm_tiles;
            const int n_tile = rem % p.n_tiles, m_tile = rem / p.n_tiles;
            const int n = n_tile * PF_TILE_N + q * 32 + lane;
            const float bias = p.bias && n < p.N ? __bfloat162float(p.bias[n]) : 0.f;
            const int rows = min(TM, p.M - m_tile * TM);
            mbar_wait(tmem_full, ti & 1);
            tc_fence_after();
            if (p.ksplit > 1) {
                // split-K: fp32 partial of this k slice -> scratch; the last slice to arrive sums all of them in slice order
                float *pbase = p.part + ((size_t)ks * p.M + (size_t)m_tile * TM) * p.N + n;
#pragma unroll 1
                for (int c = 0; c < TM / 32; ++c) {
                    uint32_t v[32];
                    tc_ld_32x32b_x32(tmem_base + ((uint32_t)(q * 32) << 16) + c * 32, v);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    if (n < p.N) {
#pragma unroll
                        for (int i = 0; i < 32; ++i)
                            if (c * 32 + i < rows) __stcg(pbase + (size_t)(c * 32 + i) * p.N, __uint_as_float(v[i]));
                    }
                }
                tc_fence_before();
                if (lane == 0) mbar_arrive(tmem_empty);
                asm volatile("fence.acq_rel.gpu;" ::: "memory");
                asm volatile("bar.sync 1, 128;" ::: "memory");
                int *flag = reinterpret_cast<int *>(smem_out);
                if (et == 0) {
                    int prev;
                    asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], 1;" : "=r"(prev) : "l"(p.tickets + rem) : "memory");
                    const int last = prev == p.ksplit - 1;
                    if (last) p.tickets[rem] = 0;          // self-reset for the next launch
                    *flag = last;
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");
                const int last = *flag;
                asm volatile("bar.sync 1, 128;" ::: "memory");   // the flag word is staging memory again
                if (last) {
                    asm volatile("fence.acq_rel.gpu;" ::: "memory");
                    // 128 threads: 32 float4 columns x 4 rows per pass, every slice's loads issued before the sums
                    const int c4 = et & 31, r0 = et >> 5;
                    const int nn = n_tile * PF_TILE_N + c4 * 4;
                    const unsigned act = __ballot_sync(0xffffffffu, nn < p.N);   // a half-full last tile: lanes 0..15
                    if (nn < p.N) {
                        float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (p.bias) {
                            b4.x = __bfloat162float(p.bias[nn]); b4.y = __bfloat162float(p.bias[nn + 1]);
                            b4.z = __bfloat162float(p.bias[nn + 2]); b4.w = __bfloat162float(p.bias[nn + 3]);
                        }
                        const size_t slice = (size_t)p.M * p.N;
                        // 8 rows per pass: all 8 loads of a slice are in flight before they are added (slice order)
#pragma unroll 1
                        for (int row0 = r0; row0 < rows; row0 += 32) {
                            float4 acc[8];
#pragma unroll
                            for (int i = 0; i < 8; ++i) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                            for (int s2 = 0; s2 < p.ksplit; ++s2) {
                                float4 v[8];
#pragma unroll
                                for (int i = 0; i < 8; ++i) {
                                    const int row = row0 + 4 * i;
                                    v[i] = row < rows ? __ldcg(reinterpret_cast<const float4 *>(
                                                            p.part + (size_t)s2 * slice + (size_t)(m_tile * TM + row) * p.N + nn))
                                                      : make_float4(0.f, 0.f, 0.f, 0.f);
                                }
#pragma unroll
                                for (int i = 0; i < 8; ++i) { acc[i].x += v[i].x; acc[i].y += v[i].y; acc[i].z += v[i].z; acc[i].w += v[i].w; }
                            }
                            if (p.swiglu) {   // columns nn.. (gate) and nn+8.. (up) sit two lanes apart
#pragma unroll
                                for (int i = 0; i < 8; ++i) {
                                    const int row = row0 + 4 * i;      // warp-uniform
                                    if (row >= rows) continue;
                                    const float4 a = make_float4(acc[i].x + b4.x, acc[i].y + b4.y, acc[i].z + b4.z, acc[i].w + b4.w);
                                    float4 o;
                                    o.x = __shfl_xor_sync(act, a.x, 2); o.y = __shfl_xor_sync(act, a.y, 2);
                                    o.z = __shfl_xor_sync(act, a.z, 2); o.w = __shfl_xor_sync(act, a.w, 2);
                                    if (!(c4 & 2)) {
                                        __nv_bfloat162 lo = __halves2bfloat162(swiglu_bf16(a.x, o.x), swiglu_bf16(a.y, o.y));
                                        __nv_bfloat162 hi = __halves2bfloat162(swiglu_bf16(a.z, o.z), swiglu_bf16(a.w, o.w));
                                        uint2 out;
                                        out.x = *reinterpret_cast<uint32_t *>(&lo); out.y = *reinterpret_cast<uint32_t *>(&hi);
                                        *reinterpret_cast<uint2 *>(p.y + (size_t)(m_tile * TM + row) * (p.N >> 1) + (nn >> 4) * 8 + (nn & 7)) = out;
                                    }
                                }
                                continue;
                            }
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                const int row = row0 + 4 * i;
                                if (row < rows) {
                                    __nv_bfloat162 lo = __floats2bfloat162_rn(acc[i].x + b4.x, acc[i].y + b4.y);
                                    __nv_bfloat162 hi = __floats2bfloat162_rn(acc[i].z + b4.z, acc[i].w + b4.w);
                                    uint2 out;
                                    out.x = *reinterpret_cast<uint32_t *>(&lo); out.y = *reinterpret_cast<uint32_t *>(&hi);
                                    *reinterpret_cast<uint2 *>(p.y + (size_t)(m_tile * TM + row) * p.N + nn) = out;
                                }
                            }
                        }
                    }
                }
                continue;
            }
            if (p.swiglu) {
                // TMEM lane l holds gate column 8 (l / 16) + l % 8 when l & 8 == 0, lane l ^ 8 the matching up column.
                // The pair splits the token rows: the gate lane finishes even rows, the up lane odd rows.
                const bool up_lane = (lane & 8) != 0;
                const int oc = (q * 2 + (lane >> 4)) * 8 + (lane & 7);
#pragma unroll 1
                for (int c = 0; c < TM / 32; ++c) {
                    uint32_t v[32];
                    tc_ld_32x32b_x32(tmem_base + ((uint32_t)(q * 32) << 16) + c * 32, v);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {
                        const float a = __uint_as_float(v[i]) + bias, b = __uint_as_float(v[i + 1]) + bias;
                        const float recv = __shfl_xor_sync(0xffffffffu, up_lane ? a : b, 8);
                        smem_out[(c * 32 + i + (up_lane ? 1 : 0)) * PF_TILE_N + oc] = up_lane ? swiglu_bf16(recv, b) : swiglu_bf16(a, recv);
                    }
                }
                tc_fence_before();
                if (lane == 0) mbar_arrive(tmem_empty);
                asm volatile("bar.sync 1, 128;" ::: "memory");
                const int I = p.N >> 1;
                __nv_bfloat16 *yb = p.y + (size_t)m_tile * TM * I + (size_t)n_tile * (PF_TILE_N / 2);
                for (int idx = et; idx < rows * 8; idx += 128) {
                    const int row = idx >> 3, c8 = idx & 7;
                    if (n_tile * (PF_TILE_N / 2) + c8 * 8 < I)
                        *reinterpret_cast<uint4 *>(yb + (size_t)row * I + c8 * 8) =
                        *reinterpret_cast<const uint4 *>(smem_out + row * PF_TILE_N + c8 * 8);
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");
                continue;
            }
#pragma unroll 1
            for (int c = 0; c < TM / 32; ++c) {
                uint32_t v[32];
                tc_ld_32x32b_x32(tmem_base + ((uint32_t)(q * 32) << 16) + c * 32, v);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int i = 0; i < 32; ++i)
                    smem_out[(c * 32 + i) * PF_TILE_N + q * 32 + lane] = __float2bfloat16_rn(__uint_as_float(v[i]) + bias);
            }
            tc_fence_before();
            if (lane == 0) mbar_arrive(tmem_empty);   // the next tile's MMAs may overwrite the accumulator
            asm volatile("bar.sync 1, 128;" ::: "memory");
            __nv_bfloat16 *ybase = p.y + (size_t)m_tile * TM * p.N + (size_t)n_tile * PF_TILE_N;
            for (int idx = et; idx < rows * 16; idx += 128) {
                const int row = idx >> 4, c16 = idx & 15;
                if (n_tile * PF_TILE_N + c16 * 8 < p.N)
                    *reinterpret_cast<uint4 *>(ybase + (size_t)row * p.N + c16 * 8) =
                    *reinterpret_cast<const uint4 *>(smem_out + row * PF_TILE_N + c16 * 8);
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");  // staging is free again
        }
    }

    tc_fence_before();
    __syncthreads();
    if constexpr (CL2) cluster_sync_all();   // the peer may still signal this CTA's barriers
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// ---- host side -----------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static std::once_flag g_encode_once;
static std::string g_encode_err;

int get_encode() {
    std::call_once(g_encode_once, [] {
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn)
            g_encode_err = std::string("cuTensorMapEncodeTiled entry point not available: ") + cudaGetErrorString(e);
        else
            g_encode = reinterpret_cast<EncodeTiledFn>(fn);
    });
    if (!g_encode) return fail(g_encode_err);
    return 0;
}

bool prefill_available() { return true; }

// x [M][K] bf16 as a 2-D tensor map: boxes of 64 k (one 128-byte swizzle row) x box_rows tokens, rows beyond M zero-filled.
int encode_x_map(CUtensorMap *map, const __nv_bfloat16 *x, int64_t K, int64_t M, int box_rows) {
    if (get_encode()) return 1;
    const cuuint64_t gdim[2] = {(cuuint64_t)K, (cuuint64_t)M};
    const cuuint64_t gstride[1] = {(cuuint64_t)K * 2};
    const cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
    const cuuint32_t estride[2] = {1, 1};
    CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<__nv_bfloat16 *>(x), gdim, gstride, box,
                          estride, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled(x) failed with CUresult " + std::to_string((int)r));
    return 0;
}

template <int TM, int GPS, bool CL2>
static int launch_pf(const CUtensorMap &map, const PrefillParams &p, int grid, int device, cudaStream_t stream) {
    using C = PfCfg<TM>;
    static cudaError_t attr_err[64];
    static int max_clusters[64];
    static bool attr_done[64] = {};
    static std::mutex mu;
    if (device < 0 || device >= 64) return fail("launch_prefill: device index out of range");
    {
        std::lock_guard<std::mutex> lk(mu);
        if (!attr_done[device]) {   // per device
            attr_err[device] = cudaFuncSetAttribute(prefill_kernel<TM, GPS, CL2>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES);
            max_clusters[device] = 0;
            if (CL2 && attr_err[device] == cudaSuccess) {
                cudaLaunchConfig_t q{};
                q.gridDim = dim3(2); q.blockDim = dim3(PF_THREADS); q.dynamicSmemBytes = C::SMEM_BYTES;
                cudaLaunchAttribute qa[1];
                qa[0].id = cudaLaunchAttributeClusterDimension;
                qa[0].val.clusterDim.x = 2; qa[0].val.clusterDim.y = 1; qa[0].val.clusterDim.z = 1;
                q.attrs = qa; q.numAttrs = 1;
                int n = 0;
                if (cudaOccupancyMaxActiveClusters(&n, prefill_kernel<TM, GPS, CL2>, &q) != cudaSuccess) { cudaGetLastError(); n = 0; }
                max_clusters[device] = n;
            }
            attr_done[device] = true;
        }
    }
    if (attr_err[device] != cudaSuccess)
        return fail(std::string("cudaFuncSetAttribute(prefill_kernel smem): ") + cudaGetErrorString(attr_err[device]));
    if (!CL2) {
        static const bool pdl = getenv("W4A16_NO_PDL") == nullptr || atoi(getenv("W4A16_NO_PDL")) == 0;
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3((unsigned)grid);
        cfg.blockDim = dim3(PF_THREADS);
        cfg.dynamicSmemBytes = C::SMEM_BYTES;
        cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = pdl ? 1 : 0;
        W4_CUDA_OK(cudaLaunchKernelEx(&cfg, prefill_kernel<TM, GPS, CL2>, map, p));
        count_launch();
        return 0;
    }
    if (max_clusters[device] < 1) return 2;   // clusters of two cannot be scheduled here: the caller falls back
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)std::min(grid & ~1, 2 * max_clusters[device]));
    cfg.blockDim = dim3(PF_THREADS);
    cfg.dynamicSmemBytes = C::SMEM_BYTES;
    cfg.stream = stream;
    static const bool pdl2 = getenv("W4A16_NO_PDL") == nullptr || atoi(getenv("W4A16_NO_PDL")) == 0;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl2 ? 2 : 1;
    W4_CUDA_OK(cudaLaunchKernelEx(&cfg, prefill_kernel<TM, GPS, CL2>, map, p));
    count_launch();
    return 0;
}

int prefill_token_tile(int64_t M) { return M > 128 ? 256 : M > 64 ? 128 : 64; }

// Split-K scratch (fp32 partials + self-resetting tickets): one per (device, STREAM).  Launches on one stream are ordered,
// so they may share it; launches on different streams get different scratch and cannot race (re-entrant for
// read-only weights).  Grown on demand outside stream capture -- inside a capture that would need a larger one the launch
// simply runs unsplit -- and an outgrown block is never freed: a CUDA graph captured earlier may hold its address.
constexpr int PF_MAX_SPLIT_TILES = SPLITK_MAX_TILES;
struct SplitScratch { int device; cudaStream_t stream; float *part = nullptr; size_t bytes = 0; int *tickets = nullptr; };
static std::vector<SplitScratch> g_split;
static std::mutex g_split_mu;
int splitk_scratch(int device, size_t need, cudaStream_t stream, float **part, int **tickets) {
    std::lock_guard<std::mutex> lock(g_split_mu);
    SplitScratch *s = nullptr;
    for (SplitScratch &c : g_split)
        if (c.device == device && c.stream == stream) { s = &c; break; }
    const bool have = s && s->bytes >= need && s->tickets;
    if (!have) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(stream, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) { *part = nullptr; return 0; }
        if (!s) {
            g_split.push_back(SplitScratch{device, stream});
            s = &g_split.back();
        }
        if (!s->tickets) {
            if (cudaMalloc(&s->tickets, PF_MAX_SPLIT_TILES * sizeof(int)) != cudaSuccess) { s->tickets = nullptr; cudaGetLastError(); return 1; }
            cudaMemset(s->tickets, 0, PF_MAX_SPLIT_TILES * sizeof(int));
        }
        if (s->bytes < need) {
            const size_t want = std::max<size_t>(need, (size_t)32 << 20);
            float *fresh = nullptr;
            if (cudaMalloc(&fresh, want) != cudaSuccess) { cudaGetLastError(); return 1; }
            s->part = fresh;          // the outgrown block stays allocated (see above)
            s->bytes = want;
        }
    }
    *part = s->part; *tickets = s->tickets;
    return 0;
}

int launch_prefill(const Linear &L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias,
                   cudaStream_t stream) {
    if (M < 1) return fail("launch_prefill: M must be >= 1");
    if (M > (int64_t)1 << 30) return fail("launch_prefill: M too large");
    if (get_encode()) return 1;
    const int gs = L.d.gs;
    if (gs != 32 && gs != 64 && gs != 128 && gs != L.d.K) return fail("launch_prefill: unsupported group size");
    const int TM = prefill_token_tile(M);
    PrefillParams p;
    p.qw = reinterpret_cast<const uint4 *>(L.qw);
    p.meta = L.meta;
    p.bias = add_bias && L.has_bias ? L.bias : nullptr;
    p.y = y;
    p.M = (int)M; p.K = (int)L.d.K; p.N = (int)L.d.N; p.KB = (int)L.d.KB; p.G = (int)L.d.G;
    p.swiglu = L.epilogue == W4A16_EPILOGUE_SWIGLU;
    p.gs_shift = gs == L.d.K && gs != 32 && gs != 64 && gs != 128 ? 31 : (gs == 32 ? 5 : gs == 64 ? 6 : 7);
    p.n_tiles = (int)((L.d.N + PF_TILE_N - 1) / PF_TILE_N);  // N % 64 == 0: the last tile may be half full
    p.m_tiles = (int)((M + TM - 1) / TM);
    // Few tokens x few output tiles would leave most SMs idle (and the weight stream latency-bound): split K so
    // that (tiles x slices) fills the machine; partial sums meet in a per-device fp32 scratch, summed in slice order.
    p.ksplit = 1; p.part = nullptr; p.tickets = nullptr;
    {
        const int kblks = (int)(L.d.K / PF_BLOCK_K), nm = p.n_tiles * p.m_tiles, sms = L.scratch->sm_count;
        // Cost model in units of one 64-k block of one CTA (~0.35 us: the TS-mode MMA chain): waves x (blocks per
        // slice + a fixed per-slice cost: pipeline fill, fp32 partial write, ticket) + the last slice's re-read of all
        // partials of its tile, which is latency-bound on 128 threads: ~0.09 units per (slice x token row), fitted to
        // qkv / o at M = 64 (split wins by 8-11 us), M = 128 (by ~4 us) and M = 256 (split LOSES 9-14 us).
        int ks = 1;
        {
            static const int forced = exp_env_int("W4A16_PF_KSPLIT", 0);
            static const int c_fix = exp_env_int("W4A16_PF_KSPLIT_FIX", 6);
            static const double c_row = exp_env_double("W4A16_PF_KSPLIT_ROW", 0.09);
            const int rows = (int)std::min<int64_t>(M, TM);
            const int cands[6] = {1, 2, 3, 4, 6, 8};
            double best = -1;
            for (int c : cands) {
                if (kblks % (2 * c) != 0 || kblks / c < 8) continue;      // a slice keeps an even number of k blocks
                const long waves = ((long)nm * c + sms - 1) / sms;
                // every split tile writes its fp32 partial (~ rows); the tile that arrives last re-reads c of them
                const double cost = c == 1 ? (double)waves * kblks
                                           : (double)waves * (kblks / c + c_fix + c_row * rows) + c_row * c * rows;
                if (best < 0 || cost < best) { best = cost; ks = c; }
            }
            if (forced > 0 && kblks % (2 * forced) == 0) ks = forced;
        }
        static const bool off = exp_env_int("W4A16_PF_NO_SPLITK", 0) != 0;
        if (ks > 1 && !off && nm <= PF_MAX_SPLIT_TILES) {
            float *part = nullptr; int *tickets = nullptr;
            if (splitk_scratch(L.device, (size_t)ks * M * L.d.N * sizeof(float), stream, &part, &tickets) == 0 && part) {
                p.ksplit = ks; p.part = part; p.tickets = tickets;
            }
        }
        p.kblks_tile = kblks / p.ksplit;
    }
    p.debug = exp_env_int("W4A16_PF_DEBUG", 0);
    static unsigned long long *dbg_buf = nullptr;
    p.dbg = nullptr;
    if (p.debug & 4) {
        if (!dbg_buf) W4_CUDA_OK(cudaMalloc(&dbg_buf, 256 * 8 * sizeof(unsigned long long)));
        W4_CUDA_OK(cudaMemsetAsync(dbg_buf, 0, 256 * 8 * sizeof(unsigned long long), stream));
        p.dbg = dbg_buf;
    }

    const int ntiles = p.n_tiles * p.m_tiles * p.ksplit;
    const int grid = std::min(ntiles, L.scratch->sm_count);
    const int gps = gs == 32 ? 2 : 1;
    // pairs of CTAs share every x tile (TMA multicast) when the two n-tiles of a pair exist for every m-tile
    static const bool no_cl2 = exp_env_int("W4A16_PF_NO_CLUSTER", 0) != 0;
    bool cl2 = TM == 256 && p.ksplit == 1 && (p.n_tiles % 2) == 0 && grid >= 2 && !no_cl2 && !(p.debug & 3);

    int rc = 2;
    for (int attempt = 0; attempt < 2 && rc == 2; ++attempt) {
        CUtensorMap map;
        const cuuint64_t gdim[2] = {(cuuint64_t)L.d.K, (cuuint64_t)M};
        const cuuint64_t gstride[1] = {(cuuint64_t)L.d.K * 2};
        const cuuint32_t box[2] = {(cuuint32_t)PF_BLOCK_K, (cuuint32_t)(cl2 ? TM / 2 : TM)};
        const cuuint32_t estride[2] = {1, 1};
        CUresult r = g_encode(&map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<__nv_bfloat16 *>(x), gdim, gstride, box,
                              estride, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled(x) failed with CUresult " + std::to_string((int)r));
        if (cl2) {
            rc = gps == 1 ? launch_pf<256, 1, true>(map, p, grid, L.device, stream) : launch_pf<256, 2, true>(map, p, grid, L.device, stream);
            if (rc == 2) cl2 = false;   // no room for clusters of two on this device: plain launch
            continue;
        }
#define W4_PF(TM_)                                                                          \
    do {                                                                                    \
        if (gps == 1) return launch_pf<TM_, 1, false>(map, p, grid, L.device, stream);      \
        return launch_pf<TM_, 2, false>(map, p, grid, L.device, stream);                    \
    } while (0)
        if (TM == 256) rc = [&] { W4_PF(256); }();
        else if (TM == 128) rc = [&] { W4_PF(128); }();
        else rc = [&] { W4_PF(64); }();
#undef W4_PF
    }
    if (rc == 0 && (p.debug & 4)) {
        static unsigned long long h[256 * 8];
        W4_CUDA_OK(cudaStreamSynchronize(stream));
        W4_CUDA_OK(cudaMemcpy(h, dbg_buf, sizeof(h), cudaMemcpyDeviceToHost));
        double cyc = 0, ns = 0, wb = 0, wa = 0, we = 0, wm = 0, wc = 0; int n = 0;
        for (int i = 0; i < grid && i < 256; ++i) { cyc += h[i*8]; ns += h[i*8+1]; wb += h[i*8+2]; wa += h[i*8+3]; we += h[i*8+4]; wm += h[i*8+6]; wc += h[i*8+7]; ++n; }
        fprintf(stderr, "[pf dbg] grid=%d tiles/cta=%llu cycles=%.0f ns=%.0f clk=%.3f GHz wait_full_b=%.1f%% wait_full_a=%.1f%% wait_tmem_empty=%.1f%% mma_issue=%.1f%% commit=%.1f%%\n",
                grid, h[5], cyc / n, ns / n, cyc / ns, 100 * wb / cyc, 100 * wa / cyc, 100 * we / cyc, 100 * wm / cyc, 100 * wc / cyc);
    }
    return rc;
}

}  // namespace w4
