// w4a16_midm_umma.cu -- the skinny GEMM for 33 <= M <= 128 tokens on the tcgen05 tensor cores.
//
// Replaces, for these batch sizes, awq_marlin_dequant_int4_bf16 + cuBLAS (awq_marlin_dequant.cu:35-69,
// marlin_cuda.rs:793-818) and marlin_tile_mma (marlin_tile_mma.cu:348-475), like the prefill kernel.
//
// Measured: legacy mma.sync tops out near 190 TFLOPS on this part, so above ~32 tokens the
// math must go through tcgen05 -- and below ~256 tokens the prefill kernel is bound by its dequant warps, not by its
// MMAs (TS mode: every MMA costs ~170 cycles whatever its N; few tiles x few tokens leave most SMs idle; x is
// re-streamed per tile).  This kernel keeps that kernel's roles but
//   - ALWAYS splits K so that (128-row tiles x k slices) covers every SM evenly, slices of uneven length where K / 128
//     does not divide; partial sums meet in fp32 scratch, the last slice of a tile to arrive sums them in slice order;
//   - keeps its k slice of x for ALL tokens resident in shared memory (TMA, 128B-swizzled [TM tokens][64 k] tiles,
//     loaded once per CTA): no x traffic per tile, and the stage ring holds weights only;
//   - dequantises into a 128B-swizzled SHARED-memory A tile ([128 n][64 k], 8-byte stores, conflict-free) and issues
//     SS-mode MMAs (both operands by descriptor): an M=128 x N=TM x K=16 MMA then costs ~TM/2 cycles, not 170;
//   - double-buffers the accumulator in TMEM (2 x TM columns) so the epilogue of a tile overlaps the next tile's MMAs.
#include <cuda.h>

#include <algorithm>
#include <cstdio>
#include <mutex>

#include "w4a16_dequant.cuh"
#include "w4a16_epilogue.cuh"
#include "w4a16_internal.cuh"
#include "w4a16_tc.cuh"

namespace w4 {

int encode_x_map(CUtensorMap *map, const __nv_bfloat16 *x, int64_t K, int64_t M, int box_rows);   // w4a16_prefill.cu
int mm_scratch(int device, size_t need, cudaStream_t stream, float4 **part, int **tickets);       // w4a16_midm.cu

constexpr int MU_THREADS = 704;       // 22 warps: 16 dequant, 4 epilogue, MMA (+ barrier init, TMA issue), TMEM alloc -> 88 registers per thread
constexpr int MU_A_STAGES = 6;       // at most; stages of 128 k: two [128 n][64 k] bf16 tiles, 32 KB.  As many as fit beside the x slice (>= 3)
constexpr int MU_A_BYTES = 2 * 128 * 128;
constexpr int MU_WPREF = 4;           // weight chunks each dequant warp keeps in flight (registers)
constexpr int MU_MAX_KBLKS = 32;      // k blocks of x resident per CTA
constexpr int MU_SMEM_X_MAX = 128 * 1024;
constexpr int MU_SMEM_TOTAL = 225 * 1024;   // dynamic shared memory limit used (227 KB minus the static barriers)

struct MuParams {
    const uint4 *qw;
    const uint8_t *meta;
    const __nv_bfloat16 *bias;
    __nv_bfloat16 *y;
    float *part;      // [ns][M][N] fp32 partial sums
    int *tickets;     // [n_tiles], self-resetting
    int M, K, N, KB, G;
    int gs_shift;     // log2(group size), 31 for per-channel
    int n_tiles;      // ceil(N / 128)
    int ns, U;        // k slices, K / 128
    int x_bytes;      // shared memory of the x slice (the A ring follows)
    int stages;       // A stages
    int swiglu;
    unsigned long long *dbg;   // W4A16_MU_DEBUG & 8: per-CTA cycle counters
    int debug;        // timing experiments (W4A16_MU_DEBUG, experiment builds only): 1 = no reduce, 2 = no partial stores
};
#ifdef W4A16_EXPERIMENT_ENV
#define MU_DBG(bits) ((p.debug & (bits)) != 0)
#else
#define MU_DBG(bits) false
#endif

template <int GPS> struct MuEntry {
    uint4 w;
    uint32_t s2[GPS];
    uint32_t zb[GPS];
};

// TM = tokens (UMMA N): 64 or 128, M <= TM.  GPS = groups per 64-k chunk: 1 (g >= 64 or per-channel), 2 (g = 32).
// grid = (cps, ns): CTA (cx, s) keeps k slice s of x resident and works on the 128-row tiles cx, cx + cps, ...
template <int TM, int GPS>
__global__ void __launch_bounds__(MU_THREADS, 1)
mu_kernel(const __grid_constant__ CUtensorMap x_map, const MuParams p) {
    constexpr int SMAX = MU_A_STAGES;
    const int S = p.stages;
    constexpr int XT = TM * 128;   // bytes of one 64-k block of x
    extern __shared__ uint8_t mu_smem_raw[];
    __shared__ uint64_t bars[MU_MAX_KBLKS + 3 * SMAX + 4];
    __shared__ uint32_t tmem_slot;
    __shared__ int last_flag;
    const uint32_t xs = (smem_u32(mu_smem_raw) + 1023u) & ~1023u;
    const uint32_t a_ring = xs + (uint32_t)p.x_bytes;
    const uint32_t xbar = smem_u32(bars), full = xbar + 8 * MU_MAX_KBLKS, empty = full + 8 * SMAX;
    const uint32_t tfull = empty + 8 * SMAX, tempty = tfull + 16, written = tempty + 16;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int s = blockIdx.y;
    const int u0 = (int)(((long long)s * p.U) / p.ns), u1 = (int)(((long long)(s + 1) * p.U) / p.ns);
    const int kb0 = 2 * u0, nkb = 2 * (u1 - u0);
    const int my_tiles = (int)blockIdx.x < p.n_tiles ? (p.n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

    if (warp == 20 && lane == 0) {
        for (int i = 0; i < nkb; ++i) mbar_init(xbar + 8 * i, 1);
        for (int i = 0; i < S; ++i) {
            mbar_init(written + 8 * i, 16); // the 16 dequant warps: 8 row blocks x 2 halves of a 128-k stage
            mbar_init(full + 8 * i, 1);    // the fence warp, once it has made their stores visible to the async proxy
            mbar_init(empty + 8 * i, 1);   // tcgen05.commit
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(tfull + 8 * b, 1);   // tcgen05.commit after the last k block of a tile
            mbar_init(tempty + 8 * b, 4);  // one arrival per epilogue warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 21) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_u32(&tmem_slot)), "r"((uint32_t)(2 * TM)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (warp == 20 && lane == 0) asm volatile("prefetch.tensormap [%0];" ::"l"(&x_map) : "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t *>(&tmem_slot);

    if (warp == 20) {
        // ===== x slice (TMA, once) + MMA issuer =====
        if (elect_one()) {
            for (int i = 0; i < nkb; ++i) {
                mbar_arrive_expect_tx(xbar + 8 * i, XT);
                tma_load_2d(xs + i * XT, &x_map, (kb0 + i) * 64, 0, xbar + 8 * i);
            }
        }
        __syncwarp();
        constexpr uint32_t idesc = make_idesc(128, TM);
        int st = 0; uint32_t ph = 0;
        const bool prof = MU_DBG(8);
        long long c_start = clock64(), w_f = 0, w_e = 0, w_x = 0, w_m = 0;
        for (int ti = 0; ti < my_tiles; ++ti) {
            const int buf = ti & 1;
            long long c0 = prof ? clock64() : 0;
            mbar_wait(tempty + 8 * buf, (uint32_t)(((ti >> 1) & 1) ^ 1));   // the epilogue has drained this accumulator
            if (prof) w_e += clock64() - c0;
            tc_fence_after();
            for (int kbl = 0; kbl < nkb; kbl += 2) {
                if (prof) c0 = clock64();
                if (ti == 0) { mbar_wait(xbar + 8 * kbl, 0); mbar_wait(xbar + 8 * kbl + 8, 0); }
                if (prof) { long long c1 = clock64(); w_x += c1 - c0; c0 = c1; }
                mbar_wait(full + 8 * st, ph);      // all 8 row blocks of this k block are in the A stage, fenced
                if (prof) { long long c1 = clock64(); w_f += c1 - c0; c0 = c1; }
                tc_fence_after();
                if (elect_one()) {
                    // one stage = 128 k = two 64-k tiles of A and of x: 8 MMAs and ONE commit (the issuing thread is the
                    // serial resource of this pipeline: ~40 cycles per MMA, ~75 per commit, plus the barrier polls)
#pragma unroll
                    for (int hk = 0; hk < 2; ++hk) {
                        const uint64_t adesc = make_b_desc(a_ring + st * MU_A_BYTES + hk * (MU_A_BYTES / 2));
                        const uint64_t bdesc = make_b_desc(xs + (kbl + hk) * XT);
#pragma unroll
                        for (int k = 0; k < 4; ++k)   // +32 bytes of k inside the swizzle atom = +2 in the address field
                            tc_mma_ss(tmem_base + buf * TM, adesc + 2 * k, bdesc + 2 * k, idesc, (uint32_t)((kbl | hk | k) != 0));
                    }
                    tc_commit(empty + 8 * st);
                    if (kbl == nkb - 2) tc_commit(tfull + 8 * buf);
                }
                __syncwarp();
                if (prof) w_m += clock64() - c0;
                if (++st == S) { st = 0; ph ^= 1; }
            }
        }
        if (prof && lane == 0) {
            unsigned long long *d = p.dbg + (blockIdx.y * gridDim.x + blockIdx.x) * 16;
            d[0] = clock64() - c_start; d[1] = w_f; d[2] = w_e; d[3] = w_x; d[4] = w_m; d[5] = my_tiles * (nkb >> 1);
        }
    } else if (warp < 16) {
        // ===== dequant warps: W4T chunk -> bf16 -> swizzled A tile in shared memory =====
        // Warp (rloc, set) fills row block rloc of the stage's 64-k tile `set`: all 16 warps work on the same 128-k stage.
        const int dq = warp, rloc = dq & 7, set = dq >> 3;
        const int g = lane >> 2, t = lane & 3;
        uint32_t a_off[4];   // row rloc*16 + g of the tile, this lane's 8 bytes of the 16-byte piece (2j + t/2) ^ g
#pragma unroll
        for (int j = 0; j < 4; ++j)
            a_off[j] = a_ring + (uint32_t)(set * (MU_A_BYTES / 2) + (rloc * 16 + g) * 128 + 8 * (t & 1) + (((2 * j + (t >> 1)) ^ g) << 4));
        const int total = my_tiles * (nkb >> 1);
        const int meta_step = p.gs_shift <= 7 ? (128 >> p.gs_shift) * META_BYTES : 0;
        const int zoff = META_ZERO_OFF - 3 * g;

        // Prefetch cursor: pointer increments only, and the loads are UNCONDITIONAL -- a predicated load with a zero
        // fallback made the compiler copy the loaded registers right behind the load, which stalled the warp for the full
        // memory latency once per chunk (ncu: 28 % of all samples on four MOVs).  N % 128 == 0 (checked by the host), so
        // every row block of a tile exists; past the last tile the cursor stops and re-reads its last chunk.
        int pf_tile = blockIdx.x, pf_left = 1 << 30, pf_wstep = 0, pf_mstep = 0;
        const uint4 *pf_w = p.qw + lane;
        const uint8_t *pf_m = p.meta + 4 * g;
        auto set_tile = [&](int tile) {
            const int R = tile * 8 + rloc;
            pf_left = nkb >> 1;
            pf_wstep = 64; pf_mstep = meta_step;
            pf_w = p.qw + ((size_t)R * p.KB + kb0 + set) * 32 + lane;
            pf_m = p.meta + ((size_t)R * p.G + (((kb0 + set) * 64) >> p.gs_shift)) * META_BYTES + 4 * g;
        };
        if (pf_tile < p.n_tiles) set_tile(pf_tile);
        auto fetch = [&](MuEntry<GPS> &e) {
            e.w = ldg_nc_v4(pf_w);
#pragma unroll
            for (int h = 0; h < GPS; ++h) {
                e.s2[h] = __ldg(reinterpret_cast<const uint32_t *>(pf_m + h * META_BYTES));
                e.zb[h] = __ldg(pf_m + h * META_BYTES + zoff);
            }
            if (--pf_left == 0) {
                pf_tile += gridDim.x;
                if (pf_tile < p.n_tiles) set_tile(pf_tile);
                else { pf_left = 1 << 30; pf_wstep = 0; pf_mstep = 0; }
            } else {
                pf_w += pf_wstep;
                pf_m += pf_mstep;
            }
        };
        MuEntry<GPS> ring[MU_WPREF];
#pragma unroll
        for (int u = 0; u < MU_WPREF; ++u) fetch(ring[u]);

        int st = 0; uint32_t ph = 0;   // stage / phase of this warp's iteration it = 0, 1, ... (one 128-k stage each)
        const bool dprof = MU_DBG(8);
        long long dq_t0 = clock64(), dq_emp = 0, dq_ld = 0;
        for (int it0 = 0; it0 < total; it0 += MU_WPREF) {
#pragma unroll
            for (int u = 0; u < MU_WPREF; ++u) {
                if (it0 + u < total) {
                    const MuEntry<GPS> &e = ring[u];
                    if (dprof) {
                        const long long c0 = clock64();
                        uint32_t tmp;
                        asm volatile("add.u32 %0, %1, %2;" : "=r"(tmp) : "r"(e.w.x), "r"(e.s2[0] + e.zb[0]));
                        asm volatile("" ::"r"(tmp));
                        dq_ld += clock64() - c0;
                    }
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
                    {
                        const long long c0 = dprof ? clock64() : 0;
                        mbar_wait(empty + 8 * st, ph ^ 1);   // the MMAs that read this stage have retired
                        if (dprof) dq_emp += clock64() - c0;
                    }
                    const uint32_t so = (uint32_t)st * MU_A_BYTES;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        st_shared_v2(a_off[j] + so, r[4 * j + 0], r[4 * j + 1]);
                        st_shared_v2(a_off[j] + so + 1024, r[4 * j + 2], r[4 * j + 3]);
                    }
                    // No proxy fence here: it compiles to MEMBAR.ALL.CTA, which waits for everything the warp has in flight
                    // -- this warp's weight prefetch (or, in the MMA warp, the MMAs it has just issued): measured, either
                    // placement capped the pipeline at one k block per memory / MMA latency.  A warp with nothing in
                    // flight (the fence warp below) acquires these arrivals, fences and hands the stage to the MMA warp.
                    __syncwarp();
                    if (lane == 0) mbar_arrive(written + 8 * st);
                    if (++st == S) { st = 0; ph ^= 1; }
                }
            }
        }
        if (dprof && lane == 0 && (warp == 0 || warp == 8)) {
            unsigned long long *d = p.dbg + (blockIdx.y * gridDim.x + blockIdx.x) * 16 + 8 + (warp >> 3) * 4;
            d[1] = clock64() - dq_t0; d[2] = dq_emp; d[3] = dq_ld;
        }
    }
    else if (warp == 21) {
        // ===== fence warp: generic-proxy stores of the dequant warps -> async proxy of the MMAs =====
        const int total = my_tiles * (nkb >> 1);
        int st = 0; uint32_t ph = 0;
        for (int it = 0; it < total; ++it) {
            mbar_wait(written + 8 * st, ph);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(full + 8 * st);
            if (++st == S) { st = 0; ph ^= 1; }
        }
    }
    else if (warp < 20) {
        // ===== epilogue warps: TMEM -> fp32 partial of this k slice; the last slice of a tile to arrive sums them =====
        const int q = warp & 3;
        const int et = threadIdx.x - 16 * 32;   // 0..127
        const int rows = p.M;                   // M <= TM
        for (int ti = 0; ti < my_tiles; ++ti) {
            const int tile = blockIdx.x + ti * gridDim.x, buf = ti & 1;
            const int n = tile * 128 + q * 32 + lane;
            mbar_wait(tfull + 8 * buf, (uint32_t)((ti >> 1) & 1));
            tc_fence_after();
            float *pbase = p.part + (size_t)s * p.M * p.N + n;
#pragma unroll 1
            for (int c = 0; c < TM / 32; ++c) {
                if (c * 32 >= rows) break;   // warp-uniform
                uint32_t v[32];
                tc_ld_32x32b_x32(tmem_base + ((uint32_t)(q * 32) << 16) + buf * TM + c * 32, v);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (n < p.N && !MU_DBG(2)) {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (c * 32 + i < rows) __stcg(pbase + (size_t)(c * 32 + i) * p.N, __uint_as_float(v[i]));
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty + 8 * buf);
            asm volatile("fence.acq_rel.gpu;" ::: "memory");
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (et == 0) {
                int prev;
                asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], 1;" : "=r"(prev) : "l"(p.tickets + tile) : "memory");
                const int last = prev == p.ns - 1;
                if (last) p.tickets[tile] = 0;          // self-reset for the next launch
                last_flag = last;
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            const int last = last_flag;
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (!last || MU_DBG(1)) continue;
            asm volatile("fence.acq_rel.gpu;" ::: "memory");
            // 128 threads: 32 float4 columns x 4 rows per pass, 8 rows per thread in flight, slices added in slice order
            const int c4 = et & 31, r0 = et >> 5;
            const int nn = tile * 128 + c4 * 4;
            const unsigned act = __ballot_sync(0xffffffffu, nn < p.N);   // a half-full last tile: lanes 0..15
            if (nn >= p.N) continue;
            float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
            if (p.bias) {
                b4.x = __bfloat162float(p.bias[nn]); b4.y = __bfloat162float(p.bias[nn + 1]);
                b4.z = __bfloat162float(p.bias[nn + 2]); b4.w = __bfloat162float(p.bias[nn + 3]);
            }
            const size_t slice = (size_t)p.M * p.N;
#pragma unroll 1
            for (int row0 = r0; row0 < rows; row0 += 32) {
                float4 acc[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int s2 = 0; s2 < p.ns; ++s2) {
                    float4 v[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int row = row0 + 4 * i;
                        v[i] = row < rows ? __ldcg(reinterpret_cast<const float4 *>(p.part + (size_t)s2 * slice + (size_t)row * p.N + nn))
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
                            *reinterpret_cast<uint2 *>(p.y + (size_t)row * (p.N >> 1) + (nn >> 4) * 8 + (nn & 7)) = out;
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
                        *reinterpret_cast<uint2 *>(p.y + (size_t)row * p.N + nn) = out;
                    }
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 21) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)(2 * TM)) : "memory");
    }
}

// ---- host side -----------------------------------------------------------------------
struct MuCfg { int tm, ns, cps; };

// Slices and CTAs per slice: the slice's x must fit shared memory; within that, the cheapest of
//     tiles per CTA x (k blocks per slice + a fixed per-tile cost) + a per-slice cost (partials written and re-read)
static MuCfg mu_pick(const Linear &L, int64_t M, int sms) {
    MuCfg c;
    c.tm = M <= 64 ? 64 : 128;
    const int U = (int)(L.d.K / 128);
    const int n_tiles = (int)((L.d.N + 127) / 128);
    const int max_units = std::min(MU_MAX_KBLKS / 2, MU_SMEM_X_MAX / (2 * c.tm * 128));
    static const int forced = exp_env_int("W4A16_MU_NS", 0);
    static const double c_tile = exp_env_double("W4A16_MU_CTILE", 6.0), c_slice = exp_env_double("W4A16_MU_CSLICE", 0.5);
    double best = -1;
    c.ns = 0; c.cps = 0;
    for (int ns = 1; ns <= U; ++ns) {
        if ((U + ns - 1) / ns > max_units) continue;
        const int cps = std::max(1, std::min(n_tiles, sms / ns));
        const int passes = (n_tiles + cps - 1) / cps;
        const double cost = passes * (2.0 * ((U + ns - 1) / ns) + c_tile) + c_slice * ns * (c.tm / 64);
        if ((forced > 0 && ns == forced) || (forced <= 0 && (best < 0 || cost < best))) { best = cost; c.ns = ns; c.cps = cps; }
        if (forced > 0 && ns == forced) break;
    }
    return c;
}

bool midm_umma_usable(const Linear &L, int64_t M) {
    static const bool off = exp_env_int("W4A16_NO_MIDM_UMMA", 0) != 0;
    static const int min_m = exp_env_int("W4A16_MU_MIN_M", 33), max_m = exp_env_int("W4A16_MU_MAX_M", 128);
    const int gs = L.d.gs;
    if (off || M < min_m || M > max_m || M > 128) return false;
    if (gs != 32 && gs != 64 && gs != 128 && gs != L.d.K) return false;
    return L.d.N % 128 == 0 && L.d.N / 128 <= 16384;
}

template <int TM, int GPS>
static int mu_launch(const CUtensorMap &map, const MuParams &p, dim3 grid, size_t smem, int device, cudaStream_t stream) {
    static cudaError_t attr_err[64];
    static bool attr_done[64] = {};
    static std::mutex mu;
    if (device < 0 || device >= 64) return fail("launch_midm_umma: device index out of range");
    {
        std::lock_guard<std::mutex> lk(mu);
        if (!attr_done[device]) {
            attr_err[device] = cudaFuncSetAttribute(mu_kernel<TM, GPS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                    MU_SMEM_TOTAL);
            attr_done[device] = true;
        }
    }
    if (attr_err[device] != cudaSuccess)
        return fail(std::string("cudaFuncSetAttribute(mu_kernel smem): ") + cudaGetErrorString(attr_err[device]));
    MuParams q = p;
    static unsigned long long *dbg_buf = nullptr;
    if (p.debug & 8) {
        if (!dbg_buf) W4_CUDA_OK(cudaMalloc(&dbg_buf, 256 * 16 * sizeof(unsigned long long)));
        W4_CUDA_OK(cudaMemsetAsync(dbg_buf, 0, 256 * 16 * sizeof(unsigned long long), stream));
        q.dbg = dbg_buf;
    }
    mu_kernel<TM, GPS><<<grid, MU_THREADS, smem, stream>>>(map, q);
    W4_CUDA_OK(cudaGetLastError());
    count_launch();
    if (p.debug & 8) {
        static unsigned long long h[256 * 16];
        W4_CUDA_OK(cudaStreamSynchronize(stream));
        W4_CUDA_OK(cudaMemcpy(h, dbg_buf, sizeof(h), cudaMemcpyDeviceToHost));
        const int n = (int)std::min(256u, grid.x * grid.y);
        double cyc = 0, wf = 0, we = 0, wx = 0, wm = 0, kb = 0, dt = 0, de = 0, dl = 0;
        for (int i = 0; i < n; ++i) { cyc += h[i*16]; wf += h[i*16+1]; we += h[i*16+2]; wx += h[i*16+3]; wm += h[i*16+4]; kb += h[i*16+5]; dt += h[i*16+9]; de += h[i*16+10]; dl += h[i*16+11]; }
        fprintf(stderr, "[mu dbg] grid=%dx%d stages/cta=%.1f mma warp: %.0f cycles (%.0f per 128-k stage): wait full %.1f%% tempty %.1f%% x %.1f%% issue+commit %.1f%% | dequant warp 0: %.0f cycles, empty wait %.1f%%, load wait %.1f%%\n",
                grid.x, grid.y, kb / n, cyc / n, cyc / kb, 100 * wf / cyc, 100 * we / cyc, 100 * wx / cyc, 100 * wm / cyc, dt / n, 100 * de / dt, 100 * dl / dt);
    }
    return 0;
}

// Returns 0 = launched, 1 = error, 2 = not available for this call (no scratch under capture): use another kernel.
int launch_midm_umma(const Linear &L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias, cudaStream_t stream) {
    if (!midm_umma_usable(L, M)) return 2;
    const MuCfg c = mu_pick(L, M, L.scratch->sm_count);
    if (c.ns < 1) return 2;
    const int gs = L.d.gs;
    MuParams p;
    p.qw = reinterpret_cast<const uint4 *>(L.qw);
    p.meta = L.meta;
    p.bias = add_bias && L.has_bias ? L.bias : nullptr;
    p.y = y;
    p.M = (int)M; p.K = (int)L.d.K; p.N = (int)L.d.N; p.KB = (int)L.d.KB; p.G = (int)L.d.G;
    p.gs_shift = gs == L.d.K && gs != 32 && gs != 64 && gs != 128 ? 31 : (gs == 32 ? 5 : gs == 64 ? 6 : 7);
    p.n_tiles = (int)((L.d.N + 127) / 128);
    p.ns = c.ns; p.U = (int)(L.d.K / 128);
    p.swiglu = L.epilogue == W4A16_EPILOGUE_SWIGLU;
    p.debug = exp_env_int("W4A16_MU_DEBUG", 0); p.dbg = nullptr;
    float4 *part4 = nullptr;
    if (mm_scratch(L.device, (size_t)c.ns * M * L.d.N * sizeof(float), stream, &part4, &p.tickets)) return 1;
    if (!part4) return 2;
    p.part = reinterpret_cast<float *>(part4);
    CUtensorMap map;
    if (encode_x_map(&map, x, L.d.K, M, c.tm)) return 1;
    const int max_units = (p.U + c.ns - 1) / c.ns;
    p.x_bytes = max_units * 2 * c.tm * 128;
    p.stages = std::min(MU_A_STAGES, (MU_SMEM_TOTAL - 1024 - p.x_bytes) / MU_A_BYTES);
    if (p.stages < 3) return 2;
    const size_t smem = (size_t)p.x_bytes + (size_t)p.stages * MU_A_BYTES + 1024;
    const dim3 grid((unsigned)c.cps, (unsigned)c.ns);
    const int gps = gs == 32 ? 2 : 1;
    if (c.tm == 64) return gps == 1 ? mu_launch<64, 1>(map, p, grid, smem, L.device, stream) : mu_launch<64, 2>(map, p, grid, smem, L.device, stream);
    return gps == 1 ? mu_launch<128, 1>(map, p, grid, smem, L.device, stream) : mu_launch<128, 2>(map, p, grid, smem, L.device, stream);
}

}  // namespace w4
