// w4a16_midm.cu -- the skinny GEMM for 11 <= M <= 48 tokens: between the GEMV kernels and the tcgen05 prefill kernel.
//
// Replaces, for these batch sizes, awq_gemv_int4_kt_mloop_m16 looped over 16-token pieces and the dequant + cuBLAS
// path above it (marlin_cuda.rs:843-953: M <= 16 -> GEMV, else awq_marlin_dequant_int4_bf16 + matmul), and
// marlin_tile_mma (marlin_tile_mma.cu:348-475) where that is selected.
//
// Why another kernel.  Measured with the prefill kernel's own counters: below ~128 tokens that
// kernel's time does not depend on the MMAs at all -- it is the same with every tcgen05.mma skipped -- but on its 16
// dequant warps, which spend ~200 issue slots per 16x64 chunk (unpack, subtract, scale, TMEM store, ring bookkeeping)
// and sit on ONE row-block-of-128 tile per CTA, so few tokens x few tiles leave most schedulers idle.  Here every warp
// owns one 16-row block x one k slice and the whole chip's 2368 warps work on dequantisation:
//   weights      : one coalesced 16-byte load per lane per chunk (the W4T word order IS the mma.m16n8k16 A fragment
//                  after LOP3 + sub + mul -- the D2 rounding grid of the prefill kernel, w4a16_dequant.cuh), PF chunks
//                  ahead in registers;
//   activations  : the CTA's k slice of x for ALL tokens sits in shared memory, put there by TMA as 128B-swizzled
//                  [64 k][8 NT tokens] tiles (no thread instructions, rows beyond M zero-filled); a lane's B fragment
//                  is ONE 8-byte LDS (k-slots 4t..4t+3 are contiguous in W4T's k order), conflict-free under the swizzle
//                  once the MMA columns are dealt to the token rows as 0,2,4,6,1,3,5,7 (mm_rho);
//   math         : mma.sync.m16n8k16 bf16, fp32 accumulators in registers, NT token tiles per A fragment;
//   split-K      : slices meet in an fp32 scratch in fragment order (one coalesced 512-byte store per token tile); the
//                  warp that arrives last at a row block's ticket sums the slices in slice order (deterministic) and
//                  writes y (+ bias / SwiGLU).  No CTA-wide barrier after set-up: warps never wait for each other.
// Issue slots per chunk at 32 tokens: ~175 (12 shifts + 48 dequant + 16 LDS + 16 MMA + ring / loop bookkeeping) for 32768
// MACs; the prefill kernel's dequant alone is more than that per chunk, sits on 16 warps of at most 96 SMs at these sizes
// and is redone per tile.  Above ~48 tokens the x slice no longer leaves room for few slices and the fp32 partials
// (slices x M x N x 4 bytes, written and re-read) outweigh the weights: the tcgen05 kernel takes over.
#include <cuda.h>

#include <algorithm>
#include <mutex>
#include <vector>

#include "w4a16_dequant.cuh"
#include "w4a16_epilogue.cuh"
#include "w4a16_internal.cuh"
#include "w4a16_tc.cuh"

namespace w4 {

int encode_x_map(CUtensorMap *map, const __nv_bfloat16 *x, int64_t K, int64_t M, int box_rows);   // w4a16_prefill.cu

constexpr int MM_WARPS = 16;
constexpr int MM_THREADS = MM_WARPS * 32;
constexpr int MM_RING = 8;            // weight chunks each warp keeps in flight: 16 warps x 8 x 512 B = 64 KB per SM
constexpr int MM_STAGE = 512 + 2 * META_BYTES;   // one chunk + the meta block(s) of its group(s)
constexpr int MM_RING_BYTES = MM_WARPS * MM_RING * MM_STAGE;
constexpr int MM_MAX_KBLKS = 144;     // mbarriers: one per 64-k block of a slice
constexpr int MM_SMEM_X_MAX = 144 * 1024;

struct MidmParams {
    const uint4 *qw;
    const uint8_t *meta;
    const __nv_bfloat16 *bias;
    __nv_bfloat16 *y;
    float4 *part;        // [ns][NR][NT][32] fp32 fragments (ns > 1)
    int *tickets;        // [NR], self-resetting
    int M, K, N, KB, G, NR;
    int gs_shift;        // log2(group size), 31 for per-channel
    int ns;              // k slices
    int cps;             // CTAs per slice (gridDim.x)
    int U;               // K / 128
    int swiglu;
    int x_bytes;         // shared memory of the x slice (the weight rings follow it)
    int debug;           // timing experiments (W4A16_MM_DEBUG, experiment builds only): 1 = no MMAs, 2 = no dequant arithmetic, 4 = no B loads
};
#ifdef W4A16_EXPERIMENT_ENV
#define MM_DBG(bits) ((p.debug & (bits)) != 0)
#else
#define MM_DBG(bits) false
#endif

__device__ __forceinline__ void mm_mma(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint2 mm_lds64(uint32_t addr) {
    uint2 r;
    asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "r"(addr));
    return r;
}

// MMA column c of a token tile is token row mm_rho(c) of that tile, not c: with the TMA 128B swizzle the 16-byte piece a
// lane reads is (2j + t/2) ^ (row & 7), so rows r and r ^ 1 meet in the same pair of pieces -- columns 0..3 (one half-warp of
// an 8-byte LDS) must sit on rows with distinct r >> 1.  Measured before: 44 % of the shared-memory wavefronts were bank
// conflicts and the pipe was ~77 % busy.
__device__ __forceinline__ int mm_rho(int c) { return c < 4 ? 2 * c : 2 * (c - 4) + 1; }

// Fragment -> y.  Lane (g, t) holds rows 16R+g (x, y) and 16R+g+8 (z, w) for MMA columns 2t (x, z) and 2t+1 (y, w).
template <int NT>
__device__ __forceinline__ void mm_store(const MidmParams &p, int R, int lane, const float4 (&v)[NT]) {
    const int g = lane >> 2, t = lane & 3;
    const int r0 = mm_rho(2 * t), r1 = mm_rho(2 * t + 1);
    const int n0 = R * 16 + g;
    const float b0 = p.bias ? __bfloat162float(p.bias[n0]) : 0.f;
    const float b1 = p.bias ? __bfloat162float(p.bias[n0 + 8]) : 0.f;
    if (p.swiglu) {
        const size_t I = (size_t)(p.N >> 1), c = (size_t)R * 8 + g;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int tok0 = nt * 8 + r0, tok1 = nt * 8 + r1;
            if (tok0 < p.M) p.y[(size_t)tok0 * I + c] = swiglu_bf16(v[nt].x + b0, v[nt].z + b1);
            if (tok1 < p.M) p.y[(size_t)tok1 * I + c] = swiglu_bf16(v[nt].y + b0, v[nt].w + b1);
        }
        return;
    }
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
        const int tok0 = nt * 8 + r0, tok1 = nt * 8 + r1;
        if (tok0 < p.M) {
            p.y[(size_t)tok0 * p.N + n0] = __float2bfloat16_rn(v[nt].x + b0);
            p.y[(size_t)tok0 * p.N + n0 + 8] = __float2bfloat16_rn(v[nt].z + b1);
        }
        if (tok1 < p.M) {
            p.y[(size_t)tok1 * p.N + n0] = __float2bfloat16_rn(v[nt].y + b0);
            p.y[(size_t)tok1 * p.N + n0 + 8] = __float2bfloat16_rn(v[nt].w + b1);
        }
    }
}

__device__ __forceinline__ void mm_cp16(uint32_t dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ uint4 mm_lds128(uint32_t addr) {
    uint4 r;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
    return r;
}
__device__ __forceinline__ uint32_t mm_lds32(uint32_t addr) {
    uint32_t r;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r) : "r"(addr));
    return r;
}
__device__ __forceinline__ uint32_t mm_lds8(uint32_t addr) {
    uint32_t r;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(r) : "r"(addr));
    return r;
}

// NT = token tiles of 8 (M <= 8 NT).  GPS = groups per 64-k chunk: 1 (g >= 64 or per-channel), 2 (g = 32).
// grid = (cps, ns): CTA (cx, s) keeps k slice s of x in shared memory; its warp w works on row blocks cx*16 + w + i*cps*16.
// Weights: every warp pulls ITS chunks (512 B + the 48-byte meta block(s) of the chunk's group(s)) through a private ring of
// MM_RING stages with per-lane 16-byte cp.async -- bulk copies execute one at a time per SM, too slow at this size
// (w4a16_gemv_imma.cu), and a register ring deep enough to cover HBM latency (~50 KB in flight per SM) does not fit beside
// the accumulators.  The ring runs across the warp's row blocks: the next block's first chunks are in flight while this
// one finishes.
template <int NT, int GPS>
__global__ void __launch_bounds__(MM_THREADS, 1)
midm_kernel(const __grid_constant__ CUtensorMap x_map, const MidmParams p) {
    constexpr int TILE = NT * 1024;   // one 64-k block of x: 8 NT token rows x 128 bytes
    extern __shared__ uint8_t mm_smem_raw[];
    __shared__ uint64_t bars[MM_MAX_KBLKS];
    const uint32_t xs = (smem_u32(mm_smem_raw) + 1023u) & ~1023u;
    const uint32_t bar0 = smem_u32(bars);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    const int s = blockIdx.y;
    const int u0 = (int)(((long long)s * p.U) / p.ns), u1 = (int)(((long long)(s + 1) * p.U) / p.ns);
    const int kb0 = 2 * u0, nkb = 2 * (u1 - u0);
    const uint32_t ring = xs + (uint32_t)p.x_bytes + (uint32_t)warp * (MM_RING * MM_STAGE);

    // ---- weight ring: producer cursor (row block, k block) runs MM_RING chunks ahead of the consumer ----
    const int R0 = blockIdx.x * MM_WARPS + warp, Rstep = gridDim.x * MM_WARPS;
    // (running pointers: the address arithmetic of a from-scratch version was a fifth of the loop's issue slots)
    int pr_R = R0, pr_kbl = 0, pr_slot = 0;
    const uint4 *pr_w = p.qw + ((size_t)R0 * p.KB + kb0) * 32 + lane;
    const size_t pr_wjump = ((size_t)Rstep * p.KB - nkb) * 32 + 32;          // last chunk of a row block -> first of the next
    const uint8_t *pr_m = p.meta + (size_t)R0 * p.G * META_BYTES + lane * 16;   // lanes < 3 GPS: the row block's meta, 16 bytes each
    const size_t pr_mjump = (size_t)Rstep * p.G * META_BYTES;
    uint32_t pr_dst = ring + lane * 16;
    auto issue = [&]() {
        if (pr_R < p.NR) {
            mm_cp16(pr_dst, pr_w);
            if (lane < 3 * GPS) mm_cp16(pr_dst + 512, pr_m + (((kb0 + pr_kbl) * 64) >> p.gs_shift) * META_BYTES);
            if (++pr_kbl == nkb) { pr_kbl = 0; pr_R += Rstep; pr_w += pr_wjump; pr_m += pr_mjump; }
            else pr_w += 32;
        }
        asm volatile("cp.async.commit_group;" ::: "memory");   // an empty group past the end keeps the wait count uniform
        pr_dst += MM_STAGE;
        if (++pr_slot == MM_RING) { pr_slot = 0; pr_dst -= MM_RING * MM_STAGE; }
    };
#pragma unroll 1
    for (int i = 0; i < MM_RING; ++i) issue();   // weights are immutable: in flight before the previous kernel has finished

    if (threadIdx.x == 0) {
        for (int i = 0; i < nkb; ++i) mbar_init(bar0 + 8 * i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&x_map) : "memory");
    }
    __syncthreads();
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");   // x may be the previous kernel's output, y its input
    if (threadIdx.x == 0) {
        for (int i = 0; i < nkb; ++i) {
            mbar_arrive_expect_tx(bar0 + 8 * i, TILE);
            tma_load_2d(xs + i * TILE, &x_map, (kb0 + i) * 64, 0, bar0 + 8 * i);
        }
    }

    // this lane's 8 bytes of a B fragment: MMA column g = token row 8 nt + rho(g), k-slots 16 j + 4 t .. + 3 -> 16-byte piece
    // (2 j + t / 2) ^ rho(g) of that row
    uint32_t xl[4];
    const int xr = mm_rho(g);
#pragma unroll
    for (int j = 0; j < 4; ++j) xl[j] = xs + (uint32_t)(xr * 128 + (((2 * j + (t >> 1)) ^ xr) << 4) + (t & 1) * 8);
    const uint32_t my_w = ring + lane * 16, my_s = ring + 512 + 4 * g, my_z = ring + 512 + META_ZERO_OFF + g;
    int slot = 0;

    for (int R = R0; R < p.NR; R += Rstep) {
        float acc[NT][4];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[nt][i] = 0.f;

#pragma unroll 1
        for (int kbl = 0; kbl < nkb; ++kbl) {
            asm volatile("cp.async.wait_group %0;" ::"n"(MM_RING - 1) : "memory");
            __syncwarp();   // the meta block was copied by lanes 0..2
            const uint32_t so = (uint32_t)slot * MM_STAGE;
            const uint4 w = mm_lds128(my_w + so);
            uint32_t s2[GPS], zb[GPS];
#pragma unroll
            for (int h = 0; h < GPS; ++h) {
                s2[h] = mm_lds32(my_s + so + h * META_BYTES);
                zb[h] = mm_lds8(my_z + so + h * META_BYTES);
            }
            const uint32_t wj[4] = {w.x, w.y, w.z, w.w};
            uint32_t a[4][4];
            if (MM_DBG(2)) {
#pragma unroll
                for (int j = 0; j < 4; ++j)
#pragma unroll
                    for (int i = 0; i < 4; ++i) a[j][i] = wj[j] + s2[0] + zb[0];
            } else
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int h = GPS == 2 ? (j >> 1) : 0;
                const uint32_t zlo = zero_magic(zb[h] & 0xFu), zhi = zero_magic(zb[h] >> 4);
                const uint32_t slo2 = prmt(s2[h], 0, 0x1010), shi2 = prmt(s2[h], 0, 0x3232);
                a[j][0] = dequant_pair(wj[j], zlo, slo2);         // row g   , k-slots 4t, 4t+1
                a[j][1] = dequant_pair(wj[j] >> 4, zhi, shi2);    // row g+8 , k-slots 4t, 4t+1
                a[j][2] = dequant_pair(wj[j] >> 8, zlo, slo2);    // row g   , k-slots 4t+2, 4t+3
                a[j][3] = dequant_pair(wj[j] >> 12, zhi, shi2);   // row g+8 , k-slots 4t+2, 4t+3
            }
            __syncwarp();   // every lane holds its words of this stage: the slot may be refilled
            issue();
            if (++slot == MM_RING) slot = 0;
            if (R == R0) mbar_wait(bar0 + 8 * kbl, 0);   // this k block of x has landed (once per launch: the first row block waits)
            const uint32_t xo = (uint32_t)kbl * TILE;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                uint2 b[NT];
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) b[nt] = MM_DBG(4) ? make_uint2(a[j][0], a[j][1]) : mm_lds64(xl[j] + xo + nt * 1024);
                if (MM_DBG(1)) {
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt) acc[nt][0] += __uint_as_float(a[j][0] ^ a[j][1] ^ a[j][2] ^ a[j][3] ^ b[nt].x ^ b[nt].y);
                } else
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) mm_mma(acc[nt], a[j][0], a[j][1], a[j][2], a[j][3], b[nt].x, b[nt].y);
            }
        }

        float4 v[NT];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) v[nt] = make_float4(acc[nt][0], acc[nt][1], acc[nt][2], acc[nt][3]);
        if (p.ns == 1) {
            mm_store<NT>(p, R, lane, v);
            continue;
        }
        // split-K: publish this slice's fragment (fire and forget: no fence, no ticket here -- a fence + atomic round
        // trip per row block stalled the warp's weight stream for longer than the block itself took)
        float4 *part = p.part + (((size_t)s * p.NR + R) * NT) * 32 + lane;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) __stcg(part + nt * 32, v[nt]);
    }
    if (p.ns == 1 || R0 >= p.NR) return;

    // ---- tickets of ALL this warp's row blocks at once (lane i takes the i-th), after one fence; the warp that arrives
    //      last at a row block sums its slices in slice order (deterministic) and writes y ----
    __threadfence();
    __syncwarp();
    const int npass = (p.NR - R0 + Rstep - 1) / Rstep;
    for (int base = 0; base < npass; base += 32) {
        const int i = base + lane;
        int prev = -1;
        if (i < npass) {
            int *tk = p.tickets + R0 + i * Rstep;
            asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], 1;" : "=r"(prev) : "l"(tk) : "memory");
            if (prev == p.ns - 1) *tk = 0;   // self-reset for the next launch
        }
        unsigned mask = __ballot_sync(0xffffffffu, prev == p.ns - 1);
        if (mask) __threadfence();
        while (mask) {
            const int R = R0 + (base + __ffs(mask) - 1) * Rstep;
            mask &= mask - 1;
            float4 v[NT];
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) v[nt] = make_float4(0.f, 0.f, 0.f, 0.f);
            const float4 *src = p.part + ((size_t)R * NT) * 32 + lane;
            const size_t sstride = (size_t)p.NR * NT * 32;
            int s2 = 0;
            for (; s2 + 2 <= p.ns; s2 += 2) {   // two slices' loads in flight, added in slice order
                float4 o0[NT], o1[NT];
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) { o0[nt] = __ldcg(src + s2 * sstride + nt * 32); o1[nt] = __ldcg(src + (s2 + 1) * sstride + nt * 32); }
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    v[nt].x += o0[nt].x; v[nt].y += o0[nt].y; v[nt].z += o0[nt].z; v[nt].w += o0[nt].w;
                    v[nt].x += o1[nt].x; v[nt].y += o1[nt].y; v[nt].z += o1[nt].z; v[nt].w += o1[nt].w;
                }
            }
            if (s2 < p.ns) {
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    const float4 o = __ldcg(src + s2 * sstride + nt * 32);
                    v[nt].x += o.x; v[nt].y += o.y; v[nt].z += o.z; v[nt].w += o.w;
                }
            }
            mm_store<NT>(p, R, lane, v);
        }
    }
}

// ---- host side -----------------------------------------------------------------------
// Split-K scratch (fp32 fragments + self-resetting tickets): one per (device, STREAM), as the prefill kernel's
// (w4a16_prefill.cu): launches on one stream are ordered and may share it.  Grown on demand outside stream capture; a
// launch that would have to grow it under capture is refused (the caller falls back to the prefill kernel), and an
// outgrown block is never freed: a CUDA graph captured earlier may hold its address.
constexpr int MM_MAX_TICKETS = 16384;   // N <= 262144
struct MmScratch { int device; cudaStream_t stream; float4 *part = nullptr; size_t bytes = 0; int *tickets = nullptr; };
static std::vector<MmScratch> g_mm_scratch;
static std::mutex g_mm_mu;
static int mm_scratch(int device, size_t need, cudaStream_t stream, float4 **part, int **tickets) {
    std::lock_guard<std::mutex> lock(g_mm_mu);
    MmScratch *sc = nullptr;
    for (MmScratch &c : g_mm_scratch)
        if (c.device == device && c.stream == stream) { sc = &c; break; }
    *part = nullptr; *tickets = nullptr;
    if (!(sc && sc->bytes >= need && sc->tickets)) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(stream, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) return 0;
        if (!sc) {
            g_mm_scratch.push_back(MmScratch{device, stream});
            sc = &g_mm_scratch.back();
        }
        if (!sc->tickets) {
            if (cudaMalloc(&sc->tickets, MM_MAX_TICKETS * sizeof(int)) != cudaSuccess) { sc->tickets = nullptr; cudaGetLastError(); return 0; }
            cudaMemset(sc->tickets, 0, MM_MAX_TICKETS * sizeof(int));
        }
        if (sc->bytes < need) {
            const size_t want = std::max<size_t>(need, (size_t)32 << 20);
            float4 *fresh = nullptr;
            if (cudaMalloc(&fresh, want) != cudaSuccess) { cudaGetLastError(); return 0; }
            sc->part = fresh;          // the outgrown block stays allocated (see above)
            sc->bytes = want;
        }
    }
    *part = sc->part; *tickets = sc->tickets;
    return 0;
}

struct MmCfg { int nt, ns, cps; };

// Slices and CTAs per slice.  A slice's x must fit shared memory (8 NT KB per 128 k); within that, the cheapest of
//     passes x (chunks per slice + fixed per-pass cost) + per-slice cost (partials written, re-read, ticket)
// in units of one chunk of one warp, passes = row blocks per warp.
static MmCfg mm_pick(const Linear &L, int64_t M, int sms) {
    MmCfg c;
    c.nt = M <= 48 ? (int)((M + 7) / 8) : 8;   // token tiles of 8: 1 .. 6, or 8
    const int U = (int)(L.d.K / 128), NR = (int)L.d.NR;
    const int max_units = std::min(MM_MAX_KBLKS / 2, MM_SMEM_X_MAX / (c.nt * 2048));
    const int groups = (NR + MM_WARPS - 1) / MM_WARPS;
    static const int forced = exp_env_int("W4A16_MM_NS", 0);
    static const double c_pass = exp_env_double("W4A16_MM_CPASS", 6.0), c_slice = exp_env_double("W4A16_MM_CSLICE", 0.5);
    double best = -1;
    c.ns = 0; c.cps = 0;
    for (int ns = 1; ns <= U; ++ns) {
        if ((U + ns - 1) / ns > max_units) continue;
        const int cps = std::max(1, std::min(groups, sms / ns));
        const int passes = (NR + cps * MM_WARPS - 1) / (cps * MM_WARPS);
        const double cost = passes * (2.0 * ((U + ns - 1) / ns) + c_pass) + c_slice * ns;
        if ((forced > 0 && ns == forced) || (forced <= 0 && (best < 0 || cost < best))) { best = cost; c.ns = ns; c.cps = cps; }
        if (forced > 0 && ns == forced) break;
    }
    return c;
}

bool midm_usable(const Linear &L, int64_t M) {
    static const bool off = exp_env_int("W4A16_NO_MIDM", 0) != 0;
    const int gs = L.d.gs;
    static const int min_m = exp_env_int("W4A16_MM_MIN_M", 11), max_m = exp_env_int("W4A16_MM_MAX_M", 48);
    if (off || M < min_m || M > max_m || M > 64) return false;
    if (gs != 32 && gs != 64 && gs != 128 && gs != L.d.K) return false;
    return L.d.NR <= MM_MAX_TICKETS;
}

template <int NT, int GPS>
static int mm_launch(const CUtensorMap &map, const MidmParams &p, dim3 grid, size_t smem, int device, cudaStream_t stream) {
    static cudaError_t attr_err[64];
    static bool attr_done[64] = {};
    static std::mutex mu;
    if (device < 0 || device >= 64) return fail("launch_midm: device index out of range");
    {
        std::lock_guard<std::mutex> lk(mu);
        if (!attr_done[device]) {
            attr_err[device] = cudaFuncSetAttribute(midm_kernel<NT, GPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, MM_SMEM_X_MAX + MM_RING_BYTES + 1024);
            attr_done[device] = true;
        }
    }
    if (attr_err[device] != cudaSuccess)
        return fail(std::string("cudaFuncSetAttribute(midm_kernel smem): ") + cudaGetErrorString(attr_err[device]));
    // Programmatic dependent launch: in a stream of linears the next kernel's CTAs take an SM the moment one frees up, set up
    // their barriers and start their weight rings (weights are immutable) while this grid drains; x, y and the split-K scratch
    // are only touched behind griddepcontrol.wait.
    static const bool pdl = getenv("W4A16_NO_PDL") == nullptr || atoi(getenv("W4A16_NO_PDL")) == 0;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(MM_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    W4_CUDA_OK(cudaLaunchKernelEx(&cfg, midm_kernel<NT, GPS>, map, p));
    count_launch();
    return 0;
}

// Returns 0 = launched, 1 = error, 2 = not available for this call (no scratch under capture): use another kernel.
int launch_midm(const Linear &L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias, cudaStream_t stream) {
    if (!midm_usable(L, M)) return 2;
    const MmCfg c = mm_pick(L, M, L.scratch->sm_count);
    if (c.ns < 1) return 2;
    const int gs = L.d.gs;
    MidmParams p;
    p.qw = reinterpret_cast<const uint4 *>(L.qw);
    p.meta = L.meta;
    p.bias = add_bias && L.has_bias ? L.bias : nullptr;
    p.y = y;
    p.M = (int)M; p.K = (int)L.d.K; p.N = (int)L.d.N; p.KB = (int)L.d.KB; p.G = (int)L.d.G; p.NR = (int)L.d.NR;
    p.gs_shift = gs == L.d.K && gs != 32 && gs != 64 && gs != 128 ? 31 : (gs == 32 ? 5 : gs == 64 ? 6 : 7);
    p.ns = c.ns; p.cps = c.cps; p.U = (int)(L.d.K / 128);
    p.swiglu = L.epilogue == W4A16_EPILOGUE_SWIGLU;
    p.debug = exp_env_int("W4A16_MM_DEBUG", 0);
    p.part = nullptr; p.tickets = nullptr;
    if (c.ns > 1) {
        const size_t need = (size_t)c.ns * L.d.NR * c.nt * 32 * sizeof(float4);
        if (mm_scratch(L.device, need, stream, &p.part, &p.tickets)) return 1;
        if (!p.part) return 2;
    }
    CUtensorMap map;
    if (encode_x_map(&map, x, L.d.K, M, 8 * c.nt)) return 1;
    const int max_units = (p.U + c.ns - 1) / c.ns;
    p.x_bytes = max_units * 2 * c.nt * 1024;
    const size_t smem = (size_t)p.x_bytes + MM_RING_BYTES + 1024;
    const dim3 grid((unsigned)c.cps, (unsigned)c.ns);
    const int gps = gs == 32 ? 2 : 1;
#define W4_MM(NT_)                                                                                   \
    do {                                                                                             \
        if (gps == 1) return mm_launch<NT_, 1>(map, p, grid, smem, L.device, stream);                \
        return mm_launch<NT_, 2>(map, p, grid, smem, L.device, stream);                              \
    } while (0)
    if (c.nt == 1) W4_MM(1);
    if (c.nt == 2) W4_MM(2);
    if (c.nt == 3) W4_MM(3);
    if (c.nt == 4) W4_MM(4);
    if (c.nt == 5) W4_MM(5);
    if (c.nt == 6) W4_MM(6);
    W4_MM(8);
#undef W4_MM
}

}  // namespace w4
