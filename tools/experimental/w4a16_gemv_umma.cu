// w4a16_gemv_umma.cu -- decode (M <= 4 tokens) on the 5th-generation tensor cores.
//
// Replaces awq_gemv_int4_kt_bf16 / kt_mloop_m4 (awq_gemv.cu:377-665) and gptq_gemm_int4_bf16_tiled
// (gptq_dequant.cu:243-318) for the shapes the serving loop runs at M = 1..4.
//
// Why tensor cores for a GEMV: on sm_100a every CUDA-core instruction costs two issue cycles per SM
// sub-partition (16-lane pipes, measured by tools/ubench/pipes.cu), and the dot product itself
// (mma.sync / dp4a + x fragment loads + per-group scaling) was 2/3 of the instructions of the CUDA-core
// kernels -- they ran out of issue slots at ~45 % of the HBM rate.  Here the SM's warps only unpack:
//     chunk (512 B) -> 1 LDS.128 + 4 x (1 SHF + 4 LOP3) -> 16 fp16x2 registers -> ONE tcgen05.st.16x256b.x4
// and tcgen05.mma does all multiply-accumulate work: A = unpacked weights in TMEM, B = the activations in
// shared memory, D = fp32 accumulators in TMEM.  The unpack does not depend on the activations, so under
// programmatic dependent launch it (and the TMA bulk copies feeding it) runs BEFORE griddepcontrol.wait.
//
// Arithmetic (same regrouping as w4a16_gemv.cu):
//     y[m,n] = sum_g s[g,n] * 2^e[g,m] * ( sum_{k in g} (c + q[k,n]) * xt[m,k]  -  (c + z[g,n]) * X[g,m] )
// with xt = x * 2^-e[g,m] exact in fp16, X the fp32 group sums of xt, c = 1024 for rows 0..7 of a 16-row
// block and 64 for rows 8..15 (fp16 magic exponents: one LOP3 per register).  tcgen05.mma accumulates each
// quantisation group into its own TMEM column block, so scales and zero points are applied once per group in
// the epilogue, in fp32.
//
// Mapping.  A CTA owns whole 16-row blocks.  Every row block is cut into 4 K-parts; (row block, part) is a
// "virtual row" that occupies one 16-lane slab of the 128 TMEM lanes, so 8 virtual rows = 2 row blocks run
// per pass, every pass keeps all four SM sub-partitions busy (TMEM lane quarter q is written by warps with
// warp % 4 == q), and the parts of a row block are distinguished by the B columns they read:
// B tile of step i = [16 columns = 4 parts x 4 token slots][64 k], column (p, m) holds xt[m] for the k-chunk
// of part p at step i.  D[slab][group][column] is only meaningful in the slab's own part columns; the other
// columns are never read.  The four parts of a row block are summed in a fixed order in the epilogue.
#include <cuda.h>

#include <algorithm>
#include <vector>

#include "w4a16_internal.cuh"

namespace w4 {

constexpr int GU_THREADS = 448;        // 14 warps: 8 unpack, TMA producer, MMA issuer, 4 epilogue
constexpr int GU_UNPACK_WARPS = 8;
constexpr int GU_N = 16;               // UMMA N: 4 parts x 4 token slots
constexpr int GU_PARTS = 4;
constexpr int GU_SLOTS = 4;            // token slots per part (M <= 4)
constexpr int GU_SS = 4;               // steps (64-k chunks) per weight ring stage
constexpr int GU_WSTAGE_BYTES = 8 * GU_SS * CHUNK_BYTES;   // 8 slabs x 4 chunks = 16 KB
constexpr int GU_MAX_WSTAGES = 6;
constexpr int GU_MAX_ASTAGES = 12;
constexpr int GU_A_STAGE_COLS = 32;    // 64 k of fp16 = 32 32-bit TMEM columns
constexpr uint32_t GU_SPIN_LIMIT = 1u << 26;

struct UmmaGemvParams {
    const uint8_t *qw;
    const uint8_t *meta;
    const __nv_bfloat16 *bias;
    const __nv_bfloat16 *x;
    __nv_bfloat16 *y;
    const int *rb_start;        // [grid + 1] first row block of every CTA
    int M, K, N, KB, G;
    int NS;                     // steps per pass = KB / 4
    int gs_eff;                 // k per accumulator column block (group size, or the part length for per-channel)
    int G_part;                 // accumulator column blocks per part
    int meta_group_stride;      // 48, or 0 for per-channel
    int n_wstages, n_astages, tmem_cols;
    int off_btiles, off_gtab, off_meta, off_part, off_wring;
    int max_rb;                 // row blocks per CTA upper bound (sizes meta / part)
    unsigned long long *dbg;
};

// ---- PTX wrappers ------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t gu_smem(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void gu_mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void gu_mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void gu_mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ uint32_t gu_mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok;
}
// A wait that cannot hang the GPU: a lost arrival traps (the launch fails loudly).
__device__ __forceinline__ void gu_mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t spins = 0;
    while (!gu_mbar_try_wait(bar, parity))
        if (++spins > GU_SPIN_LIMIT) __trap();
}
__device__ __forceinline__ void gu_bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ uint32_t gu_elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred;
}
__device__ __forceinline__ void gu_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void gu_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void gu_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void gu_mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void gu_st_16x256b_x4(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.16x256b.x4.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
__device__ __forceinline__ void gu_ld_32x32b_x16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ uint32_t gu_lop3_and_or(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
// K-major SWIZZLE_128B shared-memory matrix descriptor: start >> 4, SBO = 1024 B (8 rows x 128 B), version 1,
// layout type 2 (same encoding as w4a16_prefill.cu::make_b_desc).
__device__ __forceinline__ uint64_t gu_b_desc(uint32_t smem_addr) {
    return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// kind::f16 instruction descriptor: D = f32 (bit 4), A = B = f16 (format 0), K-major, N >> 3 at 17, M >> 4 at 24.
__host__ __device__ constexpr uint32_t gu_idesc(int umma_m, int umma_n) {
    return (1u << 4) | ((uint32_t)(umma_n >> 3) << 17) | ((uint32_t)(umma_m >> 4) << 24);
}

// Barrier block at the start of dynamic shared memory.
struct GuBars {
    uint64_t full_w[GU_MAX_WSTAGES];    // weight ring stage landed (TMA complete_tx)
    uint64_t empty_w[GU_MAX_WSTAGES];   // all 8 unpack warps are done with the stage
    uint64_t a_full[GU_MAX_ASTAGES];    // all 8 slabs of an A stage are in TMEM
    uint64_t a_empty[GU_MAX_ASTAGES];   // the MMAs that read the A stage have retired (tcgen05.commit)
    uint64_t d_full, d_empty;           // accumulators of a pass complete / drained by the epilogue
    uint64_t meta_full;
    uint32_t tmem_base;
    uint32_t pad;
};

template <int GA>   // k per accumulator column block: 32, 64 or 128
__global__ void __launch_bounds__(GU_THREADS, 2) gemv_umma_kernel(const UmmaGemvParams p) {
    extern __shared__ uint8_t gu_smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(gu_smem_raw) + 1023) & ~(uintptr_t)1023);
    GuBars *bars = reinterpret_cast<GuBars *>(smem);
    uint8_t *btiles = smem + p.off_btiles;                              // [NS][16 rows][128 B], 128B swizzle
    float2 *gtab = reinterpret_cast<float2 *>(smem + p.off_gtab);       // [K/GA][4] {X, 2^e}
    const uint8_t *meta_s = smem + p.off_meta;                          // [rb][G][48]
    float *part = reinterpret_cast<float *>(smem + p.off_part);         // [vrow][16 rows][4 tokens]
    const uint8_t *wring = smem + p.off_wring;                          // [stage][slab][GU_SS][512]
    const uint32_t wring_s = gu_smem(wring);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int cta = blockIdx.x;
    const int Ra = __ldg(p.rb_start + cta), Rb = __ldg(p.rb_start + cta + 1);
    const int r = Rb - Ra;
    const int nv = r * GU_PARTS;                         // virtual rows of this CTA
    const int n_pass = (nv + 7) >> 3;
    const int NS = p.NS;
    const int total_steps = n_pass * NS;
    const int S = p.n_wstages, T = p.n_astages;
    const uint32_t bar_full_w = gu_smem(bars->full_w), bar_empty_w = gu_smem(bars->empty_w);
    const uint32_t bar_a_full = gu_smem(bars->a_full), bar_a_empty = gu_smem(bars->a_empty);
    const uint32_t bar_d_full = gu_smem(&bars->d_full), bar_d_empty = gu_smem(&bars->d_empty);
    const uint32_t bar_meta = gu_smem(&bars->meta_full);
    const int D_COLS = p.G_part * GU_N;                  // accumulator columns, then the A ring

    auto stamp = [&](int i) {
        if (p.dbg && tid == 0) {
            unsigned long long tt;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tt));
            p.dbg[(size_t)cta * 8 + i] = tt;
        }
    };

    // ---- set-up: barriers, TMEM, first weight stages (nothing here depends on the previous kernel) ----
    if (tid == 0) {
        for (int s = 0; s < S; ++s) { gu_mbar_init(bar_full_w + 8 * s, 1); gu_mbar_init(bar_empty_w + 8 * s, GU_UNPACK_WARPS); }
        for (int t = 0; t < T; ++t) { gu_mbar_init(bar_a_full + 8 * t, GU_UNPACK_WARPS); gu_mbar_init(bar_a_empty + 8 * t, 1); }
        gu_mbar_init(bar_d_full, 1);
        gu_mbar_init(bar_d_empty, 4);
        gu_mbar_init(bar_meta, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 9) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(gu_smem(&bars->tmem_base)), "r"((uint32_t)p.tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    {   // zero the B tiles: rows of unused token slots must be finite
        uint4 *bt = reinterpret_cast<uint4 *>(btiles);
        const int n16 = NS * GU_N * 8;
        for (int i = tid; i < n16; i += GU_THREADS) bt[i] = make_uint4(0, 0, 0, 0);
    }
    gu_fence_before();
    __syncthreads();
    gu_fence_after();
    const uint32_t tmem_base = bars->tmem_base;
    stamp(0);

    // weight ring stage j (global index over passes): 8 slabs x GU_SS consecutive chunks of one virtual row each
    auto issue_wstage = [&](int j) {
        const int steps_per_pass_stage = NS / GU_SS;
        const int pass = j / steps_per_pass_stage, sj = j - pass * steps_per_pass_stage;
        const int slot = j % S;
        const int act = min(8, nv - pass * 8);
        const uint32_t bar = bar_full_w + 8 * slot;
        gu_mbar_expect_tx(bar, (uint32_t)act * GU_SS * CHUNK_BYTES);
        for (int sl = 0; sl < act; ++sl) {
            const int v = pass * 8 + sl;
            const int R = Ra + (v >> 2), part_i = v & 3;
            const size_t chunk = (size_t)R * p.KB + (size_t)part_i * NS + (size_t)sj * GU_SS;
            gu_bulk_g2s(wring_s + (uint32_t)(slot * GU_WSTAGE_BYTES + sl * GU_SS * CHUNK_BYTES),
                        p.qw + chunk * CHUNK_BYTES, GU_SS * CHUNK_BYTES, bar);
        }
    };
    const int total_wstages = total_steps / GU_SS;
    if (warp == 8 && lane == 0) {
        const uint32_t mbytes = (uint32_t)r * (p.meta_group_stride ? (uint32_t)p.G * META_BYTES : META_BYTES);
        gu_mbar_expect_tx(bar_meta, mbytes);
        gu_bulk_g2s(gu_smem(meta_s), p.meta + (size_t)Ra * (p.meta_group_stride ? (size_t)p.G * META_BYTES : META_BYTES),
                    mbytes, bar_meta);
        const int first = min(S, total_wstages);
        for (int j = 0; j < first; ++j) issue_wstage(j);
    }

    // ---- unpack warps: W4T chunk -> fp16 fragments -> TMEM (A operand); independent of the activations ----
    int u_step = 0, u_i = 0, u_pass = 0;          // global step, step in pass, pass
    int u_wslot = 0; uint32_t u_wpar = 0;         // weight ring slot / parity of the stage holding this step
    int u_t = 0; uint32_t u_apar = 0;             // A ring stage / parity
    const int slab = 2 * (warp & 3) + ((warp >> 2) & 1);                 // unpack warp w writes TMEM lanes 16*slab..+15
    const uint32_t a_lane_addr = tmem_base + ((uint32_t)(slab * 16) << 16) + (uint32_t)D_COLS;
    auto unpack_steps = [&](int limit) {
        while (u_step < limit) {
            const int sub = u_i & (GU_SS - 1);
            if (sub == 0) gu_mbar_wait(bar_full_w + 8 * u_wslot, u_wpar);
            gu_mbar_wait(bar_a_empty + 8 * u_t, u_apar ^ 1u);
            gu_fence_after();
            if (u_pass * 8 + slab < nv) {
                const uint4 w4 = *(reinterpret_cast<const uint4 *>(wring + u_wslot * GU_WSTAGE_BYTES +
                                                                     (slab * GU_SS + sub) * CHUNK_BYTES) + lane);
                const uint32_t wj[4] = {w4.x, w4.y, w4.z, w4.w};
                uint32_t rr[16];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint32_t w = wj[j], v = w >> 8;
                    rr[4 * j + 0] = gu_lop3_and_or(w, 0x000F000Fu, 0x64006400u);   // row g  , k 4t, 4t+1  : 1024 + q
                    rr[4 * j + 1] = gu_lop3_and_or(v, 0x000F000Fu, 0x64006400u);   // row g  , k 4t+2, 4t+3
                    rr[4 * j + 2] = gu_lop3_and_or(w, 0x00F000F0u, 0x54005400u);   // row g+8, k 4t, 4t+1  :   64 + q
                    rr[4 * j + 3] = gu_lop3_and_or(v, 0x00F000F0u, 0x54005400u);   // row g+8, k 4t+2, 4t+3
                }
                gu_st_16x256b_x4(a_lane_addr + (uint32_t)(u_t * GU_A_STAGE_COLS), rr);
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            }
            gu_fence_before();
            __syncwarp();
            if (lane == 0) {
                gu_mbar_arrive(bar_a_full + 8 * u_t);
                if (sub == GU_SS - 1) gu_mbar_arrive(bar_empty_w + 8 * u_wslot);
            }
            ++u_step;
            if (sub == GU_SS - 1 && ++u_wslot == S) { u_wslot = 0; u_wpar ^= 1u; }
            if (++u_t == T) { u_t = 0; u_apar ^= 1u; }
            if (++u_i == NS) { u_i = 0; ++u_pass; }
        }
    };
    // fill the A ring before the dependency resolves (only from the weight stages already requested)
    if (warp < GU_UNPACK_WARPS) unpack_steps(min(min(T, S * GU_SS), total_steps));

    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    stamp(1);

    // ---- stage the activations: per-(group, token) power-of-two scale, fp16, UMMA B tiles (K-major, 128B swizzle) ----
    // piece c = 8 consecutive k of token m: k-chunk kc = c/8 -> part = kc / NS, step i = kc % NS, tile row = part*4 + m,
    // 16-byte unit (c % 8) of the 128-byte row, XOR-swizzled with (row & 7).
    {
        const int ppr = p.K >> 3;
        for (int m = 0; m < p.M; ++m) {
            for (int c0 = tid - lane; c0 < ppr; c0 += GU_THREADS) {   // warp-uniform trip count: the shuffles need all lanes
                const int c = c0 + lane;
                const bool live = c < ppr;
                uint4 v = make_uint4(0, 0, 0, 0);
                if (live) v = __ldg(reinterpret_cast<const uint4 *>(p.x + (size_t)m * p.K) + c);
                const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
                uint32_t mx = 0;
#pragma unroll
                for (int i = 0; i < 4; ++i) mx = max(mx, max(w4[i] & 0x7FFFu, (w4[i] >> 16) & 0x7FFFu));
#pragma unroll
                for (int o = GA / 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
                int E = (int)(mx >> 7);                                  // biased exponent of the group's max |x| (bf16 bits)
                E = min(max(E, 14), 254);
                const float scale = __int_as_float((267 - E) << 23);     // max|x| of the group lands in [2^13, 2^14)
                uint32_t h2[4];
                float sum = 0.f;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const __half2 hh = __floats2half2_rn(__uint_as_float(w4[i] << 16) * scale,
                                                         __uint_as_float(w4[i] & 0xFFFF0000u) * scale);
                    h2[i] = *reinterpret_cast<const uint32_t *>(&hh);
                    const float2 back = __half22float2(hh);
                    sum += back.x + back.y;
                }
#pragma unroll
                for (int o = GA / 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
                if (live) {
                    const int kc = c >> 3, part_i = kc / NS, i = kc - part_i * NS;
                    const int row = part_i * GU_SLOTS + m;
                    uint8_t *dst = btiles + (size_t)i * (GU_N * 128) + (row >> 3) * 1024 + (row & 7) * 128 +
                                   (((c & 7) ^ (row & 7)) << 4);
                    *reinterpret_cast<uint4 *>(dst) = make_uint4(h2[0], h2[1], h2[2], h2[3]);
                    if ((c & (GA / 8 - 1)) == 0)
                        gtab[(size_t)(c / (GA / 8)) * GU_SLOTS + m] = make_float2(sum, __int_as_float((E - 13) << 23));
                }
            }
        }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // B tiles: generic-proxy writes -> tensor-core reads
    __syncthreads();
    stamp(2);

    if (warp < GU_UNPACK_WARPS) {
        unpack_steps(total_steps);
    } else if (warp == 8) {
        // ===== TMA producer: refill weight ring slots as the unpack warps release them =====
        if (lane == 0) {
            int slot = 0; uint32_t par = 0;
            for (int j = S; j < total_wstages; ++j) {
                gu_mbar_wait(bar_empty_w + 8 * slot, par);
                issue_wstage(j);
                if (++slot == S) { slot = 0; par ^= 1u; }
            }
        }
    } else if (warp == 9) {
        // ===== MMA issuer: one elected lane drives the tensor core =====
        constexpr uint32_t idesc = gu_idesc(128, GU_N);
        const uint32_t btiles_s = gu_smem(btiles);
        int t = 0; uint32_t apar = 0;
        for (int pass = 0; pass < n_pass; ++pass) {
            if (pass > 0) {   // the epilogue has drained the previous pass's accumulators
                gu_mbar_wait(bar_d_empty, (uint32_t)(pass - 1) & 1u);
                gu_fence_after();
            }
            for (int i = 0; i < NS; ++i) {
                gu_mbar_wait(bar_a_full + 8 * t, apar);
                gu_fence_after();
                if (gu_elect_one()) {
                    const uint64_t bdesc = gu_b_desc(btiles_s + i * (GU_N * 128));
                    const uint32_t a_tmem = tmem_base + (uint32_t)D_COLS + (uint32_t)(t * GU_A_STAGE_COLS);
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) {
                        const int koff = i * 64 + kk * 16;                  // k offset inside the part
                        const int blk = koff / GA;
                        gu_mma_ts(tmem_base + (uint32_t)(blk * GU_N), a_tmem + kk * 8, bdesc + 2 * kk, idesc,
                                  (uint32_t)((koff & (GA - 1)) != 0));
                    }
                    gu_commit(bar_a_empty + 8 * t);
                    if (i == NS - 1) gu_commit(bar_d_full);
                }
                __syncwarp();
                if (++t == T) { t = 0; apar ^= 1u; }
            }
        }
    } else {
        // ===== epilogue warps 10..13: thread = TMEM lane = one output row of one virtual row =====
        const int q = warp & 3;
        const int L = q * 32 + lane;                  // TMEM lane
        const int sl = L >> 4, row16 = L & 15;
        gu_mbar_wait(bar_meta, 0);
        const float c_row = row16 < 8 ? 1024.f : 64.f;
        for (int pass = 0; pass < n_pass; ++pass) {
            const int v = pass * 8 + sl;
            const bool act = v < nv;
            const int rb = act ? (v >> 2) : 0, part_i = v & 3;
            float yp[GU_SLOTS] = {0.f, 0.f, 0.f, 0.f};
            gu_mbar_wait(bar_d_full, (uint32_t)pass & 1u);
            gu_fence_after();
            for (int blk = 0; blk < p.G_part; ++blk) {
                uint32_t d[16];
                gu_ld_32x32b_x16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(blk * GU_N), d);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (act) {
                    const int kblk = part_i * p.G_part + blk;                         // GA-block index along K
                    const int grp = p.meta_group_stride ? (kblk * GA) / p.gs_eff : 0;  // quantisation group
                    const uint8_t *mb = meta_s + ((size_t)rb * (p.meta_group_stride ? p.G : 1) + grp) * META_BYTES;
                    const float s = __uint_as_float((uint32_t)*reinterpret_cast<const uint16_t *>(mb + 4 * (row16 & 7) + 2 * (row16 >> 3)) << 16);
                    const uint32_t zb = mb[META_ZERO_OFF + (row16 & 7)];
                    const float z = (float)((zb >> (4 * (row16 >> 3))) & 0xFu) + c_row;
#pragma unroll
                    for (int m = 0; m < GU_SLOTS; ++m) {
                        const float2 xe = gtab[(size_t)kblk * GU_SLOTS + m];
                        const float dv = __uint_as_float(part_i == 0 ? d[m] : part_i == 1 ? d[4 + m] : part_i == 2 ? d[8 + m] : d[12 + m]);
                        yp[m] = fmaf(s * xe.y, fmaf(-z, xe.x, dv), yp[m]);
                    }
                }
            }
            gu_fence_before();
            __syncwarp();
            if (lane == 0) gu_mbar_arrive(bar_d_empty);
            if (act) *reinterpret_cast<float4 *>(part + ((size_t)v * 16 + row16) * GU_SLOTS) = make_float4(yp[0], yp[1], yp[2], yp[3]);
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        // combine the 4 parts of every row block in part order, add bias, store
        const int et = tid - 10 * 32;
        for (int it = et; it < r * 16; it += 128) {
            const int rb = it >> 4, rw = it & 15;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int pi = 0; pi < GU_PARTS; ++pi) {
                const float4 o = *reinterpret_cast<const float4 *>(part + ((size_t)(rb * 4 + pi) * 16 + rw) * GU_SLOTS);
                acc.x += o.x; acc.y += o.y; acc.z += o.z; acc.w += o.w;
            }
            const int n = (Ra + rb) * 16 + rw;
            const float b = p.bias ? __bfloat162float(p.bias[n]) : 0.f;
            const float vals[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
            for (int m = 0; m < GU_SLOTS; ++m)
                if (m < p.M) p.y[(size_t)m * p.N + n] = __float2bfloat16_rn(vals[m] + b);
        }
    }
    stamp(3);

    gu_fence_before();
    __syncthreads();
    if (warp == 9) {
        gu_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)p.tmem_cols) : "memory");
    }
    stamp(5);
}

// ---- host side ---------------------------------------------------------------------------------------
static int gu_env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

int gemv_umma_build_table(Linear &L, int sm_count) {
    L.gu_grid = 0;
    if (L.d.NR > (int64_t)1 << 24) return 0;
    const int grid = (int)std::min<int64_t>(L.d.NR, sm_count);
    std::vector<int> start(grid + 1);
    for (int i = 0; i <= grid; ++i) start[i] = (int)((int64_t)i * L.d.NR / grid);
    cudaFree(L.gu_rb_start); L.gu_rb_start = nullptr;
    W4_CUDA_OK(cudaMalloc(&L.gu_rb_start, start.size() * sizeof(int)));
    W4_CUDA_OK(cudaMemcpy(L.gu_rb_start, start.data(), start.size() * sizeof(int), cudaMemcpyHostToDevice));
    L.gu_grid = grid;
    L.gu_max_rb = (int)((L.d.NR + grid - 1) / grid);
    return 0;
}

struct UmmaPlan {
    bool ok;
    int GA, NS, G_part, S, T, tmem_cols;
    int off_btiles, off_gtab, off_meta, off_part, off_wring, smem_bytes;
    bool per_channel;
};

static UmmaPlan gemv_umma_plan(const Linear &L, int64_t M) {
    UmmaPlan pl{};
    pl.ok = false;
    if (M < 1 || M > GU_SLOTS || L.gu_grid <= 0 || !L.gu_rb_start) return pl;
    if (!gu_env_int("W4A16_UMMA_GEMV", 0)) return pl;   // opt-in: measured slower than the dp4a stream kernel at M <= 4
    const int gs = L.d.gs;
    pl.per_channel = gs == L.d.K && gs != 32 && gs != 64 && gs != 128;
    if (!pl.per_channel && gs != 32 && gs != 64 && gs != 128) return pl;
    if (L.d.K % 1024) return pl;                                   // 4 parts x whole ring stages of 4 chunks
    pl.GA = pl.per_channel ? 128 : gs;
    pl.NS = (int)(L.d.KB / GU_PARTS);
    pl.G_part = (int)(L.d.K / GU_PARTS / pl.GA);
    const int d_cols = pl.G_part * GU_N;
    if (d_cols + 2 * GU_A_STAGE_COLS > 512) return pl;
    pl.tmem_cols = d_cols + 4 * GU_A_STAGE_COLS <= 256 ? 256 : 512;
    if (gu_env_int("W4A16_UMMA_TMEM", 0) == 512) pl.tmem_cols = 512;
    pl.T = std::min(GU_MAX_ASTAGES, (pl.tmem_cols - d_cols) / GU_A_STAGE_COLS);
    int64_t off = 512;                                             // barriers
    off = (off + 1023) & ~(int64_t)1023;
    pl.off_btiles = (int)off; off += (int64_t)pl.NS * GU_N * 128;
    pl.off_gtab = (int)off; off += (L.d.K / pl.GA) * GU_SLOTS * 8;
    off = (off + 15) & ~(int64_t)15;
    pl.off_meta = (int)off; off += (int64_t)L.gu_max_rb * (pl.per_channel ? 1 : L.d.G) * META_BYTES;
    off = (off + 15) & ~(int64_t)15;
    pl.off_part = (int)off; off += (int64_t)L.gu_max_rb * GU_PARTS * 16 * GU_SLOTS * 4;
    off = (off + 127) & ~(int64_t)127;
    const int64_t fixed = off + 1024;                              // + slack for the manual 1 KB alignment
    int64_t budget = pl.tmem_cols == 256 ? 112 * 1024 : 220 * 1024;
    int64_t S = (budget - fixed) / GU_WSTAGE_BYTES;
    if (S < 3) { budget = 220 * 1024; S = (budget - fixed) / GU_WSTAGE_BYTES; }
    if (S < 2) return pl;
    S = std::min<int64_t>(S, std::min(GU_MAX_WSTAGES, gu_env_int("W4A16_UMMA_WSTAGES", GU_MAX_WSTAGES)));
    pl.S = (int)S;
    pl.off_wring = (int)off; off += S * GU_WSTAGE_BYTES;
    pl.smem_bytes = (int)off + 1024;
    pl.ok = true;
    return pl;
}

bool gemv_umma_usable(const Linear &L, int64_t M) { return gemv_umma_plan(L, M).ok; }

unsigned long long *gemv_debug_slot();   // w4a16_gemv.cu: next slot of the shared stamp ring, or null

template <int GA>
static int gu_launch_t(const UmmaGemvParams &p, const UmmaPlan &pl, int grid, cudaStream_t stream, bool pdl) {
    static bool attr_set = false;
    if (!attr_set) {
        W4_CUDA_OK(cudaFuncSetAttribute(gemv_umma_kernel<GA>, cudaFuncAttributeMaxDynamicSharedMemorySize, 221 * 1024));
        attr_set = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(GU_THREADS);
    cfg.dynamicSmemBytes = (size_t)pl.smem_bytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    W4_CUDA_OK(cudaLaunchKernelEx(&cfg, gemv_umma_kernel<GA>, p));
    count_launch();
    return 0;
}

int launch_gemv_umma(const Linear &L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, bool add_bias,
                     cudaStream_t stream) {
    const UmmaPlan pl = gemv_umma_plan(L, M);
    if (!pl.ok) return fail("launch_gemv_umma: shape not supported");
    UmmaGemvParams p;
    p.qw = reinterpret_cast<const uint8_t *>(L.qw);
    p.meta = L.meta;
    p.bias = add_bias && L.has_bias ? L.bias : nullptr;
    p.x = x; p.y = y;
    p.rb_start = L.gu_rb_start;
    p.M = (int)M; p.K = (int)L.d.K; p.N = (int)L.d.N; p.KB = (int)L.d.KB; p.G = (int)L.d.G;
    p.NS = pl.NS;
    p.gs_eff = pl.per_channel ? (int)L.d.K : L.d.gs;
    p.G_part = pl.G_part;
    p.meta_group_stride = pl.per_channel ? 0 : META_BYTES;
    p.n_wstages = pl.S; p.n_astages = pl.T; p.tmem_cols = pl.tmem_cols;
    p.off_btiles = pl.off_btiles; p.off_gtab = pl.off_gtab; p.off_meta = pl.off_meta; p.off_part = pl.off_part;
    p.off_wring = pl.off_wring;
    p.max_rb = L.gu_max_rb;
    p.dbg = gemv_debug_slot();
    const bool pdl = gu_env_int("W4A16_NO_PDL", 0) == 0;
    if (pl.GA == 128) return gu_launch_t<128>(p, pl, L.gu_grid, stream, pdl);
    if (pl.GA == 64) return gu_launch_t<64>(p, pl, L.gu_grid, stream, pdl);
    return gu_launch_t<32>(p, pl, L.gu_grid, stream, pdl);
}

}  // namespace w4
