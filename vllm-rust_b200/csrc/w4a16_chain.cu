// w4a16_chain.cu -- a CHAIN of decode linears (M <= 4 tokens) executed by ONE persistent launch.
//
// Why: launched one by one, every decode linear pays a dependency hand-over (0.65 us), the activation load + staging
// (1.0-1.6 us) and the drain of its shared-memory ring + combine (1.0-1.6 us), and while it pays them HBM idles unless
// the NEXT linear's resident CTA still has ring space to fill -- which is 96 KB per SM, one hop ahead, and qkv / o are
// smaller than that (profiles/r2_decode_timeline.md: 25.4 us per Qwen3-8B layer against 15.3 us at copy speed).  The
// dependency itself cannot be removed (linear i+1 reads what linear i wrote), but the weight stream does not depend on
// anything: here ONE CTA per SM owns ONE ring of 5-6 stages of 33.5 KB for the whole chain, a producer warp keeps it
// full ACROSS the boundaries between linears (weights of linear i+1, i+2, ... arrive while linear i still computes and
// while the grid meets at the barrier), and the consumer warps run the int8 tensor-core GEMV of w4a16_gemv_imma2.cu on
// whatever stage comes next.  Between consecutive linears: combine the warps' partial sums, write y, grid-wide barrier
// (one release-add + acquire-poll per CTA on a monotonic counter), stage the next x.  HBM never waits for any of it as
// long as the ring (28 MB chip-wide) outlasts the barrier.
//
// Replaces, for the decode step, the per-linear launches of the reference's layer loop (qwen3_quantized.rs:704-709
// attention projections, :532-551 gate/up -> silu*mul -> down; each one a marlin_gemm_pooled launch,
// marlin_cuda.rs:1655-1753) and its CUDA-graph replay of that loop (engine/cuda_graph.rs).  Arithmetic, W4T stream and
// the activation quantisation are those of gemv_imma_kernel / gemv_imma2_kernel (see there).
#include <algorithm>
#include <mutex>

#include "w4a16_gemv_common.cuh"

namespace w4 {

constexpr int CH_CONS_WARPS = 16;                     // consumer warps: 2 units each = one 32-unit stage per iteration
constexpr int CH_UPW = 2;
constexpr int CH_CONS_THREADS = CH_CONS_WARPS * 32;   // 512
constexpr int CH_THREADS = CH_CONS_THREADS + 128;     // + the producer warp group (one working warp; setmaxnreg moves whole groups of 4)
constexpr int CH_REGS_PRODUCER = 24, CH_REGS_CONSUMER = 112;   // after the role split (launched with 96 each)
constexpr int CH_MAX_STAGES = 8;
constexpr int CH_SMEM_MAX = 227 * 1024;
constexpr int CH_MAX_OPS = 4096;
constexpr int CH_DBG_WORDS = 8;                       // stamps per (CTA, op): op start, staged, streamed, y written, hint seen, x valid, -, -

struct ChainOp {
    const uint8_t *qw, *meta;
    const __nv_bfloat16 *bias, *x;
    __nv_bfloat16 *y;
    const int *rb;             // [grid + 1]: CTA c owns row blocks [rb[c], rb[c + 1])
    int M, K, N, UPR;
    int spr;                   // K / 4096 when K is a multiple of 4096 (a stage = a whole 4096-k slab of one row block), else 0
    int swiglu, hoist;
    int barrier_before;        // a grid-wide barrier separates this linear from the one before it (hazards the mailboxes do not cover)
    // Mailboxes: y of a linear that a later linear of the chain reads is ALSO stored into a chain-owned copy whose
    // words hold the sentinel 0xFFFF until written; the reader polls that copy instead of waiting at a barrier.
    unsigned short *mbox_out;        // parity-0 copy of this linear's y ([M][out width]) or null
    const unsigned short *mbox_in;   // parity-0 address of this linear's x inside its producer's mailbox, or null (read op.x)
    // Arrival hint of a mailbox: every CTA adds 1 (relaxed, no fence) once its part of y is on its way; the reader's
    // thread 0 polls this ONE word instead of 148 x 512 threads hammering the mailbox lines the writers still have to
    // reach, and only then are the pieces read -- and still validated against the sentinel, because the hint orders nothing.
    unsigned long long *arrive_out;        // this linear's own counter (null without a mailbox)
    const unsigned long long *arrive_in;   // the producer's counter (null without mbox_in)
    // Prologue (w4a16_chain_prologue): h = x (+ residual), residual := h, the linear consumes rms_norm(h) * norm_w.
    const __nv_bfloat16 *norm_w;           // [K] gamma, or null: no prologue
    const __nv_bfloat16 *res_in;           // [M, K] residual read from the caller's memory, or null
    __nv_bfloat16 *res_out, *normed_out;   // [M, K] h / the normalised row, written by CTA 0; may be null
    float eps;
    int add_running;                       // h = x + the residual the previous prologue of the chain left in shared memory
    int exchange;                          // row-parallel shard with a peer group: 1-based index among the chain's all-reduces, else 0
    const int32_t *perm;                   // host-side bookkeeping only (act-order permutation before it becomes scatter / ingest)
    // The (single) reader of this linear's mailbox is an act-order linear: columns [sc_off, sc_off + sc_len) of the mailbox
    // are stored in THE READER'S tiled order (column sc_off + c at sc_off + scatter[c]), so that it reads a gathered row
    // sequentially -- the gather costs one index load per produced element instead of 148 CTAs gathering the whole row.
    const int32_t *scatter;
    int sc_off, sc_len;
    // An act-order linear whose x does NOT come from a mailbox it owns alone (the chain's input, a shared mailbox, a
    // buffer behind a barrier): the grid gathers x[:, perm] ONCE, every CTA a slice, into this chain-owned mailbox
    // (ingest_perm = the permutation, ingest_src = where x really is), and then reads it like any other mailbox.
    const int32_t *ingest_perm;
    const unsigned short *ingest_src;      // parity-0 address if ingest_src_mbox, else the caller's buffer
    int ingest_src_mbox;
    const unsigned long long *ingest_src_arrive;
    unsigned short *ingest;                // parity-0 address of the gathered copy (== mbox_in of this linear)
    unsigned long long *ingest_arrive;     // (== arrive_in of this linear)
};

struct ChainParams {
    const ChainOp *ops;
    int n_ops;
    unsigned long long *counter;   // grid-barrier arrivals, monotonic over the life of the chain
    unsigned int *runs;            // launches completed so far (read by every CTA at start, advanced by CTA 0 at the end)
    int n_barriers;                // grid barriers per run
    long long mbox_stride;         // elements between the two parities of every mailbox (run r uses parity r & 1)
    int S, slot_stride;            // ring stages, bytes per stage slot
    int l2pf;                      // stages the L2 prefetcher runs ahead of the ring's producer (0: off)
    int off_xs, off_gtab, off_part, off_ring;
    int off_res, off_red;          // prologue: running residual bf16 [M][K], reduction scratch (floats); -1 / unused without one
    unsigned long long *dbg;       // [n_ops][grid][CH_DBG_WORDS] %globaltimer stamps, or null
    PeerView peer;                 // tensor-parallel group (w4a16_chain_set_peer); world <= 1: none
    int n_exchange;                // all-reduces per run
};

__device__ __forceinline__ void ch_imma(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ch_bar_consumers() { asm volatile("bar.sync 1, %0;" ::"n"(CH_CONS_THREADS) : "memory"); }
__device__ __forceinline__ uint32_t ch_mbar_test(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok;
}
__device__ __forceinline__ void ch_mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// Activation staging of one linear by the 512 consumer threads: identical to gemv_imma2_kernel's (per-(group, token)
// power-of-two scale, two signed-byte planes in B-fragment order, integer group sums, Inf / NaN poison the group), but x
// is read with ld.global.cg: another CTA of THIS launch wrote it a barrier ago.
// reductions over an aligned group of lanes (one instruction instead of a shuffle tree)
__device__ __forceinline__ uint32_t ch_redux_max(uint32_t v, uint32_t mask) {
    uint32_t r;
    asm volatile("redux.sync.max.u32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(mask));
    return r;
}
__device__ __forceinline__ int ch_redux_add(int v, uint32_t mask) {
    int r;
    asm volatile("redux.sync.add.s32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(mask));
    return r;
}
// Group = GL aligned lanes.  redux.sync with a partial member mask goes through a slow divergent path (the uniform
// datapath reduces whole warps), so 16-lane groups use two FULL-warp reductions over inputs masked to one half each.
template <int GL>
__device__ __forceinline__ uint32_t ch_group_max(uint32_t v, uint32_t gmask, int lane) {
    if constexpr (GL == 16) {
        const uint32_t lo = ch_redux_max(lane < 16 ? v : 0u, 0xFFFFFFFFu), hi = ch_redux_max(lane < 16 ? 0u : v, 0xFFFFFFFFu);
        return lane < 16 ? lo : hi;
    } else {
        return ch_redux_max(v, gmask);
    }
}
template <int GL>
__device__ __forceinline__ int ch_group_add(int v, uint32_t gmask, int lane) {
    if constexpr (GL == 16) {
        const int lo = ch_redux_add(lane < 16 ? v : 0, 0xFFFFFFFFu), hi = ch_redux_add(lane < 16 ? 0 : v, 0xFFFFFFFFu);
        return lane < 16 ? lo : hi;
    } else {
        return ch_redux_add(v, gmask);
    }
}
// column of a mailbox row where output column `col` goes (see ChainOp::scatter)
__device__ __forceinline__ int ch_mbox_col(const int32_t *scatter, int sc_off, int sc_len, int col) {
    const int c = col - sc_off;
    return scatter && c >= 0 && c < sc_len ? sc_off + __ldg(scatter + c) : col;
}
// a value can never be the sentinel: 0xFFFF (a NaN) is stored as the canonical NaN 0x7FFF
__device__ __forceinline__ void ch_mbox_store(unsigned short *p, __nv_bfloat16 v) {
    unsigned short b = __bfloat16_as_ushort(v);
    if (b == 0xFFFFu) b = 0x7FFFu;
    asm volatile("st.relaxed.gpu.global.u16 [%0], %1;" ::"l"(p), "h"(b) : "memory");
}
__device__ __forceinline__ uint4 ch_ld_relaxed(const uint4 *p) {
    uint4 v;
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
// any 16-bit half of the four words equal to 0xFFFF (the mailbox sentinel)?
__device__ __forceinline__ bool ch_has_sentinel(const uint4 &v) {
    uint32_t hit = 0;
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) { const uint32_t t = ~w[q]; hit |= (t - 0x00010001u) & ~t & 0x80008000u; }
    return hit != 0;
}

// One 16-byte piece of x: straight from the caller's buffer, or from a mailbox (re-read until no value is the sentinel).
__device__ __forceinline__ uint4 ch_load_piece(const uint4 *p, bool poll) {
    if (!poll) return __ldcg(p);
    uint4 v = ch_ld_relaxed(p);
    uint32_t spins = 0;
    while (ch_has_sentinel(v)) {
        v = ch_ld_relaxed(p);
        if (++spins > GV_SPIN_LIMIT) __trap();
    }
    return v;
}
__device__ __forceinline__ float ch_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float ch_hi(uint32_t w) { return __uint_as_float(w & 0xFFFF0000u); }
__device__ __forceinline__ uint32_t ch_pack(float lo, float hi) {
    const __nv_bfloat162 b = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t *>(&b);
}

// Prologue, first pass (the reference's add_bf16 + rms_norm_bf16 between the linears of a layer,
// qwen3_quantized.rs:1043-1075, activations.cu:99-113, layernorm.cu:59-126): h = bf16(f32(x) + f32(residual)) into the
// shared-memory residual (and the caller's residual_out, by CTA 0), sum of squares of the ROUNDED h per token in fp32,
// inv_rms[m] = rsqrt(sum / K + eps) into red[0..MR).  The second pass is the ordinary staging reading h from shared
// memory and scaling it (ch_stage_batch, `norm`).
template <int MR>
__device__ __forceinline__ void ch_prologue_pass1(const ChainOp &op, const uint4 *x16, bool poll, uint4 *res, float *red, int tid,
                                                  int lane, int warp, bool first_cta) {
    const int ppr = op.K >> 3, total = op.M * ppr;
    float ss[MR];
#pragma unroll
    for (int m = 0; m < MR; ++m) ss[m] = 0.f;
    const uint4 *rin = reinterpret_cast<const uint4 *>(op.res_in);
    uint4 *rout = reinterpret_cast<uint4 *>(op.res_out);
    for (int i = tid; i < total; i += CH_CONS_THREADS) {
        uint4 v = ch_load_piece(x16 + i, poll);
        if (rin || op.add_running) {
            const uint4 r = rin ? __ldcg(rin + i) : res[i];
            uint32_t *vw = &v.x;
            const uint32_t *rw = &r.x;
#pragma unroll
            for (int q = 0; q < 4; ++q) vw[q] = ch_pack(ch_lo(vw[q]) + ch_lo(rw[q]), ch_hi(vw[q]) + ch_hi(rw[q]));
        }
        res[i] = v;
        if (rout && first_cta) rout[i] = v;
        float s = 0.f;
        const uint32_t *vw = &v.x;
#pragma unroll
        for (int q = 0; q < 4; ++q) { const float a = ch_lo(vw[q]), b = ch_hi(vw[q]); s = fmaf(a, a, s); s = fmaf(b, b, s); }
        const int m = MR == 1 ? 0 : (i >= ppr) + (MR > 2 ? (i >= 2 * ppr) + (i >= 3 * ppr) : 0);
#pragma unroll
        for (int k = 0; k < MR; ++k) ss[k] += k == m ? s : 0.f;
    }
#pragma unroll
    for (int m = 0; m < MR; ++m) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ss[m] += __shfl_xor_sync(0xffffffffu, ss[m], o);
        if (lane == 0) red[MR + warp * MR + m] = ss[m];
    }
    ch_bar_consumers();
    if (tid < MR) {
        float t = 0.f;
        for (int w = 0; w < CH_CONS_WARPS; ++w) t += red[MR + w * MR + tid];   // fixed order: deterministic
        red[tid] = rsqrtf(t / (float)op.K + op.eps);
    }
    ch_bar_consumers();
}

// One batch of NB 16-byte pieces per thread (pieces base + r * 512 + tid).  poll: x is a mailbox (see ChainOp): every
// piece is re-read until none of its eight values is the sentinel.  The arithmetic of the NB pieces runs side by side
// (independent dependency chains), the shared-memory stores follow.
// MODE 0: x from the caller's buffer / a mailbox; 1: x16 is the shared-memory residual of a prologue, scaled by
// inv_rms * gamma on the way.
template <int MR, int MPU, int NB, int MODE>
__device__ __forceinline__ void ch_stage_batch(const uint4 *x16, uint8_t *xs, int2 *gtab, int ppr, int total, int base, int tid, int lane,
                                               bool poll, unsigned long long *dbg_valid, const uint4 *norm_w, const float *inv_rms,
                                               uint4 *normed_out) {
    constexpr int COLS = 2 * MR;
    constexpr int GL = (GV_UNIT_K / MPU) / 8;
    const uint32_t gmask = (GL >= 32 ? 0xFFFFFFFFu : ((1u << GL) - 1u)) << (lane & ~(GL - 1));
    uint4 v[NB];
#pragma unroll
    for (int r = 0; r < NB; ++r) {
        const int i = base + r * CH_CONS_THREADS + tid;
        if constexpr (MODE != 0) v[r] = i < total ? x16[i] : make_uint4(0, 0, 0, 0);
        else v[r] = i < total ? (poll ? ch_ld_relaxed(x16 + i) : __ldcg(x16 + i)) : make_uint4(0, 0, 0, 0);
    }
    if constexpr (MODE == 1) {   // x16 is the shared-memory residual h of the prologue: the linear consumes bf16(h * inv_rms * gamma)
#pragma unroll
        for (int r = 0; r < NB; ++r) {
            const int i = base + r * CH_CONS_THREADS + tid;
            if (i < total) {
                const int m = MR == 1 ? 0 : (i >= ppr) + (MR > 2 ? (i >= 2 * ppr) + (i >= 3 * ppr) : 0);
                const uint4 gw = __ldg(norm_w + (i - m * ppr));
                const float inv = inv_rms[m];
                uint32_t *vw = &v[r].x;
                const uint32_t *g4 = &gw.x;
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    vw[q] = ch_pack(ch_lo(vw[q]) * inv * ch_lo(g4[q]), ch_hi(vw[q]) * inv * ch_hi(g4[q]));
                if (normed_out) normed_out[i] = v[r];
            }
        }
    }
    if (MODE == 0 && poll) {
        uint32_t spins = 0;
        for (;;) {
            bool bad[NB], any = false;
#pragma unroll
            for (int r = 0; r < NB; ++r) {
                bad[r] = base + r * CH_CONS_THREADS + tid < total && ch_has_sentinel(v[r]);
                any |= bad[r];
            }
            if (!any) break;
#pragma unroll
            for (int r = 0; r < NB; ++r)
                if (bad[r]) v[r] = ch_ld_relaxed(x16 + base + r * CH_CONS_THREADS + tid);
            if (++spins > GV_SPIN_LIMIT) __trap();
        }
        if (dbg_valid && tid == 0 && base == 0) {
            unsigned long long tt;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tt));
            dbg_valid[0] = tt;
            dbg_valid[1] = spins;
        }
        __syncwarp();
    }
    if (dbg_valid) {   // profiling only: when has EVERY thread of the CTA got its pieces?  (word 7 of the stamps)
        ch_bar_consumers();
        if (tid == 0) {
            unsigned long long tt;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tt));
            dbg_valid[2] = tt;
        }
    }
    uint32_t pkA[NB][4], mxA[NB];
    int sumA[NB], EA[NB];
#pragma unroll
    for (int r = 0; r < NB; ++r) {
        const uint32_t w4[4] = {v[r].x, v[r].y, v[r].z, v[r].w};
        uint32_t mx = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) mx = max(mx, max(w4[q] & 0x7FFFu, (w4[q] >> 16) & 0x7FFFu));
        mxA[r] = mx;
    }
#pragma unroll
    for (int r = 0; r < NB; ++r) mxA[r] = ch_group_max<GL>(mxA[r], gmask, lane);   // over the GL lanes that hold the group
#pragma unroll
    for (int r = 0; r < NB; ++r) {
        const uint32_t w4[4] = {v[r].x, v[r].y, v[r].z, v[r].w};
        int E = (int)(mxA[r] >> 7);
        E = min(max(E, 14), 254);
        EA[r] = E;
        const float scale = __int_as_float((267 - E) << 23);     // max|x| lands in [2^13, 2^14)
        int sum = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float lo = fmaf(__uint_as_float(w4[q] << 16), scale, 12582912.f);
            const float hi = fmaf(__uint_as_float(w4[q] & 0xFFFF0000u), scale, 12582912.f);
            pkA[r][q] = __byte_perm(__float_as_uint(lo), __float_as_uint(hi), 0x5410);
            sum = __dp2a_lo((int)pkA[r][q], 0x0101, sum);
        }
        sumA[r] = sum;
    }
#pragma unroll
    for (int r = 0; r < NB; ++r) sumA[r] = ch_group_add<GL>(sumA[r], gmask, lane);
#pragma unroll
    for (int r = 0; r < NB; ++r) {
        const int i = base + r * CH_CONS_THREADS + tid;
        const uint32_t *pk = pkA[r];
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
            if constexpr (MR >= 2) {   // rotate the order of the four stores per lane group: 16-way -> 4-way bank conflicts
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
                gtab[(size_t)(c / GL) * MR + m] = make_int2(sumA[r], mxA[r] >= 0x7F80u ? 0x7FC00000 : (EA[r] - 13) << 23);
        }
    }
}

// Activation staging of one linear by the 512 consumer threads: the arithmetic of gemv_imma_kernel (per-(group, token)
// power-of-two scale, two signed-byte planes in B-fragment order, integer group sums, Inf / NaN poison the group).
template <int MR, int MPU, int MODE = 0>
__device__ __forceinline__ void ch_stage_x(const __nv_bfloat16 *x, uint8_t *xs, int2 *gtab, int M, int K, int tid, int lane, bool poll,
                                           unsigned long long *dbg_valid, const uint4 *norm_w = nullptr, const float *inv_rms = nullptr,
                                           uint4 *normed_out = nullptr) {
    const int ppr = K >> 3, total = M * ppr;
    const uint4 *x16 = reinterpret_cast<const uint4 *>(x);
    int base = 0;
    for (; total - base > 2 * CH_CONS_THREADS; base += 3 * CH_CONS_THREADS)   // (the batch guards pieces past the end)
        ch_stage_batch<MR, MPU, 3, MODE>(x16, xs, gtab, ppr, total, base, tid, lane, poll, dbg_valid, norm_w, inv_rms, normed_out);
    const int left = (total - base + CH_CONS_THREADS - 1) / CH_CONS_THREADS;   // 0..2 (warp-uniform)
    if (left == 2) ch_stage_batch<MR, MPU, 2, MODE>(x16, xs, gtab, ppr, total, base, tid, lane, poll, dbg_valid, norm_w, inv_rms, normed_out);
    else if (left == 1) ch_stage_batch<MR, MPU, 1, MODE>(x16, xs, gtab, ppr, total, base, tid, lane, poll, dbg_valid, norm_w, inv_rms, normed_out);
}

// The row blocks of one linear, stage after stage from the ring; every warp takes two units (wi, wi + 16) of every
// stage, so a stage passes through the CTA in the latency of 8 IMMAs.  {slot, parity} = the CTA's next stage.
template <int MR, int MPU, bool HOIST>
__device__ __forceinline__ void ch_row_blocks(const ChainParams &p, const ChainOp &op, int n_slots, int wi, int lane,
                                              uint32_t ring, uint32_t bar_full, uint32_t bar_empty, uint32_t xs_u32, uint32_t gtab_u32,
                                              float2 *part, int &slot, uint32_t &parity) {
    constexpr int COLS = 2 * MR;
    constexpr int IPG = 4 / MPU;
    constexpr uint32_t STAGE_W = 32 * GV_UNIT_BYTES;
    constexpr uint32_t XU = 4 * 4 * COLS * 8;  // bytes of B fragments per unit
    constexpr uint32_t GU = MPU * MR * 8;      // bytes of group table per unit
    const int g = lane >> 2, t = lane & 3;
    const uint32_t xs_l = xs_u32 + (uint32_t)((t * COLS + (g & (COLS - 1))) * 8);
    const uint32_t gt_l = gtab_u32 + (uint32_t)((t & (MR - 1)) * 8);
    const bool b_lane = COLS >= 8 || g < COLS;   // only the lanes of the data columns need real B values

    uint32_t bq[HOIST ? CH_UPW : 1][4][2];
    if constexpr (HOIST) {   // K = 4096: the warp meets the same four k-units in every row block
#pragma unroll
        for (int j = 0; j < CH_UPW; ++j) {
            const uint32_t xk = xs_l + (uint32_t)(wi + CH_CONS_WARPS * j) * XU;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                bq[j][i][0] = 0; bq[j][i][1] = 0;
                if (b_lane) asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(bq[j][i][0]), "=r"(bq[j][i][1]) : "r"(xk + (uint32_t)(i * 4 * COLS * 8)));
            }
        }
    }
    const uint32_t offW = (uint32_t)(wi * GV_UNIT_BYTES + lane * 16);
    const uint32_t offM = STAGE_W + (uint32_t)(wi * MPU * META_BYTES);
    const int spr = op.spr;
    for (int s = 0; s < n_slots; ++s) {
        float y0 = 0.f, y1 = 0.f;                  // rows g / g+8 of this row block, token slot t
        for (int m = 0; m < spr; ++m) {
            const uint32_t sb = ring + (uint32_t)slot * (uint32_t)p.slot_stride;
            gv_mbar_wait(bar_full + 8 * slot, parity);
            uint32_t w[CH_UPW][8], s2[CH_UPW][MPU], zb[CH_UPW][MPU];
#pragma unroll
            for (int j = 0; j < CH_UPW; ++j) {
                const uint32_t sw = sb + offW + (uint32_t)(CH_CONS_WARPS * j * GV_UNIT_BYTES), sm = sb + offM + (uint32_t)(CH_CONS_WARPS * j * MPU * META_BYTES);
                asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(w[j][0]), "=r"(w[j][1]), "=r"(w[j][2]), "=r"(w[j][3]) : "r"(sw));
                asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4+512];" : "=r"(w[j][4]), "=r"(w[j][5]), "=r"(w[j][6]), "=r"(w[j][7]) : "r"(sw));
#pragma unroll
                for (int h = 0; h < MPU; ++h) {
                    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(s2[j][h]) : "r"(sm + h * META_BYTES + 4 * g));
                    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(zb[j][h]) : "r"(sm + h * META_BYTES + META_ZERO_OFF + g));
                }
            }
            __syncwarp();                       // every lane holds its words: the slot may be refilled
            if (lane == 0) ch_mbar_arrive(bar_empty + 8 * slot);
            int acc[CH_UPW][MPU][4];
#pragma unroll
            for (int j = 0; j < CH_UPW; ++j)
#pragma unroll
                for (int h = 0; h < MPU; ++h)
#pragma unroll
                    for (int q = 0; q < 4; ++q) acc[j][h][q] = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
#pragma unroll
                for (int j = 0; j < CH_UPW; ++j) {
                    const uint32_t wa = w[j][2 * i], wb = w[j][2 * i + 1];
                    uint32_t b0, b1;
                    if constexpr (HOIST) {
                        b0 = bq[j][i][0]; b1 = bq[j][i][1];
                    } else {
                        b0 = 0; b1 = 0;
                        if (b_lane)
                            asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(b0), "=r"(b1)
                                         : "r"(xs_l + (uint32_t)(32 * m + wi + CH_CONS_WARPS * j) * XU + (uint32_t)(i * 4 * COLS * 8)));
                    }
                    // rows g: the low nibbles; rows g+8: the RAW bytes 16 q[g+8] + q[g] (rows g are subtracted afterwards)
                    ch_imma(acc[j][i / IPG], wa & 0x0F0F0F0Fu, wa, wb & 0x0F0F0F0Fu, wb, b0, b1);
                }
            }
#pragma unroll
            for (int j = 0; j < CH_UPW; ++j) {
#pragma unroll
                for (int h = 0; h < MPU; ++h) {
                    int X, e2b;      // HOIST: m == 0 always
                    asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(X), "=r"(e2b)
                                 : "r"(gt_l + (uint32_t)(32 * m + wi + CH_CONS_WARPS * j) * GU + (uint32_t)(h * MR * 8)));
                    const float e2 = __int_as_float(e2b);
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
            if (++slot == p.S) { slot = 0; parity ^= 1u; }
        }
        if (t < MR) part[(((size_t)s * CH_CONS_WARPS + wi) * 8 + g) * MR + t] = make_float2(y0, y1);
    }
}

// General path (any K % 128 == 0): the CTA's share of the linear is the unit range [R0 * UPR, R1 * UPR) of the W4T stream
// (a unit = 128 k of one 16-row block), cut into stages of 32 consecutive units whatever the row blocks are; the last
// stage may be short.  Warp wi takes units wi and wi + 16 of every stage; a unit's row block and k position follow from
// its index, its partial sums are added to the warp's entry of that row block in `part` (zeroed by the caller) whenever
// the row block changes -- for K >= 4096 that is once per several stages.
template <int MR, int MPU>
__device__ __forceinline__ void ch_row_blocks_general(const ChainParams &p, const ChainOp &op, int n_slots, int wi, int lane,
                                              uint32_t ring, uint32_t bar_full, uint32_t bar_empty, uint32_t xs_u32, uint32_t gtab_u32,
                                              float2 *part, int &slot, uint32_t &parity) {
    constexpr int COLS = 2 * MR;
    constexpr int IPG = 4 / MPU;
    constexpr uint32_t STAGE_W = 32 * GV_UNIT_BYTES;
    constexpr uint32_t XU = 4 * 4 * COLS * 8;  // bytes of B fragments per unit
    constexpr uint32_t GU = MPU * MR * 8;      // bytes of group table per unit
    const int g = lane >> 2, t = lane & 3;
    const uint32_t xs_l = xs_u32 + (uint32_t)((t * COLS + (g & (COLS - 1))) * 8);
    const uint32_t gt_l = gtab_u32 + (uint32_t)((t & (MR - 1)) * 8);
    const bool b_lane = COLS >= 8 || g < COLS;   // only the lanes of the data columns need real B values
    const uint32_t offW = (uint32_t)(wi * GV_UNIT_BYTES + lane * 16);
    const uint32_t offM = STAGE_W + (uint32_t)(wi * MPU * META_BYTES);
    const int UPR = op.UPR, n_units = n_slots * UPR, nst = (n_units + 31) >> 5;
    float y0 = 0.f, y1 = 0.f;                  // rows g / g+8 of row block `cur`, token slot t
    int cur = -1;
    auto flush = [&]() {
        if (cur >= 0 && t < MR) {
            float2 *e = part + (((size_t)cur * CH_CONS_WARPS + wi) * 8 + g) * MR + t;
            float2 v = *e;
            v.x += y0; v.y += y1;
            *e = v;
        }
    };
    for (int s = 0; s < nst; ++s) {
        const uint32_t sb = ring + (uint32_t)slot * (uint32_t)p.slot_stride;
        gv_mbar_wait(bar_full + 8 * slot, parity);
        uint32_t w[CH_UPW][8], s2[CH_UPW][MPU], zb[CH_UPW][MPU];
        bool valid[CH_UPW];
#pragma unroll
        for (int j = 0; j < CH_UPW; ++j) {
            valid[j] = 32 * s + wi + CH_CONS_WARPS * j < n_units;      // warp-uniform
            if (valid[j]) {
                const uint32_t sw = sb + offW + (uint32_t)(CH_CONS_WARPS * j * GV_UNIT_BYTES), sm = sb + offM + (uint32_t)(CH_CONS_WARPS * j * MPU * META_BYTES);
                asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(w[j][0]), "=r"(w[j][1]), "=r"(w[j][2]), "=r"(w[j][3]) : "r"(sw));
                asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4+512];" : "=r"(w[j][4]), "=r"(w[j][5]), "=r"(w[j][6]), "=r"(w[j][7]) : "r"(sw));
#pragma unroll
                for (int h = 0; h < MPU; ++h) {
                    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(s2[j][h]) : "r"(sm + h * META_BYTES + 4 * g));
                    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(zb[j][h]) : "r"(sm + h * META_BYTES + META_ZERO_OFF + g));
                }
            }
        }
        __syncwarp();                       // every lane holds its words: the slot may be refilled
        if (lane == 0) ch_mbar_arrive(bar_empty + 8 * slot);
        int rl[CH_UPW], ku[CH_UPW];          // row block (slot) and k position of each unit
        int acc[CH_UPW][MPU][4];
#pragma unroll
        for (int j = 0; j < CH_UPW; ++j) {
            const int ul = 32 * s + wi + CH_CONS_WARPS * j;          // unit index inside the CTA's range
            rl[j] = ul / UPR;
            ku[j] = ul - rl[j] * UPR;
#pragma unroll
            for (int h = 0; h < MPU; ++h)
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[j][h][q] = 0;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
#pragma unroll
            for (int j = 0; j < CH_UPW; ++j) {                        // the units' IMMA chains interleaved
                if (!valid[j]) continue;
                const uint32_t wa = w[j][2 * i], wb = w[j][2 * i + 1];
                uint32_t b0 = 0, b1 = 0;
                if (b_lane)
                    asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(b0), "=r"(b1) : "r"(xs_l + (uint32_t)ku[j] * XU + (uint32_t)(i * 4 * COLS * 8)));
                // rows g: the low nibbles; rows g+8: the RAW bytes 16 q[g+8] + q[g] (rows g are subtracted afterwards)
                ch_imma(acc[j][i / IPG], wa & 0x0F0F0F0Fu, wa, wb & 0x0F0F0F0Fu, wb, b0, b1);
            }
        }
#pragma unroll
        for (int j = 0; j < CH_UPW; ++j) {
            if (!valid[j]) continue;
            if (rl[j] != cur) { flush(); cur = rl[j]; y0 = 0.f; y1 = 0.f; }
#pragma unroll
            for (int h = 0; h < MPU; ++h) {
                int X, e2b;
                asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(X), "=r"(e2b) : "r"(gt_l + (uint32_t)ku[j] * GU + (uint32_t)(h * MR * 8)));
                const float e2 = __int_as_float(e2b);
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
        if (++slot == p.S) { slot = 0; parity ^= 1u; }
    }
    flush();
}

// K = 4096, group size 128 (one stage = one row block; the warp meets the same two k-units in every stage, so its B
// fragments live in registers): NS stages per call.  Two stages side by side double the independent IMMA chains and the
// loads in flight per warp and halve the barrier round trips -- a stage passes through the CTA in ~0.5 us when taken one
// by one, which is what a linear whose weights already sit in the ring (qkv, o, the head of gate_up) is bound by.
template <int MR, int NS>
__device__ __forceinline__ void ch_hoist_stages(const ChainParams &p, int wi, int lane, uint32_t ring, uint32_t bar_full, uint32_t bar_empty,
                                                uint32_t gt_l, const uint32_t (&bq)[CH_UPW][4][2], float2 *part, int s, int &slot,
                                                uint32_t &parity) {
    constexpr uint32_t STAGE_W = 32 * GV_UNIT_BYTES;
    constexpr uint32_t GU = MR * 8;
    const int g = lane >> 2, t = lane & 3;
    const uint32_t offW = (uint32_t)(wi * GV_UNIT_BYTES + lane * 16);
    const uint32_t offM = STAGE_W + (uint32_t)(wi * META_BYTES);
    uint32_t w[NS][CH_UPW][8], s2[NS][CH_UPW], zb[NS][CH_UPW];
#pragma unroll
    for (int n = 0; n < NS; ++n) {
        const uint32_t sb = ring + (uint32_t)slot * (uint32_t)p.slot_stride;
        if (n == 0) gv_mbar_wait(bar_full + 8 * slot, parity);      // the caller has seen the later stages complete
#pragma unroll
        for (int j = 0; j < CH_UPW; ++j) {
            const uint32_t sw = sb + offW + (uint32_t)(CH_CONS_WARPS * j * GV_UNIT_BYTES), sm = sb + offM + (uint32_t)(CH_CONS_WARPS * j * META_BYTES);
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(w[n][j][0]), "=r"(w[n][j][1]), "=r"(w[n][j][2]), "=r"(w[n][j][3]) : "r"(sw));
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4+512];" : "=r"(w[n][j][4]), "=r"(w[n][j][5]), "=r"(w[n][j][6]), "=r"(w[n][j][7]) : "r"(sw));
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(s2[n][j]) : "r"(sm + 4 * g));
            asm volatile("ld.shared.u8 %0, [%1];" : "=r"(zb[n][j]) : "r"(sm + META_ZERO_OFF + g));
        }
        __syncwarp();
        if (lane == 0) ch_mbar_arrive(bar_empty + 8 * slot);
        if (++slot == p.S) { slot = 0; parity ^= 1u; }
    }
    int acc[NS][CH_UPW][4];
#pragma unroll
    for (int n = 0; n < NS; ++n)
#pragma unroll
        for (int j = 0; j < CH_UPW; ++j)
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[n][j][q] = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int n = 0; n < NS; ++n)
#pragma unroll
            for (int j = 0; j < CH_UPW; ++j) {
                const uint32_t wa = w[n][j][2 * i], wb = w[n][j][2 * i + 1];
                ch_imma(acc[n][j], wa & 0x0F0F0F0Fu, wa, wb & 0x0F0F0F0Fu, wb, bq[j][i][0], bq[j][i][1]);
            }
    int X[CH_UPW];
    float e2[CH_UPW];
#pragma unroll
    for (int j = 0; j < CH_UPW; ++j) {
        int e2b;
        asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(X[j]), "=r"(e2b) : "r"(gt_l + (uint32_t)(wi + CH_CONS_WARPS * j) * GU));
        e2[j] = __int_as_float(e2b);
    }
#pragma unroll
    for (int n = 0; n < NS; ++n) {
        float y0 = 0.f, y1 = 0.f;
#pragma unroll
        for (int j = 0; j < CH_UPW; ++j) {
            const int tg = acc[n][j][0] * 128 + acc[n][j][1];
            const int tr = acc[n][j][2] * 128 + acc[n][j][3];
            const int ug = tg - (int)(zb[n][j] & 0x0Fu) * X[j];
            const int ug8 = tr - tg - (int)(zb[n][j] & 0xF0u) * X[j];          // 16 * (row g+8 - z8 * X)
            const float slo = __uint_as_float(s2[n][j] << 16) * e2[j];
            const float shi = __uint_as_float(s2[n][j] & 0xFFFF0000u) * (e2[j] * 0.0625f);
            y0 = fmaf(slo, (float)ug, y0);
            y1 = fmaf(shi, (float)ug8, y1);
        }
        if (t < MR) part[(((size_t)(s + n) * CH_CONS_WARPS + wi) * 8 + g) * MR + t] = make_float2(y0, y1);
    }
}

template <int MR>
__device__ __forceinline__ void ch_row_blocks_hoist(const ChainParams &p, int n_slots, int wi, int lane, uint32_t ring, uint32_t bar_full,
                                                    uint32_t bar_empty, uint32_t xs_u32, uint32_t gtab_u32, float2 *part, int &slot,
                                                    uint32_t &parity) {
    constexpr int COLS = 2 * MR;
    constexpr uint32_t XU = 4 * 4 * COLS * 8;
    const int g = lane >> 2, t = lane & 3;
    const uint32_t xs_l = xs_u32 + (uint32_t)((t * COLS + (g & (COLS - 1))) * 8);
    const uint32_t gt_l = gtab_u32 + (uint32_t)((t & (MR - 1)) * 8);
    const bool b_lane = COLS >= 8 || g < COLS;
    uint32_t bq[CH_UPW][4][2];
#pragma unroll
    for (int j = 0; j < CH_UPW; ++j) {
        const uint32_t xk = xs_l + (uint32_t)(wi + CH_CONS_WARPS * j) * XU;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            bq[j][i][0] = 0; bq[j][i][1] = 0;
            if (b_lane) asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(bq[j][i][0]), "=r"(bq[j][i][1]) : "r"(xk + (uint32_t)(i * 4 * COLS * 8)));
        }
    }
    int s = 0;
    while (s < n_slots) {
        bool two = false;
        if (s + 1 < n_slots) {               // is the stage after this one in the ring already?  (warp-uniform answer)
            int nslot = slot + 1;
            uint32_t npar = parity;
            if (nslot == p.S) { nslot = 0; npar ^= 1u; }
            two = __all_sync(0xffffffffu, ch_mbar_test(bar_full + 8 * nslot, npar) != 0);   // test_wait: never suspends
        }
        if (two) {
            ch_hoist_stages<MR, 2>(p, wi, lane, ring, bar_full, bar_empty, gt_l, bq, part, s, slot, parity);
            s += 2;
        } else {
            ch_hoist_stages<MR, 1>(p, wi, lane, ring, bar_full, bar_empty, gt_l, bq, part, s, slot, parity);
            s += 1;
        }
    }
}

// PRO: the chain has prologues (w4a16_chain_prologue); a separate instantiation, so that chains of plain linears carry
// none of it (merely compiling the prologue into the one kernel cost the plain step 6 %: 0.826 -> 0.875 ms).
// GEN: some linear of the chain has an in_features that is not a multiple of 4096 (its stages cross row-block boundaries);
// again a separate instantiation: with the general loop compiled in, the Qwen3-8B step lost 14 % (0.732 -> 0.849 ms).
// TP: row-parallel linears of the chain all-reduce their partial sums over the peers' symmetric buffers inside their
// epilogue (the exchange of gemv_imma_kernel<.., FUSED>, w4a16_gemv_imma.cu: every rank stores its fp32 partials into
// slot [epoch % 3][rank] of every rank, polls its own slots until no word is the sentinel, sums in rank order).
template <int MR, int MPU, bool PRO, bool GEN, bool TP>
__global__ void __launch_bounds__(CH_THREADS, 1) chain_kernel(const ChainParams p) {
    extern __shared__ __align__(128) uint8_t ch_smem[];
    uint8_t *xs = ch_smem + p.off_xs;
    int2 *gtab = reinterpret_cast<int2 *>(ch_smem + p.off_gtab);
    float2 *part = reinterpret_cast<float2 *>(ch_smem + p.off_part);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int cta = blockIdx.x, grid = gridDim.x;
    const uint32_t bar_full = gv_smem_u32(ch_smem), bar_empty = bar_full + 8 * CH_MAX_STAGES;
    const uint32_t ring = gv_smem_u32(ch_smem + p.off_ring);
    constexpr uint32_t STAGE_W = 32 * GV_UNIT_BYTES, STAGE_M = 32 * MPU * META_BYTES;

    if (tid == 0) {
        for (int s = 0; s < p.S; ++s) { gv_mbar_init(bar_full + 8 * s, 1); gv_mbar_init(bar_empty + 8 * s, CH_CONS_WARPS); }
        *reinterpret_cast<volatile int *>(ch_smem + 2 * 8 * CH_MAX_STAGES) = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    if (warp >= CH_CONS_WARPS) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(CH_REGS_PRODUCER));
        volatile int *issued = reinterpret_cast<volatile int *>(ch_smem + 2 * 8 * CH_MAX_STAGES);   // stages the producer has issued
        if (warp == CH_CONS_WARPS + 1 && lane == 0 && p.l2pf > 0) {
            // ---- L2 prefetcher: the ring is all the shared memory there is (25 MB chip-wide), and while the short linears
            //      (qkv, o) sit in their hand-overs it is full and HBM idles.  This warp asks L2 for the `l2pf` stages beyond
            //      the one the producer issued last, paced by the producer's progress so that only a few MB are outstanding
            //      chip-wide (one-shot prefetches of tens of MB are dropped, profiles/r2_l2_prefetch_negative.md). ----
            int q = 0;
            for (int o = 0; o < p.n_ops; ++o) {
                const ChainOp *op = p.ops + o;
                const int R0 = __ldg(op->rb + cta), R1 = __ldg(op->rb + cta + 1);
                const int n_units = (R1 - R0) * op->UPR, nst = (n_units + 31) >> 5;
                const size_t unit0 = (size_t)R0 * op->UPR;
                const uint8_t *w = op->qw + unit0 * GV_UNIT_BYTES;
                const uint8_t *m = op->meta + unit0 * (MPU * META_BYTES);
                for (int s = 0; s < nst; ++s, ++q) {
                    const uint32_t cnt = (uint32_t)min(32, n_units - 32 * s);
                    uint32_t spins = 0;
                    while (q >= *issued + p.l2pf) {
                        __nanosleep(64);
                        if (++spins > GV_SPIN_LIMIT) __trap();
                    }
                    if (q >= *issued) {        // still ahead of the ring: otherwise the real copy is already on its way
                        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(w), "r"(cnt * GV_UNIT_BYTES) : "memory");
                        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(m), "r"(cnt * (MPU * META_BYTES)) : "memory");
                    }
                    w += STAGE_W; m += STAGE_M;
                }
            }
            return;
        }
        if (warp > CH_CONS_WARPS) return;      // the group's other warps only hand their registers back
        // ---- producer: the CTA's share of every linear's W4T stream, in chain order, as fast as slots come free.
        //      Weights are immutable, so this neither waits for the previous kernel nor for the grid barriers. ----
        if (lane == 0) {
            int slot = 0, count = 0;
            uint32_t parity = 1u;                    // first round: slots are free (wait on the phase before the first)
            for (int o = 0; o < p.n_ops; ++o) {
                const ChainOp *op = p.ops + o;
                const int *rb = op->rb;
                const int R0 = __ldg(rb + cta), R1 = __ldg(rb + cta + 1);
                const int n_units = (R1 - R0) * op->UPR, nst = (n_units + 31) >> 5;
                const size_t unit0 = (size_t)R0 * op->UPR;
                const uint8_t *w = op->qw + unit0 * GV_UNIT_BYTES;
                const uint8_t *m = op->meta + unit0 * (MPU * META_BYTES);
                for (int s = 0; s < nst; ++s) {
                    gv_mbar_wait(bar_empty + 8 * slot, parity);
                    const uint32_t bar = bar_full + 8 * slot, dst = ring + (uint32_t)slot * (uint32_t)p.slot_stride;
                    const uint32_t cnt = (uint32_t)min(32, n_units - 32 * s);          // the last stage of a linear may be short
                    const uint32_t bw = cnt * GV_UNIT_BYTES, bm = cnt * (MPU * META_BYTES);
                    gv_mbar_expect_tx(bar, bw + bm);
                    gv_bulk_g2s(dst, w, bw, bar);
                    gv_bulk_g2s(dst + STAGE_W, m, bm, bar);
                    w += STAGE_W; m += STAGE_M;
                    if (++slot == p.S) { slot = 0; parity ^= 1u; }
                    *issued = ++count;
                }
            }
        }
        return;
    }

    // ---- consumers ----
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(CH_REGS_CONSUMER));
    const int wi = warp;
    auto stamp = [&](int o, int i) {
        if (p.dbg && tid == 0) {
            unsigned long long tt;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tt));
            p.dbg[((size_t)o * grid + cta) * CH_DBG_WORDS + i] = tt;
        }
    };
    asm volatile("griddepcontrol.wait;" ::: "memory");   // the first x may come from the previous kernel; and the
    // previous run of this chain (which may have released this launch early) has stored `runs` only now
    const unsigned int run = p.n_ops > 1 ? __ldcg(p.runs) : 0u;
    uint32_t ep_base = 0u;
    if constexpr (TP) ep_base = __ldcg(p.peer.epoch_ptr);      // stored by the last CTA of the previous launch on this rank
    unsigned long long bar_count = (unsigned long long)run * (unsigned long long)p.n_barriers;   // barriers passed so far
    const long long mb_cur = (run & 1u) ? p.mbox_stride : 0, mb_nxt = (run & 1u) ? 0 : p.mbox_stride;
    int slot = 0;                                         // ring position of the CTA's first stage of the current linear
    uint32_t parity = 0u;
    for (int o = 0; o < p.n_ops; ++o) {
        // (fetching the next descriptor into shared memory while this linear combines was measured: 0.825 -> 0.864 ms per
        //  step -- the warp that fetches reaches the end-of-linear barrier late)
        const ChainOp op = p.ops[o];
        const int R0 = __ldg(op.rb + cta), n_slots = __ldg(op.rb + cta + 1) - R0;
        if (op.barrier_before) {
            // grid barrier: everything the linears before this one wrote is visible, everything they read has been read
            // (the CTA barrier below the combine ordered every thread's stores before thread 0's gpu-scope release)
            if (tid == 0) {
                ++bar_count;
                const unsigned long long target = bar_count * (unsigned long long)grid;
                asm volatile("red.release.gpu.global.add.u64 [%0], 1;" ::"l"(p.counter) : "memory");
                unsigned long long seen;
                uint32_t spins = 0;
                do {
                    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(seen) : "l"(p.counter) : "memory");
                    if (++spins > GV_SPIN_LIMIT) __trap();
                } while (seen < target);
            }
            stamp(o, 4);
            ch_bar_consumers();
        }
        stamp(o, 0);
        if (o + 1 == p.n_ops) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
        if constexpr (GEN) {
            if (op.ingest) {
                // cooperative act-order gather: this CTA's slice of x[:, perm] into the ingest mailbox of this run
                const unsigned short *src = op.ingest_src + (op.ingest_src_mbox ? mb_cur : 0);
                if (op.ingest_src_mbox) {      // the source is another linear's mailbox: wait for its arrival hint first
                    if (tid == 0) {
                        const unsigned long long target = ((unsigned long long)run + 1ull) * (unsigned long long)grid;
                        unsigned long long seen;
                        uint32_t spins = 0;
                        do {
                            asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(seen) : "l"(op.ingest_src_arrive) : "memory");
                            if (++spins > GV_SPIN_LIMIT) __trap();
                        } while (seen < target);
                    }
                    ch_bar_consumers();
                }
                uint4 *dst = reinterpret_cast<uint4 *>(op.ingest + mb_cur);
                const int ppr = op.K >> 3, total = op.M * ppr;
                for (int i = cta + tid * grid; i < total; i += grid * CH_CONS_THREADS) {
                    const int m = i / ppr, c = i - m * ppr;
                    const int4 p0 = __ldg(reinterpret_cast<const int4 *>(op.ingest_perm + 8 * c)), p1 = __ldg(reinterpret_cast<const int4 *>(op.ingest_perm + 8 * c) + 1);
                    const int idx[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
                    const unsigned short *row = src + (size_t)m * op.K;
                    unsigned short e[8];
                    uint32_t spins = 0;
#pragma unroll
                    for (int j = 0; j < 8; ++j) asm volatile("ld.relaxed.gpu.global.u16 %0, [%1];" : "=h"(e[j]) : "l"(row + idx[j]) : "memory");
                    for (;;) {
                        bool any = false;
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            if (op.ingest_src_mbox && e[j] == 0xFFFFu) {
                                any = true;
                                asm volatile("ld.relaxed.gpu.global.u16 %0, [%1];" : "=h"(e[j]) : "l"(row + idx[j]) : "memory");
                            }
                        if (!any) break;
                        if (++spins > GV_SPIN_LIMIT) __trap();
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j) if (e[j] == 0xFFFFu) e[j] = 0x7FFFu;       // never store the sentinel as data
                    uint4 v;
                    v.x = (uint32_t)e[0] | ((uint32_t)e[1] << 16); v.y = (uint32_t)e[2] | ((uint32_t)e[3] << 16);
                    v.z = (uint32_t)e[4] | ((uint32_t)e[5] << 16); v.w = (uint32_t)e[6] | ((uint32_t)e[7] << 16);
                    asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(dst + i), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
                }
                if (op.ingest_src_mbox) {      // what was read from the shared mailbox becomes the sentinel again in its OTHER parity
                    uint4 *other = reinterpret_cast<uint4 *>(const_cast<unsigned short *>(op.ingest_src) + mb_nxt);
                    for (int i = cta + tid * grid; i < total; i += grid * CH_CONS_THREADS)
                        other[i] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
                }
                ch_bar_consumers();
                if (tid == 0) asm volatile("red.relaxed.gpu.global.add.u64 [%0], 1;" ::"l"(op.ingest_arrive) : "memory");
            }
        }
        const bool poll = op.mbox_in != nullptr;
        if (poll) {
            if (tid == 0) {
                const unsigned long long target = ((unsigned long long)run + 1ull) * (unsigned long long)grid;
                unsigned long long seen;
                uint32_t spins = 0;
                do {
                    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(seen) : "l"(op.arrive_in) : "memory");
                    if (++spins > GV_SPIN_LIMIT) __trap();
                } while (seen < target);
            }
            stamp(o, 4);
            ch_bar_consumers();
        }
        const __nv_bfloat16 *xsrc = poll ? reinterpret_cast<const __nv_bfloat16 *>(op.mbox_in + mb_cur) : op.x;
        bool plain = true;
        if constexpr (PRO) {
          if (op.norm_w) {
            plain = false;
            uint4 *res = reinterpret_cast<uint4 *>(ch_smem + p.off_res);
            float *red = reinterpret_cast<float *>(ch_smem + p.off_red);
            ch_prologue_pass1<MR>(op, reinterpret_cast<const uint4 *>(xsrc), poll, res, red, tid, lane, warp, cta == 0);
            ch_stage_x<MR, MPU, 1>(reinterpret_cast<const __nv_bfloat16 *>(res), xs, gtab, op.M, op.K, tid, lane, false, nullptr,
                                reinterpret_cast<const uint4 *>(op.norm_w), red,
                                cta == 0 ? reinterpret_cast<uint4 *>(op.normed_out) : nullptr);
          }
        }
        if (plain) {
            ch_stage_x<MR, MPU>(xsrc, xs, gtab, op.M, op.K, tid, lane, poll,
                                p.dbg ? p.dbg + ((size_t)o * grid + cta) * CH_DBG_WORDS + 5 : nullptr);
        }
        if (poll) {
            // put the sentinel back into the OTHER parity of what was just read: its last readers finished a run ago, its
            // next writers run a run from now (they cannot start before every CTA has passed this point)
            uint4 *dst = reinterpret_cast<uint4 *>(const_cast<unsigned short *>(op.mbox_in) + mb_nxt);
            const int pieces = (op.M * op.K) >> 3;
            for (int i = cta + tid * grid; i < pieces; i += grid * CH_CONS_THREADS)
                dst[i] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
        }
        if (GEN && !(MPU == 1 && op.hoist) && op.spr == 0)      // the general path accumulates into `part`
            for (int i = tid; i < n_slots * CH_CONS_WARPS * 8 * MR; i += CH_CONS_THREADS) part[i] = make_float2(0.f, 0.f);
        ch_bar_consumers();
        stamp(o, 1);
        if (MPU == 1 && op.hoist)
            ch_row_blocks_hoist<MR>(p, n_slots, wi, lane, ring, bar_full, bar_empty, gv_smem_u32(xs), gv_smem_u32(gtab), part, slot, parity);
        else if (!GEN || op.spr > 0)
            ch_row_blocks<MR, MPU, false>(p, op, n_slots, wi, lane, ring, bar_full, bar_empty, gv_smem_u32(xs), gv_smem_u32(gtab), part, slot, parity);
        else if constexpr (GEN)
            ch_row_blocks_general<MR, MPU>(p, op, n_slots, wi, lane, ring, bar_full, bar_empty, gv_smem_u32(xs), gv_smem_u32(gtab), part, slot, parity);
        ch_bar_consumers();
        stamp(o, 2);
        // combine the 16 warps of every row block (fixed tree: deterministic) and write y
        for (int i = tid; i < n_slots * 32 * MR; i += CH_CONS_THREADS) {      // multiples of 32: whole warps
            const int q = i & 3, e = (i >> 2) & (8 * MR - 1), sl = i / (32 * MR);
            const float2 *src = part + ((size_t)sl * CH_CONS_WARPS + 4 * q) * 8 * MR + e;
            float2 v = src[0];
#pragma unroll
            for (int k = 1; k < 4; ++k) { const float2 o2 = src[k * 8 * MR]; v.x += o2.x; v.y += o2.y; }
            v.x += __shfl_xor_sync(0xffffffffu, v.x, 1); v.y += __shfl_xor_sync(0xffffffffu, v.y, 1);
            v.x += __shfl_xor_sync(0xffffffffu, v.x, 2); v.y += __shfl_xor_sync(0xffffffffu, v.y, 2);
            const int tok = e % MR, n0 = (R0 + sl) * 16 + e / MR;
            if constexpr (TP) {
                if (op.exchange) {          // this rank's partials into slot [epoch % 3][rank] of EVERY rank (lane q serves peers q, q + 4)
                    uint32_t bx = __float_as_uint(v.x), by = __float_as_uint(v.y);
                    if (bx == PEER_EMPTY) bx = 0u;
                    if (by == PEER_EMPTY) by = 0u;
                    const size_t cur_off = (size_t)((ep_base + (uint32_t)op.exchange) % 3u) * ((size_t)p.peer.world * p.peer.cap);
                    for (int r = q; r < p.peer.world && tok < op.M; r += 4) {
                        uint32_t *dst = reinterpret_cast<uint32_t *>(p.peer.base[r] + cur_off + (size_t)p.peer.rank * p.peer.cap);
                        asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(dst + (size_t)n0 * PEER_TOK + tok), "r"(bx) : "memory");
                        asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(dst + (size_t)(n0 + 8) * PEER_TOK + tok), "r"(by) : "memory");
                    }
                    continue;
                }
            }
            if (q == 0 && tok < op.M) {
                const float b0 = op.bias ? __bfloat162float(op.bias[n0]) : 0.f;
                const float b1 = op.bias ? __bfloat162float(op.bias[n0 + 8]) : 0.f;
                unsigned short *mb = op.mbox_out ? op.mbox_out + mb_cur : nullptr;
                if (op.swiglu) {   // rows g / g+8 are (gate, up) of output column 8 R + g
                    const int col = (n0 >> 4) * 8 + (n0 & 7);
                    const size_t row = (size_t)tok * (op.N >> 1);
                    const __nv_bfloat16 r = swiglu_bf16(v.x + b0, v.y + b1);
                    op.y[row + col] = r;
                    if (mb) ch_mbox_store(mb + row + ch_mbox_col(op.scatter, op.sc_off, op.sc_len, col), r);
                } else {
                    const size_t at = (size_t)tok * op.N + n0;
                    const __nv_bfloat16 r0 = __float2bfloat16_rn(v.x + b0), r1 = __float2bfloat16_rn(v.y + b1);
                    op.y[at] = r0;
                    op.y[at + 8] = r1;
                    if (mb) {
                        const size_t row = (size_t)tok * op.N;
                        ch_mbox_store(mb + row + ch_mbox_col(op.scatter, op.sc_off, op.sc_len, n0), r0);
                        ch_mbox_store(mb + row + ch_mbox_col(op.scatter, op.sc_off, op.sc_len, n0 + 8), r1);
                    }
                }
            }
        }
        if constexpr (TP) {
            if (op.exchange) {
                // poll this rank's slots until every source's word has arrived, sum in rank order (bit-identical on every
                // rank), + bias, bf16, into y and the mailbox; the consumed words become the sentinel again
                const size_t cur_off = (size_t)((ep_base + (uint32_t)op.exchange) % 3u) * ((size_t)p.peer.world * p.peer.cap);
                const float *own = p.peer.base[p.peer.rank & (PEER_MAX_WORLD - 1)];
                unsigned short *mb = op.mbox_out ? op.mbox_out + mb_cur : nullptr;
                unsigned long long t0;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
                for (int it = tid; it < n_slots * 16 * MR; it += CH_CONS_THREADS) {
                    const int tok = it % MR, n = R0 * 16 + it / MR;
                    if (tok >= op.M) continue;
                    const uint32_t *srcw = reinterpret_cast<const uint32_t *>(own + cur_off) + (size_t)n * PEER_TOK + tok;
                    uint32_t wv[PEER_MAX_WORLD];
                    uint32_t spins = 0;
                    for (;;) {
                        bool all = true;
#pragma unroll
                        for (int r = 0; r < PEER_MAX_WORLD; ++r) {
                            wv[r] = 0u;
                            if (r < p.peer.world)
                                asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(wv[r]) : "l"(srcw + (size_t)r * p.peer.cap) : "memory");
                        }
#pragma unroll
                        for (int r = 0; r < PEER_MAX_WORLD; ++r) all = all && wv[r] != PEER_EMPTY;
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
                            acc += __uint_as_float(wv[r]);
                            const_cast<uint32_t *>(srcw)[(size_t)r * p.peer.cap] = PEER_EMPTY;
                        }
                    }
                    const float bsv = op.bias ? __bfloat162float(op.bias[n]) : 0.f;
                    const __nv_bfloat16 r = __float2bfloat16_rn(acc + bsv);
                    op.y[(size_t)tok * op.N + n] = r;
                    if (mb) ch_mbox_store(mb + (size_t)tok * op.N + ch_mbox_col(op.scatter, op.sc_off, op.sc_len, n), r);
                }
            }
        }
        ch_bar_consumers();   // part / xs are free again; this CTA's y stores precede whatever thread 0 releases next
        if (op.arrive_out && tid == 0) asm volatile("red.relaxed.gpu.global.add.u64 [%0], 1;" ::"l"(op.arrive_out) : "memory");
        stamp(o, 3);
    }
    if (p.n_ops > 1 && cta == 0 && tid == 0) *p.runs = run + 1u;   // everybody read `runs` before the first barrier
    if constexpr (TP) {
        // the last CTA of this launch to finish advances the rank's exchange epoch (every CTA read it after its
        // griddepcontrol.wait; the next launch reads it after ITS wait)
        if (tid == 0 && p.n_exchange > 0) {
            __threadfence();
            const int prev = atomicAdd(p.peer.ticket, 1);
            if (prev == (int)gridDim.x - 1) {
                *p.peer.ticket = 0;
                *p.peer.epoch_ptr = ep_base + (uint32_t)p.n_exchange;
                __threadfence();
            }
        }
    }
}

// ---- host side ------------------------------------------------------------------------------------------------------
struct Chain {
    int device = 0, n = 0;
    int64_t M = 0;
    std::vector<Linear *> lin;
    std::vector<const __nv_bfloat16 *> x;
    std::vector<__nv_bfloat16 *> y;
    std::vector<w4a16_chain_prologue> pro;   // per linear; norm_weight == null: none
    bool has_prologue = false;
    bool general_k = false;        // some in_features is not a multiple of 4096
    std::vector<ChainOp> h_ops;    // host copy of the device table (w4a16_chain_set_peer edits it)
    std::vector<int> dep, barrier; // per linear: producer whose mailbox it reads (-1: none), grid barrier in front of it
    bool persistent = false;       // the chain kernel takes it; otherwise the linears are launched one by one
    std::string why_not;
    ChainOp *d_ops = nullptr;
    int *d_rb = nullptr;
    unsigned long long *d_counter = nullptr;
    unsigned int *d_runs = nullptr;
    unsigned short *d_mbox = nullptr;
    std::vector<int32_t *> d_scatter;      // inverse act-order permutations handed to the producers of act-order linears
    unsigned long long *d_arrive = nullptr;
    int n_mailboxes = 0, n_barriers = 0;
    unsigned long long *d_dbg = nullptr;
    ChainParams prm{};
    int MR = 1, mpu = 1, smem = 0, grid = 0;
};

template <int MR, int MPU, bool PRO, bool GEN, bool TP>
static int chain_launch_t(const Chain &c, cudaStream_t stream, bool pdl) {
    static bool attr_set[64] = {};
    if (c.device >= 0 && c.device < 64 && !attr_set[c.device]) {
        W4_CUDA_OK(cudaFuncSetAttribute(chain_kernel<MR, MPU, PRO, GEN, TP>, cudaFuncAttributeMaxDynamicSharedMemorySize, CH_SMEM_MAX));
        attr_set[c.device] = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)c.grid);
    cfg.blockDim = dim3(CH_THREADS);
    cfg.dynamicSmemBytes = (size_t)c.smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    W4_CUDA_OK(cudaLaunchKernelEx(&cfg, chain_kernel<MR, MPU, PRO, GEN, TP>, c.prm));
    count_launch();
    return 0;
}

static int chain_launch(const Chain &c, cudaStream_t stream) {
    const bool pdl = gv_env_int("W4A16_NO_PDL", 0) == 0;
#define W4_CH4(MR_, MPU_)                                                                          \
    do {                                                                                           \
        if (c.prm.n_exchange > 0) {   /* tensor-parallel chains: no prologues (rejected at set_peer) */ \
            if (c.general_k) return chain_launch_t<MR_, MPU_, false, true, true>(c, stream, pdl);  \
            return chain_launch_t<MR_, MPU_, false, false, true>(c, stream, pdl);                  \
        }                                                                                          \
        if (c.has_prologue) {                                                                      \
            if (c.general_k) return chain_launch_t<MR_, MPU_, true, true, false>(c, stream, pdl);  \
            return chain_launch_t<MR_, MPU_, true, false, false>(c, stream, pdl);                  \
        }                                                                                          \
        if (c.general_k) return chain_launch_t<MR_, MPU_, false, true, false>(c, stream, pdl);     \
        return chain_launch_t<MR_, MPU_, false, false, false>(c, stream, pdl);                     \
    } while (0)
#define W4_CH(MR_)                        \
    do {                                  \
        if (c.mpu == 1) W4_CH4(MR_, 1);   \
        if (c.mpu == 2) W4_CH4(MR_, 2);   \
        W4_CH4(MR_, 4);                   \
    } while (0)
    switch (c.MR) {
        case 1: W4_CH(1);
        case 2: W4_CH(2);
        default: W4_CH(4);
    }
#undef W4_CH
#undef W4_CH4
}

static void chain_free(Chain *c) {
    if (!c) return;
    cudaFree(c->d_ops); cudaFree(c->d_rb); cudaFree(c->d_counter); cudaFree(c->d_runs); cudaFree(c->d_mbox); cudaFree(c->d_arrive); cudaFree(c->d_dbg);
    for (int32_t *q : c->d_scatter) cudaFree(q);
    delete c;
}

// Decides whether the persistent kernel takes the chain and builds its device tables.
static int chain_plan(Chain *c) {
    c->persistent = false;
    if (gv_env_int("W4A16_NO_CHAIN_KERNEL", 0) && !c->has_prologue) { c->why_not = "disabled by W4A16_NO_CHAIN_KERNEL"; return 0; }
    if (c->M > 4) { c->why_not = "more than 4 tokens"; return 0; }
    if (c->n > CH_MAX_OPS) { c->why_not = "too many linears"; return 0; }
    DeviceScratch *ds = nullptr;
    if (get_device_scratch(c->device, &ds)) return 1;
    const int grid = ds->sm_count;
    int gs = 0;
    int64_t maxK = 0, max_slots = 1;
    bool has_perm = false;
    for (Linear *L : c->lin) {
        if (L->d.gs != 32 && L->d.gs != 64 && L->d.gs != 128) { c->why_not = "group size not 32 / 64 / 128"; return 0; }
        if (gs && L->d.gs != gs) { c->why_not = "mixed group sizes"; return 0; }
        if (L->perm && !L->perm_external) c->general_k = has_perm = true;    // the gather lives in the GEN instantiation
        if (L->d.NR * (L->d.K / GV_UNIT_K) >= ((int64_t)1 << 30)) { c->why_not = "linear too large"; return 0; }
        gs = L->d.gs;
        maxK = std::max(maxK, L->d.K);
        if (L->d.K % (32 * GV_UNIT_K)) c->general_k = true;
        max_slots = std::max<int64_t>(max_slots, (L->d.NR + grid - 1) / grid);
    }
    c->mpu = 128 / gs;
    c->MR = c->M <= 1 ? 1 : c->M <= 2 ? 2 : 4;
    int64_t off = 2 * 8 * CH_MAX_STAGES + 16;        // mbarriers + the producer's progress counter
    off = (off + 127) & ~(int64_t)127;
    const int off_xs = (int)off; off += maxK * 2 * c->MR;
    const int off_gtab = (int)off; off += (maxK / gs) * c->MR * 8;
    off = (off + 15) & ~(int64_t)15;
    const int off_part = (int)off; off += max_slots * CH_CONS_WARPS * 8 * c->MR * 8;
    off = (off + 127) & ~(int64_t)127;
    int off_res = -1, off_red = -1;
    if (c->has_prologue) {
        int64_t proK = 0;
        for (int i = 0; i < c->n; ++i)
            if (c->pro[i].norm_weight) proK = std::max(proK, c->lin[i]->d.K);
        off_res = (int)off; off += c->M * proK * 2;
        off_red = (int)off; off += (CH_CONS_WARPS + 1) * 4 * sizeof(float);
        off = (off + 127) & ~(int64_t)127;
    }
    const int64_t stride = 32 * (GV_UNIT_BYTES + c->mpu * META_BYTES);
    int64_t S = (CH_SMEM_MAX - off) / stride;
    if (S < 2) { c->why_not = "activations leave no room for a ring of 2 stages"; return 0; }
    S = std::min<int64_t>(S, CH_MAX_STAGES);
    const int off_ring = (int)off; off += S * stride;
    c->smem = (int)off;
    c->grid = grid;
    {   // every CTA must be resident at once (the hand-overs wait for all of them): one CTA per SM has to fit
        int per_sm = 0;
        cudaFuncSetAttribute(chain_kernel<1, 1, false, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, CH_SMEM_MAX);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, chain_kernel<1, 1, false, false, false>, CH_THREADS, (size_t)c->smem) != cudaSuccess ||
            per_sm < 1) {
            cudaGetLastError();
            c->why_not = "one CTA of the chain kernel does not fit an SM of this device";
            return 0;
        }
    }

    // ---- dependencies between the linears (byte ranges of the caller's buffers) ----
    // RAW: x[i] inside the y of its latest writer j -> mailbox (data-as-flag hand-over, no barrier).  Everything else that
    // needs an order -- a partial overlap, a write over something an earlier linear reads from the caller's memory, two
    // linears writing overlapping ranges with different row partitions -- gets a grid barrier in front of linear i.
    const int n = c->n;
    const int64_t M = c->M;
    auto width = [&](int i) { Linear *L = c->lin[i]; return L->epilogue == W4A16_EPILOGUE_SWIGLU ? L->d.N / 2 : L->d.N; };
    auto lo_x = [&](int i) { return reinterpret_cast<uintptr_t>(c->x[i]); };
    auto hi_x = [&](int i) { return lo_x(i) + (uintptr_t)(M * c->lin[i]->d.K * 2); };
    auto lo_y = [&](int i) { return reinterpret_cast<uintptr_t>(c->y[i]); };
    auto hi_y = [&](int i) { return lo_y(i) + (uintptr_t)(M * width(i) * 2); };
    auto overlap = [](uintptr_t a0, uintptr_t a1, uintptr_t b0, uintptr_t b1) { return a0 < b1 && b0 < a1; };
    std::vector<int> dep(n, -1), barrier(n, 0);
    const bool no_mbox = exp_env_int("W4A16_CHAIN_NO_MAILBOX", 0) != 0;
    for (int i = 1; i < n; ++i) {
        for (int j = i - 1; j >= 0; --j) {
            if (!overlap(lo_x(i), hi_x(i), lo_y(j), hi_y(j))) continue;
            const bool inside = lo_x(i) >= lo_y(j) && hi_x(i) <= hi_y(j) && ((lo_x(i) - lo_y(j)) & 15) == 0 &&
                                (M == 1 || (lo_x(i) == lo_y(j) && c->lin[i]->d.K == width(j)));
            if (inside && !no_mbox) dep[i] = j; else barrier[i] = 1;
            break;                                   // the latest writer decides
        }
        for (int j = 0; j < i && !barrier[i]; ++j) {
            if (overlap(lo_y(i), hi_y(i), lo_y(j), hi_y(j)) &&
                !(lo_y(i) == lo_y(j) && c->lin[i]->d.N == c->lin[j]->d.N && width(i) == width(j)))
                barrier[i] = 1;                      // same rows from the same CTA in program order would be fine
            if (overlap(lo_y(i), hi_y(i), lo_x(j), hi_x(j)) && dep[j] < 0) barrier[i] = 1;   // j reads the caller's memory
        }
        // prologue buffers: residual_in is read from the caller's memory, residual_out / normed_out are written by CTA 0
        if (c->has_prologue && !barrier[i]) {
            const uintptr_t hb = (uintptr_t)(M * c->lin[i]->d.K * 2);
            auto rng = [&](const void *q, uintptr_t &a, uintptr_t &b) { a = reinterpret_cast<uintptr_t>(q); b = a + (q ? hb : 0); };
            uintptr_t ri0, ri1, ro0, ro1, no0, no1;
            rng(c->pro[i].norm_weight ? c->pro[i].residual_in : nullptr, ri0, ri1);
            rng(c->pro[i].norm_weight ? c->pro[i].residual_out : nullptr, ro0, ro1);
            rng(c->pro[i].norm_weight ? c->pro[i].normed_out : nullptr, no0, no1);
            for (int j = 0; j < i && !barrier[i]; ++j) {
                const uintptr_t hj = (uintptr_t)(M * c->lin[j]->d.K * 2);
                const void *wj[2] = {c->pro[j].norm_weight ? c->pro[j].residual_out : nullptr, c->pro[j].norm_weight ? c->pro[j].normed_out : nullptr};
                for (const void *w : wj) {
                    if (!w) continue;
                    const uintptr_t w0 = reinterpret_cast<uintptr_t>(w), w1 = w0 + hj;
                    if (overlap(lo_x(i), hi_x(i), w0, w1) || (ri1 > ri0 && overlap(ri0, ri1, w0, w1))) barrier[i] = 1;   // reads what CTA 0 wrote
                }
                if (ri1 > ri0 && overlap(ri0, ri1, lo_y(j), hi_y(j))) barrier[i] = 1;
                if ((ro1 > ro0 && overlap(ro0, ro1, lo_x(j), hi_x(j)) && dep[j] < 0) ||
                    (no1 > no0 && overlap(no0, no1, lo_x(j), hi_x(j)) && dep[j] < 0))
                    barrier[i] = 1;
            }
        }
        if (barrier[i]) dep[i] = -1;                 // behind a barrier the caller's buffer itself is read
    }
    std::vector<int64_t> mbox_off(n, -1);
    int64_t mbox_total = 0;
    for (int i = 0; i < n; ++i)
        if (dep[i] >= 0 && mbox_off[dep[i]] < 0) {
            mbox_off[dep[i]] = mbox_total;
            mbox_total += (M * width(dep[i]) + 7) & ~(int64_t)7;
            ++c->n_mailboxes;
        }
    // act-order linears: the producer scatters when this linear is the only reader of its mailbox; otherwise the grid
    // gathers once into an ingest mailbox of the linear's own
    std::vector<int> readers(n, 0), use_scatter(n, 0);
    std::vector<int64_t> ingest_off(n, -1);
    for (int i = 0; i < n; ++i)
        if (dep[i] >= 0) ++readers[dep[i]];
    for (int i = 0; i < n; ++i) {
        Linear *L = c->lin[i];
        if (!(L->perm && !L->perm_external)) continue;
        if (dep[i] >= 0 && readers[dep[i]] == 1) { use_scatter[i] = 1; continue; }
        ingest_off[i] = mbox_total;
        mbox_total += (M * L->d.K + 7) & ~(int64_t)7;
    }
    if (mbox_total) {
        W4_CUDA_OK(cudaMalloc(&c->d_arrive, (size_t)2 * n * 16 * sizeof(unsigned long long)));     // one 128-byte line per linear (+ ingest)
        W4_CUDA_OK(cudaMemset(c->d_arrive, 0, (size_t)2 * n * 16 * sizeof(unsigned long long)));
        W4_CUDA_OK(cudaMalloc(&c->d_mbox, (size_t)mbox_total * 2 * sizeof(unsigned short)));
        W4_CUDA_OK(cudaMemset(c->d_mbox, 0xFF, (size_t)mbox_total * 2 * sizeof(unsigned short)));
    }

    std::vector<int> rb((size_t)c->n * (grid + 1));
    std::vector<ChainOp> ops(c->n);
    W4_CUDA_OK(cudaMalloc(&c->d_rb, rb.size() * sizeof(int)));
    for (int i = 0; i < c->n; ++i) {
        Linear *L = c->lin[i];
        for (int k = 0; k <= grid; ++k) rb[(size_t)i * (grid + 1) + k] = (int)((int64_t)k * L->d.NR / grid);
        ChainOp &o = ops[i];
        o.qw = reinterpret_cast<const uint8_t *>(L->qw);
        o.meta = L->meta;
        const bool add_bias = L->tp_mode != W4A16_TP_ROW || L->tp_size == 1;
        o.bias = add_bias && L->has_bias ? L->bias : nullptr;
        o.x = c->x[i]; o.y = c->y[i];
        o.rb = c->d_rb + (size_t)i * (grid + 1);
        o.M = (int)c->M; o.K = (int)L->d.K; o.N = (int)L->d.N;
        o.UPR = (int)(L->d.K / GV_UNIT_K);
        o.spr = o.UPR % 32 == 0 ? o.UPR / 32 : 0;
        o.swiglu = L->epilogue == W4A16_EPILOGUE_SWIGLU;
        o.hoist = o.spr == 1 && c->mpu == 1;
        o.barrier_before = barrier[i];
        c->n_barriers += barrier[i];
        o.mbox_out = mbox_off[i] >= 0 ? c->d_mbox + mbox_off[i] : nullptr;
        o.mbox_in = dep[i] >= 0 ? c->d_mbox + mbox_off[dep[i]] + (lo_x(i) - lo_y(dep[i])) / 2 : nullptr;
        o.arrive_out = mbox_off[i] >= 0 ? c->d_arrive + (size_t)i * 16 : nullptr;
        o.arrive_in = dep[i] >= 0 ? c->d_arrive + (size_t)dep[i] * 16 : nullptr;
        const w4a16_chain_prologue &pr = c->pro[i];
        o.norm_w = reinterpret_cast<const __nv_bfloat16 *>(pr.norm_weight);
        o.res_in = pr.norm_weight ? reinterpret_cast<const __nv_bfloat16 *>(pr.residual_in) : nullptr;
        o.res_out = pr.norm_weight ? reinterpret_cast<__nv_bfloat16 *>(pr.residual_out) : nullptr;
        o.normed_out = pr.norm_weight ? reinterpret_cast<__nv_bfloat16 *>(pr.normed_out) : nullptr;
        o.eps = pr.eps;
        o.add_running = pr.norm_weight && !pr.residual_in && pr.add_running_residual;
        o.exchange = 0;
        o.perm = L->perm && !L->perm_external ? L->perm : nullptr;
        o.scatter = nullptr; o.sc_off = 0; o.sc_len = 0;
        o.ingest_perm = nullptr; o.ingest_src = nullptr; o.ingest_src_mbox = 0; o.ingest_src_arrive = nullptr;
        o.ingest = nullptr; o.ingest_arrive = nullptr;
        if (ingest_off[i] >= 0) {
            o.ingest_perm = o.perm;
            o.perm = nullptr;
            o.ingest_src_mbox = o.mbox_in != nullptr;
            o.ingest_src = o.ingest_src_mbox ? o.mbox_in : reinterpret_cast<const unsigned short *>(o.x);
            o.ingest_src_arrive = o.arrive_in;
            o.ingest = c->d_mbox + ingest_off[i];
            o.ingest_arrive = c->d_arrive + (size_t)(n + i) * 16;
            o.mbox_in = o.ingest;                 // from here on the linear reads its ingest mailbox like any other
            o.arrive_in = o.ingest_arrive;
        }
    }
    // act-order readers: where a mailbox has exactly one reader and that reader gathers, its producer scatters instead
    {
        for (int i = 0; i < n; ++i) {
            if (!use_scatter[i] || !ops[i].perm) continue;
            const int64_t K = c->lin[i]->d.K;
            std::vector<int32_t> perm((size_t)K), inv((size_t)K);
            W4_CUDA_OK(cudaMemcpy(perm.data(), ops[i].perm, (size_t)K * sizeof(int32_t), cudaMemcpyDeviceToHost));
            for (int64_t t = 0; t < K; ++t) inv[(size_t)perm[(size_t)t]] = (int32_t)t;      // tiled row t reads column perm[t]
            int32_t *d = nullptr;
            W4_CUDA_OK(cudaMalloc(&d, (size_t)K * sizeof(int32_t)));
            c->d_scatter.push_back(d);
            W4_CUDA_OK(cudaMemcpy(d, inv.data(), (size_t)K * sizeof(int32_t), cudaMemcpyHostToDevice));
            ChainOp &prod = ops[dep[i]];
            prod.scatter = d;
            prod.sc_off = (int)((lo_x(i) - lo_y(dep[i])) / 2);      // M = 1: any aligned sub-range; M > 1: the whole row (offset 0)
            prod.sc_len = (int)K;
            ops[i].perm = nullptr;                                  // the mailbox row arrives gathered
        }
    }
    W4_CUDA_OK(cudaMemcpy(c->d_rb, rb.data(), rb.size() * sizeof(int), cudaMemcpyHostToDevice));
    W4_CUDA_OK(cudaMalloc(&c->d_ops, ops.size() * sizeof(ChainOp)));
    W4_CUDA_OK(cudaMemcpy(c->d_ops, ops.data(), ops.size() * sizeof(ChainOp), cudaMemcpyHostToDevice));
    c->h_ops = ops; c->dep = dep; c->barrier = barrier;
    W4_CUDA_OK(cudaMalloc(&c->d_counter, sizeof(unsigned long long)));
    W4_CUDA_OK(cudaMemset(c->d_counter, 0, sizeof(unsigned long long)));
    W4_CUDA_OK(cudaMalloc(&c->d_runs, sizeof(unsigned int)));
    W4_CUDA_OK(cudaMemset(c->d_runs, 0, sizeof(unsigned int)));
    if (gv_env_int("W4A16_GV_DEBUG", 0)) {
        const size_t words = (size_t)c->n * grid * CH_DBG_WORDS;
        W4_CUDA_OK(cudaMalloc(&c->d_dbg, words * sizeof(unsigned long long)));
        W4_CUDA_OK(cudaMemset(c->d_dbg, 0, words * sizeof(unsigned long long)));
    }
    ChainParams &p = c->prm;
    p.ops = c->d_ops; p.n_ops = c->n;
    p.counter = c->d_counter; p.runs = c->d_runs;
    p.n_barriers = c->n_barriers; p.mbox_stride = mbox_total;
    p.S = (int)S; p.slot_stride = (int)stride;
    p.l2pf = std::max(0, exp_env_int("W4A16_CHAIN_L2PF", 6));   // measured on the Qwen3-8B step: 0 / 2 / 4 / 6 / 8 / 12 / 16 stages -> 0.776 / 0.756 / 0.737 / 0.732 / 0.728 / 0.847 / 1.013 ms
    p.off_xs = off_xs; p.off_gtab = off_gtab; p.off_part = off_part; p.off_ring = off_ring;
    p.off_res = off_res; p.off_red = off_red;
    p.dbg = c->d_dbg;
    p.peer = PeerView{}; p.n_exchange = 0;
    c->persistent = true;
    return 0;
}

int forward_linear(Linear *L, const __nv_bfloat16 *x, __nv_bfloat16 *y, int64_t M, cudaStream_t stream);   // w4a16_api.cu
int peer_make_view(void *peer, PeerView *v);                                                             // w4a16_api.cu

}  // namespace w4

using namespace w4;

extern "C" {

int w4a16_chain_create(w4a16_chain_t *out, int n, const w4a16_handle_t *handles, const void *const *x, void *const *y, int64_t M) {
    return w4a16_chain_create_ex(out, n, handles, x, y, M, nullptr);
}

int w4a16_chain_create_ex(w4a16_chain_t *out, int n, const w4a16_handle_t *handles, const void *const *x, void *const *y, int64_t M,
                          const w4a16_chain_prologue *prologues) {
    if (!out) return fail("chain_create: null out");
    *out = nullptr;
    if (n < 1 || !handles || !x || !y) return fail("chain_create: need at least one linear and its x / y pointers");
    if (M < 1) return fail("chain_create: M must be >= 1");
    Chain *c = new Chain();
    c->n = n; c->M = M;
    for (int i = 0; i < n; ++i) {
        Linear *L = reinterpret_cast<Linear *>(handles[i]);
        if (!L) { chain_free(c); return fail("chain_create: null handle at position " + std::to_string(i)); }
        if (!L->loaded) { chain_free(c); return fail("chain_create: weights not loaded (call load_weights first) at position " + std::to_string(i)); }
        if (i == 0) c->device = L->device;
        if (L->device != c->device) { chain_free(c); return fail("chain_create: all linears must live on one device"); }
        if (!x[i] || !y[i]) { chain_free(c); return fail("chain_create: null x or y at position " + std::to_string(i)); }
        if ((reinterpret_cast<uintptr_t>(x[i]) & 15) || (reinterpret_cast<uintptr_t>(y[i]) & 15)) {
            chain_free(c);
            return fail("chain_create: x and y must be 16-byte aligned");
        }
        c->lin.push_back(L);
        c->x.push_back(reinterpret_cast<const __nv_bfloat16 *>(x[i]));
        c->y.push_back(reinterpret_cast<__nv_bfloat16 *>(y[i]));
        w4a16_chain_prologue pr{};
        if (prologues) pr = prologues[i];
        if (pr.norm_weight) {
            c->has_prologue = true;
            const void *ptrs[4] = {pr.norm_weight, pr.residual_in, pr.residual_out, pr.normed_out};
            for (const void *q : ptrs)
                if (reinterpret_cast<uintptr_t>(q) & 15) { chain_free(c); return fail("chain_create: prologue buffers must be 16-byte aligned"); }
            if (!(pr.eps >= 0.f)) { chain_free(c); return fail("chain_create: prologue eps must be >= 0"); }
            if (L->perm && !L->perm_external) { chain_free(c); return fail("chain_create: a prologue on an act-order linear is not supported"); }
        }
        c->pro.push_back(pr);
    }
    if (c->has_prologue) {
        // the running residual lives in shared memory between the prologues of one launch: it needs one width and a first
        // prologue that does not ask for it
        int64_t k_run = -1;
        bool have_running = false;
        for (int i = 0; i < n; ++i) {
            const w4a16_chain_prologue &pr = c->pro[i];
            if (!pr.norm_weight) continue;
            if (!pr.residual_in && pr.add_running_residual) {
                if (!have_running) { chain_free(c); return fail("chain_create: add_running_residual at position " + std::to_string(i) + " but no earlier prologue in this chain"); }
                if (c->lin[i]->d.K != k_run) { chain_free(c); return fail("chain_create: the running residual changes width at position " + std::to_string(i)); }
            }
            have_running = true;
            k_run = c->lin[i]->d.K;
        }
        if (M > 4) { chain_free(c); return fail("chain_create: prologues need M <= 4 (the persistent kernel)"); }
    }
    int prev = 0;
    cudaGetDevice(&prev);
    cudaSetDevice(c->device);
    const int rc = chain_plan(c);
    cudaSetDevice(prev);
    if (rc) { chain_free(c); return 1; }
    if (c->has_prologue && !c->persistent) {
        const std::string why = c->why_not;
        chain_free(c);
        return fail("chain_create: prologues need the persistent kernel, which does not take this chain (" + why + ")");
    }
    *out = reinterpret_cast<w4a16_chain_t>(c);
    return 0;
}

int w4a16_chain_set_peer(w4a16_chain_t chain, w4a16_peer_t peer) {
    Chain *c = reinterpret_cast<Chain *>(chain);
    if (!c) return fail("chain_set_peer: null chain");
    if (!c->persistent) return fail("chain_set_peer: the persistent kernel does not take this chain (" + c->why_not + "); use w4a16_forward_row_fused per linear");
    if (c->has_prologue) return fail("chain_set_peer: prologues and the fused all-reduce are not combined yet");
    PeerView v{};
    if (peer_make_view(peer, &v)) return 1;
    int n_ex = 0, prev_ex = -1;
    std::vector<ChainOp> ops = c->h_ops;
    std::vector<int> barrier = c->barrier;
    for (int i = 0; i < c->n; ++i) {
        Linear *L = c->lin[i];
        ops[i].exchange = 0;
        if (L->tp_mode != W4A16_TP_ROW || L->tp_size <= 1) continue;
        if (L->tp_size != v.world || L->tp_rank != v.rank) return fail("chain_set_peer: peer group does not match the shard at position " + std::to_string(i));
        if (L->epilogue != W4A16_EPILOGUE_NONE) return fail("chain_set_peer: a row-parallel linear cannot carry an epilogue");
        if ((uint64_t)L->d.N * PEER_TOK > v.cap) return fail("chain_set_peer: peer buffers too small for out_features " + std::to_string(L->d.N));
        ops[i].exchange = ++n_ex;
        ops[i].bias = L->has_bias ? L->bias : nullptr;          // added after the reduce (parallel_layers.rs:305-311)
        // Three rotating buffers are enough only if no CTA of this rank can start this all-reduce before every CTA has
        // finished the previous one: that holds when a chain of hand-overs (mailboxes or a barrier) links the two.
        if (prev_ex >= 0) {
            bool linked = false;
            for (int k = i; k > prev_ex;) {
                if (barrier[k]) { linked = true; break; }
                if (c->dep[k] < 0) break;
                k = c->dep[k];
                if (k == prev_ex) { linked = true; break; }
                if (k < prev_ex) break;
            }
            if (!linked) barrier[i] = 1;
        }
        prev_ex = i;
    }
    int n_bar = 0;
    for (int i = 0; i < c->n; ++i) { ops[i].barrier_before = barrier[i]; n_bar += barrier[i]; }
    int prev = 0;
    cudaGetDevice(&prev);
    if (prev != c->device) cudaSetDevice(c->device);
    const cudaError_t e = cudaMemcpy(c->d_ops, ops.data(), ops.size() * sizeof(ChainOp), cudaMemcpyHostToDevice);
    if (prev != c->device) cudaSetDevice(prev);
    if (e != cudaSuccess) return fail(std::string("chain_set_peer: ") + cudaGetErrorString(e));
    c->h_ops = ops; c->barrier = barrier;
    c->n_barriers = n_bar;
    c->prm.n_barriers = n_bar;
    c->prm.peer = v;
    c->prm.n_exchange = n_ex;
    return 0;
}

int w4a16_chain_run(w4a16_chain_t chain, w4a16_stream_t stream) {
    Chain *c = reinterpret_cast<Chain *>(chain);
    if (!c) return fail("chain_run: null chain");
    int prev = 0;
    cudaGetDevice(&prev);
    if (prev != c->device) cudaSetDevice(c->device);
    int rc = 0;
    if (c->persistent) {
        rc = chain_launch(*c, reinterpret_cast<cudaStream_t>(stream));
    } else {
        for (int i = 0; i < c->n && !rc; ++i)
            rc = forward_linear(c->lin[i], c->x[i], c->y[i], c->M, reinterpret_cast<cudaStream_t>(stream));
    }
    if (prev != c->device) cudaSetDevice(prev);
    return rc;
}

int w4a16_chain_run_host(w4a16_chain_t chain, const void *x0_host, void *y_last_host, w4a16_stream_t stream) {
    Chain *c = reinterpret_cast<Chain *>(chain);
    if (!c) return fail("chain_run_host: null chain");
    if (!x0_host || !y_last_host) return fail("chain_run_host: null host buffer");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    Linear *first = c->lin.front(), *last = c->lin.back();
    const size_t xb = (size_t)c->M * first->d.K * sizeof(__nv_bfloat16);
    const size_t yb = (size_t)c->M * (last->epilogue == W4A16_EPILOGUE_SWIGLU ? last->d.N / 2 : last->d.N) * sizeof(__nv_bfloat16);
    int prev = 0;
    cudaGetDevice(&prev);
    if (prev != c->device) cudaSetDevice(c->device);
    int rc = 0;
    if (cudaMemcpyAsync(const_cast<__nv_bfloat16 *>(c->x.front()), x0_host, xb, cudaMemcpyHostToDevice, st) != cudaSuccess)
        rc = fail("chain_run_host: host-to-device copy of x failed");
    if (!rc) rc = w4a16_chain_run(chain, stream);
    if (!rc && cudaMemcpyAsync(y_last_host, c->y.back(), yb, cudaMemcpyDeviceToHost, st) != cudaSuccess)
        rc = fail("chain_run_host: device-to-host copy of y failed");
    if (!rc && cudaStreamSynchronize(st) != cudaSuccess) rc = fail("chain_run_host: stream synchronise failed");
    if (prev != c->device) cudaSetDevice(prev);
    return rc;
}

const char *w4a16_chain_kernel(w4a16_chain_t chain) {
    Chain *c = reinterpret_cast<Chain *>(chain);
    if (!c) return "";
    static thread_local std::string s;
    if (c->persistent) {
        s = "chain_imma (" + std::to_string(c->n_mailboxes) + " mailboxes, " + std::to_string(c->n_barriers) + " barriers" +
            (c->prm.n_exchange ? ", " + std::to_string(c->prm.n_exchange) + " fused all-reduces" : std::string()) + ")";
        return s.c_str();
    }
    s = "per_linear (" + c->why_not + ")";
    return s.c_str();
}

int w4a16_chain_destroy(w4a16_chain_t chain) {
    chain_free(reinterpret_cast<Chain *>(chain));
    return 0;
}

int w4a16_chain_debug_stamps(w4a16_chain_t chain, uint64_t *out, int64_t words) {
    Chain *c = reinterpret_cast<Chain *>(chain);
    if (!c || !c->d_dbg) return fail("chain_debug_stamps: no stamps (set W4A16_GV_DEBUG=1 before chain_create)");
    const int64_t have = (int64_t)c->n * c->grid * CH_DBG_WORDS;
    if (words < have) return fail("chain_debug_stamps: buffer too small, need " + std::to_string(have) + " words");
    W4_CUDA_OK(cudaDeviceSynchronize());
    W4_CUDA_OK(cudaMemcpy(out, c->d_dbg, (size_t)have * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    return 0;
}

}  // extern "C"
