// w4a16_repack.cu -- load-time repack of AWQ / GPTQ checkpoints into the W4T layout,
// plus the bit-exact parity hooks (unpack, dequant, scales) that read it back.
//
// Replaces, in one pass each:
//   awq_to_gptq_qweight_int4 + .t().contiguous()   (marlin_gemm.cu:491-517, awq_marlin.rs:621-635)
//   awq_marlin_repack_int4 on qzeros               (marlin_gemm.cu:431-459, awq_marlin.rs:638-644)
//   scales F16 -> BF16                             (awq_marlin.rs:649-656)
//   awq_to_marlin_tile_repack_cpu                  (awq_marlin.rs:276-366; the opt-in tile_b copy)
// Integer work, bit-exact; tests/test_parity_gpu.py round-trips it through
// w4a16_unpack_into against the oracle's unpack of the checkpoint tensors.
#include "w4a16_internal.cuh"
#include "w4a16_dequant.cuh"

namespace w4 {

__device__ __constant__ int kAwqShift[8] = {0, 16, 4, 20, 8, 24, 12, 28};  // awq_marlin.rs:372

struct RepackParams {
    const uint32_t *qweight;
    const uint32_t *qzeros;
    const void *scales;
    int scales_dtype, format;
    int64_t K_full, N_full;
    int64_t k_off, n_off;
    int gs;       // local effective group size
    int gs_ckpt;  // checkpoint group size (K_full for per-channel)
    const int32_t *perm;       // local tiled row -> local checkpoint row (null = identity)
    const int32_t *grp_map;    // local group -> checkpoint group (null = (k_off + grp*gs)/gs_ckpt)
    Dims d;
    uint32_t *qw;
    uint8_t *meta;
    // column parts (fused sibling projections): this launch fills destination row blocks
    // [dst_R0, dst_R0 + NR_part) from a checkpoint of its own (n_off = 0, N_full = the part's width).
    // half = 0 / 1 (gate / up of a SwiGLU pair): the part's column c becomes row 16 (c / 8) + 8 half + c % 8, i.e. it fills
    // one half (rows g or rows g+8) of EVERY row block, read-modify-write on the words both halves share; -1 = plain.
    int64_t dst_R0, NR_part;
    int half;
};

__device__ __forceinline__ uint32_t src_nibble(const RepackParams &p, int64_t k_local, int64_t n_local) {
    int64_t row = p.k_off + (p.perm ? (int64_t)p.perm[k_local] : k_local);
    int64_t col = p.n_off + n_local;
    if (p.format == 0) {  // AWQ [K, N/8], interleaved along N (awq.rs:38)
        uint32_t w = p.qweight[row * (p.N_full >> 3) + (col >> 3)];
        return (w >> kAwqShift[col & 7]) & 0xFu;
    }
    uint32_t w = p.qweight[(row >> 3) * p.N_full + col];  // GPTQ [K/8, N], sequential along K
    return (w >> (4 * (int)(row & 7))) & 0xFu;
}

__global__ void repack_qweight_kernel(RepackParams p) {
    int64_t total = p.NR_part * p.d.KB * CHUNK_WORDS;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        int64_t chunk = idx >> 7;
        int wi = (int)(idx & 127);
        int lane = wi >> 2, j = wi & 3;
        int64_t R = chunk / p.d.KB, kb = chunk - R * p.d.KB;   // R: row block within the part
        uint32_t out = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            int kk, nn;
            nibble_coord(lane, j, i, kk, nn);
            if (p.half < 0) out |= src_nibble(p, kb * CHUNK_K + kk, R * CHUNK_N + nn) << (4 * i);
            else if ((i & 1) == p.half) out |= src_nibble(p, kb * CHUNK_K + kk, R * 8 + (nn & 7)) << (4 * i);
        }
        uint32_t *dst = p.qw + ((p.dst_R0 + R) * p.d.KB + kb) * CHUNK_WORDS + wi;
        if (p.half < 0) *dst = out;
        else *dst = (*dst & (p.half ? 0x0F0F0F0Fu : 0xF0F0F0F0u)) | out;
    }
}

__device__ __forceinline__ uint16_t load_scale_bf16(const void *s, int dtype, int64_t idx) {
    float f;
    if (dtype == 1) return reinterpret_cast<const uint16_t *>(s)[idx];
    if (dtype == 0) f = __half2float(reinterpret_cast<const __half *>(s)[idx]);
    else f = reinterpret_cast<const float *>(s)[idx];
    return __bfloat16_as_ushort(__float2bfloat16_rn(f));
}

__global__ void repack_meta_kernel(RepackParams p) {
    int64_t total = p.NR_part * p.d.G * 8;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        int g = (int)(idx & 7);
        int64_t blk = idx >> 3;
        int64_t R = blk / p.d.G, grp = blk - R * p.d.G;
        int64_t cg = p.grp_map ? (int64_t)p.grp_map[grp] : (p.k_off + grp * p.gs) / p.gs_ckpt;
        uint8_t *m = p.meta + ((p.dst_R0 + R) * p.d.G + grp) * META_BYTES;
        uint32_t zz[2];
        uint16_t ss[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (p.half >= 0 && h != p.half) {   // the other half of the pair keeps what it holds
                ss[h] = (uint16_t)(reinterpret_cast<uint32_t *>(m)[g] >> (16 * h));
                zz[h] = (m[META_ZERO_OFF + g] >> (4 * h)) & 0xFu;
                continue;
            }
            int64_t col = p.half < 0 ? p.n_off + R * CHUNK_N + g + 8 * h : R * 8 + g;
            ss[h] = load_scale_bf16(p.scales, p.scales_dtype, cg * p.N_full + col);
            if (p.format == 2) {
                zz[h] = 8u;  // symmetric: W = (q - 8) * s  (marlin_gemm.cu:43-48)
            } else {
                uint32_t w = p.qzeros[cg * (p.N_full >> 3) + (col >> 3)];
                int sh = p.format == 0 ? kAwqShift[col & 7] : 4 * (int)(col & 7);
                zz[h] = (w >> sh) & 0xFu;
            }
        }
        reinterpret_cast<uint32_t *>(m)[g] = (uint32_t)ss[0] | ((uint32_t)ss[1] << 16);
        m[META_ZERO_OFF + g] = (uint8_t)(zz[0] | (zz[1] << 4));
        if (g == 0) reinterpret_cast<uint64_t *>(m)[5] = 0ull;  // pad bytes 40..47
    }
}

int launch_repack(const Linear &L, const RepackSrc &src, cudaStream_t stream) {
    RepackParams p;
    p.qweight = src.qweight; p.qzeros = src.qzeros; p.scales = src.scales;
    p.scales_dtype = src.scales_dtype; p.format = src.format;
    p.K_full = src.K_full; p.N_full = src.N_full; p.k_off = src.k_off; p.n_off = src.n_off;
    p.gs = L.d.gs; p.gs_ckpt = src.gs;
    p.perm = src.perm; p.grp_map = src.row_group;
    p.d = L.d; p.qw = L.qw; p.meta = L.meta;
    p.dst_R0 = src.dst_R0; p.NR_part = src.NR_part > 0 ? src.NR_part : L.d.NR; p.half = src.half;
    int64_t words = p.NR_part * L.d.KB * CHUNK_WORDS;
    int blocks = (int)((words + 255) / 256 < 148 * 16 ? (words + 255) / 256 : 148 * 16);
    repack_qweight_kernel<<<blocks, 256, 0, stream>>>(p);
    count_launch();
    int64_t metas = p.NR_part * L.d.G * 8;
    int mblocks = (int)((metas + 255) / 256 < 148 * 8 ? (metas + 255) / 256 : 148 * 8);
    repack_meta_kernel<<<mblocks, 256, 0, stream>>>(p);
    count_launch();
    W4_CUDA_OK(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------
// parity hooks
// ---------------------------------------------------------------------------
struct HookParams {
    Dims d;
    const uint32_t *qw;
    const uint8_t *meta;
    const int32_t *perm;     // local tiled row -> local checkpoint row, or null
    const int32_t *grp_map;  // local group -> output group row, or null (identity)
    int pair;                // SwiGLU handle: the hooks answer in [gate | up] column order (col_of_row / row_of_col)
};

// SwiGLU pair layout: gate column c sits in tiled row 16 (c / 8) + c % 8, up column c eight rows further.
__device__ __forceinline__ int64_t row_of_col(const HookParams &p, int64_t n) {
    if (!p.pair) return n;
    const int64_t I = p.d.N >> 1, c = n >= I ? n - I : n;
    return 16 * (c >> 3) + (n >= I ? 8 : 0) + (c & 7);
}
__device__ __forceinline__ int64_t col_of_row(const HookParams &p, int64_t r) {
    if (!p.pair) return r;
    return ((r >> 4) << 3) + (r & 7) + ((r & 8) ? (p.d.N >> 1) : 0);
}

__global__ void unpack_q_kernel(HookParams p, uint8_t *q) {
    int64_t total = p.d.K * p.d.N;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        int64_t k = idx / p.d.N, n = idx - k * p.d.N;
        NibblePos np = nibble_pos(p.d, k, row_of_col(p, n));
        int64_t ko = p.perm ? (int64_t)p.perm[k] : k;
        q[ko * p.d.N + n] = (uint8_t)((p.qw[np.word] >> np.shift) & 0xFu);
    }
}

__global__ void unpack_z_kernel(HookParams p, uint8_t *z, __nv_bfloat16 *s) {
    int64_t total = p.d.G * p.d.N;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        int64_t grp = idx / p.d.N, n = idx - grp * p.d.N;
        const int64_t nt = row_of_col(p, n);
        int64_t R = nt >> 4;
        int r = (int)(nt & 15), g = r & 7, h = r >> 3;
        const uint8_t *m = p.meta + meta_offset(p.d, R, grp);
        int64_t go = p.grp_map ? (int64_t)p.grp_map[grp] : grp;
        if (z) z[go * p.d.N + n] = (uint8_t)((m[META_ZERO_OFF + g] >> (4 * h)) & 0xF);
        if (s) {
            uint32_t pr = reinterpret_cast<const uint32_t *>(m)[g];
            s[go * p.d.N + n] = __ushort_as_bfloat16((unsigned short)(h ? (pr >> 16) : (pr & 0xFFFFu)));
        }
    }
}

// Dense [K, N] bf16 through the SAME word-level primitive the prefill kernel uses
// (w4a16_dequant.cuh): (128+q) - (128+z) in bf16 is exact, times the bf16 scale with
// one rounding == bf16_rn(f32(q - z) * f32(scale)) of awq_marlin_dequant.cu:66-67.
__global__ void dequant_dense_kernel(HookParams p, __nv_bfloat16 *w) {
    int64_t total = p.d.NR * p.d.KB * CHUNK_WORDS;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        int64_t chunk = idx >> 7;
        int wi = (int)(idx & 127);
        int lane = wi >> 2, j = wi & 3, g = lane >> 2;
        int64_t R = chunk / p.d.KB, kb = chunk - R * p.d.KB;
        uint32_t word = p.qw[idx];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            int kk, nn;
            nibble_coord(lane, j, r, kk, nn);  // low half of pair r; the high half is k+1
            int64_t k = kb * CHUNK_K + kk;
            int64_t grp = k / p.d.gs;
            const uint8_t *m = p.meta + meta_offset(p.d, R, grp);
            int h = r & 1;
            uint32_t pr = reinterpret_cast<const uint32_t *>(m)[g];
            uint32_t s16 = h ? (pr >> 16) : (pr & 0xFFFFu);
            uint32_t z4 = (m[META_ZERO_OFF + g] >> (4 * h)) & 0xFu;
            uint32_t v = dequant_pair(word >> (4 * r), zero_magic(z4), s16 | (s16 << 16));
            int64_t n = col_of_row(p, R * CHUNK_N + nn);
            int64_t k0 = p.perm ? (int64_t)p.perm[k] : k;
            int64_t k1 = p.perm ? (int64_t)p.perm[k + 1] : k + 1;
            w[k0 * p.d.N + n] = __ushort_as_bfloat16((unsigned short)(v & 0xFFFFu));
            w[k1 * p.d.N + n] = __ushort_as_bfloat16((unsigned short)(v >> 16));
        }
    }
}

static int grid_for(int64_t n) {
    int64_t b = (n + 255) / 256;
    return (int)(b < 148 * 16 ? (b > 0 ? b : 1) : 148 * 16);
}

// grp_map for the hooks lives next to perm on the handle (Linear::perm + K entries).
static HookParams hook_params(const Linear &L) {
    HookParams p;
    p.d = L.d; p.qw = L.qw; p.meta = L.meta;
    p.perm = L.perm_external ? nullptr : L.perm;   // an externally permuted ROW shard answers in tiled row order
    p.grp_map = p.perm ? L.perm + L.d.K : nullptr;
    p.pair = L.epilogue == W4A16_EPILOGUE_SWIGLU;
    return p;
}

int launch_unpack(const Linear &L, uint8_t *q, uint8_t *z, cudaStream_t stream) {
    HookParams p = hook_params(L);
    if (q) { unpack_q_kernel<<<grid_for(L.d.K * L.d.N), 256, 0, stream>>>(p, q); count_launch(); }
    if (z) { unpack_z_kernel<<<grid_for(L.d.G * L.d.N), 256, 0, stream>>>(p, z, nullptr); count_launch(); }
    W4_CUDA_OK(cudaGetLastError());
    return 0;
}

int launch_scales_out(const Linear &L, __nv_bfloat16 *s, cudaStream_t stream) {
    HookParams p = hook_params(L);
    unpack_z_kernel<<<grid_for(L.d.G * L.d.N), 256, 0, stream>>>(p, nullptr, s);
    count_launch();
    W4_CUDA_OK(cudaGetLastError());
    return 0;
}

int launch_dequant(const Linear &L, __nv_bfloat16 *w, cudaStream_t stream) {
    HookParams p = hook_params(L);
    dequant_dense_kernel<<<grid_for(L.d.NR * L.d.KB * CHUNK_WORDS), 256, 0, stream>>>(p, w);
    count_launch();
    W4_CUDA_OK(cudaGetLastError());
    return 0;
}

// The reference's OTHER dequant grid ("D1", SURVEY 8a hazard i): awq_dequantize_cpu (awq.rs:56-166) multiplies
// (q - z) as f32 by the checkpoint's ORIGINAL scale (f16 in AWQ checkpoints, not the bf16-rounded resident copy) and
// casts the matrix to F16 (awq.rs:163-165).  The resident layout keeps only bf16 scales (S1), so the hook takes the
// checkpoint's scales tensor again; q and z come from the resident layout.  Parity hook only: no kernel consumes D1.
__global__ void dequant_d1_kernel(HookParams p, const void *scales, int dtype, int64_t N_full, int64_t n_off,
                                  int64_t k_off, int gs_ckpt, __half *w) {
    int64_t total = p.d.K * p.d.N;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        int64_t k = idx / p.d.N, n = idx - k * p.d.N;        // k = tiled row
        const int64_t nt = row_of_col(p, n);
        NibblePos np = nibble_pos(p.d, k, nt);
        int q = (int)((p.qw[np.word] >> np.shift) & 0xFu);
        int64_t grp = k / p.d.gs;
        int r = (int)(nt & 15);
        const uint8_t *m = p.meta + meta_offset(p.d, nt >> 4, grp);
        int z = (m[META_ZERO_OFF + (r & 7)] >> (4 * (r >> 3))) & 0xF;
        int64_t cg = p.grp_map ? (int64_t)p.grp_map[grp] : (k_off + grp * p.d.gs) / gs_ckpt;
        int64_t si = cg * N_full + n_off + n;
        float sc = dtype == 0 ? __half2float(reinterpret_cast<const __half *>(scales)[si])
                 : dtype == 1 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16 *>(scales)[si])
                              : reinterpret_cast<const float *>(scales)[si];
        int64_t ko = p.perm ? (int64_t)p.perm[k] : k;
        w[ko * p.d.N + n] = __float2half_rn((float)(q - z) * sc);
    }
}

int launch_dequant_d1(const Linear &L, const void *scales, int scales_dtype, __half *w, cudaStream_t stream) {
    HookParams p = hook_params(L);
    const int gs_ckpt = L.group_size_cfg > 0 ? L.group_size_cfg : (int)L.K_full;
    dequant_d1_kernel<<<grid_for(L.d.K * L.d.N), 256, 0, stream>>>(p, scales, scales_dtype, L.N_full, L.n_off, L.k_off,
                                                                  gs_ckpt, w);
    count_launch();
    W4_CUDA_OK(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------
// small helpers around the forward
// ---------------------------------------------------------------------------
// half < 0: dst[i] = src[i]; half = 0 / 1: source element i lands at row 16 (i / 8) + 8 half + i % 8 (SwiGLU pair layout)
__global__ void convert_bias_kernel(const void *src, int dtype, __nv_bfloat16 *dst, int64_t n, int half) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t o = half < 0 ? i : 16 * (i >> 3) + 8 * half + (i & 7);
    dst[o] = __ushort_as_bfloat16(load_scale_bf16(src, dtype, i));
}

int launch_convert_bias(const void *src, int dtype, __nv_bfloat16 *dst, int64_t n, cudaStream_t stream, int half) {
    convert_bias_kernel<<<(int)((n + 255) / 256), 256, 0, stream>>>(src, dtype, dst, n, half);
    count_launch();
    W4_CUDA_OK(cudaGetLastError());
    return 0;
}

// act-order: xg[m][i] = x[m][perm[i]]  (8 tiled rows = one 16-byte store per thread)
__global__ void gather_x_kernel(const __nv_bfloat16 *__restrict__ x, const int32_t *__restrict__ perm,
                                __nv_bfloat16 *__restrict__ xg, int64_t M, int64_t K) {
    int64_t per_row = K >> 3;
    int64_t total = M * per_row;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        int64_t m = idx / per_row, c = idx - m * per_row;
        const __nv_bfloat16 *xr = x + m * K;
        __align__(16) __nv_bfloat16 v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = xr[perm[c * 8 + e]];
        *reinterpret_cast<uint4 *>(xg + m * K + c * 8) = *reinterpret_cast<const uint4 *>(v);
    }
}

int launch_gather_x(const __nv_bfloat16 *x, const int32_t *perm, __nv_bfloat16 *xg, int64_t M, int64_t K,
                    cudaStream_t stream) {
    gather_x_kernel<<<grid_for(M * (K >> 3)), 256, 0, stream>>>(x, perm, xg, M, K);
    count_launch();
    W4_CUDA_OK(cudaGetLastError());
    return 0;
}

// y += bias (after the all-reduce of a row-parallel shard, parallel_layers.rs:301-304)
__global__ void add_bias_kernel(__nv_bfloat16 *y, const __nv_bfloat16 *bias, int64_t M, int64_t N) {
    int64_t total = M * N;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        int64_t n = idx % N;
        y[idx] = __float2bfloat16_rn(__bfloat162float(y[idx]) + __bfloat162float(bias[n]));
    }
}

int launch_add_bias(__nv_bfloat16 *y, const __nv_bfloat16 *bias, int64_t M, int64_t N, cudaStream_t stream) {
    add_bias_kernel<<<grid_for(M * N), 256, 0, stream>>>(y, bias, M, N);
    count_launch();
    W4_CUDA_OK(cudaGetLastError());
    return 0;
}

}  // namespace w4
