/*
 * w4a16_oracle.c -- CPU restatement of vllm-rust's W4A16 quantized-linear path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library, and only as the checker or as the
 * timed CPU baseline.  The product path (vllm-rust_b200/) never links it.
 *
 * Parity status: PINNED.  Every function below restates a reference function
 * (cited file:line under /root/reference) and is checked in
 * tests/test_oracle_kat.py against the reference's own known-answer tests
 * (SURVEY.md section 8c, K-1 .. K-9).  On the GPU box the oracle is checked a
 * second time against the reference's own CUDA kernels, compiled unmodified
 * from /root/reference into oracle/_ref/ (tests/test_ref_kernels_gpu.py).
 * Two sub-paths stay unpinned and say so where they are defined:
 *   - oracle_awq_linear_forward (F1): the reference's fp32 GEMM is the `gemm`
 *     crate 0.18.2 (Cargo.lock:1309-1330), not vendored; restated as a plain
 *     sequential fp32 dot product.
 *   - g_idx / act-order (G4): the reference loads g_idx and never uses it.
 *
 * Plain C99, no dependencies.  Build: make -C oracle
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>

/* Minimal static-schedule parallel-for (libgomp is not in this image). */
typedef void (*par_body_fn)(long begin, long end, void *ctx);
typedef struct { par_body_fn fn; void *ctx; long begin, end; } par_task_t;
static void *par_trampoline(void *p) { par_task_t *t = (par_task_t *)p; t->fn(t->begin, t->end, t->ctx); return NULL; }
static void par_for(long n, int threads, par_body_fn fn, void *ctx) {
    if (threads <= 1 || n <= 1) { fn(0, n, ctx); return; }
    if (threads > 256) threads = 256;
    if ((long)threads > n) threads = (int)n;
    pthread_t tid[256]; par_task_t task[256];
    for (int t = 0; t < threads; ++t) {
        task[t].fn = fn; task[t].ctx = ctx;
        task[t].begin = n * t / threads; task[t].end = n * (t + 1) / threads;
        if (t > 0) pthread_create(&tid[t], NULL, par_trampoline, &task[t]);
    }
    par_trampoline(&task[0]);
    for (int t = 1; t < threads; ++t) pthread_join(tid[t], NULL);
}

/* ------------------------------------------------------------------------ */
/* 16-bit float helpers (round-to-nearest-even, like half::{f16,bf16}::from_f32
 * used by candle's to_dtype and like __float2bfloat16 on the device).       */
/* ------------------------------------------------------------------------ */

static inline uint32_t f32_bits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float bits_f32(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

uint16_t oracle_f32_to_bf16(float f) {
    uint32_t u = f32_bits(f);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x0040u); /* quiet NaN */
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}

float oracle_bf16_to_f32(uint16_t h) { return bits_f32((uint32_t)h << 16); }

uint16_t oracle_f32_to_f16(float f) {
    uint32_t x = f32_bits(f);
    uint32_t sign = (x >> 16) & 0x8000u;
    uint32_t absx = x & 0x7fffffffu;
    if (absx > 0x7f800000u) return (uint16_t)(sign | 0x7e00u);       /* NaN  */
    if (absx >= 0x47800000u) return (uint16_t)(sign | 0x7c00u);      /* >= 65536 -> inf (65520 handled by rounding below) */
    if (absx < 0x33000000u) return (uint16_t)sign;                   /* < 2^-25 -> 0 */
    int32_t exp = (int32_t)(absx >> 23) - 127;
    uint32_t man = (absx & 0x7fffffu) | 0x800000u;                   /* 24-bit significand */
    uint32_t shift;
    uint32_t base;
    if (exp < -14) {            /* subnormal half: value = man * 2^(exp-23), unit = 2^-24 */
        shift = (uint32_t)(-exp - 1);   /* 13 + (-14 - exp) */
        base = 0;
    } else {
        shift = 13;
        base = (uint32_t)(exp + 15) << 10;
        man &= 0x7fffffu;
    }
    uint32_t q = man >> shift;
    uint32_t rem = man & ((1u << shift) - 1u);
    uint32_t half = 1u << (shift - 1);
    uint32_t h = base + q;
    if (rem > half || (rem == half && (q & 1u))) h += 1;             /* RNE; carries into exponent correctly */
    return (uint16_t)(sign | h);
}

float oracle_f16_to_f32(uint16_t h) {
    uint32_t sign = ((uint32_t)h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1fu;
    uint32_t man = h & 0x3ffu;
    if (exp == 0) {
        if (man == 0) return bits_f32(sign);
        float v = (float)man * 5.9604644775390625e-08f; /* 2^-24 */
        return sign ? -v : v;
    }
    if (exp == 31) return bits_f32(sign | 0x7f800000u | (man << 13));
    return bits_f32(sign | ((exp + 112u) << 23) | (man << 13));
}

void oracle_f32_to_bf16_array(const float *in, uint16_t *out, size_t n) {
    for (size_t i = 0; i < n; ++i) out[i] = oracle_f32_to_bf16(in[i]);
}
void oracle_f16_to_bf16_array(const uint16_t *in, uint16_t *out, size_t n) {
    /* S1: scales F16 -> BF16 at load, awq_marlin.rs:649-656 (RNE cast). */
    for (size_t i = 0; i < n; ++i) out[i] = oracle_f32_to_bf16(oracle_f16_to_f32(in[i]));
}

/* ------------------------------------------------------------------------ */
/* U1: AWQ nibble interleave.                                                */
/* awq.rs:38  AWQ_UNDO_PACK = [0,4,1,5,2,6,3,7]                              */
/* awq_marlin.rs:372  AWQ_UNDO_PACK_SHIFTS = [0,16,4,20,8,24,12,28]          */
/* ------------------------------------------------------------------------ */
static const uint32_t AWQ_UNDO_PACK[8] = {0, 4, 1, 5, 2, 6, 3, 7};
static const uint32_t AWQ_UNDO_PACK_SHIFTS[8] = {0, 16, 4, 20, 8, 24, 12, 28};

/* Unpack AWQ-packed words [rows, cols/8] into one byte per nibble [rows, cols].
 * Used for both qweight [K, N/8] and qzeros [G, N/8] (awq.rs:142-161).      */
void oracle_awq_unpack(const uint32_t *packed, uint8_t *out, size_t rows, size_t cols) {
    size_t pc = cols / 8;
    for (size_t r = 0; r < rows; ++r)
        for (size_t p = 0; p < pc; ++p) {
            uint32_t w = packed[r * pc + p];
            for (int lane = 0; lane < 8; ++lane)
                out[r * cols + p * 8 + lane] = (uint8_t)((w >> (AWQ_UNDO_PACK[lane] * 4)) & 0xF);
        }
}

/* GPTQ qweight [K/8, N], nibble i of a word = k%8 (gptq_dequant.cu:6-14, 76-80). */
void oracle_gptq_unpack_qweight(const uint32_t *qweight, uint8_t *out, size_t K, size_t N) {
    for (size_t k = 0; k < K; ++k)
        for (size_t n = 0; n < N; ++n)
            out[k * N + n] = (uint8_t)((qweight[(k / 8) * N + n] >> ((k % 8) * 4)) & 0xF);
}

/* Sequential nibbles along the last axis: GPTQ qzeros [G, N/8]
 * (gptq_dequant.cu:68-73), also the AwqMarlin resident qzeros.             */
void oracle_seq_unpack(const uint32_t *packed, uint8_t *out, size_t rows, size_t cols) {
    size_t pc = cols / 8;
    for (size_t r = 0; r < rows; ++r)
        for (size_t c = 0; c < cols; ++c)
            out[r * cols + c] = (uint8_t)((packed[r * pc + c / 8] >> ((c % 8) * 4)) & 0xF);
}

/* ------------------------------------------------------------------------ */
/* D1: awq_dequantize_cpu, awq.rs:56-166.                                    */
/* weight[k,n] = (i32 diff) as f32 * f32(scale); whole matrix cast to F16.   */
/* scales are passed as f32 (to_dtype(F32) of an f16/bf16 tensor is exact).  */
/* group_size == 0 means one group (awq.rs:101-110).                         */
/* ------------------------------------------------------------------------ */
void oracle_awq_dequantize_cpu_f32(const uint32_t *qweight, const float *scales,
                                   const uint32_t *qzeros, size_t group_size, size_t K, size_t N,
                                   float *weight) {
    const size_t packed_out = N / 8;
    size_t num_groups = group_size == 0 ? 1 : (K + group_size - 1) / group_size;
    size_t egs = group_size == 0 ? (K > 0 ? K : 1) : group_size;
    for (size_t in_idx = 0; in_idx < K; ++in_idx) {
        size_t group = in_idx / egs;
        if (group > num_groups - 1) group = num_groups - 1;
        size_t qz_row = group * packed_out, sc_row = group * N;
        size_t qw_row = in_idx * packed_out, w_row = in_idx * N;
        for (size_t op = 0; op < packed_out; ++op) {
            uint32_t w_word = qweight[qw_row + op];
            uint32_t z_word = qzeros[qz_row + op];
            size_t out_base = op * 8;
            for (int lane = 0; lane < 8; ++lane) {
                uint32_t sh = AWQ_UNDO_PACK[lane] * 4;
                int32_t w_val = (int32_t)((w_word >> sh) & 0xF);
                int32_t z_val = (int32_t)((z_word >> sh) & 0xF);
                size_t out_idx = out_base + (size_t)lane;
                weight[w_row + out_idx] = (float)(w_val - z_val) * scales[sc_row + out_idx];
            }
        }
    }
}

void oracle_awq_dequantize_cpu(const uint32_t *qweight, const float *scales, const uint32_t *qzeros,
                               size_t group_size, size_t K, size_t N, float *scratch_f32,
                               uint16_t *weight_f16) {
    oracle_awq_dequantize_cpu_f32(qweight, scales, qzeros, group_size, K, N, scratch_f32);
    for (size_t i = 0; i < K * N; ++i) weight_f16[i] = oracle_f32_to_f16(scratch_f32[i]); /* awq.rs:163-165 */
}

/* ------------------------------------------------------------------------ */
/* F1: AwqLinear::forward, awq.rs:522-567.                                   */
/* x -> f32, W(f16) -> f32, x2d.matmul(W), + bias, cast back to x dtype.     */
/* The dequant is re-run on EVERY forward (awq.rs:538).  PARITY UNPINNED for */
/* the GEMM's accumulation order (gemm crate, not vendored): restated as a   */
/* sequential fp32 dot product.  x / y / bias are bf16 bit patterns.         */
/* threads <= 1: faithful single-thread.  threads > 1: OpenMP over rows of W */
/* in the dequant and over output columns in the GEMM ("all-cores" variant). */
/* ------------------------------------------------------------------------ */
typedef struct {
    const uint32_t *qweight, *qzeros; const float *scales; float *w;
    size_t group_size, K, N;
} f1_deq_ctx;
static void f1_deq_rows(long b, long e, void *p) {
    f1_deq_ctx *c = (f1_deq_ctx *)p;
    const size_t packed_out = c->N / 8;
    size_t num_groups = c->group_size == 0 ? 1 : (c->K + c->group_size - 1) / c->group_size;
    size_t egs = c->group_size == 0 ? (c->K > 0 ? c->K : 1) : c->group_size;
    for (long in_idx = b; in_idx < e; ++in_idx) {
        size_t group = (size_t)in_idx / egs;
        if (group > num_groups - 1) group = num_groups - 1;
        for (size_t op = 0; op < packed_out; ++op) {
            uint32_t w_word = c->qweight[(size_t)in_idx * packed_out + op];
            uint32_t z_word = c->qzeros[group * packed_out + op];
            for (int lane = 0; lane < 8; ++lane) {
                uint32_t sh = AWQ_UNDO_PACK[lane] * 4;
                int32_t d = (int32_t)((w_word >> sh) & 0xF) - (int32_t)((z_word >> sh) & 0xF);
                size_t o = op * 8 + (size_t)lane;
                float w = (float)d * c->scales[group * c->N + o];
                /* dequantize() returns F16 (awq.rs:163-165); forward casts it back to F32 (awq.rs:540) */
                c->w[(size_t)in_idx * c->N + o] = oracle_f16_to_f32(oracle_f32_to_f16(w));
            }
        }
    }
}
typedef struct {
    const float *xr, *w; const uint16_t *bias; uint16_t *y; size_t K, N;
} f1_gemm_ctx;
static void f1_gemm_panels(long b, long e, void *p) {
    f1_gemm_ctx *c = (f1_gemm_ctx *)p;
    for (long nb = b; nb < e; ++nb) {
        /* 64-column panels, k outer: the same per-column summation order as a
         * scalar k loop (sequential fp32 adds), but streams W row-wise.       */
        float acc[64];
        size_t n0 = (size_t)nb * 64, nn = c->N - n0 < 64 ? c->N - n0 : 64;
        for (size_t j = 0; j < nn; ++j) acc[j] = 0.0f;
        for (size_t k = 0; k < c->K; ++k) {
            float xv = c->xr[k];
            const float *wr = c->w + k * c->N + n0;
            for (size_t j = 0; j < nn; ++j) acc[j] += xv * wr[j];
        }
        for (size_t j = 0; j < nn; ++j) {
            float v = acc[j];
            if (c->bias) v += oracle_bf16_to_f32(c->bias[n0 + j]);
            c->y[n0 + j] = oracle_f32_to_bf16(v);
        }
    }
}
void oracle_awq_linear_forward(const uint16_t *x_bf16, size_t M, const uint32_t *qweight,
                               const float *scales, const uint32_t *qzeros,
                               const uint16_t *bias_bf16, size_t group_size, size_t K, size_t N,
                               float *scratch_w /* K*N */, float *scratch_x /* M*K */,
                               uint16_t *y_bf16, int threads) {
    f1_deq_ctx dc = {qweight, qzeros, scales, scratch_w, group_size, K, N};
    par_for((long)K, threads, f1_deq_rows, &dc);
    for (size_t i = 0; i < M * K; ++i) scratch_x[i] = oracle_bf16_to_f32(x_bf16[i]);
    for (size_t m = 0; m < M; ++m) {
        f1_gemm_ctx gc = {scratch_x + m * K, scratch_w, bias_bf16, y_bf16 + m * N, K, N};
        par_for((long)((N + 63) / 64), threads, f1_gemm_panels, &gc);
    }
}

/* ------------------------------------------------------------------------ */
/* R1: awq_to_gptq_u32, awq_marlin.rs:380-391; repack_awq_nibbles :126-151;  */
/*     CUDA twin awq_marlin_repack_int4, marlin_gemm.cu:431-459.             */
/* ------------------------------------------------------------------------ */
uint32_t oracle_awq_to_gptq_u32(uint32_t w) {
    uint32_t n0 = w & 0xF, n1 = (w >> 4) & 0xF, n2 = (w >> 8) & 0xF, n3 = (w >> 12) & 0xF;
    uint32_t n4 = (w >> 16) & 0xF, n5 = (w >> 20) & 0xF, n6 = (w >> 24) & 0xF, n7 = (w >> 28) & 0xF;
    return n0 | (n4 << 4) | (n1 << 8) | (n5 << 12) | (n2 << 16) | (n6 << 20) | (n3 << 24) | (n7 << 28);
}

void oracle_repack_awq_nibbles(const uint32_t *in, uint32_t *out, size_t nwords) {
    for (size_t i = 0; i < nwords; ++i) out[i] = oracle_awq_to_gptq_u32(in[i]);
}

/* R2: awq_to_gptq_qweight, awq_marlin.rs:170-240 (CPU loop :222-239);       */
/*     CUDA twin awq_to_gptq_qweight_int4, marlin_gemm.cu:491-517.           */
int oracle_awq_to_gptq_qweight(const uint32_t *awq, uint32_t *out, size_t K, size_t N) {
    if (K % 8 != 0 || N % 8 != 0) return -1;
    size_t packed_n = N / 8, packed_k = K / 8;
    for (size_t n = 0; n < N; ++n) {
        uint32_t undo_shift = AWQ_UNDO_PACK_SHIFTS[n % 8];
        size_t n_block = n / 8;
        for (size_t kb = 0; kb < packed_k; ++kb) {
            uint32_t word = 0;
            for (size_t i = 0; i < 8; ++i) {
                uint32_t aw = awq[(kb * 8 + i) * packed_n + n_block];
                word |= ((aw >> undo_shift) & 0xF) << (i * 4);
            }
            out[kb * N + n] = word;
        }
    }
    return 0;
}

/* R3: .t().contiguous() -> [N, K/8], awq_marlin.rs:635.                     */
void oracle_transpose_u32(const uint32_t *in, uint32_t *out, size_t rows, size_t cols) {
    for (size_t r = 0; r < rows; ++r)
        for (size_t c = 0; c < cols; ++c) out[c * rows + r] = in[r * cols + c];
}

/* R4: awq_to_marlin_tile_repack_cpu, awq_marlin.rs:276-366.                 */
int oracle_awq_to_marlin_tile_repack(const uint32_t *words, uint32_t *out, size_t size_k,
                                     size_t size_n) {
    static const size_t PACK_IDX[8] = {0, 2, 4, 6, 1, 3, 5, 7};
    static const size_t TC_OFFSETS[4] = {0, 1, 8, 9};
    if (size_k % 16 != 0 || size_n % 64 != 0) return -1;
    size_t packed_n = size_n / 8, k_tiles = size_k / 16, n_tiles = size_n / 64;
    size_t out_n_per_row = n_tiles * 128;
    for (size_t kt = 0; kt < k_tiles; ++kt) {
        size_t first_k = kt * 16;
        for (size_t nt = 0; nt < n_tiles; ++nt) {
            size_t first_n_packed = (nt * 64) / 8;
            size_t base = kt * out_n_per_row + nt * 128;
            for (size_t warp_id = 0; warp_id < 4; ++warp_id)
                for (size_t th_id = 0; th_id < 32; ++th_id) {
                    size_t tc_col = th_id / 4, tc_row = (th_id % 4) * 2;
                    size_t cur_n = warp_id * 16 + tc_col;
                    size_t cur_n_packed = cur_n / 8, cur_n_pos = cur_n % 8;
                    uint32_t undo_shift = AWQ_UNDO_PACK[cur_n_pos] * 4;
                    uint32_t vals[8];
                    for (size_t i = 0; i < 4; ++i) {
                        size_t cur_elem = tc_row + TC_OFFSETS[i];
                        size_t row_base = (first_k + cur_elem) * packed_n + first_n_packed;
                        vals[i] = (words[row_base + cur_n_packed] >> undo_shift) & 0xF;
                        vals[4 + i] = (words[row_base + cur_n_packed + 1] >> undo_shift) & 0xF;
                    }
                    uint32_t res = 0;
                    for (size_t j = 0; j < 8; ++j) res |= vals[PACK_IDX[j]] << (j * 4);
                    out[base + th_id * 4 + warp_id] = res;
                }
        }
    }
    return 0;
}

/* Independent position decoder used by the reference's round-trip test,
 * awq_marlin.rs:875-887 (marlin_tile_decode_position).                      */
void oracle_marlin_tile_decode_position(size_t k_in_tile, size_t n_in_tile, size_t *u32_idx,
                                        uint32_t *bit_offset) {
    static const uint32_t INV_PACK_IDX[8] = {0, 4, 1, 5, 2, 6, 3, 7};
    size_t m_n = n_in_tile % 16, warp_id = n_in_tile / 16;
    size_t tc_col = m_n < 8 ? m_n : m_n - 8, hi = m_n < 8 ? 0 : 4;
    size_t tc_row = ((k_in_tile / 2) % 4) * 2;
    size_t i_inner = (k_in_tile % 2) + (k_in_tile / 8) * 2;
    size_t th_id = tc_row / 2 + tc_col * 4;
    *u32_idx = th_id * 4 + warp_id;
    *bit_offset = INV_PACK_IDX[i_inner + hi] * 4;
}

/* ------------------------------------------------------------------------ */
/* D2: awq_marlin_dequant_int4_bf16, awq_marlin_dequant.cu:35-69 and         */
/*     gptq_dequant_int4_bf16, gptq_dequant.cu:48-87 (same arithmetic):      */
/*     out[k,n] = bf16_rn( (float)(q - z) * f32(bf16 scale) ).               */
/* Inputs here are already unpacked nibbles (one byte each) so the same      */
/* function serves the AWQ-Marlin resident layout and the GPTQ layout; the   */
/* unpack functions above carry the layout.  g_idx == NULL => group = k/gs   */
/* (what the reference does, even when desc_act is set: gptq.rs:349-366);    */
/* g_idx != NULL => canonical act-order semantics (UNPINNED).  */
/* zero_mode: 0 = use z as stored (G1), 1 = symmetric q-8, z ignored (G3,    */
/* marlin_gemm.cu:43-48, 67-187).                                            */
/* ------------------------------------------------------------------------ */
void oracle_dequant_bf16(const uint8_t *q, const uint8_t *z, const uint16_t *scales_bf16,
                         const int32_t *g_idx, size_t group_size, size_t K, size_t N,
                         int zero_mode, uint16_t *out_bf16) {
    size_t egs = group_size == 0 ? K : group_size;
    for (size_t k = 0; k < K; ++k) {
        size_t g = g_idx ? (size_t)g_idx[k] : k / egs;
        for (size_t n = 0; n < N; ++n) {
            int32_t zero = zero_mode == 1 ? 8 : (int32_t)z[g * N + n];
            float scale = oracle_bf16_to_f32(scales_bf16[g * N + n]);
            float wfp = (float)((int32_t)q[k * N + n] - zero) * scale;
            out_bf16[k * N + n] = oracle_f32_to_bf16(wfp);
        }
    }
}

/* ------------------------------------------------------------------------ */
/* GEMM oracle: the in-test scalar fp32 reference, marlin_cuda.rs:2091-2112. */
/* Sequential over k, fp32: acc += x[m,k] * ((q - z) as f32 * scale).        */
/* x_f32 and scales_f32 are the bf16-ROUNDED values (marlin_cuda.rs:2083-    */
/* 2089).  Same unpacked-nibble convention as oracle_dequant_bf16.           */
/* Also the arithmetic of gptq_gemm_int4_bf16_tiled (gptq_dequant.cu:243-    */
/* 318) and awq_gemv_int4_kt_bf16 (awq_gemv.cu:571-665) up to summation      */
/* order.  bias (f32, may be NULL) is added after the reduction.             */
/* ------------------------------------------------------------------------ */
typedef struct {
    const float *x, *scales, *bias; const uint8_t *q, *z; const int32_t *g_idx; float *y;
    size_t M, K, N, egs; int zero_mode;
} gemm_ctx;
static void gemm_cols(long b, long e, void *p) {
    gemm_ctx *c = (gemm_ctx *)p;
    for (long ni = b; ni < e; ++ni)
        for (size_t mi = 0; mi < c->M; ++mi) {
            float acc = 0.0f;
            for (size_t k = 0; k < c->K; ++k) {
                size_t g = c->g_idx ? (size_t)c->g_idx[k] : k / c->egs;
                int32_t zero = c->zero_mode == 1 ? 8 : (int32_t)c->z[g * c->N + (size_t)ni];
                float w_val = (float)((int32_t)c->q[k * c->N + (size_t)ni] - zero) * c->scales[g * c->N + (size_t)ni];
                acc += c->x[mi * c->K + k] * w_val;
            }
            if (c->bias) acc += c->bias[ni];
            c->y[mi * c->N + (size_t)ni] = acc;
        }
}
void oracle_gemm_f32(const float *x_f32, size_t M, const uint8_t *q, const uint8_t *z,
                     const float *scales_f32, const int32_t *g_idx, const float *bias,
                     size_t group_size, size_t K, size_t N, int zero_mode, float *y_f32,
                     int threads) {
    gemm_ctx c = {x_f32, scales_f32, bias, q, z, g_idx, y_f32, M, K, N,
                  group_size == 0 ? K : group_size, zero_mode};
    par_for((long)N, threads, gemm_cols, &c);
}

/* ------------------------------------------------------------------------ */
/* The reference tests' LCG (awq_marlin.rs:1019-1023, marlin_cuda.rs:2054-   */
/* 2089): x = x*1664525 + 1013904223; next(m) = state % m.                   */
/* Generation order of K-6: qweight words (8 nibbles each, sequential along  */
/* K, layout [K/8, N]), qzeros words ([G, N/8] sequential), scales           */
/* ((next(2000) - 1000) * 1e-4), inputs ((next(2000) - 1000) * 1e-3), both   */
/* bf16-rounded.                                                             */
/* ------------------------------------------------------------------------ */
static inline uint32_t lcg_next(uint32_t *s, uint32_t modulus) {
    *s = *s * 1664525u + 1013904223u;
    return *s % modulus;
}

void oracle_lcg_words(uint32_t seed, uint32_t *out, size_t n) {
    uint32_t s = seed;
    for (size_t i = 0; i < n; ++i) { s = s * 1664525u + 1013904223u; out[i] = s; }
}

void oracle_lcg_gemm_case(uint32_t seed, size_t M, size_t K, size_t N, size_t group_size,
                          uint32_t *qweight_gptq /* [K/8,N] */, uint32_t *qzeros_seq /* [G,N/8] */,
                          float *scales_bf16r /* [G,N] */, float *inputs_bf16r /* [M,K] */) {
    uint32_t s = seed;
    size_t G = K / group_size;
    for (size_t i = 0; i < (K / 8) * N; ++i) {
        uint32_t w = 0;
        for (int j = 0; j < 8; ++j) w |= (lcg_next(&s, 16) & 0xF) << (j * 4);
        qweight_gptq[i] = w;
    }
    for (size_t i = 0; i < G * (N / 8); ++i) {
        uint32_t w = 0;
        for (int j = 0; j < 8; ++j) w |= (lcg_next(&s, 16) & 0xF) << (j * 4);
        qzeros_seq[i] = w;
    }
    for (size_t i = 0; i < G * N; ++i) {
        float v = ((float)lcg_next(&s, 2000) - 1000.0f) * 1e-4f;
        scales_bf16r[i] = oracle_bf16_to_f32(oracle_f32_to_bf16(v));
    }
    for (size_t i = 0; i < M * K; ++i) {
        float v = ((float)lcg_next(&s, 2000) - 1000.0f) * 1e-3f;
        inputs_bf16r[i] = oracle_bf16_to_f32(oracle_f32_to_bf16(v));
    }
}

int oracle_max_threads(void) {
    long n = sysconf(_SC_NPROCESSORS_ONLN);
    return n > 0 ? (int)n : 1;
}
