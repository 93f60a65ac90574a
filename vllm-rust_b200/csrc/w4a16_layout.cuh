// w4a16_layout.cuh -- the resident ("W4T") layout every kernel in this library reads.
//
// It replaces the reference's three resident forms (AwqMarlinLinear's [N,K/8]
// transposed qweight + sequential qzeros + bf16 scales, awq_marlin.rs:621-656, and
// the opt-in Marlin tile_b, awq_marlin.rs:276-366) with ONE tiled layout that both
// the decode kernel (mma.sync fragments straight from a 16-byte load) and the
// prefill kernel (tcgen05, dequant into swizzled smem) consume, and that is a
// pure linear stream for cp.async.bulk (TMA bulk copy): no tensor map needed.
//
//   qw    : u32 [N/16][K/64][32 lanes][4]            (a "chunk" = 16 n x 64 k = 512 B)
//   meta  : 48 B  [N/16][G]                          (scales + zeros of 16 rows, one group)
//
// Chunk (R, kb), word (lane = g*4 + t, j), nibble i (bits 4i..4i+3) holds q[k][n] with
//   n = 16R + g + 8*(i & 1)
//   k = 64kb + 16j + 4t + 2*((i >> 1) & 1) + (i >> 2)
// so that  ((word >> 4r) & 0x000F000F) | 0x43004300  is, for r = 0..3, exactly the
// bf16x2 A-fragment register a_r of mma.m16n8k16 (value 128+q) for the 16 k of word j
// (k-slot s = 4t+{0,1,2,3} of that MMA is k = 16j + s: thread t's B fragment is one
// contiguous 8-byte piece of the activation row), and at the same time the register
// quadruple {r0,r2,r1,r3} of a tcgen05.st.16x256b fragment (row g cols 2t,2t+1; row g+8
// cols 2t,2t+1), which is how the prefill kernel puts dequantised weights into TMEM
// as the A operand of tcgen05.mma with NO k permutation of the activations.
//
// Meta block (R, grp), 48 bytes:
//   [ 4g .. 4g+3 ]  bf16 scale of row 16R+g, bf16 scale of row 16R+g+8      (g = 0..7)
//   [ 32 + g     ]  zero nibbles: z[16R+g] | z[16R+g+8] << 4
//   [ 40 .. 47   ]  pad (keeps blocks 16-byte aligned for bulk copies)
// Symmetric GPTQ (q-8) stores z = 8, so kernels never branch on the format.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define W4_HD __host__ __device__ __forceinline__
#else
#define W4_HD inline
#endif

namespace w4 {

constexpr int CHUNK_N = 16;
constexpr int CHUNK_K = 64;
constexpr int CHUNK_WORDS = 128;   // u32 per chunk
constexpr int CHUNK_BYTES = 512;
constexpr int META_BYTES = 48;
constexpr int META_ZERO_OFF = 32;

struct Dims {
    int64_t K;       // local in_features  (multiple of 128)
    int64_t N;       // local out_features (multiple of 64)
    int gs;          // effective group size (K for per-channel)
    int64_t KB;      // K / 64
    int64_t NR;      // N / 16
    int64_t G;       // K / gs
};

W4_HD int64_t chunk_index(const Dims &d, int64_t R, int64_t kb) { return R * d.KB + kb; }

// (k, n) -> position of its nibble.
struct NibblePos { int64_t word; int shift; };
W4_HD NibblePos nibble_pos(const Dims &d, int64_t k, int64_t n) {
    int64_t R = n >> 4; int r = (int)(n & 15); int g = r & 7; int i0 = r >> 3;
    int64_t kb = k >> 6; int kk = (int)(k & 63);
    int j = kk >> 4; int t = (kk & 15) >> 2; int i1 = (kk & 3) >> 1; int i2 = kk & 1;
    int i = i0 | (i1 << 1) | (i2 << 2);
    int lane = g * 4 + t;
    NibblePos p;
    p.word = chunk_index(d, R, kb) * CHUNK_WORDS + lane * 4 + j;
    p.shift = 4 * i;
    return p;
}

// word (within chunk: lane, j) and nibble i -> (k offset within chunk, n offset within chunk)
W4_HD void nibble_coord(int lane, int j, int i, int &k_in_chunk, int &n_in_chunk) {
    int g = lane >> 2, t = lane & 3;
    n_in_chunk = g + 8 * (i & 1);
    k_in_chunk = 16 * j + 4 * t + 2 * ((i >> 1) & 1) + (i >> 2);
}

W4_HD int64_t meta_offset(const Dims &d, int64_t R, int64_t grp) { return (R * d.G + grp) * META_BYTES; }

W4_HD int64_t qw_bytes(const Dims &d) { return d.NR * d.KB * CHUNK_BYTES; }
W4_HD int64_t meta_bytes(const Dims &d) { return d.NR * d.G * META_BYTES; }

}  // namespace w4
