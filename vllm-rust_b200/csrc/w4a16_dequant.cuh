// w4a16_dequant.cuh -- word-level int4 -> bf16 primitives shared by every kernel.
//
// The arithmetic to reproduce is the reference's dequant,
//     out = bf16_rn( (float)(q - z) * f32(bf16 scale) )      awq_marlin_dequant.cu:66-67
//                                                            gptq_dequant.cu:83-86
// Here it is done on packed bf16x2 pairs:
//     (w & 0x000F000F) | 0x43004300   ->  two bf16 values 128+q   (exact: 7-bit mantissa, 128+15 < 256)
//     sub.rn.bf16x2 with (128+z)      ->  q - z                   (exact: |q-z| <= 15)
//     mul.rn.bf16x2 with the scale    ->  one rounding of the exact product
// (q - z) has <= 4 significant bits and a bf16 scale 8, so f32(q-z)*f32(s) is exact in
// fp32 and its bf16_rn is the same single rounding.  tests/test_oracle_kat.py proves the
// identity exhaustively for all 31 differences x all finite bf16 scales; the GPU parity
// test checks w4a16_dequant_into bit-for-bit against the oracle.
#pragma once
#include <stdint.h>

namespace w4 {

// (a & b) | c in one LOP3 (truth table 0xEA).
__device__ __forceinline__ uint32_t lop3_and_or(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}

// bf16x2 {128+q_lo, 128+q_hi} from nibbles at bits 0..3 and 16..19.
__device__ __forceinline__ uint32_t q_magic(uint32_t w) { return lop3_and_or(w, 0x000F000Fu, 0x43004300u); }

// bf16x2 {128+z, 128+z}
__device__ __forceinline__ uint32_t zero_magic(uint32_t z4) { return 0x43004300u | z4 | (z4 << 16); }

__device__ __forceinline__ uint32_t bf16x2_sub(uint32_t a, uint32_t b) {
    uint32_t r;
    asm("sub.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}

__device__ __forceinline__ uint32_t bf16x2_mul(uint32_t a, uint32_t b) {
    uint32_t r;
    asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}

// bf16x2 {q_lo - z, q_hi - z}: exact integers, the decode kernel's MMA operand.
__device__ __forceinline__ uint32_t centered_pair(uint32_t w_shifted, uint32_t zmagic) {
    return bf16x2_sub(q_magic(w_shifted), zmagic);
}

// bf16x2 {(q_lo - z) * s, (q_hi - z) * s}: the dequantised weights (D2 grid).
__device__ __forceinline__ uint32_t dequant_pair(uint32_t w_shifted, uint32_t zmagic, uint32_t scale2) {
    return bf16x2_mul(centered_pair(w_shifted, zmagic), scale2);
}

}  // namespace w4
