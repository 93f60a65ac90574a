// w4a16_epilogue.cuh -- the SwiGLU epilogue shared by the forward kernels.
//
// A SwiGLU handle interleaves gate_proj and up_proj: tiled row 16 R + g is gate column 8 R + g, row 16 R + g + 8 is up
// column 8 R + g, so every kernel that owns a row-block pair (rows g / g+8 of one MMA fragment) owns a complete
// (gate, up) pair.  The arithmetic is the reference's fused_swiglu_bf16 (swiglu.cu:17-19, 40-45) as ITS build compiles
// it (nvcc -O3 --use_fast_math, build.rs:335-339): expf -> ex2.approx.ftz of x * -log2(e), the division -> rcp.approx.ftz
// and a multiply, everything flush-to-zero; spelled in PTX here because this library is not built with fast-math.
// Inputs are the two projections rounded to bf16 first -- the tensors the reference's separate linears hand to that
// kernel (qwen3_quantized.rs:532-551) -- so the result is bit-identical to linear, linear, fused_swiglu
// (tests/test_fused_gpu.py checks it against the reference kernel itself, oracle/_ref).
#pragma once
#include <cuda_bf16.h>

namespace w4 {

__device__ __forceinline__ float silu_mul_fastmath(float g, float u) {
    float t, e, d, r, s, o;
    asm("mul.ftz.f32 %0, %1, 0fBFB8AA3B;" : "=f"(t) : "f"(g));   // -g * log2(e)
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(t));
    asm("add.ftz.f32 %0, %1, 0f3F800000;" : "=f"(d) : "f"(e));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d));
    asm("mul.ftz.f32 %0, %1, %2;" : "=f"(s) : "f"(g), "f"(r));
    asm("mul.ftz.f32 %0, %1, %2;" : "=f"(o) : "f"(s), "f"(u));
    return o;
}

// gate / up: the fp32 accumulators with their bias already added
__device__ __forceinline__ __nv_bfloat16 swiglu_bf16(float gate, float up) {
    const float g = __bfloat162float(__float2bfloat16_rn(gate));
    const float u = __bfloat162float(__float2bfloat16_rn(up));
    return __float2bfloat16_rn(silu_mul_fastmath(g, u));
}

}  // namespace w4
