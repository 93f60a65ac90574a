// Host launcher around the REFERENCE's own fused SwiGLU kernel, compiled from /root/reference (never copied):
// crates/core/kernels/swiglu.cu :: fused_swiglu_bf16 (grid = tokens, block = min(hidden, 1024); launched by
// fused_swiglu, swiglu.rs, from QuantizedSwiGluMlp::forward, qwen3_quantized.rs:532-551).
// Test infrastructure only (oracle/_ref).
#define REF_STR2(x) #x
#define REF_STR(x) REF_STR2(x)
#include REF_STR(REF_KERNEL_DIR/swiglu.cu)

extern "C" int ref_fused_swiglu_bf16(void *out, const void *gate, const void *up, int tokens, int hidden, void *stream) {
    const int block = hidden < 1024 ? hidden : 1024;
    fused_swiglu_bf16<<<tokens, block, 0, (cudaStream_t)stream>>>((__nv_bfloat16 *)out, (const __nv_bfloat16 *)gate,
                                                                 (const __nv_bfloat16 *)up, hidden);
    return (int)cudaGetLastError();
}
