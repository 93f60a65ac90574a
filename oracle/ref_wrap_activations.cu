// Host launcher around the REFERENCE's own separate-tensor SiLU*mul kernel, compiled from /root/reference (never copied):
// crates/core/kernels/activations.cu :: silu_and_mul_separate_bf16 (:148-164; grid = tokens, one block per token) --
// the kernel QuantizedSwiGluMlp::forward reaches through silu_and_mul_separate_pooled (qwen3_quantized.rs:532-551,
// cuda_kernels.rs:2883).  Test infrastructure only (oracle/_ref).
#define REF_STR2(x) #x
#define REF_STR(x) REF_STR2(x)
#include REF_STR(REF_KERNEL_DIR/activations.cu)

extern "C" int ref_silu_and_mul_separate_bf16(void *out, const void *gate, const void *up, int tokens, int d, void *stream) {
    const int block = d < 1024 ? d : 1024;
    silu_and_mul_separate_bf16<<<tokens, block, 0, (cudaStream_t)stream>>>((__nv_bfloat16 *)out, (const __nv_bfloat16 *)gate,
                                                                          (const __nv_bfloat16 *)up, d);
    return (int)cudaGetLastError();
}

// add_bf16 (:99-113): the residual add between the linears (half_add_pooled, qwen3_quantized.rs:1062-1068)
extern "C" int ref_add_bf16(void *out, const void *a, const void *b, int tokens, int hidden, void *stream) {
    const int block = hidden < 1024 ? hidden : 1024;
    add_bf16<<<tokens, block, 0, (cudaStream_t)stream>>>((__nv_bfloat16 *)out, (const __nv_bfloat16 *)a, (const __nv_bfloat16 *)b, hidden);
    return (int)cudaGetLastError();
}
