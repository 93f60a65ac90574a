// Host launchers around the REFERENCE's own RMSNorm kernels, compiled from /root/reference (never copied):
// crates/core/kernels/layernorm.cu :: rms_norm_bf16 (:59-126) and fused_add_rms_norm_bf16 (:233-270); grid = tokens,
// block = min(hidden / 4, 1024) as the reference launches them (layers/normalization.rs).  Test infrastructure only.
#define REF_STR2(x) #x
#define REF_STR(x) REF_STR2(x)
#include REF_STR(REF_KERNEL_DIR/layernorm.cu)

extern "C" int ref_rms_norm_bf16(void *out, const void *in, const void *weight, float eps, int tokens, int hidden, void *stream) {
    int block = hidden / 4;
    if (block > 1024) block = 1024;
    if (block < 32) block = 32;
    rms_norm_bf16<<<tokens, block, 0, (cudaStream_t)stream>>>((__nv_bfloat16 *)out, (const __nv_bfloat16 *)in,
                                                             (const __nv_bfloat16 *)weight, eps, hidden);
    return (int)cudaGetLastError();
}
