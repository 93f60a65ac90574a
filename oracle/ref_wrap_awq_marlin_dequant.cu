// Host launcher around the REFERENCE's own kernel, compiled from where it lies under /root/reference
// (never copied into this repo): crates/core/kernels/awq_marlin_dequant.cu, launched with the geometry of
// marlin_cuda.rs:700-735 (32 x 8 threads, grid = (N/32, K/8)).  Test infrastructure only (oracle/_ref).
#define REF_STR2(x) #x
#define REF_STR(x) REF_STR2(x)
#include REF_STR(REF_KERNEL_DIR/awq_marlin_dequant.cu)

extern "C" int ref_awq_marlin_dequant_int4_bf16(void *out, const void *qweight_t, const void *scales_bf16,
                                                const void *qzeros_seq, int K, int N, int group_size, void *stream) {
    dim3 block(32, 8), grid((N + 31) / 32, (K + 7) / 8);
    awq_marlin_dequant_int4_bf16<<<grid, block, 0, (cudaStream_t)stream>>>(
        (__nv_bfloat16 *)out, (const uint32_t *)qweight_t, (const __nv_bfloat16 *)scales_bf16,
        (const uint32_t *)qzeros_seq, K, N, group_size, K / group_size);
    return (int)cudaGetLastError();
}
