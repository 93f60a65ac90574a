// Host launchers around the REFERENCE's own GPTQ kernels, compiled from /root/reference (never copied):
// crates/core/kernels/gptq_dequant.cu :: gptq_dequant_int4_bf16 and gptq_gemm_int4_bf16_tiled
// (16 x 16 tiles, gptq_dequant.cu:239-318).  Test infrastructure only (oracle/_ref).
#define REF_STR2(x) #x
#define REF_STR(x) REF_STR2(x)
#include REF_STR(REF_KERNEL_DIR/gptq_dequant.cu)

extern "C" int ref_gptq_dequant_int4_bf16(void *out, const void *qweight, const void *scales_bf16, const void *qzeros,
                                          int K, int N, int group_size, void *stream) {
    dim3 block(16, 16), grid((N + 15) / 16, (K + 15) / 16);
    gptq_dequant_int4_bf16<<<grid, block, 0, (cudaStream_t)stream>>>(
        (__nv_bfloat16 *)out, (const uint32_t *)qweight, (const __nv_bfloat16 *)scales_bf16, (const uint32_t *)qzeros,
        K, N, group_size, K / group_size);
    return (int)cudaGetLastError();
}

extern "C" int ref_gptq_gemm_int4_bf16_tiled(void *out, const void *x, const void *qweight, const void *scales_bf16,
                                             const void *qzeros, const void *bias, int M, int N, int K,
                                             int group_size, void *stream) {
    dim3 block(BLOCK_N, BLOCK_M), grid((N + BLOCK_N - 1) / BLOCK_N, (M + BLOCK_M - 1) / BLOCK_M);
    gptq_gemm_int4_bf16_tiled<<<grid, block, 0, (cudaStream_t)stream>>>(
        (__nv_bfloat16 *)out, (const __nv_bfloat16 *)x, (const uint32_t *)qweight, (const __nv_bfloat16 *)scales_bf16,
        (const uint32_t *)qzeros, (const __nv_bfloat16 *)bias, M, N, K, group_size, K / group_size, bias != nullptr);
    return (int)cudaGetLastError();
}
