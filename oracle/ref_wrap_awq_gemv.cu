// Host launcher around the REFERENCE's own decode kernel, compiled from /root/reference (never copied):
// crates/core/kernels/awq_gemv.cu :: awq_gemv_int4_kt_bf16, launched like marlin_cuda.rs:571-601
// (grid = N/32, 256 threads, K*2 bytes of dynamic shared memory).  Test infrastructure only (oracle/_ref).
#define REF_STR2(x) #x
#define REF_STR(x) REF_STR2(x)
#include REF_STR(REF_KERNEL_DIR/awq_gemv.cu)

extern "C" int ref_awq_gemv_int4_kt_bf16(void *out, const void *x, const void *qweight_kt, const void *scales_bf16,
                                         const void *qzeros_seq, const void *bias, int M, int N, int K,
                                         int group_size, void *stream) {
    if (N % N_TILE || K % 8 || K * 2 > 47 * 1024) return -1;
    awq_gemv_int4_kt_bf16<<<N / N_TILE, SPLIT_KT_BLOCK_THREADS, K * 2, (cudaStream_t)stream>>>(
        (__nv_bfloat16 *)out, (const __nv_bfloat16 *)x, (const unsigned int *)qweight_kt,
        (const __nv_bfloat16 *)scales_bf16, (const unsigned int *)qzeros_seq, (const __nv_bfloat16 *)bias, M, N, K,
        group_size, 1, bias != nullptr);
    return (int)cudaGetLastError();
}
