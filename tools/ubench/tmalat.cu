// TMA bulk-copy completion timing (sm_100a): each CTA issues S copies of B bytes at t0, then thread 0 polls the S
// mbarriers in order and records clock64 at each completion; a second pass over the (already complete) barriers by
// all threads measures the cost of a satisfied wait + LDS + __syncthreads iteration.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile("{\n\t.reg .pred p;\n\tW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@!p bra W;\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__global__ void __launch_bounds__(512) k(const uint8_t *src, int S, int B, long long *out, unsigned *sink) {
    extern __shared__ __align__(128) uint8_t sm[];
    const uint32_t bar0 = smem_u32(sm), ring = smem_u32(sm + 256);
    const int tid = threadIdx.x;
    const uint8_t *base = src + (size_t)blockIdx.x * S * B;
    if (tid == 0) {
        for (int s = 0; s < S; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar0 + 8 * s), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    long long *o = out + (size_t)blockIdx.x * 64;
    if (tid == 0) {
        const long long t0 = clock64();
        for (int s = 0; s < S; ++s) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar0 + 8 * s), "r"(B) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(ring + s * B), "l"(base + (size_t)s * B), "r"(B), "r"(bar0 + 8 * s) : "memory");
        }
        o[0] = clock64() - t0;
        for (int s = 0; s < S; ++s) { mbar_wait(bar0 + 8 * s, 0); o[1 + s] = clock64() - t0; }
    }
    __syncthreads();
    unsigned acc = 0;
    const long long t1 = clock64();
    for (int s = 0; s < S; ++s) {
        mbar_wait(bar0 + 8 * s, 0);
        acc ^= *reinterpret_cast<const unsigned *>(sm + 256 + s * B + ((tid * 16) % B));
        __syncthreads();
    }
    const long long t2 = clock64();
    if (tid == 0) o[40] = t2 - t1;
    if (acc == 0x12345678u) sink[0] = acc;
}
int main() {
    uint8_t *pool; long long *out; unsigned *sink;
    const size_t total = 512u << 20;
    cudaMalloc(&pool, total); cudaMalloc(&out, 148 * 64 * 8); cudaMalloc(&sink, 4);
    cudaMemset(pool, 1, total);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    size_t off = 0;
    for (int grid : {1, 148})
        for (int B : {1024, 4096, 16384, 32768})
            for (int S : {1, 6}) {
                for (int rep = 0; rep < 2; ++rep) {
                    off = (off + (size_t)148 * S * B + (64 << 20)) % (total - (size_t)148 * S * B);
                    off &= ~(size_t)127;
                    k<<<grid, 512, S * B + 256>>>(pool + off, S, B, out, sink);
                    cudaDeviceSynchronize();
                }
                long long h[64 * 148];
                cudaMemcpy(h, out, sizeof h, cudaMemcpyDeviceToHost);
                const int c = grid > 1 ? 77 : 0;
                printf("grid=%3d B=%5d S=%d  issue %lld  complete:", grid, B, S, h[c * 64]);
                for (int s = 0; s < S; ++s) printf(" %lld", h[c * 64 + 1 + s]);
                printf("  | satisfied-wait loop: %lld cycles for %d iterations\n", h[c * 64 + 40], S);
            }
    return 0;
}
