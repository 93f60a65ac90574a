// Streaming-read microbenchmark (sm_100a): how fast can 148 CTAs pull a weight-sized buffer out of HBM,
//   mode 0: cp.async.bulk (UBLKCP) global -> shared ring of S stages x B bytes, one elected thread per CTA
//   mode 1: same, but every warp issues its own slice of each stage (B / warps bytes per copy)
//   mode 2: ld.global.nc.v4 by all threads, U loads in flight per thread, results xor-reduced
// Buffers rotate over > 2x L2 so every launch reads HBM.  Prints GB/s per configuration.
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "W: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@!p bra W;\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// mode 0 / 1
__global__ void __launch_bounds__(512) bulk_kernel(const uint8_t *src, size_t bytes_per_cta, int S, int B, int split,
                                                    unsigned *sink) {
    extern __shared__ __align__(128) uint8_t sm[];
    uint64_t *bars = reinterpret_cast<uint64_t *>(sm);
    const uint32_t bar0 = smem_u32(bars), ring = smem_u32(sm + 128);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, nw = blockDim.x >> 5;
    const uint8_t *base = src + (size_t)blockIdx.x * bytes_per_cta;
    const int n = (int)(bytes_per_cta / B);
    if (tid == 0) {
        for (int s = 0; s < S; ++s) mbar_init(bar0 + 8 * s, split ? nw : 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto issue = [&](int j) {   // called by one thread (split = 0) or lane 0 of every warp (split = 1)
        const int s = j % S;
        if (!split) {
            mbar_expect_tx(bar0 + 8 * s, B);
            bulk_g2s(ring + s * B, base + (size_t)j * B, B, bar0 + 8 * s);
        } else {
            const int piece = B / nw;
            mbar_expect_tx(bar0 + 8 * s, piece);
            bulk_g2s(ring + s * B + warp * piece, base + (size_t)j * B + warp * piece, piece, bar0 + 8 * s);
        }
    };
    if (split ? lane == 0 : tid == 0)
        for (int j = 0; j < min(S, n); ++j) issue(j);
    unsigned acc = 0;
    for (int j = 0; j < n; ++j) {
        const int s = j % S;
        mbar_wait(bar0 + 8 * s, (j / S) & 1);
        acc ^= *reinterpret_cast<const unsigned *>(sm + 128 + s * B + ((tid * 16) % B));
        __syncthreads();   // everyone is done with slot s
        if (j + S < n && (split ? lane == 0 : tid == 0)) issue(j + S);
    }
    if (acc == 0x12345678u) sink[0] = acc;
}

// mode 2
template <int U>
__global__ void __launch_bounds__(512) ldg_kernel(const uint4 *src, size_t vec_per_cta, unsigned *sink) {
    const uint4 *base = src + (size_t)blockIdx.x * vec_per_cta;
    unsigned acc = 0;
    for (size_t i = threadIdx.x; i < vec_per_cta; i += (size_t)blockDim.x * U) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t k = i + (size_t)u * blockDim.x;
            if (k < vec_per_cta)
                asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                             : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(base + k));
            else v[u] = make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
    if (acc == 0x12345678u) sink[0] = acc;
}

int main(int argc, char **argv) {
    const size_t MB = 1 << 20;
    const size_t total = 416 * MB;   // rotating pool, > 3x L2
    uint8_t *pool; unsigned *sink;
    cudaMalloc(&pool, total); cudaMalloc(&sink, 4);
    cudaMemset(pool, 1, total);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaFuncSetAttribute(bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    const size_t sizes[] = {13 * MB, 52 * MB};
    for (size_t sz : sizes) {
        const int copies = (int)(total / sz);
        auto time_it = [&](auto launch, const char *name) {
            for (int i = 0; i < copies; ++i) launch(pool + (size_t)i * sz);
            cudaDeviceSynchronize();
            cudaEventRecord(e0);
            const int reps = 3;
            for (int r = 0; r < reps; ++r)
                for (int i = 0; i < copies; ++i) launch(pool + (size_t)i * sz);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            cudaError_t e = cudaGetLastError();
            const double us = ms * 1e3 / (reps * copies);
            printf("%6.0f MB  %-44s %7.2f us/launch  %7.1f GB/s %s\n", sz / 1048576.0, name, us, sz / us / 1e3,
                   e == cudaSuccess ? "" : cudaGetErrorString(e));
        };
        char name[128];
        for (int grid : {148, 296}) {
            for (int B : {4096, 8192, 16384, 32768}) {
                for (int S : {3, 6, 12}) {
                    if ((size_t)S * B + 128 > (grid == 148 ? 200u : 100u) * 1024) continue;
                    for (int split : {0, 1}) {
                        size_t per = sz / grid / B * B;
                        snprintf(name, sizeof name, "bulk grid=%d B=%d S=%d split=%d (%d KB in flight/SM)", grid, B, S, split,
                                 S * B * (grid / 148) / 1024);
                        time_it([&](uint8_t *p) { bulk_kernel<<<grid, 512, S * B + 128>>>(p, per, S, B, split, sink); }, name);
                    }
                }
            }
        }
        for (int grid : {148, 296, 592}) {
            size_t vec = sz / grid / 16;
            snprintf(name, sizeof name, "ldg.v4 grid=%d U=2", grid);
            time_it([&](uint8_t *p) { ldg_kernel<2><<<grid, 512>>>((const uint4 *)p, vec, sink); }, name);
            snprintf(name, sizeof name, "ldg.v4 grid=%d U=4", grid);
            time_it([&](uint8_t *p) { ldg_kernel<4><<<grid, 512>>>((const uint4 *)p, vec, sink); }, name);
            snprintf(name, sizeof name, "ldg.v4 grid=%d U=8", grid);
            time_it([&](uint8_t *p) { ldg_kernel<8><<<grid, 512>>>((const uint4 *)p, vec, sink); }, name);
        }
        printf("\n");
    }
    return 0;
}
