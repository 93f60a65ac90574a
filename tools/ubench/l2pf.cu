// l2pf.cu -- does an L2 prefetch make a later streaming read faster on B200?
//   flush L2 (memset 1 GB) -> [prefetch X MB with method m] -> [delay] -> timed streaming read of the same X MB
// methods: 0 none (cold), 1 prefetch.global.L2 per 128 B by all threads, 2 cp.async.bulk.prefetch.L2 32 KB chunks by one
// thread per CTA, 3 a plain read pass (data certainly in L2 afterwards, as far as capacity allows), 4 like 2 with 4 KB chunks
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l2pf l2pf.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void stream_read(const uint4 *p, size_t n16, unsigned *sink) {
    // each CTA reads a contiguous slice, 8 loads in flight per thread
    const size_t per = (n16 + gridDim.x - 1) / gridDim.x;
    const size_t lo = blockIdx.x * per, hi = min(lo + per, n16);
    unsigned acc = 0;
    for (size_t i = lo + threadIdx.x; i < hi; i += 8 * blockDim.x) {
        uint4 v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const size_t j = i + (size_t)u * blockDim.x;
            v[u] = j < hi ? __ldcg(p + j) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
    if (acc == 0x12345678u) *sink = acc;
}
__global__ void pf_lsu(const char *p, size_t bytes) {
    const size_t per = ((bytes + gridDim.x - 1) / gridDim.x + 127) & ~(size_t)127;
    const size_t lo = blockIdx.x * per, hi = min(lo + per, bytes);
    for (size_t off = lo + (size_t)threadIdx.x * 128; off < hi; off += (size_t)blockDim.x * 128)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(p + off));
}
__global__ void pf_bulk(const char *p, size_t bytes, unsigned chunk) {
    const size_t per = ((bytes + gridDim.x - 1) / gridDim.x + 127) & ~(size_t)127;
    const size_t lo = blockIdx.x * per, hi = min(lo + per, bytes);
    if (threadIdx.x == 0)
        for (size_t off = lo; off < hi; off += chunk) {
            const unsigned n = (unsigned)min((size_t)chunk, hi - off);
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p + off), "r"(n) : "memory");
        }
}
// a REAL load of 4 bytes per `stride` bytes with an L2 prefetch-size hint: the line cannot be dropped like a prefetch
template <int STRIDE>
__global__ void pf_ld(const char *p, size_t bytes, unsigned *sink) {
    const size_t per = ((bytes + gridDim.x - 1) / gridDim.x + 255) & ~(size_t)255;
    const size_t lo = blockIdx.x * per, hi = min(lo + per, bytes);
    unsigned acc = 0;
    for (size_t off = lo + (size_t)threadIdx.x * STRIDE; off < hi; off += (size_t)blockDim.x * STRIDE) {
        unsigned v;
        if (STRIDE == 256) asm volatile("ld.global.L2::256B.b32 %0, [%1];" : "=r"(v) : "l"(p + off));
        else asm volatile("ld.global.L2::128B.b32 %0, [%1];" : "=r"(v) : "l"(p + off));
        acc ^= v;
    }
    if (acc == 0x12345678u) *sink = acc;
}
__global__ void spin(long long cycles) {
    const long long t0 = clock64();
    while (clock64() - t0 < cycles) {}
}

int main() {
    const size_t flush_bytes = (size_t)1 << 30;
    char *flush, *w;
    unsigned *sink;
    CK(cudaMalloc(&flush, flush_bytes));
    CK(cudaMalloc(&w, (size_t)160 << 20));
    CK(cudaMalloc(&sink, 4));
    CK(cudaMemset(w, 1, (size_t)160 << 20));
    cudaEvent_t e0, e1, p0, p1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1)); CK(cudaEventCreate(&p0)); CK(cudaEventCreate(&p1));
    const int mbs[] = {8, 16, 32, 48, 64, 96, 128};
    const char *names[] = {"cold", "prefetch.global.L2", "bulk.prefetch 32KB", "read pass", "bulk.prefetch 4KB", "ld.L2::256B 4B/256B", "ld.L2::128B 4B/128B"};
    for (int mb : mbs) {
        const size_t bytes = (size_t)mb << 20;
        for (int m = 0; m < 7; ++m) {
            float best = 1e9f, bestp = 0;
            for (int rep = 0; rep < 3; ++rep) {
                CK(cudaMemsetAsync(flush, rep, flush_bytes));
                CK(cudaEventRecord(p0));
                if (m == 1) pf_lsu<<<148, 512>>>(w, bytes);
                if (m == 2) pf_bulk<<<148, 64>>>(w, bytes, 32768u);
                if (m == 4) pf_bulk<<<148, 64>>>(w, bytes, 4096u);
                if (m == 5) pf_ld<256><<<148, 512>>>(w, bytes, sink);
                if (m == 6) pf_ld<128><<<148, 512>>>(w, bytes, sink);
                if (m == 3) stream_read<<<148, 512>>>((const uint4 *)w, bytes / 16, sink);
                CK(cudaEventRecord(p1));
                spin<<<1, 1>>>(60000);   // ~30 us: let the prefetches land
                CK(cudaEventRecord(e0));
                stream_read<<<148, 512>>>((const uint4 *)w, bytes / 16, sink);
                CK(cudaEventRecord(e1));
                CK(cudaDeviceSynchronize());
                float ms, msp;
                CK(cudaEventElapsedTime(&ms, e0, e1));
                CK(cudaEventElapsedTime(&msp, p0, p1));
                if (ms < best) { best = ms; bestp = msp; }
            }
            printf("%4d MB  %-22s read %8.2f us  %7.1f GB/s   (prefetch kernel %7.2f us)\n", mb, names[m], best * 1e3,
                   bytes / (best * 1e-3) / 1e9, bestp * 1e3);
        }
    }
    return 0;
}
