// PDL + bulk-copy prefetch microbenchmark (sm_100a).  A chain of N kernels in a CUDA graph, each streaming its own
// buffer (rotating, > L2) through a shared-memory ring with cp.async.bulk; the first S stages are issued BEFORE
// griddepcontrol.wait.  Prints the period per launch and, for the last launches, per-CTA time stamps
// {start, after wait, first stage landed, last stage landed, end}.  Knobs: bytes per launch, stage size, ring depth,
// threads per CTA, `spin_ns` of fake post-wait work before consuming (models activation staging).
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <algorithm>
#include <vector>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "W: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@!p bra W;\n\t}" ::"r"(bar), "r"(parity) : "memory");
}

__global__ void __launch_bounds__(512) stream_kernel(const uint8_t *src, size_t bytes_per_cta, int S, int B, int pre_wait, int wait_mode, int interleave,
                                                      unsigned spin_ns, unsigned *sink, unsigned long long *stamps) {
    extern __shared__ __align__(128) uint8_t sm[];
    const uint32_t bar0 = smem_u32(sm), ring = smem_u32(sm + 256);
    const int tid = threadIdx.x;
    const uint8_t *base = src + (size_t)blockIdx.x * bytes_per_cta;
    const int n = (int)(bytes_per_cta / B);
    auto issue = [&](int j) {
        const int s = j % S;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar0 + 8 * s), "r"(B) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(ring + s * B), "l"(interleave ? src + ((size_t)j * gridDim.x + blockIdx.x) * B : base + (size_t)j * B), "r"(B), "r"(bar0 + 8 * s) : "memory");
    };
    unsigned long long *st = stamps ? stamps + (size_t)blockIdx.x * 8 : nullptr;
    if (tid == 0) {
        for (int s = 0; s < S; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar0 + 8 * s), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        if (pre_wait) for (int j = 0; j < min(S, n); ++j) issue(j);
        if (st) st[0] = gtime();
    }
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (tid == 0) {
        if (st) st[1] = gtime();
        if (!pre_wait) for (int j = 0; j < min(S, n); ++j) issue(j);
    }
    if (spin_ns) { const unsigned long long t0 = gtime(); while (gtime() - t0 < spin_ns) {} }
    __syncthreads();
    unsigned acc = 0;
    for (int j = 0; j < n; ++j) {
        const int s = j % S;
        // wait_mode 0: every thread polls; 1: lane 0 of every warp polls, then __syncwarp; 2: thread 0 polls, then CTA barrier
        if (wait_mode == 0) mbar_wait(bar0 + 8 * s, (j / S) & 1);
        else if (wait_mode == 1) { if ((tid & 31) == 0) mbar_wait(bar0 + 8 * s, (j / S) & 1); __syncwarp(); }
        else { if (tid == 0) mbar_wait(bar0 + 8 * s, (j / S) & 1); __syncthreads(); }
        if (st && tid == 0 && j == 0) st[2] = gtime();
        if (st && tid == 0 && j == n - 1) st[3] = gtime();
        acc ^= *reinterpret_cast<const unsigned *>(sm + 256 + s * B + ((tid * 16) % B));
        __syncthreads();
        if (j + S < n && tid == 0) issue(j + S);
    }
    if (acc == 0x12345678u) sink[0] = acc;
    if (tid == 0) { sink[1 + blockIdx.x] = acc; if (st) st[4] = gtime(); }
}

// LSU variant: every warp owns a private ring of D slots x U bytes filled with cp.async (16 B per lane per
// instruction); warp w takes units w, w+16, ... of the CTA's contiguous range.  No mbarriers, no CTA barriers.
template <int D>
__global__ void __launch_bounds__(512) lsu_kernel(const uint8_t *src, size_t bytes_per_cta, int U, int pre_wait, unsigned spin_ns, unsigned *sink,
                                                   unsigned long long *stamps) {
    extern __shared__ __align__(128) uint8_t sm[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint8_t *base = src + (size_t)blockIdx.x * bytes_per_cta;
    const int n = (int)(bytes_per_cta / ((size_t)U * 16));    // units per warp
    const uint32_t ring = smem_u32(sm) + warp * D * U;
    auto issue = [&](int j) {
        if (j < n) {
            const uint8_t *g = base + ((size_t)j * 16 + warp) * U + lane * 16;
            const uint32_t d = ring + (j % D) * U + lane * 16;
            for (int o = 0; o < U; o += 512)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d + o), "l"(g + o) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    unsigned long long *st = stamps ? stamps + (size_t)blockIdx.x * 8 : nullptr;
    if (pre_wait) for (int j = 0; j < D; ++j) issue(j);
    if (st && tid == 0) st[0] = gtime();
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (st && tid == 0) st[1] = gtime();
    if (!pre_wait) for (int j = 0; j < D; ++j) issue(j);
    if (spin_ns) { const unsigned long long t0 = gtime(); while (gtime() - t0 < spin_ns) {} __syncthreads(); }
    unsigned acc = 0;
    for (int j = 0; j < n; ++j) {
        asm volatile("cp.async.wait_group %0;" ::"n"(D - 1) : "memory");
        __syncwarp();
        if (st && tid == 0 && j == 0) st[2] = gtime();
        if (st && tid == 0 && j == n - 1) st[3] = gtime();
        const uint4 v = *reinterpret_cast<const uint4 *>(sm + warp * D * U + (j % D) * U + lane * 16);
        acc ^= v.x ^ v.y ^ v.z ^ v.w;
        __syncwarp();
        issue(j + D);
    }
    if (acc == 0x12345678u) sink[0] = acc;
    __syncthreads();
    if (tid == 0) { sink[1 + blockIdx.x] = acc; if (st) st[4] = gtime(); }
}

int main(int argc, char **argv) {
    const size_t MB = 1 << 20;
    const size_t total = 416 * MB;
    uint8_t *pool; unsigned *sink; unsigned long long *stamps;
    cudaMalloc(&pool, total); cudaMalloc(&sink, 4096); cudaMalloc(&stamps, 4 * 1024 * 8 * 8);
    cudaMemset(pool, 1, total);
    cudaStream_t st; cudaStreamCreate(&st);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    struct Cfg { size_t sz; int B, S, pre, pdl; unsigned spin; int threads; int wm = 0; int il = 0; int lsuD = 0; int U = 1024; };
    std::vector<Cfg> cfgs;
    for (int il : {0, 1}) {
        cfgs.push_back({13 * MB, 8192, 11, 1, 1, 0, 512, 0, il});
        cfgs.push_back({13 * MB + 148 * 8192, 16384, 6, 1, 1, 0, 512, 0, il});
        cfgs.push_back({14 * MB, 32768, 3, 1, 1, 0, 512, 0, il});
        cfgs.push_back({52 * MB, 8192, 12, 1, 1, 0, 512, 0, il});
        cfgs.push_back({52 * MB, 16384, 6, 1, 1, 0, 512, 0, il});
        cfgs.push_back({52 * MB, 32768, 3, 1, 1, 0, 512, 0, il});
        cfgs.push_back({52 * MB, 16384, 12, 1, 1, 0, 512, 0, il});
        cfgs.push_back({52 * MB, 32768, 6, 1, 1, 0, 512, 0, il});
    }
    cfgs.clear();
    for (size_t sz : {13 * MB + 148 * 8192, 52 * MB})
        for (unsigned spin : {0u, 1000u, 2000u, 3000u})
            for (int pre : {0, 1}) { Cfg c{sz, 16384, 6, pre, 1, spin, 512}; c.lsuD = 6; c.U = 1024; cfgs.push_back(c); }
    cudaFuncSetAttribute(lsu_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(lsu_kernel<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    for (const Cfg &c : cfgs) {
        const int copies = (int)(total / c.sz);
        const int N = std::max(copies, 16);
        const size_t per = c.sz / 148 / c.B * c.B;
        cudaMemsetAsync(stamps, 0, 4 * 1024 * 8 * 8, st);
        cudaGraph_t g; cudaGraphExec_t ge;
        cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
        for (int i = 0; i < N; ++i) {
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3(148); cfg.blockDim = dim3(c.threads); cfg.stream = st;
            cfg.dynamicSmemBytes = (size_t)c.S * c.B + 256;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr; cfg.numAttrs = c.pdl ? 1 : 0;
            unsigned long long *sp = i >= N - 3 ? stamps + (size_t)(i - (N - 3)) * 1024 * 8 : nullptr;
            if (c.lsuD) {
                cfg.dynamicSmemBytes = (size_t)16 * c.lsuD * c.U;
                const uint8_t *sp0 = pool + (size_t)(i % copies) * c.sz;
                if (c.lsuD == 3) cudaLaunchKernelEx(&cfg, lsu_kernel<3>, sp0, per, c.U, c.pre, c.spin, sink, sp);
                else cudaLaunchKernelEx(&cfg, lsu_kernel<6>, sp0, per, c.U, c.pre, c.spin, sink, sp);
            } else
            cudaLaunchKernelEx(&cfg, stream_kernel, (const uint8_t *)(pool + (size_t)(i % copies) * c.sz), per, c.S, c.B, c.pre,
                               c.wm, c.il, c.spin, sink, sp);
        }
        cudaStreamEndCapture(st, &g);
        cudaGraphInstantiate(&ge, g, 0);
        cudaGraphLaunch(ge, st); cudaStreamSynchronize(st);
        cudaEventRecord(e0, st);
        const int reps = 5;
        for (int r = 0; r < reps; ++r) cudaGraphLaunch(ge, st);
        cudaEventRecord(e1, st); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        cudaError_t e = cudaGetLastError();
        const double us = ms * 1e3 / (reps * N);
        printf("%s %5.1f MB/launch B=%5d S=%2d thr=%3d pre_wait=%d pdl=%d spin=%u : %7.2f us/launch %7.1f GB/s %s\n", c.lsuD ? "LSU" : "TMA", per * 148 / 1048576.0, c.B,
               c.S, c.threads, c.pre, c.pdl, c.spin, us, per * 148 / us / 1e3, e == cudaSuccess ? "" : cudaGetErrorString(e));
        fflush(stdout);
        std::vector<unsigned long long> h(3 * 1024 * 8);
        cudaMemcpy(h.data(), stamps, h.size() * 8, cudaMemcpyDeviceToHost);
        unsigned long long t0 = ~0ull;
        for (int b = 0; b < 148; ++b) t0 = std::min(t0, h[b * 8]);
        const char *nm[5] = {"start", "waited", "first-in", "last-in", "end"};
        for (int l = 1; l < 3; ++l) {
            printf("    launch %d:", l);
            for (int k = 0; k < 5; ++k) {
                std::vector<double> v;
                for (int b = 0; b < 148; ++b) v.push_back((double)(h[((size_t)l * 1024 + b) * 8 + k] - t0) / 1e3);
                std::sort(v.begin(), v.end());
                printf("  %s %.2f/%.2f/%.2f", nm[k], v[0], v[74], v[147]);
            }
            printf("\n");
        }
        cudaGraphExecDestroy(ge); cudaGraphDestroy(g);
    }
    return 0;
}
