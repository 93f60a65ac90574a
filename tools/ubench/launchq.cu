// Launch-period microbenchmark (sm_100a): a kernel whose CTAs spin for `d` ns, launched back to back in one stream.
// Prints the average period per launch against d for plain launches, programmatic dependent launches
// (griddepcontrol.wait at the top / at the bottom of the kernel) and the same inside a CUDA graph.
// Shows the fixed cost and any quantisation of kernel-to-kernel hand-over.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

// mode 0: no griddepcontrol; 1: launch_dependents first, wait, then spin; 2: launch_dependents, spin, then wait
__global__ void spin_kernel(unsigned long long d_ns, int mode, unsigned *sink) {
    if (mode) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (mode == 1) asm volatile("griddepcontrol.wait;" ::: "memory");
    unsigned long long t0, t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    do { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); } while (t - t0 < d_ns);
    if (mode == 2) asm volatile("griddepcontrol.wait;" ::: "memory");
    if (t == 1) sink[0] = 1;
    if (threadIdx.x == 0 && blockIdx.x == 0) sink[1] = (unsigned)t;   // a dependent write
}

static void launch(unsigned long long d, int mode, bool pdl, unsigned *sink, cudaStream_t st, int grid, int block) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, spin_kernel, d, mode, sink);
}

int main() {
    unsigned *sink; cudaMalloc(&sink, 8);
    cudaStream_t st; cudaStreamCreate(&st);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int N = 400;
    struct Cfg { const char *name; int mode; bool pdl; bool graph; int grid; int block; };
    const Cfg cfgs[] = {
        {"plain stream, grid 148x512", 0, false, false, 148, 512},
        {"plain graph,  grid 148x512", 0, false, true, 148, 512},
        {"PDL wait-first graph, 148x512", 1, true, true, 148, 512},
        {"PDL wait-last  graph, 148x512", 2, true, true, 148, 512},
        {"PDL wait-last  stream, 148x512", 2, true, false, 148, 512},
        {"plain graph,  grid 1x32", 0, false, true, 1, 32},
    };
    for (const Cfg &c : cfgs) {
        printf("%s\n", c.name);
        for (unsigned long long d = 0; d <= 9000; d += 500) {
            cudaGraphExec_t ge = nullptr;
            if (c.graph) {
                cudaGraph_t g;
                cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
                for (int i = 0; i < N; ++i) launch(d, c.mode, c.pdl, sink, st, c.grid, c.block);
                cudaStreamEndCapture(st, &g);
                cudaGraphInstantiate(&ge, g, 0);
                cudaGraphLaunch(ge, st);
                cudaStreamSynchronize(st);
            } else {
                for (int i = 0; i < N; ++i) launch(d, c.mode, c.pdl, sink, st, c.grid, c.block);
                cudaStreamSynchronize(st);
            }
            cudaEventRecord(e0, st);
            if (c.graph) cudaGraphLaunch(ge, st);
            else for (int i = 0; i < N; ++i) launch(d, c.mode, c.pdl, sink, st, c.grid, c.block);
            cudaEventRecord(e1, st);
            cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            cudaError_t e = cudaGetLastError();
            printf("   spin %5llu ns -> period %7.3f us  (overhead %6.3f us) %s\n", d, ms * 1e3 / N, ms * 1e3 / N - d / 1e3,
                   e == cudaSuccess ? "" : cudaGetErrorString(e));
            if (ge) cudaGraphExecDestroy(ge);
        }
    }
    return 0;
}
