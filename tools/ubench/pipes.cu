// Throughput microbenchmark of the issue pipes that matter to the decode kernel (sm_100a).
// One CTA per SM, `warps` warps, each runs `iters` x 16 independent chains of one instruction.
// Prints warp-instructions per cycle per SM (SMSP rate x 4).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

enum Op { FFMA = 0, HFMA2, HFMA2_BF16, LOP3, IDP4A, HMMA_F16, HMMA_BF16, IMAD, SHF, PRMT, I2F, HMMA_F16_F16ACC, IMMA_S8, MIX_LOP_IDP, MIX_LOP_IMAD, MIX_LOP_FFMA, MIX_IDP_FFMA, MIX_LOP_HFMA2, MIX_LOP_HMMA, LDS64, LDS128, NOPS };
const char *names[] = {"FFMA", "HFMA2.f16", "HFMA2.bf16", "LOP3", "IDP.4A", "HMMA.16816 f16->f32", "HMMA.16816 bf16->f32",
                       "IMAD", "SHF", "PRMT", "I2F", "HMMA.16816 f16->f16", "IMMA.16832 s8", "mix LOP3+IDP", "mix LOP3+IMAD", "mix LOP3+FFMA", "mix IDP+FFMA", "mix LOP3+HFMA2", "mix 5xLOP3+HMMA", "LDS.64", "LDS.128"};

template <int OP>
__global__ void bench(unsigned long long *out, int iters, uint32_t seed) {
    uint32_t r[16];
    float f[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) { r[i] = seed * (i + 1) + threadIdx.x; f[i] = (float)(r[i] & 255); }
    __shared__ uint32_t smbuf[4096];
    if (threadIdx.x == 0) smbuf[0] = seed;
    float c4[8][4];
    uint32_t h4[8][2];
    int ci[4][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) { for (int j = 0; j < 4; ++j) c4[i][j] = 0.f; h4[i][0] = h4[i][1] = 0; }
#pragma unroll
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) ci[i][j] = 0;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            if (OP == FFMA) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(f[(i + 1) & 15] * 0.f + 1.0001f), "f"(0.5f));
            if (OP == HFMA2) asm volatile("fma.rn.f16x2 %0, %0, %1, %1;" : "+r"(r[i]) : "r"(seed));
            if (OP == HFMA2_BF16) asm volatile("fma.rn.bf16x2 %0, %0, %1, %1;" : "+r"(r[i]) : "r"(seed));
            if (OP == LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(r[i]) : "r"(seed), "r"(0x64006400u));
            if (OP == IDP4A) asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(r[i]) : "r"(seed), "r"(seed ^ 0x55u));
            if (OP == IMAD) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(seed), "r"(seed ^ 3u));
            if (OP == SHF) asm volatile("shf.r.wrap.b32 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(seed), "r"(5u));
            if (OP == PRMT) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(seed), "r"(0x3120u));
            if (OP == I2F) asm volatile("cvt.rn.f32.s32 %0, %1;" : "=f"(f[i]) : "r"(r[i] + it));
            if (OP == MIX_LOP_IDP) { if (i & 1) asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(r[i]) : "r"(seed), "r"(0x64006400u)); else asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(r[i]) : "r"(seed), "r"(seed ^ 0x55u)); }
            if (OP == MIX_LOP_IMAD) { if (i & 1) asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(r[i]) : "r"(seed), "r"(0x64006400u)); else asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(seed), "r"(seed ^ 3u)); }
            if (OP == MIX_LOP_FFMA) { if (i & 1) asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(r[i]) : "r"(seed), "r"(0x64006400u)); else asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(1.0001f), "f"(0.5f)); }
            if (OP == MIX_IDP_FFMA) { if (i & 1) asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(r[i]) : "r"(seed), "r"(seed ^ 0x55u)); else asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(1.0001f), "f"(0.5f)); }
            if (OP == MIX_LOP_HFMA2) { if (i & 1) asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(r[i]) : "r"(seed), "r"(0x64006400u)); else asm volatile("fma.rn.f16x2 %0, %0, %1, %1;" : "+r"(r[i]) : "r"(seed)); }
            if (OP == MIX_LOP_HMMA) { if (i < 10) asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(r[i]) : "r"(seed), "r"(0x64006400u)); else if (i < 12)
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(c4[i-10][0]), "+f"(c4[i-10][1]), "+f"(c4[i-10][2]), "+f"(c4[i-10][3])
                             : "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])); }
            if (OP == LDS64) { uint2 v; asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"((r[i] & 0x3F8u))); r[(i+1)&15] ^= v.x & v.y & 1u; }
            if (OP == LDS128) { uint4 v; asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(((threadIdx.x & 31) * 16 + i * 512) & 0x3FF0u)); r[i] ^= v.x & v.y & v.z & v.w & 1u; }
            if (OP == HMMA_F16 && i < 8)
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(c4[i][0]), "+f"(c4[i][1]), "+f"(c4[i][2]), "+f"(c4[i][3])
                             : "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]));
            if (OP == HMMA_BF16 && i < 8)
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(c4[i][0]), "+f"(c4[i][1]), "+f"(c4[i][2]), "+f"(c4[i][3])
                             : "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]));
            if (OP == HMMA_F16_F16ACC && i < 8)
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f16.f16.f16.f16 {%0,%1}, {%2,%3,%4,%5}, {%6,%7}, {%0,%1};"
                             : "+r"(h4[i][0]), "+r"(h4[i][1])
                             : "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]));
            if (OP == IMMA_S8 && i < 4)
                asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+r"(ci[i][0]), "+r"(ci[i][1]), "+r"(ci[i][2]), "+r"(ci[i][3])
                             : "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]));
        }
    }
    long long t1 = clock64();
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) acc ^= r[i] ^ __float_as_uint(f[i]);
#pragma unroll
    for (int i = 0; i < 8; ++i) acc ^= __float_as_uint(c4[i][0] + c4[i][1] + c4[i][2] + c4[i][3]) ^ h4[i][0] ^ h4[i][1];
#pragma unroll
    for (int i = 0; i < 4; ++i) acc ^= (uint32_t)(ci[i][0] + ci[i][1] + ci[i][2] + ci[i][3]);
    if (threadIdx.x == 0) { out[blockIdx.x * 2] = (unsigned long long)(t1 - t0); out[blockIdx.x * 2 + 1] = acc; }
}

template <int OP>
void run(int warps, int per_iter) {
    unsigned long long *d, h[2];
    cudaMalloc(&d, 2 * 148 * sizeof(unsigned long long));
    const int iters = 2000;
    bench<OP><<<148, warps * 32>>>(d, iters, 12345u);
    bench<OP><<<148, warps * 32>>>(d, iters, 12345u);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    double cyc = (double)h[0];
    double inst = (double)iters * per_iter * warps;
    printf("%-24s warps=%2d  %.3f warp-inst/cycle/SM  (%.2f cycles per inst per SMSP)%s\n", names[OP], warps, inst / cyc,
           cyc / (inst / 4.0), e == cudaSuccess ? "" : cudaGetErrorString(e));
    cudaFree(d);
}

int main() {
    for (int warps : {8, 16}) {
        run<FFMA>(warps, 16); run<HFMA2>(warps, 16); run<HFMA2_BF16>(warps, 16); run<LOP3>(warps, 16); run<IMAD>(warps, 16);
        run<SHF>(warps, 16); run<PRMT>(warps, 16); run<I2F>(warps, 16); run<IDP4A>(warps, 16);
        run<HMMA_F16>(warps, 8); run<HMMA_BF16>(warps, 8); run<HMMA_F16_F16ACC>(warps, 8); run<IMMA_S8>(warps, 4);
        run<MIX_LOP_IDP>(warps, 16); run<MIX_LOP_IMAD>(warps, 16); run<MIX_LOP_FFMA>(warps, 16); run<MIX_IDP_FFMA>(warps, 16); run<MIX_LOP_HFMA2>(warps, 16);
        run<MIX_LOP_HMMA>(warps, 12); run<LDS64>(warps, 16);
        printf("\n");
    }
    return 0;
}
