#include <cstdio>
#include <cstdint>
// throughput of the legacy int8 tensor-core MMA (mma.sync m16n8k32 s8) on sm_100a
__device__ __forceinline__ void imma(int (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2])
{
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
#define ITERS 8192
template <int NACC>
__global__ void k(int* out, uint32_t seed)
{
    uint32_t a[4], b[NACC][2];
    int d[NACC][4] = {};
    for (int i = 0; i < 4; ++i) a[i] = seed * (threadIdx.x + i + 1);
    for (int i = 0; i < NACC; ++i) { b[i][0] = seed ^ (threadIdx.x * 77 + i); b[i][1] = b[i][0] * 3; }
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int j = 0; j < NACC; ++j) imma(d[j], a, b[j]);
    }
    int s = 0;
    for (int j = 0; j < NACC; ++j) for (int i = 0; i < 4; ++i) s += d[j][i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int NACC>
void run(int warps_per_cta, int ctas_per_sm)
{
    int* d; cudaMalloc(&d, 148 * 16 * 1024 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<NACC><<<148 * ctas_per_sm, warps_per_cta * 32>>>(d, 12345);
    cudaEventRecord(e0);
    k<NACC><<<148 * ctas_per_sm, warps_per_cta * 32>>>(d, 12345);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double mmas = (double)148 * ctas_per_sm * warps_per_cta * ITERS * NACC;
    printf("NACC %d warps/SM %2d: %.3f ms  %.4f imma/clk/SMSP  %.1f TMAC/s  (%s)\n", NACC, warps_per_cta * ctas_per_sm, ms,
           mmas / (ms * 1e-3) / (148 * 4) / 1.965e9, mmas * 4096 / (ms * 1e-3) / 1e12, cudaGetErrorString(cudaGetLastError()));
    cudaFree(d);
}
int main()
{
    run<1>(4, 1); run<4>(4, 1); run<8>(4, 1); run<8>(8, 1); run<8>(8, 2); run<8>(8, 4);
    return 0;
}
