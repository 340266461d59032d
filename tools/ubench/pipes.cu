// Which pipe do the min/max flavours issue on?  Throughput of dependent-free streams, alone and interleaved.
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdint>
#define ITERS 4096
#define NACC 8
template <int MODE>
__global__ void k(uint32_t* out, uint32_t seed)
{
    uint32_t a[NACC], b[NACC];
    for (int i = 0; i < NACC; ++i) { a[i] = seed * (threadIdx.x + i + 1); b[i] = seed ^ (threadIdx.x * 77 + i); }
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i) {
            if (MODE == 0 || MODE == 2 || MODE == 3 || MODE == 5) a[i] = __vimax3_u16x2(a[i], b[i], a[(i + 1) % NACC]);   // ALU
            if (MODE == 1 || MODE == 2) {                                                       // half2 min
                __half2 x = *reinterpret_cast<__half2*>(&b[i]), y = *reinterpret_cast<__half2*>(&b[(i + 3) % NACC]);
                __half2 z = __hmin2(x, y);
                b[i] = *reinterpret_cast<uint32_t*>(&z) + 1;
            }
            if (MODE == 3) b[i] = b[i] * 3 + a[i];                                              // IMAD (fma pipe)
            if (MODE == 4 || MODE == 5) {                                                       // FMNMX3
                float x = __uint_as_float(b[i]), y = __uint_as_float(b[(i + 3) % NACC]), z = __uint_as_float(b[(i + 5) % NACC]);
                float r;
                asm volatile("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(x), "f"(y), "f"(z));
                b[i] = __float_as_uint(r);
            }
        }
    }
    uint32_t s = 0;
    for (int i = 0; i < NACC; ++i) s += a[i] ^ b[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE>
void run(const char* name, int ops_per_iter)
{
    uint32_t* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<148 * 8, 256>>>(d, 12345);
    cudaEventRecord(e0);
    k<MODE><<<148 * 8, 256>>>(d, 12345);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double winst = (double)148 * 8 * 8 * ITERS * NACC * ops_per_iter;   // warp instructions of interest
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("%-28s %.3f ms  %.3f winst/clk/SMSP (at %d MHz nominal)\n", name, ms, winst / (ms * 1e-3) / (148 * 4) / (clk * 1e3), clk / 1000);
    cudaFree(d);
}
int main()
{
    run<0>("VIMNMX3.U16x2", 1);
    run<1>("HMNMX2 (+IADD)", 2);
    run<2>("VIMNMX3 + HMNMX2 + IADD", 3);
    run<3>("VIMNMX3 + IMAD", 2);
    run<4>("FMNMX3", 1);
    run<5>("VIMNMX3 + FMNMX3", 2);
    return 0;
}
