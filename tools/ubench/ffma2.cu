// ffma2.cu -- does the packed fp32 FMA of sm_100 (fma.rn.f32x2 -> FFMA2) free issue slots?
//
// The level kernel is issue-bound and its Gaussian is ~25 % FFMA.  FFMA2 does two fp32 FMAs per thread in one
// instruction; what matters here is whether a warp-level FFMA2 costs ONE issue slot (then the ALU work of the same
// warp scheduler fits in the slots it frees) or two.  Modes, register-only independent chains on every SM:
//   0  FFMA  only              1  FFMA2 only (same number of scalar FMAs)
//   2  FFMA  + as many LOP3    3  FFMA2 + as many LOP3 as mode 2 (same scalar FMAs, same LOP3 count)
//   4  LOP3 only
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2 tools/ubench/ffma2.cu && ./ffma2
// One JSON line: scalar FMA/s and time per mode.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define CH 8
#define ITERS 4096

__device__ __forceinline__ unsigned long long pk(float a, float b)
{
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c)
{
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, float seed, uint32_t useed)
{
    float x[2 * CH];
    unsigned long long p[CH];
    uint32_t u[2 * CH];
    const float m = 1.0f + seed * 1e-9f, a = seed * 1e-7f;
    const unsigned long long m2 = pk(m, m), a2 = pk(a, a);
#pragma unroll
    for (int c = 0; c < 2 * CH; ++c) { x[c] = seed + threadIdx.x + c; u[c] = useed * (threadIdx.x + 1) + c; }
#pragma unroll
    for (int c = 0; c < CH; ++c) p[c] = pk(x[2 * c], x[2 * c + 1]);
#pragma unroll 1
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int c = 0; c < CH; ++c) {
            if (MODE == 0 || MODE == 2) { x[2 * c] = fmaf(x[2 * c], m, a); x[2 * c + 1] = fmaf(x[2 * c + 1], m, a); }
            if (MODE == 1 || MODE == 3) p[c] = fma2(p[c], m2, a2);
            if (MODE >= 2) { u[2 * c] = (u[2 * c] & useed) ^ 0x9e3779b9u; u[2 * c + 1] = (u[2 * c + 1] & useed) ^ 0x7f4a7c15u; }
        }
    }
    float r = 0;
#pragma unroll
    for (int c = 0; c < 2 * CH; ++c) r += x[c] + (float)u[c];
#pragma unroll
    for (int c = 0; c < CH; ++c) { float lo, hi; asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[c])); r += lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int MODE>
static float run(int grid, float* d_out)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int r = 0; r < 300; ++r) k<MODE><<<grid, 256>>>(d_out, 1.5f, 0xffff00ffu);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 20; ++r) {
        cudaEventRecord(e0);
        k<MODE><<<grid, 256>>>(d_out, 1.5f + r, 0xffff00ffu + r);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best;
}

int main()
{
    cudaDeviceProp pr;
    cudaGetDeviceProperties(&pr, 0);
    const int grid = pr.multiProcessorCount * 4;
    float* d_out;
    cudaMalloc(&d_out, sizeof(float) * grid * 256);
    const float t0 = run<0>(grid, d_out), t1 = run<1>(grid, d_out), t2 = run<2>(grid, d_out), t3 = run<3>(grid, d_out),
                t4 = run<4>(grid, d_out);
    const double fmas = (double)grid * 256 * 2 * CH * ITERS;
    printf("{\"ubench\": \"ffma2\", \"sms\": %d, \"scalar_fma_per_launch\": %.0f, \"ms\": {\"ffma\": %.4f, \"ffma2\": %.4f, "
           "\"ffma_lop3\": %.4f, \"ffma2_lop3\": %.4f, \"lop3\": %.4f}, \"tflops\": {\"ffma\": %.2f, \"ffma2\": %.2f}}\n",
           pr.multiProcessorCount, fmas, t0, t1, t2, t3, t4, 2 * fmas / (t0 * 1e-3) / 1e12, 2 * fmas / (t1 * 1e-3) / 1e12);
    return 0;
}
