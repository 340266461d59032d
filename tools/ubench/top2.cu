// Throughput of top-2 update formulations over streams of 32 "dot products" per iteration (values via LDS.128).
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdint>
#define ITERS 2048
__device__ __forceinline__ void top2_s32(int& k1, int& k2, int x, int y)
{
    const int lo = min(x, y), hi = max(x, y);
    const int n2 = __vimin3_s32(max(k1, lo), k2, hi);
    k1 = min(k1, lo); k2 = n2;
}
__device__ __forceinline__ void top2_u16(uint32_t& k1, uint32_t& k2, uint32_t x, uint32_t y)
{
    const uint32_t lo = __vminu2(x, y), hi = __vmaxu2(x, y);
    const uint32_t n2 = __vimin3_u16x2(__vmaxu2(k1, lo), k2, hi);
    k1 = __vminu2(k1, lo); k2 = n2;
}
__device__ __forceinline__ uint32_t hmn(uint32_t a, uint32_t b) { __half2 r = __hmin2(*(__half2*)&a, *(__half2*)&b); return *(uint32_t*)&r; }
__device__ __forceinline__ uint32_t hmx(uint32_t a, uint32_t b) { __half2 r = __hmax2(*(__half2*)&a, *(__half2*)&b); return *(uint32_t*)&r; }
__device__ __forceinline__ void top2_h2(uint32_t& k1, uint32_t& k2, uint32_t x, uint32_t y)
{
    const uint32_t lo = hmn(x, y), hi = hmx(x, y);
    const uint32_t n2 = hmn(hmn(hmx(k1, lo), k2), hi);
    k1 = hmn(k1, lo); k2 = n2;
}
template <int MODE>
__global__ void __launch_bounds__(512, 1) k(uint32_t* out, int seed)
{
    __shared__ __align__(16) int32_t s_v[64 * 32];
    __shared__ __align__(16) int32_t s_c[64];
    for (int i = threadIdx.x; i < 64 * 32; i += blockDim.x) s_v[i] = (i * seed) & 127;
    if (threadIdx.x < 64) s_c[threadIdx.x] = (threadIdx.x * 977 + seed) & 0x3fff;
    __syncthreads();
    int k1[4], k2[4];
    uint32_t p1[4], p2[4];
    for (int c = 0; c < 4; ++c) { k1[c] = k2[c] = 0x7fffffff; p1[c] = p2[c] = 0x7bff7bffu; }
    for (int it = 0; it < ITERS; ++it) {
        const int4* vp = (const int4*)(s_v + (it & 63) * 32);
        int v[32];
#pragma unroll
        for (int j = 0; j < 8; ++j) { int4 t = vp[j]; v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w; }
        if (MODE == 0) {
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
                const int4 cc = *(const int4*)(s_c + j), cd = *(const int4*)(s_c + j + 4);
                top2_s32(k1[0], k2[0], v[j] * -131072 + cc.x, v[j + 1] * -131072 + cc.y);
                top2_s32(k1[1], k2[1], v[j + 2] * -131072 + cc.z, v[j + 3] * -131072 + cc.w);
                top2_s32(k1[2], k2[2], v[j + 4] * -131072 + cd.x, v[j + 5] * -131072 + cd.y);
                top2_s32(k1[3], k2[3], v[j + 6] * -131072 + cd.z, v[j + 7] * -131072 + cd.w);
            }
        } else {
            // packed: element pair (j, j+1) -> one register of 16-bit keys c - 128 * v
            uint32_t pk[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const uint32_t c2 = ((const uint32_t*)s_c)[j];
                const uint32_t t = (uint32_t)v[2 * j] * (uint32_t)(-128) + c2;
                pk[j] = (uint32_t)v[2 * j + 1] * (uint32_t)(-128 * 65536) + t;
            }
#pragma unroll
            for (int j = 0; j < 16; j += 4) {
                if (MODE == 1) { top2_u16(p1[0], p2[0], pk[j], pk[j + 1]); top2_u16(p1[1], p2[1], pk[j + 2], pk[j + 3]); }
                if (MODE == 2) { top2_h2(p1[0], p2[0], pk[j], pk[j + 1]); top2_h2(p1[1], p2[1], pk[j + 2], pk[j + 3]); }
                if (MODE == 3) { top2_u16(p1[0], p2[0], pk[j], pk[j + 1]); top2_h2(p1[1], p2[1], pk[j + 2], pk[j + 3]); }
            }
        }
    }
    uint32_t s = 0;
    for (int c = 0; c < 4; ++c) s += k1[c] ^ k2[c] ^ p1[c] ^ p2[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE>
void run(const char* name)
{
    uint32_t* d; cudaMalloc(&d, 148 * 512 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<148, 512>>>(d, 12345);
    cudaEventRecord(e0);
    k<MODE><<<148, 512>>>(d, 12345);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double elems = (double)148 * 512 * ITERS * 32;
    printf("%-34s %.3f ms  %.2f T elements/s   %.3f thread-elements/clk/SMSP\n", name, ms, elems / (ms * 1e-3) / 1e12,
           elems / (ms * 1e-3) / (148 * 4) / 1.965e9);
    cudaFree(d);
}
int main()
{
    run<0>("s32 keys (current)");
    run<1>("u16x2 packed, ALU");
    run<2>("f16x2 packed, HMNMX2");
    run<3>("u16x2 + f16x2 chains mixed");
    return 0;
}
