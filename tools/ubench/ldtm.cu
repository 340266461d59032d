// Throughput of tcgen05.ld (TMEM -> registers) per SM: up to 16 warps (4 per lane quadrant) read 32x32b.x32 chunks
// in a loop; reports TMEM bytes / cycle / SM for plain and .pack::16b loads and for 1 / 2 loads in flight per wait.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ldtm tools/ubench/ldtm.cu && ./ldtm
#include <cstdint>
#include <cstdio>
#define ITERS 4096

#define LD32_BODY(OPC)                                                                                                   \
    asm volatile(OPC " {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "                         \
                     "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"           \
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),       \
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), \
                   "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]),            \
                   "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]),            \
                   "=r"(v[30]), "=r"(v[31])                                                                              \
                 : "r"(taddr))

__device__ __forceinline__ void ld32(uint32_t (&v)[32], uint32_t taddr) { LD32_BODY("tcgen05.ld.sync.aligned.32x32b.x32.b32"); }
// 32 registers = 64 TMEM columns, the low 16 bits of two adjacent columns packed per register
__device__ __forceinline__ void ld32_pack(uint32_t (&v)[32], uint32_t taddr) { LD32_BODY("tcgen05.ld.sync.aligned.32x32b.x32.pack::16b.b32"); }

// MODE 0: one x32 load per wait; 1: two x32 loads per wait; 2: one packed load (64 columns) per wait
template <int MODE>
__global__ void __launch_bounds__(512, 1) k(uint32_t* out, long long* cycles, int warps)
{
    __shared__ uint32_t s_tmem;
    const int wid = threadIdx.x >> 5;
    if (wid == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_tmem)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = s_tmem;
    uint32_t acc = 0;
    const long long t0 = clock64();
    if (wid < warps) {
        const uint32_t lane_base = tmem + (((uint32_t)(wid & 3) * 32u) << 16);
        for (int it = 0; it < ITERS; ++it) {
            const uint32_t col = (uint32_t)((it * 64 + (wid >> 2) * 128) & 448);
            uint32_t va[32], vb[32];
            if (MODE == 0) {
                ld32(va, lane_base + col);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int j = 0; j < 32; ++j) acc ^= va[j];
            } else if (MODE == 1) {
                ld32(va, lane_base + col);
                ld32(vb, lane_base + col + 32);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int j = 0; j < 32; ++j) acc ^= va[j] + vb[j];
            } else {
                ld32_pack(va, lane_base + col);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int j = 0; j < 32; ++j) acc ^= va[j];
            }
        }
    }
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (wid == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
}

template <int MODE>
void run(const char* name, int warps, int cols_per_iter)
{
    uint32_t* d; long long* c;
    cudaMalloc(&d, 148 * 512 * 4); cudaMalloc(&c, 148 * 8);
    k<MODE><<<148, 512>>>(d, c, warps);
    k<MODE><<<148, 512>>>(d, c, warps);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, c, sizeof(h), cudaMemcpyDeviceToHost);
    double cyc = 0;
    for (int i = 0; i < 148; ++i) cyc += (double)h[i] / 148;
    // TMEM cells read per SM: warps x 32 lanes x columns x 4 B per iteration
    const double bytes = (double)warps * 32 * cols_per_iter * 4 * ITERS;
    printf("%-30s warps %2d  %9.0f cycles  %7.1f TMEM B/clk/SM  %6.1f cycles per iteration  (%s)\n", name, warps, cyc,
           bytes / cyc, cyc / ITERS, cudaGetErrorString(e));
    cudaFree(d); cudaFree(c);
}

int main()
{
    for (int w : {1, 4, 8, 16}) {
        run<0>("x32, one per wait", w, 32);
        run<1>("x32, two per wait", w, 64);
        run<2>("x32.pack::16b, one per wait", w, 64);
    }
    return 0;
}
