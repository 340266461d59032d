// popc.cu -- measured peak of the POPC pipe: register-only independent __popc chains on every SM.
//
// SURVEY.md 8(d) / BASELINE.md ask for `popc_peak_measured` as the denominator of the matching kernel's
// popc-pipe fraction (the XOR + POPC formulation: 8 popcounts per 256-bit descriptor pair); the paper figure
// 16 lanes/clk/SM was "unverified for cc 10.0 -- measure".
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o popc tools/ubench/popc.cu && ./popc
// Prints one JSON line: popc32 per second (all SMs), lanes per clock per SM at the measured SM clock, and the
// same for the XOR + POPC + ADD triple the kernel issues per word (POPC shares nothing with the ALU pipe if the
// triple runs at the POPC-only rate).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define CHAINS 8
#define ITERS 4096

template <int MODE>
__global__ void __launch_bounds__(256) k(uint32_t* out, uint32_t seed, long long* clk)
{
    uint32_t x[CHAINS], acc[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) { x[c] = seed * (threadIdx.x + 1) + c * 0x9e3779b9u + blockIdx.x; acc[c] = 0; }
    const long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) {
            if (MODE == 0) {
                x[c] = __popc(x[c]) + 0x01010101u * 0;      // dependent chain of POPC per register: 8 independent chains
                x[c] ^= 0xa5a5a5a5u;                         // keeps the value from collapsing (LOP3 on the ALU pipe)
            } else {
                acc[c] += __popc(x[c] ^ seed);               // the kernel's XOR + POPC + ADD triple
                x[c] += 0x9e3779b9u;
            }
        }
    }
    const long long t1 = clock64();
    uint32_t r = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) r += x[c] + acc[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

template <int MODE>
static double run(int sms, int ctas_per_sm, uint32_t* d_out, long long* d_clk, double* cyc_per_sm)
{
    const int grid = sms * ctas_per_sm;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    // warm up for about a second so that the SM clock has ramped to its boost state
    for (int r = 0; r < 600; ++r) k<MODE><<<grid, 256>>>(d_out, 12345u, d_clk);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 20; ++r) {
        cudaEventRecord(e0);
        k<MODE><<<grid, 256>>>(d_out, 12345u + r, d_clk);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    long long h_clk[4096];
    cudaMemcpy(h_clk, d_clk, sizeof(long long) * grid, cudaMemcpyDeviceToHost);
    double c = 0;
    for (int i = 0; i < grid; ++i) c += (double)h_clk[i];
    *cyc_per_sm = c / grid;                                   // cycles one CTA was resident (all CTAs co-resident)
    const double popc = (double)grid * 256 * CHAINS * ITERS;
    return popc / (best * 1e-3);
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount, ctas = 4;          // 4 x 256 threads per SM: 32 warps, 8 per scheduler
    uint32_t* d_out;
    long long* d_clk;
    cudaMalloc(&d_out, sizeof(uint32_t) * sms * ctas * 256);
    cudaMalloc(&d_clk, sizeof(long long) * sms * ctas);
    double cyc0, cyc1;
    const double r0 = run<0>(sms, ctas, d_out, d_clk, &cyc0);
    const double r1 = run<1>(sms, ctas, d_out, d_clk, &cyc1);
    const double per_cta = 256.0 * CHAINS * ITERS;
    // lanes per clock per SM from the SM's own cycle counter: popc issued by the SM's resident CTAs / cycles
    const double lanes0 = per_cta * ctas / cyc0, lanes1 = per_cta * ctas / cyc1;
    printf("{\"device\": \"%s\", \"sms\": %d, \"sm_mhz_during_run\": %.0f, \"popc32_per_s\": %.4e, \"popc_lanes_per_clk_per_sm\": %.2f, "
           "\"xor_popc_add_per_s\": %.4e, \"xor_popc_add_lanes_per_clk_per_sm\": %.2f, "
           "\"pairs256_per_s_at_8_popc\": %.4e}\n",
           p.name, sms, r0 / (lanes0 * sms) / 1e6, r0, lanes0, r1, lanes1, (r1 < r0 ? r1 : r0) / 8.0);
    return 0;
}
