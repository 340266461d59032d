#include <cstdio>
#include <cstdint>
// throughput + semantics probe of the 1-bit tensor-core MMA on sm_100a
__device__ __forceinline__ void mma_b1_and(int (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2])
{
    asm volatile("mma.sync.aligned.m16n8k256.row.col.s32.b1.b1.s32.and.popc {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
#ifdef TRY_XOR
__device__ __forceinline__ void mma_b1_xor(int (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2])
{
    asm volatile("mma.sync.aligned.m16n8k256.row.col.s32.b1.b1.s32.xor.popc {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
#endif
#define ITERS 8192
__global__ void k(int* out, uint32_t seed)
{
    uint32_t a[4], b[4][2];
    int d[4][4] = {};
    for (int i = 0; i < 4; ++i) { a[i] = seed * (threadIdx.x + i + 1); b[i][0] = seed ^ (threadIdx.x * 77 + i); b[i][1] = b[i][0] * 3; }
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
#ifdef TRY_XOR
            mma_b1_xor(d[j], a, b[j]);
#else
            mma_b1_and(d[j], a, b[j]);
#endif
        }
    }
    int s = 0;
    for (int j = 0; j < 4; ++j) for (int i = 0; i < 4; ++i) s += d[j][i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// semantics: one warp, A = 16 rows x 256 bits, B = 8 cols x 256 bits, D[i][j] = popc(A_i & B_j)
__global__ void sem(const uint32_t* A, const uint32_t* B, int* D)
{
    const int lane = threadIdx.x, g = lane >> 2, t = lane & 3;
    // fragment layout of m16n8k256 b1 (as m16n8k32 s8 with 8x wider k): a0:(row g, k-words t*... )
    uint32_t a[4], b[2];
    a[0] = A[g * 8 + t];           // row g,   k words [0..3]  -> word t
    a[1] = A[(g + 8) * 8 + t];     // row g+8, k words [0..3]
    a[2] = A[g * 8 + 4 + t];       // row g,   k words [4..7]
    a[3] = A[(g + 8) * 8 + 4 + t];
    b[0] = B[g * 8 + t];           // col g, k words [0..3]
    b[1] = B[g * 8 + 4 + t];
    int d[4] = {0, 0, 0, 0};
    mma_b1_and(d, a, b);
    D[g * 8 + 2 * t] = d[0]; D[g * 8 + 2 * t + 1] = d[1];
    D[(g + 8) * 8 + 2 * t] = d[2]; D[(g + 8) * 8 + 2 * t + 1] = d[3];
}
int main()
{
    int* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<<<148 * 8, 256>>>(d, 12345);
    cudaEventRecord(e0);
    k<<<148 * 8, 256>>>(d, 12345);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double mmas = (double)148 * 8 * 8 * ITERS * 4;
    printf("b1 mma m16n8k256: %.3f ms, %.4f mma/clk/SMSP @1965MHz, %.1f Tpairs/s (256-bit pairs)\n", ms,
           mmas / (ms * 1e-3) / (148 * 4) / 1.965e9, mmas * 128 / (ms * 1e-3) / 1e12);
    printf("err: %s\n", cudaGetErrorString(cudaGetLastError()));
    uint32_t hA[128], hB[64]; int hD[128];
    for (int i = 0; i < 128; ++i) hA[i] = 0x9E3779B9u * (i + 1) ^ (i << 13);
    for (int i = 0; i < 64; ++i) hB[i] = 0x85EBCA6Bu * (i + 7) ^ (i << 7);
    uint32_t *dA, *dB; int* dD; cudaMalloc(&dA, 512); cudaMalloc(&dB, 256); cudaMalloc(&dD, 512);
    cudaMemcpy(dA, hA, 512, cudaMemcpyHostToDevice); cudaMemcpy(dB, hB, 256, cudaMemcpyHostToDevice);
    sem<<<1, 32>>>(dA, dB, dD); cudaMemcpy(hD, dD, 512, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int i = 0; i < 16; ++i) for (int j = 0; j < 8; ++j) {
        int r = 0; for (int w = 0; w < 8; ++w) r += __builtin_popcount(hA[i * 8 + w] & hB[j * 8 + w]);
        if (r != hD[i * 8 + j]) ++bad;
    }
    printf("semantics: %d of 128 mismatches (err %s)\n", bad, cudaGetErrorString(cudaGetLastError()));
    return 0;
}
