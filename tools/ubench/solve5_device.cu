// solve5_device.cu -- the five-point solver (csrc/geom_5pt.cuh) run on the device and on the host on the same minimal
// samples; prints the number of solutions and E[0] of each, per sample, for both.  They must agree.
//
// Why it exists: with polish_5pt inlined into solve_5pt_in_basis, cicc 12.9 at -O1 and above generated device code in
// which the root list of the degree-10 polynomial was overwritten (entries 1.. read back as 0 after the first polish),
// so the kernels silently lost solutions; -Xcicc -O0, -G and the host build were correct.  Build both ways to see it:
//   nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -fmad=false -O3 -o solve5 tools/ubench/solve5_device.cu
//   (define VS_NOINLINE as empty in csrc/geom_solvers.cuh to get the inlined, broken variant)
// Input file: int32 n, then n x 5 x 2 doubles (view-1 points, normalised), then the same for view 2.
#include <stdio.h>
#include <stdlib.h>

#include "../../visual-slam_b200/csrc/geom_5pt.cuh"

__global__ void k(const double* x1, const double* x2, int nsamp, double* E, int* ne)
{
    const int it = blockIdx.x * blockDim.x + threadIdx.x;
    if (it >= nsamp) return;
    double a[5][2], b[5][2], out[10][9];
    for (int j = 0; j < 5; ++j) {
        a[j][0] = x1[it * 10 + 2 * j]; a[j][1] = x1[it * 10 + 2 * j + 1];
        b[j][0] = x2[it * 10 + 2 * j]; b[j][1] = x2[it * 10 + 2 * j + 1];
    }
    const int n = solve_5pt(a, b, out);
    for (int q = 0; q < n; ++q)
        for (int e = 0; e < 9; ++e) E[(it * 10 + q) * 9 + e] = out[q][e];
    ne[it] = n;
}

int main(int argc, char** argv)
{
    FILE* f = argc > 1 ? fopen(argv[1], "rb") : nullptr;
    int nsamp = 0;
    if (!f || fread(&nsamp, 4, 1, f) != 1 || nsamp < 1) { fprintf(stderr, "usage: solve5 samples.bin\n"); return 2; }
    double* h1 = (double*)malloc(nsamp * 80);
    double* h2 = (double*)malloc(nsamp * 80);
    if (fread(h1, 80, nsamp, f) != (size_t)nsamp || fread(h2, 80, nsamp, f) != (size_t)nsamp) return 2;
    double *d1, *d2, *dE;
    int* dn;
    cudaMalloc(&d1, nsamp * 80); cudaMalloc(&d2, nsamp * 80); cudaMalloc(&dE, nsamp * 720); cudaMalloc(&dn, nsamp * 4);
    cudaMemcpy(d1, h1, nsamp * 80, cudaMemcpyHostToDevice);
    cudaMemcpy(d2, h2, nsamp * 80, cudaMemcpyHostToDevice);
    k<<<(nsamp + 31) / 32, 32>>>(d1, d2, nsamp, dE, dn);
    int* hn = (int*)malloc(nsamp * 4);
    double* hE = (double*)malloc(nsamp * 720);
    cudaError_t e = cudaMemcpy(hn, dn, nsamp * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(hE, dE, nsamp * 720, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) { fprintf(stderr, "%s\n", cudaGetErrorString(e)); return 1; }
    int bad = 0;
    for (int i = 0; i < nsamp; ++i) {
        double a[5][2], b[5][2], out[10][9];
        for (int j = 0; j < 5; ++j) {
            a[j][0] = h1[i * 10 + 2 * j]; a[j][1] = h1[i * 10 + 2 * j + 1];
            b[j][0] = h2[i * 10 + 2 * j]; b[j][1] = h2[i * 10 + 2 * j + 1];
        }
        const int n = solve_5pt(a, b, out);
        bool same = n == hn[i];
        for (int q = 0; same && q < n; ++q) same = fabs(out[q][0] - hE[(i * 10 + q) * 9]) < 1e-9;
        printf("sample %d: device %d host %d %s\n", i, hn[i], n, same ? "ok" : "MISMATCH");
        if (!same && argc > 2) {
            for (int q = 0; q < hn[i]; ++q) printf("   dev  %d: %.12f %.12f %.12f\n", q, hE[(i * 10 + q) * 9], hE[(i * 10 + q) * 9 + 1], hE[(i * 10 + q) * 9 + 2]);
            for (int q = 0; q < n; ++q) printf("   host %d: %.12f %.12f %.12f\n", q, out[q][0], out[q][1], out[q][2]);
        }
        bad += !same;
    }
    printf("%d of %d samples differ\n", bad, nsamp);
    return bad != 0;
}
