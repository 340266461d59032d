#include <stdio.h>
#include <stdint.h>
#include <stdlib.h>
#include <algorithm>
#define RAW_S 80
#ifndef VARIANT
#define VARIANT 0
#endif
__host__ __device__ inline bool run9(uint32_t m){ m |= m << 16; m &= m >> 1; m &= m >> 2; m &= m >> 4; m &= m >> 1; return m != 0; }
#ifndef __CUDA_ARCH__
using std::min; using std::max;
#endif
__host__ __device__ inline int fast_score(const uint8_t* p, int thr, int* dbg)
{
    const int off[16] = {3 * RAW_S, 3 * RAW_S + 1, 2 * RAW_S + 2, RAW_S + 3, 3, -RAW_S + 3,
                         -2 * RAW_S + 2, -3 * RAW_S + 1, -3 * RAW_S, -3 * RAW_S - 1,
                         -2 * RAW_S - 2, -RAW_S - 3, -3, RAW_S - 3, 2 * RAW_S - 2, 3 * RAW_S - 1};
    const int c = p[0];
    int v[16];
    uint32_t bm = 0, dm = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        v[k] = p[off[k]];
        bm |= (uint32_t)(v[k] > c + thr) << k;
        dm |= (uint32_t)(v[k] < c - thr) << k;
    }
    dbg[0] = bm; dbg[1] = dm;
    if (!run9(bm) && !run9(dm)) return 0;
#if VARIANT==1
    int d[16], e[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) { d[k] = c - v[k]; e[k] = v[k] - c; }
    int l3[16], m3[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        l3[k] = min(d[k], min(d[(k + 1) & 15], d[(k + 2) & 15]));
        m3[k] = min(e[k], min(e[(k + 1) & 15], e[(k + 2) & 15]));
    }
    int bst = -256;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        int mn = min(l3[k], min(l3[(k + 3) & 15], l3[(k + 6) & 15]));
        int mm = min(m3[k], min(m3[(k + 3) & 15], m3[(k + 6) & 15]));
        bst = max(bst, max(mn, mm));
    }
    return bst > thr ? bst - 1 : 0;
#elif VARIANT==2
    int bst = -256;
    for (int k = 0; k < 16; ++k) {
        int mn = 1000, mx = -1000;
        for (int j = 0; j < 9; ++j) { int dd = c - v[(k + j) & 15]; mn = min(mn, dd); mx = max(mx, dd); }
        bst = max(bst, max(mn, -mx));
    }
    return bst > thr ? bst - 1 : 0;
#endif
    int d[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) d[k] = c - v[k];
    int lo3[16], hi3[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        lo3[k] = min(d[k], min(d[(k + 1) & 15], d[(k + 2) & 15]));
        hi3[k] = max(d[k], max(d[(k + 1) & 15], d[(k + 2) & 15]));
    }
    int best = -256;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        int mn = min(lo3[k], min(lo3[(k + 3) & 15], lo3[(k + 6) & 15]));
        int mx = max(hi3[k], max(hi3[(k + 3) & 15], hi3[(k + 6) & 15]));
        best = max(best, max(mn, -mx));
    }
    return best > thr ? best - 1 : 0;
}
__global__ void k(const uint8_t* img, int* out, int* dbg, int n){
    int i = blockIdx.x*blockDim.x+threadIdx.x; if (i>=n) return;
    int y = 3 + i / 74, x = 3 + i % 74;
    out[i] = fast_score(img + y*RAW_S + x, 7, dbg + 2*i);
}
int main(){
    const int H=80; uint8_t* h=(uint8_t*)malloc(H*RAW_S); srand(1);
    for(int i=0;i<H*RAW_S;i++){ int y=i/RAW_S,x=i%RAW_S; h[i]= (uint8_t)((((x/9+y/7)&1)?200:60) + rand()%30); }
    int n=74*74; uint8_t* d; int *o,*g; cudaMalloc(&d,H*RAW_S); cudaMalloc(&o,n*4); cudaMalloc(&g,n*8);
    cudaMemcpy(d,h,H*RAW_S,cudaMemcpyHostToDevice);
    k<<<(n+127)/128,128>>>(d,o,g,n);
    int* ho=(int*)malloc(n*4); int* hg=(int*)malloc(n*8); cudaMemcpy(ho,o,n*4,cudaMemcpyDeviceToHost); cudaMemcpy(hg,g,n*8,cudaMemcpyDeviceToHost);
    int bad=0, nz=0;
    for(int i=0;i<n;i++){ int y=3+i/74,x=3+i%74; int dbg[2]; int r=fast_score(h+y*RAW_S+x,7,dbg); if(r) nz++; if(r!=ho[i]){ if(bad<10) printf("x%d y%d host %d dev %d bm %x/%x dm %x/%x\n",x,y,r,ho[i],dbg[0],hg[2*i],dbg[1],hg[2*i+1]); bad++; } }
    printf("bad %d of %d (nonzero %d) err=%s\n",bad,n,nz,cudaGetErrorString(cudaGetLastError()));
}
