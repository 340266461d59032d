#include <stdio.h>
#include <stdint.h>
#include <stdlib.h>
#include <algorithm>
#define RAW_S 80
__global__ void k(const uint8_t* img, int* out, int* dbg, int n){
    int i = blockIdx.x*blockDim.x+threadIdx.x; if (i>=n) return;
    int y = 3 + i / 74, x = 3 + i % 74;
    const uint8_t* p = img + y*RAW_S + x;
    const int off[16] = {3 * RAW_S, 3 * RAW_S + 1, 2 * RAW_S + 2, RAW_S + 3, 3, -RAW_S + 3,
                         -2 * RAW_S + 2, -3 * RAW_S + 1, -3 * RAW_S, -3 * RAW_S - 1,
                         -2 * RAW_S - 2, -RAW_S - 3, -3, RAW_S - 3, 2 * RAW_S - 2, 3 * RAW_S - 1};
    const int c = p[0];
    int v[16];
#pragma unroll
    for (int k2 = 0; k2 < 16; ++k2) v[k2] = p[off[k2]];
    int bst = -256;
    int* g = dbg + 48*i;
#pragma unroll
    for (int k2 = 0; k2 < 16; ++k2) {
        int mn = 1000, mx = -1000;
#pragma unroll
        for (int j = 0; j < 9; ++j) { int dd = c - v[(k2 + j) & 15]; mn = min(mn, dd); mx = max(mx, dd); }
        g[k2] = mn; g[16+k2] = mx; g[32+k2] = c - v[k2];
        bst = max(bst, max(mn, -mx));
    }
    out[i] = bst;
}
int main(){
    const int H=80; uint8_t* h=(uint8_t*)malloc(H*RAW_S); srand(1);
    for(int i=0;i<H*RAW_S;i++){ int y=i/RAW_S,x=i%RAW_S; h[i]= (uint8_t)((((x/9+y/7)&1)?200:60) + rand()%30); }
    int n=74*74; uint8_t* d; int *o,*g; cudaMalloc(&d,H*RAW_S); cudaMalloc(&o,n*4); cudaMalloc(&g,n*48*4);
    cudaMemcpy(d,h,H*RAW_S,cudaMemcpyHostToDevice);
    k<<<(n+127)/128,128>>>(d,o,g,n);
    int* ho=(int*)malloc(n*4); int* hg=(int*)malloc(n*48*4); cudaMemcpy(ho,o,n*4,cudaMemcpyDeviceToHost); cudaMemcpy(hg,g,n*48*4,cudaMemcpyDeviceToHost);
    int i = 0*74 + 12; // x15 y3
    printf("dev bst %d\n d :", ho[i]); for(int k2=0;k2<16;k2++) printf(" %d", hg[48*i+32+k2]);
    printf("\n mn:"); for(int k2=0;k2<16;k2++) printf(" %d", hg[48*i+k2]);
    printf("\n mx:"); for(int k2=0;k2<16;k2++) printf(" %d", hg[48*i+16+k2]); printf("\n");
}
