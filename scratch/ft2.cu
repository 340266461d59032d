#include <stdio.h>
#include <stdint.h>
#include <stdlib.h>
#define RAW_S 80
__global__ void k(const uint8_t* img, int x, int y){
    const uint8_t* p = img + y*RAW_S + x;
    const int off[16] = {3 * RAW_S, 3 * RAW_S + 1, 2 * RAW_S + 2, RAW_S + 3, 3, -RAW_S + 3,
                         -2 * RAW_S + 2, -3 * RAW_S + 1, -3 * RAW_S, -3 * RAW_S - 1,
                         -2 * RAW_S - 2, -RAW_S - 3, -3, RAW_S - 3, 2 * RAW_S - 2, 3 * RAW_S - 1};
    int c = p[0];
    printf("c=%d\n", c);
    for (int k2 = 0; k2 < 16; ++k2) printf("k%d off %d v %d\n", k2, off[k2], (int)p[off[k2]]);
}
int main(){
    const int H=80; uint8_t* h=(uint8_t*)malloc(H*RAW_S); srand(1);
    for(int i=0;i<H*RAW_S;i++){ int y=i/RAW_S,x=i%RAW_S; h[i]= (uint8_t)((((x/9+y/7)&1)?200:60) + rand()%30); }
    uint8_t* d; cudaMalloc(&d,H*RAW_S); cudaMemcpy(d,h,H*RAW_S,cudaMemcpyHostToDevice);
    int x=13,y=3;
    const int off[16] = {3 * RAW_S, 3 * RAW_S + 1, 2 * RAW_S + 2, RAW_S + 3, 3, -RAW_S + 3,
                         -2 * RAW_S + 2, -3 * RAW_S + 1, -3 * RAW_S, -3 * RAW_S - 1,
                         -2 * RAW_S - 2, -RAW_S - 3, -3, RAW_S - 3, 2 * RAW_S - 2, 3 * RAW_S - 1};
    printf("host c=%d:", h[y*RAW_S+x]); for(int k2=0;k2<16;k2++) printf(" %d", h[y*RAW_S+x+off[k2]]); printf("\n");
    k<<<1,1>>>(d,x,y); cudaDeviceSynchronize();
}
