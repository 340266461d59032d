#include <stdio.h>
__global__ void k(const int* in, int* out, int n){
    int i = blockIdx.x*blockDim.x+threadIdx.x; if(i>=n) return;
    int a=in[3*i],b=in[3*i+1],c=in[3*i+2];
    out[4*i]=min(a,min(b,c)); out[4*i+1]=max(a,max(b,c));
    out[4*i+2]=max(min(a,b),-max(b,c)); out[4*i+3]=min(max(a,b),c);
}
int main(){
    const int n=1<<12; int h[3*n]; srand(3); for(int i=0;i<3*n;i++) h[i]=rand()%511-255;
    int *d,*o; cudaMalloc(&d,sizeof(h)); cudaMalloc(&o,n*16); cudaMemcpy(d,h,sizeof(h),cudaMemcpyHostToDevice);
    k<<<n/128,128>>>(d,o,n); static int ho[4*n]; cudaMemcpy(ho,o,n*16,cudaMemcpyDeviceToHost);
    int bad[4]={0,0,0,0};
    for(int i=0;i<n;i++){int a=h[3*i],b=h[3*i+1],c=h[3*i+2];
      int r[4]={std::min(a,std::min(b,c)),std::max(a,std::max(b,c)),std::max(std::min(a,b),-std::max(b,c)),std::min(std::max(a,b),c)};
      for(int j=0;j<4;j++) if(r[j]!=ho[4*i+j]){ if(bad[j]<3) printf("j%d a%d b%d c%d host %d dev %d\n",j,a,b,c,r[j],ho[4*i+j]); bad[j]++; }}
    printf("bad %d %d %d %d\n",bad[0],bad[1],bad[2],bad[3]);
}
