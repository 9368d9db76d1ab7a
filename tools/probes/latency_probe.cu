// DEV: dependent-chain latencies on this GPU (one warp, clock64 around 512 dependent operations): DFMA, DMUL, DADD, SHFL(64-bit), LDS.64 pointer chase,
// DFMA fed by a shuffle (the triangular-solve step), 1/x and sqrt in fp64.   nvcc -arch=sm_100a -O3 -o /tmp/latency_probe latency_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
#define N 512
__global__ void probe(double* out,long long* cyc,double a,double b){
    __shared__ double sm[64]; __shared__ int nxt[64];
    const int lane=threadIdx.x; sm[lane]=a+lane; sm[lane+32]=b; nxt[lane]=(lane+1)&31; nxt[lane+32]=0; __syncwarp();
    double x=a+lane*1e-9; long long t0,t1;
    t0=clock64();
    #pragma unroll 16
    for(int i=0;i<N;i++) x=fma(x,b,a);
    t1=clock64(); if(lane==0) cyc[0]=t1-t0;
    t0=clock64();
    #pragma unroll 16
    for(int i=0;i<N;i++) x=x*b;
    t1=clock64(); if(lane==0) cyc[1]=t1-t0;
    t0=clock64();
    #pragma unroll 16
    for(int i=0;i<N;i++) x=x+a;
    t1=clock64(); if(lane==0) cyc[2]=t1-t0;
    t0=clock64();
    #pragma unroll 16
    for(int i=0;i<N;i++) x=__shfl_sync(0xffffffffu,x,(i+1)&31);
    t1=clock64(); if(lane==0) cyc[3]=t1-t0;
    int p=lane;
    t0=clock64();
    #pragma unroll 16
    for(int i=0;i<N;i++) p=nxt[p];
    t1=clock64(); if(lane==0) cyc[4]=t1-t0;
    x+=p;
    t0=clock64();
    #pragma unroll 16
    for(int i=0;i<N;i++){ const double ak=__shfl_sync(0xffffffffu,x*b,i&31); x=fma(-a,ak,x); }
    t1=clock64(); if(lane==0) cyc[5]=t1-t0;
    t0=clock64();
    #pragma unroll 4
    for(int i=0;i<64;i++) x=1.0/x+a;
    t1=clock64(); if(lane==0) cyc[6]=t1-t0;
    t0=clock64();
    #pragma unroll 4
    for(int i=0;i<64;i++) x=sqrt(x)+a;
    t1=clock64(); if(lane==0) cyc[7]=t1-t0;
    // dot product the way Qt_times runs it: unroll 4, two shared loads per term
    double s=0.0;
    t0=clock64();
    #pragma unroll 4
    for(int i=0;i<N;i++) s+=sm[(i+lane)&31]*sm[32+(i&31)];
    t1=clock64(); if(lane==0) cyc[8]=t1-t0;
    t0=clock64();
    #pragma unroll 16
    for(int i=0;i<N;i++) s+=sm[(i+lane)&31]*sm[32+(i&31)];
    t1=clock64(); if(lane==0) cyc[9]=t1-t0;
    out[lane]=x+s;
}
int main(){
    double* out; long long* cyc; cudaMalloc(&out,32*8); cudaMalloc(&cyc,16*8);
    for(int w=0;w<2;w++) probe<<<1,32>>>(out,cyc,1.0000001,0.9999999);
    long long h[16]; cudaMemcpy(h,cyc,16*8,cudaMemcpyDeviceToHost);
    const char* nm[]={"DFMA chain","DMUL chain","DADD chain","SHFL(f64) chain","LDS pointer chase","DMUL+SHFL+DFMA (trisolve step)","1/x + add","sqrt + add","dot, unroll 4","dot, unroll 16"};
    for(int i=0;i<10;i++) printf("%-32s %7.1f cycles per link\n",nm[i],(double)h[i]/((i==6||i==7)?64:N));
    return cudaGetLastError()!=cudaSuccess;
}
