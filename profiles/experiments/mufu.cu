#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ float rsq(float v){ float r; asm volatile("rsqrt.approx.ftz.f32 %0, %1;":"=f"(r):"f"(v)); return r; }
__device__ __forceinline__ float ex2(float v){ float r; asm volatile("ex2.approx.ftz.f32 %0, %1;":"=f"(r):"f"(v)); return r; }
__device__ __forceinline__ float rcp(float v){ float r; asm volatile("rcp.approx.ftz.f32 %0, %1;":"=f"(r):"f"(v)); return r; }
__device__ __forceinline__ float lg2(float v){ float r; asm volatile("lg2.approx.ftz.f32 %0, %1;":"=f"(r):"f"(v)); return r; }
#define KERN(name, OP) extern "C" __global__ void name(float* out, int iters){ int i=blockIdx.x*blockDim.x+threadIdx.x; \
  float a=1.5f+i*1e-6f,b=2.5f+i*1e-6f,c=3.5f,d=4.5f,e=5.5f,f=6.5f,g=7.5f,h=8.5f; \
  for(int k=0;k<iters;k++){ a=OP(a);b=OP(b);c=OP(c);d=OP(d);e=OP(e);f=OP(f);g=OP(g);h=OP(h);} out[i]=a+b+c+d+e+f+g+h; }
KERN(kRSQ, rsq) KERN(kEX2, ex2) KERN(kRCP, rcp) KERN(kLG2, lg2)
// mixed: 1 MUFU + 12 FFMA per step (like n-body), independent chains
extern "C" __global__ void kMIX(float* out, int iters){ int i=blockIdx.x*blockDim.x+threadIdx.x;
  float a=1.5f+i*1e-6f, x0=a,x1=a+1,x2=a+2,x3=a+3,x4=a+4,x5=a+5,x6=a+6,x7=a+7,x8=a+8,x9=a+9,x10=a+10,x11=a+11; float c1=0.999f,c2=0.001f;
  for(int k=0;k<iters;k++){ a=rsq(a)+1.0f; // 1 MUFU + 1 FADD
    x0=__fmaf_rn(x0,c1,c2);x1=__fmaf_rn(x1,c1,c2);x2=__fmaf_rn(x2,c1,c2);x3=__fmaf_rn(x3,c1,c2);x4=__fmaf_rn(x4,c1,c2);x5=__fmaf_rn(x5,c1,c2);
    x6=__fmaf_rn(x6,c1,c2);x7=__fmaf_rn(x7,c1,c2);x8=__fmaf_rn(x8,c1,c2);x9=__fmaf_rn(x9,c1,c2);x10=__fmaf_rn(x10,c1,c2); }
  out[i]=a+x0+x1+x2+x3+x4+x5+x6+x7+x8+x9+x10+x11; }
typedef void(*kern)(float*,int);
int main(){ int sms,clk; cudaDeviceGetAttribute(&sms,cudaDevAttrMultiProcessorCount,0); cudaDeviceGetAttribute(&clk,cudaDevAttrClockRate,0);
  float* out; int grid=sms*8, th=256; cudaMalloc(&out,grid*th*4);
  struct{const char*n;kern k;int ops;} ks[]={{"MUFU.RSQ",kRSQ,8},{"MUFU.EX2",kEX2,8},{"MUFU.RCP",kRCP,8},{"MUFU.LG2",kLG2,8},{"1 RSQ+12 FP",kMIX,1}};
  for(auto&e:ks){ e.k<<<grid,th>>>(out,100); cudaDeviceSynchronize(); cudaEvent_t a,b; cudaEventCreate(&a);cudaEventCreate(&b);
    int iters=20000; cudaEventRecord(a); e.k<<<grid,th>>>(out,iters); cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms,a,b);
    double ops=(double)grid*th*iters*e.ops; double per=ops/(ms*1e-3)/(sms*(double)clk*1e3);
    printf("%-12s %8.3f ms  ops/SM/clk %6.2f   cycles per warp-op per SMSP %6.2f\n",e.n,ms,per,32.0/(per/4) ); }
  return 0; }
