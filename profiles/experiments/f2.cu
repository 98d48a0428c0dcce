#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b){ u64 r; asm("mov.b64 %0, {%1,%2};":"=l"(r):"f"(a),"f"(b)); return r;}
__device__ __forceinline__ void upk(u64 v, float&a, float&b){ asm("mov.b64 {%0,%1}, %2;":"=f"(a),"=f"(b):"l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a,u64 b,u64 c){ u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;":"=l"(d):"l"(a),"l"(b),"l"(c)); return d;}
__device__ __forceinline__ u64 mul2(u64 a,u64 b){ u64 d; asm volatile("mul.rn.f32x2 %0, %1, %2;":"=l"(d):"l"(a),"l"(b)); return d;}
__device__ __forceinline__ u64 add2(u64 a,u64 b){ u64 d; asm volatile("add.rn.f32x2 %0, %1, %2;":"=l"(d):"l"(a),"l"(b)); return d;}
#define INIT int i=blockIdx.x*blockDim.x+threadIdx.x; float f=1.0f+i*1e-7f; u64 a0=pk(f,f+1),a1=pk(f+2,f+3),a2=pk(f+4,f+5),a3=pk(f+6,f+7),a4=pk(f+8,f+9),a5=pk(f+10,f+11),a6=pk(f+12,f+13),a7=pk(f+14,f+15); u64 c1=pk(0.999f,0.998f), c2=pk(0.001f,0.002f);
#define FIN float x,y,s=0; upk(a0,x,y);s+=x+y; upk(a1,x,y);s+=x+y; upk(a2,x,y);s+=x+y; upk(a3,x,y);s+=x+y; upk(a4,x,y);s+=x+y; upk(a5,x,y);s+=x+y; upk(a6,x,y);s+=x+y; upk(a7,x,y);s+=x+y; out[i]=s;
// 1: fma2(a,a,c2): 2 distinct pairs
extern "C" __global__ void k1(float* out,int it){ INIT for(int k=0;k<it;k++){ a0=fma2(a0,a0,c2);a1=fma2(a1,a1,c2);a2=fma2(a2,a2,c2);a3=fma2(a3,a3,c2);a4=fma2(a4,a4,c2);a5=fma2(a5,a5,c2);a6=fma2(a6,a6,c2);a7=fma2(a7,a7,c2);} FIN }
// 2: fma2(a,c1,c2): 3 distinct, two shared across instrs
extern "C" __global__ void k2(float* out,int it){ INIT for(int k=0;k<it;k++){ a0=fma2(a0,c1,c2);a1=fma2(a1,c1,c2);a2=fma2(a2,c1,c2);a3=fma2(a3,c1,c2);a4=fma2(a4,c1,c2);a5=fma2(a5,c1,c2);a6=fma2(a6,c1,c2);a7=fma2(a7,c1,c2);} FIN }
// 3: fma2(a_k, a_{k+1}, a_{k+2}) 3 distinct changing
extern "C" __global__ void k3(float* out,int it){ INIT for(int k=0;k<it;k++){ a0=fma2(a1,a2,a0);a1=fma2(a2,a3,a1);a2=fma2(a3,a4,a2);a3=fma2(a4,a5,a3);a4=fma2(a5,a6,a4);a5=fma2(a6,a7,a5);a6=fma2(a7,a0,a6);a7=fma2(a0,a1,a7);} FIN }
// 4: mul2(a,c1): 2 distinct
extern "C" __global__ void k4(float* out,int it){ INIT for(int k=0;k<it;k++){ a0=mul2(a0,c1);a1=mul2(a1,c1);a2=mul2(a2,c1);a3=mul2(a3,c1);a4=mul2(a4,c1);a5=mul2(a5,c1);a6=mul2(a6,c1);a7=mul2(a7,c1);} FIN }
// 5: add2(a,c2)
extern "C" __global__ void k5(float* out,int it){ INIT for(int k=0;k<it;k++){ a0=add2(a0,c2);a1=add2(a1,c2);a2=add2(a2,c2);a3=add2(a3,c2);a4=add2(a4,c2);a5=add2(a5,c2);a6=add2(a6,c2);a7=add2(a7,c2);} FIN }
// 6: acc pattern: A_k = fma2(D_k, S, A_k) with S shared across 3, D changing
extern "C" __global__ void k6(float* out,int it){ INIT for(int k=0;k<it;k++){ a0=fma2(a3,a6,a0);a1=fma2(a4,a6,a1);a2=fma2(a5,a6,a2); a0=fma2(a3,a7,a0);a1=fma2(a4,a7,a1);a2=fma2(a5,a7,a2); a3=add2(a3,c2); a4=add2(a4,c2);} FIN }
typedef void(*kern)(float*,int);
int main(){ int sms,clk; cudaDeviceGetAttribute(&sms,cudaDevAttrMultiProcessorCount,0); cudaDeviceGetAttribute(&clk,cudaDevAttrClockRate,0);
  float* out; int grid=sms*8, th=256; cudaMalloc(&out,grid*th*4);
  struct{const char*n;kern k;int ops;} ks[]={{"fma2(a,a,c)",k1,8},{"fma2(a,c1,c2)",k2,8},{"fma2 3 distinct",k3,8},{"mul2(a,c)",k4,8},{"add2(a,c)",k5,8},{"acc pattern",k6,8}};
  for(auto&e:ks){ e.k<<<grid,th>>>(out,100); cudaDeviceSynchronize(); cudaEvent_t a,b; cudaEventCreate(&a);cudaEventCreate(&b);
    int iters=20000; cudaEventRecord(a); e.k<<<grid,th>>>(out,iters); cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms,a,b);
    double ops=(double)grid*th*iters*e.ops; double per=ops/(ms*1e-3)/(sms*(double)clk*1e3);
    printf("%-16s %8.3f ms  packed-ops/SM/clk %6.2f   cycles per packed warp-op per SMSP %5.2f\n",e.n,ms,per,32.0/(per/4)); }
  return 0; }
