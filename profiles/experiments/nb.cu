// micro-benchmark: n-body inner-loop formulations on B200
#include <cstdio>
#include <cuda_runtime.h>
#include <vector>
#include <cmath>
typedef unsigned long long u64;
#define TILE 256
__device__ __forceinline__ float rsq(float v){ float r; asm("rsqrt.approx.ftz.f32 %0, %1;":"=f"(r):"f"(v)); return r; }
__device__ __forceinline__ u64 pk(float a, float b){ u64 r; asm("mov.b64 %0, {%1,%2};":"=l"(r):"f"(a),"f"(b)); return r;}
__device__ __forceinline__ void upk(u64 v, float&a, float&b){ asm("mov.b64 {%0,%1}, %2;":"=f"(a),"=f"(b):"l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a,u64 b,u64 c){ u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;":"=l"(d):"l"(a),"l"(b),"l"(c)); return d;}
__device__ __forceinline__ u64 mul2(u64 a,u64 b){ u64 d; asm("mul.rn.f32x2 %0, %1, %2;":"=l"(d):"l"(a),"l"(b)); return d;}
__device__ __forceinline__ u64 sub2(u64 a,u64 b){ u64 d; asm("sub.rn.f32x2 %0, %1, %2;":"=l"(d):"l"(a),"l"(b)); return d;}

// A: current kernel (1 body per thread, scalar)
extern "C" __global__ void nA(const float4* pos, float4* acc, int n, float eps2){
  __shared__ float4 tile[TILE];
  int i = blockIdx.x*blockDim.x+threadIdx.x; float4 pi=pos[i]; float ax=0,ay=0,az=0;
  for(int base=0;base<n;base+=TILE){ tile[threadIdx.x]=pos[base+threadIdx.x]; __syncthreads();
#pragma unroll 8
    for(int t=0;t<TILE;t++){ float4 pj=tile[t]; float dx=pj.x-pi.x,dy=pj.y-pi.y,dz=pj.z-pi.z;
      float r2=__fmaf_rn(dz,dz,__fmaf_rn(dy,dy,__fmaf_rn(dx,dx,eps2))); float inv=rsq(r2); float s=pj.w*(inv*inv*inv);
      ax=__fmaf_rn(dx,s,ax); ay=__fmaf_rn(dy,s,ay); az=__fmaf_rn(dz,s,az);} __syncthreads(); }
  acc[i]=make_float4(ax,ay,az,0); }
// B: 2 bodies per thread (i and i+half), scalar
extern "C" __global__ void nB(const float4* pos, float4* acc, int n, float eps2){
  __shared__ float4 tile[TILE];
  int i = blockIdx.x*blockDim.x+threadIdx.x; int i2=i+n/2; float4 p0=pos[i], p1=pos[i2]; float ax=0,ay=0,az=0,bx=0,by=0,bz=0;
  for(int base=0;base<n;base+=TILE){ tile[threadIdx.x]=pos[base+threadIdx.x]; __syncthreads();
#pragma unroll 4
    for(int t=0;t<TILE;t++){ float4 pj=tile[t];
      { float dx=pj.x-p0.x,dy=pj.y-p0.y,dz=pj.z-p0.z; float r2=__fmaf_rn(dz,dz,__fmaf_rn(dy,dy,__fmaf_rn(dx,dx,eps2))); float inv=rsq(r2); float s=pj.w*(inv*inv*inv);
        ax=__fmaf_rn(dx,s,ax); ay=__fmaf_rn(dy,s,ay); az=__fmaf_rn(dz,s,az);} 
      { float dx=pj.x-p1.x,dy=pj.y-p1.y,dz=pj.z-p1.z; float r2=__fmaf_rn(dz,dz,__fmaf_rn(dy,dy,__fmaf_rn(dx,dx,eps2))); float inv=rsq(r2); float s=pj.w*(inv*inv*inv);
        bx=__fmaf_rn(dx,s,bx); by=__fmaf_rn(dy,s,by); bz=__fmaf_rn(dz,s,bz);} }
    __syncthreads(); }
  acc[i]=make_float4(ax,ay,az,0); acc[i2]=make_float4(bx,by,bz,0); }
// D: packed f32x2 over pairs of j
extern "C" __global__ void nD(const float4* pos, float4* acc, int n, float eps2){
  __shared__ float4 sA[TILE/2]; __shared__ float4 sB[TILE/2];
  int i = blockIdx.x*blockDim.x+threadIdx.x; float4 pi=pos[i];
  u64 PIX=pk(pi.x,pi.x), PIY=pk(pi.y,pi.y), PIZ=pk(pi.z,pi.z), EPS=pk(eps2,eps2), AX=pk(0,0),AY=pk(0,0),AZ=pk(0,0);
  float* fA=(float*)sA; float* fB=(float*)sB;
  for(int base=0;base<n;base+=TILE){ float4 pj=pos[base+threadIdx.x]; int p=threadIdx.x>>1,h=threadIdx.x&1;
    fA[p*4+h]=pj.x; fA[p*4+2+h]=pj.y; fB[p*4+h]=pj.z; fB[p*4+2+h]=pj.w; __syncthreads();
#pragma unroll 4
    for(int t=0;t<TILE/2;t++){ float4 a=sA[t], b=sB[t];
      u64 DX=sub2(pk(a.x,a.y),PIX), DY=sub2(pk(a.z,a.w),PIY), DZ=sub2(pk(b.x,b.y),PIZ);
      u64 R2=fma2(DZ,DZ,fma2(DY,DY,fma2(DX,DX,EPS))); float r0,r1; upk(R2,r0,r1);
      u64 INV=pk(rsq(r0),rsq(r1)); u64 S=mul2(pk(b.z,b.w),mul2(mul2(INV,INV),INV));
      AX=fma2(DX,S,AX); AY=fma2(DY,S,AY); AZ=fma2(DZ,S,AZ);} __syncthreads(); }
  float a0,a1,b0,b1,c0,c1; upk(AX,a0,a1); upk(AY,b0,b1); upk(AZ,c0,c1);
  acc[i]=make_float4(a0+a1,b0+b1,c0+c1,0); }
// E: packed over j AND 2 bodies per thread
extern "C" __global__ void nE(const float4* pos, float4* acc, int n, float eps2){
  __shared__ float4 sA[TILE/2]; __shared__ float4 sB[TILE/2];
  int i = blockIdx.x*blockDim.x+threadIdx.x; int i2=i+n/2; float4 pi=pos[i], qi=pos[i2];
  u64 PIX=pk(pi.x,pi.x), PIY=pk(pi.y,pi.y), PIZ=pk(pi.z,pi.z), QIX=pk(qi.x,qi.x), QIY=pk(qi.y,qi.y), QIZ=pk(qi.z,qi.z), EPS=pk(eps2,eps2);
  u64 AX=pk(0,0),AY=pk(0,0),AZ=pk(0,0),BX=pk(0,0),BY=pk(0,0),BZ=pk(0,0);
  float* fA=(float*)sA; float* fB=(float*)sB;
  for(int base=0;base<n;base+=TILE){ float4 pj=pos[base+threadIdx.x]; int p=threadIdx.x>>1,h=threadIdx.x&1;
    fA[p*4+h]=pj.x; fA[p*4+2+h]=pj.y; fB[p*4+h]=pj.z; fB[p*4+2+h]=pj.w; __syncthreads();
#pragma unroll 2
    for(int t=0;t<TILE/2;t++){ float4 a=sA[t], b=sB[t]; u64 PX=pk(a.x,a.y),PY=pk(a.z,a.w),PZ=pk(b.x,b.y),PM=pk(b.z,b.w);
      { u64 DX=sub2(PX,PIX), DY=sub2(PY,PIY), DZ=sub2(PZ,PIZ); u64 R2=fma2(DZ,DZ,fma2(DY,DY,fma2(DX,DX,EPS))); float r0,r1; upk(R2,r0,r1);
        u64 INV=pk(rsq(r0),rsq(r1)); u64 S=mul2(PM,mul2(mul2(INV,INV),INV)); AX=fma2(DX,S,AX); AY=fma2(DY,S,AY); AZ=fma2(DZ,S,AZ);} 
      { u64 DX=sub2(PX,QIX), DY=sub2(PY,QIY), DZ=sub2(PZ,QIZ); u64 R2=fma2(DZ,DZ,fma2(DY,DY,fma2(DX,DX,EPS))); float r0,r1; upk(R2,r0,r1);
        u64 INV=pk(rsq(r0),rsq(r1)); u64 S=mul2(PM,mul2(mul2(INV,INV),INV)); BX=fma2(DX,S,BX); BY=fma2(DY,S,BY); BZ=fma2(DZ,S,BZ);} }
    __syncthreads(); }
  float a0,a1,b0,b1,c0,c1; upk(AX,a0,a1); upk(AY,b0,b1); upk(AZ,c0,c1); acc[i]=make_float4(a0+a1,b0+b1,c0+c1,0);
  upk(BX,a0,a1); upk(BY,b0,b1); upk(BZ,c0,c1); acc[i2]=make_float4(a0+a1,b0+b1,c0+c1,0); }
typedef void(*kern)(const float4*,float4*,int,float);
int main(){
  int n=131072; std::vector<float4> h(n); unsigned s=12345; auto rnd=[&](){ s=s*1664525u+1013904223u; return (s>>8)*(1.0f/16777216.0f)*30.f-15.f; };
  for(auto&p:h){p.x=rnd();p.y=rnd();p.z=rnd();p.w=1.f;}
  float4 *d,*acc; cudaMalloc(&d,n*16); cudaMalloc(&acc,n*16); cudaMemcpy(d,h.data(),n*16,cudaMemcpyHostToDevice);
  int sms,clk; cudaDeviceGetAttribute(&sms,cudaDevAttrMultiProcessorCount,0); cudaDeviceGetAttribute(&clk,cudaDevAttrClockRate,0);
  struct {const char*n; kern k; int per;} ks[]={{"A scalar 1/thr",nA,1},{"B scalar 2/thr",nB,2},{"D packed-j",nD,1},{"E packed-j 2/thr",nE,2}};
  std::vector<float4> ref(n), out(n);
  for(int v=0;v<4;v++){ auto&e=ks[v]; int grid=n/e.per/256;
    e.k<<<grid,256>>>(d,acc,n,1e-4f); cudaDeviceSynchronize();
    cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b); cudaEventRecord(a); e.k<<<grid,256>>>(d,acc,n,1e-4f); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms,a,b); cudaMemcpy(out.data(),acc,n*16,cudaMemcpyDeviceToHost); if(v==0) ref=out;
    double maxrel=0; for(int i=0;i<n;i++){ double dd=fabs(out[i].x-ref[i].x)/fmax(1e-3,fabs(ref[i].x)); if(dd>maxrel)maxrel=dd; }
    double inter=(double)n*n; double ginter=inter/ms/1e6; double frac=ginter*1e9*20/(2.0*128*sms*clk*1e3);
    printf("%-18s %7.3f ms %8.1f GInter/s  frac(20flop) %.3f  slots/inter %.2f  maxrel-vs-A %.2e  %s\n",e.n,ms,ginter,frac, 128.0*sms*clk*1e3/(ginter*1e9), maxrel, cudaGetErrorString(cudaGetLastError())); }
  return 0; }
