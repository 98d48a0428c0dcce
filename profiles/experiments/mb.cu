// micro-benchmark: which Mandelbrot inner-loop formulation issues fastest on B200?
#include <cstdio>
#include <cuda_runtime.h>
#define K 16
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b){ u64 r; asm("mov.b64 %0, {%1,%2};":"=l"(r):"f"(a),"f"(b)); return r;}
__device__ __forceinline__ void upk(u64 v, float&a, float&b){ asm("mov.b64 {%0,%1}, %2;":"=f"(a),"=f"(b):"l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a,u64 b,u64 c){ u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;":"=l"(d):"l"(a),"l"(b),"l"(c)); return d;}
__device__ __forceinline__ u64 mul2(u64 a,u64 b){ u64 d; asm("mul.rn.f32x2 %0, %1, %2;":"=l"(d):"l"(a),"l"(b)); return d;}
__device__ __forceinline__ u64 sub2(u64 a,u64 b){ u64 d; asm("sub.rn.f32x2 %0, %1, %2;":"=l"(d):"l"(a),"l"(b)); return d;}

#define PROLOG int i = blockIdx.x*blockDim.x+threadIdx.x; float cx = -0.1f + 1e-6f*(i&1023), cy = 0.1f+1e-6f*(i>>10); float x=0.f,y=0.f;
// A: old textbook form: y' = fma(x+x, y, cy)
extern "C" __global__ void kA(float* out, int blocks){ PROLOG
  for(int b=0;b<blocks;b++){
#pragma unroll
    for(int u=0;u<K;u++){ float yy=__fmul_rn(y,y); float t=__fsub_rn(cx,yy); float x2=__fadd_rn(x,x); y=__fmaf_rn(x2,y,cy); x=__fmaf_rn(x,x,t);} 
    if(!(__fmaf_rn(x,x,y*y)<=4.f)) break; }
  out[i]=x+y; }
// B: new form: w=x*y; y'=fma(2,w,cy)
extern "C" __global__ void kB(float* out, int blocks){ PROLOG
  for(int b=0;b<blocks;b++){
#pragma unroll
    for(int u=0;u<K;u++){ float yy=__fmul_rn(y,y); float w=__fmul_rn(x,y); float t=__fsub_rn(cx,yy); y=__fmaf_rn(2.f,w,cy); x=__fmaf_rn(x,x,t);} 
    if(!(__fmaf_rn(x,x,y*y)<=4.f)) break; }
  out[i]=x+y; }
// C: new form, 2 pixels per thread interleaved (ILP 2)
extern "C" __global__ void kC(float* out, int blocks){ PROLOG float x1=0.f,y1=0.f, cx1=cx+0.01f, cy1=cy-0.01f;
  for(int b=0;b<blocks;b++){
#pragma unroll
    for(int u=0;u<K;u++){ float yy=__fmul_rn(y,y); float w=__fmul_rn(x,y); float t=__fsub_rn(cx,yy); y=__fmaf_rn(2.f,w,cy); x=__fmaf_rn(x,x,t);
                          float yy1=__fmul_rn(y1,y1); float w1=__fmul_rn(x1,y1); float t1=__fsub_rn(cx1,yy1); y1=__fmaf_rn(2.f,w1,cy1); x1=__fmaf_rn(x1,x1,t1);} 
    if(!(__fmaf_rn(x,x,y*y)<=4.f) || !(__fmaf_rn(x1,x1,y1*y1)<=4.f)) break; }
  out[i]=x+y+x1+y1; }
// D: packed f32x2, 2 pixels per thread
extern "C" __global__ void kD(float* out, int blocks){ PROLOG
  u64 X=pk(0.f,0.f), Y=pk(0.f,0.f), CX=pk(cx,cx+0.01f), CY=pk(cy,cy-0.01f), TWO=pk(2.f,2.f);
  for(int b=0;b<blocks;b++){
#pragma unroll
    for(int u=0;u<K;u++){ u64 YY=mul2(Y,Y); u64 Wv=mul2(X,Y); u64 T=sub2(CX,YY); Y=fma2(TWO,Wv,CY); X=fma2(X,X,T);} 
    float a,bq,c,d; upk(X,a,bq); upk(Y,c,d);
    if(!(__fmaf_rn(a,a,c*c)<=4.f) || !(__fmaf_rn(bq,bq,d*d)<=4.f)) break; }
  float a,bq,c,d; upk(X,a,bq); upk(Y,c,d); out[i]=a+bq+c+d; }
// E: pure FFMA peak probe: independent chains, 2 distinct regs per instr
extern "C" __global__ void kE(float* out, int blocks){ PROLOG float a0=x,a1=y,a2=cx,a3=cy,a4=1.f,a5=2.f,a6=3.f,a7=4.f;
  for(int b=0;b<blocks;b++){
#pragma unroll
    for(int u=0;u<K;u++){ a0=__fmaf_rn(a0,a0,cx); a1=__fmaf_rn(a1,a1,cy); a2=__fmaf_rn(a2,a2,cx); a3=__fmaf_rn(a3,a3,cy); a4=__fmaf_rn(a4,a4,cx);} 
  }
  out[i]=a0+a1+a2+a3+a4+a5+a6+a7; }
// F: pure FFMA with 3 distinct regs
extern "C" __global__ void kF(float* out, int blocks){ PROLOG float a0=x,a1=y,a2=cx,a3=cy,a4=1.f;
  for(int b=0;b<blocks;b++){
#pragma unroll
    for(int u=0;u<K;u++){ a0=__fmaf_rn(a0,cx,cy); a1=__fmaf_rn(a1,cx,cy); a2=__fmaf_rn(a2,cx,cy); a3=__fmaf_rn(a3,cx,cy); a4=__fmaf_rn(a4,cx,cy);} 
  }
  out[i]=a0+a1+a2+a3+a4; }

// G: 4-op form: t=fma(-y,y,cx); w=x*y; y'=fma(2,w,cy); x'=fma(x,x,t)   K=16
#define BODY4 { float t=__fmaf_rn(-y,y,cx); float w=__fmul_rn(x,y); y=__fmaf_rn(2.f,w,cy); x=__fmaf_rn(x,x,t);} 
extern "C" __global__ void kG(float* out, int blocks){ PROLOG
  for(int b=0;b<blocks;b++){
#pragma unroll
    for(int u=0;u<16;u++) BODY4
    if(!(__fmaf_rn(x,x,y*y)<=4.f)) break; }
  out[i]=x+y; }
// H: 4-op, K=32 (blocks/2)
extern "C" __global__ void kH(float* out, int blocks){ PROLOG
  for(int b=0;b<blocks/2;b++){
#pragma unroll
    for(int u=0;u<32;u++) BODY4
    if(!(__fmaf_rn(x,x,y*y)<=4.f)) break; }
  out[i]=x+y; }
// J: 4-op ILP2
extern "C" __global__ void kJ(float* out, int blocks){ PROLOG float x1=0.f,y1=0.f, cx1=cx+0.01f, cy1=cy-0.01f;
  for(int b=0;b<blocks;b++){
#pragma unroll
    for(int u=0;u<16;u++){ BODY4 { float t=__fmaf_rn(-y1,y1,cx1); float w=__fmul_rn(x1,y1); y1=__fmaf_rn(2.f,w,cy1); x1=__fmaf_rn(x1,x1,t);} }
    if(!(__fmaf_rn(x,x,y*y)<=4.f) || !(__fmaf_rn(x1,x1,y1*y1)<=4.f)) break; }
  out[i]=x+y+x1+y1; }
// L: 4-op, K=16, keep intermediates for replay (like the production kernel): store x_j,y_j of the block
extern "C" __global__ void kL(float* out, int blocks){ PROLOG int cnt=0;
  for(int b=0;b<blocks;b++){
    float xs[16], ys[16];
#pragma unroll
    for(int u=0;u<16;u++){ xs[u]=x; ys[u]=y; BODY4 }
    if(!(__fmaf_rn(x,x,y*y)<=4.f)) {
#pragma unroll
      for(int u=0;u<16;u++){ if(__fmaf_rn(xs[u],xs[u],ys[u]*ys[u])>4.f) break; cnt++; }
      break; }
    cnt+=16; }
  out[i]=x+y+cnt; }
typedef void (*kern)(float*,int);
int main(){
  int sms=0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  int clk=0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  float* out; int threads=256, grid=sms*8; cudaMalloc(&out, sizeof(float)*threads*grid);
  struct {const char* n; kern k; int pix;} ks[]={{"A textbook",kA,1},{"B 2-reg form",kB,1},{"C 2-reg ILP2",kC,2},{"D f32x2",kD,2},{"E ffma 2reg",kE,1},{"G 4op K16",kG,1},{"H 4op K32",kH,1},{"J 4op ILP2",kJ,2},{"L 4op keep",kL,1},{"F ffma 3reg",kF,1}};
  int blocks=20000;
  for(auto& e: ks){
    e.k<<<grid,threads>>>(out, 100); cudaDeviceSynchronize();
    cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a); e.k<<<grid,threads>>>(out, blocks); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms,a,b);
    double iters=(double)grid*threads*e.pix*blocks*K;
    double perclk = iters/(ms*1e-3)/(sms*(double)clk*1e3); // lane-iterations per SM per clock
    printf("%-14s %8.2f ms  %7.2f Giter/s  lane-iter/SM/clk %6.2f  -> slots/iter %5.2f  frac(8flop) %5.3f  err=%s\n", e.n, ms, iters/ms/1e6, perclk, 128.0/perclk, perclk*8/256.0, cudaGetErrorString(cudaGetLastError()));
  }
  return 0; }
