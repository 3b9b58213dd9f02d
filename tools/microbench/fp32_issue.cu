// microbenchmark: issue/pipe throughput of scalar FMUL/FADD/FFMA vs packed FFMA2 on sm_100a
#include <cuda_runtime.h>
#include <cstdio>
typedef unsigned long long u64;
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c){ u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
template<int MODE> __global__ void k(float* out, int iters, float s, float t){
  float a[8]; u64 p[8];
  for(int i=0;i<8;i++){ a[i]=threadIdx.x*0.001f+i; p[i]=((u64)__float_as_uint(a[i])<<32)|__float_as_uint(a[i]+1.f); }
  u64 ps = ((u64)__float_as_uint(s)<<32)|__float_as_uint(s), pt=((u64)__float_as_uint(t)<<32)|__float_as_uint(t);
  for(int it=0; it<iters; ++it){
    #pragma unroll
    for(int i=0;i<8;i++){
      if(MODE==0){ a[i]=__fmul_rn(a[i],s); a[i]=__fadd_rn(a[i],t); }         // FMUL+FADD : 2 instr, 2 flop
      if(MODE==1){ a[i]=__fmaf_rn(a[i],s,t); a[i]=__fmaf_rn(a[i],t,s);}        // 2 FFMA
      if(MODE==2){ p[i]=fma2(p[i],ps,pt); p[i]=fma2(p[i],pt,ps);}              // 2 FFMA2 (4 lanes-ops)
      if(MODE==3){ a[i]=__fmul_rn(a[i],s); a[i]=__fmul_rn(a[i],t);}           // 2 FMUL
      if(MODE==4){ a[i]=__fadd_rn(a[i],s); a[i]=__fadd_rn(a[i],t);}           // 2 FADD
    }
  }
  float r=0; for(int i=0;i<8;i++){ r+=a[i]+__uint_as_float((unsigned)p[i])+__uint_as_float((unsigned)(p[i]>>32)); }
  out[blockIdx.x*blockDim.x+threadIdx.x]=r;
}
template<int MODE> void run(const char* name){
  float* d; cudaMalloc(&d, 148*8*256*4);
  cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  int iters=20000; k<MODE><<<148*8,256>>>(d,100,1.0001f,0.0001f);
  cudaEventRecord(e0); k<MODE><<<148*8,256>>>(d,iters,1.0001f,0.0001f); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms,e0,e1);
  double instr = (double)148*8*256/32 * iters * 16;   // warp-instructions
  printf("%-12s %.3f ms  %.2f warp-instr/clk/SM (at 1.9GHz)\n", name, ms, instr/(ms*1e-3)/148/1.9e9);
  cudaFree(d);
}
int main(){ run<0>("FMUL+FADD"); run<1>("FFMA"); run<2>("FFMA2"); run<3>("FMUL"); run<4>("FADD"); return 0; }
