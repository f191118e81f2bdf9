// micro-benchmark: MUFU.EX2 throughput per SM vs warps per scheduler, alone and mixed with FADD/F2FP
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
__device__ __forceinline__ float ex2(float x){float y; asm volatile("ex2.approx.ftz.f32 %0, %1;":"=f"(y):"f"(x)); return y;}
template<int MODE>
__global__ void k(float* out, int iters, long long* cyc){
  float a[16];
  for(int i=0;i<16;++i) a[i]=threadIdx.x*0.001f+i*0.01f;
  float s0=0,s1=0; unsigned pk=0;
  long long t0=clock64();
  for(int it=0;it<iters;++it){
#pragma unroll
    for(int i=0;i<16;++i){
      if(MODE==0){ a[i]=ex2(a[i]); }
      else if(MODE==1){ float p=ex2(a[i]-1.0f); s0+=p; a[i]=p*0.5f; }
      else { float p0=ex2(a[i]-1.0f); s0+=p0; if(i&1){ __nv_bfloat162 h=__floats2bfloat162_rn(p0,a[i-1]); pk^=*reinterpret_cast<unsigned*>(&h);} a[i]=a[i]*0.999f+0.001f; }
    }
  }
  long long t1=clock64();
  float r=s0+s1+__uint_as_float(pk&0xff); for(int i=0;i<16;++i) r+=a[i];
  out[blockIdx.x*blockDim.x+threadIdx.x]=r;
  if(threadIdx.x==0) cyc[blockIdx.x]=t1-t0;
}
int main(){
  float* out; long long* cyc; cudaMalloc(&out,148*1024*4); cudaMalloc(&cyc,148*8);
  const int iters=2000;
  for(int mode=0;mode<3;++mode) for(int warps: {4,8,16,32}){
    for(int rep=0;rep<2;++rep){
      if(mode==0) k<0><<<148,warps*32>>>(out,iters,cyc); else if(mode==1) k<1><<<148,warps*32>>>(out,iters,cyc); else k<2><<<148,warps*32>>>(out,iters,cyc);
      cudaDeviceSynchronize();
    }
    long long h[148]; cudaMemcpy(h,cyc,148*8,cudaMemcpyDeviceToHost);
    double c=h[0]; double mufu=(double)iters*16*warps;  // warp-level MUFU per SM
    printf("mode %d warps/SM %2d: %.0f cycles, %.2f cycles per warp-MUFU per SM (=> %.1f lanes/clk/SM)\n",mode,warps,c,c/mufu,32.0*mufu/c);
  }
  return 0;
}
