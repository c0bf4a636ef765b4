// FP32-pipe microbenchmark for B200 (sm_100a): scalar FFMA vs packed FFMA2 issue/throughput,
// and the mixed FFMA2 + SHF + broadcast-LDS pattern of the ray-sphere cull loop.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp32_peak fp32_peak.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b){ u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk(u64 v, float& a, float& b){ asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c){ u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }

template<int CH> __global__ void k_ffma(float* out, float a, float b, int iters){
  float acc[CH];
  #pragma unroll
  for(int i=0;i<CH;i++) acc[i] = threadIdx.x*0.001f + i;
  float x = a + threadIdx.x*1e-6f, y = b;
  for(int it=0; it<iters; ++it){
    #pragma unroll
    for(int u=0;u<8;u++){
      #pragma unroll
      for(int i=0;i<CH;i++) acc[i] = fmaf(acc[i], x, y);
    }
  }
  float s=0; 
  #pragma unroll
  for(int i=0;i<CH;i++) s+=acc[i];
  out[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
template<int CH> __global__ void k_ffma2(float* out, float a, float b, int iters){
  u64 acc[CH];
  #pragma unroll
  for(int i=0;i<CH;i++) acc[i] = pk(threadIdx.x*0.001f + i, i*0.5f);
  u64 x = pk(a + threadIdx.x*1e-6f, a), y = pk(b, b+1e-3f);
  for(int it=0; it<iters; ++it){
    #pragma unroll
    for(int u=0;u<8;u++){
      #pragma unroll
      for(int i=0;i<CH;i++) acc[i] = ffma2(acc[i], x, y);
    }
  }
  float s=0; 
  #pragma unroll
  for(int i=0;i<CH;i++){ float lo,hi; upk(acc[i],lo,hi); s+=lo+hi; }
  out[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
// cull-loop pattern: per sphere pair: 2 LDS.128 + 1 LDS.64 (broadcast), 10 FFMA2-class ops, 4 SHF
__global__ void k_mix(float* out, const float* sph, int npairs, int iters, float ox, float oy, float oz){
  extern __shared__ float sm[];
  for(int i=threadIdx.x;i<npairs*12;i+=blockDim.x) sm[i]=sph[i];
  __syncthreads();
  float t = threadIdx.x*1e-3f;
  u64 m2ox=pk(ox+t,ox+t), m2oy=pk(oy+t,oy+t), m2oz=pk(oz+t,oz+t), oo=pk(t,t), dx=pk(0.6f+t,0.6f+t), dy=pk(0.5f,0.5f), dz=pk(0.3f-t,0.3f-t), od=pk(1.f+t,1.f+t), oL=pk(2.f+t,2.f+t);
  u64 one = pk(1.f,1.f), mone = pk(-1.f,-1.f);
  unsigned mc=0, ms=0, tot=0;
  for(int it=0; it<iters; ++it){
    #pragma unroll 4
    for(int p=0;p<npairs;p++){
      const float4 A = *reinterpret_cast<const float4*>(sm + p*12);      // cx0 cx1 cy0 cy1
      const float4 B = *reinterpret_cast<const float4*>(sm + p*12 + 4);  // cz0 cz1 w0 w1
      const float2 C = *reinterpret_cast<const float2*>(sm + p*12 + 8);  // cl0 cl1
      u64 cx=pk(A.x,A.y), cy=pk(A.z,A.w), cz=pk(B.x,B.y), w=pk(B.z,B.w), cl=pk(C.x,C.y);
      u64 c = ffma2(w, one, oo);            // FADD2
      c = ffma2(m2oz, cz, c); c = ffma2(m2oy, cy, c); c = ffma2(m2ox, cx, c);
      u64 b1 = ffma2(dz, cz, od); b1 = ffma2(dy, cy, b1); b1 = ffma2(dx, cx, b1);
      u64 nc = ffma2(c, mone, pk(0.f,0.f));   // stand-in (neg folded in real code)
      u64 d1 = ffma2(b1, b1, nc);
      u64 b2 = ffma2(cl, mone, oL);
      u64 d2 = ffma2(b2, b2, nc);
      unsigned lo,hi;
      asm("mov.b64 {%0,%1}, %2;" : "=r"(lo), "=r"(hi) : "l"(d1));
      mc = __funnelshift_l(lo, mc, 1); mc = __funnelshift_l(hi, mc, 1);
      asm("mov.b64 {%0,%1}, %2;" : "=r"(lo), "=r"(hi) : "l"(d2));
      ms = __funnelshift_l(lo, ms, 1); ms = __funnelshift_l(hi, ms, 1);
      if ((p & 15) == 15) { tot += __popc(mc) + __popc(ms); }
    }
  }
  out[blockIdx.x*blockDim.x+threadIdx.x]=__uint_as_float(tot + mc + ms);
}
template<class F> float timeit(F f){ cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1); f(); cudaDeviceSynchronize(); cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms,e0,e1); return ms; }
int main(){
  cudaDeviceProp p; cudaGetDeviceProperties(&p,0); int sms=p.multiProcessorCount; int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("device %s SMs %d clockRate %d kHz\n", p.name, sms, clk);
  float* out; cudaMalloc(&out, sizeof(float)*sms*8*1024);
  const int iters=20000;
  for(int threads : {128,256,512,1024}) for(int bps : {1,2}) {
    if (threads*bps>2048) continue;
    int grid=sms*bps;
    { float ms=timeit([&]{k_ffma<8><<<grid,threads>>>(out,1.0001f,0.5f,iters);}); double fma=(double)grid*threads*iters*8.0*8; printf("FFMA  ch8  thr %4d bps %d: %.3f ms  %.2f TFLOP/s  (%.1f lane-FMA/clk/SM @1965)\n",threads,bps,ms,2*fma/ms*1e-9, fma/(ms*1e-3)/sms/1.965e9); }
    { float ms=timeit([&]{k_ffma2<8><<<grid,threads>>>(out,1.0001f,0.5f,iters);}); double fma=(double)grid*threads*iters*8.0*8*2; printf("FFMA2 ch8  thr %4d bps %d: %.3f ms  %.2f TFLOP/s  (%.1f lane-FMA/clk/SM @1965)\n",threads,bps,ms,2*fma/ms*1e-9, fma/(ms*1e-3)/sms/1.965e9); }
  }
  { int threads=512, grid=sms; float ms=timeit([&]{k_ffma<4><<<grid,threads>>>(out,1.0001f,0.5f,iters);}); double fma=(double)grid*threads*iters*8.0*4; printf("FFMA  ch4  thr 512: %.2f TFLOP/s\n",2*fma/ms*1e-9);
    ms=timeit([&]{k_ffma2<4><<<grid,threads>>>(out,1.0001f,0.5f,iters);}); fma=(double)grid*threads*iters*8.0*4*2; printf("FFMA2 ch4  thr 512: %.2f TFLOP/s\n",2*fma/ms*1e-9);
    ms=timeit([&]{k_ffma<16><<<grid,threads>>>(out,1.0001f,0.5f,iters);}); fma=(double)grid*threads*iters*8.0*16; printf("FFMA  ch16 thr 512: %.2f TFLOP/s\n",2*fma/ms*1e-9);
    ms=timeit([&]{k_ffma2<16><<<grid,threads>>>(out,1.0001f,0.5f,iters);}); fma=(double)grid*threads*iters*8.0*16*2; printf("FFMA2 ch16 thr 512: %.2f TFLOP/s\n",2*fma/ms*1e-9); }
  // mixed pattern
  int npairs=256; float* sph; cudaMalloc(&sph, npairs*12*sizeof(float)); { float* h=(float*)malloc(npairs*12*4); for(int i=0;i<npairs*12;i++) h[i]=(i%7)*0.37f-1.f; cudaMemcpy(sph,h,npairs*48,cudaMemcpyHostToDevice);} 
  for(int threads : {128,256,512}) for(int bps : {1,2,3,4}) { if (threads*bps>2048) continue; int grid=sms*bps; int it=200;
    float ms=timeit([&]{k_mix<<<grid,threads,npairs*48>>>(out,sph,npairs,it,0.1f,0.2f,0.3f);}); double tests=(double)grid*threads*it*npairs*2*2; // 2 spheres x 2 queries
    {cudaError_t e=cudaGetLastError(); if(e) printf("ERR %s\n",cudaGetErrorString(e));} printf("MIX thr %4d bps %d: %.3f ms  %.3e tests/s  -> %.2f clk/test/lane-slot @1965 (per SMSP-lane)\n",threads,bps,ms,tests/(ms*1e-3), (double)sms*128*1.965e9/(tests/(ms*1e-3))); }
  return 0;
}
