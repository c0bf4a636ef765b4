// Cull-loop microbenchmark: R rays per thread share each sphere-pair load (LDS traffic per test ~ 1/R).
// Finds whether the fused closest+shadow cull loop is bound by the FMA pipe or by the shared-memory data pipe.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b){ u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk(u64 v, unsigned& lo, unsigned& hi){ asm("mov.b64 {%0,%1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v)); }
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c){ u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 fadd2(u64 a, u64 b){ u64 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
struct Ray { float ox2, oy2, oz2, ndx, ndy, ndz, noo, od, oL; unsigned mc, ms; };
template<int R, int COMBINED> __global__ void __launch_bounds__(512) k(float* out, const float* sph, int npairs, int iters){
  extern __shared__ __align__(16) float sm[];
  float4* sA = (float4*)sm; float4* sB = sA + npairs; float2* sC = (float2*)(sB + npairs);
  for(int i=threadIdx.x;i<npairs*10;i+=blockDim.x) sm[i]=sph[i];
  __syncthreads();
  Ray r[R]; unsigned tot=0;
  #pragma unroll
  for(int j=0;j<R;j++){ float t = threadIdx.x*1e-3f + j*0.37f; r[j] = Ray{0.1f+t,0.2f+t,0.3f+t,0.6f+t,0.5f,0.3f-t,-t,1.f+t,2.f+t,0u,0u}; }
  for(int it=0; it<iters; ++it){
    for(int g=0; g<npairs/16; ++g){
      #pragma unroll
      for(int j=0;j<R;j++){ r[j].mc=0; r[j].ms=0; }
      #pragma unroll
      for(int q=0;q<16;q++){
        const float4 A=sA[g*16+q], B=sB[g*16+q]; const float2 C=sC[g*16+q];
        const u64 cx=pk(A.x,A.y), cy=pk(A.z,A.w), cz=pk(B.x,B.y), w=pk(B.z,B.w), cl=pk(C.x,C.y);
        #pragma unroll
        for(int j=0;j<R;j++){
          u64 c = fadd2(w, pk(r[j].noo,r[j].noo));
          c = ffma2(pk(r[j].oz2,r[j].oz2), cz, c); c = ffma2(pk(r[j].oy2,r[j].oy2), cy, c); c = ffma2(pk(r[j].ox2,r[j].ox2), cx, c);
          u64 b1 = ffma2(pk(r[j].ndz,r[j].ndz), cz, pk(r[j].od,r[j].od)); b1 = ffma2(pk(r[j].ndy,r[j].ndy), cy, b1); b1 = ffma2(pk(r[j].ndx,r[j].ndx), cx, b1);
          u64 b2 = fadd2(cl, pk(r[j].oL,r[j].oL));
          unsigned lo,hi;
          if (COMBINED) {
            unsigned b1l,b1h,b2l,b2h; upk(b1,b1l,b1h); upk(b2,b2l,b2h);
            u64 e = pk(fmaxf(fabsf(__uint_as_float(b1l)),fabsf(__uint_as_float(b2l))), fmaxf(fabsf(__uint_as_float(b1h)),fabsf(__uint_as_float(b2h))));
            u64 d = ffma2(e,e,c); upk(d,lo,hi);
            r[j].mc = __funnelshift_l(lo, r[j].mc, 1); r[j].mc = __funnelshift_l(hi, r[j].mc, 1);
          } else {
            u64 d1 = ffma2(b1,b1,c), d2 = ffma2(b2,b2,c);
            upk(d1,lo,hi); r[j].mc = __funnelshift_l(lo, r[j].mc, 1); r[j].mc = __funnelshift_l(hi, r[j].mc, 1);
            upk(d2,lo,hi); r[j].ms = __funnelshift_l(lo, r[j].ms, 1); r[j].ms = __funnelshift_l(hi, r[j].ms, 1);
          }
        }
      }
      #pragma unroll
      for(int j=0;j<R;j++){ tot += __popc(r[j].mc) + __popc(r[j].ms); r[j].noo += 1e-6f * (float)(tot & 1); }
    }
  }
  out[blockIdx.x*blockDim.x+threadIdx.x]=__uint_as_float(tot);
}
template<class F> float timeit(F f){ cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1); f(); cudaDeviceSynchronize(); cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms,e0,e1); return ms; }
int main(){
  cudaDeviceProp p; cudaGetDeviceProperties(&p,0); int sms=p.multiProcessorCount;
  float* out; cudaMalloc(&out, sizeof(float)*sms*8*1024);
  int npairs=256; float* sph; cudaMalloc(&sph, npairs*10*sizeof(float)); { float* h=(float*)malloc(npairs*40); for(int i=0;i<npairs*10;i++) h[i]=(i%7)*0.37f-1.f; cudaMemcpy(sph,h,npairs*40,cudaMemcpyHostToDevice);}
  int it=100;
  #define RUN(R,C) for(int threads : {128,256,512}) for(int bps : {1,2,4}) { if (threads*bps>1024) continue; int grid=sms*bps; \
    float ms=timeit([&]{k<R,C><<<grid,threads,npairs*40>>>(out,sph,npairs,it);}); double tests=(double)grid*threads*it*npairs*2*2*R; \
    cudaError_t e=cudaGetLastError(); if(e) printf("ERR %s\n",cudaGetErrorString(e)); \
    printf("R %d comb %d thr %4d bps %d: %.3f ms  %.3e slot-tests/s  -> %.2f lane-clk/test @1965\n",R,C,threads,bps,ms,tests/(ms*1e-3), (double)sms*128*1.965e9/(tests/(ms*1e-3))); }
  RUN(1,0) RUN(2,0) RUN(3,0) RUN(4,0) RUN(1,1) RUN(2,1)
  return 0;
}
