// Variants of the fused closest+shadow cull loop to find the binding limit (issue vs FMA pipe vs LSU).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b){ u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c){ u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 fadd2(u64 a, u64 b){ u64 d; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
// MODE 0: packed FFMA2 + SHF; 1: packed, no SHF (xor accumulate every 8th); 2: scalar FFMA + SHF; 3: scalar no SHF
// 4: packed, combined mask via max(|b1|,|b2|); 5: scalar with FSETP+predicated-or instead of SHF
template<int MODE> __global__ void __launch_bounds__(512) k_mix(float* out, const float* sph, int npairs, int iters, float ox, float oy, float oz){
  extern __shared__ __align__(16) float sm[];
  for(int i=threadIdx.x;i<npairs*12;i+=blockDim.x) sm[i]=sph[i];
  __syncthreads();
  float t = threadIdx.x*1e-3f;
  float m2ox=ox+t, m2oy=oy+t, m2oz=oz+t, oo=t, dx=0.6f+t, dy=0.5f, dz=0.3f-t, od=1.f+t, oL=2.f+t;
  unsigned mc=0, ms=0, tot=0;
  for(int it=0; it<iters; ++it){
    #pragma unroll 4
    for(int p=0;p<npairs;p++){
      const float4 A = *reinterpret_cast<const float4*>(sm + p*12);      // cx0 cx1 cy0 cy1
      const float4 B = *reinterpret_cast<const float4*>(sm + p*12 + 4);  // cz0 cz1 w0 w1
      const float2 C = *reinterpret_cast<const float2*>(sm + p*12 + 8);  // cl0 cl1
      if (MODE==0 || MODE==1 || MODE==4) {
        u64 cx=pk(A.x,A.y), cy=pk(A.z,A.w), cz=pk(B.x,B.y), w=pk(B.z,B.w), cl=pk(C.x,C.y);
        u64 c = fadd2(w, pk(oo,oo));
        c = ffma2(pk(m2oz,m2oz), cz, c); c = ffma2(pk(m2oy,m2oy), cy, c); c = ffma2(pk(m2ox,m2ox), cx, c);   // c = -(c')
        u64 b1 = ffma2(pk(dz,dz), cz, pk(od,od)); b1 = ffma2(pk(dy,dy), cy, b1); b1 = ffma2(pk(dx,dx), cx, b1);
        u64 b2 = fadd2(cl, pk(oL,oL));
        unsigned lo,hi;
        if (MODE==4) {
          float b1l,b1h,b2l,b2h; asm("mov.b64 {%0,%1}, %2;" : "=f"(b1l), "=f"(b1h) : "l"(b1)); asm("mov.b64 {%0,%1}, %2;" : "=f"(b2l), "=f"(b2h) : "l"(b2));
          u64 e = pk(fmaxf(fabsf(b1l),fabsf(b2l)), fmaxf(fabsf(b1h),fabsf(b2h)));
          u64 d = ffma2(e,e,c);
          asm("mov.b64 {%0,%1}, %2;" : "=r"(lo), "=r"(hi) : "l"(d));
          mc = __funnelshift_l(lo, mc, 1); mc = __funnelshift_l(hi, mc, 1);
        } else {
          u64 d1 = ffma2(b1, b1, c);
          u64 d2 = ffma2(b2, b2, c);
          if (MODE==0) {
            asm("mov.b64 {%0,%1}, %2;" : "=r"(lo), "=r"(hi) : "l"(d1));
            mc = __funnelshift_l(lo, mc, 1); mc = __funnelshift_l(hi, mc, 1);
            asm("mov.b64 {%0,%1}, %2;" : "=r"(lo), "=r"(hi) : "l"(d2));
            ms = __funnelshift_l(lo, ms, 1); ms = __funnelshift_l(hi, ms, 1);
          } else {
            unsigned a0,a1,a2,a3;
            asm("mov.b64 {%0,%1}, %2;" : "=r"(a0), "=r"(a1) : "l"(d1)); asm("mov.b64 {%0,%1}, %2;" : "=r"(a2), "=r"(a3) : "l"(d2));
            if ((p&7)==7) { mc ^= a0 ^ a1; ms ^= a2 ^ a3; } else { asm volatile("" :: "r"(a0),"r"(a1),"r"(a2),"r"(a3)); }
          }
        }
      } else {
        float c0 = A.x*0+B.z + oo, c1 = B.w + oo;
        c0 = fmaf(m2oz,B.x,c0); c1 = fmaf(m2oz,B.y,c1); c0 = fmaf(m2oy,A.z,c0); c1 = fmaf(m2oy,A.w,c1); c0 = fmaf(m2ox,A.x,c0); c1 = fmaf(m2ox,A.y,c1);
        float b10 = fmaf(dz,B.x,od), b11 = fmaf(dz,B.y,od); b10 = fmaf(dy,A.z,b10); b11=fmaf(dy,A.w,b11); b10=fmaf(dx,A.x,b10); b11=fmaf(dx,A.y,b11);
        float b20 = C.x + oL, b21 = C.y + oL;
        float d10 = fmaf(b10,b10,c0), d11 = fmaf(b11,b11,c1), d20 = fmaf(b20,b20,c0), d21 = fmaf(b21,b21,c1);
        if (MODE==2) {
          mc = __funnelshift_l(__float_as_uint(d10), mc, 1); mc = __funnelshift_l(__float_as_uint(d11), mc, 1);
          ms = __funnelshift_l(__float_as_uint(d20), ms, 1); ms = __funnelshift_l(__float_as_uint(d21), ms, 1);
        } else if (MODE==5) {
          if (d10 > 0.f) mc |= 1u << ((2*p)&31); if (d11 > 0.f) mc |= 1u << ((2*p+1)&31);
          if (d20 > 0.f) ms |= 1u << ((2*p)&31); if (d21 > 0.f) ms |= 1u << ((2*p+1)&31);
        } else {
          if ((p&7)==7) { mc ^= __float_as_uint(d10) ^ __float_as_uint(d11); ms ^= __float_as_uint(d20) ^ __float_as_uint(d21); } else { asm volatile("" :: "f"(d10),"f"(d11),"f"(d20),"f"(d21)); }
        }
      }
      if ((p & 15) == 15) { tot += __popc(mc) + __popc(ms); }
    }
  }
  out[blockIdx.x*blockDim.x+threadIdx.x]=__uint_as_float(tot + mc + ms);
}
template<class F> float timeit(F f){ cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1); f(); cudaDeviceSynchronize(); cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms,e0,e1); return ms; }
int main(){
  cudaDeviceProp p; cudaGetDeviceProperties(&p,0); int sms=p.multiProcessorCount;
  float* out; cudaMalloc(&out, sizeof(float)*sms*8*1024);
  int npairs=256; float* sph; cudaMalloc(&sph, npairs*12*sizeof(float)); { float* h=(float*)malloc(npairs*12*4); for(int i=0;i<npairs*12;i++) h[i]=(i%7)*0.37f-1.f; cudaMemcpy(sph,h,npairs*48,cudaMemcpyHostToDevice);} 
  int it=200;
  #define RUN(M) for(int threads : {256,512}) for(int bps : {2,4}) { if (threads*bps>2048) continue; int grid=sms*bps; \
    float ms=timeit([&]{k_mix<M><<<grid,threads,npairs*48>>>(out,sph,npairs,it,0.1f,0.2f,0.3f);}); double tests=(double)grid*threads*it*npairs*2*2; \
    cudaError_t e=cudaGetLastError(); if(e) printf("ERR %s\n",cudaGetErrorString(e)); \
    printf("MODE %d thr %4d bps %d: %.3f ms  %.3e tests/s  -> %.2f lane-clk/test @1965\n",M,threads,bps,ms,tests/(ms*1e-3), (double)sms*128*1.965e9/(tests/(ms*1e-3))); }
  RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5)
  return 0;
}
