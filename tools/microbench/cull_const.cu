// Where should the sphere list live for the converged cull loop?  Every lane reads the SAME pair, so besides shared memory
// (LDS, broadcast) the constant bank (uniform loads / constant operands) is a candidate (north_star: "staged in
// shared/constant memory").  SRC 0: shared memory, 1: __constant__ arrays indexed by the (warp-uniform) loop counter.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b){ u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk(u64 v, unsigned& lo, unsigned& hi){ asm("mov.b64 {%0,%1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v)); }
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c){ u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 fadd2(u64 a, u64 b){ u64 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
#define NP 256
__constant__ float4 cA[NP]; __constant__ float4 cB[NP]; __constant__ float2 cC[NP];
template<int SRC> __global__ void __launch_bounds__(256) k(float* out, const float* sph, int npairs, int iters){
  extern __shared__ __align__(16) float sm[];
  float4* sA = (float4*)sm; float4* sB = sA + npairs; float2* sC = (float2*)(sB + npairs);
  if (SRC == 0 || SRC >= 3) { for(int i=threadIdx.x;i<npairs*10;i+=blockDim.x) sm[i]=sph[i]; __syncthreads(); }
  const float t = threadIdx.x*1e-3f;
  float ox2=0.1f+t, oy2=0.2f+t, oz2=0.3f+t, ndx=0.6f+t, ndy=0.5f, ndz=0.3f-t, noo=-t, od=1.f+t, oL=2.f+t;
  unsigned tot=0;
  if (SRC == 3 || SRC == 4) {   // shared memory, every instruction a volatile asm in a hand-written order: the loads of pair q+D are
                                // issued between the FP2 operations of pair q (D = 2 for SRC 3, 4 for SRC 4)
    constexpr int D = SRC == 3 ? 2 : 4;
    const unsigned baseA = (unsigned)__cvta_generic_to_shared(sA), baseB = (unsigned)__cvta_generic_to_shared(sB), baseC = (unsigned)__cvta_generic_to_shared(sC);
    #define VLDS128(v, addr) asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr))
    #define VLDS64(v, addr) asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr))
    #define VFFMA2(d, a, b, c) asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c))
    #define VFADD2(d, a, b) asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b))
    const u64 Pnoo = pk(noo,noo), Poz2 = pk(oz2,oz2), Poy2 = pk(oy2,oy2), Pox2 = pk(ox2,ox2), Pndz = pk(ndz,ndz), Pndy = pk(ndy,ndy), Pndx = pk(ndx,ndx), Pod = pk(od,od), PoL = pk(oL,oL);
    for(int it=0; it<iters; ++it){
      for(int g=0; g<npairs/16; ++g){
        unsigned mc=0, ms=0;
        float4 A[D], B[D]; float2 C[D];
        const unsigned ga = baseA + g * 256, gb = baseB + g * 256, gc = baseC + g * 128;
        #pragma unroll
        for (int q = 0; q < D; ++q) { VLDS128(A[q], ga + q * 16); VLDS128(B[q], gb + q * 16); VLDS64(C[q], gc + q * 8); }
        #pragma unroll
        for(int q=0;q<16;q++){
          const float4 a = A[q % D], b = B[q % D]; const float2 cc = C[q % D];
          const u64 cx=pk(a.x,a.y), cy=pk(a.z,a.w), cz=pk(b.x,b.y), w=pk(b.z,b.w), cl=pk(cc.x,cc.y);
          u64 c, b1, b2, d1, d2;
          VFADD2(c, w, Pnoo); VFFMA2(b1, Pndz, cz, Pod);
          if (q + D < 16) VLDS128(A[q % D], ga + (q + D) * 16);
          VFFMA2(c, Poz2, cz, c); VFFMA2(b1, Pndy, cy, b1);
          if (q + D < 16) VLDS128(B[q % D], gb + (q + D) * 16);
          VFFMA2(c, Poy2, cy, c); VFFMA2(b1, Pndx, cx, b1);
          if (q + D < 16) VLDS64(C[q % D], gc + (q + D) * 8);
          VFFMA2(c, Pox2, cx, c); VFADD2(b2, cl, PoL);
          VFFMA2(d1, b1, b1, c); VFFMA2(d2, b2, b2, c);
          unsigned lo,hi;
          upk(d1,lo,hi); mc = __funnelshift_l(lo, mc, 1); mc = __funnelshift_l(hi, mc, 1);
          upk(d2,lo,hi); ms = __funnelshift_l(lo, ms, 1); ms = __funnelshift_l(hi, ms, 1);
        }
        tot += __popc(mc) + __popc(ms);
      }
    }
    out[blockIdx.x*blockDim.x+threadIdx.x]=__uint_as_float(tot);
    return;
  }
  if (SRC == 2) {            // sphere data held in registers (4 pairs, reused): the FP2 + SHF mix of the loop without any load
    float4 rA[4], rB[4]; float2 rC[4];
    #pragma unroll
    for (int q = 0; q < 4; ++q) { rA[q] = ((const float4*)sph)[q]; rB[q] = ((const float4*)sph)[npairs + q]; rC[q] = ((const float2*)sph)[4 * npairs + q]; }
    for(int it=0; it<iters; ++it){
      for(int g=0; g<npairs/16; ++g){
        unsigned mc=0, ms=0;
        #pragma unroll
        for(int q=0;q<16;q++){
          const float4 A = rA[q & 3], B = rB[q & 3]; const float2 C = rC[q & 3];
          const u64 cx=pk(A.x,A.y), cy=pk(A.z,A.w), cz=pk(B.x,B.y), w=pk(B.z,B.w), cl=pk(C.x,C.y);
          u64 c = fadd2(w, pk(noo,noo));
          c = ffma2(pk(oz2,oz2), cz, c); c = ffma2(pk(oy2,oy2), cy, c); c = ffma2(pk(ox2,ox2), cx, c);
          u64 b1 = ffma2(pk(ndz,ndz), cz, pk(od,od)); b1 = ffma2(pk(ndy,ndy), cy, b1); b1 = ffma2(pk(ndx,ndx), cx, b1);
          u64 b2 = fadd2(cl, pk(oL,oL));
          u64 d1 = ffma2(b1,b1,c), d2 = ffma2(b2,b2,c);
          unsigned lo,hi;
          upk(d1,lo,hi); mc = __funnelshift_l(lo, mc, 1); mc = __funnelshift_l(hi, mc, 1);
          upk(d2,lo,hi); ms = __funnelshift_l(lo, ms, 1); ms = __funnelshift_l(hi, ms, 1);
          noo += 1e-9f; od += 1e-9f;          // keep the iterations distinct (2 scalar FADD per pair)
        }
        tot += __popc(mc) + __popc(ms); noo += 1e-6f * (float)(tot & 1);
      }
    }
    out[blockIdx.x*blockDim.x+threadIdx.x]=__uint_as_float(tot);
    return;
  }
  for(int it=0; it<iters; ++it){
    for(int g=0; g<npairs/16; ++g){
      unsigned mc=0, ms=0;
      #pragma unroll
      for(int q=0;q<16;q++){
        const float4 A = SRC ? cA[g*16+q] : sA[g*16+q], B = SRC ? cB[g*16+q] : sB[g*16+q]; const float2 C = SRC ? cC[g*16+q] : sC[g*16+q];
        const u64 cx=pk(A.x,A.y), cy=pk(A.z,A.w), cz=pk(B.x,B.y), w=pk(B.z,B.w), cl=pk(C.x,C.y);
        u64 c = fadd2(w, pk(noo,noo));
        c = ffma2(pk(oz2,oz2), cz, c); c = ffma2(pk(oy2,oy2), cy, c); c = ffma2(pk(ox2,ox2), cx, c);
        u64 b1 = ffma2(pk(ndz,ndz), cz, pk(od,od)); b1 = ffma2(pk(ndy,ndy), cy, b1); b1 = ffma2(pk(ndx,ndx), cx, b1);
        u64 b2 = fadd2(cl, pk(oL,oL));
        u64 d1 = ffma2(b1,b1,c), d2 = ffma2(b2,b2,c);
        unsigned lo,hi;
        upk(d1,lo,hi); mc = __funnelshift_l(lo, mc, 1); mc = __funnelshift_l(hi, mc, 1);
        upk(d2,lo,hi); ms = __funnelshift_l(lo, ms, 1); ms = __funnelshift_l(hi, ms, 1);
      }
      tot += __popc(mc) + __popc(ms); noo += 1e-6f * (float)(tot & 1);
    }
  }
  out[blockIdx.x*blockDim.x+threadIdx.x]=__uint_as_float(tot);
}
int main(){
  cudaDeviceProp p; cudaGetDeviceProperties(&p,0); int sms=p.multiProcessorCount;
  float* out; cudaMalloc(&out, sizeof(float)*sms*8*1024);
  const int npairs=NP; float* sph; cudaMalloc(&sph, npairs*10*sizeof(float));
  float* h=(float*)malloc(npairs*40); for(int i=0;i<npairs*10;i++) h[i]=(i%7)*0.37f-1.f; cudaMemcpy(sph,h,npairs*40,cudaMemcpyHostToDevice);
  cudaMemcpyToSymbol(cA, h, npairs*16); cudaMemcpyToSymbol(cB, h+npairs*4, npairs*16); cudaMemcpyToSymbol(cC, h+npairs*8, npairs*8);
  const int it=100;
  for (int src = 0; src < 5; ++src) for (int bps : {2, 3, 4}) {
    const int grid = sms*bps, threads = 256;
    cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    auto f = [&]{ if (src == 3) k<3><<<grid,threads,npairs*40>>>(out,sph,npairs,it); else if (src == 4) k<4><<<grid,threads,npairs*40>>>(out,sph,npairs,it); else if (src == 2) k<2><<<grid,threads,0>>>(out,sph,npairs,it); else if (src) k<1><<<grid,threads,0>>>(out,sph,npairs,it); else k<0><<<grid,threads,npairs*40>>>(out,sph,npairs,it); };
    f(); cudaDeviceSynchronize(); cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms,e0,e1);
    const double tests=(double)grid*threads*it*npairs*2*2;
    printf("SRC %d (%s) thr %d bps %d: %.3f ms  %.3e tests/s  -> %.2f clk per pair per SMSP @1965\n", src, src == 4 ? "shared, manual order D=4" : src == 3 ? "shared, manual order D=2" : src == 2 ? "registers" : src ? "constant" : "shared", threads, bps, ms, tests/(ms*1e-3),
           ms*1e-3*1.965e9 / ((double)it*npairs*(threads/32.0*bps/4.0)));
  }
  return 0;
}
