// Does FFMA2 (fma.rn.f32x2) block the SMSP issue port for its second pipe cycle?
//   A: 8 independent FFMA2 chains                      B: A + one independent SHF per FFMA2 (ALU pipe)
//   C: 8 independent scalar FFMA chains                D: C + one SHF per FFMA
//   E: A + one LDS.128 (broadcast) per 4 FFMA2         F: 16 independent scalar FFMA chains (2x the FMAs of C, = the FMAs of A)
// If B takes as long as A, the ALU instruction issues in the shadow of the FFMA2 and the cull loop is FMA-pipe bound;
// if B ~ 1.5 A, an FFMA2 costs two issue slots and the loop is issue bound (SURVEY 8(d) roofline discussion in DESIGN.md).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b){ u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c){ u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
template<int MODE> __global__ void __launch_bounds__(256) k(float* out, int iters, float s, unsigned sh){
  __shared__ float4 sm[64];
  if (threadIdx.x < 64) sm[threadIdx.x] = make_float4(s, s, s, s);
  __syncthreads();
  const float t = threadIdx.x * 1e-3f;
  u64 a[8]; float f[16]; unsigned m[8];
  #pragma unroll
  for (int j = 0; j < 8; ++j) { a[j] = pk(t + j, t - j); m[j] = threadIdx.x + j; }
  #pragma unroll
  for (int j = 0; j < 16; ++j) f[j] = t + j;
  const u64 mul = pk(s, s), add = pk(1e-7f, 1e-7f);
  float4 acc4 = make_float4(0, 0, 0, 0);
  for (int it = 0; it < iters; ++it) {
    #pragma unroll
    for (int r = 0; r < 8; ++r) {
      #pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (MODE == 0 || MODE == 1 || MODE == 4 || MODE == 6 || MODE == 7) a[j] = ffma2(a[j], mul, add);
        if (MODE == 2 || MODE == 3) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[j]) : "f"(s), "f"(1e-7f));
        if (MODE == 5) { asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[j]) : "f"(s), "f"(1e-7f)); asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[j + 8]) : "f"(s), "f"(1e-7f)); }
        if (MODE == 1 || MODE == 3) asm volatile("shf.l.wrap.b32 %0, %0, %1, %2;" : "+r"(m[j]) : "r"(m[(j + 1) & 7]), "r"(sh));
        if (MODE == 6 && (j & 1) == 1) { float2 v; asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"((unsigned)__cvta_generic_to_shared(&sm[(r * 4 + (j >> 1)) & 63]))); acc4.x += v.x; }
        if (MODE == 7 && (j & 3) == 3) { float4 v; asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"((unsigned)__cvta_generic_to_shared(&sm[(r * 2 + (j >> 2)) & 63]))); asm volatile("" :: "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)); }
        if (MODE == 4 && (j & 3) == 3) { float4 v; asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"((unsigned)__cvta_generic_to_shared(&sm[(r * 2 + (j >> 2)) & 63]))); acc4.x += v.x; }
      }
    }
  }
  float o = acc4.x;
  #pragma unroll
  for (int j = 0; j < 8; ++j) { unsigned lo, hi; asm("mov.b64 {%0,%1}, %2;" : "=r"(lo), "=r"(hi) : "l"(a[j])); o += __uint_as_float(lo) + __uint_as_float(hi) + (float)m[j]; }
  #pragma unroll
  for (int j = 0; j < 16; ++j) o += f[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = o;
}
template<int MODE> void run(const char* name, float* out, int sms, int iters, double fma_per_thread_iter){
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<sms * 4, 256>>>(out, iters, 0.999f, 1u); cudaDeviceSynchronize();
  cudaEventRecord(e0); k<MODE><<<sms * 4, 256>>>(out, iters, 0.999f, 1u); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double warps_per_smsp = 4.0 * 256 / 32 / 4;           // 8 warps per SMSP
  const double slots = 64.0 * iters * warps_per_smsp;          // FP instructions per SMSP
  const double clk = ms * 1e-3 * 1.965e9;
  printf("%-44s %8.3f ms  %.3f clk per FP instruction per SMSP  %.1f TFLOP/s\n", name, ms, clk / slots, fma_per_thread_iter * iters * 2.0 * sms * 4 * 256 / (ms * 1e-3) / 1e12);
}
int main(){
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0); const int sms = p.multiProcessorCount;
  float* out; cudaMalloc(&out, sizeof(float) * sms * 4 * 256);
  const int it = 20000;
  run<0>("A  FFMA2 x64", out, sms, it, 128);
  run<1>("B  FFMA2 x64 + SHF x64", out, sms, it, 128);
  run<2>("C  FFMA  x64", out, sms, it, 64);
  run<3>("D  FFMA  x64 + SHF x64", out, sms, it, 64);
  run<4>("E  FFMA2 x64 + LDS.128 x16", out, sms, it, 128);
  run<5>("F  FFMA  x128 (clk per PAIR of FFMA)", out, sms, it, 128);
  run<6>("G  FFMA2 x64 + LDS.64 x32 (+FADD each)", out, sms, it, 128);
  run<7>("H  FFMA2 x64 + LDS.128 x16, result unused", out, sms, it, 128);
  return 0;
}
