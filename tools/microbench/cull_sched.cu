// Cull-loop scheduling microbenchmark (round 2).  Question: the fused cull loop issues 10 FFMA2/FADD2 + 4 SHF + 2.5 LDS per
// sphere pair and runs at ~28.6 clk per pair although the FMA pipe needs only 20 (an FFMA2 holds it for 2 cycles) -- are the
// idle issue cycles caused by SHF / LDS instructions that sit right behind their producers (so they are never READY when
// the FMA pipe is busy and a non-FMA instruction could fill the slot)?  Variants move the sign-bit shifts one or two pairs
// behind the FMAs that produce them and the loads one or two pairs ahead of their consumers, in closest+shadow (CS) and
// closest-only (C) form.  Same occupancy as the render kernels: 128-thread CTAs, 7 per SM (28 KB dynamic shared memory each).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk2(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk2(u64 v, uint32_t& lo, uint32_t& hi) { asm("mov.b64 {%0,%1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v)); }
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 fadd2(u64 a, u64 b) { u64 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 bc2(float x) { return pk2(x, x); }
struct CullRay { float ox2, oy2, oz2, ndx, ndy, ndz, noo, od, oL; };
#define SHIFT2(m, v) { uint32_t lo_, hi_; upk2(v, lo_, hi_); m = __funnelshift_l(lo_, m, 1); m = __funnelshift_l(hi_, m, 1); }

template <bool SH> __device__ __forceinline__ void fma_part(const CullRay& r, const float4 A, const float4 B, const float2 C, u64& d1, u64& d2) {
    const u64 cx = pk2(A.x, A.y), cy = pk2(A.z, A.w), cz = pk2(B.x, B.y);
    u64 nc2 = fadd2(pk2(B.z, B.w), bc2(r.noo));
    nc2 = ffma2(bc2(r.oz2), cz, nc2); nc2 = ffma2(bc2(r.oy2), cy, nc2); nc2 = ffma2(bc2(r.ox2), cx, nc2);
    u64 b1 = ffma2(bc2(r.ndz), cz, bc2(r.od)); b1 = ffma2(bc2(r.ndy), cy, b1); b1 = ffma2(bc2(r.ndx), cx, b1);
    d1 = ffma2(b1, b1, nc2);
    if (SH) { const u64 b2 = fadd2(pk2(C.x, C.y), bc2(r.oL)); d2 = ffma2(b2, b2, nc2); }
}
// DELAY: pairs between the FMAs and the shifts of their sign bits (0 = as in round 1); AHEAD: pairs between a load and its use
template <bool SH, int DELAY, int AHEAD> __global__ void __launch_bounds__(896) k(const CullRay* rays, const float4* g, int np, int iters, uint32_t* out) {
    extern __shared__ __align__(16) unsigned char smem[];
    float4* sA = reinterpret_cast<float4*>(smem); float4* sB = sA + np; float2* sC = reinterpret_cast<float2*>(sB + np);
    for (int i = threadIdx.x; i < np * 5 / 2; i += blockDim.x) sA[i] = g[i];
    __syncthreads();
    CullRay r = rays[blockIdx.x * blockDim.x + threadIdx.x];
    uint32_t acc = 0;
    for (int it = 0; it < iters; ++it) {
        for (int gi = 0; gi < np / 16; ++gi) {
            uint32_t mc = 0, ms = 0;
            u64 q1[DELAY + 1], q2[DELAY + 1];
            float4 A[AHEAD + 1], B[AHEAD + 1]; float2 C[AHEAD + 1];
#pragma unroll
            for (int a = 0; a < AHEAD; ++a) { A[a] = sA[gi * 16 + a]; B[a] = sB[gi * 16 + a]; if (SH) C[a] = sC[gi * 16 + a]; }
#pragma unroll
            for (int q = 0; q < 16 + DELAY; ++q) {
                if (q + AHEAD < 16) { A[(q + AHEAD) % (AHEAD + 1)] = sA[gi * 16 + q + AHEAD]; B[(q + AHEAD) % (AHEAD + 1)] = sB[gi * 16 + q + AHEAD]; if (SH) C[(q + AHEAD) % (AHEAD + 1)] = sC[gi * 16 + q + AHEAD]; }
                if (q >= DELAY) { SHIFT2(mc, q1[(q - DELAY) % (DELAY + 1)]); if (SH) SHIFT2(ms, q2[(q - DELAY) % (DELAY + 1)]); }
                if (q < 16) fma_part<SH>(r, A[q % (AHEAD + 1)], B[q % (AHEAD + 1)], C[q % (AHEAD + 1)], q1[q % (DELAY + 1)], q2[q % (DELAY + 1)]);
            }
            acc += __popc(~mc) + 3 * __popc(~ms);
        }
        r.noo += 1e-7f * static_cast<float>(acc & 1u);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
// R independent closest-only rays per lane share each loaded sphere pair (stage 1 with R frames of a pixel per lane)
template <int R> __global__ void __launch_bounds__(896) k_r(const CullRay* rays, const float4* g, int np, int iters, uint32_t* out) {
    extern __shared__ __align__(16) unsigned char smem[];
    float4* sA = reinterpret_cast<float4*>(smem); float4* sB = sA + np;
    for (int i = threadIdx.x; i < np * 5 / 2; i += blockDim.x) sA[i] = g[i];
    __syncthreads();
    CullRay r[R];
#pragma unroll
    for (int j = 0; j < R; ++j) { r[j] = rays[blockIdx.x * blockDim.x + threadIdx.x]; r[j].od += 0.37f * j; r[j].ndx -= 0.11f * j; }
    uint32_t acc = 0;
    for (int it = 0; it < iters; ++it) {
        for (int gi = 0; gi < np / 16; ++gi) {
            uint32_t mc[R];
#pragma unroll
            for (int j = 0; j < R; ++j) mc[j] = 0;
#pragma unroll
            for (int q = 0; q < 16; ++q) {
                const float4 A = sA[gi * 16 + q], B = sB[gi * 16 + q];
#pragma unroll
                for (int j = 0; j < R; ++j) { u64 d1, d2; fma_part<false>(r[j], A, B, make_float2(0.f, 0.f), d1, d2); SHIFT2(mc[j], d1); }
            }
#pragma unroll
            for (int j = 0; j < R; ++j) acc += (j + 1) * __popc(~mc[j]);
        }
        r[0].noo += 1e-7f * static_cast<float>(acc & 1u);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
// What do the sign-collecting funnel shifts cost?  Fused loop (closest + shadow) with the per-value SHF replaced by
//   COLLECT 0: SHF.L.W (as in the kernels)   1: one LOP3 xor per value (same count, no shift)   2: F2FP pack of the two halves to f16x2
//   (signs land in bits 15 / 31 of ONE word) + one xor per PAIR   3: nothing per value: a packed FADD2 accumulates the discriminants
template <int COLLECT> __global__ void __launch_bounds__(896) k_c(const CullRay* rays, const float4* g, int np, int iters, uint32_t* out) {
    extern __shared__ __align__(16) unsigned char smem[];
    float4* sA = reinterpret_cast<float4*>(smem); float4* sB = sA + np;
    for (int i = threadIdx.x; i < np * 5 / 2; i += blockDim.x) sA[i] = g[i];
    __syncthreads();
    CullRay r = rays[blockIdx.x * blockDim.x + threadIdx.x];
    uint32_t acc = 0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll 1
        for (int gi = 0; gi < np / 16; ++gi) {
            uint32_t mc = 0, ms = 0; u64 sc = 0, ss = 0;
#pragma unroll
            for (int q = 0; q < 16; ++q) {
                const float4 A = sA[gi * 16 + q], B = sB[gi * 16 + q];
                const u64 cx = pk2(A.x, A.y), cy = pk2(A.z, A.w), cz = pk2(B.x, B.y);
                u64 nc2 = fadd2(pk2(B.z, B.w), bc2(r.noo));
                nc2 = ffma2(bc2(r.oz2), cz, nc2); nc2 = ffma2(bc2(r.oy2), cy, nc2); nc2 = ffma2(bc2(r.ox2), cx, nc2);
                u64 b1 = ffma2(bc2(r.ndz), cz, bc2(r.od)); b1 = ffma2(bc2(r.ndy), cy, b1); b1 = ffma2(bc2(r.ndx), cx, b1);
                const u64 d1 = ffma2(b1, b1, nc2);
                const u64 b2 = ffma2(bc2(-1.0f), cz, bc2(r.oL));
                const u64 d2 = ffma2(b2, b2, nc2);
                uint32_t l1, h1, l2, h2; upk2(d1, l1, h1); upk2(d2, l2, h2);
                if (COLLECT == 0) { mc = __funnelshift_l(l1, mc, 1); mc = __funnelshift_l(h1, mc, 1); ms = __funnelshift_l(l2, ms, 1); ms = __funnelshift_l(h2, ms, 1); }
                else if (COLLECT == 1) { mc ^= l1; mc ^= h1 >> 1; ms ^= l2; ms ^= h2 >> 1; }
                else if (COLLECT == 2) {
                    uint32_t p1, p2;
                    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(p1) : "f"(__uint_as_float(h1)), "f"(__uint_as_float(l1)));
                    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(p2) : "f"(__uint_as_float(h2)), "f"(__uint_as_float(l2)));
                    mc = __funnelshift_l(p1 & 0x80008000u, mc, 1) ^ p1; ms = __funnelshift_l(p2 & 0x80008000u, ms, 1) ^ p2;
                } else { sc = fadd2(sc, d1); ss = fadd2(ss, d2); }
            }
            if (COLLECT == 3) { uint32_t a0, a1, a2, a3; upk2(sc, a0, a1); upk2(ss, a2, a3); mc = a0 ^ a1; ms = a2 ^ a3; }
            acc += __popc(~mc) + 3 * __popc(~ms);
        }
        r.noo += 1e-7f * static_cast<float>(acc & 1u);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <class F> float timeit(F f) { cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b); f(); cudaDeviceSynchronize(); cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms, a, b); return ms; }
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount, np = 256, tpb = 128, bps = 7, iters = 40, grid = sms * bps;
    CullRay* rays; cudaMalloc(&rays, sizeof(CullRay) * grid * tpb);
    { CullRay* h = new CullRay[grid * tpb]; for (int i = 0; i < grid * tpb; ++i) { const float t = (i % 977) * 1e-3f; h[i] = CullRay{0.1f + t, 0.2f + t, 0.3f + t, 0.6f + t, 0.5f, 0.3f - t, -t, 1.f + t, 2.f + t}; }
      cudaMemcpy(rays, h, sizeof(CullRay) * grid * tpb, cudaMemcpyHostToDevice); delete[] h; }
    float4* g; cudaMalloc(&g, np * 40); { float* h = new float[np * 10]; for (int i = 0; i < np * 10; ++i) h[i] = (i % 7) * 0.37f - 1.f; cudaMemcpy(g, h, np * 40, cudaMemcpyHostToDevice); delete[] h; }
    uint32_t* out; cudaMalloc(&out, 4 * grid * tpb);
    const size_t smem = 28 * 1024;
#define RUN(SH, D, A) { cudaFuncSetAttribute(k<SH, D, A>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        const float ms = timeit([&] { k<SH, D, A><<<grid, tpb, smem>>>(rays, g, np, iters, out); }); const cudaError_t e = cudaGetLastError(); \
        const double pairs = (double)grid * tpb / 32.0 * iters * np;   /* warp-level pair evaluations */ \
        printf("%s delay %d ahead %d: %8.3f ms  %6.2f clk per pair per SMSP @%d MHz %s\n", SH ? "CS" : "C ", D, A, ms, ms * 1e-3 * p.clockRate * 1e3 / (pairs / (sms * 4)), p.clockRate / 1000, e ? cudaGetErrorString(e) : ""); }
    RUN(true, 0, 0) RUN(true, 1, 0) RUN(true, 2, 0) RUN(true, 0, 1) RUN(true, 1, 1) RUN(true, 2, 1) RUN(true, 1, 2) RUN(true, 2, 2) RUN(true, 4, 2)
    RUN(false, 0, 0) RUN(false, 1, 0) RUN(false, 2, 0) RUN(false, 1, 1) RUN(false, 2, 1) RUN(false, 2, 2) RUN(false, 4, 2)
#define RUNR(R) { cudaFuncSetAttribute(k_r<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        const float ms = timeit([&] { k_r<R><<<grid, tpb, smem>>>(rays, g, np, iters, out); }); const cudaError_t e = cudaGetLastError(); \
        const double pairs = (double)grid * tpb / 32.0 * iters * np * R;   /* warp-level pair evaluations per ray */ \
        printf("C  %d rays per lane: %8.3f ms  %6.2f clk per pair and ray per SMSP %s\n", R, ms, ms * 1e-3 * p.clockRate * 1e3 / (pairs / (sms * 4)), e ? cudaGetErrorString(e) : ""); }
    RUNR(1) RUNR(2) RUNR(3) RUNR(4)
#define RUNC(Cc) { cudaFuncSetAttribute(k_c<Cc>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        const float ms = timeit([&] { k_c<Cc><<<grid, tpb, smem>>>(rays, g, np, iters, out); }); const cudaError_t e = cudaGetLastError(); \
        const double pairs = (double)grid * tpb / 32.0 * iters * np; \
        printf("CS light-aligned (2 LDS), sign collection %d: %8.3f ms  %6.2f clk per pair per SMSP %s\n", Cc, ms, ms * 1e-3 * p.clockRate * 1e3 / (pairs / (sms * 4)), e ? cudaGetErrorString(e) : ""); }
    RUNC(0) RUNC(1) RUNC(2) RUNC(3)
    return 0;
}
