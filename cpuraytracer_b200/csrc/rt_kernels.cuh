// rt_kernels.cuh -- the sm_100a kernels of the `src/spheres` hot path.
//   rt_render_kernel   K1-K4 fused persistent megakernel: ray generation (spheres-app.cpp:132-161 +
//                      camera.cpp:30-48), closest hit + shadow query (ray-tracing.cpp:42-84,
//                      light.cpp:13-18), scatter (material.cpp:20-164), direct light (light.cpp:11-42),
//                      integrator (spheres-app.cpp:238-257) and per-pixel accumulation (:177-184).
//   rt_combine_kernel  adds the per-chunk partial sums to the running sum in a fixed order.
//   rt_resolve_kernel  K5: mean, ACES, gamma, BGRA8 (spheres-app.cpp:186-214).
//   rt_primary_kernel  K6: primary-hit sphere / material id buffers.
#pragma once
#include "rt_device.cuh"
#include "rt_bvh.cuh"

namespace rt {

struct RenderParams {
    // scene (global memory; the cull arrays are staged into shared memory by every CTA)
    const float4* cullA; const float4* cullB;   // pair arrays of the cull in the light-aligned frame of THIS launch's light (rt_device.cuh, SceneView)
    float rot[9];                // that frame: rows e1, e2, L (identity without a light)
    const DevMat* mats;
    const float4* sphere_geom;   // (cx, cy, cz, radius) per sphere (padded to n_pairs_pad * 2), world space: exact tests, hit records
    int n_spheres, n_pairs_pad, n_groups;
    int n_lights;
    float Lx, Ly, Lz;            // normalised light direction (light.cpp:7)
    float lcx, lcy, lcz;         // luminance * colour
    float skyx, skyy, skyz;      // sky luminance * colour (material.cpp:172-175)
    DevCamera cam;
    // frame
    int W, H, tiles_x, tiles_y;
    u64 first_sample;            // m_sampleCount of the first frame
    int spp, chunk_spp, n_chunks, max_depth;
    uint32_t flags, seed;
    float exposure;              // 2^m_exposure (spheres-app.cpp:174)
    const float2* jitter;        // [spp]  Halton(2,3) frame jitter (spheres-app.cpp:140)
    const float2* lens;          // [spp + W + H - 1]  HaltonSampleDisk(n+i+j; 4,5) (spheres-app.cpp:152)
    float4* hdr;                 // running sum, m_backbufferHdr
    float4* partial;             // [n_chunks][W*H] when n_chunks > 1
    u64* work_counter;
    u64* stats;                  // [0] closest rays [1] shadow rays [2] warp passes [3] paths [4] shadow rays answered without a query
                                 // [5] sphere tests evaluated for rays that asked for them (both query modes)
    AccelParams accel;           // RT_FLAG_ACCEL: box hierarchy image (rt_bvh.cuh)
    float iLx, iLy, iLz;         // safe_inv of the light direction (slab test of the shadow ray)
    int primary_pass;            // 1: this launch owns the sample count and the path / closest-ray statistics; 0: an additional light's pass over the same paths
    u64* sampler;                // RT_FLAG_HALTON: [n_materials][2] = (m_reflectionProbabilitySampleIndex, m_sampleIndex), material.h:34,50-51,67
    // two-stage schedule (rt_first_hit_kernel + rt_path_kernel): one launch pair renders the batch of frames
    // [first_sample + batch_s0, first_sample + batch_s0 + batch_n); chunk_spp / n_chunks then split the BATCH into work items
    uint16_t* hits;              // [W*H][hit_stride]: slot + 1 of the sphere the primary ray of (pixel, batch sample) hits, 0 = sky
    int hit_stride;              // samples per row: batch_n rounded up to a multiple of 16 (rows are whole 32-byte sectors)
    int batch_s0, batch_n;
    int n_ladder;                // stage 2: > 0: chunk c of the batch covers samples [ladder[c], ladder[c + 1]) (halving sizes, see rt_b200.cu);
    int ladder[10];              // 0: every chunk has chunk_spp samples.  Chunk 0 is accumulated straight into hdr, chunk c >= 1 goes through
                                 // partial[c - 1].  All boundaries are multiples of 8 (record alignment)
    int overwrite;               // 1: this launch starts the running sum (no read of hdr): RT_FLAG_OVERWRITE, first batch of the call
};

#define RT_FLAG_JITTER_OFF_ 1u
#define RT_FLAG_LENS_OFF_ 2u
#define RT_FLAG_WAVEFRONT_ 4u
#define RT_FLAG_ACCEL_ 8u
#define RT_FLAG_HALTON_ 16u
#define RT_FLAG_OVERWRITE_ 32u
#define RT_FLAG_DEBUG_NO_CULL_ 0x100u
#define RT_FLAG_DEBUG_ONE_KERNEL_ 0x200u   /* round-1 schedule: primary rays inside the path megakernel (kept for the measured comparison) */

enum LaneState { ST_NEED_ITEM = 0, ST_NEW_PATH = 1, ST_IN_PATH = 2, ST_DONE = 3 };

// Shared-memory layout of one CTA: [cullA np*16][cullB np*16][geom 2np*16][candidate lists 2*CAP*blockDim*4].
// `lists` is this thread's column of the word-interleaved candidate lists (entry e at lists[e * blockDim.x]).
#ifndef RT_MAX_TPB
#define RT_MAX_TPB 896   /* launch bound of the brute-force kernels: 72 registers, no spills -> 7 CTAs x 128 threads = 28 warps per SM (r01t: +1 % over 768 / 80 registers) */
#endif
#ifdef RT_MIN_BLOCKS
#define RT_RENDER_BOUNDS __launch_bounds__(RT_MAX_TPB, RT_MIN_BLOCKS)
#else
#define RT_RENDER_BOUNDS __launch_bounds__(RT_MAX_TPB)
#endif
// GLOBAL_SCENE (MODE_CULL_GLOBAL): a sphere list too large for one CTA's shared memory is read in place from global memory
// (broadcast loads through L1 / L2; every lane of a warp reads the same address) and shared memory holds the candidate lists only.
template <bool GLOBAL_SCENE = false>
__device__ __forceinline__ SceneView stage_scene(const RenderParams& p, uint32_t*& lists, unsigned char* smem) {
    const int np = p.n_pairs_pad;
    SceneView sv;
    sv.rot = p.rot;
    if constexpr (GLOBAL_SCENE) {
        sv.A = p.cullA; sv.B = p.cullB; sv.G = p.sphere_geom;
        lists = reinterpret_cast<uint32_t*>(smem) + threadIdx.x;
        return sv;
    }
    float4* sA = reinterpret_cast<float4*>(smem);
    float4* sB = sA + np;
    float4* sG = sB + np;
    lists = reinterpret_cast<uint32_t*>(sG + 2 * np) + threadIdx.x;
    for (int i = threadIdx.x; i < np; i += blockDim.x) { sA[i] = p.cullA[i]; sB[i] = p.cullB[i]; }
    for (int i = threadIdx.x; i < 2 * np; i += blockDim.x) sG[i] = p.sphere_geom[i];
    __syncthreads();
    sv.A = sA; sv.B = sB; sv.G = sG;
    return sv;
}
// (cx, cy, cz, r) of the sphere in `slot`: brute force -> the world-space copy; hierarchy -> the image's pair arrays + radii
template <int MODE_>
__device__ __forceinline__ float4 sphere_at(const SceneView& sv, const float4* sA, const float4* sB, const float* sR, int slot) {
    if constexpr (MODE_ == 2) {
        const float* fa = reinterpret_cast<const float*>(sA) + (slot >> 1) * 4 + (slot & 1);
        return make_float4(fa[0], fa[2], reinterpret_cast<const float*>(sB)[(slot >> 1) * 4 + (slot & 1)], sR[slot]);
    } else return sv.G[slot];
}
#if RT_FLAT_EXACT
// This warp's queue, behind the candidate lists: [best keys 32 x 8 B][tasks RT_WQ_CAP x 4 B][occlusion word, 8 B] per warp.
__device__ __forceinline__ WarpQueue warp_queue(const uint32_t* lists_col0) {
    unsigned char* base = reinterpret_cast<unsigned char*>(const_cast<uint32_t*>(lists_col0) + 2 * RT_LIST_CAP * blockDim.x)
                        + static_cast<size_t>(threadIdx.x >> 5) * RT_WQ_BYTES_PER_WARP;
    WarpQueue q;
    q.best = reinterpret_cast<u64*>(base);
    q.tasks = reinterpret_cast<uint32_t*>(base + 32 * 8);
    q.occ = reinterpret_cast<uint32_t*>(base + 32 * 8 + RT_WQ_CAP * 4);
    return q;
}
#endif

// Unshadowed DirectionalLight::Shade (light.cpp:20-40) for one light; the occlusion test is resolved by the
// next fused pass (the shadow ray shares its origin with the scattered ray).
// `tex` = texture_rgb(m, u, v), evaluated once per hit and shared with scatter().
__device__ __forceinline__ V3 direct_unshadowed(const RenderParams& p, const DevMat& m, V3 pos, V3 n, V3 tex, float n_dot_l) {
    V3 albedo = mk3(0.f, 0.f, 0.f), f0 = mk3(0.04f, 0.04f, 0.04f);            // material.h:27-29,43-45,60-62
    if (m.kind == 0) albedo = tex;
    else if (m.kind == 1) f0 = tex;
    const float s = m.smoothness;
    const V3 L = mk3(p.Lx, p.Ly, p.Lz);
    const V3 rin = mk3(p.lcx * n_dot_l, p.lcy * n_dot_l, p.lcz * n_dot_l);     // :27
    const V3 view = normalize3(mk3(p.cam.ox, p.cam.oy, p.cam.oz) - pos);       // :33 (camera origin, spheres-app.cpp:250)
    const V3 half = normalize3(L + view);                                      // :34
    const float n_dot_h = saturate1(dot3(n, half));                            // :35
    const float n_dot_v = saturate1(dot3(view, n));                            // :36
    const float pw5 = pow5(1.0f - n_dot_v);
    const float pw = pow_lobe(n_dot_h, s);                                     // :38
    const float k = s + 8.f;
    const V3 refl = mk3(f0.x + (1.0f - f0.x) * pw5, f0.y + (1.0f - f0.y) * pw5, f0.z + (1.0f - f0.z) * pw5);   // :37
    const V3 spec = mk3(((refl.x * 0.125f) * k) * pw, ((refl.y * 0.125f) * k) * pw, ((refl.z * 0.125f) * k) * pw);
    return rin * (albedo + spec);                                              // :40
}

// Material::Scatter, material.cpp:20-164.  Returns false when the path ends here.
// The mirror direction normalize(reflect(d, n)) is what all three materials return on their reflecting branch (:35, :89, :155);
// it is evaluated once in front of the material switch (the warp would otherwise run up to three copies of it per pass).
__device__ __forceinline__ bool scatter(const DevMat& m, V3 d, V3 pos, V3 n, V3 tex,
                                        float xi, float u1, float u2, V3& atten, V3& out_d) {
    const V3 mirror = normalize3(reflect3(d, n));
    if (m.kind == 0) {                                                         // DielectricOpaque
        const float ndv = dot3(neg3(d), n);
        if (!(ndv > 0.f)) return false;                                        // :22
        const float refl = schlick(0.04f, saturate1(ndv));                     // :25-27
        if (refl > xi) {                                                       // :29-30
            atten = mk3(1.f, 1.f, 1.f);
            out_d = mirror;                                                    // :35
            return true;
        }
        atten = tex;                                                           // :41
        const V3 dir = hemisphere_from_u(u1, u2);                              // :44
        const V3 up = fabsf(n.x) < 0.5f ? mk3(1.f, 0.f, 0.f) : mk3(0.f, 1.f, 0.f);   // :50
        const V3 b1 = cross3(up, n);                                           // :51
        const V3 b2 = cross3(n, b1);                                           // :52
        out_d = normalize3((dir.x * b1 + dir.y * b2) + dir.z * n);             // :55-56
        return true;
    }
    if (m.kind == 1) {                                                         // Metal
        const float ndv = dot3(neg3(d), n);
        if (!(ndv > 0.f)) return false;                                        // :74
        // :77-85: reflectance = f0 + (1 - f0) (1 - n.v)^5 on all FOUR lanes of the texture colour, reflect when ANY lane exceeds
        // the random number.  Lane w is the texture's alpha: with alpha == 1 that lane is exactly 1 > xi, i.e. always.
        const float alpha = texture_alpha(m, 0.5f * n.x + 0.5f, 0.5f * n.z + 0.5f);
        if (alpha != 1.0f) {
            const float nv = saturate1(ndv);
            if (!(schlick(tex.x, nv) > xi || schlick(tex.y, nv) > xi || schlick(tex.z, nv) > xi || schlick(alpha, nv) > xi)) return false;
        }
        atten = tex;                                                           // :87
        out_d = mirror;                                                        // :89
        return true;
    }
    atten = mk3(1.f, 1.f, 1.f);                                                // DielectricTransparent :114
    V3 outward; float eta, cos_i;
    const float dn = dot3(d, n);
    if (dn > 0.f) { outward = neg3(n); eta = m.ior; cos_i = dn; }              // :121-127
    else { outward = n; eta = 1.0f / m.ior; cos_i = dot3(d, neg3(n)); }        // :128-134 (exact reciprocal)
    const V3 refr = refract3(d, outward, eta);                                 // :137
    const bool can_refract = (refr.x != 0.f) || (refr.y != 0.f) || (refr.z != 0.f);
    const float prob = can_refract ? fresnel_term(cos_i, m.ior) : 1.0f;        // :140-149
    out_d = (prob > xi) ? mirror : normalize3(refr);                           // :153-161
    return true;
}

// RT_FLAG_HALTON: the reference's sampler -- one post-incremented Halton index per material instance and purpose
// (material.cpp:29,44,82,151), drawn in the order and under the conditions Scatter draws them.  The counters live in global
// memory and are shared by every path, exactly as the reference's std::atomic members are shared by its worker threads, so
// (like the reference) the image is reproducible only when paths run one after the other (a 1x1 framebuffer here).
__device__ __forceinline__ void draw_halton(const DevMat& m, u64* sampler, V3 d, V3 n, float& xi, float& u1, float& u2) {
    u64* c = sampler + 2 * m.mat_index;
    xi = 0.f; u1 = 0.f; u2 = 0.f;
    if (m.kind == 0) {
        const float ndv = dot3(neg3(d), n);
        if (!(ndv > 0.f)) return;                                              // :22
        xi = halton_sample(atomicAdd(&c[0], 1ull), 3);                         // :29
        if (!(schlick(0.04f, saturate1(ndv)) > xi)) {                          // :30
            const u64 k = atomicAdd(&c[1], 1ull);                              // :44
            u1 = halton_sample(k, 5); u2 = halton_sample(k, 7);
        }
    } else if (m.kind == 1) {
        if (!(dot3(neg3(d), n) > 0.f)) return;                                 // :74
        xi = halton_sample(atomicAdd(&c[0], 1ull), 3);                         // :82
    } else {
        xi = halton_sample(atomicAdd(&c[1], 1ull), 7);                         // :151
    }
}

// Maps a linear work index to (chunk, pixel).  Pixels are enumerated in 8x4 tiles so the 32 lanes of a warp
// start on a compact screen tile (coherent primary rays); chunks are the outer loop.
__device__ __forceinline__ bool item_to_pixel(const RenderParams& p, u64 item, int& chunk, int& px, int& py) {
    const u64 per_chunk = static_cast<u64>(p.tiles_x) * p.tiles_y * 32ull;
    chunk = static_cast<int>(item / per_chunk);
    const uint32_t r = static_cast<uint32_t>(item % per_chunk);
    const uint32_t tile = r >> 5, in = r & 31u;
    px = static_cast<int>((tile % p.tiles_x) * 8u + (in & 7u));
    py = static_cast<int>((tile / p.tiles_x) * 4u + (in >> 3));
    return px < p.W && py < p.H;
}

// MODE: 0 = brute force with the conservative cull (default), 1 = brute force, every sphere through the exact test
// (debug: proves the cull is result-neutral), 2 = RT_FLAG_ACCEL box hierarchy in front of the same cull (rt_bvh.cuh).
enum { MODE_CULL = 0, MODE_NO_CULL = 1, MODE_ACCEL = 2, MODE_CULL_GLOBAL = 3 };   // 3 = MODE_CULL with the sphere list left in global memory (stage_scene)
#ifndef RT_ACCEL_MAX_TPB
#define RT_ACCEL_MAX_TPB 640
#endif
/* the hierarchy walk keeps two sets of slab constants live: 128 registers per thread */
template <int MODE, bool HALTON = false, int GSHIFT = 5>
__global__ void __launch_bounds__(MODE == MODE_ACCEL ? RT_ACCEL_MAX_TPB : RT_MAX_TPB) rt_render_kernel(const RenderParams p) {
    extern __shared__ __align__(256) unsigned char smem[];
    float4 *sA = nullptr, *sB = nullptr; float* sR = nullptr; uint32_t* lists = nullptr; SceneView sv{nullptr, nullptr, nullptr, nullptr};
    AccelSmem av{};
    if constexpr (MODE == MODE_ACCEL) {
        av = stage_accel(p.accel, smem);
        sA = const_cast<float4*>(av.sA); sB = const_cast<float4*>(av.sB); sR = const_cast<float*>(av.sR);
    } else sv = stage_scene(p, lists, smem);
    const int lstride = static_cast<int>(blockDim.x);
    WarpQueue wq{nullptr, nullptr, nullptr};
#if RT_FLAT_EXACT
    if constexpr (MODE == MODE_CULL) wq = warp_queue(lists - threadIdx.x);
#endif
    const V3 invL = mk3(p.iLx, p.iLy, p.iLz);
    AccelCounters acnt; acnt.group_visits = 0u;

    const unsigned FULL = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31u;
    const u64 n_items = static_cast<u64>(p.tiles_x) * p.tiles_y * 32ull * static_cast<u64>(p.n_chunks);
    const V3 Ldir = mk3(p.Lx, p.Ly, p.Lz);
    const V3 sky = mk3(p.skyx, p.skyy, p.skyz);
    const bool has_light = p.n_lights > 0;

    int st = ST_NEED_ITEM;
    int pix = 0, px = 0, py = 0, chunk = 0, s_cur = 0, s_end = 0, depth = 0;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    V3 o = mk3(0.f, 0.f, 0.f), d = mk3(0.f, 0.f, 1.f);
    V3 T = mk3(1.f, 1.f, 1.f), Lsum = mk3(0.f, 0.f, 0.f), pend = mk3(0.f, 0.f, 0.f);
    bool want_closest = false, want_shadow = false;
    uint32_t n_closest = 0, n_shadow = 0, n_pass = 0, n_paths = 0, n_skip = 0;

    for (;;) {
        // ---- work distribution: warp-aggregated fetch from the global queue ----
        const unsigned need = __ballot_sync(FULL, st == ST_NEED_ITEM);
        if (need) {
            const int leader = __ffs(need) - 1;
            u64 base = 0;
            if (static_cast<int>(lane) == leader) base = atomicAdd(p.work_counter, static_cast<u64>(__popc(need)));
            base = __shfl_sync(FULL, base, leader);
            if (st == ST_NEED_ITEM) {
                const u64 item = base + __popc(need & ((1u << lane) - 1u));
                if (item >= n_items) st = ST_DONE;
                else if (item_to_pixel(p, item, chunk, px, py)) {
                    pix = py * p.W + px;
                    s_cur = chunk * p.chunk_spp;
                    s_end = min(p.spp, s_cur + p.chunk_spp);
                    acc = (p.n_chunks == 1) ? p.hdr[pix] : make_float4(0.f, 0.f, 0.f, 0.f);
                    st = ST_NEW_PATH;
                }
            }
        }
        if (__all_sync(FULL, st == ST_DONE)) break;

        // ---- K1: primary ray (GenerateRays + Camera::GetRay) ----
        if (st == ST_NEW_PATH) {
            float jx = 0.f, jy = 0.f, lx = 0.f, ly = 0.f;
            if (!(p.flags & RT_FLAG_JITTER_OFF_)) { const float2 j = __ldg(&p.jitter[s_cur]); jx = j.x; jy = j.y; }
            if (!(p.flags & RT_FLAG_LENS_OFF_)) { const float2 l = __ldg(&p.lens[s_cur + px + py]); lx = l.x; ly = l.y; }
            const float u = div_rn(static_cast<float>(px) + jx, static_cast<float>(p.W));   // spheres-app.cpp:149
            const float v = div_rn(static_cast<float>(py) + jy, static_cast<float>(p.H));   // :150
            camera_get_ray(p.cam, u, v, lx, ly, o, d);
            T = mk3(1.f, 1.f, 1.f); Lsum = mk3(0.f, 0.f, 0.f);
            depth = 0; want_closest = true; want_shadow = false;
            st = ST_IN_PATH;
        }
        const bool active = (st == ST_IN_PATH);

        // ---- K2+K3: fused closest-hit / shadow query over every sphere ----
        QueryResult q;
        if constexpr (MODE == MODE_ACCEL) q = fused_query_accel<GSHIFT>(av, p.accel, lstride, o, d, Ldir, invL, active && want_closest, active && want_shadow, acnt);
        else q = fused_query<MODE == MODE_NO_CULL>(sv, lists, lstride, p.n_groups, p.n_spheres, o, d, Ldir, active && want_closest, active && want_shadow, wq);
        ++n_pass;

        // ---- K4: integrate ----
        if (active) {
            if (want_shadow) { ++n_shadow; if (!q.occluded) Lsum = Lsum + pend; }        // light.cpp:15-19
            want_shadow = false;
            if (want_closest) {
                ++n_closest;
                if (q.idx < 0) {                                                         // spheres-app.cpp:255
                    Lsum = Lsum + T * sky;
                    want_closest = false;
                } else {
                    const int idx = q.idx, slot = (MODE == MODE_ACCEL) ? q.slot : q.idx;
                    const float4 sg = sphere_at<MODE>(sv, sA, sB, sR, slot);
                    const V3 ctr = mk3(sg.x, sg.y, sg.z);
                    const float radius = sg.w;
                    const V3 pos = mk3(q.t * d.x + o.x, q.t * d.y + o.y, q.t * d.z + o.z);   // ray-tracing.cpp:61
                    const V3 n = (pos - ctr) / radius;                                   // :62
                    const float tu = 0.5f * n.x + 0.5f, tv = 0.5f * n.z + 0.5f;          // :36-37
                    const DevMat m = p.mats[idx];
                    const V3 tex = texture_rgb(m, tu, tv);                               // texture.cpp:8-33, once per hit
                    const bool cont = depth < p.max_depth;                               // spheres-app.cpp:247
                    V3 atten = mk3(0.f, 0.f, 0.f), nd = d;
                    bool scattered = false;
                    if (cont || HALTON) {                                                // the reference scatters before it tests the depth (:246-247)
                        float xi, u1, u2;
                        if constexpr (HALTON) draw_halton(m, p.sampler, d, n, xi, u1, u2);
                        else rng_counter(static_cast<uint32_t>(pix), p.first_sample + static_cast<u64>(s_cur), static_cast<uint32_t>(depth), p.seed, xi, u1, u2);
                        scattered = scatter(m, d, pos, n, tex, xi, u1, u2, atten, nd);      // :246
                    }
                    if (has_light) {                                                     // :250
                        const float n_dot_l = saturate1(dot3(n, Ldir));                  // light.cpp:26
                        if (n_dot_l > 0.f) { pend = T * direct_unshadowed(p, m, pos, n, tex, n_dot_l); want_shadow = true; }
                        else { ++n_shadow; ++n_skip; }   // facing away: radianceIn == 0 (light.cpp:27), so the term is 0 whether or not the
                                             // reference's shadow ray (:13-15) is occluded; the ray is counted, not traced
                    }
                    o = pos;
                    if (cont && scattered) { T = T * atten; d = nd; ++depth; }           // :251
                    else want_closest = false;
                }
            }
            if (!want_closest && !want_shadow) {                                         // path complete
                acc.x += Lsum.x * p.exposure; acc.y += Lsum.y * p.exposure; acc.z += Lsum.z * p.exposure;   // :182-183
                if (p.primary_pass) { acc.w += 1.0f; ++n_paths; }
                if (++s_cur < s_end) st = ST_NEW_PATH;
                else {
                    if (p.n_chunks == 1) p.hdr[pix] = acc;
                    else p.partial[static_cast<size_t>(chunk) * (static_cast<size_t>(p.W) * p.H) + pix] = acc;
                    st = ST_NEED_ITEM;
                }
            }
        }
    }

    // ---- stats: warp reduce, one atomic per warp ----
    u64 c0 = p.primary_pass ? n_closest : 0u, c1 = n_shadow, c3 = n_paths, c4 = n_skip;
    // sphere tests evaluated for rays that asked for them: every sphere per query (brute force), or the big spheres plus
    // 2^GSHIFT per visited group (box hierarchy)
    u64 c5 = (MODE == MODE_ACCEL)
        ? static_cast<u64>(p.accel.n_big_pairs * 2) * (n_closest + n_shadow - n_skip) + static_cast<u64>(1 << GSHIFT) * acnt.group_visits
        : static_cast<u64>(p.n_spheres) * (n_closest + n_shadow - n_skip);
    for (int off = 16; off > 0; off >>= 1) {
        c0 += __shfl_down_sync(FULL, c0, off); c1 += __shfl_down_sync(FULL, c1, off); c3 += __shfl_down_sync(FULL, c3, off);
        c4 += __shfl_down_sync(FULL, c4, off); c5 += __shfl_down_sync(FULL, c5, off);
    }
    if (lane == 0) {
        atomicAdd(&p.stats[0], c0); atomicAdd(&p.stats[1], c1); atomicAdd(&p.stats[2], static_cast<u64>(n_pass)); atomicAdd(&p.stats[3], c3);
        atomicAdd(&p.stats[4], c4); atomicAdd(&p.stats[5], c5);
    }
}

// ------------------------------------------------------------------------------------------------------------------
// Two-stage schedule (default since round 2).  Measured on the one-kernel megakernel above (profiles/r01y_*): 42 % of all
// fused passes carry a PRIMARY ray -- a closest-only query whose shadow half idles -- and after every pass only ~55 % of
// the lanes have a hit to shade (most paths end in the sky), so the shading code runs at ~12 of 32 threads.  Stage 1 traces
// all primary rays of a batch of frames on their own: warps of 8x4 pixel tiles, coherent rays, closest-only cull (8 instead
// of 10 packed FP ops per sphere pair, no per-light array), no shading code, and leaves one 16-bit record per (pixel, sample):
// the slot of the sphere hit, or 0 for the sky.  Stage 2 is the path megakernel with regeneration FROM those records: sky
// samples are accumulated without a query, and a regenerated path arrives with a hit in hand, so the single shading block
// of the loop serves the query's hits AND the regenerated primary hits (~30 of 32 lanes busy instead of ~17).
// Arithmetic per path is unchanged (same device functions, same order), so the bits are those of the one-kernel schedule.
// ------------------------------------------------------------------------------------------------------------------

// First-hit records are written once and read once: RT_REC_STREAM = 1 marks both sides as streaming (evict-first), so that the
// records do not push the running-sum / partial-sum lines out of L2 before their half-sector stores have merged.
#ifndef RT_REC_STREAM
#define RT_REC_STREAM 0
#endif
#if RT_REC_STREAM
#define RT_REC_LOAD(ptr) __ldcs(ptr)
#define RT_REC_STORE(ptr, v) __stcs(ptr, v)
#else
#define RT_REC_LOAD(ptr) __ldg(ptr)
#define RT_REC_STORE(ptr, v) (*(ptr) = (v))
#endif
#ifndef RT_FIRST_RAYS
#define RT_FIRST_RAYS 4    /* frames per lane and query of the brute-force first-hit stage (1, 2 or 4) */
#endif
// Stage 1.  Work item of a WARP = (8x4 pixel tile, group of 16 consecutive batch samples); static round-robin over the grid.
#ifndef RT_FIRST_MAX_TPB
#define RT_FIRST_MAX_TPB 512   /* brute-force first-hit stage with RT_FIRST_RAYS rays per lane: up to 128 registers per thread */
#endif
template <int MODE, int GSHIFT = 5>
__global__ void __launch_bounds__(MODE == MODE_ACCEL ? RT_ACCEL_MAX_TPB : (MODE == MODE_NO_CULL || RT_FIRST_RAYS == 1) ? RT_MAX_TPB : RT_FIRST_MAX_TPB)
rt_first_hit_kernel(const RenderParams p) {
    extern __shared__ __align__(256) unsigned char smem[];
    float4 *sA = nullptr, *sB = nullptr; float* sR = nullptr; uint32_t* lists = nullptr; SceneView sv{nullptr, nullptr, nullptr, nullptr};
    AccelSmem av{};
    if constexpr (MODE == MODE_ACCEL) av = stage_accel(p.accel, smem);
    else sv = stage_scene<MODE == MODE_CULL_GLOBAL>(p, lists, smem);
    AccelCounters acnt; acnt.group_visits = 0u;
    const unsigned FULL = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31u;
    const V3 Ldir = mk3(p.Lx, p.Ly, p.Lz);
    const u64 n_tiles = static_cast<u64>(p.tiles_x) * p.tiles_y;
    const u64 n_items = n_tiles * static_cast<u64>(p.hit_stride >> 4);
    const u64 n_warps = (static_cast<u64>(gridDim.x) * blockDim.x) >> 5;
    uint32_t n_closest = 0, n_pass = 0;
    for (u64 item = (static_cast<u64>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5; item < n_items; item += n_warps) {
        const uint32_t tile = static_cast<uint32_t>(item % n_tiles);
        const int g = static_cast<int>(item / n_tiles);
        const int px = static_cast<int>((tile % p.tiles_x) * 8u + (lane & 7u));
        const int py = static_cast<int>((tile / p.tiles_x) * 4u + (lane >> 3));
        const bool valid = px < p.W && py < p.H;
        u64 rec0 = 0ull, rec1 = 0ull, rec2 = 0ull, rec3 = 0ull;      // 16 records of this lane's pixel = one 32-byte sector
        // R frames of the pixel per query: the brute-force cull shares every sphere-pair load between R rays (closest_query_multi)
        constexpr int R = (MODE == MODE_CULL || MODE == MODE_CULL_GLOBAL) ? RT_FIRST_RAYS : 1;
#pragma unroll 1
        for (int k = 0; k < 16; k += R) {
            if (g * 16 + k >= p.batch_n) break;              // warp-uniform
            V3 o[R], d[R]; bool want[R]; QueryResult q[R];
#pragma unroll
            for (int j = 0; j < R; ++j) {
                const int sb = g * 16 + k + j;               // sample index inside the batch
                const int s = p.batch_s0 + sb;               // index into the per-call Halton tables
                want[j] = valid && sb < p.batch_n;
                o[j] = mk3(0.f, 0.f, 0.f); d[j] = mk3(0.f, 0.f, 1.f);
                if (want[j]) {
                    float jx = 0.f, jy = 0.f, lx = 0.f, ly = 0.f;
                    if (!(p.flags & RT_FLAG_JITTER_OFF_)) { const float2 jj = __ldg(&p.jitter[s]); jx = jj.x; jy = jj.y; }
                    if (!(p.flags & RT_FLAG_LENS_OFF_)) { const float2 l = __ldg(&p.lens[s + px + py]); lx = l.x; ly = l.y; }
                    const float u = div_rn(static_cast<float>(px) + jx, static_cast<float>(p.W));   // spheres-app.cpp:149
                    const float v = div_rn(static_cast<float>(py) + jy, static_cast<float>(p.H));   // :150
                    camera_get_ray(p.cam, u, v, lx, ly, o[j], d[j]);
                    ++n_closest;
                }
                if (sb < p.batch_n) ++n_pass;
            }
            if constexpr (R == 1) {
                if constexpr (MODE == MODE_ACCEL) q[0] = fused_query_accel<GSHIFT, RAYS_CLOSEST>(av, p.accel, static_cast<int>(blockDim.x), o[0], d[0], Ldir, mk3(p.iLx, p.iLy, p.iLz), want[0], false, acnt);
                else q[0] = fused_query<MODE == MODE_NO_CULL, RAYS_CLOSEST>(sv, lists, static_cast<int>(blockDim.x), p.n_groups, p.n_spheres, o[0], d[0], Ldir, want[0], false);
            } else closest_query_multi<R>(sv, lists, static_cast<int>(blockDim.x), p.n_groups, p.n_spheres, o, d, want, q);
#pragma unroll
            for (int j = 0; j < R; ++j) {
                const int kk = k + j;
                const u64 v16 = ((want[j] && q[j].idx >= 0) ? static_cast<u64>(((MODE == MODE_ACCEL) ? q[j].slot : q[j].idx) + 1) : 0ull) << (16 * (kk & 3));
                if (kk < 4) rec0 |= v16; else if (kk < 8) rec1 |= v16; else if (kk < 12) rec2 |= v16; else rec3 |= v16;
            }
        }
        if (valid) {                                         // two 16-byte stores into one aligned 32-byte sector: no partial-sector write reaches HBM
            const size_t pix = static_cast<size_t>(py) * p.W + px;
            uint4* dst = reinterpret_cast<uint4*>(p.hits + pix * p.hit_stride + static_cast<size_t>(g) * 16);
            RT_REC_STORE(dst, make_uint4(static_cast<uint32_t>(rec0), static_cast<uint32_t>(rec0 >> 32), static_cast<uint32_t>(rec1), static_cast<uint32_t>(rec1 >> 32)));
            RT_REC_STORE(dst + 1, make_uint4(static_cast<uint32_t>(rec2), static_cast<uint32_t>(rec2 >> 32), static_cast<uint32_t>(rec3), static_cast<uint32_t>(rec3 >> 32)));
        }
    }
    u64 c0 = n_closest;
    u64 c5 = (MODE == MODE_ACCEL) ? static_cast<u64>(p.accel.n_big_pairs * 2) * n_closest + static_cast<u64>(1 << GSHIFT) * acnt.group_visits
                                  : static_cast<u64>(p.n_spheres) * n_closest;
    for (int off = 16; off > 0; off >>= 1) { c0 += __shfl_down_sync(FULL, c0, off); c5 += __shfl_down_sync(FULL, c5, off); }
    if (lane == 0) { atomicAdd(&p.stats[0], c0); atomicAdd(&p.stats[2], static_cast<u64>(n_pass)); atomicAdd(&p.stats[5], c5); }
}

// Stage 2.  One thread = one path at a time; work item of a lane = (pixel, chunk of the batch's samples), handed out by a
// warp-aggregated atomicAdd as in the one-kernel schedule.  Lane states:
//   NEED_ITEM -> SCAN (walk the pixel's first-hit records in sample order: a sky sample is accumulated on the spot, a hit
//   sample becomes a path) -> HIT_NEW (primary ray regenerated, t recomputed by the same exact test stage 1 ran) -> SHADE
//   (K4 at the vertex in hand) -> QUERY (fused closest + shadow pass) -> SHADE (hit) | SCAN (path complete) ... -> DONE.
// DIRECT_ONLY (max_depth == 0, BASELINE configs[1]): no lane ever asks for a closest hit here, the pass is shadow-only.
enum PathLaneState { PS_NEED_ITEM = 0, PS_SCAN = 1, PS_HIT_NEW = 2, PS_SHADE = 3, PS_QUERY = 4, PS_DONE = 5 };
template <int MODE, bool HALTON = false, int GSHIFT = 5, bool DIRECT_ONLY = false>
__global__ void __launch_bounds__(MODE == MODE_ACCEL ? RT_ACCEL_MAX_TPB : RT_MAX_TPB) rt_path_kernel(const RenderParams p) {
    extern __shared__ __align__(256) unsigned char smem[];
    float4 *sA = nullptr, *sB = nullptr; float* sR = nullptr; uint32_t* lists = nullptr; SceneView sv{nullptr, nullptr, nullptr, nullptr};
    AccelSmem av{};
    if constexpr (MODE == MODE_ACCEL) {
        av = stage_accel(p.accel, smem);
        sA = const_cast<float4*>(av.sA); sB = const_cast<float4*>(av.sB); sR = const_cast<float*>(av.sR);
    } else sv = stage_scene<MODE == MODE_CULL_GLOBAL>(p, lists, smem);
    const int lstride = static_cast<int>(blockDim.x);
    const V3 invL = mk3(p.iLx, p.iLy, p.iLz);
    AccelCounters acnt; acnt.group_visits = 0u;
    constexpr int kRays = DIRECT_ONLY ? RAYS_SHADOW : RAYS_BOTH;

    const unsigned FULL = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31u;
    const u64 n_items = static_cast<u64>(p.tiles_x) * p.tiles_y * 32ull * static_cast<u64>(p.n_chunks);
    const size_t n_pix = static_cast<size_t>(p.W) * p.H;
    const V3 Ldir = mk3(p.Lx, p.Ly, p.Lz);
    const V3 sky = mk3(p.skyx, p.skyy, p.skyz);
    const bool has_light = p.n_lights > 0;

    int st = PS_NEED_ITEM;
    int pix = 0, px = 0, py = 0, chunk = 0, s_cur = 0, s_end = 0, depth = 0, hit_slot = 0;
    float hit_t = 0.f;
    uint4 rec = make_uint4(0u, 0u, 0u, 0u);                 // first-hit records of batch samples [s_cur & ~7, +8) of this pixel
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    V3 o = mk3(0.f, 0.f, 0.f), d = mk3(0.f, 0.f, 1.f);
    V3 T = mk3(1.f, 1.f, 1.f), Lsum = mk3(0.f, 0.f, 0.f), pend = mk3(0.f, 0.f, 0.f);
    bool want_closest = false, want_shadow = false;
    uint32_t n_closest = 0, n_shadow = 0, n_pass = 0, n_paths = 0, n_skip = 0;

    for (;;) {
        // ---- regeneration: work items (warp-aggregated fetch) and the walk over the pixel's first-hit records ----
        for (;;) {
            const unsigned need = __ballot_sync(FULL, st == PS_NEED_ITEM);
            if (need) {
                const int leader = __ffs(need) - 1;
                u64 base = 0;
                if (static_cast<int>(lane) == leader) base = atomicAdd(p.work_counter, static_cast<u64>(__popc(need)));
                base = __shfl_sync(FULL, base, leader);
                if (st == PS_NEED_ITEM) {
                    const u64 item = base + __popc(need & ((1u << lane) - 1u));
                    if (item >= n_items) st = PS_DONE;
                    else if (item_to_pixel(p, item, chunk, px, py)) {
                        pix = py * p.W + px;
                        // chunk 0 continues the pixel's running sum in place; later chunks start partial sums (multiples of 8: records stay aligned)
                        if (p.n_ladder) { s_cur = p.ladder[chunk]; s_end = min(p.batch_n, p.ladder[chunk + 1]); }
                        else { s_cur = chunk * p.chunk_spp; s_end = min(p.batch_n, s_cur + p.chunk_spp); }
                        acc = (chunk == 0 && !p.overwrite) ? p.hdr[pix] : make_float4(0.f, 0.f, 0.f, 0.f);
                        st = PS_SCAN;
                    }
                }
            }
            if (st == PS_SCAN) {
                for (;;) {
                    if (s_cur >= s_end) {
                        if (chunk == 0) p.hdr[pix] = acc;
                        else p.partial[static_cast<size_t>(chunk - 1) * n_pix + pix] = acc;
                        st = PS_NEED_ITEM;
                        break;
                    }
                    if ((s_cur & 7) == 0) rec = RT_REC_LOAD(reinterpret_cast<const uint4*>(p.hits + static_cast<size_t>(pix) * p.hit_stride + s_cur));
                    const int f = s_cur & 7;
                    const uint32_t w = (f < 2) ? rec.x : (f < 4) ? rec.y : (f < 6) ? rec.z : rec.w;
                    const uint32_t v16 = (w >> (16 * (f & 1))) & 0xffffu;
                    if (v16) { hit_slot = static_cast<int>(v16) - 1; st = PS_HIT_NEW; break; }
                    // the primary ray left the scene: Lsum = T * sky with T = 1 (spheres-app.cpp:255), then :182-183
                    acc.x += sky.x * p.exposure; acc.y += sky.y * p.exposure; acc.z += sky.z * p.exposure;
                    if (p.primary_pass) { acc.w += 1.0f; ++n_paths; }
                    ++s_cur;
                }
            }
            if (!__any_sync(FULL, st == PS_NEED_ITEM)) break;
        }
        if (__all_sync(FULL, st == PS_DONE)) break;

        // ---- K1 again for the regenerated paths: the primary ray of (pixel, sample) and the t stage 1 found on that sphere ----
        if (st == PS_HIT_NEW) {
            const int s = p.batch_s0 + s_cur;
            float jx = 0.f, jy = 0.f, lx = 0.f, ly = 0.f;
            if (!(p.flags & RT_FLAG_JITTER_OFF_)) { const float2 j = __ldg(&p.jitter[s]); jx = j.x; jy = j.y; }
            if (!(p.flags & RT_FLAG_LENS_OFF_)) { const float2 l = __ldg(&p.lens[s + px + py]); lx = l.x; ly = l.y; }
            const float u = div_rn(static_cast<float>(px) + jx, static_cast<float>(p.W));
            const float v = div_rn(static_cast<float>(py) + jy, static_cast<float>(p.H));
            camera_get_ray(p.cam, u, v, lx, ly, o, d);
            const float4 sg = sphere_at<MODE>(sv, sA, sB, sR, hit_slot);
            const float a = dot3(d, d);
            hit_t = 0.f;
            sphere_exact(sg.x, sg.y, sg.z, sg.w, o, d, a, rcp_refined(a), false, hit_t);
            T = mk3(1.f, 1.f, 1.f); Lsum = mk3(0.f, 0.f, 0.f);
            depth = 0; want_shadow = false;
            st = PS_SHADE;
        }

        // ---- K4 at the vertex in hand (a query hit of the last pass, or a regenerated primary hit) ----
        if (st == PS_SHADE) {
            const int slot = hit_slot;
            int idx = slot;
            if constexpr (MODE == MODE_ACCEL) idx = av.sI[slot];
            const float4 sg = sphere_at<MODE>(sv, sA, sB, sR, slot);
            const V3 ctr = mk3(sg.x, sg.y, sg.z);
            const float radius = sg.w;
            const V3 pos = mk3(hit_t * d.x + o.x, hit_t * d.y + o.y, hit_t * d.z + o.z);   // ray-tracing.cpp:61
            const V3 n = (pos - ctr) / radius;                                   // :62
            const float tu = 0.5f * n.x + 0.5f, tv = 0.5f * n.z + 0.5f;          // :36-37
            const DevMat m = p.mats[idx];
            const V3 tex = texture_rgb(m, tu, tv);                               // texture.cpp:8-33, once per hit
            const bool cont = depth < p.max_depth;                               // spheres-app.cpp:247
            V3 atten = mk3(0.f, 0.f, 0.f), nd = d;
            bool scattered = false;
            if (cont || HALTON) {                                                // the reference scatters before it tests the depth (:246-247)
                float xi, u1, u2;
                if constexpr (HALTON) draw_halton(m, p.sampler, d, n, xi, u1, u2);
                else rng_counter(static_cast<uint32_t>(pix), p.first_sample + static_cast<u64>(p.batch_s0 + s_cur), static_cast<uint32_t>(depth), p.seed, xi, u1, u2);
                scattered = scatter(m, d, pos, n, tex, xi, u1, u2, atten, nd);      // :246
            }
            want_shadow = false;
            if (has_light) {                                                     // :250
                const float n_dot_l = saturate1(dot3(n, Ldir));                  // light.cpp:26
                if (n_dot_l > 0.f) { pend = T * direct_unshadowed(p, m, pos, n, tex, n_dot_l); want_shadow = true; }
                else { ++n_shadow; ++n_skip; }   // facing away: radianceIn == 0 (light.cpp:27), so the term is 0 whether or not the
                                     // reference's shadow ray (:13-15) is occluded; the ray is counted, not traced
            }
            o = pos;
            want_closest = cont && scattered;
            if (want_closest) { T = T * atten; d = nd; ++depth; }                // :251
            st = PS_QUERY;
        }

        // ---- K2+K3: fused closest-hit / shadow query ----
        const bool active = (st == PS_QUERY);
        QueryResult q;
        if constexpr (MODE == MODE_ACCEL) q = fused_query_accel<GSHIFT, kRays>(av, p.accel, lstride, o, d, Ldir, invL, active && want_closest, active && want_shadow, acnt);
        else q = fused_query<MODE == MODE_NO_CULL, kRays>(sv, lists, lstride, p.n_groups, p.n_spheres, o, d, Ldir, active && want_closest, active && want_shadow);
        ++n_pass;

        // ---- resolve the pass ----
        if (active) {
            if (want_shadow) { ++n_shadow; if (!q.occluded) Lsum = Lsum + pend; }        // light.cpp:15-19
            want_shadow = false;
            bool complete = true;
            if (want_closest) {
                ++n_closest;
                if (q.idx < 0) Lsum = Lsum + T * sky;                                    // spheres-app.cpp:255
                else { hit_slot = (MODE == MODE_ACCEL) ? q.slot : q.idx; hit_t = q.t; st = PS_SHADE; complete = false; }
                want_closest = false;
            }
            if (complete) {
                acc.x += Lsum.x * p.exposure; acc.y += Lsum.y * p.exposure; acc.z += Lsum.z * p.exposure;   // :182-183
                if (p.primary_pass) { acc.w += 1.0f; ++n_paths; }
                ++s_cur;
                st = PS_SCAN;
            }
        }
    }

    // ---- stats: warp reduce, one atomic per warp (primary rays were counted by stage 1) ----
    u64 c0 = p.primary_pass ? n_closest : 0u, c1 = n_shadow, c3 = n_paths, c4 = n_skip;
    u64 c5 = (MODE == MODE_ACCEL)
        ? static_cast<u64>(p.accel.n_big_pairs * 2) * (n_closest + n_shadow - n_skip) + static_cast<u64>(1 << GSHIFT) * acnt.group_visits
        : static_cast<u64>(p.n_spheres) * (n_closest + n_shadow - n_skip);
    for (int off = 16; off > 0; off >>= 1) {
        c0 += __shfl_down_sync(FULL, c0, off); c1 += __shfl_down_sync(FULL, c1, off); c3 += __shfl_down_sync(FULL, c3, off);
        c4 += __shfl_down_sync(FULL, c4, off); c5 += __shfl_down_sync(FULL, c5, off);
    }
    if (lane == 0) {
        atomicAdd(&p.stats[0], c0); atomicAdd(&p.stats[1], c1); atomicAdd(&p.stats[2], static_cast<u64>(n_pass)); atomicAdd(&p.stats[3], c3);
        atomicAdd(&p.stats[4], c4); atomicAdd(&p.stats[5], c5);
    }
}

// Per-call Halton tables on the device (no host table, no pageable copy, no stream synchronisation in rt_render):
// jitter[s] = Halton2D(first + s; 2, 3) (spheres-app.cpp:140), lens[k] = HaltonSampleDisk(first + k; 4, 5) (:152,
// quasi-random.cpp:52-61).  Same operations in the same order as the host restatement they replace (IEEE division, no
// contraction), so the primary rays are bit-identical to the reference's for every frame index.
__global__ void rt_tables_kernel(u64 first_sample, int n_jitter, int n_lens, float2* __restrict__ jitter, float2* __restrict__ lens) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_jitter) { const u64 n = first_sample + static_cast<u64>(i); jitter[i] = make_float2(halton_sample(n, 2), halton_sample(n, 3)); }
    if (i < n_lens) {
        const u64 n = first_sample + static_cast<u64>(i);
        const float theta = RT_TWO_PI * halton_sample(n, 4);                 // quasi-random.cpp:54
        const float r = halton_sample(n, 5);                                 // :55
        float sn, cs; sincos_pinned(theta, &sn, &cs);
        lens[i] = make_float2(r * cs, r * sn);                               // :57-60
    }
}

// hdr[p] = ((hdr[p] + partial[0][p]) + partial[1][p]) + ...   (fixed order => deterministic)
__global__ void rt_combine_kernel(float4* __restrict__ hdr, const float4* __restrict__ partial, int n_chunks, size_t n_pixels, int overwrite) {
    const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
    if (i >= n_pixels) return;
    float4 a = overwrite ? make_float4(0.f, 0.f, 0.f, 0.f) : hdr[i];
    for (int c = 0; c < n_chunks; ++c) {
        const float4 b = partial[static_cast<size_t>(c) * n_pixels + i];
        a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
    hdr[i] = a;
}

// K5 resolve, spheres-app.cpp:186-214.  16 B read + 4 B write per pixel (HBM-bound).
// powf(x, 1/2.2) followed by XMStoreColor's x255 + round-to-nearest-even, for x in [0,1].  The SFU form ex2(g * lg2(x)) is
// within ~1e-6 relative of powf (MUFU.LG2 / MUFU.EX2: <= 2^-21 absolute on the exponent, 2^-22 relative on the result), i.e.
// within 3e-4 of powf * 255; the rounded code can therefore differ only when the scaled value sits within 2e-3 of a
// half-integer, and exactly those values (0.4 % of the channels) are recomputed with powf.  Output == the powf-only kernel.
__device__ __forceinline__ uint32_t gamma_store_channel(float x, float inv_gamma) {
    if (!(x > 0.0f)) return 0u;                                              // powf(0, g) = 0 (x is already saturated: never negative / NaN-free)
    float lg, v;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg) : "f"(x));                   // a denormal x is flushed: its power scales to code 0 either way
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(v) : "f"(inv_gamma * lg));
    v = v * 255.0f;
    if (fabsf(fabsf(v - rintf(v)) - 0.5f) < 2e-3f) v = powf(x, inv_gamma) * 255.0f;   // spheres-app.cpp:208 as written
    v = fminf(v, 255.0f);
    return static_cast<uint32_t>(__float2int_rn(v));                         // :211 XMStoreColor
}
__device__ __forceinline__ uint32_t resolve_pixel(const float4 h, float n, float rn) {
    const float a = 2.51f, b = 0.03f, c = 2.43f, dd = 0.59f, e = 0.14f;
    const float inv_gamma = 1 / 2.2f;
    const float ch[3] = {h.x, h.y, h.z};
    uint32_t out[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float x = div_by(ch[k], n, rn);                                      // :201
        x = saturate1(div_rn(x * (a * x + b), x * (c * x + dd) + e));        // :205 (denominator >= 0.14)
        out[k] = gamma_store_channel(x, inv_gamma);                          // :208, :211
    }
    return 0xFF000000u | (out[0] << 16) | (out[1] << 8) | out[2];
}
// Four pixels per thread and iteration, loads issued before any arithmetic (4 x 16 B in flight per thread), streaming
// cache hints on both sides (every byte is touched once).
__global__ void __launch_bounds__(256) rt_resolve_kernel(const float4* __restrict__ hdr, uint32_t* __restrict__ ldr, float n, size_t n_pixels) {
    const float rn = rcp_refined(n);
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n_pixels; i += 4 * stride) {
        float4 h[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) { const size_t j = i + u * stride; if (j < n_pixels) h[u] = __ldcs(&hdr[j]); }
#pragma unroll
        for (int u = 0; u < 4; ++u) { const size_t j = i + u * stride; if (j < n_pixels) __stcs(&ldr[j], resolve_pixel(h[u], n, rn)); }
    }
}

// K6: primary-hit id buffers (jitter/lens per flags) for frame `first_sample`.
template <int MODE, int GSHIFT = 5>
__global__ void __launch_bounds__(MODE == MODE_ACCEL ? RT_ACCEL_MAX_TPB : RT_MAX_TPB) rt_primary_kernel(const RenderParams p, int32_t* __restrict__ sphere_idx, int32_t* __restrict__ material_id) {
    extern __shared__ __align__(256) unsigned char smem[];
    float4 *sA = nullptr, *sB = nullptr; float* sR = nullptr; uint32_t* lists = nullptr; SceneView sv{nullptr, nullptr, nullptr, nullptr};
    AccelSmem av{};
    if constexpr (MODE == MODE_ACCEL) av = stage_accel(p.accel, smem);
    else sv = stage_scene<MODE == MODE_CULL_GLOBAL>(p, lists, smem);
    AccelCounters acnt; acnt.group_visits = 0u;
    const u64 n_items = static_cast<u64>(p.tiles_x) * p.tiles_y * 32ull;
    for (u64 item = blockIdx.x * static_cast<u64>(blockDim.x) + threadIdx.x; item < ((n_items + 31ull) & ~31ull); item += static_cast<u64>(gridDim.x) * blockDim.x) {
        int chunk, px, py;
        const bool valid = item < n_items && item_to_pixel(p, item, chunk, px, py);
        V3 o = mk3(0.f, 0.f, 0.f), d = mk3(0.f, 0.f, 1.f);
        if (valid) {
            float jx = 0.f, jy = 0.f, lx = 0.f, ly = 0.f;
            if (!(p.flags & RT_FLAG_JITTER_OFF_)) { const float2 j = __ldg(&p.jitter[0]); jx = j.x; jy = j.y; }
            if (!(p.flags & RT_FLAG_LENS_OFF_)) { const float2 l = __ldg(&p.lens[px + py]); lx = l.x; ly = l.y; }
            const float u = div_rn(static_cast<float>(px) + jx, static_cast<float>(p.W));
            const float v = div_rn(static_cast<float>(py) + jy, static_cast<float>(p.H));
            camera_get_ray(p.cam, u, v, lx, ly, o, d);
        }
        QueryResult q;
        if constexpr (MODE == MODE_ACCEL) q = fused_query_accel<GSHIFT>(av, p.accel, static_cast<int>(blockDim.x), o, d, mk3(p.Lx, p.Ly, p.Lz), mk3(p.iLx, p.iLy, p.iLz), valid, false, acnt);
        else q = fused_query<MODE == MODE_NO_CULL>(sv, lists, static_cast<int>(blockDim.x), p.n_groups, p.n_spheres, o, d, mk3(p.Lx, p.Ly, p.Lz), valid, false);
        if (valid) {
            const int pix = py * p.W + px;
            if (sphere_idx) sphere_idx[pix] = q.idx;
            if (material_id) material_id[pix] = q.idx >= 0 ? p.mats[q.idx].kind : 3;
        }
    }
}

}  // namespace rt
