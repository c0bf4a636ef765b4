// rt_wavefront.cuh -- WAVEFRONT layout of the same render loop, kept beside the megakernel so the choice between the two
// is a measurement (BASELINE.json north_star), not an assumption.  Same device functions, same work items, same chunked
// accumulation => the same bits as rt_render_kernel; only the schedule differs:
//   path state lives in HBM as SoA float4 / int4 arrays (coalesced 16-byte loads and stores, one slot per thread),
//   wf_intersect_kernel  = K2+K3 for every live slot (fused closest + occlusion query; no shading code in the kernel),
//   wf_shade_kernel      = K4 integrate + accumulation + regeneration (next sample / next work item) + K1 primary ray.
// A slot whose path ends regenerates in place, so all P slots stay live until the work queue drains (no compaction pass
// is needed to keep the intersect kernel's warps full; the shade kernel still diverges by material like the megakernel).
#pragma once
#include "rt_kernels.cuh"

namespace rt {

struct WfState {
    float4* o;        // origin.xyz, w = t of the last query
    float4* d;        // direction.xyz, w = bit pattern of the last query's sphere index
    float4* thr;      // throughput T.xyz
    float4* rad;      // radiance sum of the current path Lsum.xyz
    float4* pend;     // unshadowed direct term waiting for its occlusion answer
    float4* acc;      // running sum of the current work item
    int4* ids;        // x = pixel, y = s_cur (sample index inside this launch), z = chunk, w = depth << 8 | flags
    int n_slots;
};
enum { WF_WANT_CLOSEST = 1, WF_WANT_SHADOW = 2, WF_LIVE = 4, WF_DONE = 8, WF_OCCLUDED = 16 };

template <bool NO_CULL>
__global__ void RT_RENDER_BOUNDS wf_intersect_kernel(const RenderParams p, const WfState s) {
    extern __shared__ __align__(256) unsigned char smem[];
    uint32_t* lists;
    const SceneView sv = stage_scene(p, lists, smem);
    const V3 Ldir = mk3(p.Lx, p.Ly, p.Lz);
    for (int slot = blockIdx.x * blockDim.x + threadIdx.x; slot < s.n_slots; slot += gridDim.x * blockDim.x) {
        const int4 id = s.ids[slot];
        const bool live = (id.w & WF_LIVE) != 0;
        if (__all_sync(0xffffffffu, !live)) continue;
        float4 o4 = s.o[slot], d4 = s.d[slot];
        const V3 o = mk3(o4.x, o4.y, o4.z), d = live ? mk3(d4.x, d4.y, d4.z) : mk3(0.f, 0.f, 1.f);
        const QueryResult q = fused_query<NO_CULL>(sv, lists, static_cast<int>(blockDim.x), p.n_groups, p.n_spheres, o, d, Ldir,
                                                   live && (id.w & WF_WANT_CLOSEST), live && (id.w & WF_WANT_SHADOW));
        if (live) {
            o4.w = q.t; d4.w = __int_as_float(q.idx);
            s.o[slot] = o4; s.d[slot] = d4;
            if (q.occluded) s.ids[slot] = make_int4(id.x, id.y, id.z, id.w | WF_OCCLUDED);
        }
    }
}

// A zero-filled ids array means "no path yet": the first launch only fetches work and generates primary rays.
__global__ void __launch_bounds__(256) wf_shade_kernel(const RenderParams p, const WfState s, unsigned int* live_out) {
    const unsigned FULL = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31u;
    const u64 n_items = static_cast<u64>(p.tiles_x) * p.tiles_y * 32ull * static_cast<u64>(p.n_chunks);
    const V3 Ldir = mk3(p.Lx, p.Ly, p.Lz);
    const V3 sky = mk3(p.skyx, p.skyy, p.skyz);
    const bool has_light = p.n_lights > 0;
    uint32_t n_closest = 0, n_shadow = 0, n_paths = 0, n_skip = 0, n_live = 0;

    for (int slot = blockIdx.x * blockDim.x + threadIdx.x; slot < s.n_slots; slot += gridDim.x * blockDim.x) {
        int4 id = s.ids[slot];
        int flags = id.w & 0xff, depth = id.w >> 8, chunk = id.z;
        int pix = id.x, s_cur = id.y, s_end = min(p.spp, (chunk + 1) * p.chunk_spp);
        if (__all_sync(FULL, (flags & WF_DONE) != 0)) continue;
        int st = (flags & WF_DONE) ? ST_DONE : ((flags & WF_LIVE) ? ST_IN_PATH : ST_NEED_ITEM);
        V3 o = mk3(0.f, 0.f, 0.f), d = mk3(0.f, 0.f, 1.f), T = mk3(1.f, 1.f, 1.f), Lsum = mk3(0.f, 0.f, 0.f), pend = mk3(0.f, 0.f, 0.f);
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        bool want_closest = false, want_shadow = false;

        // ---- K4: integrate the answers of the last intersect pass (same statements as the megakernel) ----
        if (st == ST_IN_PATH) {
            const float4 o4 = s.o[slot], d4 = s.d[slot], t4 = s.thr[slot], r4 = s.rad[slot], p4 = s.pend[slot];
            acc = s.acc[slot];
            o = mk3(o4.x, o4.y, o4.z); d = mk3(d4.x, d4.y, d4.z); T = mk3(t4.x, t4.y, t4.z); Lsum = mk3(r4.x, r4.y, r4.z); pend = mk3(p4.x, p4.y, p4.z);
            const float qt = o4.w; const int qidx = __float_as_int(d4.w);
            want_closest = (flags & WF_WANT_CLOSEST) != 0; want_shadow = (flags & WF_WANT_SHADOW) != 0;
            if (want_shadow) { ++n_shadow; if (!(flags & WF_OCCLUDED)) Lsum = Lsum + pend; }
            want_shadow = false;
            if (want_closest) {
                ++n_closest;
                if (qidx < 0) { Lsum = Lsum + T * sky; want_closest = false; }
                else {
                    const int idx = qidx;
                    const float4 sc = p.sphere_geom[idx];                                   // centre.xyz, radius
                    const V3 ctr = mk3(sc.x, sc.y, sc.z);
                    const float radius = sc.w;
                    const V3 pos = mk3(qt * d.x + o.x, qt * d.y + o.y, qt * d.z + o.z);
                    const V3 n = (pos - ctr) / radius;
                    const float tu = 0.5f * n.x + 0.5f, tv = 0.5f * n.z + 0.5f;
                    const DevMat m = p.mats[idx];
                    const V3 tex = texture_rgb(m, tu, tv);
                    const bool cont = depth < p.max_depth;
                    V3 atten = mk3(0.f, 0.f, 0.f), nd = d;
                    bool scattered = false;
                    if (cont) {
                        float xi, u1, u2;
                        rng_counter(static_cast<uint32_t>(pix), p.first_sample + static_cast<u64>(s_cur), static_cast<uint32_t>(depth), p.seed, xi, u1, u2);
                        scattered = scatter(m, d, pos, n, tex, xi, u1, u2, atten, nd);
                    }
                    if (has_light) {
                        const float n_dot_l = saturate1(dot3(n, Ldir));
                        if (n_dot_l > 0.f) { pend = T * direct_unshadowed(p, m, pos, n, tex, n_dot_l); want_shadow = true; }
                        else { ++n_shadow; ++n_skip; }
                    }
                    o = pos;
                    if (cont && scattered) { T = T * atten; d = nd; ++depth; }
                    else want_closest = false;
                }
            }
            if (!want_closest && !want_shadow) {
                acc.x += Lsum.x * p.exposure; acc.y += Lsum.y * p.exposure; acc.z += Lsum.z * p.exposure;
                acc.w += 1.0f;
                ++n_paths;
                if (++s_cur < s_end) st = ST_NEW_PATH;
                else {
                    if (p.n_chunks == 1) p.hdr[pix] = acc;
                    else p.partial[static_cast<size_t>(chunk) * (static_cast<size_t>(p.W) * p.H) + pix] = acc;
                    st = ST_NEED_ITEM;
                }
            }
        }
        // ---- regeneration: next work item (warp-aggregated fetch) ----
        const unsigned need = __ballot_sync(FULL, st == ST_NEED_ITEM);
        if (need) {
            const int leader = __ffs(need) - 1;
            u64 base = 0;
            if (static_cast<int>(lane) == leader) base = atomicAdd(p.work_counter, static_cast<u64>(__popc(need)));
            base = __shfl_sync(FULL, base, leader);
            if (st == ST_NEED_ITEM) {
                const u64 item = base + __popc(need & ((1u << lane) - 1u));
                int px, py;
                if (item >= n_items) st = ST_DONE;
                else if (item_to_pixel(p, item, chunk, px, py)) {
                    pix = py * p.W + px;
                    s_cur = chunk * p.chunk_spp;
                    s_end = min(p.spp, s_cur + p.chunk_spp);
                    acc = (p.n_chunks == 1) ? p.hdr[pix] : make_float4(0.f, 0.f, 0.f, 0.f);
                    st = ST_NEW_PATH;
                }
                // an item outside the image (ragged tile) leaves the slot in NEED_ITEM: it fetches again next pass
            }
        }
        // ---- K1: primary ray of the next sample ----
        if (st == ST_NEW_PATH) {
            const int px = pix % p.W, py = pix / p.W;
            float jx = 0.f, jy = 0.f, lx = 0.f, ly = 0.f;
            if (!(p.flags & RT_FLAG_JITTER_OFF_)) { const float2 j = __ldg(&p.jitter[s_cur]); jx = j.x; jy = j.y; }
            if (!(p.flags & RT_FLAG_LENS_OFF_)) { const float2 l = __ldg(&p.lens[s_cur + px + py]); lx = l.x; ly = l.y; }
            const float u = div_rn(static_cast<float>(px) + jx, static_cast<float>(p.W));
            const float v = div_rn(static_cast<float>(py) + jy, static_cast<float>(p.H));
            camera_get_ray(p.cam, u, v, lx, ly, o, d);
            T = mk3(1.f, 1.f, 1.f); Lsum = mk3(0.f, 0.f, 0.f);
            depth = 0; want_closest = true; want_shadow = false;
            st = ST_IN_PATH;
        }
        // ---- write the slot back ----
        int nf = 0;
        if (st == ST_IN_PATH) {
            nf = WF_LIVE | (want_closest ? WF_WANT_CLOSEST : 0) | (want_shadow ? WF_WANT_SHADOW : 0);
            s.o[slot] = make_float4(o.x, o.y, o.z, 0.f);
            s.d[slot] = make_float4(d.x, d.y, d.z, 0.f);
            s.thr[slot] = make_float4(T.x, T.y, T.z, 0.f);
            s.rad[slot] = make_float4(Lsum.x, Lsum.y, Lsum.z, 0.f);
            s.pend[slot] = make_float4(pend.x, pend.y, pend.z, 0.f);
            s.acc[slot] = acc;
            ++n_live;
        } else if (st == ST_DONE) nf = WF_DONE;
        else ++n_live;                                   // NEED_ITEM after a ragged-tile item: still has to drain the queue
        s.ids[slot] = make_int4(pix, s_cur, chunk, (depth << 8) | nf);
    }
    u64 c0 = n_closest, c1 = n_shadow, c3 = n_paths, c4 = n_skip; unsigned c5 = n_live;
    for (int off = 16; off > 0; off >>= 1) {
        c0 += __shfl_down_sync(FULL, c0, off); c1 += __shfl_down_sync(FULL, c1, off); c3 += __shfl_down_sync(FULL, c3, off);
        c4 += __shfl_down_sync(FULL, c4, off); c5 += __shfl_down_sync(FULL, c5, off);
    }
    if (lane == 0) {
        if (c0) atomicAdd(&p.stats[0], c0);
        if (c1) atomicAdd(&p.stats[1], c1);
        if (c3) atomicAdd(&p.stats[3], c3);
        if (c4) atomicAdd(&p.stats[4], c4);
        if (c0 + c1 - c4) atomicAdd(&p.stats[5], static_cast<u64>(p.n_spheres) * (c0 + c1 - c4));
        if (live_out && c5) atomicAdd(live_out, c5);
    }
}

}  // namespace rt
