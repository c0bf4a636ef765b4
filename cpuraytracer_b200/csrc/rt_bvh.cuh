// rt_bvh.cuh -- RT_FLAG_ACCEL: a two-level bounding-box hierarchy in shared memory in front of the same cull + exact tests
// (the reference's own closest-hit query walks a BVH, ray-tracing.cpp:107-214).
//
// Brute force (rt_device.cuh, fused_query) stays the parity ground truth; this path must return the SAME bits:
//   * the spheres are re-ordered on the host (rt_b200.cu, build_accel): "big" spheres first (radius > 4 x median -- the
//     r = 1000 ground would bloat every box that contains it), then the rest in Morton order of their centres;
//     8, 16 or 32 consecutive slots (AccelParams::gshift, the finest size that fits the fan-out) form a GROUP (the unit of the
//     candidate masks; the big spheres' groups always have 32), 8 consecutive groups an L2 NODE;
//   * a group is visited when the closest ray or the shadow ray touches its (padded) bounding box; visited groups go
//     through the SAME conservative FFMA2 discriminant cull as brute force, and every candidate through the SAME exact
//     Sphere::Intersect restatement (sphere_exact), so an accepted hit has the reference's t bit for bit;
//   * the winner is chosen by `closer()` (smaller t, then smaller ORIGINAL index), which does not depend on visiting order.
// So the only way to differ from brute force is to skip a group that holds a sphere the exact test accepts.  The box test
// is made conservative against that with a per-ray pad (see box_ray()); tests/test_gpu_parity.py compares whole renders,
// hit buffers and ray counters of both modes bit for bit.
//
// Schedule of one query (per lane; lanes of a warp run the same phase together):
//   T  top             converged loop over the n2 L2 boxes                       -> bit mask m2 of touched L2 nodes
//   M  middle          per lane: pop an L2 node, test its 8 group boxes          -> one byte per L2 node in shared memory
//   G  groups          the big spheres first (every lane, same addresses: broadcast LDS), then per lane: pop a touched group,
//                      cull its 4 / 8 / 16 sphere pairs for both rays (pair order XOR-swizzled by the lane id so that the
//                      quarter-warps' LDS.128 hit 8 different bank groups)       -> candidate-mask entries
//   E  exact           per lane: walk the entries, exact test per candidate
#pragma once
#include "rt_device.cuh"

namespace rt {

// Everything the kernel needs to find the sections of the shared-memory image (built on the host, copied verbatim).
struct AccelParams {
    const uint4* blob;       // device copy of the image
    int blob_bytes;          // multiple of 16
    int off_B, off_R, off_I, off_box, off_top;   // byte offsets of the sections (A starts at 0)
    int n_pairs;             // sphere pairs in the image (big groups padded to 32 slots, small groups to 2^gshift)
    int n_big_groups;        // groups that are always culled
    int n_big_pairs;         // sphere pairs of the big groups that hold spheres (the rest of the last big group is padding)
    int n_groups;            // small groups (Morton order), 2^gshift slots each
    int gshift;              // log2 of the slots per small group: 3, 4 or 5 (chosen per scene by build_accel; kernel template argument)
    int n2;                  // L2 nodes = ceil(n_groups / 8), <= 32
    int gmask_words;         // per-thread scratch words holding one byte per L2 node
    float csx, csy, csz, cs_radius;   // bounding sphere of the small spheres (centres + radii)
    float r_min, r_max;               // smallest / largest radius among the small spheres
    float c_max;                      // largest |coordinate| of any small-sphere box face
};

#ifndef RT_ACCEL_ENTRY_CAP
#define RT_ACCEL_ENTRY_CAP 8          /* candidate-mask entries (2 words each) per lane; a full list is drained (E) before G goes on */
#endif
#ifndef RT_ACCEL_FLAT_NODES
#define RT_ACCEL_FLAT_NODES 4         /* up to this many L2 nodes the top level is skipped (every node counts as touched) */
#endif
#define RT_ACCEL_GSHIFT_MAX_GROUPS 256 /* 32 L2 nodes x 8 groups */
#define RT_ACCEL_NODE_FLOATS 52       /* 8 children x (cx,cy,cz,hx,hy,hz) SoA = 48 floats, padded to 13 float4 (bank skew) */

struct AccelSmem {
    const float4* sA; const float4* sB; const float* sR; const uint16_t* sI;
    const float* box;        // [n2][RT_ACCEL_NODE_FLOATS]
    const float4* top;       // [n2][2]: (cx,cy,cz,hx) (hy,hz,-,-)
    uint32_t* scratch;       // this thread's column: word w at scratch[w * blockDim.x]
};

__device__ __forceinline__ AccelSmem stage_accel(const AccelParams& a, unsigned char* smem) {
    uint4* dst = reinterpret_cast<uint4*>(smem);
    const int n16 = a.blob_bytes >> 4;
    for (int i = threadIdx.x; i < n16; i += blockDim.x) dst[i] = a.blob[i];
    __syncthreads();
    AccelSmem s;
    s.sA = reinterpret_cast<const float4*>(smem);
    s.sB = reinterpret_cast<const float4*>(smem + a.off_B);
    s.sR = reinterpret_cast<const float*>(smem + a.off_R);
    s.sI = reinterpret_cast<const uint16_t*>(smem + a.off_I);
    s.box = reinterpret_cast<const float*>(smem + a.off_box);
    s.top = reinterpret_cast<const float4*>(smem + a.off_top);
    s.scratch = reinterpret_cast<uint32_t*>(smem + a.blob_bytes) + threadIdx.x;
    // the XOR-swizzled group loads (fused_query_accel, G) need the A and B sections on 256-byte boundaries
    if ((static_cast<uint32_t>(__cvta_generic_to_shared(smem)) | static_cast<uint32_t>(a.off_B)) & 255u) __trap();
    return s;
}

__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v; asm("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

// Conservative discriminant signs of one sphere pair for both rays; as cull_pair() but (o-c).L is evaluated with three
// FMAs from the centre instead of being read from a per-light array (the image carries no light-dependent data).
struct CullRay2 { float ox2, oy2, oz2, ndx, ndy, ndz, nLx, nLy, nLz, noo, od, oL; };
template <int RAYS>
__device__ __forceinline__ void cull_pair2(const CullRay2& r, const float4 A, const float4 B, uint32_t& mc, uint32_t& ms) {
    const u64 cx = pk2(A.x, A.y), cy = pk2(A.z, A.w), cz = pk2(B.x, B.y);
    u64 nc2 = fadd2(pk2(B.z, B.w), bc2(r.noo));
    nc2 = ffma2(bc2(r.oz2), cz, nc2); nc2 = ffma2(bc2(r.oy2), cy, nc2); nc2 = ffma2(bc2(r.ox2), cx, nc2);
    uint32_t lo, hi;
    if constexpr ((RAYS & RAYS_CLOSEST) != 0) {
        u64 b1 = ffma2(bc2(r.ndz), cz, bc2(r.od)); b1 = ffma2(bc2(r.ndy), cy, b1); b1 = ffma2(bc2(r.ndx), cx, b1);
        const u64 d1 = ffma2(b1, b1, nc2);
#if RT_CULL_DIRFILTER
        uint32_t blo, bhi, nlo, nhi;                         // as cull_pair(): centre behind the origin and origin outside = miss
        upk2(nc2, nlo, nhi);
        upk2(d1, lo, hi); upk2(b1, blo, bhi); lo |= ~blo & nlo; hi |= ~bhi & nhi;
#else
        upk2(d1, lo, hi);
#endif
        mc = __funnelshift_l(lo, mc, 1); mc = __funnelshift_l(hi, mc, 1);
    }
    if constexpr ((RAYS & RAYS_SHADOW) != 0) {
        u64 b2 = ffma2(bc2(r.nLz), cz, bc2(r.oL)); b2 = ffma2(bc2(r.nLy), cy, b2); b2 = ffma2(bc2(r.nLx), cx, b2);
        const u64 d2 = ffma2(b2, b2, nc2);
#if RT_CULL_DIRFILTER
        uint32_t blo, bhi, nlo, nhi;
        upk2(nc2, nlo, nhi);
        upk2(d2, lo, hi); upk2(b2, blo, bhi); lo |= ~blo & nlo; hi |= ~bhi & nhi;
#else
        upk2(d2, lo, hi);
#endif
        ms = __funnelshift_l(lo, ms, 1); ms = __funnelshift_l(hi, ms, 1);
    }
}

// Slab test constants of one ray.  A box (centre c, half extent h) is touched when, per axis k,
//     tn_k = (c_k - o_k) inv_k - (h_k + pad) |inv_k|,   tf_k = (c_k - o_k) inv_k + (h_k + pad) |inv_k|,
//     max_k tn_k <= min_k tf_k   and   min_k tf_k >= -t_slack.
// `pad` makes the test conservative against everything the exact test can accept:
//   (1) Sphere::Intersect in binary32 accepts a ray whose true distance p to the centre satisfies p^2 < r^2 + E, where E
//       bounds the rounding of disc = b*b - a*c (ray-tracing.cpp:44-50): |error| <= 16 ulp-units of |oc|^2 + r^2.  With the
//       ray's distance to any small sphere <= Dmax = |o - Cs| + Rs this is E <= 2^-18 (Dmax^2 + r_max^2) (4x margin), i.e. the
//       line passes within sqrt(r^2 + E) - r <= min(E / 2r_min, sqrt(E)) of the sphere: first pad term (x1.5);
//   (2) the slab arithmetic itself (one rounding of o_k*inv_k, one of the FMA) moves a plane by at most
//       2^-23 (|o_k| + |c_k|) in world units: second pad term 2^-19 (|o|_1 + c_max).
// The accepted root exceeds the 0.001 bias (ray-tracing.cpp:58,71) and the far root of the inflated sphere is not smaller
// than the computed one (up to the rounding of b, <= 2^-22 Dmax), hence tf >= -t_slack with t_slack = 2^-16 Dmax.
// Directions with |d_k| < 1e-18 use 1e-18 (both slabs then span the whole axis unless the origin is outside them).
struct BoxRay { float ix, iy, iz, ax, ay, az, nx, ny, nz, fx, fy, fz; };
__device__ __forceinline__ float safe_inv(float d) {
    const float dd = fabsf(d) < 1e-18f ? copysignf(1e-18f, d) : d;
    return div_rn(1.0f, dd);          // |dd| in [1e-18, ~1]: inside div_rn's range, so this is the correctly rounded 1/dd, without the
                                      // range check and slow path nvcc attaches to `1.0f / dd`
}
__device__ __forceinline__ BoxRay box_ray(V3 o, float ix, float iy, float iz, float pad) {
    BoxRay r;
    r.ix = ix; r.iy = iy; r.iz = iz;
    r.ax = fabsf(ix); r.ay = fabsf(iy); r.az = fabsf(iz);
    const float nox = -(o.x * ix), noy = -(o.y * iy), noz = -(o.z * iz);
    const float ex = pad * r.ax, ey = pad * r.ay, ez = pad * r.az;
    r.nx = nox - ex; r.ny = noy - ey; r.nz = noz - ez;
    r.fx = nox + ex; r.fy = noy + ey; r.fz = noz + ez;
    return r;
}
__device__ __forceinline__ bool box_touched(const BoxRay& r, float cx, float cy, float cz, float hx, float hy, float hz, float neg_t_slack) {
    const float tnx = fmaf(-hx, r.ax, fmaf(cx, r.ix, r.nx)), tfx = fmaf(hx, r.ax, fmaf(cx, r.ix, r.fx));
    const float tny = fmaf(-hy, r.ay, fmaf(cy, r.iy, r.ny)), tfy = fmaf(hy, r.ay, fmaf(cy, r.iy, r.fy));
    const float tnz = fmaf(-hz, r.az, fmaf(cz, r.iz, r.nz)), tfz = fmaf(hz, r.az, fmaf(cz, r.iz, r.fz));
    const float tn = fmaxf(fmaxf(tnx, tny), tnz);
    const float tf = fminf(fminf(tfx, tfy), tfz);
    return fmaxf(tn, neg_t_slack) <= tf;              // == (tn <= tf && tf >= neg_t_slack): no NaN reaches this point (finite boxes,
                                                      // |inv| <= 1e18; an empty child has h = -1e30, i.e. tn = +inf, tf = -inf)
}

struct AccelCounters { uint32_t group_visits; };

// Same contract as fused_query(): closest hit along (o,d) and occlusion along (o,L).  res.idx is the ORIGINAL sphere
// index, res.slot the position in the re-ordered image (for the hit record's centre / radius).
template <int GSHIFT, int RAYS = RAYS_BOTH>
__device__ __forceinline__ QueryResult fused_query_accel(const AccelSmem& s, const AccelParams& a, int lstride,
                                                         V3 o, V3 d, V3 L, V3 invL, bool want_closest, bool want_shadow, AccelCounters& cnt) {
    QueryResult res; res.t = 0.f; res.idx = -1; res.slot = -1; res.occluded = false;
    const unsigned lane = threadIdx.x & 31u;
    const float oo = fmaf(o.x, o.x, fmaf(o.y, o.y, o.z * o.z));
    CullRay2 cr;
    cr.ox2 = o.x + o.x; cr.oy2 = o.y + o.y; cr.oz2 = o.z + o.z;
    cr.ndx = -d.x; cr.ndy = -d.y; cr.ndz = -d.z;
    cr.nLx = -L.x; cr.nLy = -L.y; cr.nLz = -L.z;
    cr.noo = fmaf(RT_CULL_K, oo, -oo);
    cr.od = fmaf(o.x, d.x, fmaf(o.y, d.y, o.z * d.z));
    cr.oL = fmaf(o.x, L.x, fmaf(o.y, L.y, o.z * L.z));
    ExactCtx ex;
    ex.sA = s.sA; ex.sB = s.sB; ex.sR = s.sR; ex.G = nullptr; ex.o = o; ex.d = d; ex.L = L;
    ex.a = dot3(d, d);
    ex.aL = dot3(L, L);
    ex.ra = rcp_refined(ex.a); ex.raL = rcp_refined(ex.aL);
    if constexpr ((RAYS & RAYS_CLOSEST) == 0) want_closest = false;
    if constexpr ((RAYS & RAYS_SHADOW) == 0) want_shadow = false;
    const uint32_t en_c = want_closest ? 0xffffffffu : 0u;
    uint32_t en_s = want_shadow ? 0xffffffffu : 0u;
    uint32_t* const gm = s.scratch;                                   // one byte per L2 node
    uint32_t* const ent = s.scratch + a.gmask_words * lstride;        // entries: (mask, slot_base | 2 rot << 24 | shadow << 31)
    const uint32_t sa32 = static_cast<uint32_t>(__cvta_generic_to_shared(s.sA)), sb32 = static_cast<uint32_t>(__cvta_generic_to_shared(s.sB));

    // ---- per-ray slab constants ----
    const float dcx = o.x - a.csx, dcy = o.y - a.csy, dcz = o.z - a.csz;
    const float dmax = sqrtf(fmaf(dcx, dcx, fmaf(dcy, dcy, dcz * dcz))) + a.cs_radius;
    const float E = 3.814697265625e-06f * fmaf(dmax, dmax, a.r_max * a.r_max);                       // 2^-18
    const float pad = 1.5f * fminf(E / (2.0f * a.r_min), sqrtf(E))
                    + 1.9073486328125e-06f * (fabsf(o.x) + fabsf(o.y) + fabsf(o.z) + a.c_max);      // 2^-19
    const float neg_t_slack = -1.52587890625e-05f * dmax;                                            // 2^-16
    // a ray this lane does not ask for can never touch a box (near planes at +inf, far planes at -inf), so the box loops need no
    // per-box masking by want_closest / want_shadow
    BoxRay rc = box_ray(o, safe_inv(d.x), safe_inv(d.y), safe_inv(d.z), pad);
    BoxRay rs = box_ray(o, invL.x, invL.y, invL.z, pad);
    if (!want_closest) { rc.nx = __int_as_float(0x7f800000); rc.fx = __int_as_float(0xff800000); }
    if (!want_shadow) { rs.nx = __int_as_float(0x7f800000); rs.fx = __int_as_float(0xff800000); }

    // ---- T: L2 boxes, converged.  Both rays are tested without a branch (a ray the lane does not want touches nothing, see above).
    // With at most RT_ACCEL_FLAT_NODES L2 nodes (<= 1024 small spheres) their boxes are so large that nearly every ray touches
    // them: the level is skipped and every lane runs M over all nodes, converged and with broadcast loads. ----
    const bool flat = a.n2 <= RT_ACCEL_FLAT_NODES;
    uint32_t m2 = flat ? ((1u << a.n2) - 1u) : 0u;
    if (!flat) {
        for (int j = 0; j < a.n2; ++j) {
            const float4 b0 = s.top[2 * j], b1 = s.top[2 * j + 1];
            const bool hc = (RAYS & RAYS_CLOSEST) && box_touched(rc, b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, neg_t_slack);
            const bool hs = (RAYS & RAYS_SHADOW) && box_touched(rs, b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, neg_t_slack);
            if (hc | hs) m2 |= 1u << j;
        }
    }
    if (!(want_closest | want_shadow)) m2 = 0u;

    // ---- M: group boxes of the touched L2 nodes ----
    uint32_t m2g = 0u;
    while (m2) {
        const int j = __ffs(m2) - 1; m2 &= m2 - 1u;
        const float4* blk = reinterpret_cast<const float4*>(s.box + j * RT_ACCEL_NODE_FLOATS);
        uint32_t byte = 0u;
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {
            const float4 cx = blk[h], cy = blk[2 + h], cz = blk[4 + h], hx = blk[6 + h], hy = blk[8 + h], hz = blk[10 + h];
            const float cxa[4] = {cx.x, cx.y, cx.z, cx.w}, cya[4] = {cy.x, cy.y, cy.z, cy.w}, cza[4] = {cz.x, cz.y, cz.z, cz.w};
            const float hxa[4] = {hx.x, hx.y, hx.z, hx.w}, hya[4] = {hy.x, hy.y, hy.z, hy.w}, hza[4] = {hz.x, hz.y, hz.z, hz.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const bool hc = (RAYS & RAYS_CLOSEST) && box_touched(rc, cxa[k], cya[k], cza[k], hxa[k], hya[k], hza[k], neg_t_slack);
                const bool hs = (RAYS & RAYS_SHADOW) && box_touched(rs, cxa[k], cya[k], cza[k], hxa[k], hya[k], hza[k], neg_t_slack);
                if (hc | hs) byte |= 1u << (h * 4 + k);
            }
        }
        if (byte) {
            m2g |= 1u << j;
            reinterpret_cast<unsigned char*>(gm + (j >> 2) * lstride)[j & 3] = static_cast<unsigned char>(byte);
        }
    }

    // ---- G + E.  One copy of each in the code: G runs until this lane's groups are exhausted or its entry list is
    // full, E drains the list; the outer loop repeats only after an overflow (rare). ----
    constexpr int kGS = 1 << GSHIFT, kPPG = kGS / 2;                  // slots / sphere pairs of one small group
    const int rot_lane = static_cast<int>(lane & static_cast<unsigned>(kPPG - 1));
    int ne = 0;
    // big groups (every lane, converged, broadcast loads): only the pairs that hold spheres -- 2 of 16 in the default scene
    for (int bg = 0; bg < a.n_big_groups; ++bg) {
        const int np_g = min(16, a.n_big_pairs - 16 * bg);
        uint32_t mc = 0u, ms = 0u;
#pragma unroll 2
        for (int q = 0; q < np_g; ++q) cull_pair2<RAYS>(cr, s.sA[bg * 16 + q], s.sB[bg * 16 + q], mc, ms);
        const uint32_t valid = 0xffffffffu << (32 - 2 * np_g);           // bit 31 = first sphere of the group, as everywhere
        const uint32_t cc = ~(mc << (32 - 2 * np_g)) & valid & en_c, cs = ~(ms << (32 - 2 * np_g)) & valid & en_s;
        if (cc) { ent[(2 * ne) * lstride] = cc; ent[(2 * ne + 1) * lstride] = static_cast<uint32_t>(bg * 32); ++ne; }
        if (cs) { ent[(2 * ne) * lstride] = cs; ent[(2 * ne + 1) * lstride] = static_cast<uint32_t>(bg * 32) | 0x80000000u; ++ne; }
    }
    uint32_t cur8 = 0u; int curj = 0;
    bool more = true;
    while (more) {
        for (;;) {
            if (ne > RT_ACCEL_ENTRY_CAP - 2) break;         // a group adds at most two entries
            if (cur8 == 0u) {
                if (m2g == 0u) { more = false; break; }
                curj = __ffs(m2g) - 1; m2g &= m2g - 1u;
                cur8 = reinterpret_cast<const unsigned char*>(gm + (curj >> 2) * lstride)[curj & 3];
            }
            const int k = __ffs(cur8) - 1; cur8 &= cur8 - 1u;
            const int slot_base = a.n_big_groups * 32 + (curj * 8 + k) * kGS, rot = rot_lane;
            cnt.group_visits += (en_c & 1u) + (en_s & 1u);
            // pair q of this step is pair (q ^ rot) of the group: an XOR on a (16 kPPG)-byte aligned group is one LOP3 per address, and
            // inside a quarter-warp (8 consecutive lanes) the 8 addresses fall into 8 different 16-byte bank groups.  The
            // candidate masks stay in visiting order; E undoes the permutation per candidate (slot = base + (bit ^ 2 rot)).
            const uint32_t pa = sa32 + static_cast<uint32_t>(slot_base) * 8u + (static_cast<uint32_t>(rot) << 4);
            const uint32_t pb = sb32 + static_cast<uint32_t>(slot_base) * 8u + (static_cast<uint32_t>(rot) << 4);
            uint32_t mc = 0u, ms = 0u;
#pragma unroll
            for (int q = 0; q < kPPG; ++q) cull_pair2<RAYS>(cr, lds128(pa ^ (q << 4)), lds128(pb ^ (q << 4)), mc, ms);
            // bit 31 = first slot of the group, as everywhere; a group smaller than 32 slots leaves the low bits clear
            const uint32_t gvalid = 0xffffffffu << (32 - kGS);
            const uint32_t cc = ~(mc << (32 - kGS)) & gvalid & en_c, cs = ~(ms << (32 - kGS)) & gvalid & en_s;
            const uint32_t meta = static_cast<uint32_t>(slot_base) | (static_cast<uint32_t>(rot) << 25);
            if (cc) { ent[(2 * ne) * lstride] = cc; ent[(2 * ne + 1) * lstride] = meta; ++ne; }
            if (cs) { ent[(2 * ne) * lstride] = cs; ent[(2 * ne + 1) * lstride] = meta | 0x80000000u; ++ne; }
        }
        int e = 0; uint32_t cur = 0u, meta = 0u;
        for (;;) {
            if (cur == 0u) {
                if (e >= ne) break;
                cur = ent[(2 * e) * lstride]; meta = ent[(2 * e + 1) * lstride]; ++e;
                if ((meta >> 31) && en_s == 0u) { cur = 0u; continue; }
            }
            const int j = __clz(cur); cur &= ~(0x80000000u >> j);
            const bool sh = (meta >> 31) != 0u;
            const int slot = static_cast<int>(meta & 0xffffffu) + (j ^ static_cast<int>((meta >> 24) & 0x7eu));
            float t;
            if (exact_candidate(ex, slot, sh, t)) {
                if (sh) { res.occluded = true; en_s = 0u; cur = 0u; }
                else { const int idx = s.sI[slot]; if (closer(t, idx, res)) { res.t = t; res.idx = idx; res.slot = slot; } }
            }
        }
        ne = 0;
    }
    return res;
}

}  // namespace rt
