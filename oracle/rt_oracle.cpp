/*
 * oracle/rt_oracle.cpp -- CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).  See rt_oracle.h.
 *
 * Scalar restatement of the SakibSaikia/CPURayTracer `src/spheres` render path with a PINNED
 * operation order: every float operation below is a single IEEE-754 binary32 operation
 * (+ - * / sqrt, or an explicit fmaf), evaluated in the order DirectXMath's SSE2 code paths
 * evaluate them (no FMA contraction: the reference is built /arch:SSE2, common-lib.vcxproj:72).
 * Compile with -ffp-contract=off (oracle/Makefile does).  Citations: /root/reference/src/...
 *
 * Deliberate, documented deviations from the reference binary (all below 2 ulp per call):
 *   D1  pow(x,5) in the Schlick terms (material.cpp:27,79; light.cpp:37) is (x*x)*(x*x)*x.
 *   D2  cos/sin of the Halton angles (quasi-random.cpp:41-49,54-60) use the pinned
 *       Cody-Waite + polynomial routine sincos_pinned() so CPU and GPU agree bit for bit.
 *   D3  XMVectorReciprocalEst(ior) (material.cpp:132, rcpps, ~12 bit, CPU-vendor specific)
 *       is the exact 1.0f/ior.
 *   D4  ORC_INTEGRATOR_ITERATIVE unrolls GetHitColor's tail recursion (spheres-app.cpp:238-257)
 *       into L += T*S; T *= a.  ORC_INTEGRATOR_RECURSIVE keeps the reference's nesting.
 *   D5  ORC_RNG_COUNTER replaces the per-material atomic Halton counters (material.h:34,50-51,67)
 *       by a stateless pcg4d hash of (pixel, sample, bounce, seed) -- BASELINE.json sanctions it.
 *       ORC_RNG_HALTON keeps the reference's counters (deterministic only with threads == 1).
 *   D6  ORC_ACCEL_BRUTE tests every sphere in index order, strict '<' so the lowest index wins
 *       ties; ORC_ACCEL_BVH restates BvhNode (ray-tracing.cpp:107-214) with a fixed-seed axis
 *       choice instead of the unseeded std::rand() (ray-tracing.cpp:123).
 */
#include "rt_oracle.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstring>
#include <limits>
#include <memory>
#include <thread>
#include <vector>
#if defined(__SSE__)
#include <xmmintrin.h>
#endif

namespace {

struct V3 { float x, y, z; };
static inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
static inline V3 v3(const float* p) { return V3{p[0], p[1], p[2]}; }
static inline V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline V3 operator*(V3 a, V3 b) { return V3{a.x * b.x, a.y * b.y, a.z * b.z}; }
static inline V3 operator*(float s, V3 a) { return V3{s * a.x, s * a.y, s * a.z}; }   /* XMVectorScale */
static inline V3 operator/(V3 a, float s) { return V3{a.x / s, a.y / s, a.z / s}; }   /* XMVectorDivide */
static inline V3 neg(V3 a) { return V3{-a.x, -a.y, -a.z}; }
static inline void st3(float* p, V3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }

/* XMVector3Dot, SSE2 path: mul, then (x + y) + z. */
static inline float dot3(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
/* XMVector3Cross: (a.y b.z - a.z b.y, a.z b.x - a.x b.z, a.x b.y - a.y b.x), unfused. */
static inline V3 cross3(V3 a, V3 b) {
    return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
/* XMVector3Normalize, SSE2 path: v / sqrt(dot); zero length -> 0; infinite length -> NaN. */
static inline V3 normalize3(V3 v) {
    const float l2 = dot3(v, v);
    const float l = std::sqrt(l2);
    if (l == 0.0f) return V3{0.f, 0.f, 0.f};
    if (l2 == std::numeric_limits<float>::infinity()) {
        const float q = std::numeric_limits<float>::quiet_NaN();
        return V3{q, q, q};
    }
    return V3{v.x / l, v.y / l, v.z / l};
}
static inline float saturate(float x) { return std::fmin(std::fmax(x, 0.0f), 1.0f); }  /* XMVectorSaturate */
/* XMVector3Reflect: I - (2 * dot(I,N)) * N, unfused. */
static inline V3 reflect3(V3 i, V3 n) {
    float d = dot3(i, n);
    d = d + d;
    return V3{i.x - d * n.x, i.y - d * n.y, i.z - d * n.z};
}
/* XMVector3RefractV, SSE2 path. Returns the zero vector on total internal reflection. */
static inline V3 refract3(V3 i, V3 n, float eta) {
    const float idn = dot3(i, n);
    float r = idn * idn;
    r = 1.0f - r;
    r = r * eta;
    r = r * eta;
    r = 1.0f - r;
    if (r <= 0.0f) return V3{0.f, 0.f, 0.f};
    r = std::sqrt(r);
    r = r + eta * idn;
    return V3{eta * i.x - r * n.x, eta * i.y - r * n.y, eta * i.z - r * n.z};
}
/* XMFresnelTerm, SSE2 path. */
static inline float fresnel_term(float c, float ior) {
    float g = ior * ior;
    float t = c * c;
    g = g - 1.0f;
    t = t + g;
    g = std::fmax(0.0f - t, t);
    g = std::sqrt(g);
    float gac = g + c, gsc = g - c;
    float res = gsc * gsc;
    t = gac * gac;
    res = res * 0.5f;
    res = res / t;
    gac = gac * c;
    gsc = gsc * c;
    gac = gac - 1.0f;
    gsc = gsc + 1.0f;
    gac = gac * gac;
    gsc = gsc * gsc;
    gac = gac / gsc;
    gac = gac + 1.0f;
    res = res * gac;
    return std::fmin(std::fmax(res, 0.0f), 1.0f);
}
/* Math mode (test-only global, constant during a render): 0 = pinned (D1-D3 below hold, CPU == GPU bit for bit),
 * 1 = reference libm: powf(x,5), libm cosf/sinf and rcpps exactly as the reference binary calls them, so that this
 * restatement can be compared bit for bit with oracle/_ref (the reference's own sources). */
static int g_math_ref = 0;
static inline float pow5(float x) {                                                            /* D1 */
    if (g_math_ref) return std::pow(x, 5.f);                                                   /* XMVectorPow = powf per lane */
    const float x2 = x * x; return (x2 * x2) * x;
}
static inline float rcp_ior(float ior) {                                                       /* D3 */
#if defined(__SSE__)
    if (g_math_ref) return _mm_cvtss_f32(_mm_rcp_ss(_mm_set_ss(ior)));                         /* XMVectorReciprocalEst = rcpps */
#endif
    return 1.0f / ior;
}
/* f0 + (1 - f0) * (1 - nDotV)^5, material.cpp:27 / light.cpp:37. */
static inline float schlick(float f0, float n_dot_v) { return f0 + (1.0f - f0) * pow5(1.0f - n_dot_v); }

/* D2: pinned sin/cos for theta in [0, 2*pi].  Two-constant Cody-Waite reduction by pi/2 with
 * explicit fused multiply-adds, cephes minimax polynomials on [-pi/4, pi/4], Horner with fmaf. */
static inline void sincos_pinned(float theta, float* s_out, float* c_out) {
    if (g_math_ref) { *s_out = std::sin(theta); *c_out = std::cos(theta); return; }              /* quasi-random.cpp:46-47,58-59 */
    const float q = std::nearbyint(theta * 0.63661975f);          /* 2/pi, round-to-nearest-even */
    float r = std::fmaf(-q, 1.57079637050628662109375f, theta);    /* pi/2 high part */
    r = std::fmaf(-q, -4.37113900018624283e-8f, r);                /* pi/2 low part  */
    const float r2 = r * r;
    float sp = std::fmaf(r2, -1.9515295891e-4f, 8.3321608736e-3f);
    sp = std::fmaf(sp, r2, -1.6666654611e-1f);
    const float s = std::fmaf(sp * r2, r, r);
    float cp = std::fmaf(r2, 2.443315711809948e-5f, -1.388731625493765e-3f);
    cp = std::fmaf(cp, r2, 4.166664568298827e-2f);
    const float c = std::fmaf(cp, r2 * r2, std::fmaf(-0.5f, r2, 1.0f));
    const int k = static_cast<int>(q) & 3;
    switch (k) {
        case 0: *s_out = s;  *c_out = c;  break;
        case 1: *s_out = c;  *c_out = -s; break;
        case 2: *s_out = -s; *c_out = -c; break;
        default: *s_out = -c; *c_out = s; break;
    }
}

/* Random::HaltonSample, quasi-random.cpp:3-16 (float accumulation, uint64 digits). */
static inline float halton(uint64_t index, uint32_t base) {
    float result = 0.f, f = 1.f;
    const float fb = static_cast<float>(base);
    while (index > 0) {
        f = f / fb;
        result += f * static_cast<float>(index % base);
        index = index / base;
    }
    return result;
}
static const float k2Pi = 2.0f * 3.141592654f;   /* 2 * XM_PI, quasi-random.cpp:42,54 */
/* Random::HaltonSampleDisk, quasi-random.cpp:52-61: theta from base1, r = base2 sample (not sqrt). */
static inline void disk_from_u(float u_theta, float u_r, float* x, float* y) {
    const float theta = k2Pi * u_theta;
    float s, c; sincos_pinned(theta, &s, &c);
    *x = u_r * c; *y = u_r * s;
}
/* Random::HaltonSampleHemisphere, quasi-random.cpp:36-50: z = u1 (uniform solid angle). */
static inline V3 hemisphere_from_u(float u1, float u2) {
    const float r = std::sqrt(1.f - u1 * u1);
    const float phi = k2Pi * u2;
    float s, c; sincos_pinned(phi, &s, &c);
    return V3{r * c, r * s, u1};
}

/* XMCOLOR(float,float,float,float) -> XMStoreColor: saturate, *255, cvtps2dq (nearest-even). */
static inline uint8_t xmcolor_quantize(float x) {
    float v = std::fmin(std::fmax(x, 0.0f), 1.0f);
    v = v * 255.0f;
    return static_cast<uint8_t>(static_cast<int>(std::nearbyint(v)));
}
/* XMLoadColor: float(k) * (1/255) with the constant rounded to float first. */
static inline float xmcolor_load(uint8_t k) { return static_cast<float>(k) * (1.0f / 255.0f); }

/* D5: pcg4d (Jarzynski & Olano 2020) of (pixel, sample_lo, bounce, seed ^ sample_hi). */
static inline void pcg4d(uint32_t v[4]) {
    for (int i = 0; i < 4; ++i) v[i] = v[i] * 1664525u + 1013904223u;
    v[0] += v[1] * v[3]; v[1] += v[2] * v[0]; v[2] += v[0] * v[1]; v[3] += v[1] * v[2];
    for (int i = 0; i < 4; ++i) v[i] ^= v[i] >> 16;
    v[0] += v[1] * v[3]; v[1] += v[2] * v[0]; v[2] += v[0] * v[1]; v[3] += v[1] * v[2];
}
static inline void rng_counter(uint32_t pixel, uint64_t sample, uint32_t bounce, uint32_t seed, float out[4]) {
    uint32_t v[4] = {pixel, static_cast<uint32_t>(sample), bounce, seed ^ static_cast<uint32_t>(sample >> 32)};
    pcg4d(v);
    for (int i = 0; i < 4; ++i) out[i] = static_cast<float>(v[i] >> 8) * (1.0f / 16777216.0f);
}

/* ---------------- camera (camera.cpp) ---------------- */
static void camera_make(const orc_camera_params* p, orc_camera* c) {
    const V3 origin = v3(p->origin), look = v3(p->look_at);
    const float theta = p->vfov_deg * 3.141592654f / 180.f;              /* camera.cpp:14 */
    const float half_h = std::tan(theta / 2.f);                          /* :15 */
    const float half_w = p->aspect * half_h;                             /* :16 */
    const V3 up = v3(0.f, 1.f, 0.f);
    const V3 w = normalize3(look - origin);                              /* :20 */
    const V3 u = normalize3(cross3(up, w));                              /* :21 */
    const V3 v = cross3(w, u);                                           /* :22 */
    st3(c->origin, origin);
    st3(c->image_origin, origin + 1.f * w);                              /* :24-25 */
    st3(c->x, half_w * u);                                               /* :26 */
    st3(c->y, half_h * v);                                               /* :27 */
    c->aperture = p->aperture; c->focus = p->focus;
}
struct Ray { V3 o, d; };
static inline Ray camera_get_ray(const orc_camera& c, float u, float v, float lx, float ly) {
    const float ndc_x = 2.f * u - 1.f;                                   /* camera.cpp:34 */
    const float ndc_y = -2.f * v + 1.f;                                  /* :35 */
    const V3 org = v3(c.origin), mx = v3(c.x), my = v3(c.y);
    const V3 p = (v3(c.image_origin) + ndc_x * mx) + ndc_y * my;         /* :38 */
    const V3 focal = org + c.focus * normalize3(p - org);                /* :39 */
    const float rdx = 0.5f * c.aperture * lx;                            /* :43 */
    const float rdy = 0.5f * c.aperture * ly;                            /* :44 */
    const V3 o2 = (org + rdx * mx) + rdy * my;                           /* :45 */
    return Ray{o2, normalize3(focal - o2)};                              /* :47 */
}

/* ---------------- geometry (ray-tracing.cpp) ---------------- */
struct Hit { float t; V3 pos, n; float u, v; int sphere; };

static inline bool sphere_intersect(const orc_sphere& s, int index, const Ray& r, Hit& h) {
    const V3 ctr = v3(s.cx, s.cy, s.cz);
    const V3 oc = r.o - ctr;                                             /* ray-tracing.cpp:44 */
    const float a = dot3(r.d, r.d);                                      /* :46 */
    const float b = dot3(oc, r.d);                                       /* :47 */
    const float c = dot3(oc, oc) - s.radius * s.radius;                  /* :48 */
    const float disc = b * b - a * c;                                    /* :50 */
    if (disc > 0.f) {                                                    /* :54 strict */
        const float sq = std::sqrt(disc);
        float t = (-b - sq) / a;                                         /* :56 */
        if (!(t > 0.001f)) {                                             /* :58 */
            t = (-b + sq) / a;                                           /* :69 */
            if (!(t > 0.001f)) return false;                             /* :71 */
        }
        h.t = t;
        h.pos = V3{t * r.d.x + r.o.x, t * r.d.y + r.o.y, t * r.d.z + r.o.z};   /* :61, mul then add */
        h.n = (h.pos - ctr) / s.radius;                                  /* :62 */
        h.u = 0.5f * h.n.x + 0.5f;                                       /* :36 (ComputeUV uses x and z) */
        h.v = 0.5f * h.n.z + 0.5f;                                       /* :37 */
        h.sphere = index;
        return true;
    }
    return false;
}

/* BVH restated from ray-tracing.cpp:96-214 + DirectXCollision BoundingBox. */
struct Box { V3 c, e; };
struct Node { Box box; int left, right; };  /* child >= 0: node index; child < 0: sphere ~child; right == INT_MIN: none */
static const int kNoChild = std::numeric_limits<int>::min();

static inline Box box_merge(const Box& a, const Box& b) {            /* BoundingBox::CreateMerged */
    const V3 amin = a.c - a.e, amax = a.c + a.e, bmin = b.c - b.e, bmax = b.c + b.e;
    const V3 mn{std::fmin(amin.x, bmin.x), std::fmin(amin.y, bmin.y), std::fmin(amin.z, bmin.z)};
    const V3 mx{std::fmax(amax.x, bmax.x), std::fmax(amax.y, bmax.y), std::fmax(amax.z, bmax.z)};
    return Box{0.5f * (mn + mx), 0.5f * (mx - mn)};
}
static inline bool box_hit(const Box& bx, const Ray& r) {             /* BoundingBox::Intersects(o, d, dist) */
    const float o[3] = {bx.c.x - r.o.x, bx.c.y - r.o.y, bx.c.z - r.o.z};
    const float d[3] = {r.d.x, r.d.y, r.d.z};
    const float e[3] = {bx.e.x, bx.e.y, bx.e.z};
    float tmin = -std::numeric_limits<float>::max(), tmax = std::numeric_limits<float>::max();
    bool reject = false;
    for (int k = 0; k < 3; ++k) {
        const bool parallel = std::fabs(d[k]) <= 1e-20f;
        const float inv = 1.0f / d[k];
        const float t1 = (o[k] - e[k]) * inv, t2 = (o[k] + e[k]) * inv;
        const float lo = parallel ? -std::numeric_limits<float>::max() : std::fmin(t1, t2);
        const float hi = parallel ? std::numeric_limits<float>::max() : std::fmax(t1, t2);
        tmin = std::fmax(tmin, lo);
        tmax = std::fmin(tmax, hi);
        const bool in_bounds = (o[k] <= e[k]) && (o[k] >= -e[k]);
        if (parallel && !in_bounds) reject = true;
    }
    if (tmin > tmax) reject = true;
    if (tmax < 0.f) reject = true;
    return !reject;
}

struct Bvh {
    std::vector<Node> nodes;
    int root = -1;
    const orc_sphere* spheres = nullptr;
    uint32_t lcg = 12345u;                                               /* D6: fixed-seed axis choice */

    Box child_box(int child) const {
        if (child >= 0) return nodes[child].box;
        const orc_sphere& s = spheres[~child];
        return Box{v3(s.cx, s.cy, s.cz), v3(s.radius, s.radius, s.radius)};   /* Sphere::GetAABB :86-94 */
    }
    int build(std::vector<int>& items, size_t begin, size_t end) {        /* BvhNode::BvhNode :107-167 */
        const size_t n = end - begin;
        Node nd{};
        if (n == 1) { nd.left = items[begin]; nd.right = kNoChild; }
        else if (n == 2) { nd.left = items[begin]; nd.right = items[begin + 1]; }
        else {
            lcg = lcg * 1103515245u + 12345u;
            const int axis = static_cast<int>((lcg >> 16) % 3u);
            std::stable_sort(items.begin() + begin, items.begin() + end, [&](int a, int b) {
                const Box ba = child_box(a), bb = child_box(b);
                const V3 ma = ba.c - ba.e, mb = bb.c - bb.e;
                return axis == 0 ? ma.x < mb.x : axis == 1 ? ma.y < mb.y : ma.z < mb.z;
            });
            nd.left = build(items, begin, begin + n / 2);
            nd.right = build(items, begin + n / 2, end);
        }
        nd.box = (nd.right != kNoChild) ? box_merge(child_box(nd.left), child_box(nd.right)) : child_box(nd.left);
        nodes.push_back(nd);
        return static_cast<int>(nodes.size()) - 1;
    }
    void make(const orc_scene* sc) {
        spheres = sc->spheres;
        nodes.clear(); nodes.reserve(2 * static_cast<size_t>(sc->n_spheres));
        std::vector<int> items(sc->n_spheres);
        for (int i = 0; i < sc->n_spheres; ++i) items[i] = ~i;
        root = sc->n_spheres > 0 ? build(items, 0, items.size()) : -1;
    }
    bool hit_child(int child, const Ray& r, Hit& h) const {
        if (child >= 0) return hit_node(child, r, h);
        return sphere_intersect(spheres[~child], ~child, r, h);
    }
    bool hit_node(int ni, const Ray& r, Hit& h) const {                   /* BvhNode::Intersect :174-214 */
        const Node& nd = nodes[ni];
        if (!box_hit(nd.box, r)) return false;
        Hit lh, rh;
        const bool l = hit_child(nd.left, r, lh);
        const bool rr = nd.right != kNoChild ? hit_child(nd.right, r, rh) : false;
        if (l && rr) { h = (lh.t < rh.t) ? lh : rh; return true; }        /* :184 strict, right wins ties */
        if (l) { h = lh; return true; }
        if (rr) { h = rh; return true; }
        return false;
    }
};

/* ---------------- shading state ---------------- */
struct Counters { std::atomic<uint64_t> sample{0}, coin{0}; };

struct World {
    const orc_scene* sc;
    orc_camera cam;
    Bvh bvh;
    int accel, rng_mode, integrator, max_depth;
    uint32_t seed;
    std::unique_ptr<Counters[]> counters;       /* per material instance, material.h:34,50-51,67 */
    std::vector<V3> light_dir;                  /* DirectionalLight::m_direction = XMVector3Normalize(dir), light.cpp:7 */
    V3 sky;
    std::atomic<uint64_t> closest_rays{0}, shadow_rays{0};
};
struct RayCount { uint64_t closest = 0, shadow = 0; };

static inline bool closest_hit(const World& w, const Ray& r, Hit& h) {
    if (w.accel == ORC_ACCEL_BVH) return w.bvh.root >= 0 && w.bvh.hit_node(w.bvh.root, r, h);
    bool any = false; Hit cur;
    for (int i = 0; i < w.sc->n_spheres; ++i)
        if (sphere_intersect(w.sc->spheres[i], i, r, cur) && (!any || cur.t < h.t)) { h = cur; any = true; }
    return any;
}
static inline void texture_eval(const orc_texture& t, float u, float v, float rgba[4]) {
    const float* c = t.c0;
    if (t.kind == ORC_TEX_CHECKER) {                                     /* texture.cpp:20-33 */
        const int iu = static_cast<int>(t.tiling * u), iv = static_cast<int>(t.tiling * v);
        c = (iu % 2 == iv % 2) ? t.c0 : t.c1;
    }
    rgba[0] = c[0]; rgba[1] = c[1]; rgba[2] = c[2]; rgba[3] = c[3];
}
static inline V3 texture_rgb(const World& w, int tex, float u, float v) {
    float c[4]; texture_eval(w.sc->textures[tex], u, v, c); return v3(c);
}

/* Material::Shade -> DirectionalLight::Shade, material.cpp:4-13, light.cpp:11-42. */
static inline V3 shade(const World& w, const Hit& h, V3 view_origin, RayCount& rc) {
    const orc_material& m = w.sc->materials[w.sc->spheres[h.sphere].material];
    V3 total{0.f, 0.f, 0.f};
    for (int li = 0; li < w.sc->n_lights; ++li) {
        const orc_light& lt = w.sc->lights[li];
        const V3 L = w.light_dir[li];                                     /* m_direction, normalised at light.cpp:7 */
        Hit dummy;
        ++rc.shadow;
        if (closest_hit(w, Ray{h.pos, L}, dummy)) continue;               /* light.cpp:13-18 (adds zero) */
        V3 albedo{0.f, 0.f, 0.f}, f0{0.04f, 0.04f, 0.04f};                /* material.h:27-29,43-45,60-62 */
        if (m.kind == ORC_MAT_OPAQUE) albedo = texture_rgb(w, m.texture, h.u, h.v);
        else if (m.kind == ORC_MAT_METAL) f0 = texture_rgb(w, m.texture, h.u, h.v);
        const float s = m.smoothness;
        const float n_dot_l = saturate(dot3(h.n, L));                     /* light.cpp:26 */
        const V3 lc = v3(lt.color);
        const V3 radiance_in{(lt.luminance * lc.x) * n_dot_l, (lt.luminance * lc.y) * n_dot_l, (lt.luminance * lc.z) * n_dot_l}; /* :27 */
        const V3 view = normalize3(view_origin - h.pos);                  /* :33 */
        const V3 half = normalize3(L + view);                             /* :34 */
        const float n_dot_h = saturate(dot3(h.n, half));                  /* :35 */
        const float n_dot_v = saturate(dot3(view, h.n));                  /* :36 */
        const float p = pow5(1.0f - n_dot_v);
        const float pw = std::pow(n_dot_h, s);                            /* :38 XMVectorPow = powf */
        const float k = s + 8.f;
        const V3 refl{f0.x + (1.0f - f0.x) * p, f0.y + (1.0f - f0.y) * p, f0.z + (1.0f - f0.z) * p};  /* :37 */
        const V3 spec{((refl.x * 0.125f) * k) * pw, ((refl.y * 0.125f) * k) * pw, ((refl.z * 0.125f) * k) * pw}; /* :38 */
        total = total + radiance_in * (albedo + spec);                    /* :40 */
    }
    return total;
}

/* Material::Scatter for the three surface materials, material.cpp:20-164. */
struct Rand3 { float xi, u1, u2; };
static inline bool scatter_with(const World& w, const Ray& r, const Hit& h, const Rand3& rnd, V3& atten, Ray& out) {
    const orc_material& m = w.sc->materials[w.sc->spheres[h.sphere].material];
    if (m.kind == ORC_MAT_OPAQUE) {
        const float ndv = dot3(neg(r.d), h.n);
        if (!(ndv > 0.f)) return false;                                   /* material.cpp:22 */
        const float refl = schlick(0.04f, saturate(ndv));                 /* :25-27 */
        if (refl > rnd.xi) {                                              /* :29-30 */
            atten = v3(1.f, 1.f, 1.f);
            out = Ray{h.pos, normalize3(reflect3(r.d, h.n))};             /* :35-36 */
            return true;
        }
        atten = texture_rgb(w, m.texture, h.u, h.v);                      /* :41 */
        const V3 dir = hemisphere_from_u(rnd.u1, rnd.u2);                 /* :44 */
        const V3 b3 = h.n;
        const V3 up = std::fabs(b3.x) < 0.5f ? v3(1.f, 0.f, 0.f) : v3(0.f, 1.f, 0.f);   /* :50 */
        const V3 b1 = cross3(up, b3);                                     /* :51 (not normalised) */
        const V3 b2 = cross3(b3, b1);                                     /* :52 */
        const V3 sd = (dir.x * b1 + dir.y * b2) + dir.z * b3;             /* :55 */
        out = Ray{h.pos, normalize3(sd)};                                 /* :56 */
        return true;
    }
    if (m.kind == ORC_MAT_METAL) {
        const float ndv = dot3(neg(r.d), h.n);
        if (!(ndv > 0.f)) return false;                                   /* :74 */
        /* :77-85: reflectance = f0 + (1 - f0) (1 - n.v)^5 on all four lanes of the texture colour; reflect when ANY lane
         * exceeds the random number.  Lane w is the texture's alpha: 1 in every scene the reference builds (then that lane is
         * exactly 1 > rand: always reflects), but any alpha is representable. */
        float c4[4]; texture_eval(w.sc->textures[m.texture], h.u, h.v, c4);
        const float nv = saturate(ndv);
        if (!(schlick(c4[0], nv) > rnd.xi || schlick(c4[1], nv) > rnd.xi || schlick(c4[2], nv) > rnd.xi || schlick(c4[3], nv) > rnd.xi)) return false;
        atten = texture_rgb(w, m.texture, h.u, h.v);                      /* :87 */
        out = Ray{h.pos, normalize3(reflect3(r.d, h.n))};                 /* :89-90 */
        return true;
    }
    /* DielectricTransparent, :111-164 */
    atten = v3(1.f, 1.f, 1.f);
    V3 outward; float eta, cos_i;
    const float dn = dot3(r.d, h.n);
    if (dn > 0.f) { outward = neg(h.n); eta = m.ior; cos_i = dn; }        /* :121-127 */
    else { outward = h.n; eta = rcp_ior(m.ior); cos_i = dot3(r.d, neg(h.n)); }   /* :128-134, D3 */
    const V3 refr = refract3(r.d, outward, eta);                          /* :137 */
    const bool can_refract = (refr.x != 0.f) || (refr.y != 0.f) || (refr.z != 0.f);   /* :138 */
    const float prob = can_refract ? fresnel_term(cos_i, m.ior) : 1.0f;   /* :140-149 */
    if (prob > rnd.xi) out = Ray{h.pos, normalize3(reflect3(r.d, h.n))};  /* :153-157 */
    else out = Ray{h.pos, normalize3(refr)};                              /* :161 */
    return true;
}
/* Draws the random numbers the way the selected RNG mode dictates, then scatters. */
static inline bool scatter(World& w, const Ray& r, const Hit& h, uint32_t pixel, uint64_t sample, uint32_t bounce, V3& atten, Ray& out) {
    Rand3 rnd{0.f, 0.f, 0.f};
    if (w.rng_mode == ORC_RNG_COUNTER) {
        float u[4]; rng_counter(pixel, sample, bounce, w.seed, u);
        rnd = Rand3{u[0], u[1], u[2]};
        return scatter_with(w, r, h, rnd, atten, out);
    }
    /* Reference ordering of the atomic post-increments (material.cpp:29,44,82,151). */
    const int mi = w.sc->spheres[h.sphere].material;
    const orc_material& m = w.sc->materials[mi];
    Counters& c = w.counters[mi];
    if (m.kind == ORC_MAT_OPAQUE) {
        if (!(dot3(neg(r.d), h.n) > 0.f)) return false;
        rnd.xi = halton(c.coin.fetch_add(1, std::memory_order_relaxed), 3);
        const float refl = schlick(0.04f, saturate(dot3(neg(r.d), h.n)));
        if (!(refl > rnd.xi)) {
            const uint64_t k = c.sample.fetch_add(1, std::memory_order_relaxed);
            rnd.u1 = halton(k, 5); rnd.u2 = halton(k, 7);
        }
    } else if (m.kind == ORC_MAT_METAL) {
        if (!(dot3(neg(r.d), h.n) > 0.f)) return false;
        rnd.xi = halton(c.coin.fetch_add(1, std::memory_order_relaxed), 3);
    } else {
        rnd.xi = halton(c.sample.fetch_add(1, std::memory_order_relaxed), 7);
    }
    return scatter_with(w, r, h, rnd, atten, out);
}

/* SpheresApp::GetHitColor, spheres-app.cpp:238-257, as written (recursive). */
static V3 hit_color_recursive(World& w, const Ray& r, int depth, uint32_t pixel, uint64_t sample, RayCount& rc) {
    Hit h;
    ++rc.closest;
    if (closest_hit(w, r, h)) {
        V3 atten{0.f, 0.f, 0.f}; Ray out{};
        const bool scattered = scatter(w, r, h, pixel, sample, static_cast<uint32_t>(depth), atten, out);   /* :246 */
        const bool recurse = depth < w.max_depth && scattered;            /* :247 */
        const V3 emit{0.f, 0.f, 0.f};                                     /* :249, surface materials emit 0 */
        const V3 direct = shade(w, h, v3(w.cam.origin), rc);              /* :250 */
        const V3 indirect = recurse ? atten * hit_color_recursive(w, out, depth + 1, pixel, sample, rc) : v3(0.f, 0.f, 0.f);
        return (emit + direct) + indirect;                                /* :249-251 */
    }
    return w.sky;                                                         /* :255 */
}
/* D4: the same integral as a loop carrying the throughput. */
static V3 hit_color_iterative(World& w, Ray r, uint32_t pixel, uint64_t sample, RayCount& rc) {
    V3 L{0.f, 0.f, 0.f}, T{1.f, 1.f, 1.f};
    for (int depth = 0;; ++depth) {
        Hit h;
        ++rc.closest;
        if (!closest_hit(w, r, h)) { L = L + T * w.sky; break; }
        V3 atten{0.f, 0.f, 0.f}; Ray out{};
        const bool cont = depth < w.max_depth;
        /* The reference scatters even at the depth limit (:246 precedes :247); with the counter RNG
         * that call has no side effect, with Halton it advances the counters -- keep it there. */
        bool scattered = false;
        if (cont || w.rng_mode == ORC_RNG_HALTON)
            scattered = scatter(w, r, h, pixel, sample, static_cast<uint32_t>(depth), atten, out);
        const V3 direct = shade(w, h, v3(w.cam.origin), rc);
        L = L + T * direct;
        if (!(cont && scattered)) break;
        T = T * atten;
        r = out;
    }
    return L;
}

static void world_lights(World& w) {
    w.light_dir.clear();
    for (int i = 0; i < w.sc->n_lights; ++i) w.light_dir.push_back(normalize3(v3(w.sc->lights[i].dir)));
}
static void world_init(World& w, const orc_scene* sc, const orc_camera_params* cp, const orc_config* cfg) {
    w.sc = sc;
    world_lights(w);
    camera_make(cp, &w.cam);
    w.accel = cfg->accel; w.rng_mode = cfg->rng_mode; w.integrator = cfg->integrator;
    w.max_depth = cfg->max_depth; w.seed = cfg->seed;
    if (w.accel == ORC_ACCEL_BVH) w.bvh.make(sc);
    w.counters.reset(new Counters[std::max(1, sc->n_materials)]);
    float c[4]; texture_eval(sc->textures[sc->sky_texture], 0.f, 0.f, c);
    w.sky = sc->sky_luminance * v3(c);                                    /* material.cpp:172-175 */
}

/* SpheresApp::GenerateRays for one pixel, spheres-app.cpp:132-161. */
static inline Ray primary_ray(const World& w, const orc_config* cfg, uint64_t n, int i, int j, float jx, float jy) {
    const float xsize = static_cast<float>(cfg->width), ysize = static_cast<float>(cfg->height);
    const float u = (static_cast<float>(i) + jx) / xsize;                 /* :149 */
    const float v = (static_cast<float>(j) + jy) / ysize;                 /* :150 */
    float lx = 0.f, ly = 0.f;
    if (!(cfg->flags & ORC_FLAG_LENS_OFF)) {
        const uint64_t k = n + static_cast<uint64_t>(i) + static_cast<uint64_t>(j);   /* :152 */
        disk_from_u(halton(k, 4), halton(k, 5), &lx, &ly);
    }
    return camera_get_ray(w.cam, u, v, lx, ly);                           /* :154 */
}

template <class F> static void parallel_rows(int height, int threads, F&& fn) {
    if (threads <= 1) { for (int j = 0; j < height; ++j) fn(j); return; }
    std::atomic<int> next{0};
    std::vector<std::thread> pool;
    pool.reserve(threads);
    for (int t = 0; t < threads; ++t)
        pool.emplace_back([&] { for (int j; (j = next.fetch_add(1)) < height;) fn(j); });
    for (auto& th : pool) th.join();
}

}  // namespace

/* ======================= exported entry points ======================= */
extern "C" {

int orc_render(const orc_scene* sc, const orc_camera_params* cp, const orc_config* cfg, float* hdr, orc_stats* stats) {
    if (!sc || !cp || !cfg || !hdr || cfg->width <= 0 || cfg->height <= 0 || cfg->spp < 0 || cfg->first_sample < 1) return 1;
    g_math_ref = cfg->math_mode;
    World w; world_init(w, sc, cp, cfg);
    int threads = cfg->threads > 0 ? cfg->threads : static_cast<int>(std::thread::hardware_concurrency());
    if (threads < 1) threads = 1;
    const float exposure = static_cast<float>(std::pow(2, cfg->exposure_stops));       /* spheres-app.cpp:174 */
    const auto t0 = std::chrono::steady_clock::now();
    const int W = cfg->width, H = cfg->height;
    for (int s = 0; s < cfg->spp; ++s) {                                  /* one DrawBitmap() per iteration */
        const uint64_t n = cfg->first_sample + static_cast<uint64_t>(s);  /* m_sampleCount after :168 */
        float jx = 0.f, jy = 0.f;
        if (!(cfg->flags & ORC_FLAG_JITTER_OFF)) { jx = halton(n, 2); jy = halton(n, 3); }   /* :140 */
        parallel_rows(H, threads, [&](int j) {
            RayCount rc;
            for (int i = 0; i < W; ++i) {
                const uint32_t id = static_cast<uint32_t>(j) * static_cast<uint32_t>(W) + static_cast<uint32_t>(i);
                const Ray r = primary_ray(w, cfg, n, i, j, jx, jy);
                const V3 c = (w.integrator == ORC_INTEGRATOR_RECURSIVE) ? hit_color_recursive(w, r, 0, id, n, rc)
                                                                         : hit_color_iterative(w, r, id, n, rc);
                float* px = hdr + 4 * static_cast<size_t>(id);
                px[0] += c.x * exposure; px[1] += c.y * exposure; px[2] += c.z * exposure;   /* :182-183 */
                px[3] += 1.0f;
            }
            w.closest_rays.fetch_add(rc.closest, std::memory_order_relaxed);
            w.shadow_rays.fetch_add(rc.shadow, std::memory_order_relaxed);
        });
    }
    if (stats) {
        stats->paths = static_cast<uint64_t>(W) * H * cfg->spp;
        stats->closest_rays = w.closest_rays.load(); stats->shadow_rays = w.shadow_rays.load();
        stats->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    }
    return 0;
}

int orc_primary_hits(const orc_scene* sc, const orc_camera_params* cp, const orc_config* cfg, int32_t* sphere_idx, int32_t* material_id) {
    if (!sc || !cp || !cfg || cfg->width <= 0 || cfg->height <= 0 || cfg->first_sample < 1) return 1;
    g_math_ref = cfg->math_mode;
    World w; world_init(w, sc, cp, cfg);
    int threads = cfg->threads > 0 ? cfg->threads : static_cast<int>(std::thread::hardware_concurrency());
    const uint64_t n = cfg->first_sample;
    float jx = 0.f, jy = 0.f;
    if (!(cfg->flags & ORC_FLAG_JITTER_OFF)) { jx = halton(n, 2); jy = halton(n, 3); }
    parallel_rows(cfg->height, threads, [&](int j) {
        for (int i = 0; i < cfg->width; ++i) {
            const size_t id = static_cast<size_t>(j) * cfg->width + i;
            Hit h;
            const bool hit = closest_hit(w, primary_ray(w, cfg, n, i, j, jx, jy), h);
            if (sphere_idx) sphere_idx[id] = hit ? h.sphere : -1;
            if (material_id) material_id[id] = hit ? sc->materials[sc->spheres[h.sphere].material].kind : ORC_MAT_SKY;
        }
    });
    return 0;
}

/* Resolve, spheres-app.cpp:186-214: mean, ACES fit, gamma 1/2.2, XMStoreColor (BGRA8). */
void orc_resolve_ldr(const float* hdr, int64_t n_pixels, uint64_t sample_count, uint32_t* bgra8) {
    const float a = 2.51f, b = 0.03f, c = 2.43f, d = 0.59f, e = 0.14f;
    const float inv_gamma = 1 / 2.2f;
    const float n = static_cast<float>(sample_count);
    for (int64_t p = 0; p < n_pixels; ++p) {
        uint32_t ch[3];
        for (int k = 0; k < 3; ++k) {
            float x = hdr[4 * p + k] / n;                                 /* :201 */
            x = saturate((x * (a * x + b)) / (x * (c * x + d) + e));     /* :205 */
            x = std::pow(x, inv_gamma);                                   /* :208 */
            ch[k] = xmcolor_quantize(x);                                  /* :211 */
        }
        bgra8[p] = 0xFF000000u | (ch[0] << 16) | (ch[1] << 8) | ch[2];    /* alpha lane: pow(w, 0) == 1 */
    }
}

void orc_set_math_mode(int mode) { g_math_ref = mode; }
float orc_halton(uint64_t index, uint32_t base) { return halton(index, base); }
void orc_halton_disk(uint64_t index, uint32_t b1, uint32_t b2, float out[2]) { disk_from_u(halton(index, b1), halton(index, b2), &out[0], &out[1]); }
void orc_halton_hemisphere(uint64_t index, uint32_t b1, uint32_t b2, float out[3]) { st3(out, hemisphere_from_u(halton(index, b1), halton(index, b2))); }
void orc_hemisphere_from_u(float u1, float u2, float out[3]) { st3(out, hemisphere_from_u(u1, u2)); }
void orc_sincos(float theta, float* s, float* c) { sincos_pinned(theta, s, c); }
void orc_camera_make(const orc_camera_params* p, orc_camera* c) { camera_make(p, c); }
void orc_camera_get_ray(const orc_camera* c, float u, float v, float lx, float ly, float o[3], float d[3]) {
    const Ray r = camera_get_ray(*c, u, v, lx, ly); st3(o, r.o); st3(d, r.d);
}
int orc_sphere_intersect(const orc_sphere* s, const float o[3], const float d[3], float* t, float pos[3], float n[3], float uv[2]) {
    Hit h;
    if (!sphere_intersect(*s, 0, Ray{v3(o), v3(d)}, h)) return 0;
    if (t) *t = h.t;
    if (pos) st3(pos, h.pos);
    if (n) st3(n, h.n);
    if (uv) { uv[0] = h.u; uv[1] = h.v; }
    return 1;
}
int orc_closest_hit(const orc_scene* sc, const float o[3], const float d[3], int accel, float* t) {
    World w; w.sc = sc; w.accel = accel;
    if (accel == ORC_ACCEL_BVH) w.bvh.make(sc);
    Hit h;
    if (!closest_hit(w, Ray{v3(o), v3(d)}, h)) return -1;
    if (t) *t = h.t;
    return h.sphere;
}
void orc_texture_eval(const orc_texture* t, float u, float v, float rgba[4]) { texture_eval(*t, u, v, rgba); }
float orc_schlick(float f0, float n_dot_v) { return schlick(f0, n_dot_v); }
float orc_fresnel_term(float cos_i, float ior) { return fresnel_term(cos_i, ior); }
void orc_reflect(const float i[3], const float n[3], float out[3]) { st3(out, reflect3(v3(i), v3(n))); }
void orc_refract(const float i[3], const float n[3], float eta, float out[3]) { st3(out, refract3(v3(i), v3(n), eta)); }
uint8_t orc_xmcolor_quantize(float x) { return xmcolor_quantize(x); }
float orc_xmcolor_load(uint8_t k) { return xmcolor_load(k); }
void orc_rng_counter(uint32_t pixel, uint64_t sample, uint32_t bounce, uint32_t seed, float out[4]) { rng_counter(pixel, sample, bounce, seed, out); }

int orc_scatter(const orc_scene* sc, int sphere, const float o[3], const float d[3], float xi, float u1, float u2,
                float atten[3], float out_o[3], float out_d[3]) {
    World w; w.sc = sc; w.accel = ORC_ACCEL_BRUTE;
    const Ray r{v3(o), v3(d)};
    Hit h;
    if (!sphere_intersect(sc->spheres[sphere], sphere, r, h)) return -1;
    V3 a{0.f, 0.f, 0.f}; Ray out{};
    const bool s = scatter_with(w, r, h, Rand3{xi, u1, u2}, a, out);
    st3(atten, a); st3(out_o, out.o); st3(out_d, out.d);
    return s ? 1 : 0;
}
int orc_shade_hit(const orc_scene* sc, const float view_origin[3], const float o[3], const float d[3], int accel, float rgb[3]) {
    World w; w.sc = sc; w.accel = accel;
    world_lights(w);
    if (accel == ORC_ACCEL_BVH) w.bvh.make(sc);
    Hit h;
    if (!closest_hit(w, Ray{v3(o), v3(d)}, h)) return -1;
    RayCount rc;
    st3(rgb, shade(w, h, v3(view_origin), rc));
    return h.sphere;
}
const char* orc_build_info(void) {
    return "rt_oracle tier-B scalar restatement; "
#if defined(__FMA__)
           "fmaf=hw-fma; "
#else
           "fmaf=libm; "
#endif
#if defined(__FAST_MATH__)
           "FAST_MATH(!); "
#endif
           "compiler " __VERSION__;
}

}  // extern "C"
