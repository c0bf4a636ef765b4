// rt_b200.cu -- C-ABI implementation (include/rt_b200.h) of the B200-native `src/spheres` hot path.
// Host side: scene recorder (the reference's constructors, flattened), camera basis, Halton tables,
// launch logic.  Device side: rt_kernels.cuh.  No CPU fallback exists: every compute entry point
// launches a CUDA kernel or fails with RT_ERR_CUDA.
//
// Build (see cpuraytracer_b200/build.py):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -Xcompiler -fPIC,-ffp-contract=off
//        -shared -o librt_b200.so rt_b200.cu
#include "../../include/rt_b200.h"
#include "rt_kernels.cuh"
#include "rt_wavefront.cuh"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

using rt::u64;

// ------------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------------
static thread_local std::string g_last_error;
static int fail(int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    g_last_error = buf;
    return code;
}
#define RT_CUDA(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) \
    return fail(e_ == cudaErrorMemoryAllocation ? RT_ERR_OUT_OF_MEMORY : RT_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)

// ------------------------------------------------------------------------------------------------
// host restatement of the scene-building arithmetic (runs once per scene/camera, not per sample)
// ------------------------------------------------------------------------------------------------
namespace host {
struct V3 { float x, y, z; };
static inline V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline V3 add(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline V3 scale(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
static inline float dot3(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }          // XMVector3Dot, SSE2 order
static inline V3 cross3(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
static inline V3 normalize3(V3 v) {                                                           // XMVector3Normalize
    const float l2 = dot3(v, v), l = std::sqrt(l2);
    if (l == 0.f) return {0.f, 0.f, 0.f};
    if (std::isinf(l2)) { const float q = std::nanf(""); return {q, q, q}; }
    return {v.x / l, v.y / l, v.z / l};
}
// XMCOLOR(float,float,float,float): saturate, *255, round-to-nearest-even (XMStoreColor); XMLoadColor: k * (1/255)
static inline float xmcolor_roundtrip(float x) {
    float v = std::fmin(std::fmax(x, 0.0f), 1.0f);
    v = v * 255.0f;
    const int k = static_cast<int>(std::nearbyint(v));
    return static_cast<float>(k) * (1.0f / 255.0f);
}
// Random::HaltonSample, quasi-random.cpp:3-16
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
// Camera::Camera, camera.cpp:3-28
static void camera_basis(const rt_camera_params& p, rt_camera_basis& c) {
    const V3 origin{p.origin[0], p.origin[1], p.origin[2]}, look{p.look_at[0], p.look_at[1], p.look_at[2]};
    const float theta = p.vfov_deg * 3.141592654f / 180.f;
    const float half_h = std::tan(theta / 2.f);
    const float half_w = p.aspect * half_h;
    const V3 up{0.f, 1.f, 0.f};
    const V3 w = normalize3(sub(look, origin));
    const V3 u = normalize3(cross3(up, w));
    const V3 v = cross3(w, u);
    const V3 io = add(origin, scale(1.f, w)), mx = scale(half_w, u), my = scale(half_h, v);
    c.origin[0] = origin.x; c.origin[1] = origin.y; c.origin[2] = origin.z;
    c.image_origin[0] = io.x; c.image_origin[1] = io.y; c.image_origin[2] = io.z;
    c.x[0] = mx.x; c.x[1] = mx.y; c.x[2] = mx.z;
    c.y[0] = my.x; c.y[1] = my.y; c.y[2] = my.z;
    c.aperture = p.aperture; c.focus = p.focus;
}
}  // namespace host

// ------------------------------------------------------------------------------------------------
// scene recorder
// ------------------------------------------------------------------------------------------------
struct rt_scene {
    std::vector<rt_sphere_desc> spheres;
    std::vector<rt_material_desc> materials;
    std::vector<rt_texture_desc> textures;
    std::vector<rt_light_desc> lights;
    float sky_luminance = 0.f;
    int32_t sky_texture = -1;
};

static bool finite3(const float* v) { return std::isfinite(v[0]) && std::isfinite(v[1]) && std::isfinite(v[2]); }

extern "C" {

const char* rt_last_error(void) { return g_last_error.c_str(); }
const char* rt_build_info(void) {
    return "cpuraytracer_b200 librt_b200: sm_100a megakernel (FFMA2 conservative cull + exact reference-order tests), "
           "nvcc " __DATE__ ", -fmad=false";
}

int rt_scene_create(rt_scene** out) {
    if (!out) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_create: out is NULL");
    *out = new rt_scene();
    return RT_OK;
}
void rt_scene_destroy(rt_scene* s) { delete s; }

int rt_scene_add_texture_const(rt_scene* s, const float rgba[4]) {
    if (!s || !rgba) return -fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_add_texture_const: NULL argument");
    rt_texture_desc t{}; t.kind = RT_TEX_CONST; t.tiling = 0.f;
    for (int k = 0; k < 4; ++k) { t.c0[k] = host::xmcolor_roundtrip(rgba[k]); t.c1[k] = t.c0[k]; }
    s->textures.push_back(t);
    return static_cast<int>(s->textures.size()) - 1;
}
int rt_scene_add_texture_checker(rt_scene* s, const float rgba0[4], const float rgba1[4], float tiling) {
    if (!s || !rgba0 || !rgba1) return -fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_add_texture_checker: NULL argument");
    if (!std::isfinite(tiling)) return -fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_add_texture_checker: tiling must be finite");
    rt_texture_desc t{}; t.kind = RT_TEX_CHECKER; t.tiling = tiling;
    for (int k = 0; k < 4; ++k) { t.c0[k] = host::xmcolor_roundtrip(rgba0[k]); t.c1[k] = host::xmcolor_roundtrip(rgba1[k]); }
    s->textures.push_back(t);
    return static_cast<int>(s->textures.size()) - 1;
}
static int add_material(rt_scene* s, int kind, int texture, float smoothness, float ior, const char* who) {
    if (!s) return -fail(RT_ERR_INVALID_ARGUMENT, "%s: scene is NULL", who);
    if (!std::isfinite(smoothness)) return -fail(RT_ERR_INVALID_ARGUMENT, "%s: smoothness must be finite", who);
    if (kind != RT_MAT_TRANSPARENT && (texture < 0 || texture >= static_cast<int>(s->textures.size())))
        return -fail(RT_ERR_INVALID_ARGUMENT, "%s: texture id %d out of range [0,%zu)", who, texture, s->textures.size());
    rt_material_desc m{}; m.kind = kind; m.texture = kind == RT_MAT_TRANSPARENT ? -1 : texture; m.smoothness = smoothness; m.ior = ior;
    s->materials.push_back(m);
    return static_cast<int>(s->materials.size()) - 1;
}
int rt_scene_add_material_opaque(rt_scene* s, int texture, float smoothness) { return add_material(s, RT_MAT_OPAQUE, texture, smoothness, 0.f, "rt_scene_add_material_opaque"); }
int rt_scene_add_material_metal(rt_scene* s, int texture, float smoothness) { return add_material(s, RT_MAT_METAL, texture, smoothness, 0.f, "rt_scene_add_material_metal"); }
int rt_scene_add_material_transparent(rt_scene* s, float smoothness, float ior) {
    if (!(ior > 0.f) || !std::isfinite(ior)) return -fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_add_material_transparent: ior must be finite and > 0");
    return add_material(s, RT_MAT_TRANSPARENT, -1, smoothness, ior, "rt_scene_add_material_transparent");
}
int rt_scene_add_sphere(rt_scene* s, const float center[3], float radius, int material) {
    if (!s || !center) return -fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_add_sphere: NULL argument");
    if (!finite3(center) || !(radius > 0.f) || !std::isfinite(radius)) return -fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_add_sphere: centre must be finite and radius > 0");
    if (material < 0 || material >= static_cast<int>(s->materials.size()))
        return -fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_add_sphere: material id %d out of range [0,%zu)", material, s->materials.size());
    s->spheres.push_back(rt_sphere_desc{center[0], center[1], center[2], radius, material});
    return static_cast<int>(s->spheres.size()) - 1;
}
int rt_scene_set_sky(rt_scene* s, float luminance, int texture) {
    if (!s) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_set_sky: scene is NULL");
    if (texture < 0 || texture >= static_cast<int>(s->textures.size())) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_set_sky: texture id %d out of range", texture);
    s->sky_luminance = luminance; s->sky_texture = texture;
    return RT_OK;
}
int rt_scene_add_directional_light(rt_scene* s, const float dir[3], const float rgba[4], float luminance) {
    if (!s || !dir || !rgba) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_add_directional_light: NULL argument");
    const host::V3 n = host::normalize3({dir[0], dir[1], dir[2]});           // light.cpp:7
    if (!(host::dot3(n, n) > 0.f)) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_add_directional_light: zero direction");
    rt_light_desc l{};
    l.dir[0] = dir[0]; l.dir[1] = dir[1]; l.dir[2] = dir[2];          // the ctor argument; normalised when bound (rt_set_scene)
    for (int k = 0; k < 4; ++k) l.color[k] = host::xmcolor_roundtrip(rgba[k]);   // light.cpp:8
    l.luminance = luminance;
    s->lights.push_back(l);
    return RT_OK;
}
int rt_scene_counts(const rt_scene* s, int32_t* ns, int32_t* nm, int32_t* nt, int32_t* nl) {
    if (!s) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_counts: scene is NULL");
    if (ns) *ns = static_cast<int32_t>(s->spheres.size());
    if (nm) *nm = static_cast<int32_t>(s->materials.size());
    if (nt) *nt = static_cast<int32_t>(s->textures.size());
    if (nl) *nl = static_cast<int32_t>(s->lights.size());
    return RT_OK;
}
int rt_scene_export(const rt_scene* s, rt_sphere_desc* sp, rt_material_desc* ma, rt_texture_desc* te, rt_light_desc* li, float* sky_lum, int32_t* sky_tex) {
    if (!s) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_export: scene is NULL");
    if (sp) std::copy(s->spheres.begin(), s->spheres.end(), sp);
    if (ma) std::copy(s->materials.begin(), s->materials.end(), ma);
    if (te) std::copy(s->textures.begin(), s->textures.end(), te);
    if (li) std::copy(s->lights.begin(), s->lights.end(), li);
    if (sky_lum) *sky_lum = s->sky_luminance;
    if (sky_tex) *sky_tex = s->sky_texture;
    return RT_OK;
}
int rt_scene_import(rt_scene* s, int32_t ns, const rt_sphere_desc* sp, int32_t nm, const rt_material_desc* ma, int32_t nt, const rt_texture_desc* te,
                    int32_t nl, const rt_light_desc* li, float sky_lum, int32_t sky_tex) {
    if (!s || ns < 0 || nm < 0 || nt < 0 || nl < 0 || (ns && !sp) || (nm && !ma) || (nt && !te) || (nl && !li))
        return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_import: bad argument");
    // the same checks the rt_scene_add_* recorders make, so that a scene file cannot carry what the recorder would refuse
    for (int i = 0; i < nt; ++i) {
        if (te[i].kind != RT_TEX_CONST && te[i].kind != RT_TEX_CHECKER) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_import: texture %d has kind %d", i, te[i].kind);
        bool ok = std::isfinite(te[i].tiling);
        for (int k = 0; k < 4; ++k) ok = ok && std::isfinite(te[i].c0[k]) && std::isfinite(te[i].c1[k]);
        if (!ok) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_import: texture %d has a non-finite colour or tiling", i);
    }
    for (int i = 0; i < nm; ++i) {
        if (ma[i].kind < RT_MAT_OPAQUE || ma[i].kind > RT_MAT_TRANSPARENT) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_import: material %d has kind %d", i, ma[i].kind);
        if (ma[i].kind != RT_MAT_TRANSPARENT && (ma[i].texture < 0 || ma[i].texture >= nt)) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_import: material %d texture out of range", i);
        if (!std::isfinite(ma[i].smoothness)) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_import: material %d has a non-finite smoothness", i);
        if (ma[i].kind == RT_MAT_TRANSPARENT && !(ma[i].ior > 0.f && std::isfinite(ma[i].ior))) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_import: material %d: ior must be finite and > 0", i);
    }
    for (int i = 0; i < ns; ++i) {
        const float ctr[3] = {sp[i].cx, sp[i].cy, sp[i].cz};
        if (sp[i].material < 0 || sp[i].material >= nm || !(sp[i].radius > 0.f) || !std::isfinite(sp[i].radius) || !finite3(ctr))
            return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_import: sphere %d invalid (material id, finite centre, finite radius > 0)", i);
    }
    for (int i = 0; i < nl; ++i) {
        const host::V3 nrm = host::normalize3({li[i].dir[0], li[i].dir[1], li[i].dir[2]});
        if (!finite3(li[i].dir) || !(host::dot3(nrm, nrm) > 0.f) || !std::isfinite(li[i].luminance) || !finite3(li[i].color))
            return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_import: light %d invalid (finite non-zero direction, finite colour and luminance)", i);
    }
    if (sky_tex < 0 || sky_tex >= nt || !std::isfinite(sky_lum)) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_import: sky texture out of range or non-finite luminance");
    s->spheres.assign(sp, sp + ns); s->materials.assign(ma, ma + nm); s->textures.assign(te, te + nt); s->lights.assign(li, li + nl);
    s->sky_luminance = sky_lum; s->sky_texture = sky_tex;
    return RT_OK;
}
int rt_camera_basis_of(const rt_camera_params* cam, rt_camera_basis* out) {
    if (!cam || !out) return fail(RT_ERR_INVALID_ARGUMENT, "rt_camera_basis_of: NULL argument");
    host::camera_basis(*cam, *out);
    return RT_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// context
// ------------------------------------------------------------------------------------------------
struct EventPair { cudaEvent_t a, b; int kind; };

struct rt_context {
    int device = 0;
    cudaStream_t stream = nullptr; bool own_stream = false;
    cudaDeviceProp prop{};
    int clock_khz = 0;
    // scene
    float4 *cullA = nullptr, *cullB = nullptr; rt::DevMat* mats = nullptr;
    float4* geom = nullptr;
    // wavefront layout (RT_FLAG_WAVEFRONT): SoA path state, allocated on first use
    rt::WfState wf{}; void* wf_block = nullptr; unsigned int* wf_live = nullptr; int wf_bps = 0;
    int n_spheres = 0, n_pad = 0;
    bool scene_set = false, camera_set = false;
    rt::RenderParams base{};
    // framebuffer
    int W = 0, H = 0;
    float4* hdr_own = nullptr; float4* hdr_ext = nullptr;
    uint32_t* ldr = nullptr;
    int32_t *ids = nullptr, *mids = nullptr;
    float4* partial = nullptr; size_t partial_bytes = 0;
    uint16_t* hits = nullptr; size_t hits_bytes = 0;   // first-hit records of one batch (two-stage schedule)
    int path_tpb = 0, path_bps = 0;                    // CTA size / resident CTAs per SM of the stage-2 kernel
    int first_tpb = 0, first_bps = 0;                  // CTA size / resident CTAs per SM of the brute-force stage-1 kernel (its own register budget)
    bool global_scene = false;                         // the sphere list does not fit one CTA's shared memory: kernels read it from global memory
    int accel_gshift_min = 3;                          // finest group size rt_set_scene tries (rt_set_accel_group_slots)
    float2 *jitter = nullptr, *lens = nullptr; size_t jitter_cap = 0, lens_cap = 0;
    u64* counters = nullptr;   // [0] work counter, [1..4] stats
    // settings
    float exposure_stops = -15.f; uint32_t seed = 0; int chunk_spp = 0;
    // instrumentation
    std::vector<EventPair> pending; std::vector<EventPair> pool;
    double render_ms = 0, resolve_ms = 0; int launches = 0;
    u64 stat_base[6] = {0, 0, 0, 0, 0, 0};
    int render_blocks_per_sm = 0, tpb = 256;
    // RT_FLAG_ACCEL: shared-memory image of the box hierarchy (rt_bvh.cuh), built by rt_set_scene
    uint4* accel_blob = nullptr; bool accel_ok = false; std::string accel_why;
    struct ExtraLight { float L[3], lc[3], iL[3], rot[9]; float4 *cullA, *cullB; };
    std::vector<ExtraLight> extra_lights; float4* extra_pairs = nullptr;   // lights 1.. (one additional launch each, see rt_render)
    u64* sampler = nullptr; size_t n_sampler = 0;   // RT_FLAG_HALTON: per-material-instance Halton indices (material.h:34,50-51,67); second half = snapshot
    int accel_tpb = 0, accel_bps = 0;

    float4* hdr() const { return hdr_ext ? hdr_ext : hdr_own; }
};

static int ctx_use(rt_context* c) {
    RT_CUDA(cudaSetDevice(c->device));
    return RT_OK;
}
static void ev_account(rt_context* c, const EventPair& e) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, e.a, e.b) == cudaSuccess) { if (e.kind == 0) c->render_ms += ms; else c->resolve_ms += ms; }
}
static EventPair ev_begin(rt_context* c, int kind) {
    // pairs whose end event has completed go back to the pool (their time is accounted now), so a long interactive loop
    // that never asks for statistics keeps a bounded number of live CUDA events
    size_t k = 0;
    while (k < c->pending.size() && cudaEventQuery(c->pending[k].b) == cudaSuccess) ++k;     // one stream: completion is in order
    if (k == 0 && c->pending.size() >= 256) { cudaEventSynchronize(c->pending[0].b); k = 1; }
    for (size_t i = 0; i < k; ++i) { ev_account(c, c->pending[i]); c->pool.push_back(c->pending[i]); }
    c->pending.erase(c->pending.begin(), c->pending.begin() + static_cast<std::ptrdiff_t>(k));
    EventPair e;
    if (!c->pool.empty()) { e = c->pool.back(); c->pool.pop_back(); }
    else { cudaEventCreate(&e.a); cudaEventCreate(&e.b); }
    e.kind = kind;
    cudaEventRecord(e.a, c->stream);
    return e;
}
static void ev_end(rt_context* c, EventPair e) { cudaEventRecord(e.b, c->stream); c->pending.push_back(e); }
static void ev_collect(rt_context* c) {
    for (auto& e : c->pending) { ev_account(c, e); c->pool.push_back(e); }
    c->pending.clear();
}
template <class T> static int dev_realloc(T*& p, size_t n) {
    if (p) { cudaFree(p); p = nullptr; }
    if (n == 0) return RT_OK;
    RT_CUDA(cudaMalloc(&p, n * sizeof(T)));
    return RT_OK;
}
static size_t scene_smem_bytes(int n_pad) { return static_cast<size_t>(n_pad / 2) * 32u + static_cast<size_t>(n_pad) * 16u; }   // pair arrays + world-space (c, r)
// scene arrays + the per-thread candidate lists (2 lists x RT_LIST_CAP words), see stage_scene()
static size_t cta_smem_bytes(int n_pad, int tpb, bool global_scene = false) {
    size_t b = (global_scene ? 0u : scene_smem_bytes(n_pad)) + static_cast<size_t>(tpb) * 2u * RT_LIST_CAP * 4u;
#if RT_FLAT_EXACT
    b += static_cast<size_t>((tpb + 31) / 32) * RT_WQ_BYTES_PER_WARP;       // per-warp candidate queue of the exact-test phase
#endif
    return b;
}

// ------------------------------------------------------------------------------------------------
// RT_FLAG_ACCEL: host build of the box hierarchy image (layout + conservativeness argument: rt_bvh.cuh).
// The reference builds its BVH in BvhNode::BvhNode (ray-tracing.cpp:107-167: median split on a random axis);
// here the order is a Morton sort, the tree is implicit (32 spheres per group, 8 groups per L2 node) and the
// image is position-independent so every CTA copies it verbatim into shared memory.
// ------------------------------------------------------------------------------------------------
static size_t accel_cta_smem_bytes(const rt::AccelParams& a, int tpb) {
    return static_cast<size_t>(a.blob_bytes) + static_cast<size_t>(tpb) * (a.gmask_words + 2 * RT_ACCEL_ENTRY_CAP) * 4u;
}
static uint32_t morton_spread10(uint32_t v) {
    v = (v | (v << 16)) & 0x030000FFu; v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u; v = (v | (v << 2)) & 0x09249249u;
    return v;
}
static float round_up(double x) { float f = static_cast<float>(x); if (static_cast<double>(f) < x) f = std::nextafter(f, INFINITY); return f; }

// The accel kernels are instantiated for 8, 16 and 32 slots per small group (template argument GSHIFT).
template <int G> static const void* accel_render_fn() { return reinterpret_cast<const void*>(&rt::rt_render_kernel<rt::MODE_ACCEL, false, G>); }
template <int G> static const void* accel_primary_fn() { return reinterpret_cast<const void*>(&rt::rt_primary_kernel<rt::MODE_ACCEL, G>); }
static const void* accel_render_kernel(int gshift) { return gshift == 3 ? accel_render_fn<3>() : gshift == 4 ? accel_render_fn<4>() : accel_render_fn<5>(); }
static const void* accel_primary_kernel(int gshift) { return gshift == 3 ? accel_primary_fn<3>() : gshift == 4 ? accel_primary_fn<4>() : accel_primary_fn<5>(); }

// Two-stage schedule: kernel of a stage for (query mode, sampler, group size, direct-only).
#define RT_FN(...) reinterpret_cast<const void*>(&__VA_ARGS__)
template <int G> static const void* accel_path_fn(bool halton, bool direct) {
    if (halton) return RT_FN(rt::rt_path_kernel<rt::MODE_ACCEL, true, G, false>);
    return direct ? RT_FN(rt::rt_path_kernel<rt::MODE_ACCEL, false, G, true>) : RT_FN(rt::rt_path_kernel<rt::MODE_ACCEL, false, G, false>);
}
static const void* path_kernel(int mode, bool halton, int gshift, bool direct) {
    if (mode == rt::MODE_ACCEL) return gshift == 3 ? accel_path_fn<3>(halton, direct) : gshift == 4 ? accel_path_fn<4>(halton, direct) : accel_path_fn<5>(halton, direct);
    if (mode == rt::MODE_NO_CULL) return RT_FN(rt::rt_path_kernel<rt::MODE_NO_CULL, false, 5, false>);
    if (mode == rt::MODE_CULL_GLOBAL) return direct ? RT_FN(rt::rt_path_kernel<rt::MODE_CULL_GLOBAL, false, 5, true>) : RT_FN(rt::rt_path_kernel<rt::MODE_CULL_GLOBAL, false, 5, false>);
    if (halton) return RT_FN(rt::rt_path_kernel<rt::MODE_CULL, true, 5, false>);
    return direct ? RT_FN(rt::rt_path_kernel<rt::MODE_CULL, false, 5, true>) : RT_FN(rt::rt_path_kernel<rt::MODE_CULL, false, 5, false>);
}
static const void* first_hit_kernel(int mode, int gshift) {
    if (mode == rt::MODE_ACCEL) return gshift == 3 ? RT_FN(rt::rt_first_hit_kernel<rt::MODE_ACCEL, 3>) : gshift == 4 ? RT_FN(rt::rt_first_hit_kernel<rt::MODE_ACCEL, 4>) : RT_FN(rt::rt_first_hit_kernel<rt::MODE_ACCEL, 5>);
    if (mode == rt::MODE_CULL_GLOBAL) return RT_FN(rt::rt_first_hit_kernel<rt::MODE_CULL_GLOBAL, 5>);
    return mode == rt::MODE_NO_CULL ? RT_FN(rt::rt_first_hit_kernel<rt::MODE_NO_CULL, 5>) : RT_FN(rt::rt_first_hit_kernel<rt::MODE_CULL, 5>);
}
// Opt every kernel of a mode in to the large dynamic shared memory carve-out.
static int optin_smem(int mode, int gshift, int optin) {
    for (int h = 0; h < 2; ++h) for (int dd = 0; dd < 2; ++dd)
        RT_CUDA(cudaFuncSetAttribute(path_kernel(mode, h != 0, gshift, dd != 0), cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    RT_CUDA(cudaFuncSetAttribute(first_hit_kernel(mode, gshift), cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    return RT_OK;
}

// Host half of build_accel: the position-independent image and its section table (no CUDA call, no context), also reachable
// through rt_scene_accel_image() so that the CPU tests can check its invariants.  false + `why` = hierarchy unavailable.
static bool make_accel_image(const rt_scene* s, rt::AccelParams& a, std::vector<unsigned char>& img, std::string& why, int gshift_min = 3) {
    a = rt::AccelParams{};
    const int n = static_cast<int>(s->spheres.size());
    if (n > 65000) { why = "more than 65000 spheres (slot -> index map and first-hit records are 16 bit)"; return false; }
    // big spheres: radius > 4 x median (at most 64, the largest first)
    std::vector<float> radii(n);
    for (int i = 0; i < n; ++i) radii[i] = s->spheres[i].radius;
    std::vector<int> big, small;
    if (n > 0) {
        std::vector<float> sorted = radii; std::sort(sorted.begin(), sorted.end());
        float thr = 4.f * sorted[n / 2];
        const int over = static_cast<int>(sorted.end() - std::upper_bound(sorted.begin(), sorted.end(), thr));
        if (over > 64) thr = sorted[n - 65];
        for (int i = 0; i < n; ++i) (radii[i] > thr ? big : small).push_back(i);
    }
    // Morton order of the small spheres' centres (10 bits per axis inside their bounding box)
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    for (int i : small) { const rt_sphere_desc& sp = s->spheres[i]; const double cc[3] = {sp.cx, sp.cy, sp.cz};
        for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], cc[k]); hi[k] = std::max(hi[k], cc[k]); } }
    std::vector<std::pair<uint32_t, int>> keyed;
    for (int i : small) {
        const rt_sphere_desc& sp = s->spheres[i]; const double cc[3] = {sp.cx, sp.cy, sp.cz};
        uint32_t q[3];
        for (int k = 0; k < 3; ++k) { const double e = hi[k] - lo[k]; q[k] = e > 0 ? static_cast<uint32_t>(std::min(1023.0, (cc[k] - lo[k]) / e * 1024.0)) : 0u; }
        keyed.push_back({morton_spread10(q[0]) | (morton_spread10(q[1]) << 1) | (morton_spread10(q[2]) << 2), i});
    }
    std::stable_sort(keyed.begin(), keyed.end(), [](const std::pair<uint32_t, int>& x, const std::pair<uint32_t, int>& y) { return x.first < y.first; });
    const int n_big_groups = (static_cast<int>(big.size()) + 31) / 32;
    // Slots per small group: the finest of 8 / 16 / 32 whose group count still fits the 32 x 8 fan-out of the two box levels
    // (a ray then culls fewer spheres per touched box; measured in profiles/r01x_accel_group_size.txt).  gshift_min
    // (rt_set_accel_group_slots: 8, 16 or 32 slots) sets the finest size tried.
    int gshift = std::min(5, std::max(3, gshift_min));
    while (gshift < 5 && (static_cast<int>(small.size()) + (1 << gshift) - 1) / (1 << gshift) > RT_ACCEL_GSHIFT_MAX_GROUPS) ++gshift;
    const int GS = 1 << gshift;
    const int n_groups = (static_cast<int>(small.size()) + GS - 1) / GS;
    const int n2 = (n_groups + 7) / 8;
    if (n2 > 32) { why = "too many small spheres for this group size (the L2 mask is one 32-bit word: at most 8192 with 32-slot groups)"; return false; }
    const int small_base = n_big_groups * 32;
    const int n_slots = small_base + n_groups * GS;
    std::vector<int> slot_orig(std::max(n_slots, 1), -1);
    for (size_t k = 0; k < big.size(); ++k) slot_orig[k] = big[k];
    for (size_t k = 0; k < keyed.size(); ++k) slot_orig[small_base + k] = keyed[k].second;
    const int np = n_slots / 2;
    // sections (A and B start on 256-byte boundaries: the XOR-swizzled group loads rely on aligned groups)
    auto align16 = [](size_t x) { return (x + 15u) & ~static_cast<size_t>(15u); };
    auto align256 = [](size_t x) { return (x + 255u) & ~static_cast<size_t>(255u); };
    const size_t off_A = 0, off_B = align256(off_A + static_cast<size_t>(np) * 16), off_R = off_B + static_cast<size_t>(np) * 16;
    const size_t off_I = align16(off_R + static_cast<size_t>(n_slots) * 4);
    const size_t off_box = align16(off_I + static_cast<size_t>(n_slots) * 2);
    const size_t off_top = off_box + static_cast<size_t>(n2) * RT_ACCEL_NODE_FLOATS * 4;
    const size_t total = std::max<size_t>(16, off_top + static_cast<size_t>(n2) * 32);
    img.assign(total, 0);
    float* A = reinterpret_cast<float*>(img.data() + off_A); float* B = reinterpret_cast<float*>(img.data() + off_B);
    float* R = reinterpret_cast<float*>(img.data() + off_R); uint16_t* I = reinterpret_cast<uint16_t*>(img.data() + off_I);
    float* BOX = reinterpret_cast<float*>(img.data() + off_box); float* TOP = reinterpret_cast<float*>(img.data() + off_top);
    const double K = static_cast<double>(RT_CULL_K);
    for (int slot = 0; slot < n_slots; ++slot) {
        const int p = slot >> 1, h = slot & 1, i = slot_orig[slot];
        float cx = 0.f, cy = 0.f, cz = 0.f, wn = RT_PAD_W, r = 0.f;
        if (i >= 0) {
            const rt_sphere_desc& sp = s->spheres[i];
            cx = sp.cx; cy = sp.cy; cz = sp.cz; r = sp.radius;
            const double cc = static_cast<double>(sp.cx) * sp.cx + static_cast<double>(sp.cy) * sp.cy + static_cast<double>(sp.cz) * sp.cz;
            const double r2 = static_cast<double>(sp.radius) * sp.radius;
            wn = static_cast<float>(r2 - cc + K * (cc + r2));
        }
        A[p * 4 + h] = cx; A[p * 4 + 2 + h] = cy; B[p * 4 + h] = cz; B[p * 4 + 2 + h] = wn;
        R[slot] = r; I[slot] = i >= 0 ? static_cast<uint16_t>(i) : 0xFFFFu;
    }
    // boxes
    struct Box { double lo[3], hi[3]; bool any; };
    auto empty_box = []() { Box b; for (int k = 0; k < 3; ++k) { b.lo[k] = 1e300; b.hi[k] = -1e300; } b.any = false; return b; };
    auto grow = [](Box& b, const Box& o) { if (!o.any) return; for (int k = 0; k < 3; ++k) { b.lo[k] = std::min(b.lo[k], o.lo[k]); b.hi[k] = std::max(b.hi[k], o.hi[k]); } b.any = true; };
    auto to_ch = [](const Box& b, float* cxyz, float* hxyz) {
        if (!b.any) { for (int k = 0; k < 3; ++k) { cxyz[k] = 0.f; hxyz[k] = -1.0e30f; } return; }
        for (int k = 0; k < 3; ++k) {
            const float cf = static_cast<float>(0.5 * (b.lo[k] + b.hi[k]));
            cxyz[k] = cf; hxyz[k] = round_up(std::max(b.hi[k] - static_cast<double>(cf), static_cast<double>(cf) - b.lo[k]));
        }
    };
    std::vector<Box> gbox(n2 * 8, empty_box());
    Box all = empty_box();
    double r_min = 1e300, r_max = 0;
    for (int g = 0; g < n_groups; ++g)
        for (int j = 0; j < GS; ++j) {
            const int i = slot_orig[small_base + g * GS + j];
            if (i < 0) continue;
            const rt_sphere_desc& sp = s->spheres[i];
            Box b; b.any = true; const double cc[3] = {sp.cx, sp.cy, sp.cz};
            for (int k = 0; k < 3; ++k) { b.lo[k] = cc[k] - sp.radius; b.hi[k] = cc[k] + sp.radius; }
            grow(gbox[g], b);
            r_min = std::min<double>(r_min, sp.radius); r_max = std::max<double>(r_max, sp.radius);
        }
    for (int j2 = 0; j2 < n2; ++j2) {
        Box node = empty_box();
        float* blk = BOX + static_cast<size_t>(j2) * RT_ACCEL_NODE_FLOATS;
        for (int k = 0; k < 8; ++k) {
            const Box& b = gbox[j2 * 8 + k];
            grow(node, b);
            float cc[3], hh[3]; to_ch(b, cc, hh);
            blk[k] = cc[0]; blk[8 + k] = cc[1]; blk[16 + k] = cc[2]; blk[24 + k] = hh[0]; blk[32 + k] = hh[1]; blk[40 + k] = hh[2];
        }
        float cc[3], hh[3]; to_ch(node, cc, hh);
        float* t = TOP + static_cast<size_t>(j2) * 8;
        t[0] = cc[0]; t[1] = cc[1]; t[2] = cc[2]; t[3] = hh[0]; t[4] = hh[1]; t[5] = hh[2];
        grow(all, node);
    }
    a.csx = a.csy = a.csz = 0.f; a.cs_radius = 0.f; a.r_min = 1.f; a.r_max = 1.f; a.c_max = 0.f;
    if (all.any) {
        const double cs[3] = {0.5 * (all.lo[0] + all.hi[0]), 0.5 * (all.lo[1] + all.hi[1]), 0.5 * (all.lo[2] + all.hi[2])};
        a.csx = static_cast<float>(cs[0]); a.csy = static_cast<float>(cs[1]); a.csz = static_cast<float>(cs[2]);
        double rad = 0, cmax = 0;
        for (int i : small) {
            const rt_sphere_desc& sp = s->spheres[i];
            const double dx = sp.cx - static_cast<double>(a.csx), dy = sp.cy - static_cast<double>(a.csy), dz = sp.cz - static_cast<double>(a.csz);
            rad = std::max(rad, std::sqrt(dx * dx + dy * dy + dz * dz) + sp.radius);
        }
        for (int k = 0; k < 3; ++k) cmax = std::max(cmax, std::max(std::fabs(all.lo[k]), std::fabs(all.hi[k])));
        a.cs_radius = round_up(rad * (1.0 + 1e-6)); a.c_max = round_up(cmax);
        a.r_min = static_cast<float>(r_min); a.r_max = static_cast<float>(r_max);
    }
    a.blob = nullptr; a.blob_bytes = static_cast<int>(total);
    a.off_B = static_cast<int>(off_B); a.off_R = static_cast<int>(off_R); a.off_I = static_cast<int>(off_I);
    a.off_box = static_cast<int>(off_box); a.off_top = static_cast<int>(off_top);
    a.n_pairs = np; a.n_big_groups = n_big_groups; a.n_big_pairs = (static_cast<int>(big.size()) + 1) / 2; a.n_groups = n_groups; a.gshift = gshift; a.n2 = n2; a.gmask_words = std::max(1, (n2 + 3) / 4);
    return true;
}

static int build_accel(rt_context* c, const rt_scene* s) {
    c->accel_ok = false; c->accel_why.clear();
    rt::AccelParams& a = c->base.accel;
    std::vector<unsigned char> img;
    if (!make_accel_image(s, a, img, c->accel_why, c->accel_gshift_min)) return RT_OK;
    if (a.n_pairs * 2 > 65534) { c->accel_why = "more than 65534 slots in the hierarchy image (first-hit records are 16 bit)"; return RT_OK; }
    const size_t total = img.size();
    const int gshift = a.gshift;
    if (int r = dev_realloc(c->accel_blob, total / 16)) return r;
    RT_CUDA(cudaMemcpyAsync(c->accel_blob, img.data(), total, cudaMemcpyHostToDevice, c->stream));
    RT_CUDA(cudaStreamSynchronize(c->stream));
    a.blob = c->accel_blob;
    // CTA size: most resident warps per SM next to one image copy per CTA
    const int optin = static_cast<int>(c->prop.sharedMemPerBlockOptin);
    RT_CUDA(cudaFuncSetAttribute(accel_render_kernel(gshift), cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    RT_CUDA(cudaFuncSetAttribute(accel_primary_kernel(gshift), cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    if (int r = optin_smem(rt::MODE_ACCEL, gshift, optin)) return r;
    int best_tpb = 0, best_bps = 0;
    for (int tpb = 128; tpb <= RT_ACCEL_MAX_TPB; tpb += 128) {
        const size_t smem = accel_cta_smem_bytes(a, tpb);
        if (smem > static_cast<size_t>(optin)) break;
        int bps = 0, bps1 = 0;                               // the CTA shape must fit the one-kernel and the two-stage kernels alike
        RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, path_kernel(rt::MODE_ACCEL, false, gshift, false), tpb, smem));
        RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps1, accel_render_kernel(gshift), tpb, smem));
        bps = std::min(bps, bps1);
        if (bps * tpb > best_bps * best_tpb) { best_tpb = tpb; best_bps = bps; }
    }
    if (best_bps < 1) { c->accel_why = "the hierarchy image does not fit in shared memory"; return RT_OK; }
    c->accel_tpb = best_tpb; c->accel_bps = best_bps; c->accel_ok = true;
    return RT_OK;
}

// ------------------------------------------------------------------------------------------------
// Pair arrays of the brute-force cull in the light-aligned frame (rt_device.cuh, SceneView): rows of `rot` = (e1, e2, L), c' = rot c,
// everything in double, one rounding to float per coordinate.  w = r^2 - |c|^2 + k (|c|^2 + r^2) is rotation invariant.  Host only
// (no CUDA call): rt_set_scene uploads the result, rt_scene_cull_image hands it to the CPU tests.
// ------------------------------------------------------------------------------------------------
static void light_frame(const host::V3& Lu, bool lit, double rot[9]) {
    if (!lit) { const double id[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}; std::copy(id, id + 9, rot); return; }
    double L[3] = {Lu.x, Lu.y, Lu.z};
    const double ln = std::sqrt(L[0] * L[0] + L[1] * L[1] + L[2] * L[2]);
    for (double& v : L) v /= ln;
    const int ax = std::fabs(L[0]) <= std::fabs(L[1]) ? (std::fabs(L[0]) <= std::fabs(L[2]) ? 0 : 2) : (std::fabs(L[1]) <= std::fabs(L[2]) ? 1 : 2);
    double u[3] = {0, 0, 0}; u[ax] = 1.0;                               // the axis least aligned with L
    double e1[3] = {u[1] * L[2] - u[2] * L[1], u[2] * L[0] - u[0] * L[2], u[0] * L[1] - u[1] * L[0]};
    const double n1 = std::sqrt(e1[0] * e1[0] + e1[1] * e1[1] + e1[2] * e1[2]);
    for (double& v : e1) v /= n1;
    const double e2[3] = {L[1] * e1[2] - L[2] * e1[1], L[2] * e1[0] - L[0] * e1[2], L[0] * e1[1] - L[1] * e1[0]};
    for (int k = 0; k < 3; ++k) { rot[k] = e1[k]; rot[3 + k] = e2[k]; rot[6 + k] = L[k]; }
}
static void fill_pairs(const rt_scene* s, int np, const double rot[9], float4* Ao, float4* Bo) {
    const int n = static_cast<int>(s->spheres.size());
    const double K = static_cast<double>(RT_CULL_K);
    for (int p = 0; p < np; ++p) {
        float v[2][4];
        for (int h = 0; h < 2; ++h) {
            const int i = 2 * p + h;
            if (i >= n) { v[h][0] = v[h][1] = v[h][2] = 0.f; v[h][3] = RT_PAD_W; continue; }
            const rt_sphere_desc& sp = s->spheres[i];
            const double cw[3] = {sp.cx, sp.cy, sp.cz};
            for (int k = 0; k < 3; ++k) v[h][k] = static_cast<float>(rot[3 * k] * cw[0] + rot[3 * k + 1] * cw[1] + rot[3 * k + 2] * cw[2]);
            const double cc = cw[0] * cw[0] + cw[1] * cw[1] + cw[2] * cw[2], r2 = static_cast<double>(sp.radius) * sp.radius;
            v[h][3] = static_cast<float>(r2 - cc + K * (cc + r2));      // -(|c|^2 - r^2) + slack
        }
        Ao[p] = make_float4(v[0][0], v[1][0], v[0][1], v[1][1]);
        Bo[p] = make_float4(v[0][2], v[1][2], v[0][3], v[1][3]);
    }
}
static void fill_geom(const rt_scene* s, int n_pad, float4* G) {
    for (int i = 0; i < n_pad; ++i) G[i] = make_float4(0.f, 0.f, 0.f, 1.f);
    for (size_t i = 0; i < s->spheres.size(); ++i) G[i] = make_float4(s->spheres[i].cx, s->spheres[i].cy, s->spheres[i].cz, s->spheres[i].radius);
}

extern "C" {

int rt_context_create(int device, rt_context** out) {
    if (!out) return fail(RT_ERR_INVALID_ARGUMENT, "rt_context_create: out is NULL");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(RT_ERR_CUDA, "rt_context_create: no CUDA device available (%s); this library has no CPU fallback", cudaGetErrorString(e));
    if (device < 0 || device >= n) return fail(RT_ERR_INVALID_ARGUMENT, "rt_context_create: device %d out of range [0,%d)", device, n);
    RT_CUDA(cudaSetDevice(device));
    rt_context* c = new rt_context();
    c->device = device;
    if (cudaGetDeviceProperties(&c->prop, device) != cudaSuccess) { delete c; return fail(RT_ERR_CUDA, "cudaGetDeviceProperties failed"); }
    if (c->prop.major < 10) {
        const int maj = c->prop.major, min = c->prop.minor; delete c;
        return fail(RT_ERR_UNSUPPORTED, "rt_context_create: device is sm_%d%d; this build contains sm_100a code only", maj, min);
    }
    cudaDeviceGetAttribute(&c->clock_khz, cudaDevAttrClockRate, device);
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) { delete c; return fail(RT_ERR_CUDA, "cudaStreamCreate failed"); }
    c->own_stream = true;
    if (cudaMalloc(&c->counters, 8 * sizeof(u64)) != cudaSuccess) { delete c; return fail(RT_ERR_OUT_OF_MEMORY, "cudaMalloc(counters) failed"); }
    cudaMemset(c->counters, 0, 8 * sizeof(u64));
    *out = c;
    return RT_OK;
}
void rt_context_destroy(rt_context* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    ev_collect(c);
    for (auto& e : c->pool) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); }
    cudaFree(c->geom); cudaFree(c->wf_block); cudaFree(c->wf_live); cudaFree(c->accel_blob); cudaFree(c->sampler); cudaFree(c->extra_pairs);
    cudaFree(c->cullA); cudaFree(c->cullB); cudaFree(c->mats);
    cudaFree(c->hdr_own); cudaFree(c->ldr); cudaFree(c->ids); cudaFree(c->mids); cudaFree(c->partial); cudaFree(c->hits);
    cudaFree(c->jitter); cudaFree(c->lens); cudaFree(c->counters);
    if (c->own_stream) cudaStreamDestroy(c->stream);
    delete c;
}
int rt_set_stream(rt_context* c, void* s) {
    if (!c) return fail(RT_ERR_INVALID_ARGUMENT, "rt_set_stream: ctx is NULL");
    if (int r = ctx_use(c)) return r;
    RT_CUDA(cudaStreamSynchronize(c->stream));
    if (c->own_stream) { cudaStreamDestroy(c->stream); c->own_stream = false; }
    c->stream = static_cast<cudaStream_t>(s);
    return RT_OK;
}
void* rt_get_stream(rt_context* c) { return c ? static_cast<void*>(c->stream) : nullptr; }
int rt_synchronize(rt_context* c) {
    if (!c) return fail(RT_ERR_INVALID_ARGUMENT, "rt_synchronize: ctx is NULL");
    if (int r = ctx_use(c)) return r;
    RT_CUDA(cudaStreamSynchronize(c->stream));
    return RT_OK;
}
int rt_set_exposure(rt_context* c, float stops) { if (!c) return fail(RT_ERR_INVALID_ARGUMENT, "rt_set_exposure: ctx is NULL"); c->exposure_stops = stops; return RT_OK; }
int rt_set_seed(rt_context* c, uint32_t seed) { if (!c) return fail(RT_ERR_INVALID_ARGUMENT, "rt_set_seed: ctx is NULL"); c->seed = seed; return RT_OK; }
int rt_sampler_reset(rt_context* c) {
    if (!c) return fail(RT_ERR_INVALID_ARGUMENT, "rt_sampler_reset: ctx is NULL");
    if (!c->sampler) return fail(RT_ERR_NOT_READY, "rt_sampler_reset: no scene set");
    if (int r = ctx_use(c)) return r;
    RT_CUDA(cudaMemsetAsync(c->sampler, 0, c->n_sampler * sizeof(u64), c->stream));
    return RT_OK;
}
static int gshift_of_slots(int slots) { return slots == 0 || slots == 8 ? 3 : slots == 16 ? 4 : slots == 32 ? 5 : -1; }
int rt_set_accel_group_slots(rt_context* c, int slots) {
    if (!c || gshift_of_slots(slots) < 0) return fail(RT_ERR_INVALID_ARGUMENT, "rt_set_accel_group_slots: slots must be 0 (default), 8, 16 or 32");
    c->accel_gshift_min = gshift_of_slots(slots);
    return RT_OK;
}
int rt_set_chunk_spp(rt_context* c, int chunk) {
    if (!c || chunk < 0) return fail(RT_ERR_INVALID_ARGUMENT, "rt_set_chunk_spp: bad argument");
    c->chunk_spp = chunk; return RT_OK;
}

int rt_set_scene(rt_context* c, const rt_scene* s) {
    if (!c || !s) return fail(RT_ERR_INVALID_ARGUMENT, "rt_set_scene: NULL argument");
    if (s->sky_texture < 0) return fail(RT_ERR_NOT_READY, "rt_set_scene: the scene has no sky (rt_scene_set_sky)");
    if (int r = ctx_use(c)) return r;
    const int n = static_cast<int>(s->spheres.size());
    const int n_pad = std::max(32, (n + 31) / 32 * 32);
    const int np = n_pad / 2;
    if (n > 65534) return fail(RT_ERR_UNSUPPORTED, "rt_set_scene: %d spheres; the first-hit records index spheres with 16 bits (at most 65534)", n);
    // a list that does not fit next to the candidate lists of a 256-thread CTA stays in global memory (slower, any size)
    c->global_scene = cta_smem_bytes(n_pad, 256) > static_cast<size_t>(c->prop.sharedMemPerBlockOptin);
    const bool has_light = !s->lights.empty();
    host::V3 Ln{0.f, 0.f, 0.f};
    if (has_light) Ln = host::normalize3({s->lights[0].dir[0], s->lights[0].dir[1], s->lights[0].dir[2]});   // light.cpp:7
    const double L[3] = {Ln.x, Ln.y, Ln.z};
    std::vector<float4> A(np), B(np);
    std::vector<rt::DevMat> M(std::max(n, 1));
    double rot0[9];
    light_frame(Ln, has_light, rot0);
    fill_pairs(s, np, rot0, A.data(), B.data());
    for (int i = 0; i < n; ++i) {
        const rt_material_desc& m = s->materials[s->spheres[i].material];
        rt::DevMat dm{}; dm.kind = m.kind; dm.smoothness = m.smoothness; dm.ior = m.ior; dm.mat_index = s->spheres[i].material;
        if (m.texture >= 0) {
            const rt_texture_desc& t = s->textures[m.texture];
            dm.tex_kind = t.kind; dm.tiling = t.tiling;
            for (int k = 0; k < 4; ++k) { dm.c0[k] = t.c0[k]; dm.c1[k] = t.c1[k]; }
        }
        M[i] = dm;
    }
    if (int r = dev_realloc(c->cullA, np)) return r;
    if (int r = dev_realloc(c->cullB, np)) return r;
    if (int r = dev_realloc(c->mats, M.size())) return r;
    c->n_sampler = 2 * std::max<size_t>(1, s->materials.size());
    if (int r = dev_realloc(c->sampler, 2 * c->n_sampler)) return r;    // [indices][snapshot taken before light 0's pass of a batch]
    RT_CUDA(cudaMemsetAsync(c->sampler, 0, 2 * c->n_sampler * sizeof(u64), c->stream));      // the reference's indices start at 0 with the scene
    std::vector<float4> G(n_pad);                                              // padded: the kernels stage 2 * n_pairs_pad entries
    fill_geom(s, n_pad, G.data());
    if (int r = dev_realloc(c->geom, G.size())) return r;
    RT_CUDA(cudaMemcpyAsync(c->geom, G.data(), G.size() * sizeof(float4), cudaMemcpyHostToDevice, c->stream));
    RT_CUDA(cudaMemcpyAsync(c->cullA, A.data(), np * sizeof(float4), cudaMemcpyHostToDevice, c->stream));
    RT_CUDA(cudaMemcpyAsync(c->cullB, B.data(), np * sizeof(float4), cudaMemcpyHostToDevice, c->stream));
    RT_CUDA(cudaMemcpyAsync(c->mats, M.data(), M.size() * sizeof(rt::DevMat), cudaMemcpyHostToDevice, c->stream));
    RT_CUDA(cudaStreamSynchronize(c->stream));
    rt::RenderParams& b = c->base;
    b.cullA = c->cullA; b.cullB = c->cullB; b.mats = c->mats; b.sphere_geom = c->geom;
    for (int k = 0; k < 9; ++k) b.rot[k] = static_cast<float>(rot0[k]);
    b.sampler = c->sampler;
    b.n_spheres = n; b.n_pairs_pad = np; b.n_groups = n_pad / 32; b.n_lights = has_light ? 1 : 0;
    b.Lx = b.Ly = b.Lz = 0.f; b.lcx = b.lcy = b.lcz = 0.f;
    if (has_light) {
        const rt_light_desc& l = s->lights[0];
        b.Lx = Ln.x; b.Ly = Ln.y; b.Lz = Ln.z;
        b.lcx = l.luminance * l.color[0]; b.lcy = l.luminance * l.color[1]; b.lcz = l.luminance * l.color[2];   // light.cpp:27
    }
    // Lights 1..: Material::Shade sums over m_lights (material.cpp:4-13).  Radiance is linear in the emitters and the counter
    // RNG makes the paths independent of the lights, so light k > 0 is one more launch over the SAME paths with only that
    // light switched on and the sky switched off, accumulated into the same running sum (rt_render).
    c->extra_lights.clear();
    if (s->lights.size() > 1) {
        const size_t ne = s->lights.size() - 1;
        if (int r = dev_realloc(c->extra_pairs, 2 * ne * np)) return r;       // per further light: its own light-aligned pair arrays [A][B]
        std::vector<float4> PE(2 * ne * np);
        for (size_t k = 1; k < s->lights.size(); ++k) {
            const rt_light_desc& l = s->lights[k];
            const host::V3 Lk = host::normalize3({l.dir[0], l.dir[1], l.dir[2]});
            rt_context::ExtraLight e{};
            e.L[0] = Lk.x; e.L[1] = Lk.y; e.L[2] = Lk.z;
            for (int q = 0; q < 3; ++q) e.lc[q] = l.luminance * l.color[q];
            double rk[9];
            light_frame(Lk, true, rk);
            for (int q = 0; q < 9; ++q) e.rot[q] = static_cast<float>(rk[q]);
            e.cullA = c->extra_pairs + 2 * (k - 1) * np; e.cullB = e.cullA + np;
            fill_pairs(s, np, rk, PE.data() + 2 * (k - 1) * np, PE.data() + 2 * (k - 1) * np + np);
            c->extra_lights.push_back(e);
        }
        RT_CUDA(cudaMemcpyAsync(c->extra_pairs, PE.data(), PE.size() * sizeof(float4), cudaMemcpyHostToDevice, c->stream));
        RT_CUDA(cudaStreamSynchronize(c->stream));
    }
    const rt_texture_desc& sky = s->textures[s->sky_texture];              // Emissive::Emit at uv = (0,0): material.cpp:172-175
    const float* sc = sky.c0;                                             // checker at (0,0): u%2 == v%2 -> colour 0
    b.skyx = s->sky_luminance * sc[0]; b.skyy = s->sky_luminance * sc[1]; b.skyz = s->sky_luminance * sc[2];
    c->n_spheres = n; c->n_pad = n_pad; c->scene_set = true;
    // opt in to large dynamic shared memory; pick the CTA size that keeps the most warps resident per SM (one scene copy per
    // CTA, so big scenes want big CTAs) and size the persistent grid to exactly fill the machine
    const int optin = static_cast<int>(c->prop.sharedMemPerBlockOptin);
    RT_CUDA(cudaFuncSetAttribute(rt::rt_render_kernel<rt::MODE_CULL>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    RT_CUDA(cudaFuncSetAttribute(rt::rt_render_kernel<rt::MODE_NO_CULL>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    RT_CUDA((cudaFuncSetAttribute(rt::rt_render_kernel<rt::MODE_CULL, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin)));
    RT_CUDA(cudaFuncSetAttribute(rt::rt_primary_kernel<rt::MODE_CULL>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    RT_CUDA(cudaFuncSetAttribute(rt::rt_primary_kernel<rt::MODE_NO_CULL>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    RT_CUDA(cudaFuncSetAttribute(rt::wf_intersect_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    RT_CUDA(cudaFuncSetAttribute(rt::wf_intersect_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    if (int r = optin_smem(rt::MODE_CULL, 5, optin)) return r;
    if (int r = optin_smem(rt::MODE_NO_CULL, 5, optin)) return r;
    if (int r = optin_smem(rt::MODE_CULL_GLOBAL, 5, optin)) return r;
    RT_CUDA(cudaFuncSetAttribute(rt::rt_primary_kernel<rt::MODE_CULL_GLOBAL>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    int best_tpb = 0, best_bps = 0;
    c->path_tpb = 0; c->path_bps = 0;
    const int scene_mode = c->global_scene ? rt::MODE_CULL_GLOBAL : rt::MODE_CULL;
    for (int tpb = 128; tpb <= RT_MAX_TPB; tpb += 128) {
        const size_t smem = cta_smem_bytes(n_pad, tpb, c->global_scene);
        if (smem > static_cast<size_t>(optin)) break;
        int bps = 0;
        if (!c->global_scene) {                              // the one-kernel schedule and the wavefront layout stage the list: small scenes only
            RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, rt::rt_render_kernel<rt::MODE_CULL>, tpb, smem));
            if (bps * tpb > best_bps * best_tpb) { best_tpb = tpb; best_bps = bps; }
        }
        RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, path_kernel(scene_mode, false, 5, false), tpb, smem));
        if (bps * tpb > c->path_bps * c->path_tpb) { c->path_tpb = tpb; c->path_bps = bps; }
    }
    if (c->global_scene) { best_tpb = c->path_tpb; best_bps = c->path_bps; }
    c->first_tpb = 0; c->first_bps = 0;
    for (int tpb = 128; tpb <= std::min(RT_MAX_TPB, RT_FIRST_RAYS == 1 ? RT_MAX_TPB : RT_FIRST_MAX_TPB); tpb += 128) {
        const size_t smem = cta_smem_bytes(n_pad, tpb, c->global_scene);
        if (smem > static_cast<size_t>(optin)) break;
        int bps = 0;
        RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, first_hit_kernel(scene_mode, 5), tpb, smem));
        if (bps * tpb > c->first_bps * c->first_tpb) { c->first_tpb = tpb; c->first_bps = bps; }
    }
    if (c->first_bps < 1) return fail(RT_ERR_UNSUPPORTED, "rt_set_scene: first-hit kernel does not fit on an SM with %zu B shared memory", cta_smem_bytes(n_pad, 128, c->global_scene));
    if (c->path_bps < 1) return fail(RT_ERR_UNSUPPORTED, "rt_set_scene: path kernel does not fit on an SM with %zu B shared memory", cta_smem_bytes(n_pad, 256));
    if (best_bps < 1) return fail(RT_ERR_UNSUPPORTED, "rt_set_scene: render kernel does not fit on an SM with %zu B shared memory", cta_smem_bytes(n_pad, 256));
    c->tpb = best_tpb;
    c->render_blocks_per_sm = best_bps;
    c->wf_bps = 1;
    if (!c->global_scene) RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c->wf_bps, rt::wf_intersect_kernel<false>, best_tpb, cta_smem_bytes(n_pad, best_tpb)));
    if (c->wf_bps < 1) c->wf_bps = 1;
    auto safe_inv = [](float d) { const float dd = std::fabs(d) < 1e-18f ? std::copysign(1e-18f, d) : d; return 1.0f / dd; };   // rt_bvh.cuh safe_inv
    b.iLx = safe_inv(b.Lx); b.iLy = safe_inv(b.Ly); b.iLz = safe_inv(b.Lz);
    for (auto& e : c->extra_lights) for (int q = 0; q < 3; ++q) e.iL[q] = safe_inv(e.L[q]);
    b.primary_pass = 1;
    return build_accel(c, s);
}

int rt_set_camera(rt_context* c, const rt_camera_params* cam) {
    if (!c || !cam) return fail(RT_ERR_INVALID_ARGUMENT, "rt_set_camera: NULL argument");
    if (!finite3(cam->origin) || !finite3(cam->look_at) || !(cam->vfov_deg > 0.f && cam->vfov_deg < 180.f) || !(cam->aspect > 0.f))
        return fail(RT_ERR_INVALID_ARGUMENT, "rt_set_camera: non-finite position or fov/aspect out of range");
    rt_camera_basis b; host::camera_basis(*cam, b);
    rt::DevCamera& d = c->base.cam;
    d.ox = b.origin[0]; d.oy = b.origin[1]; d.oz = b.origin[2];
    d.xx = b.x[0]; d.xy = b.x[1]; d.xz = b.x[2];
    d.yx = b.y[0]; d.yy = b.y[1]; d.yz = b.y[2];
    d.ix = b.image_origin[0]; d.iy = b.image_origin[1]; d.iz = b.image_origin[2];
    d.aperture = b.aperture; d.focus = b.focus;
    c->camera_set = true;
    return RT_OK;
}

int rt_framebuffer_resize(rt_context* c, int W, int H) {
    if (!c || W <= 0 || H <= 0 || static_cast<int64_t>(W) * H > (1ll << 30)) return fail(RT_ERR_INVALID_ARGUMENT, "rt_framebuffer_resize: bad size %dx%d", W, H);
    if (int r = ctx_use(c)) return r;
    RT_CUDA(cudaStreamSynchronize(c->stream));
    const size_t n = static_cast<size_t>(W) * H;
    if (int r = dev_realloc(c->hdr_own, n)) return r;
    if (int r = dev_realloc(c->ldr, n)) return r;
    if (int r = dev_realloc(c->ids, n)) return r;
    if (int r = dev_realloc(c->mids, n)) return r;
    c->hdr_ext = nullptr;
    c->W = W; c->H = H;
    return rt_framebuffer_clear(c);
}
int rt_framebuffer_clear(rt_context* c) {
    if (!c || !c->hdr()) return fail(RT_ERR_NOT_READY, "rt_framebuffer_clear: no framebuffer");
    if (int r = ctx_use(c)) return r;
    const size_t n = static_cast<size_t>(c->W) * c->H;
    RT_CUDA(cudaMemsetAsync(c->hdr(), 0, n * sizeof(float4), c->stream));     // InitBuffers, app.cpp:112-119
    RT_CUDA(cudaMemsetAsync(c->ldr, 0, n * sizeof(uint32_t), c->stream));
    return RT_OK;
}
void* rt_hdr_device_ptr(rt_context* c) { return c ? c->hdr() : nullptr; }
int rt_hdr_bind_external(rt_context* c, void* p) {
    if (!c) return fail(RT_ERR_INVALID_ARGUMENT, "rt_hdr_bind_external: ctx is NULL");
    if (!c->hdr_own) return fail(RT_ERR_NOT_READY, "rt_hdr_bind_external: call rt_framebuffer_resize first");
    if (p && (reinterpret_cast<uintptr_t>(p) & 15u)) return fail(RT_ERR_INVALID_ARGUMENT, "rt_hdr_bind_external: pointer must be 16-byte aligned");
    c->hdr_ext = static_cast<float4*>(p);
    return RT_OK;
}

// Fills RenderParams for a launch and uploads the per-frame Halton tables.
static int prepare_frames(rt_context* c, u64 first_sample, int spp, int max_depth, uint32_t flags, rt::RenderParams& p) {
    p = c->base;
    p.W = c->W; p.H = c->H; p.tiles_x = (c->W + 7) / 8; p.tiles_y = (c->H + 3) / 4;
    p.first_sample = first_sample; p.spp = spp; p.max_depth = max_depth; p.flags = flags; p.seed = c->seed;
    p.exposure = static_cast<float>(std::pow(2, c->exposure_stops));          // spheres-app.cpp:174
    p.hdr = c->hdr(); p.work_counter = c->counters; p.stats = c->counters + 1;
    // frame jitter Halton2D(n; 2,3) and lens samples HaltonSampleDisk(n+i+j; 4,5), spheres-app.cpp:140,152: built on the device,
    // on the context's stream (no host table, no copy, no synchronisation: rt_render stays asynchronous)
    const size_t n_lens = static_cast<size_t>(spp) + c->W + c->H;
    if (c->jitter_cap < static_cast<size_t>(spp) || c->lens_cap < n_lens) {
        RT_CUDA(cudaStreamSynchronize(c->stream));              // a previous launch may still read the tables being replaced (growth only)
        if (c->jitter_cap < static_cast<size_t>(spp)) { const size_t cap = std::max<size_t>(spp, 2 * c->jitter_cap); if (int r = dev_realloc(c->jitter, cap)) return r; c->jitter_cap = cap; }
        if (c->lens_cap < n_lens) { const size_t cap = std::max(n_lens, 2 * c->lens_cap); if (int r = dev_realloc(c->lens, cap)) return r; c->lens_cap = cap; }
    }
    rt::rt_tables_kernel<<<static_cast<unsigned>((n_lens + 255) / 256), 256, 0, c->stream>>>(first_sample, spp, static_cast<int>(n_lens), c->jitter, c->lens);
    ++c->launches;
    p.jitter = c->jitter; p.lens = c->lens;
    return RT_OK;
}

// Wavefront schedule of one rt_render call: intersect / shade kernels over SoA path state until the work queue drains.
static int render_wavefront(rt_context* c, const rt::RenderParams& p, uint32_t flags) {
    const u64 n_items = static_cast<u64>(p.tiles_x) * p.tiles_y * 32ull * static_cast<u64>(p.n_chunks);
    const int want = static_cast<int>(std::min<u64>(n_items, 1ull << 21));
    if (c->wf.n_slots != want) {
        if (c->wf_block) { cudaFree(c->wf_block); c->wf_block = nullptr; c->wf.n_slots = 0; }
        const size_t per = static_cast<size_t>(want) * 16;                       // every array is 16 B per slot
        RT_CUDA(cudaMalloc(&c->wf_block, per * 7));
        char* b = static_cast<char*>(c->wf_block);
        c->wf.o = reinterpret_cast<float4*>(b); c->wf.d = reinterpret_cast<float4*>(b + per); c->wf.thr = reinterpret_cast<float4*>(b + 2 * per);
        c->wf.rad = reinterpret_cast<float4*>(b + 3 * per); c->wf.pend = reinterpret_cast<float4*>(b + 4 * per);
        c->wf.acc = reinterpret_cast<float4*>(b + 5 * per); c->wf.ids = reinterpret_cast<int4*>(b + 6 * per);
        c->wf.n_slots = want;
    }
    if (!c->wf_live) RT_CUDA(cudaMalloc(&c->wf_live, sizeof(unsigned int)));
    RT_CUDA(cudaMemsetAsync(c->wf.ids, 0, static_cast<size_t>(want) * 16, c->stream));       // all slots: "no path yet"
    const size_t smem = cta_smem_bytes(c->n_pad, c->tpb);
    const int grid_i = c->prop.multiProcessorCount * c->wf_bps;
    const int grid_s = c->prop.multiProcessorCount * 8;
    rt::wf_shade_kernel<<<grid_s, 256, 0, c->stream>>>(p, c->wf, nullptr);                     // fetch work + primary rays
    ++c->launches;
    const int kBatch = 16;
    for (int round = 0; round < (1 << 20); ++round) {
        for (int k = 0; k < kBatch; ++k) {
            if (flags & RT_FLAG_DEBUG_NO_CULL_) rt::wf_intersect_kernel<true><<<grid_i, c->tpb, smem, c->stream>>>(p, c->wf);
            else rt::wf_intersect_kernel<false><<<grid_i, c->tpb, smem, c->stream>>>(p, c->wf);
            const bool last = (k == kBatch - 1);
            if (last) RT_CUDA(cudaMemsetAsync(c->wf_live, 0, sizeof(unsigned int), c->stream));
            rt::wf_shade_kernel<<<grid_s, 256, 0, c->stream>>>(p, c->wf, last ? c->wf_live : nullptr);
            c->launches += 2;
        }
        unsigned int live = 0;
        RT_CUDA(cudaMemcpyAsync(&live, c->wf_live, sizeof live, cudaMemcpyDeviceToHost, c->stream));
        RT_CUDA(cudaStreamSynchronize(c->stream));
        if (live == 0) return RT_OK;
    }
    return fail(RT_ERR_CUDA, "rt_render (wavefront): the work queue did not drain");
}

// Switches the launch parameters to light k (k >= 1): that light alone, sky off, same paths; sample counts and the path /
// closest-ray statistics come from light 0's launches.
static void select_extra_light(const rt_context* c, int k, rt::RenderParams& p) {
    const rt_context::ExtraLight& e = c->extra_lights[k - 1];
    p.Lx = e.L[0]; p.Ly = e.L[1]; p.Lz = e.L[2]; p.lcx = e.lc[0]; p.lcy = e.lc[1]; p.lcz = e.lc[2];
    p.iLx = e.iL[0]; p.iLy = e.iL[1]; p.iLz = e.iL[2]; p.cullA = e.cullA; p.cullB = e.cullB;
    for (int q = 0; q < 9; ++q) p.rot[q] = e.rot[q];
    p.skyx = p.skyy = p.skyz = 0.f; p.primary_pass = 0;
}
static int ensure_partial(rt_context* c, size_t n_pix, int n_chunks) {
    const size_t need = n_pix * static_cast<size_t>(n_chunks) * sizeof(float4);
    if (c->partial_bytes < need) {
        if (c->partial) { RT_CUDA(cudaStreamSynchronize(c->stream)); cudaFree(c->partial); c->partial = nullptr; c->partial_bytes = 0; }
        RT_CUDA(cudaMalloc(&c->partial, need)); c->partial_bytes = need;
    }
    return RT_OK;
}

// Round-1 schedule (RT_FLAG_DEBUG_ONE_KERNEL, RT_FLAG_WAVEFRONT): one launch per light over the whole sample range.
static int render_one_kernel(rt_context* c, rt::RenderParams& p, int spp, uint32_t flags) {
    int chunk = c->chunk_spp;
    if (chunk <= 0) { const int n_chunks = spp <= 16 ? 1 : std::min(16, spp / 8); chunk = (spp + n_chunks - 1) / n_chunks; }
    chunk = std::min(chunk, spp);
    p.chunk_spp = chunk; p.n_chunks = (spp + chunk - 1) / chunk;
    const size_t n_pix = static_cast<size_t>(c->W) * c->H;
    p.partial = nullptr;
    if (p.n_chunks > 1) { if (int r = ensure_partial(c, n_pix, p.n_chunks)) return r; p.partial = c->partial; }
    const size_t smem = cta_smem_bytes(c->n_pad, c->tpb);
    const int grid = c->prop.multiProcessorCount * c->render_blocks_per_sm;
    const int n_launch_lights = 1 + static_cast<int>(c->extra_lights.size());
    for (int k = 0; k < n_launch_lights; ++k) {
        if (k > 0) select_extra_light(c, k, p);
        RT_CUDA(cudaMemsetAsync(c->counters, 0, sizeof(u64), c->stream));
        if (flags & RT_FLAG_WAVEFRONT) { if (int r = render_wavefront(c, p, flags)) return r; }
        else if (flags & RT_FLAG_ACCEL) {
            void* args[] = {const_cast<rt::RenderParams*>(&p)};
            RT_CUDA(cudaLaunchKernel(accel_render_kernel(p.accel.gshift), dim3(c->prop.multiProcessorCount * c->accel_bps), dim3(c->accel_tpb), args,
                                     accel_cta_smem_bytes(p.accel, c->accel_tpb), c->stream));
            ++c->launches;
        } else if (flags & RT_FLAG_HALTON) {
            rt::rt_render_kernel<rt::MODE_CULL, true><<<grid, c->tpb, smem, c->stream>>>(p);
            ++c->launches;
        } else {
            if (flags & RT_FLAG_DEBUG_NO_CULL_) rt::rt_render_kernel<rt::MODE_NO_CULL><<<grid, c->tpb, smem, c->stream>>>(p);
            else rt::rt_render_kernel<rt::MODE_CULL><<<grid, c->tpb, smem, c->stream>>>(p);
            ++c->launches;
        }
        if (p.n_chunks > 1) {
            rt::rt_combine_kernel<<<static_cast<unsigned>((n_pix + 255) / 256), 256, 0, c->stream>>>(p.hdr, p.partial, p.n_chunks, n_pix, 0);
            ++c->launches;
        }
    }
    return RT_OK;
}

#ifndef RT_LADDER_GRAN
#define RT_LADDER_GRAN 8
#endif
#ifndef RT_BATCH_CAP_PIXELS
#define RT_BATCH_CAP_PIXELS (4u << 20)
#endif
#ifndef RT_BATCH_CAP_MIB
#define RT_BATCH_CAP_MIB 512   /* first-hit records of one batch of an image up to 4 Mpixel (larger images: 256 MiB or two running-sum buffers) */
#endif
// Two-stage schedule (default): per batch of frames one first-hit launch, then one path launch per light.
static int render_two_stage(rt_context* c, rt::RenderParams& p, int spp, int max_depth, uint32_t flags) {
    const size_t n_pix = static_cast<size_t>(c->W) * c->H;
    const bool accel = (flags & RT_FLAG_ACCEL) != 0, halton = (flags & RT_FLAG_HALTON) != 0;
    const int mode = accel ? rt::MODE_ACCEL : (flags & RT_FLAG_DEBUG_NO_CULL_) ? rt::MODE_NO_CULL : c->global_scene ? rt::MODE_CULL_GLOBAL : rt::MODE_CULL;
    const int gshift = accel ? p.accel.gshift : 5;
    const bool direct = max_depth == 0 && !halton && mode != rt::MODE_NO_CULL;
    const int tpb = accel ? c->accel_tpb : c->path_tpb, bps = accel ? c->accel_bps : c->path_bps;
    const size_t smem = accel ? accel_cta_smem_bytes(p.accel, tpb) : cta_smem_bytes(c->n_pad, tpb, c->global_scene);
    const int grid = c->prop.multiProcessorCount * bps;
    const void* k_first = first_hit_kernel(mode, gshift);
    const void* k_path = path_kernel(mode, halton, gshift, direct);
    // stage 1 of the brute-force query runs several rays per lane with its own register budget, hence its own CTA shape
    const bool own1 = mode == rt::MODE_CULL || mode == rt::MODE_CULL_GLOBAL;
    const int tpb1 = own1 ? c->first_tpb : tpb, grid1 = own1 ? c->prop.multiProcessorCount * c->first_bps : grid;
    const size_t smem1 = own1 ? cta_smem_bytes(c->n_pad, tpb1, c->global_scene) : smem;
    // batch: as many frames as fit first-hit records of 512 MiB (2 B per pixel and frame; images above 4 Mpixel: 256 MiB or two running-sum
    // buffers, whichever is more), a multiple of 16, at most 4096: 128 frames at 1920x1080, 16 at 3840x2160 -- the records are the only
    // per-batch allocation
    // (measured at 1920x1080, 256 frames per call: 64 / 128 / 256 frames per batch = 295.5 / 289.9 / 292.5 ms -- profiles/r02/r02x_batch_size_probe.txt)
    const size_t cap_bytes = n_pix <= RT_BATCH_CAP_PIXELS ? (static_cast<size_t>(RT_BATCH_CAP_MIB) << 20) : std::max<size_t>(256ull << 20, n_pix * 32);
    int batch = static_cast<int>(std::min<size_t>(4096, std::max<size_t>(16, (cap_bytes / (n_pix * 2)) & ~static_cast<size_t>(15))));
    batch = std::min(batch, (spp + 15) & ~15);
    const size_t hits_need = n_pix * static_cast<size_t>(batch) * sizeof(uint16_t);
    if (c->hits_bytes < hits_need) {
        if (c->hits) { RT_CUDA(cudaStreamSynchronize(c->stream)); cudaFree(c->hits); c->hits = nullptr; c->hits_bytes = 0; }
        RT_CUDA(cudaMalloc(&c->hits, hits_need)); c->hits_bytes = hits_need;
    }
    p.hits = c->hits;
    const int n_launch_lights = 1 + static_cast<int>(c->extra_lights.size());
    const rt::RenderParams base = p;
    const u64 lanes = static_cast<u64>(grid) * tpb;
    for (int s0 = 0; s0 < spp; s0 += batch) {
        p = base;
        p.batch_s0 = s0; p.batch_n = std::min(batch, spp - s0); p.hit_stride = (p.batch_n + 15) & ~15;
        // work items of stage 2: (pixel, chunk of the batch's samples), handed out chunk 0 first.  Chunk 0 adds to the pixel's running
        // sum in place (the reference's sequential order); chunks 1.. produce partial sums that rt_combine_kernel adds in fixed order.
        // The persistent kernel's tail is one item long, so large images get chunks of halving size (one partial buffer per chunk after
        // the first); small images are split evenly so that every lane sees >= ~4 items.
        int chunk = 8, n_chunks = 1;
        p.n_ladder = 0;
        const int n8 = (p.batch_n + 7) / 8;                     // 8-sample groups that hold samples
        if (c->chunk_spp > 0) { chunk = (c->chunk_spp + 7) & ~7; n_chunks = (p.batch_n + chunk - 1) / chunk; }
        else if (4 * lanes > n_pix) {
            const u64 want = (4 * lanes) / n_pix;
            const int nc = static_cast<int>(std::min<u64>(want, static_cast<u64>(n8)));
            chunk = (((p.batch_n + nc - 1) / nc) + 7) & ~7; n_chunks = (p.batch_n + chunk - 1) / chunk;
        } else {
            // a ladder of halving chunks, e.g. 64 frames = 32 + 16 + 8 + 8: the frames still queued behind a chunk take the whole image
            // as long as that chunk's heaviest item takes one lane (a glass pixel: ~0.4 ms per sample; the image: n_pix x 0.63 ns per
            // sample, i.e. more from ~0.6 Mpixel up), so the machine stays full while long items finish and the final tail is 8 samples
            // RT_LADDER_GRAN: frames per rung unit.  8 = finest (tail of 8 frames; the last two rungs share a 32-byte record sector);
            // 16 = a record sector belongs to exactly one work item (64 frames = 32 + 16 + 16, one partial buffer less).
            int g[9], n = 0, left = (p.batch_n + RT_LADDER_GRAN - 1) / RT_LADDER_GRAN;
            while (left > 1 && n < 7) { const int take = (left + 1) / 2; g[n++] = take; left -= take; }
            if (left > 0) g[n++] = left;
            p.n_ladder = n; n_chunks = n;
            p.ladder[0] = 0;
            for (int i = 0; i < n; ++i) p.ladder[i + 1] = p.ladder[i] + RT_LADDER_GRAN * g[i];
        }
        p.chunk_spp = chunk; p.n_chunks = n_chunks;
        p.partial = nullptr;
        if (n_chunks > 1) { if (int r = ensure_partial(c, n_pix, n_chunks - 1)) return r; p.partial = c->partial; }
        void* args[] = {&p};
        RT_CUDA(cudaLaunchKernel(k_first, dim3(grid1), dim3(tpb1), args, smem1, c->stream));
        ++c->launches;
        // RT_FLAG_HALTON with several lights: the reference draws once per vertex and loops over the lights there (material.cpp:4-13);
        // here light k re-walks the batch's paths, so the materials' indices are rewound to where light 0's pass found them and
        // every pass draws the same numbers (exactly so when paths run one after another, as with the reference itself)
        const bool rewind = halton && n_launch_lights > 1;
        if (rewind) RT_CUDA(cudaMemcpyAsync(c->sampler + c->n_sampler, c->sampler, c->n_sampler * sizeof(u64), cudaMemcpyDeviceToDevice, c->stream));
        for (int k = 0; k < n_launch_lights; ++k) {
            if (k > 0) select_extra_light(c, k, p);
            if (k > 0 && rewind) RT_CUDA(cudaMemcpyAsync(c->sampler, c->sampler + c->n_sampler, c->n_sampler * sizeof(u64), cudaMemcpyDeviceToDevice, c->stream));
            p.overwrite = ((flags & RT_FLAG_OVERWRITE) && s0 == 0 && k == 0) ? 1 : 0;
            RT_CUDA(cudaMemsetAsync(c->counters, 0, sizeof(u64), c->stream));
            RT_CUDA(cudaLaunchKernel(k_path, dim3(grid), dim3(tpb), args, smem, c->stream));
            ++c->launches;
            if (p.n_chunks > 1) {
                rt::rt_combine_kernel<<<static_cast<unsigned>((n_pix + 255) / 256), 256, 0, c->stream>>>(p.hdr, p.partial, p.n_chunks - 1, n_pix, 0);
                ++c->launches;
            }
        }
    }
    return RT_OK;
}

int rt_render(rt_context* c, uint64_t first_sample, int spp, int max_depth, uint32_t flags) {
    if (!c) return fail(RT_ERR_INVALID_ARGUMENT, "rt_render: ctx is NULL");
    if (!c->scene_set || !c->camera_set || !c->hdr()) return fail(RT_ERR_NOT_READY, "rt_render: scene, camera and framebuffer must be set first");
    if (first_sample < 1 || spp < 0 || max_depth < 0) return fail(RT_ERR_INVALID_ARGUMENT, "rt_render: first_sample >= 1, spp >= 0, max_depth >= 0 required");
    if (flags & ~(RT_FLAG_JITTER_OFF | RT_FLAG_LENS_OFF | RT_FLAG_WAVEFRONT | RT_FLAG_ACCEL | RT_FLAG_HALTON | RT_FLAG_OVERWRITE | RT_FLAG_DEBUG_NO_CULL_ | RT_FLAG_DEBUG_ONE_KERNEL_))
        return fail(RT_ERR_INVALID_ARGUMENT, "rt_render: unknown flag bits 0x%x", flags);
    const bool one_kernel = (flags & (RT_FLAG_WAVEFRONT | RT_FLAG_DEBUG_ONE_KERNEL_)) != 0;
    if ((flags & RT_FLAG_HALTON) && (flags & (RT_FLAG_WAVEFRONT | RT_FLAG_DEBUG_NO_CULL_))) return fail(RT_ERR_INVALID_ARGUMENT, "rt_render: RT_FLAG_HALTON cannot be combined with the wavefront layout or the no-cull debug mode");
    if ((flags & RT_FLAG_HALTON) && (flags & RT_FLAG_ACCEL) && one_kernel) return fail(RT_ERR_INVALID_ARGUMENT, "rt_render: RT_FLAG_HALTON with RT_FLAG_ACCEL runs on the two-stage schedule only");
    if ((flags & RT_FLAG_ACCEL) && (flags & (RT_FLAG_WAVEFRONT | RT_FLAG_DEBUG_NO_CULL_))) return fail(RT_ERR_INVALID_ARGUMENT, "rt_render: RT_FLAG_ACCEL cannot be combined with the wavefront layout or the no-cull debug mode");
    if ((flags & RT_FLAG_OVERWRITE) && one_kernel) return fail(RT_ERR_INVALID_ARGUMENT, "rt_render: RT_FLAG_OVERWRITE runs on the two-stage schedule only");
    if ((flags & RT_FLAG_ACCEL) && !c->accel_ok) return fail(RT_ERR_UNSUPPORTED, "rt_render: RT_FLAG_ACCEL unavailable for this scene: %s", c->accel_why.c_str());
    if ((flags & RT_FLAG_WAVEFRONT) && max_depth >= (1 << 22)) return fail(RT_ERR_INVALID_ARGUMENT, "rt_render: the wavefront layout packs the depth into 23 bits");
    if (!c->extra_lights.empty() && (flags & RT_FLAG_WAVEFRONT))
        return fail(RT_ERR_UNSUPPORTED, "rt_render: scenes with more than one light do not run on the wavefront layout");
    if (c->global_scene && !(flags & RT_FLAG_ACCEL) && (one_kernel || (flags & (RT_FLAG_HALTON | RT_FLAG_DEBUG_NO_CULL_))))
        return fail(RT_ERR_UNSUPPORTED, "rt_render: %d spheres do not fit in shared memory; such scenes run on the default two-stage schedule with the counter RNG only", c->n_spheres);
    if (spp == 0) return RT_OK;
    if (int r = ctx_use(c)) return r;
    rt::RenderParams p;
    if (int r = prepare_frames(c, first_sample, spp, max_depth, flags, p)) return r;
    EventPair ev = ev_begin(c, 0);
    const int rc = one_kernel ? render_one_kernel(c, p, spp, flags) : render_two_stage(c, p, spp, max_depth, flags);
    ev_end(c, ev);                                          // on every path: the pair goes back to the pool with the next collection
    if (rc) return rc;
    RT_CUDA(cudaGetLastError());
    return RT_OK;
}

// ------------------------------------------------------------------------------------------------
// rt_render_static: the two-stage render with ALL of its scratch memory (per-call Halton tables, first-hit records) in a
// file-scope device array instead of cudaMalloc'ed buffers.  No counterpart in the reference; added
// for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
// The array is held by every GPU that loads the library until the process exits.  Every call on a device shares it and it keeps
// the last call's contents: nothing in it has to start at zero (the tables and the records of every pixel / frame the path stage
// reads are rewritten by this call's own table and first-hit launches first), but two calls whose kernels overlap on the GPU
// (different streams / contexts on one device) race on it.
// ------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_static_scratch[RT_STATIC_SCRATCH_BYTES];
static unsigned char* g_static_scratch_ptr[64];          // per device: the array's DEVICE address (the symbol itself is a host-side stand-in).
                                                         // Two host threads racing through the first call on a device store the same value.
static int static_scratch_of_current_device(int device, unsigned char** out) {
    if (device < 0 || device >= 64) return fail(RT_ERR_UNSUPPORTED, "rt_render_static: device %d out of the supported range [0,64)", device);
    if (!g_static_scratch_ptr[device]) {                 // once per device, with that device current (ctx_use ran)
        void* ptr = nullptr;
        RT_CUDA(cudaGetSymbolAddress(&ptr, g_static_scratch));
        if (!ptr) return fail(RT_ERR_CUDA, "rt_render_static: cudaGetSymbolAddress returned a null address");
        g_static_scratch_ptr[device] = static_cast<unsigned char*>(ptr);
    }
    *out = g_static_scratch_ptr[device];
    return RT_OK;
}
size_t rt_render_static_scratch_bytes(int width, int height, int spp) {
    if (width <= 0 || height <= 0 || spp < 0) return 0;
    auto a256 = [](size_t x) { return (x + 255u) & ~static_cast<size_t>(255u); };
    const size_t n_pix = static_cast<size_t>(width) * height;
    return a256(static_cast<size_t>(spp) * sizeof(float2)) + a256((static_cast<size_t>(spp) + width + height) * sizeof(float2))
         + n_pix * RT_STATIC_BATCH * sizeof(uint16_t);
}
int rt_render_static(rt_context* c, uint64_t first_sample, int spp, int max_depth, uint32_t flags) {
    if (!c) return fail(RT_ERR_INVALID_ARGUMENT, "rt_render_static: ctx is NULL");
    if (!c->scene_set || !c->camera_set || !c->hdr()) return fail(RT_ERR_NOT_READY, "rt_render_static: scene, camera and framebuffer must be set first");
    if (first_sample < 1 || spp < 0 || max_depth < 0) return fail(RT_ERR_INVALID_ARGUMENT, "rt_render_static: first_sample >= 1, spp >= 0, max_depth >= 0 required");
    if (flags & ~(RT_FLAG_JITTER_OFF | RT_FLAG_LENS_OFF | RT_FLAG_ACCEL | RT_FLAG_OVERWRITE))
        return fail(RT_ERR_INVALID_ARGUMENT, "rt_render_static: only RT_FLAG_JITTER_OFF, RT_FLAG_LENS_OFF, RT_FLAG_ACCEL and RT_FLAG_OVERWRITE are supported (flags 0x%x)", flags);
    if ((flags & RT_FLAG_ACCEL) && !c->accel_ok) return fail(RT_ERR_UNSUPPORTED, "rt_render_static: RT_FLAG_ACCEL unavailable for this scene: %s", c->accel_why.c_str());
    if (c->global_scene && !(flags & RT_FLAG_ACCEL)) return fail(RT_ERR_UNSUPPORTED, "rt_render_static: %d spheres do not fit in shared memory; use rt_render", c->n_spheres);
    const size_t need = rt_render_static_scratch_bytes(c->W, c->H, spp);
    if (need > static_cast<size_t>(RT_STATIC_SCRATCH_BYTES))
        return fail(RT_ERR_UNSUPPORTED, "rt_render_static: %dx%d with %d spp needs %zu B of scratch, the static array holds %zu B "
                    "(32 B per pixel + 16 B per frame + 8 B per column and row); use rt_render, or fewer frames per call", c->W, c->H, spp, need,
                    static_cast<size_t>(RT_STATIC_SCRATCH_BYTES));
    if (spp == 0) return RT_OK;
    if (int r = ctx_use(c)) return r;
    unsigned char* scratch = nullptr;
    if (int r = static_scratch_of_current_device(c->device, &scratch)) return r;
    auto a256 = [](size_t x) { return (x + 255u) & ~static_cast<size_t>(255u); };
    const size_t n_pix = static_cast<size_t>(c->W) * c->H;
    const size_t n_lens = static_cast<size_t>(spp) + c->W + c->H;
    float2* jitter = reinterpret_cast<float2*>(scratch);
    float2* lens = reinterpret_cast<float2*>(scratch + a256(static_cast<size_t>(spp) * sizeof(float2)));
    uint16_t* hits = reinterpret_cast<uint16_t*>(reinterpret_cast<unsigned char*>(lens) + a256(n_lens * sizeof(float2)));
    // launch parameters: as prepare_frames + render_two_stage build them, with the scratch pointers above, batches of RT_STATIC_BATCH
    // frames and ONE work item per pixel and batch (sequential per-pixel sums, the reference's order; no partial buffers)
    rt::RenderParams p = c->base;
    p.W = c->W; p.H = c->H; p.tiles_x = (c->W + 7) / 8; p.tiles_y = (c->H + 3) / 4;
    p.first_sample = first_sample; p.spp = spp; p.max_depth = max_depth; p.flags = flags; p.seed = c->seed;
    p.exposure = static_cast<float>(std::pow(2, c->exposure_stops));
    p.hdr = c->hdr(); p.work_counter = c->counters; p.stats = c->counters + 1;
    p.jitter = jitter; p.lens = lens; p.hits = hits; p.partial = nullptr;
    rt::rt_tables_kernel<<<static_cast<unsigned>((n_lens + 255) / 256), 256, 0, c->stream>>>(first_sample, spp, static_cast<int>(n_lens), jitter, lens);
    ++c->launches;
    const bool accel = (flags & RT_FLAG_ACCEL) != 0;
    const int mode = accel ? rt::MODE_ACCEL : rt::MODE_CULL;
    const int gshift = accel ? p.accel.gshift : 5;
    const int tpb = accel ? c->accel_tpb : c->path_tpb, bps = accel ? c->accel_bps : c->path_bps;
    const size_t smem = accel ? accel_cta_smem_bytes(p.accel, tpb) : cta_smem_bytes(c->n_pad, tpb, false);
    const int grid = c->prop.multiProcessorCount * bps;
    const void* k_first = first_hit_kernel(mode, gshift);
    const void* k_path = path_kernel(mode, false, gshift, max_depth == 0);
    const int tpb1 = accel ? tpb : c->first_tpb, grid1 = accel ? grid : c->prop.multiProcessorCount * c->first_bps;
    const size_t smem1 = accel ? smem : cta_smem_bytes(c->n_pad, tpb1, false);
    const int n_launch_lights = 1 + static_cast<int>(c->extra_lights.size());
    const rt::RenderParams base = p;
    for (int s0 = 0; s0 < spp; s0 += RT_STATIC_BATCH) {
        p = base;
        p.batch_s0 = s0; p.batch_n = std::min<int>(RT_STATIC_BATCH, spp - s0); p.hit_stride = RT_STATIC_BATCH;
        p.n_ladder = 0; p.chunk_spp = RT_STATIC_BATCH; p.n_chunks = 1;
        void* args[] = {&p};
        RT_CUDA(cudaLaunchKernel(k_first, dim3(grid1), dim3(tpb1), args, smem1, c->stream));
        ++c->launches;
        for (int k = 0; k < n_launch_lights; ++k) {
            if (k > 0) select_extra_light(c, k, p);
            p.overwrite = ((flags & RT_FLAG_OVERWRITE) && s0 == 0 && k == 0) ? 1 : 0;
            RT_CUDA(cudaMemsetAsync(c->counters, 0, sizeof(u64), c->stream));
            RT_CUDA(cudaLaunchKernel(k_path, dim3(grid), dim3(tpb), args, smem, c->stream));
            ++c->launches;
        }
    }
    RT_CUDA(cudaGetLastError());
    return RT_OK;
}

int rt_read_hdr_sum(rt_context* c, float* rgba, size_t n_pixels) {
    if (!c || !rgba) return fail(RT_ERR_INVALID_ARGUMENT, "rt_read_hdr_sum: NULL argument");
    if (!c->hdr()) return fail(RT_ERR_NOT_READY, "rt_read_hdr_sum: no framebuffer");
    if (n_pixels != static_cast<size_t>(c->W) * c->H) return fail(RT_ERR_INVALID_ARGUMENT, "rt_read_hdr_sum: n_pixels %zu != %d*%d", n_pixels, c->W, c->H);
    if (int r = ctx_use(c)) return r;
    RT_CUDA(cudaMemcpyAsync(rgba, c->hdr(), n_pixels * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
    RT_CUDA(cudaStreamSynchronize(c->stream));
    return RT_OK;
}
int rt_write_hdr_sum(rt_context* c, const float* rgba, size_t n_pixels) {
    if (!c || !rgba) return fail(RT_ERR_INVALID_ARGUMENT, "rt_write_hdr_sum: NULL argument");
    if (!c->hdr()) return fail(RT_ERR_NOT_READY, "rt_write_hdr_sum: no framebuffer");
    if (n_pixels != static_cast<size_t>(c->W) * c->H) return fail(RT_ERR_INVALID_ARGUMENT, "rt_write_hdr_sum: n_pixels mismatch");
    if (int r = ctx_use(c)) return r;
    RT_CUDA(cudaMemcpyAsync(c->hdr(), rgba, n_pixels * sizeof(float4), cudaMemcpyHostToDevice, c->stream));
    RT_CUDA(cudaStreamSynchronize(c->stream));
    return RT_OK;
}
int rt_resolve_ldr(rt_context* c, uint64_t sample_count, uint32_t* bgra8, size_t n_pixels) {
    if (!c) return fail(RT_ERR_INVALID_ARGUMENT, "rt_resolve_ldr: ctx is NULL");
    if (!c->hdr()) return fail(RT_ERR_NOT_READY, "rt_resolve_ldr: no framebuffer");
    if (sample_count < 1) return fail(RT_ERR_INVALID_ARGUMENT, "rt_resolve_ldr: sample_count must be >= 1");
    const size_t n = static_cast<size_t>(c->W) * c->H;
    if (bgra8 && n_pixels != n) return fail(RT_ERR_INVALID_ARGUMENT, "rt_resolve_ldr: n_pixels mismatch");
    if (int r = ctx_use(c)) return r;
    const int grid = c->prop.multiProcessorCount * 8;
    EventPair ev = ev_begin(c, 1);
    rt::rt_resolve_kernel<<<grid, 256, 0, c->stream>>>(c->hdr(), c->ldr, static_cast<float>(sample_count), n);
    ev_end(c, ev);
    ++c->launches;
    RT_CUDA(cudaGetLastError());
    if (bgra8) {
        RT_CUDA(cudaMemcpyAsync(bgra8, c->ldr, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
        RT_CUDA(cudaStreamSynchronize(c->stream));
    }
    return RT_OK;
}
int rt_primary_hits(rt_context* c, uint64_t sample, uint32_t flags, int32_t* sphere_idx, int32_t* material_id, size_t n_pixels) {
    if (!c) return fail(RT_ERR_INVALID_ARGUMENT, "rt_primary_hits: ctx is NULL");
    if (!c->scene_set || !c->camera_set || !c->hdr()) return fail(RT_ERR_NOT_READY, "rt_primary_hits: scene, camera and framebuffer must be set first");
    const size_t n = static_cast<size_t>(c->W) * c->H;
    if (n_pixels != n || sample < 1) return fail(RT_ERR_INVALID_ARGUMENT, "rt_primary_hits: n_pixels mismatch or sample < 1");
    if (int r = ctx_use(c)) return r;
    rt::RenderParams p;
    if (int r = prepare_frames(c, sample, 1, 0, flags, p)) return r;
    p.chunk_spp = 1; p.n_chunks = 1;
    const size_t smem = cta_smem_bytes(c->n_pad, c->tpb, c->global_scene);
    const int grid = c->prop.multiProcessorCount * std::max(1, c->render_blocks_per_sm);
    if (c->global_scene && (flags & RT_FLAG_DEBUG_NO_CULL_)) return fail(RT_ERR_UNSUPPORTED, "rt_primary_hits: the no-cull debug mode stages the sphere list in shared memory (small scenes only)");
    if ((flags & RT_FLAG_ACCEL) && !c->accel_ok) return fail(RT_ERR_UNSUPPORTED, "rt_primary_hits: RT_FLAG_ACCEL unavailable for this scene: %s", c->accel_why.c_str());
    if (flags & RT_FLAG_ACCEL) {
        int32_t* a_ids = c->ids; int32_t* a_mids = c->mids;
        void* args[] = {&p, &a_ids, &a_mids};
        RT_CUDA(cudaLaunchKernel(accel_primary_kernel(p.accel.gshift), dim3(c->prop.multiProcessorCount * c->accel_bps), dim3(c->accel_tpb), args,
                                 accel_cta_smem_bytes(p.accel, c->accel_tpb), c->stream));
    }
    else if (flags & RT_FLAG_DEBUG_NO_CULL_) rt::rt_primary_kernel<rt::MODE_NO_CULL><<<grid, c->tpb, smem, c->stream>>>(p, c->ids, c->mids);
    else if (c->global_scene) rt::rt_primary_kernel<rt::MODE_CULL_GLOBAL><<<grid, c->tpb, smem, c->stream>>>(p, c->ids, c->mids);
    else rt::rt_primary_kernel<rt::MODE_CULL><<<grid, c->tpb, smem, c->stream>>>(p, c->ids, c->mids);
    ++c->launches;
    RT_CUDA(cudaGetLastError());
    if (sphere_idx) RT_CUDA(cudaMemcpyAsync(sphere_idx, c->ids, n * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    if (material_id) RT_CUDA(cudaMemcpyAsync(material_id, c->mids, n * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    RT_CUDA(cudaStreamSynchronize(c->stream));
    return RT_OK;
}
int rt_scene_accel_image(const rt_scene* s, void* buf, size_t capacity, rt_accel_image_info* info) { return rt_scene_accel_image_slots(s, 0, buf, capacity, info); }
int rt_scene_accel_image_slots(const rt_scene* s, int min_group_slots, void* buf, size_t capacity, rt_accel_image_info* info) {
    if (!s || !info) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_accel_image: NULL argument");
    if (gshift_of_slots(min_group_slots) < 0) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_accel_image: min_group_slots must be 0, 8, 16 or 32");
    rt::AccelParams a; std::vector<unsigned char> img; std::string why;
    if (!make_accel_image(s, a, img, why, gshift_of_slots(min_group_slots))) return fail(RT_ERR_UNSUPPORTED, "rt_scene_accel_image: RT_FLAG_ACCEL unavailable for this scene: %s", why.c_str());
    *info = rt_accel_image_info{};
    info->bytes = img.size();
    info->off_B = a.off_B; info->off_R = a.off_R; info->off_I = a.off_I; info->off_box = a.off_box; info->off_top = a.off_top;
    info->n_pairs = a.n_pairs; info->n_big_groups = a.n_big_groups; info->n_big_pairs = a.n_big_pairs; info->n_groups = a.n_groups;
    info->group_slots = 1 << a.gshift; info->n_l2_nodes = a.n2; info->node_floats = RT_ACCEL_NODE_FLOATS;
    info->bound_center[0] = a.csx; info->bound_center[1] = a.csy; info->bound_center[2] = a.csz; info->bound_radius = a.cs_radius;
    info->r_min = a.r_min; info->r_max = a.r_max; info->c_max = a.c_max;
    if (buf) {
        if (capacity < img.size()) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_accel_image: buffer of %zu bytes, image needs %zu", capacity, img.size());
        std::memcpy(buf, img.data(), img.size());
    }
    return RT_OK;
}
int rt_scene_cull_image(const rt_scene* s, int light_index, void* buf, size_t capacity, rt_cull_image_info* info) {
    if (!s || !info) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_cull_image: NULL argument");
    if (light_index >= static_cast<int>(s->lights.size())) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_cull_image: light %d of %zu", light_index, s->lights.size());
    const int n = static_cast<int>(s->spheres.size());
    const int n_pad = std::max(32, (n + 31) / 32 * 32), np = n_pad / 2;
    double rot[9];
    host::V3 Ln{0.f, 0.f, 0.f};
    if (light_index >= 0) Ln = host::normalize3({s->lights[light_index].dir[0], s->lights[light_index].dir[1], s->lights[light_index].dir[2]});
    light_frame(Ln, light_index >= 0, rot);
    *info = rt_cull_image_info{};
    info->bytes = static_cast<size_t>(np) * 32u + static_cast<size_t>(n_pad) * 16u;
    info->n_spheres = n; info->n_pairs = np; info->off_B = np * 16; info->off_G = np * 32;
    for (int k = 0; k < 9; ++k) info->rot[k] = static_cast<float>(rot[k]);
    info->light_dir[0] = Ln.x; info->light_dir[1] = Ln.y; info->light_dir[2] = Ln.z;
    if (buf) {
        if (capacity < info->bytes) return fail(RT_ERR_INVALID_ARGUMENT, "rt_scene_cull_image: buffer of %zu bytes, image needs %zu", capacity, info->bytes);
        unsigned char* b = static_cast<unsigned char*>(buf);
        fill_pairs(s, np, rot, reinterpret_cast<float4*>(b), reinterpret_cast<float4*>(b + info->off_B));
        fill_geom(s, n_pad, reinterpret_cast<float4*>(b + info->off_G));
    }
    return RT_OK;
}
int rt_get_stats(rt_context* c, rt_stats* out) {
    if (!c || !out) return fail(RT_ERR_INVALID_ARGUMENT, "rt_get_stats: NULL argument");
    if (int r = ctx_use(c)) return r;
    RT_CUDA(cudaStreamSynchronize(c->stream));
    ev_collect(c);
    u64 h[7] = {0, 0, 0, 0, 0, 0, 0};
    RT_CUDA(cudaMemcpy(h, c->counters, 7 * sizeof(u64), cudaMemcpyDeviceToHost));
    memset(out, 0, sizeof *out);
    out->closest_rays = h[1] - c->stat_base[0]; out->shadow_rays = h[2] - c->stat_base[1];
    out->passes = h[3] - c->stat_base[2]; out->paths = h[4] - c->stat_base[3];
    // tests the device evaluated: shadow rays whose light term is identically zero are answered without a query
    out->sphere_tests = h[6] - c->stat_base[5];
    out->render_ms = c->render_ms; out->resolve_ms = c->resolve_ms;
    out->sm_count = c->prop.multiProcessorCount; out->clock_khz = c->clock_khz; out->launches = c->launches;
    return RT_OK;
}
int rt_stats_reset(rt_context* c) {
    if (!c) return fail(RT_ERR_INVALID_ARGUMENT, "rt_stats_reset: ctx is NULL");
    if (int r = ctx_use(c)) return r;
    RT_CUDA(cudaStreamSynchronize(c->stream));
    ev_collect(c);
    u64 h[7];
    RT_CUDA(cudaMemcpy(h, c->counters, 7 * sizeof(u64), cudaMemcpyDeviceToHost));
    for (int k = 0; k < 6; ++k) c->stat_base[k] = h[k + 1];
    c->render_ms = c->resolve_ms = 0; c->launches = 0;
    return RT_OK;
}

}  // extern "C"
