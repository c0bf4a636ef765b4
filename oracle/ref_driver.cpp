// oracle/ref_driver.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE ("tier A" of the oracle).
//
// Headless driver around the reference's OWN src/common-lib translation units (camera, light, material, quasi-random,
// ray-tracing, texture), which oracle/Makefile compiles unmodified and in place from /root/reference against the
// DirectXMath stand-in in dxm_shim/.  Only the parts of src/spheres/spheres-app.cpp that are welded to HWND / Direct2D
// are restated here (scene construction from a flat description, GenerateRays :132-161, GetHitColor :238-257, the
// accumulate and resolve lambdas of DrawBitmap :177-214); every Intersect / Scatter / Shade / Emit / Evaluate /
// GetRay / Halton* call below executes reference code.  The result pins rt_oracle.cpp: tests/test_oracle_vs_ref.py
// requires the two to agree bit for bit when the oracle runs in ORC_MATH_REFERENCE mode.
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <map>
#include <mutex>
#include <thread>

#include "camera.h"
#include "light.h"
#include "material.h"
#include "quasi-random.h"
#include "ray-tracing.h"
#include "texture.h"

#include "rt_oracle.h"

namespace {

// Ray counters are INSTRUMENTATION the reference does not have (spheres-app.cpp:224-236, :124-128).  They are plain
// thread-local increments -- no shared cache line is touched per ray -- and every worker adds its totals to the world once
// per job (per frame), so the timed region carries no contention the reference would not have.
thread_local uint64_t tl_closest = 0, tl_shadow = 0;
#ifdef REF_SHARED_ATOMIC_COUNTERS
// Round-1 instrumentation (two adjacent std::atomic counters bumped by every ray of every thread), kept ONLY so that the
// A/B in profiles/ (r02_cpu_baseline_ab.txt) can be re-measured: `make -C oracle ref_ab` builds _ref/libref_r01counters.so.
std::atomic<uint64_t> g_closest_shared{0}, g_shadow_shared{0};
#define REF_COUNT_CLOSEST() (g_closest_shared.fetch_add(1, std::memory_order_relaxed), ++tl_closest)
#define REF_COUNT_SHADOW()  (g_shadow_shared.fetch_add(1, std::memory_order_relaxed), ++tl_shadow)
#else
#define REF_COUNT_CLOSEST() (++tl_closest)
#define REF_COUNT_SHADOW()  (++tl_shadow)
#endif

// Persistent worker pool: the reference's std::for_each(std::execution::par, ...) (spheres-app.cpp:177-184) runs on the
// runtime's pool, which outlives the frame; so does this one.  run(n, fn) executes fn(k) for k in [0, n), work handed out
// by one atomic per item (items are rows or ray-buffer blocks: thousands of rays each), and returns when all are done.
class Pool {
public:
    explicit Pool(int threads) : n_threads_(threads < 1 ? 1 : threads) {
        for (int t = 1; t < n_threads_; ++t) workers_.emplace_back([this] { loop(); });
    }
    ~Pool() {
        { std::lock_guard<std::mutex> g(m_); stop_ = true; ++generation_; }
        cv_.notify_all();
        for (auto& th : workers_) th.join();
    }
    int threads() const { return n_threads_; }
    void run(int n, const std::function<void(int)>& fn, std::atomic<uint64_t>& closest, std::atomic<uint64_t>& shadow) {
        {
            std::lock_guard<std::mutex> g(m_);
            fn_ = &fn; n_ = n; next_.store(0); pending_ = static_cast<int>(workers_.size()); closest_ = &closest; shadow_ = &shadow;
            ++generation_;
        }
        cv_.notify_all();
        tl_closest = tl_shadow = 0;                                         // the caller may have used the unit-level entry points before
        work();                                                             // the calling thread is worker 0
        std::unique_lock<std::mutex> g(m_);
        done_.wait(g, [this] { return pending_ == 0; });
    }
private:
    void work() {
        for (int k; (k = next_.fetch_add(1, std::memory_order_relaxed)) < n_;) (*fn_)(k);
        closest_->fetch_add(tl_closest, std::memory_order_relaxed); shadow_->fetch_add(tl_shadow, std::memory_order_relaxed);
        tl_closest = tl_shadow = 0;
    }
    void loop() {
        uint64_t seen = 0;
        for (;;) {
            { std::unique_lock<std::mutex> g(m_); cv_.wait(g, [&] { return generation_ != seen; }); seen = generation_; if (stop_) return; }
            work();
            { std::lock_guard<std::mutex> g(m_); if (--pending_ == 0) done_.notify_one(); }
        }
    }
    int n_threads_;
    std::vector<std::thread> workers_;
    std::mutex m_; std::condition_variable cv_, done_;
    uint64_t generation_ = 0; bool stop_ = false; int pending_ = 0;
    const std::function<void(int)>* fn_ = nullptr; int n_ = 0; std::atomic<int> next_{0};
    std::atomic<uint64_t>* closest_ = nullptr; std::atomic<uint64_t>* shadow_ = nullptr;
};

XMVECTOR v3(const float* p) { return XMVectorSet(p[0], p[1], p[2], 0.f); }
void st3(float* p, FXMVECTOR v) { XMFLOAT3 f; XMStoreFloat3(&f, v); p[0] = f.x; p[1] = f.y; p[2] = f.z; }
// The flat scene stores post-quantisation floats k/255; rebuild the XMCOLOR the reference constructors take.
XMCOLOR color_of(const float* c) { return XMCOLOR(c[0], c[1], c[2], c[3]); }

struct RefWorld {
    std::vector<std::unique_ptr<Texture>> textures;
    std::vector<std::unique_ptr<Hitable>> scene;
    std::vector<std::unique_ptr<Light>> lights;
    std::unique_ptr<BvhNode> bvh;
    std::unique_ptr<Material> sky;
    std::unique_ptr<Camera> camera;
    std::map<const Material*, int> sphere_of;      // Payload carries the material pointer only
    std::vector<int> kind_of;                      // material kind per sphere index
    std::vector<Sphere*> spheres;                  // non-owning, insertion order (the BVH owns them afterwards)
    std::atomic<uint64_t> closest{0}, shadow{0};   // totals; written once per worker and job (see tl_closest above)
    int max_depth = 50;
    std::unique_ptr<Pool> pool;
    std::vector<std::pair<Ray, int>> ray_buffer;   // ORC_FLAG_REF_DRAWBITMAP: GenerateRays' buffer, spheres-app.cpp:132-161

    // InitScene (spheres-app.cpp:51-130) from a flat description instead of the unseeded generator.
    void build(const orc_scene* s, bool with_bvh) {
        std::srand(1);                                                      // BvhNode picks axes with std::rand() (ray-tracing.cpp:123)
        for (int i = 0; i < s->n_textures; ++i) {
            const orc_texture& t = s->textures[i];
            if (t.kind == ORC_TEX_CHECKER) textures.push_back(std::make_unique<CheckerTexture>(color_of(t.c0), color_of(t.c1), t.tiling));
            else textures.push_back(std::make_unique<ConstTexture>(color_of(t.c0)));
        }
        scene.reserve(s->n_spheres);
        for (int i = 0; i < s->n_spheres; ++i) {
            const orc_sphere& sp = s->spheres[i];
            const orc_material& m = s->materials[sp.material];
            std::unique_ptr<Material> mat;
            const XMVECTOR smooth = XMVectorReplicate(m.smoothness);
            if (m.kind == ORC_MAT_OPAQUE) mat = std::make_unique<DielectricOpaque>(textures[m.texture].get(), smooth);
            else if (m.kind == ORC_MAT_METAL) mat = std::make_unique<Metal>(textures[m.texture].get(), smooth);
            else mat = std::make_unique<DielectricTransparent>(smooth, m.ior);
            sphere_of[mat.get()] = i;
            kind_of.push_back(m.kind);
            auto sphere = std::make_unique<Sphere>(XMVectorSet(sp.cx, sp.cy, sp.cz, 1.f), sp.radius, std::move(mat));
            spheres.push_back(sphere.get());
            scene.push_back(std::move(sphere));
        }
        if (with_bvh && !scene.empty()) bvh = std::make_unique<BvhNode>(scene.begin(), scene.end());   // :117
        if (s->sky_texture >= 0) sky = std::make_unique<Emissive>(s->sky_luminance, textures[s->sky_texture].get());   // :120-121
        auto occlusion = [this](const Ray& ray) -> bool {                   // :124-128
            Payload dummy{};
            REF_COUNT_SHADOW();
            return bvh && bvh->Intersect(ray, dummy);
        };
        for (int i = 0; i < s->n_lights; ++i) {
            const orc_light& l = s->lights[i];
            lights.push_back(std::make_unique<DirectionalLight>(v3(l.dir), color_of(l.color), l.luminance, occlusion));   // :129
        }
    }
    void set_camera(const orc_camera_params* p) {                           // InitCamera, :35-49
        camera = std::make_unique<Camera>(XMVectorSet(p->origin[0], p->origin[1], p->origin[2], 1.f),
                                          XMVectorSet(p->look_at[0], p->look_at[1], p->look_at[2], 1.f),
                                          p->vfov_deg, p->aspect, p->focus, p->aperture);
    }
    bool closest_hit(const Ray& ray, Payload& hit) {                        // GetClosestIntersection, :224-236
        REF_COUNT_CLOSEST();
        return bvh && bvh->Intersect(ray, hit);
    }
    // GetHitColor, :238-257: emit + direct + (recurse ? attenuation * indirect : 0); sky on a miss.
    XMVECTOR hit_color(const Ray& ray, int depth) {
        Payload hit{};
        if (!closest_hit(ray, hit)) return sky->Emit(Payload{});
        XMVECTOR attenuation;
        Ray scattered;
        const bool is_scattered = hit.material->Scatter(ray, hit, attenuation, scattered);
        const bool recurse = depth < max_depth && is_scattered;
        const XMVECTOR emitted = hit.material->Emit(hit);
        const XMVECTOR direct = hit.material->Shade(hit, lights, camera->GetOrigin());
        const XMVECTOR indirect = recurse ? attenuation * hit_color(scattered, depth + 1) : XM_Zero;
        return emitted + direct + indirect;
    }
    // One element of GenerateRays' double loop, :144-158.
    Ray primary_ray(const orc_config* cfg, uint64_t n, int i, int j, XMFLOAT2 jitter) const {
        const auto xsize = static_cast<float>(cfg->width), ysize = static_cast<float>(cfg->height);
        XMFLOAT2 uv;
        uv.x = static_cast<float>(i + jitter.x) / xsize;
        uv.y = static_cast<float>(j + jitter.y) / ysize;
        XMFLOAT2 lens(0.f, 0.f);
        if (!(cfg->flags & ORC_FLAG_LENS_OFF)) lens = Random::HaltonSampleDisk(n + i + j, 4, 5);
        return camera->GetRay(uv, lens);
    }
};

int thread_count(int requested) {
    const int t = requested > 0 ? requested : static_cast<int>(std::thread::hardware_concurrency());
    return t < 1 ? 1 : t;
}

}  // namespace

struct ref_world { RefWorld w; };

extern "C" {

void ref_resolve_ldr(const float* hdr, int64_t n_pixels, uint64_t sample_count, uint32_t* bgra8);

// A persistent world = SpheresApp after InitScene / InitCamera: scene, BVH, per-material Halton counters (material.h:34,
// 50-51,67) and the worker pool live across frames, as they do in the reference between DrawBitmap calls.
ref_world* ref_world_create(const orc_scene* sc, const orc_camera_params* cp, int threads) {
    if (!sc || !cp || sc->sky_texture < 0) return nullptr;
    ref_world* h = new ref_world();
    h->w.build(sc, true);
    h->w.set_camera(cp);
    h->w.pool = std::make_unique<Pool>(thread_count(threads));
    return h;
}
void ref_world_destroy(ref_world* h) { delete h; }
int ref_world_threads(const ref_world* h) { return h && h->w.pool ? h->w.pool->threads() : 0; }

// `spp` x DrawBitmap (spheres-app.cpp:163-184) on the reference's classes, continuing the world's sampler state.  rng /
// accel / integrator fields of the config are ignored: this build IS the reference configuration (atomic Halton counters,
// BVH, recursion).  Default schedule: rows handed to the pool, primary ray generated by the worker that traces it (the
// kinder schedule for the CPU).  ORC_FLAG_REF_DRAWBITMAP: the reference's own frame structure -- GenerateRays fills the ray
// buffer SERIALLY (:132-161), the trace runs in parallel over the buffer (:177-184), then the parallel resolve pass
// (:186-214) -- i.e. what SpheresApp::OnRender times (:20-26) minus the Direct2D blit.
int ref_world_render(ref_world* h, const orc_config* cfg, float* hdr, orc_stats* stats) {
    if (!h || !cfg || !hdr || cfg->width <= 0 || cfg->height <= 0 || cfg->spp < 0 || cfg->first_sample < 1) return 1;
    RefWorld& w = h->w;
    w.max_depth = cfg->max_depth;
    w.closest.store(0); w.shadow.store(0);
    const float exposure = std::pow(2, cfg->exposure_stops);                // :174
    const int W = cfg->width, H = cfg->height;
    std::vector<uint32_t> ldr;
    const auto t0 = std::chrono::steady_clock::now();
    for (int s = 0; s < cfg->spp; ++s) {
        const uint64_t n = cfg->first_sample + static_cast<uint64_t>(s);    // m_sampleCount after :168
        XMFLOAT2 jitter(0.f, 0.f);
        if (!(cfg->flags & ORC_FLAG_JITTER_OFF)) jitter = Random::HaltonSample2D(n, 2, 3);   // :140
        auto accumulate = [&](const Ray& ray, size_t id) {
            const XMVECTOR c = w.hit_color(ray, 0) * exposure;              // :183
            float* px = hdr + 4 * id;
            XMFLOAT3 f; XMStoreFloat3(&f, c);
            px[0] += f.x; px[1] += f.y; px[2] += f.z; px[3] += 1.0f;
        };
        if (cfg->flags & ORC_FLAG_REF_DRAWBITMAP) {
            w.ray_buffer.clear();
            w.ray_buffer.reserve(static_cast<size_t>(W) * H);               // :135
            int ray_id = 0;
            for (int j = 0; j < H; ++j)
                for (int i = 0; i < W; ++i) w.ray_buffer.push_back(std::make_pair(w.primary_ray(cfg, n, i, j, jitter), ray_id++));   // :144-158
            const int block = 1024, n_blocks = (static_cast<int>(w.ray_buffer.size()) + block - 1) / block;
            w.pool->run(n_blocks, [&](int k) {
                const size_t e = std::min(w.ray_buffer.size(), static_cast<size_t>(k + 1) * block);
                for (size_t r = static_cast<size_t>(k) * block; r < e; ++r) accumulate(w.ray_buffer[r].first, static_cast<size_t>(w.ray_buffer[r].second));
            }, w.closest, w.shadow);
            ldr.resize(static_cast<size_t>(W) * H);
            w.pool->run(H, [&](int j) { ref_resolve_ldr(hdr + 4 * static_cast<size_t>(j) * W, W, n, ldr.data() + static_cast<size_t>(j) * W); }, w.closest, w.shadow);
        } else {
            w.pool->run(H, [&](int j) {
                for (int i = 0; i < W; ++i) accumulate(w.primary_ray(cfg, n, i, j, jitter), static_cast<size_t>(j) * W + i);
            }, w.closest, w.shadow);
        }
    }
    if (stats) {
        stats->paths = static_cast<uint64_t>(W) * H * cfg->spp;
        stats->closest_rays = w.closest.load(); stats->shadow_rays = w.shadow.load();
        stats->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    }
    return 0;
}

// One-shot form: a fresh world (sampler indices at 0) for this call only.
int ref_render(const orc_scene* sc, const orc_camera_params* cp, const orc_config* cfg, float* hdr, orc_stats* stats) {
    if (!sc || !cp || !cfg || !hdr || cfg->width <= 0 || cfg->height <= 0 || cfg->spp < 0 || cfg->first_sample < 1 || sc->sky_texture < 0) return 1;
    ref_world* h = ref_world_create(sc, cp, cfg->threads);
    if (!h) return 1;
    const int rc = ref_world_render(h, cfg, hdr, stats);
    ref_world_destroy(h);
    return rc;
}

int ref_primary_hits(const orc_scene* sc, const orc_camera_params* cp, const orc_config* cfg, int32_t* sphere_idx, int32_t* material_id) {
    if (!sc || !cp || !cfg || cfg->width <= 0 || cfg->height <= 0 || cfg->first_sample < 1) return 1;
    RefWorld w;
    w.build(sc, true);
    w.set_camera(cp);
    const uint64_t n = cfg->first_sample;
    XMFLOAT2 jitter(0.f, 0.f);
    if (!(cfg->flags & ORC_FLAG_JITTER_OFF)) jitter = Random::HaltonSample2D(n, 2, 3);
    Pool pool(thread_count(cfg->threads));
    pool.run(cfg->height, [&](int j) {
        for (int i = 0; i < cfg->width; ++i) {
            Payload hit{};
            const bool ok = w.closest_hit(w.primary_ray(cfg, n, i, j, jitter), hit);
            const int idx = ok ? w.sphere_of.at(hit.material) : -1;
            const size_t id = static_cast<size_t>(j) * cfg->width + i;
            if (sphere_idx) sphere_idx[id] = idx;
            if (material_id) material_id[id] = ok ? w.kind_of[idx] : ORC_MAT_SKY;
        }
    }, w.closest, w.shadow);
    return 0;
}

// Resolve lambda of DrawBitmap, spheres-app.cpp:186-214.
void ref_resolve_ldr(const float* hdr, int64_t n_pixels, uint64_t sample_count, uint32_t* bgra8) {
    static const float a = 2.51f, b = 0.03f, c = 2.43f, d = 0.59f, e = 0.14f;
    static XMVECTORF32 invGamma{1 / 2.2f, 1 / 2.2f, 1 / 2.2f};
    for (int64_t p = 0; p < n_pixels; ++p) {
        const XMVECTOR hdrColor = XMVectorSet(hdr[4 * p], hdr[4 * p + 1], hdr[4 * p + 2], hdr[4 * p + 3]);
        XMVECTOR color = hdrColor / static_cast<float>(sample_count);
        color = XMVectorSaturate((color * (a * color + XMVectorReplicate(b))) / (color * (c * color + XMVectorReplicate(d)) + XMVectorReplicate(e)));
        color = XMVectorPow(color, invGamma);
        XMCOLOR out;
        XMStoreColor(&out, color);
        bgra8[p] = out.c;
    }
}

float ref_halton(uint64_t index, uint32_t base) { return Random::HaltonSample(index, base); }
void ref_halton_disk(uint64_t index, uint32_t b1, uint32_t b2, float out[2]) { const XMFLOAT2 r = Random::HaltonSampleDisk(index, b1, b2); out[0] = r.x; out[1] = r.y; }
void ref_halton_hemisphere(uint64_t index, uint32_t b1, uint32_t b2, float out[3]) { const XMFLOAT3 r = Random::HaltonSampleHemisphere(index, b1, b2); out[0] = r.x; out[1] = r.y; out[2] = r.z; }
void ref_camera_get_ray(const orc_camera_params* cp, float u, float v, float lx, float ly, float o[3], float d[3]) {
    RefWorld w; w.set_camera(cp);
    const Ray r = w.camera->GetRay(XMFLOAT2(u, v), XMFLOAT2(lx, ly));
    st3(o, r.origin); st3(d, r.direction);
}
int ref_sphere_intersect(const orc_sphere* s, const float o[3], const float d[3], float* t, float pos[3], float n[3], float uv[2]) {
    Sphere sphere(XMVectorSet(s->cx, s->cy, s->cz, 1.f), s->radius, nullptr);
    Payload hit{};
    if (!sphere.Intersect(Ray(v3(o), v3(d)), hit)) return 0;
    if (t) *t = XMVectorGetX(hit.t);
    if (pos) st3(pos, hit.pos);
    if (n) st3(n, hit.normal);
    if (uv) { uv[0] = hit.uv.x; uv[1] = hit.uv.y; }
    return 1;
}
int ref_closest_hit(const orc_scene* sc, const float o[3], const float d[3], float* t) {
    RefWorld w; w.build(sc, true);
    Payload hit{};
    if (!w.closest_hit(Ray(v3(o), v3(d)), hit)) return -1;
    if (t) *t = XMVectorGetX(hit.t);
    return w.sphere_of.at(hit.material);
}
void ref_texture_eval(const orc_texture* t, float u, float v, float rgba[4]) {
    std::unique_ptr<Texture> tex;
    if (t->kind == ORC_TEX_CHECKER) tex = std::make_unique<CheckerTexture>(color_of(t->c0), color_of(t->c1), t->tiling);
    else tex = std::make_unique<ConstTexture>(color_of(t->c0));
    const XMVECTOR c = tex->Evaluate(XMFLOAT2(u, v));
    rgba[0] = XMVectorGetX(c); rgba[1] = XMVectorGetY(c); rgba[2] = XMVectorGetZ(c); rgba[3] = XMVectorGetW(c);
}
float ref_fresnel_term(float cos_i, float ior) { return XMVectorGetX(XMFresnelTerm(XMVectorReplicate(cos_i), XMVectorReplicate(ior))); }
void ref_reflect(const float i[3], const float n[3], float out[3]) { st3(out, XMVector3Reflect(v3(i), v3(n))); }
void ref_refract(const float i[3], const float n[3], float eta, float out[3]) { st3(out, XMVector3RefractV(v3(i), v3(n), XMVectorReplicate(eta))); }
uint8_t ref_xmcolor_quantize(float x) { return static_cast<uint8_t>((XMCOLOR(x, 0.f, 0.f, 0.f).c >> 16) & 0xFF); }
float ref_xmcolor_load(uint8_t k) { const XMCOLOR c(static_cast<uint32_t>(k) << 16); return XMVectorGetX(XMLoadColor(&c)); }

// `count` consecutive Scatter calls on sphere `sphere`'s material with the rays (o[k], d[k]); the material's atomic
// Halton counters advance exactly as in the reference.  Outputs per call: scattered flag, attenuation, out ray.
int ref_scatter_sequence(const orc_scene* sc, int sphere, int count, const float* o, const float* d, int32_t* scattered, float* atten, float* out_o, float* out_d) {
    RefWorld w; w.build(sc, false);
    Sphere* sp = w.spheres.at(sphere);
    for (int k = 0; k < count; ++k) {
        const Ray ray(v3(o + 3 * k), v3(d + 3 * k));
        Payload hit{};
        scattered[k] = -1;
        if (!sp->Intersect(ray, hit)) continue;
        XMVECTOR a = XM_Zero;
        Ray out(XM_Zero, XM_Zero);
        scattered[k] = hit.material->Scatter(ray, hit, a, out) ? 1 : 0;
        st3(atten + 3 * k, a); st3(out_o + 3 * k, out.origin); st3(out_d + 3 * k, out.direction);
    }
    return 0;
}
// Material::Shade at the closest hit of (o, d), as GetHitColor calls it (spheres-app.cpp:250).
int ref_shade_hit(const orc_scene* sc, const float view_origin[3], const float o[3], const float d[3], float rgb[3]) {
    RefWorld w; w.build(sc, true);
    Payload hit{};
    if (!w.closest_hit(Ray(v3(o), v3(d)), hit)) return -1;
    st3(rgb, hit.material->Shade(hit, w.lights, XMVectorSet(view_origin[0], view_origin[1], view_origin[2], 1.f)));
    return w.sphere_of.at(hit.material);
}
const char* ref_build_info(void) {
    return "tier-A: reference camera/light/material/quasi-random/ray-tracing/texture .cpp compiled unmodified against oracle/dxm_shim; g++ " __VERSION__
           " -O2 (x86-64 SSE2 baseline, -ffp-contract=off); ray counters: "
#ifdef REF_SHARED_ATOMIC_COUNTERS
           "shared atomics (round-1 instrumentation, A/B build)";
#else
           "per-thread";
#endif
}

}  // extern "C"
