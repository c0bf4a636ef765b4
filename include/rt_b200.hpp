// rt_b200.hpp -- C++17 host mirror of the SakibSaikia/CPURayTracer `src/common-lib` class surface over the C ABI
// (include/rt_b200.h).  Header-only; link against cpuraytracer_b200/csrc/librt_b200.so.
//
// The classes keep the reference's names and constructor argument order so InitScene / InitCamera
// (spheres-app.cpp:35-130) read the same; they only *record* into a flat scene description.  What cannot cross to
// the device is flattened: virtual Scatter/Shade/Emit/Evaluate become enum kind + parameters, the DirectionalLight
// occlusion callback (light.h:15) is accepted and ignored (occlusion == the scene itself, spheres-app.cpp:124-128),
// XMVECTOR becomes a plain 4-float struct.  All rendering happens in the CUDA library; nothing here computes radiance.
// Errors surface as rtb200::Error carrying the rt_status code (the reference only asserts, app.cpp:45-109).
#pragma once
#include "rt_b200.h"

#include <cmath>
#include <cstdint>
#include <functional>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace rtb200 {

struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};
inline void check(int rc) { if (rc != RT_OK) throw Error(rc, rt_last_error()); }
inline int check_id(int v) { if (v < 0) throw Error(-v, rt_last_error()); return v; }

// ---- the slice of DirectXMath the reference's constructors use -------------------------------------------------
struct XMFLOAT2 { float x, y; };
struct XMFLOAT3 { float x, y, z; };
struct XMVECTOR { float x, y, z, w; };
inline XMVECTOR XMVectorSet(float x, float y, float z, float w) { return XMVECTOR{x, y, z, w}; }
inline XMVECTOR XMVectorReplicate(float v) { return XMVECTOR{v, v, v, v}; }
// DirectX::PackedVector::XMCOLOR(float r, g, b, a): kept as floats here, quantised to 8 bit inside the library exactly
// as the XMCOLOR ctor + XMLoadColor round trip does (texture.cpp:5,16-17; light.cpp:8).
struct XMCOLOR { float r, g, b, a; XMCOLOR(float r_, float g_, float b_, float a_) : r(r_), g(g_), b(b_), a(a_) {} };

struct Ray { XMVECTOR origin, direction; };                     // ray-tracing.h:15-24

// ---- textures (texture.h:6-30) ----------------------------------------------------------------------------------
class Texture {
public:
    virtual ~Texture() = default;
    virtual int Record(rt_scene* s) const = 0;                  // returns the texture id inside `s`
};
class ConstTexture : public Texture {
public:
    explicit ConstTexture(const XMCOLOR& color) : m_color(color) {}
    int Record(rt_scene* s) const override {
        const float c[4] = {m_color.r, m_color.g, m_color.b, m_color.a};
        return check_id(rt_scene_add_texture_const(s, c));
    }
private:
    XMCOLOR m_color;
};
class CheckerTexture : public Texture {
public:
    CheckerTexture(const XMCOLOR& color0, const XMCOLOR& color1, float tiling) : m_c0(color0), m_c1(color1), m_tilingScale(tiling) {}
    int Record(rt_scene* s) const override {
        const float a[4] = {m_c0.r, m_c0.g, m_c0.b, m_c0.a}, b[4] = {m_c1.r, m_c1.g, m_c1.b, m_c1.a};
        return check_id(rt_scene_add_texture_checker(s, a, b, m_tilingScale));
    }
private:
    XMCOLOR m_c0, m_c1;
    float m_tilingScale;
};

// Records every Texture object once per scene (materials share textures by pointer, spheres-app.cpp:57-62).
class TextureTable {
public:
    int IdOf(rt_scene* s, const Texture* t) {
        if (!t) throw Error(RT_ERR_INVALID_ARGUMENT, "null Texture");
        auto it = m_ids.find(t);
        if (it != m_ids.end()) return it->second;
        const int id = t->Record(s);
        m_ids.emplace(t, id);
        return id;
    }
private:
    std::map<const Texture*, int> m_ids;
};

// ---- materials (material.h:9-84).  The reference passes smoothness as a replicated XMVECTOR; both forms are accepted.
class Material {
public:
    virtual ~Material() = default;
    virtual int Record(rt_scene* s, TextureTable& tex) const = 0;   // returns the material id inside `s`
};
class Metal : public Material {
public:
    Metal(const Texture* reflectance, float smoothness) : m_reflectance(reflectance), m_smoothness(smoothness) {}
    Metal(const Texture* reflectance, const XMVECTOR& smoothness) : Metal(reflectance, smoothness.x) {}
    int Record(rt_scene* s, TextureTable& tex) const override { return check_id(rt_scene_add_material_metal(s, tex.IdOf(s, m_reflectance), m_smoothness)); }
private:
    const Texture* m_reflectance;
    float m_smoothness;
};
class DielectricOpaque : public Material {
public:
    DielectricOpaque(const Texture* albedo, float smoothness) : m_albedo(albedo), m_smoothness(smoothness) {}
    DielectricOpaque(const Texture* albedo, const XMVECTOR& smoothness) : DielectricOpaque(albedo, smoothness.x) {}
    int Record(rt_scene* s, TextureTable& tex) const override { return check_id(rt_scene_add_material_opaque(s, tex.IdOf(s, m_albedo), m_smoothness)); }
private:
    const Texture* m_albedo;
    float m_smoothness;
};
class DielectricTransparent : public Material {
public:
    DielectricTransparent(float smoothness, float ior) : m_smoothness(smoothness), m_ior(ior) {}
    DielectricTransparent(const XMVECTOR& smoothness, float ior) : DielectricTransparent(smoothness.x, ior) {}
    int Record(rt_scene* s, TextureTable&) const override { return check_id(rt_scene_add_material_transparent(s, m_smoothness, m_ior)); }
private:
    float m_smoothness, m_ior;
};
class Emissive : public Material {                                 // the sky only (spheres-app.cpp:120-121,255)
public:
    Emissive(float luminance, const Texture* color) : m_color(color), m_luminance(luminance) {}
    int Record(rt_scene*, TextureTable&) const override { throw Error(RT_ERR_UNSUPPORTED, "Emissive is the sky material: pass it to Scene::SetSky, not to a Sphere"); }
    float Luminance() const { return m_luminance; }
    const Texture* Color() const { return m_color; }
private:
    const Texture* m_color;
    float m_luminance;
};

// ---- hitables / lights / camera --------------------------------------------------------------------------------
struct Sphere {                                                    // ray-tracing.h:53-65
    XMVECTOR center;
    float radius;
    std::unique_ptr<Material> material;
    Sphere(const XMVECTOR& c, float r, std::unique_ptr<Material>&& mat) noexcept : center(c), radius(r), material(std::move(mat)) {}
};
class DirectionalLight {                                           // light.h:12-22
public:
    DirectionalLight(const XMVECTOR& dir, const XMCOLOR& color, float luminance, std::function<bool(const Ray&)> lightOcclusionTest = {})
        : m_direction(dir), m_color(color), m_luminance(luminance) { (void)lightOcclusionTest; }
    const XMVECTOR& Direction() const { return m_direction; }
    const XMCOLOR& Color() const { return m_color; }
    float Luminance() const { return m_luminance; }
private:
    XMVECTOR m_direction;
    XMCOLOR m_color;
    float m_luminance;
};
class Camera {                                                     // camera.h:6-20
public:
    Camera(XMVECTOR origin, XMVECTOR lookAt, float verticalFOV, float aspectRatio, float focalLength, float aperture) {
        m_params = rt_camera_params{{origin.x, origin.y, origin.z}, {lookAt.x, lookAt.y, lookAt.z}, verticalFOV, aspectRatio, focalLength, aperture};
        check(rt_camera_basis_of(&m_params, &m_basis));            // Camera::Camera, camera.cpp:3-28
    }
    XMVECTOR GetOrigin() const { return XMVECTOR{m_basis.origin[0], m_basis.origin[1], m_basis.origin[2], 0.f}; }
    const rt_camera_params& Params() const { return m_params; }
    const rt_camera_basis& Basis() const { return m_basis; }       // m_origin, m_x, m_y, m_originImagePlane
private:
    rt_camera_params m_params;
    rt_camera_basis m_basis;
};

// ---- scene recorder: m_scene / m_textures / m_lights / m_skyMaterial of SpheresApp (spheres-app.h:35-41) ----------
class Scene {
public:
    Scene() { check(rt_scene_create(&m_h)); }
    ~Scene() { rt_scene_destroy(m_h); }
    Scene(const Scene&) = delete;
    Scene& operator=(const Scene&) = delete;
    int Add(const Sphere& s) {
        if (!s.material) throw Error(RT_ERR_INVALID_ARGUMENT, "Sphere without a material");
        const float c[3] = {s.center.x, s.center.y, s.center.z};
        return check_id(rt_scene_add_sphere(m_h, c, s.radius, s.material->Record(m_h, m_tex)));
    }
    void SetSky(const Emissive& sky) { check(rt_scene_set_sky(m_h, sky.Luminance(), m_tex.IdOf(m_h, sky.Color()))); }
    void AddLight(const DirectionalLight& l) {
        const float d[3] = {l.Direction().x, l.Direction().y, l.Direction().z};
        const float c[4] = {l.Color().r, l.Color().g, l.Color().b, l.Color().a};
        check(rt_scene_add_directional_light(m_h, d, c, l.Luminance()));
    }
    struct Flat {
        std::vector<rt_sphere_desc> spheres; std::vector<rt_material_desc> materials;
        std::vector<rt_texture_desc> textures; std::vector<rt_light_desc> lights;
        float sky_luminance = 0.f; int32_t sky_texture = -1;
    };
    Flat Export() const {
        Flat f; int32_t ns, nm, nt, nl;
        check(rt_scene_counts(m_h, &ns, &nm, &nt, &nl));
        f.spheres.resize(ns); f.materials.resize(nm); f.textures.resize(nt); f.lights.resize(nl);
        check(rt_scene_export(m_h, f.spheres.data(), f.materials.data(), f.textures.data(), f.lights.data(), &f.sky_luminance, &f.sky_texture));
        return f;
    }
    void Import(const Flat& f) {
        check(rt_scene_import(m_h, static_cast<int32_t>(f.spheres.size()), f.spheres.data(), static_cast<int32_t>(f.materials.size()), f.materials.data(),
                              static_cast<int32_t>(f.textures.size()), f.textures.data(), static_cast<int32_t>(f.lights.size()), f.lights.data(),
                              f.sky_luminance, f.sky_texture));
    }
    rt_scene* Handle() const { return m_h; }
private:
    rt_scene* m_h = nullptr;
    TextureTable m_tex;
};

// ---- one GPU's render loop: the replacement of SpheresApp::DrawBitmap + m_backbufferHdr/Ldr ------------------------
class Renderer {
public:
    Renderer(int device, int width, int height) : m_w(width), m_h(height) {
        check(rt_context_create(device, &m_ctx));
        const int rc = rt_framebuffer_resize(m_ctx, width, height);   // InitBuffers, app.cpp:112-119
        if (rc != RT_OK) { const std::string msg = rt_last_error(); rt_context_destroy(m_ctx); throw Error(rc, msg); }
    }
    ~Renderer() { rt_context_destroy(m_ctx); }
    Renderer(const Renderer&) = delete;
    Renderer& operator=(const Renderer&) = delete;

    void SetScene(const Scene& s) { check(rt_set_scene(m_ctx, s.Handle())); }
    void SetCamera(const Camera& c) { check(rt_set_camera(m_ctx, &c.Params())); }
    void SetExposure(float stops) { check(rt_set_exposure(m_ctx, stops)); }          // m_exposure, spheres-app.cpp:48
    void SetSeed(uint32_t seed) { check(rt_set_seed(m_ctx, seed)); }
    void Clear() { check(rt_framebuffer_clear(m_ctx)); m_sampleCount = 0; }

    // render(spp, max_depth): `spp` x DrawBitmap with k_recursionDepth = maxDepth (spheres-app.cpp:163-184, spheres-app.h:9).
    // Returns the number of primary rays traced, as DrawBitmap does (:221).  Asynchronous on the context's stream.
    size_t Render(int spp, int maxDepth, uint32_t flags = 0) { return RenderFrom(m_sampleCount + 1, spp, maxDepth, flags); }
    size_t RenderFrom(uint64_t firstSample, int spp, int maxDepth, uint32_t flags = 0) {
        check(rt_render(m_ctx, firstSample, spp, maxDepth, flags));
        m_sampleCount += static_cast<uint64_t>(spp);
        return static_cast<size_t>(m_w) * m_h * static_cast<size_t>(spp);
    }
    // The same render without a single allocation (scratch in the library's static device array; rt_b200.h: rt_render_static -- no
    // counterpart in the reference).  Throws rtb200::Error(RT_ERR_UNSUPPORTED) when the job does not fit the array.
    size_t RenderStatic(int spp, int maxDepth, uint32_t flags = 0) {
        check(rt_render_static(m_ctx, m_sampleCount + 1, spp, maxDepth, flags));
        m_sampleCount += static_cast<uint64_t>(spp);
        return static_cast<size_t>(m_w) * m_h * static_cast<size_t>(spp);
    }
    void Synchronize() { check(rt_synchronize(m_ctx)); }

    std::vector<float> ReadHdrSum() {                                // m_backbufferHdr: float4 running sum, w = samples
        std::vector<float> out(static_cast<size_t>(m_w) * m_h * 4);
        check(rt_read_hdr_sum(m_ctx, out.data(), static_cast<size_t>(m_w) * m_h));
        return out;
    }
    void WriteHdrSum(const std::vector<float>& rgba, uint64_t sampleCount) {   // resume from a checkpoint
        if (rgba.size() != static_cast<size_t>(m_w) * m_h * 4) throw Error(RT_ERR_INVALID_ARGUMENT, "WriteHdrSum: size mismatch");
        check(rt_write_hdr_sum(m_ctx, rgba.data(), static_cast<size_t>(m_w) * m_h));
        m_sampleCount = sampleCount;
    }
    std::vector<uint32_t> ResolveLdr() { return ResolveLdr(m_sampleCount); }    // m_backbufferLdr, 0xAARRGGBB (spheres-app.cpp:186-214)
    std::vector<uint32_t> ResolveLdr(uint64_t sampleCount) {
        std::vector<uint32_t> out(static_cast<size_t>(m_w) * m_h);
        check(rt_resolve_ldr(m_ctx, sampleCount, out.data(), out.size()));
        return out;
    }
    void PrimaryHits(uint64_t sample, uint32_t flags, std::vector<int32_t>& sphereIdx, std::vector<int32_t>& materialId) {
        sphereIdx.resize(static_cast<size_t>(m_w) * m_h); materialId.resize(sphereIdx.size());
        check(rt_primary_hits(m_ctx, sample, flags, sphereIdx.data(), materialId.data(), sphereIdx.size()));
    }
    rt_stats Stats() { rt_stats s; check(rt_get_stats(m_ctx, &s)); return s; }
    void ResetStats() { check(rt_stats_reset(m_ctx)); }

    int Width() const { return m_w; }
    int Height() const { return m_h; }
    uint64_t SampleCount() const { return m_sampleCount; }          // m_sampleCount, spheres-app.h:42
    rt_context* Handle() const { return m_ctx; }
private:
    rt_context* m_ctx = nullptr;
    int m_w, m_h;
    uint64_t m_sampleCount = 0;
};

}  // namespace rtb200
