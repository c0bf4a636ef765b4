// spheres_main.cpp -- headless replacement of the reference's `src/spheres` executable (main.cpp:5-10 + SpheresApp,
// spheres-app.{h,cpp}) on top of the B200 library: same InitCamera / InitScene recipe, same OnRender loop
// (DrawBitmap + DisplayStats), but frames are rendered by librt_b200.so on 1..N GPUs and the result sink is a file
// (PPM / PFM / EXR) instead of a Direct2D window.  Host code only: every radiance value comes from the CUDA library.
//
//   spheres [--width W --height H] [--spp N] [--max-depth D] [--aperture A] [--focus F] [--exposure S]
//           [--scene d488|s4096|small | --scene-file F | --ranlux-seed S] [--gpus G] [--seed K] [--jitter-off] [--lens-off] [--accel]
//           [--frames-per-call K] [--out PREFIX] [--format ppm,pfm,exr] [--preview-every N]
//           [--checkpoint F.exr] [--resume F.exr] [--dump-scene F] [--primary-hits F] [--selftest-sinks DIR] [--quiet]
#include "../include/rt_b200.hpp"
#include "../include/rt_b200_group.h"
#include "../include/rt_image_io.hpp"

#include <chrono>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <iostream>
#include <random>
#include <sstream>
#include <unistd.h>

using namespace rtb200;

static FILE* g_json_out = stdout;   // the process's original stdout; fd 1 itself is pointed at stderr (NCCL prints its banner there)

namespace AppSettings {                          // spheres-app.h:5-13 (compile-time there, run-time here)
struct Settings {
    int backbufferWidth = 1280, backbufferHeight = 720, recursionDepth = 50;
    float verticalFov = 25.f, aperture = 0.4f, focus = -1.f /* <0: |origin - lookAt| */, exposure = -15.f;
    int spp = 16, gpus = 1, framesPerCall = 0, previewEvery = 0;
    uint32_t seed = 0, flags = 0;
    long long ranluxSeed = -1;                   // >= 0: replay std::ranlux24_base(seed) as the reference's InitScene does
    std::string scene = "d488", sceneFile, out = "spheres", format = "ppm", checkpoint, resume, dumpScene, primaryHits, selftest;
    bool quiet = false;
};
}

// Portable replacement of `ranlux24_base + uniform_real_distribution<float>` seeded from std::random_device
// (spheres-app.cpp:53-55, unreproducible): SplitMix64, u = top 24 bits * 2^-24.  Identical to cpuraytracer_b200/scenes.py.
struct SplitMix64 {
    uint64_t s;
    explicit SplitMix64(uint64_t seed) : s(seed) {}
    uint64_t next() { s += 0x9E3779B97F4A7C15ull; uint64_t z = s; z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31); }
    float uniform() { return static_cast<float>(next() >> 40) * (1.0f / 16777216.0f); }
};

static void check_group(int rc) { if (rc != RT_OK) throw Error(rc, rt_group_last_error()); }

class SpheresApp {
public:
    explicit SpheresApp(const AppSettings::Settings& s) : m_cfg(s) {}
    ~SpheresApp() { rt_group_destroy(m_group); }

    void OnInitialize() {
        InitCamera();
        if (!m_cfg.sceneFile.empty()) LoadSceneFile(m_cfg.sceneFile);
        else if (m_cfg.ranluxSeed >= 0) {
            // the reference's own generator (spheres-app.cpp:53-55) for a known std::random_device value: std::ranlux24_base's
            // output is fixed by the C++ standard; MSVC's uniform_real_distribution<float>(0,1) maps one draw x to x * 2^-24
            std::ranlux24_base generator(static_cast<uint32_t>(m_cfg.ranluxSeed));
            InitScene(-11, 11, [&generator]() { return static_cast<float>(generator()) * (1.0f / 16777216.0f); }, -1);
        }
        else if (m_cfg.scene == "d488") InitScene(-11, 11, 1, -1);          // spheres-app.cpp:64-66
        else if (m_cfg.scene == "s4096") InitScene(-32, 32, 2, 4092);       // N = 4096
        else if (m_cfg.scene == "small") InitScene(-3, 3, 3, -1);
        else throw Error(RT_ERR_INVALID_ARGUMENT, "unknown --scene " + m_cfg.scene);
    }

    void InitCamera() {                                                     // spheres-app.cpp:35-49
        const XMVECTOR camOrigin = XMVectorSet(12.f, 2.f, -2.5f, 1.f);
        const XMVECTOR camLookAt = XMVectorSet(0, 1, 0, 1.f);
        const float dx = camOrigin.x - camLookAt.x, dy = camOrigin.y - camLookAt.y, dz = camOrigin.z - camLookAt.z;
        const float length = std::sqrt((dx * dx + dy * dy) + dz * dz);       // XMVector3Length
        const float aspect = static_cast<float>(static_cast<double>(m_cfg.backbufferWidth) / m_cfg.backbufferHeight);
        m_camera = std::make_unique<Camera>(camOrigin, camLookAt, m_cfg.verticalFov, aspect, m_cfg.focus > 0.f ? m_cfg.focus : length, m_cfg.aperture);
        m_exposure = m_cfg.exposure;
    }

    void InitScene(int gridLo, int gridHi, uint64_t seed, int maxSmall) {
        SplitMix64 generator(seed);
        InitScene(gridLo, gridHi, [&generator]() { return generator.uniform(); }, maxSmall);
    }
    void InitScene(int gridLo, int gridHi, const std::function<float()>& uniformDist, int maxSmall) {   // spheres-app.cpp:51-130
        // Floor
        m_textures.push_back(std::make_unique<CheckerTexture>(XMCOLOR{0.9f, 0.9f, 0.9f, 1.f}, XMCOLOR{0.2f, 0.3f, 0.1f, 1.f}, 2500.f));
        m_scene.push_back(std::make_unique<Sphere>(XMVectorSet(0, -1000, 0, 0), 1000.f, std::make_unique<DielectricOpaque>(m_textures.back().get(), XMVectorReplicate(16.f))));
        // Random small spheres
        int count = 0;
        for (int a = gridLo; a < gridHi; ++a) {
            for (int b = gridLo; b < gridHi; ++b) {
                if (maxSmall >= 0 && count >= maxSmall) break;
                const float chooseMat = uniformDist();
                const float cx = static_cast<float>(a) + 0.9f * uniformDist();
                const float cz = static_cast<float>(b) + 0.9f * uniformDist();
                const XMVECTOR center = XMVectorSet(cx, 0.2f, cz, 0.f);
                if (chooseMat < 0.8f) {
                    const float r0 = uniformDist(), r1 = uniformDist(), g0 = uniformDist(), g1 = uniformDist(), b0 = uniformDist(), b1 = uniformDist();
                    m_textures.push_back(std::make_unique<ConstTexture>(XMCOLOR{r0 * r1, g0 * g1, b0 * b1, 1.f}));
                    const float smoothness = 8.f * (4.f + uniformDist());
                    m_scene.push_back(std::make_unique<Sphere>(center, 0.2f, std::make_unique<DielectricOpaque>(m_textures.back().get(), XMVectorReplicate(smoothness))));
                } else if (chooseMat < 0.95f) {
                    const float r = 0.5f * (1.f + uniformDist()), g = 0.5f * (1.f + uniformDist()), bl = 0.5f * (1.f + uniformDist());
                    m_textures.push_back(std::make_unique<ConstTexture>(XMCOLOR{r, g, bl, 1.f}));
                    m_scene.push_back(std::make_unique<Sphere>(center, 0.2f, std::make_unique<Metal>(m_textures.back().get(), XMVectorReplicate(0.f))));
                } else {
                    const float smoothness = 8.f * (4.f + uniformDist());
                    m_scene.push_back(std::make_unique<Sphere>(center, 0.2f, std::make_unique<DielectricTransparent>(XMVectorReplicate(smoothness), 1.5f)));
                }
                ++count;
            }
        }
        // Large spheres
        m_scene.push_back(std::make_unique<Sphere>(XMVectorSet(0, 1, 0, 0), 1.f, std::make_unique<DielectricTransparent>(XMVectorReplicate(16.f), 1.5f)));
        m_textures.push_back(std::make_unique<ConstTexture>(XMCOLOR{0.4f, 0.2f, 0.1f, 1.f}));
        m_scene.push_back(std::make_unique<Sphere>(XMVectorSet(-4, 1, 0, 0), 1.f, std::make_unique<DielectricOpaque>(m_textures.back().get(), XMVectorReplicate(16.f))));
        m_textures.push_back(std::make_unique<ConstTexture>(XMCOLOR{0.7f, 0.6f, 0.5f, 1.f}));
        m_scene.push_back(std::make_unique<Sphere>(XMVectorSet(4, 1, 0, 0), 1.f, std::make_unique<Metal>(m_textures.back().get(), XMVectorReplicate(0.f))));
        // (the reference builds its BVH here, spheres-app.cpp:117; the device tests every sphere)
        for (const auto& s : m_scene) m_recorded.Add(*s);
        // Sky
        m_textures.push_back(std::make_unique<ConstTexture>(XMCOLOR{0.85f, 0.91f, 0.98f, 1.f}));
        m_skyMaterial = std::make_unique<Emissive>(8000.f, m_textures.back().get());
        m_recorded.SetSky(*m_skyMaterial);
        // Sun: the occlusion callback of the reference (spheres-app.cpp:124-128) is the scene itself on the device
        m_lights.push_back(std::make_unique<DirectionalLight>(XMVectorSet(1.f, 1.f, 1.f, 0.f), XMCOLOR{1.f, 0.97f, 0.88f, 1.f}, 40000.f));
        m_recorded.AddLight(*m_lights.back());
    }

    // ---- flat scene file "RTSC" v1: the bytes of rt_scene_export ----
    void DumpSceneFile(const std::string& path) const {
        const Scene::Flat f = m_recorded.Export();
        rtio::detail::Buf o;
        o.raw("RTSC", 4); o.i32(1);
        o.i32(static_cast<int32_t>(f.spheres.size())); o.i32(static_cast<int32_t>(f.materials.size()));
        o.i32(static_cast<int32_t>(f.textures.size())); o.i32(static_cast<int32_t>(f.lights.size()));
        o.f32(f.sky_luminance); o.i32(f.sky_texture);
        o.raw(f.spheres.data(), f.spheres.size() * sizeof(rt_sphere_desc));
        o.raw(f.materials.data(), f.materials.size() * sizeof(rt_material_desc));
        o.raw(f.textures.data(), f.textures.size() * sizeof(rt_texture_desc));
        o.raw(f.lights.data(), f.lights.size() * sizeof(rt_light_desc));
        rtio::write_all(path, o.b);
    }
    void LoadSceneFile(const std::string& path) {
        const std::vector<unsigned char> b = rtio::read_all(path);
        if (b.size() < 32 || std::memcmp(b.data(), "RTSC", 4) != 0) throw Error(RT_ERR_INVALID_ARGUMENT, path + ": not an RTSC scene file");
        int32_t h[5]; std::memcpy(h, &b[4], 20);
        if (h[0] != 1 || h[1] < 0 || h[2] < 0 || h[3] < 0 || h[4] < 0) throw Error(RT_ERR_INVALID_ARGUMENT, path + ": unsupported RTSC version/counts");
        Scene::Flat f;
        std::memcpy(&f.sky_luminance, &b[24], 4); std::memcpy(&f.sky_texture, &b[28], 4);
        size_t p = 32;
        auto take = [&](auto& vec, int32_t n) {
            using T = typename std::remove_reference<decltype(vec)>::type::value_type;
            if (p + static_cast<size_t>(n) * sizeof(T) > b.size()) throw Error(RT_ERR_INVALID_ARGUMENT, path + ": truncated RTSC file");
            vec.resize(n); std::memcpy(vec.data(), &b[p], static_cast<size_t>(n) * sizeof(T)); p += static_cast<size_t>(n) * sizeof(T);
        };
        take(f.spheres, h[1]); take(f.materials, h[2]); take(f.textures, h[3]); take(f.lights, h[4]);
        m_recorded.Import(f);
    }

    void CreateDevices() {
        check_group(rt_group_create(nullptr, m_cfg.gpus, m_cfg.backbufferWidth, m_cfg.backbufferHeight, &m_group));
        check_group(rt_group_set_scene(m_group, m_recorded.Handle()));
        check_group(rt_group_set_camera(m_group, &m_camera->Params()));
        check_group(rt_group_set_exposure(m_group, m_exposure));
        check_group(rt_group_set_seed(m_group, m_cfg.seed));
    }
    rt_context* Root() const { return rt_group_member(m_group, 0); }

    // SpheresApp::DrawBitmap x frames (spheres-app.cpp:163-222): returns the number of primary rays traced
    size_t DrawBitmap(int frames) {
        check_group(rt_group_render(m_group, m_sampleCount + 1, frames, m_cfg.recursionDepth, m_cfg.flags));
        check_group(rt_group_synchronize(m_group));
        m_sampleCount += static_cast<size_t>(frames);
        return static_cast<size_t>(m_cfg.backbufferWidth) * m_cfg.backbufferHeight * static_cast<size_t>(frames);
    }
    void DisplayStats(size_t rayCount, double timeElapsedUs) {             // spheres-app.cpp:259-272 (window title -> stderr)
        m_totalTimeInSeconds += timeElapsedUs * 1e-6;
        if (m_cfg.quiet) return;
        const double mraysPerSecond = static_cast<double>(rayCount) / timeElapsedUs;
        std::fprintf(stderr, "Demo\t | Mrays/s: %f\t | spp: %zu\t | Time (seconds): %f\n", mraysPerSecond, m_sampleCount, m_totalTimeInSeconds);
    }

    std::vector<float> ReadHdrSum() {
        std::vector<float> hdr(static_cast<size_t>(m_cfg.backbufferWidth) * m_cfg.backbufferHeight * 4);
        check(rt_read_hdr_sum(Root(), hdr.data(), hdr.size() / 4));
        return hdr;
    }
    void Resume(const std::string& path) {
        std::vector<float> hdr; int w, h; uint64_t n;
        rtio::read_exr_sum(path, hdr, w, h, n);
        if (w != m_cfg.backbufferWidth || h != m_cfg.backbufferHeight) throw Error(RT_ERR_INVALID_ARGUMENT, path + ": checkpoint resolution differs from --width/--height");
        check_group(rt_group_clear(m_group));
        check(rt_write_hdr_sum(Root(), hdr.data(), hdr.size() / 4));
        m_sampleCount = n;
    }
    void WriteOutputs(const std::string& prefix) {
        const int w = m_cfg.backbufferWidth, h = m_cfg.backbufferHeight;
        std::stringstream ss(m_cfg.format);
        std::string fmt;
        std::vector<float> hdr;
        while (std::getline(ss, fmt, ',')) {
            if (fmt == "ppm") {                                            // resolve: spheres-app.cpp:186-214
                std::vector<uint32_t> ldr(static_cast<size_t>(w) * h);
                check(rt_resolve_ldr(Root(), m_sampleCount, ldr.data(), ldr.size()));
                rtio::write_ppm(prefix + ".ppm", ldr.data(), w, h);
            } else if (fmt == "pfm" || fmt == "exr") {
                if (hdr.empty()) hdr = ReadHdrSum();
                if (fmt == "pfm") rtio::write_pfm(prefix + ".pfm", hdr.data(), w, h, m_sampleCount);
                else rtio::write_exr_sum(prefix + ".exr", hdr.data(), w, h, m_sampleCount);
            } else if (!fmt.empty()) throw Error(RT_ERR_INVALID_ARGUMENT, "unknown --format " + fmt);
        }
    }
    void WritePrimaryHits(const std::string& path) {                       // int32 sphere index plane then material id plane
        const size_t n = static_cast<size_t>(m_cfg.backbufferWidth) * m_cfg.backbufferHeight;
        std::vector<int32_t> ids(n), mids(n);
        check(rt_primary_hits(Root(), 1, m_cfg.flags, ids.data(), mids.data(), n));
        rtio::detail::Buf o; o.raw(ids.data(), n * 4); o.raw(mids.data(), n * 4);
        rtio::write_all(path, o.b);
    }

    // RayTracingApp::Run (app.cpp:56-76) without the message pump: render until the requested spp is reached
    int Run() {
        if (!m_cfg.resume.empty()) Resume(m_cfg.resume);
        if (!m_cfg.primaryHits.empty()) WritePrimaryHits(m_cfg.primaryHits);
        const size_t target = m_sampleCount + static_cast<size_t>(m_cfg.spp);
        int per_call = m_cfg.framesPerCall > 0 ? m_cfg.framesPerCall : std::max(1, m_cfg.spp);
        if (m_cfg.previewEvery > 0) per_call = std::min(per_call, m_cfg.previewEvery);
        size_t next_preview = m_cfg.previewEvery > 0 ? m_sampleCount + m_cfg.previewEvery : ~size_t(0);
        size_t paths = 0;
        while (m_sampleCount < target) {
            const int frames = static_cast<int>(std::min<size_t>(per_call, target - m_sampleCount));
            const auto start = std::chrono::high_resolution_clock::now();
            const size_t rayCount = DrawBitmap(frames);                    // OnRender, spheres-app.cpp:10-33
            const auto stop = std::chrono::high_resolution_clock::now();
            DisplayStats(rayCount, std::chrono::duration<double, std::micro>(stop - start).count());
            paths += rayCount;
            if (m_sampleCount >= next_preview && m_sampleCount < target) { WriteOutputs(m_cfg.out + ".preview"); next_preview += m_cfg.previewEvery; }
        }
        WriteOutputs(m_cfg.out);
        if (!m_cfg.checkpoint.empty()) { const std::vector<float> hdr = ReadHdrSum(); rtio::write_exr_sum(m_cfg.checkpoint, hdr.data(), m_cfg.backbufferWidth, m_cfg.backbufferHeight, m_sampleCount); }
        rt_stats st{};
        uint64_t closest = 0, shadow = 0;
        for (int i = 0; i < rt_group_size(m_group); ++i) { check(rt_get_stats(rt_group_member(m_group, i), &st)); closest += st.closest_rays; shadow += st.shadow_rays; }
        std::fprintf(g_json_out, "{\"scene\": \"%s\", \"width\": %d, \"height\": %d, \"spp\": %zu, \"max_depth\": %d, \"gpus\": %d, \"paths\": %zu, \"closest_rays\": %llu, "
                    "\"shadow_rays\": %llu, \"seconds\": %.6f, \"path_samples_per_s\": %.6e, \"rays_per_s\": %.6e}\n",
                    m_cfg.sceneFile.empty() ? m_cfg.scene.c_str() : m_cfg.sceneFile.c_str(), m_cfg.backbufferWidth, m_cfg.backbufferHeight, m_sampleCount, m_cfg.recursionDepth,
                    m_cfg.gpus, paths, static_cast<unsigned long long>(closest), static_cast<unsigned long long>(shadow), m_totalTimeInSeconds,
                    m_totalTimeInSeconds > 0 ? paths / m_totalTimeInSeconds : 0.0, m_totalTimeInSeconds > 0 ? (closest + shadow) / m_totalTimeInSeconds : 0.0);
        return 0;
    }

private:
    AppSettings::Settings m_cfg;
    std::unique_ptr<Camera> m_camera;
    std::vector<std::unique_ptr<Sphere>> m_scene;
    std::vector<std::unique_ptr<Texture>> m_textures;
    std::vector<std::unique_ptr<DirectionalLight>> m_lights;
    std::unique_ptr<Emissive> m_skyMaterial;
    Scene m_recorded;                  // what m_bvh is to the reference: the structure the renderer consumes
    rt_group* m_group = nullptr;
    float m_exposure = -15.f;
    size_t m_sampleCount = 0;
    double m_totalTimeInSeconds = 0;
};

// Writes a synthetic gradient through every sink and reads the EXR back (no GPU needed): exercised by tests/test_cpp_host.py
static int selftest_sinks(const std::string& dir) {
    const int w = 7, h = 5; const uint64_t n = 3;
    std::vector<float> hdr(static_cast<size_t>(w) * h * 4);
    std::vector<uint32_t> ldr(static_cast<size_t>(w) * h);
    for (int y = 0; y < h; ++y) for (int x = 0; x < w; ++x) {
        float* p = &hdr[(static_cast<size_t>(y) * w + x) * 4];
        p[0] = 0.25f * x; p[1] = 0.5f * y; p[2] = 0.125f * (x + y); p[3] = static_cast<float>(n);
        ldr[static_cast<size_t>(y) * w + x] = 0xFF000000u | (static_cast<uint32_t>(x * 30) << 16) | (static_cast<uint32_t>(y * 50) << 8) | static_cast<uint32_t>(x + y);
    }
    rtio::write_ppm(dir + "/t.ppm", ldr.data(), w, h);
    rtio::write_pfm(dir + "/t.pfm", hdr.data(), w, h, n);
    rtio::write_exr_sum(dir + "/t.exr", hdr.data(), w, h, n);
    std::vector<float> back; int bw, bh; uint64_t bn;
    rtio::read_exr_sum(dir + "/t.exr", back, bw, bh, bn);
    if (bw != w || bh != h || bn != n || back != hdr) { std::fprintf(stderr, "EXR round trip mismatch\n"); return 1; }
    std::printf("sinks ok\n");
    return 0;
}

int main(int argc, char** argv) {
    AppSettings::Settings s;
    try {
        for (int i = 1; i < argc; ++i) {
            const std::string a = argv[i];
            auto val = [&]() -> std::string { if (i + 1 >= argc) throw Error(RT_ERR_INVALID_ARGUMENT, a + " needs a value"); return argv[++i]; };
            if (a == "--width") s.backbufferWidth = std::stoi(val());
            else if (a == "--height") s.backbufferHeight = std::stoi(val());
            else if (a == "--spp") s.spp = std::stoi(val());
            else if (a == "--max-depth") s.recursionDepth = std::stoi(val());
            else if (a == "--aperture") s.aperture = std::stof(val());
            else if (a == "--focus") s.focus = std::stof(val());
            else if (a == "--vfov") s.verticalFov = std::stof(val());
            else if (a == "--exposure") s.exposure = std::stof(val());
            else if (a == "--scene") s.scene = val();
            else if (a == "--scene-file") s.sceneFile = val();
            else if (a == "--ranlux-seed") s.ranluxSeed = std::stoll(val());
            else if (a == "--gpus") s.gpus = std::stoi(val());
            else if (a == "--seed") s.seed = static_cast<uint32_t>(std::stoul(val()));
            else if (a == "--jitter-off") s.flags |= RT_FLAG_JITTER_OFF;
            else if (a == "--lens-off") s.flags |= RT_FLAG_LENS_OFF;
            else if (a == "--accel") s.flags |= RT_FLAG_ACCEL;
            else if (a == "--frames-per-call") s.framesPerCall = std::stoi(val());
            else if (a == "--preview-every") s.previewEvery = std::stoi(val());
            else if (a == "--out") s.out = val();
            else if (a == "--format") s.format = val();
            else if (a == "--checkpoint") s.checkpoint = val();
            else if (a == "--resume") s.resume = val();
            else if (a == "--dump-scene") s.dumpScene = val();
            else if (a == "--primary-hits") s.primaryHits = val();
            else if (a == "--selftest-sinks") s.selftest = val();
            else if (a == "--quiet") s.quiet = true;
            else if (a == "--help" || a == "-h") { std::printf("see the header comment of apps/spheres_main.cpp\n"); return 0; }
            else throw Error(RT_ERR_INVALID_ARGUMENT, "unknown option " + a);
        }
        if (!s.selftest.empty()) return selftest_sinks(s.selftest);
        if (s.backbufferWidth <= 0 || s.backbufferHeight <= 0 || s.spp < 0 || s.recursionDepth < 0 || s.gpus < 1) throw Error(RT_ERR_INVALID_ARGUMENT, "bad size / spp / depth / gpus");
        std::fflush(stdout);
        g_json_out = fdopen(dup(1), "w");
        dup2(2, 1);
        SpheresApp app(s);
        app.OnInitialize();
        if (!s.dumpScene.empty()) { app.DumpSceneFile(s.dumpScene); if (s.spp == 0) return 0; }
        app.CreateDevices();               // fails loudly without a CUDA device: there is no CPU fallback
        const int rc = app.Run();
        std::fflush(g_json_out);
        return rc;
    } catch (const std::exception& e) {
        std::fprintf(stderr, "spheres: error: %s\n", e.what());
        return 2;
    }
}
