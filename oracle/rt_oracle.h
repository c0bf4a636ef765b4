/*
 * oracle/rt_oracle.h -- CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 *
 * A scalar C++ restatement of the SakibSaikia/CPURayTracer `src/spheres` render path, written
 * from the reference's behaviour (file:line citations are relative to /root/reference/).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
 * load this library.  The product (cpuraytracer_b200/, include/rt_b200.h) never links, imports
 * or calls anything in this directory.
 *
 * PARITY PIN STATUS: the reference ships no tests or golden vectors, and its arithmetic lives in
 * DirectXMath (Windows SDK 10.0.16299.0, not vendored).  This restatement is pinned three ways:
 *   (1) tests/test_oracle_kat.py: known-answer tests derived by hand from the cited reference lines;
 *   (2) tests/test_oracle_vs_ref.py: oracle/_ref (oracle/Makefile) compiles the reference's OWN camera / light / material /
 *       quasi-random / ray-tracing / texture .cpp files, unmodified and in place, against a from-scratch DirectXMath
 *       stand-in (dxm_shim/); in ORC_MATH_REFERENCE mode this restatement agrees with that build BIT FOR BIT, function by
 *       function and for whole depth-50 renders of the default scene;
 *   (3) tests/golden/*.npz: outputs of that reference build, committed with the script that made them.
 * What remains unpinned is the DirectXMath stand-in itself (restated from memory of the SSE2 code paths):
 * "parity unpinned at the DirectXMath boundary".
 */
#ifndef RT_ORACLE_H
#define RT_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---- flat scene description, version 1 (identical byte layout to include/rt_b200.h) ---- */
typedef struct orc_sphere   { float cx, cy, cz, radius; int32_t material; } orc_sphere;          /* 20 B */
typedef struct orc_material { int32_t kind; int32_t texture; float smoothness; float ior; } orc_material; /* 16 B */
typedef struct orc_texture  { int32_t kind; float tiling; float c0[4]; float c1[4]; } orc_texture; /* 40 B */
typedef struct orc_light    { float dir[3]; float color[4]; float luminance; } orc_light;        /* 32 B; dir as passed to the DirectionalLight ctor (normalised by the consumer, light.cpp:7) */

enum { ORC_MAT_OPAQUE = 0, ORC_MAT_METAL = 1, ORC_MAT_TRANSPARENT = 2, ORC_MAT_SKY = 3 };
enum { ORC_TEX_CONST = 0, ORC_TEX_CHECKER = 1 };

typedef struct orc_scene {
    int32_t n_spheres;   const orc_sphere*   spheres;
    int32_t n_materials; const orc_material* materials;
    int32_t n_textures;  const orc_texture*  textures;
    int32_t n_lights;    const orc_light*    lights;
    float   sky_luminance; int32_t sky_texture;
} orc_scene;

/* Camera as the reference constructs it (camera.h:9): the oracle derives the basis itself. */
typedef struct orc_camera_params {
    float origin[3]; float look_at[3]; float vfov_deg; float aspect; float focus; float aperture;
} orc_camera_params;
typedef struct orc_camera { float origin[3], x[3], y[3], image_origin[3]; float aperture, focus; } orc_camera;

enum { ORC_FLAG_JITTER_OFF = 1u, ORC_FLAG_LENS_OFF = 2u,
       ORC_FLAG_REF_DRAWBITMAP = 4u };  /* tier A (ref_driver.cpp) only: the reference's frame structure -- serial GenerateRays, parallel trace, parallel resolve */
enum { ORC_RNG_COUNTER = 0, ORC_RNG_HALTON = 1 };     /* HALTON = the reference's per-material atomic counters */
enum { ORC_ACCEL_BRUTE = 0, ORC_ACCEL_BVH = 1 };      /* BVH = the reference's BvhNode restated */
enum { ORC_INTEGRATOR_ITERATIVE = 0, ORC_INTEGRATOR_RECURSIVE = 1 };
/* PINNED: pow5 by multiplication, pinned sincos, exact 1/ior (CPU == GPU bit for bit).  REFERENCE: powf, libm sinf/cosf and
 * rcpps as the reference binary calls them (comparable bit for bit with oracle/_ref; x86 only). */
enum { ORC_MATH_PINNED = 0, ORC_MATH_REFERENCE = 1 };

typedef struct orc_config {
    int32_t  width, height;
    uint64_t first_sample;      /* m_sampleCount value of the first frame (>= 1), spheres-app.cpp:168 */
    int32_t  spp;               /* number of consecutive DrawBitmap() frames */
    int32_t  max_depth;         /* AppSettings::k_recursionDepth, spheres-app.h:9 */
    uint32_t flags;
    int32_t  rng_mode, accel, integrator;
    int32_t  threads;           /* 0 = hardware_concurrency */
    float    exposure_stops;    /* m_exposure, spheres-app.cpp:48 (-15) */
    uint32_t seed;              /* counter RNG only */
    int32_t  math_mode;         /* ORC_MATH_PINNED (default) or ORC_MATH_REFERENCE */
} orc_config;

typedef struct orc_stats { uint64_t paths, closest_rays, shadow_rays; double seconds; } orc_stats;

/* ---- whole-path entry points ---- */
/* hdr: W*H float4 running sum (in/out, spheres-app.cpp:177-184); w lane accumulates the sample count. */
int  orc_render(const orc_scene*, const orc_camera_params*, const orc_config*, float* hdr_rgba, orc_stats* stats);
int  orc_primary_hits(const orc_scene*, const orc_camera_params*, const orc_config*, int32_t* sphere_idx, int32_t* material_id);
void orc_resolve_ldr(const float* hdr_rgba, int64_t n_pixels, uint64_t sample_count, uint32_t* bgra8);

/* ---- unit-level entry points for known-answer tests ---- */
void     orc_set_math_mode(int mode);                                                  /* for the unit-level entry points below */
float    orc_halton(uint64_t index, uint32_t base);                                   /* quasi-random.cpp:3-16 */
void     orc_halton_disk(uint64_t index, uint32_t b1, uint32_t b2, float out[2]);      /* quasi-random.cpp:52-61 */
void     orc_halton_hemisphere(uint64_t index, uint32_t b1, uint32_t b2, float out[3]);/* quasi-random.cpp:36-50 */
void     orc_hemisphere_from_u(float u1, float u2, float out[3]);
void     orc_sincos(float theta, float* s, float* c);                                 /* pinned sin/cos */
void     orc_camera_make(const orc_camera_params*, orc_camera*);                      /* camera.cpp:3-28 */
void     orc_camera_get_ray(const orc_camera*, float u, float v, float lx, float ly, float o[3], float d[3]); /* camera.cpp:30-48 */
int      orc_sphere_intersect(const orc_sphere*, const float o[3], const float d[3], float* t, float pos[3], float n[3], float uv[2]); /* ray-tracing.cpp:42-84 */
int      orc_closest_hit(const orc_scene*, const float o[3], const float d[3], int accel, float* t);  /* returns sphere index or -1 */
void     orc_texture_eval(const orc_texture*, float u, float v, float rgba[4]);        /* texture.cpp:8-33 */
float    orc_schlick(float f0, float n_dot_v);                                         /* material.cpp:25-27 */
float    orc_fresnel_term(float cos_i, float ior);                                     /* XMFresnelTerm */
void     orc_reflect(const float i[3], const float n[3], float out[3]);               /* XMVector3Reflect */
void     orc_refract(const float i[3], const float n[3], float eta, float out[3]);    /* XMVector3RefractV */
uint8_t  orc_xmcolor_quantize(float x);                                                /* XMCOLOR(float x4) ctor */
float    orc_xmcolor_load(uint8_t k);                                                  /* XMLoadColor */
void     orc_rng_counter(uint32_t pixel, uint64_t sample, uint32_t bounce, uint32_t seed, float out[4]);
/* scatter one vertex with explicit random numbers: returns 1 if scattered. */
int      orc_scatter(const orc_scene*, int sphere, const float o[3], const float d[3], float xi, float u1, float u2,
                     float atten[3], float out_o[3], float out_d[3]);
/* direct light at the closest hit of ray (o,d) as Material::Shade returns it (material.cpp:4-13). */
int      orc_shade_hit(const orc_scene*, const float view_origin[3], const float o[3], const float d[3], int accel, float rgb[3]);
const char* orc_build_info(void);

#ifdef __cplusplus
}
#endif
#endif
