/*
 * rt_b200.h -- C ABI of the B200-native `src/spheres` path-tracing hot path.
 *
 * Drop-in boundary for SakibSaikia/CPURayTracer (citations relative to the reference tree):
 * the reference has no FFI; its boundary is (i) the src/common-lib class surface that
 * SpheresApp::InitScene / InitCamera construct (spheres-app.cpp:35-130) and (ii) the single
 * call SpheresApp::DrawBitmap (spheres-app.cpp:163-222) that consumes them.  This header
 * exports both as plain C: opaque handles, plain pointers and sizes, int status codes.
 * include/rt_b200.hpp layers the reference's C++ class names on top of it.
 *
 * Every function returns RT_OK (0) or an rt_status error; rt_last_error() holds the message.
 * There is NO CPU fallback: if no CUDA device / kernel image is usable the call fails loudly.
 * A context is not re-entrant: drive it from one host thread (as the reference's UI thread does).
 */
#ifndef RT_B200_H
#define RT_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define RT_API __declspec(dllexport)
#else
#define RT_API __attribute__((visibility("default")))
#endif

typedef enum rt_status {
    RT_OK = 0,
    RT_ERR_INVALID_ARGUMENT = 1,
    RT_ERR_CUDA = 2,
    RT_ERR_OUT_OF_MEMORY = 3,
    RT_ERR_NOT_READY = 4,       /* scene / camera / framebuffer not set */
    RT_ERR_UNSUPPORTED = 5
} rt_status;

/* ---- flat scene description, version 1 (what rt_scene_export returns) ---------------------
 * Colours are the floats AFTER the reference's 8-bit XMCOLOR round trip
 * (XMCOLOR(float x4) ctor -> XMLoadColor; texture.cpp:5,16-17, light.cpp:8). */
typedef struct rt_sphere_desc   { float cx, cy, cz, radius; int32_t material; } rt_sphere_desc;            /* ray-tracing.h:53-58 */
typedef struct rt_material_desc { int32_t kind; int32_t texture; float smoothness; float ior; } rt_material_desc; /* material.h */
typedef struct rt_texture_desc  { int32_t kind; float tiling; float c0[4]; float c1[4]; } rt_texture_desc;    /* texture.h */
typedef struct rt_light_desc    { float dir[3]; float color[4]; float luminance; } rt_light_desc;            /* light.h:15-21; dir as given to the ctor (consumers normalise it, light.cpp:7) */

/* material ids (also the values written to the primary-hit material buffer) */
enum { RT_MAT_OPAQUE = 0,        /* DielectricOpaque,       material.h:37-52 */
       RT_MAT_METAL = 1,         /* Metal,                  material.h:21-35 */
       RT_MAT_TRANSPARENT = 2,   /* DielectricTransparent,  material.h:54-68 */
       RT_MAT_SKY = 3 };         /* Emissive sky on a miss, material.h:70-84, spheres-app.cpp:255 */
enum { RT_TEX_CONST = 0, RT_TEX_CHECKER = 1 };   /* texture.h:12-30 */

/* Camera(origin, lookAt, verticalFOV, aspectRatio, focalLength, aperture), camera.h:9 */
typedef struct rt_camera_params {
    float origin[3]; float look_at[3]; float vfov_deg; float aspect; float focus; float aperture;
} rt_camera_params;
/* what the constructor derives (camera.cpp:14-27): m_origin, m_x, m_y, m_originImagePlane */
typedef struct rt_camera_basis { float origin[3], x[3], y[3], image_origin[3]; float aperture, focus; } rt_camera_basis;

/* render flags */
enum { RT_FLAG_JITTER_OFF = 1u,   /* jitterOffset = 0 (spheres-app.cpp:140)              */
       RT_FLAG_LENS_OFF = 2u,     /* lens offset = 0  (spheres-app.cpp:152): pinhole rays */
       RT_FLAG_WAVEFRONT = 4u,    /* schedule the same work as separate intersect / shade kernels over SoA path state in HBM
                                     (same bits as the default megakernel; kept for the measured layout comparison) */
       RT_FLAG_ACCEL = 8u };      /* closest-hit / occlusion queries go through a bounding-box hierarchy (what BvhNode::Intersect,
                                     ray-tracing.cpp:174-214, does in the reference) before the same cull + exact tests; returns the
                                     same bits as the default brute-force query, only faster on large scenes */
/* RT_FLAG_HALTON (16u): draw the scatter randoms from the reference's sampler -- one post-incremented Halton index per material
 * instance (m_reflectionProbabilitySampleIndex / m_sampleIndex, material.h:34,50-51,67; material.cpp:29,44,82,151) -- instead
 * of the stateless counter RNG.  The indices are shared by all paths (device atomics), start at 0 in rt_set_scene and persist
 * across rt_render calls, as the reference's members do; like the reference, the result is reproducible only when paths run
 * one after another (1x1 framebuffer), and statistically equivalent otherwise.  Works with the brute-force query and with
 * RT_FLAG_ACCEL, with one light or several (lights 1.. re-walk the batch's paths with the indices rewound, so every light
 * sees the numbers light 0 saw -- the reference draws once per vertex and loops over its lights there, material.cpp:4-13). */
enum { RT_FLAG_HALTON = 16u };
/* RT_FLAG_OVERWRITE (32u): the call STARTS the running sum instead of adding to it -- the framebuffer's previous content is not
 * read (saves the clear; used by ranks >= 1 of a sharded render, whose buffers hold only the current call's share). */
enum { RT_FLAG_OVERWRITE = 32u };
/* RT_FLAG_DEBUG_ONE_KERNEL (0x200u): the round-1 schedule -- primary rays traced inside the path megakernel instead of by the
 * first-hit stage.  Same arithmetic per path; kept so that the choice of schedule stays a measurement (bench.py, profiles/). */
enum { RT_FLAG_DEBUG_ONE_KERNEL = 0x200u };

typedef struct rt_stats {
    uint64_t paths;            /* path samples traced since the last rt_stats_reset            */
    uint64_t closest_rays;     /* closest-hit queries (GetClosestIntersection, spheres-app.cpp:224) */
    uint64_t shadow_rays;      /* occlusion queries (light.cpp:13-15)                          */
    uint64_t sphere_tests;     /* n_spheres * (closest + shadow queries the device evaluated)   */
    uint64_t passes;           /* warp-level fused passes executed (diagnostic)                */
    double   render_ms;        /* CUDA-event time of the render kernels since the last reset   */
    double   resolve_ms;       /* CUDA-event time of the resolve kernels                       */
    int32_t  sm_count;         /* multiProcessorCount of the device                            */
    int32_t  clock_khz;        /* cudaDevAttrClockRate                                          */
    int32_t  launches;         /* kernels launched since the last reset                        */
    int32_t  reserved;
} rt_stats;

typedef struct rt_context rt_context;   /* one GPU, one stream, one framebuffer */
typedef struct rt_scene   rt_scene;     /* host-side scene recorder            */

/* ---- context ---- */
RT_API int  rt_context_create(int device, rt_context** out);
RT_API void rt_context_destroy(rt_context* ctx);
RT_API const char* rt_last_error(void);                        /* thread-local message of the last failure */
RT_API int  rt_set_stream(rt_context* ctx, void* cuda_stream); /* borrow a cudaStream_t (e.g. torch's current stream) */
RT_API void* rt_get_stream(rt_context* ctx);                   /* the cudaStream_t the context launches on (e.g. to enqueue an ncclReduce behind the render) */
RT_API int  rt_synchronize(rt_context* ctx);

/* ---- scene recorder: mirrors the reference constructors ---- */
RT_API int  rt_scene_create(rt_scene** out);
RT_API void rt_scene_destroy(rt_scene* scene);
/* ConstTexture(XMCOLOR) texture.cpp:3-6 -- rgba in [0,1] is quantised to 8 bit exactly as XMCOLOR does. Returns id >= 0, or -rt_status. */
RT_API int  rt_scene_add_texture_const(rt_scene* s, const float rgba[4]);
/* CheckerTexture(XMCOLOR, XMCOLOR, tiling) texture.cpp:13-18 */
RT_API int  rt_scene_add_texture_checker(rt_scene* s, const float rgba0[4], const float rgba1[4], float tiling);
/* DielectricOpaque(albedo, smoothness) material.cpp:15-18;  Metal(reflectance, smoothness) :67-70;
 * DielectricTransparent(smoothness, ior) :105-109.  Return material id >= 0, or -rt_status. */
RT_API int  rt_scene_add_material_opaque(rt_scene* s, int texture, float smoothness);
RT_API int  rt_scene_add_material_metal(rt_scene* s, int texture, float smoothness);
RT_API int  rt_scene_add_material_transparent(rt_scene* s, float smoothness, float ior);
/* Sphere(center, radius, material) ray-tracing.cpp:21-24. Returns the sphere index (insertion order). */
RT_API int  rt_scene_add_sphere(rt_scene* s, const float center[3], float radius, int material);
/* Emissive(luminance, texture) as m_skyMaterial, spheres-app.cpp:120-121 */
RT_API int  rt_scene_set_sky(rt_scene* s, float luminance, int texture);
/* DirectionalLight(dir, XMCOLOR, luminance, occlusionTest) light.cpp:4-9; the occlusion callback is the scene itself.
 * May be called several times (m_lights, spheres-app.cpp:129): the first light is fused into the query pass, each further
 * light adds one launch over the same paths per rt_render (counter RNG only). */
RT_API int  rt_scene_add_directional_light(rt_scene* s, const float dir[3], const float rgba[4], float luminance);
/* flat access (byte-identical input for an external checker) */
RT_API int  rt_scene_counts(const rt_scene* s, int32_t* n_spheres, int32_t* n_materials, int32_t* n_textures, int32_t* n_lights);
RT_API int  rt_scene_export(const rt_scene* s, rt_sphere_desc* spheres, rt_material_desc* materials, rt_texture_desc* textures,
                            rt_light_desc* lights, float* sky_luminance, int32_t* sky_texture);
/* load already-quantised flat arrays (e.g. a frozen scene file); replaces the recorder's content */
RT_API int  rt_scene_import(rt_scene* s, int32_t n_spheres, const rt_sphere_desc* spheres, int32_t n_materials, const rt_material_desc* materials,
                            int32_t n_textures, const rt_texture_desc* textures, int32_t n_lights, const rt_light_desc* lights,
                            float sky_luminance, int32_t sky_texture);

/* ---- binding to a context (uploads to HBM) ---- */
RT_API int  rt_set_scene(rt_context* ctx, const rt_scene* scene);            /* InitScene: spheres-app.cpp:51-130 */
RT_API int  rt_set_camera(rt_context* ctx, const rt_camera_params* cam);     /* InitCamera: spheres-app.cpp:35-49 */
RT_API int  rt_camera_basis_of(const rt_camera_params* cam, rt_camera_basis* out);   /* Camera::Camera, camera.cpp:3-28 (host only) */
RT_API int  rt_set_exposure(rt_context* ctx, float stops);                   /* m_exposure, default -15: spheres-app.cpp:48 */
RT_API int  rt_set_seed(rt_context* ctx, uint32_t seed);                     /* counter-RNG stream selector, default 0 */
RT_API int  rt_sampler_reset(rt_context* ctx);                                /* RT_FLAG_HALTON: set every material's Halton indices back to 0 */
RT_API int  rt_set_chunk_spp(rt_context* ctx, int chunk_spp);                /* 0 = auto; samples per work item (load balance); rounded up to a multiple of 8 */
/* RT_FLAG_ACCEL: finest group size rt_set_scene may choose for the box hierarchy: 8, 16 or 32 slots per group (0 = default 8).
 * Takes effect at the next rt_set_scene.  A tuning knob; every size returns the same bits. */
RT_API int  rt_set_accel_group_slots(rt_context* ctx, int slots);
/* InitBuffers (app.cpp:112-119): allocate W*H float4 running-sum + BGRA8 buffers, zero-filled. */
RT_API int  rt_framebuffer_resize(rt_context* ctx, int width, int height);
RT_API int  rt_framebuffer_clear(rt_context* ctx);

/* ---- the hot path ----
 * rt_render == `spp` consecutive SpheresApp::DrawBitmap() calls (spheres-app.cpp:163-184) starting at
 * m_sampleCount == first_sample (>= 1), with AppSettings::k_recursionDepth == max_depth.
 * Adds to the running sum; asynchronous on the context's stream. */
RT_API int  rt_render(rt_context* ctx, uint64_t first_sample, int spp, int max_depth, uint32_t flags);
/* rt_render_static -- NO COUNTERPART IN THE REFERENCE; added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same contract and same result as
 * rt_render on its default two-stage schedule (flags: RT_FLAG_JITTER_OFF, RT_FLAG_LENS_OFF, RT_FLAG_ACCEL, RT_FLAG_OVERWRITE only;
 * counter RNG; any number of lights), but the call allocates NOTHING: its scratch memory -- the per-call Halton tables and the
 * first-hit records the path kernel regenerates from -- lives in a file-scope `static __device__` array of RT_STATIC_SCRATCH_BYTES
 * (64 MiB, held by every GPU that loads the library until the process exits) whose device address the library takes once per
 * device with cudaGetSymbolAddress.  Frames are rendered in batches of RT_STATIC_BATCH (16), one work item per pixel and batch
 * (sequential per-pixel sums).  Inputs (scene, camera) and the output (the running-sum framebuffer) are the context's, set up
 * beforehand with the calls above; rt_render_static itself calls no cudaMalloc / cudaFree / cudaEventCreate and records no timing
 * events (rt_stats.render_ms does not include it).
 * LIMIT: rt_render_static_scratch_bytes(W, H, spp) = 32 B per pixel + 16 B per frame + 8 B per column and row (each table rounded
 * up to 256 B) must be <= RT_STATIC_SCRATCH_BYTES -- e.g. 1920x1080 with up to 47,000 frames per call, not 3840x2160 -- else
 * RT_ERR_UNSUPPORTED with the figures in rt_last_error().
 * SHARING: every call on a device uses the same array, which keeps the last call's contents; nothing in it needs clearing (each
 * call rewrites what it reads), but calls whose kernels OVERLAP on one GPU (two contexts / streams) race on it: serialise them. */
#define RT_STATIC_SCRATCH_BYTES (64u << 20)
#define RT_STATIC_BATCH 16
RT_API int    rt_render_static(rt_context* ctx, uint64_t first_sample, int spp, int max_depth, uint32_t flags);
RT_API size_t rt_render_static_scratch_bytes(int width, int height, int spp);   /* host only; 0 for invalid arguments */
/* m_backbufferHdr (app.h:26): float4 per pixel, row-major j*W+i, row 0 = top; xyz = sum of exposure-scaled
 * samples, w = number of samples accumulated. */
RT_API int  rt_read_hdr_sum(rt_context* ctx, float* rgba, size_t n_pixels);
RT_API int  rt_write_hdr_sum(rt_context* ctx, const float* rgba, size_t n_pixels);   /* resume from a checkpoint */
/* Resolve (spheres-app.cpp:186-214): sum / sample_count, ACES fit, gamma 1/2.2, XMStoreColor -> 0xAARRGGBB.
 * bgra8 may be NULL to only run the kernel (result stays in HBM). */
RT_API int  rt_resolve_ldr(rt_context* ctx, uint64_t sample_count, uint32_t* bgra8, size_t n_pixels);
/* Primary-hit id buffers for frame `sample`: sphere index (-1 = miss) and material id (RT_MAT_*). */
RT_API int  rt_primary_hits(rt_context* ctx, uint64_t sample, uint32_t flags, int32_t* sphere_idx, int32_t* material_id, size_t n_pixels);

/* ---- inspection: the hierarchy image RT_FLAG_ACCEL stages into shared memory ----
 * What rt_set_scene builds in place of the reference's BvhNode tree (ray-tracing.cpp:107-167): host-only, needs no device.
 * Layout (all offsets in bytes from the start of the image; "slot" = position of a sphere in the re-ordered list, two slots = one pair):
 *   [0, off_B)        float4 per pair (cx0, cx1, cy0, cy1)         [off_B, off_R)   float4 per pair (cz0, cz1, w0, w1)
 *   [off_R, off_I)    float radius per slot                         [off_I, off_box) uint16 original sphere index per slot (0xFFFF = padding)
 *   [off_box, off_top) per L2 node `node_floats` floats: 8 children x (cx, cy, cz, hx, hy, hz) SoA (centre / half extent; empty: h = -1e30)
 *   [off_top, bytes)  per L2 node 8 floats (cx, cy, cz, hx, hy, hz, -, -)
 * Slots [0, 32 n_big_groups) hold the big spheres, then n_groups small groups of group_slots slots in Morton order; 8 groups = one L2 node. */
typedef struct rt_accel_image_info {
    size_t  bytes;
    int32_t off_B, off_R, off_I, off_box, off_top;
    int32_t n_pairs, n_big_groups, n_big_pairs, n_groups, group_slots, n_l2_nodes, node_floats;
    float   bound_center[3], bound_radius;     /* sphere around all small spheres (pad of the slab test) */
    float   r_min, r_max, c_max;               /* smallest / largest small-sphere radius, largest |coordinate| of a box face */
} rt_accel_image_info;
/* Fills `info`; copies the image into buf when buf != NULL (capacity >= info->bytes).  RT_ERR_UNSUPPORTED when the scene has no hierarchy. */
RT_API int  rt_scene_accel_image(const rt_scene* scene, void* buf, size_t capacity, rt_accel_image_info* info);
/* Same with the finest group size to try (8, 16 or 32 slots; 0 = default), as rt_set_accel_group_slots sets it for a context. */
RT_API int  rt_scene_accel_image_slots(const rt_scene* scene, int min_group_slots, void* buf, size_t capacity, rt_accel_image_info* info);

/* ---- inspection: what the brute-force query stages (host only, needs no device) ----
 * [0, off_B) float4 per pair (cx'0, cx'1, cy'0, cy'1); [off_B, off_G) float4 per pair (cz'0, cz'1, w0, w1); [off_G, bytes) float4 per
 * sphere (cx, cy, cz, r) in world space.  c' = rot c: the centre in the light-aligned frame of light `light_index` (rows of rot =
 * e1, e2, L; -1 = no light: identity), w = r^2 - |c|^2 + 2^-16 (|c|^2 + r^2); padding pairs have w = -1e30.  No counterpart in the
 * reference (its BVH is what these arrays replace, ray-tracing.cpp:107-214); used by the CPU tests to model the conservative cull. */
typedef struct rt_cull_image_info {
    size_t  bytes;
    int32_t n_spheres, n_pairs, off_B, off_G;
    float   rot[9];
    float   light_dir[3];      /* the normalised direction (light.cpp:7); 0 without a light */
} rt_cull_image_info;
RT_API int  rt_scene_cull_image(const rt_scene* scene, int light_index, void* buf, size_t capacity, rt_cull_image_info* info);

/* ---- interop / instrumentation ---- */
RT_API void* rt_hdr_device_ptr(rt_context* ctx);                   /* float4[W*H] in HBM (e.g. for an NCCL reduce) */
RT_API int  rt_hdr_bind_external(rt_context* ctx, void* device_ptr); /* render into a caller-owned device buffer (NULL = own buffer) */
RT_API int  rt_get_stats(rt_context* ctx, rt_stats* out);          /* synchronises the stream */
RT_API int  rt_stats_reset(rt_context* ctx);
RT_API const char* rt_build_info(void);

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_H */
