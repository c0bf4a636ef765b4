/*
 * rt_b200_group.h -- single-process multi-GPU front end of the `src/spheres` render loop (librt_b200_group.so).
 *
 * The path shards by SAMPLE RANGE: frame jitter and lens samples are pure functions of the frame
 * index (spheres-app.cpp:140,152) and the counter RNG is keyed on (pixel, frame, bounce), so GPU g of G renders frames
 * [first + g*spp/G, ...) of EVERY pixel and the union is exactly the one-GPU sample set.  The only exchange is one
 * ncclReduce(sum, float, root = member 0) of the W*H float4 running-sum buffers per rt_group_render call, enqueued on
 * each member's render stream behind its kernel (no host synchronisation in between).
 *
 * One rt_context (include/rt_b200.h) per device; member 0 owns the job's running sum (m_backbufferHdr, app.h:26).
 * The torch.distributed path (cpuraytracer_b200/distributed.py, one process per GPU) is the multi-process equivalent.
 * Not re-entrant: drive a group from one host thread.
 */
#ifndef RT_B200_GROUP_H
#define RT_B200_GROUP_H
#include "rt_b200.h"
#ifdef __cplusplus
extern "C" {
#endif

enum { RT_ERR_NCCL = 6 };

typedef struct rt_group rt_group;

/* devices == NULL: use devices 0..n-1.  Creates the contexts, W*H framebuffers and one NCCL communicator per device. */
RT_API int  rt_group_create(const int* devices, int n, int width, int height, rt_group** out);
RT_API void rt_group_destroy(rt_group* g);
RT_API const char* rt_group_last_error(void);                           /* message of the last failing rt_group_* call (thread-local) */
RT_API int  rt_group_size(const rt_group* g);
RT_API rt_context* rt_group_member(rt_group* g, int i);                 /* member 0 holds the reduced result */
RT_API int  rt_group_set_scene(rt_group* g, const rt_scene* scene);      /* InitScene on every member */
RT_API int  rt_group_set_camera(rt_group* g, const rt_camera_params* cam);
RT_API int  rt_group_set_exposure(rt_group* g, float stops);
RT_API int  rt_group_set_seed(rt_group* g, uint32_t seed);
RT_API int  rt_group_clear(rt_group* g);
/* Which frames member `rank` of `world` renders out of [first_sample, first_sample + spp): contiguous, disjoint,
 * sizes differ by at most one (identical to cpuraytracer_b200.distributed.shard_samples). */
RT_API int  rt_group_shard(uint64_t first_sample, int spp, int rank, int world, uint64_t* my_first, int* my_spp);
/* `spp` x DrawBitmap split over the members, then the sum-reduce onto member 0.  Asynchronous; rt_group_synchronize
 * (or any rt_read_* on member 0) waits for it. */
RT_API int  rt_group_render(rt_group* g, uint64_t first_sample, int spp, int max_depth, uint32_t flags);
RT_API int  rt_group_synchronize(rt_group* g);
/* milliseconds of the last rt_group_render: max over members of (render kernel + reduce) by CUDA events */
RT_API int  rt_group_last_ms(rt_group* g, double* render_ms_max, double* reduce_ms_max);

#ifdef __cplusplus
}
#endif
#endif
