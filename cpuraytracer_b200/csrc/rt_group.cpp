// rt_group.cpp -- include/rt_b200_group.h: single-process multi-GPU sample-range sharding + ncclReduce (host code only;
// the kernels live in librt_b200.so).  Build: g++ -std=c++17 -shared -fPIC rt_group.cpp -lrt_b200 -lcudart_static -ldl.
// NCCL is bound with dlopen("libnccl.so.2") when the first multi-GPU group is created, not at link time: a process that
// also hosts PyTorch must end up with ONE libnccl.so.2 (torch bundles a newer one than /usr/lib's), and a one-GPU group
// needs none.
#include "../../include/rt_b200_group.h"

#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>   // types and enums only; the entry points are resolved at run time (below)

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <string>
#include <vector>

static thread_local std::string g_group_error;
static int gfail(int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    g_group_error = buf;
    return code;
}
// a failure inside librt_b200 keeps its own message; copy it so rt_group_last_error() is the one place to look
static int forward(int rc) { if (rc != RT_OK) g_group_error = rt_last_error(); return rc; }
#define G_CUDA(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) return gfail(RT_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e_)); } while (0)
#define G_NCCL(expr) do { ncclResult_t r_ = (expr); if (r_ != ncclSuccess) return gfail(RT_ERR_NCCL, "%s failed: %s", #expr, nccl_rt::GetErrorString(r_)); } while (0)

namespace nccl_rt {
static void* handle = nullptr;
static decltype(&::ncclCommInitAll) CommInitAll = nullptr;
static decltype(&::ncclCommDestroy) CommDestroy = nullptr;
static decltype(&::ncclGroupStart) GroupStart = nullptr;
static decltype(&::ncclGroupEnd) GroupEnd = nullptr;
static decltype(&::ncclReduce) Reduce = nullptr;
static decltype(&::ncclGetErrorString) GetErrorString = nullptr;
static int load() {
    if (handle) return RT_OK;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
    if (!h) return gfail(RT_ERR_NCCL, "multi-GPU groups need NCCL: dlopen(libnccl.so.2) failed: %s", dlerror());
#define RT_SYM(name) name = reinterpret_cast<decltype(name)>(dlsym(h, "nccl" #name)); if (!name) { dlclose(h); return gfail(RT_ERR_NCCL, "libnccl lacks nccl" #name); }
    RT_SYM(CommInitAll) RT_SYM(CommDestroy) RT_SYM(GroupStart) RT_SYM(GroupEnd) RT_SYM(Reduce) RT_SYM(GetErrorString)
#undef RT_SYM
    handle = h;
    return RT_OK;
}
}  // namespace nccl_rt

struct Member {
    int device = 0;
    rt_context* ctx = nullptr;
    ncclComm_t comm = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr, e2 = nullptr;
};
struct rt_group {
    std::vector<Member> m;
    int W = 0, H = 0;
    bool timed = false;
};

extern "C" {

RT_API const char* rt_group_last_error(void) { return g_group_error.c_str(); }

int rt_group_shard(uint64_t first_sample, int spp, int rank, int world, uint64_t* my_first, int* my_spp) {
    if (spp < 0 || world < 1 || rank < 0 || rank >= world || first_sample < 1 || !my_first || !my_spp)
        return gfail(RT_ERR_INVALID_ARGUMENT, "rt_group_shard: bad argument");
    const int q = spp / world, r = spp % world;
    *my_spp = q + (rank < r ? 1 : 0);
    *my_first = first_sample + static_cast<uint64_t>(rank) * q + static_cast<uint64_t>(std::min(rank, r));
    return RT_OK;
}

void rt_group_destroy(rt_group* g) {
    if (!g) return;
    for (Member& mb : g->m) {
        cudaSetDevice(mb.device);
        if (mb.ctx) rt_synchronize(mb.ctx);
        if (mb.comm) nccl_rt::CommDestroy(mb.comm);
        if (mb.e0) cudaEventDestroy(mb.e0);
        if (mb.e1) cudaEventDestroy(mb.e1);
        if (mb.e2) cudaEventDestroy(mb.e2);
        if (mb.ctx) rt_context_destroy(mb.ctx);
    }
    delete g;
}

int rt_group_create(const int* devices, int n, int width, int height, rt_group** out) {
    if (!out || n < 1) return gfail(RT_ERR_INVALID_ARGUMENT, "rt_group_create: bad argument");
    *out = nullptr;
    rt_group* g = new rt_group();
    g->W = width; g->H = height;
    g->m.resize(n);
    std::vector<int> devs(n);
    for (int i = 0; i < n; ++i) devs[i] = devices ? devices[i] : i;
    for (int i = 0; i < n; ++i) {
        Member& mb = g->m[i];
        mb.device = devs[i];
        int rc = rt_context_create(mb.device, &mb.ctx);
        if (rc == RT_OK) rc = rt_framebuffer_resize(mb.ctx, width, height);
        if (rc != RT_OK) { forward(rc); rt_group_destroy(g); return rc; }
        cudaSetDevice(mb.device);
        if (cudaEventCreate(&mb.e0) != cudaSuccess || cudaEventCreate(&mb.e1) != cudaSuccess || cudaEventCreate(&mb.e2) != cudaSuccess) {
            rt_group_destroy(g); return gfail(RT_ERR_CUDA, "rt_group_create: cudaEventCreate failed");
        }
    }
    if (n > 1) {
        if (int rc = nccl_rt::load()) { rt_group_destroy(g); return rc; }
        std::vector<ncclComm_t> comms(n);
        const ncclResult_t r = nccl_rt::CommInitAll(comms.data(), n, devs.data());
        if (r != ncclSuccess) { rt_group_destroy(g); return gfail(RT_ERR_NCCL, "ncclCommInitAll(%d devices) failed: %s", n, nccl_rt::GetErrorString(r)); }
        for (int i = 0; i < n; ++i) g->m[i].comm = comms[i];
        // NCCL connects its transports at the first collective of a given size class (~0.3 s for 8 ranks in one process).  Pay
        // that here, with the very reduce rt_group_render issues: the framebuffers were just cleared, so the sum is zero as well.
        const size_t count = static_cast<size_t>(width) * height * 4;
        void* root_buf = rt_hdr_device_ptr(g->m[0].ctx);
        ncclResult_t wr = nccl_rt::GroupStart();
        for (int i = 0; i < n && wr == ncclSuccess; ++i)
            wr = nccl_rt::Reduce(rt_hdr_device_ptr(g->m[i].ctx), i == 0 ? root_buf : nullptr, count, ncclFloat, ncclSum, 0, g->m[i].comm,
                                 static_cast<cudaStream_t>(rt_get_stream(g->m[i].ctx)));
        const ncclResult_t we = nccl_rt::GroupEnd();
        if (wr == ncclSuccess) wr = we;
        if (wr != ncclSuccess) { rt_group_destroy(g); return gfail(RT_ERR_NCCL, "rt_group_create: warm-up ncclReduce failed: %s", nccl_rt::GetErrorString(wr)); }
        for (int i = 0; i < n; ++i) { const int rc = rt_synchronize(g->m[i].ctx); if (rc != RT_OK) { forward(rc); rt_group_destroy(g); return rc; } }
    }
    *out = g;
    return RT_OK;
}

int rt_group_size(const rt_group* g) { return g ? static_cast<int>(g->m.size()) : 0; }
rt_context* rt_group_member(rt_group* g, int i) { return (g && i >= 0 && i < static_cast<int>(g->m.size())) ? g->m[i].ctx : nullptr; }

#define FOR_MEMBERS(call) do { if (!g) return gfail(RT_ERR_INVALID_ARGUMENT, "group is NULL"); \
    for (Member& mb : g->m) { const int rc_ = (call); if (rc_ != RT_OK) return forward(rc_); } return RT_OK; } while (0)
int rt_group_set_scene(rt_group* g, const rt_scene* scene) { FOR_MEMBERS(rt_set_scene(mb.ctx, scene)); }
int rt_group_set_camera(rt_group* g, const rt_camera_params* cam) { FOR_MEMBERS(rt_set_camera(mb.ctx, cam)); }
int rt_group_set_exposure(rt_group* g, float stops) { FOR_MEMBERS(rt_set_exposure(mb.ctx, stops)); }
int rt_group_set_seed(rt_group* g, uint32_t seed) { FOR_MEMBERS(rt_set_seed(mb.ctx, seed)); }
int rt_group_clear(rt_group* g) { FOR_MEMBERS(rt_framebuffer_clear(mb.ctx)); }
int rt_group_synchronize(rt_group* g) { FOR_MEMBERS(rt_synchronize(mb.ctx)); }

int rt_group_render(rt_group* g, uint64_t first_sample, int spp, int max_depth, uint32_t flags) {
    if (!g) return gfail(RT_ERR_INVALID_ARGUMENT, "rt_group_render: group is NULL");
    const int n = static_cast<int>(g->m.size());
    // 1. every member renders its disjoint frame range into its own running sum.  Members 1.. START their sum with this call
    //    (RT_FLAG_OVERWRITE: no clear, the old content is not read; a clear where the schedule cannot overwrite or the member
    //    has no frames) so that after the reduce member 0 holds (previous total + this call), exactly what one GPU would hold.
    for (int i = 0; i < n; ++i) {
        Member& mb = g->m[i];
        uint64_t my_first; int my_spp;
        if (int rc = rt_group_shard(first_sample, spp, i, n, &my_first, &my_spp)) return rc;
        G_CUDA(cudaSetDevice(mb.device));
        cudaStream_t st = static_cast<cudaStream_t>(rt_get_stream(mb.ctx));
        const bool can_overwrite = i > 0 && my_spp > 0 && !(flags & (RT_FLAG_WAVEFRONT | RT_FLAG_DEBUG_ONE_KERNEL));
        if (i > 0 && !can_overwrite) { if (int rc = rt_framebuffer_clear(mb.ctx)) return forward(rc); }
        G_CUDA(cudaEventRecord(mb.e0, st));
        if (my_spp > 0) { if (int rc = rt_render(mb.ctx, my_first, my_spp, max_depth, flags | (can_overwrite ? RT_FLAG_OVERWRITE : 0u))) return forward(rc); }
        G_CUDA(cudaEventRecord(mb.e1, st));
    }
    // 2. one ncclReduce(sum) of the float4 buffers onto member 0, on the render streams (in place on the root)
    if (n > 1) {
        const size_t count = static_cast<size_t>(g->W) * g->H * 4;
        void* root_buf = rt_hdr_device_ptr(g->m[0].ctx);
        G_NCCL(nccl_rt::GroupStart());
        for (int i = 0; i < n; ++i) {
            Member& mb = g->m[i];
            cudaStream_t st = static_cast<cudaStream_t>(rt_get_stream(mb.ctx));
            const void* send = rt_hdr_device_ptr(mb.ctx);
            const ncclResult_t r = nccl_rt::Reduce(send, i == 0 ? root_buf : nullptr, count, ncclFloat, ncclSum, 0, mb.comm, st);
            if (r != ncclSuccess) { nccl_rt::GroupEnd(); return gfail(RT_ERR_NCCL, "ncclReduce failed on member %d: %s", i, nccl_rt::GetErrorString(r)); }
        }
        G_NCCL(nccl_rt::GroupEnd());
    }
    for (Member& mb : g->m) {
        G_CUDA(cudaSetDevice(mb.device));
        G_CUDA(cudaEventRecord(mb.e2, static_cast<cudaStream_t>(rt_get_stream(mb.ctx))));
    }
    g->timed = true;
    return RT_OK;
}

int rt_group_last_ms(rt_group* g, double* render_ms_max, double* reduce_ms_max) {
    if (!g || !g->timed) return gfail(RT_ERR_NOT_READY, "rt_group_last_ms: no rt_group_render yet");
    double rmax = 0, dmax = 0;
    for (Member& mb : g->m) {
        G_CUDA(cudaSetDevice(mb.device));
        G_CUDA(cudaEventSynchronize(mb.e2));
        float a = 0.f, b = 0.f;
        G_CUDA(cudaEventElapsedTime(&a, mb.e0, mb.e1));
        G_CUDA(cudaEventElapsedTime(&b, mb.e1, mb.e2));
        rmax = std::max(rmax, static_cast<double>(a)); dmax = std::max(dmax, static_cast<double>(b));
    }
    if (render_ms_max) *render_ms_max = rmax;
    if (reduce_ms_max) *reduce_ms_max = dmax;
    return RT_OK;
}

}  // extern "C"
