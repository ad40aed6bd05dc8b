// yb_noalloc.cu — yb200_trace_batch_static: the ray-batch hook of yb200_trace_batch (closest-hit / any-hit through the
// persistent traversal kernels of yb_trace_fast.cuh) for callers that must not allocate: code under CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in the
// reference.
//
// Everything the existing entry takes from the allocator - the ray-fetch counter of the persistent kernel and, for HOST
// pointers, the device staging of rays and results - lives in ONE file-scope device array of fixed size
// (YB200_STATIC_SCRATCH_BYTES, about 3 MiB; every GPU that loads the library holds it until exit).  The entry point calls
// no cudaMalloc / cudaFree / allocating library routine; it rejects batches that do not fit.
//
// The array is shared by every call on a GPU (all contexts, all streams) and keeps the previous call's contents: the
// counter is cleared on the stream before each launch, the staging areas are overwritten before they are read.  Calls
// that overlap on the GPU (two streams, two contexts, two threads) RACE on it - serialise them.
#include "../../include/yb200.h"
#include "yb_ctx.h"
#ifndef YB_HOSTEMU
#include "yb_trace_fast.cuh"
#include <mutex>
#endif

using namespace yb;

#ifndef YB_HOSTEMU
namespace {

// layout of the scratch array: [0, 256) ray-fetch counter (first 4 bytes) | rays of a host batch | results of a host batch
enum : size_t {
    kCounterBytes = 256,
    kRayBytes = 32,      // {from, tmin, dir, tmax}
    kOutBytesMax = 16,   // closest: {prim, t, b1, b2}; shadow: int32
    kRaysOff = kCounterBytes,
    kOutOff = kRaysOff + (size_t)YB200_STATIC_MAX_RAYS_HOST * kRayBytes,
    kScratchBytes = kOutOff + (size_t)YB200_STATIC_MAX_RAYS_HOST * kOutBytesMax,
};
static_assert(kScratchBytes == (size_t)YB200_STATIC_SCRATCH_BYTES, "include/yb200.h states the size of the static scratch");

static __device__ __align__(256) unsigned char g_static_scratch[kScratchBytes];

// In host code `g_static_scratch` is only a handle for the runtime, not a device address: the address is looked up once per
// device, with that device current, and kept here.
enum { kMaxDevices = 64 };
std::mutex g_scratch_mutex;
unsigned char *g_scratch_ptr[kMaxDevices];

// device `dev` must be current
unsigned char *static_scratch(int dev)
{
    if (dev < 0 || dev >= kMaxDevices) throw RtError("trace_batch_static: device index above 63");
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (!g_scratch_ptr[dev]) {
        unsigned char *p = 0;
        YB_CUDA_CHECK(cudaGetSymbolAddress((void **)&p, g_static_scratch));
        g_scratch_ptr[dev] = p;
    }
    return g_scratch_ptr[dev];
}

struct DeviceGuard {      // make the context's device current for the call, put the caller's back afterwards
    int prev; bool switched;
    explicit DeviceGuard(int dev) : prev(-1), switched(false)
    {
        YB_CUDA_CHECK(cudaGetDevice(&prev));
        if (prev != dev) { YB_CUDA_CHECK(cudaSetDevice(dev)); switched = true; }
    }
    ~DeviceGuard() { if (switched) cudaSetDevice(prev); }
};

}  // namespace
#endif

extern "C" int yb200_trace_batch_static(yb200_ctx *ctx, const float *rays, size_t n, int mode, void *out, int on_device)
{
    Context &c = ctx->c;
#ifdef YB_HOSTEMU
    (void)rays; (void)n; (void)mode; (void)out; (void)on_device;
    return c.fail("trace_batch_static: device-only entry point (not part of the host emulator)") ? 1 : 0;
#else
    try {
        if (!c.sceneReady) return c.fail("trace_batch_static: call update_scene first") ? 1 : 0;
        if (mode != YB200_TRACE_CLOSEST && mode != YB200_TRACE_SHADOW)
            return c.fail("trace_batch_static: mode must be YB200_TRACE_CLOSEST or YB200_TRACE_SHADOW (transparent shadows go through yb200_trace_batch)") ? 1 : 0;
        if (!on_device && n > (size_t)YB200_STATIC_MAX_RAYS_HOST)
            return c.fail("trace_batch_static: " + std::to_string(n) + " rays in host memory exceed the static staging area (YB200_STATIC_MAX_RAYS_HOST = "
                          + std::to_string((size_t)YB200_STATIC_MAX_RAYS_HOST) + "); split the batch or use yb200_trace_batch") ? 1 : 0;
        if (on_device && n > (size_t)YB200_STATIC_MAX_RAYS_DEVICE)
            return c.fail("trace_batch_static: " + std::to_string(n) + " rays exceed one launch of the 32-bit ray counter (YB200_STATIC_MAX_RAYS_DEVICE = "
                          + std::to_string((size_t)YB200_STATIC_MAX_RAYS_DEVICE) + "); split the batch or use yb200_trace_batch") ? 1 : 0;
        if (n == 0) return 1;
        if (!rays || !out) return c.fail("trace_batch_static: null ray or result pointer") ? 1 : 0;
        DeviceGuard guard(c.device);
        unsigned char *scratch = static_scratch(c.device);
        const size_t osz = mode == YB200_TRACE_SHADOW ? 4 : 16;
        const char *d_rays = (const char *)rays;
        void *d_out = out;
        if (!on_device) {
            d_rays = (const char *)(scratch + kRaysOff);
            d_out = scratch + kOutOff;
            rt_h2d(scratch + kRaysOff, rays, n * kRayBytes, c.stream);
        }
        FastJob j;
        j.nodes = c.sc.nodes; j.tri_v = c.sc.tri_v; j.rays = d_rays; j.stride = (uint32_t)kRayBytes; j.n = (uint32_t)n;
        j.order = 0; j.nPtr = 0; j.counterZeroed = false; j.counter = (uint32_t *)scratch; j.out = d_out; j.acc = 0; j.nNodes = c.sc.nNodes;
        j.surf.tri_v = 0; j.surf.tri_n = 0; j.surf.materials = 0; j.surf.prim_ng = 0; j.shadowDepth = 0;
        j.oneBits = 0x3F800000u; j.burst = 6; j.refillIdle = YB_REFILL_IDLE; j.triThreshold = 1;
        // launch_trace_fast clears the counter on the stream (cudaMemsetAsync) before the kernel: the last call's value must not survive
        if (mode == YB200_TRACE_CLOSEST) launch_trace_fast<false, OUT_CLOSEST_PRIM>(j, c.stream);
        else launch_trace_fast<true, OUT_SHADOW_FLAG>(j, c.stream);
        if (!on_device) rt_d2h(out, d_out, n * osz, c.stream);       // copies and waits for the stream
        if (mode == YB200_TRACE_CLOSEST) c.stats.rays_closest = n; else c.stats.rays_shadow = n;
        return 1;
    } catch (const std::exception &e) { c.fail(e.what()); return 0; }
#endif
}
