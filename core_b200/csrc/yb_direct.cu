// yb_direct.cu — staged direct lighting: mcIntegrator_t::doLightEstimation (mcintegrator.cc:73-195) as its own wavefront stage.
//
// Why: on scenes with several material classes and light types (BASELINE configs[3]) the fully inlined shade kernel is bound by
// instruction fetch - a task walks ~3.7 k warp instructions (eval + pdf + sample of its material for every light sample, the MIS
// loop, path spawn, glossy fan-out), far more than the 32 KB instruction cache of an SM holds.  Here one launch
// evaluates the jobs of ONE (light type, material class) group: both types are template parameters, every dispatch switch of
// yb_lights.cuh / yb_bsdf.cuh folds, and all lanes of all warps of the launch run the same short loop.
//
// A job names a task of the current wave and a light.  The surface point, material data and sample identity are recomputed from
// the task and its hit exactly as the shade kernel computes them (same functions, same inputs: same bits), then
// doLightEstimation<LT, FIX> pushes the shadow rays of the estimate onto the wave's shadow queue.
#include "yb_shade_common.cuh"
namespace yb {
#define YB_SHADE_NS shade_direct
#define YB_EXT 0
#define YB_DIRECT_ONLY 1
#include "yb_shade.inc"

namespace shade_direct {

template <int LT, int FIX> YB_D void direct_job(const RenderArgs &a, const LightJob J)
{
    const Task T = a.tasks[J.task];
    const Hit H = a.hits[J.task];
    const uint32_t l = J.lightKey & 0x00ffffffu;
    const SampleId id = sample_id(a, T.sid);
    Surf sp;
    tri_getSurface(a.sc, H.tri, T.ray.from + H.t * T.ray.dir, H.b1, H.b2, sp);
    const Material &mat = *sp.mat;
    MatDat md; uint32_t bsdfs;
    mat_initBSDF<0, FIX>(mat, md, bsdfs);
    const v3 wo = -T.ray.dir;
    // estimateAllDirectLight for an integrate() task, estimateOneDirectLight (weight x number of lights) for a path vertex
    const col3 weight = task_kind(T.bits) == TASK_INTEGRATE ? T.thr : T.thr * (float)a.sc.nLights;
    doLightEstimation<LT, FIX>(shadow_q(a), a.sc.lights[l], l, sp, md, wo, id.pixelSample, id.samplingOffs, T.rayDivision, weight, T.sid);
}

#ifndef YB_HOSTEMU
// point / directional lights (one shadow ray, no loops) fit 64 registers; the sampled lights take 80 with a few spilled words
template <int LT, int FIX> __global__ void __launch_bounds__(256, (LT == LIGHT_POINT || LT == LIGHT_DIRECTIONAL) ? 4 : 3) k_direct_kernel(uint32_t base, uint32_t n, const RenderArgs a)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) direct_job<LT, FIX>(a, a.ljobsSorted[base + i]);
}
template <int LT, int FIX> static void k_direct(uint32_t base, uint32_t n, const RenderArgs &a, rt_stream s)
{
    uint32_t blocks = (n + 255u) / 256u;
    if (blocks > 148u * 32u) blocks = 148u * 32u;
    k_direct_kernel<LT, FIX><<<blocks, 256, 0, s>>>(base, n, a);
    YB_CUDA_CHECK(cudaGetLastError());
    g_kernel_launches++;
}
#else
template <int LT, int FIX> static void k_direct(uint32_t base, uint32_t n, const RenderArgs &a, rt_stream)
{
    for (uint32_t i = 0; i < n; ++i) direct_job<LT, FIX>(a, a.ljobsSorted[base + i]);
    g_kernel_launches++;
}
#endif

template <int LT> static void direct_by_material(uint32_t mclass, uint32_t base, uint32_t n, const RenderArgs &a, rt_stream s)
{
    switch (mclass) {
    case 0: k_direct<LT, MAT_SHINYDIFFUSE>(base, n, a, s); break;
    case 1: k_direct<LT, MAT_GLOSSY>(base, n, a, s); break;
    case 2: k_direct<LT, MAT_COATED_GLOSSY>(base, n, a, s); break;
    default: k_direct<LT, -1>(base, n, a, s); break;
    }
}

}  // namespace shade_direct

void launch_direct(uint32_t key, uint32_t base, uint32_t n, const RenderArgs &a, rt_stream s)
{
    if (n == 0) return;
    const uint32_t lt = (key >> 1) / YB_LJOB_MCLASSES, mc = (key >> 1) % YB_LJOB_MCLASSES;
    switch ((int)lt) {
    case LIGHT_POINT: shade_direct::direct_by_material<LIGHT_POINT>(mc, base, n, a, s); break;
    case LIGHT_SUN: shade_direct::direct_by_material<LIGHT_SUN>(mc, base, n, a, s); break;
    case LIGHT_AREA: shade_direct::direct_by_material<LIGHT_AREA>(mc, base, n, a, s); break;
    case LIGHT_DIRECTIONAL: shade_direct::direct_by_material<LIGHT_DIRECTIONAL>(mc, base, n, a, s); break;
    case LIGHT_SPOT: shade_direct::direct_by_material<LIGHT_SPOT>(mc, base, n, a, s); break;
    case LIGHT_SPHERE: shade_direct::direct_by_material<LIGHT_SPHERE>(mc, base, n, a, s); break;
    default: throw RtError("launch_direct: light type outside the staged set");
    }
}

}  // namespace yb
