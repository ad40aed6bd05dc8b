// yb_shade_common.cuh — what the wavefront driver (yb_render.cu) and the three builds of the shade kernel
// (yb_shade_base.cu / yb_shade_ext.cu / yb_shade_nodes.cu, each its own translation unit so that they compile in parallel)
// share: the kernel argument block, camera-sample identity, QMC tables, sample accumulators and the task / shadow queues.
#pragma once
#include "yb_scene.h"
#include "yb_trace.cuh"

namespace yb {

#if defined(__CUDACC__) && !defined(YB_HOSTEMU)
#define YB_TABLE static __device__ __constant__
#define YB_TABLE_DIVERGENT static __device__
#else
#define YB_TABLE static
#define YB_TABLE_DIVERGENT static
#endif
#include "yb_faure.inc"
#if defined(__CUDA_ARCH__)
#define YB_LDG_U8(p) __ldg(p)
#else
#define YB_LDG_U8(p) (*(p))
#endif

#define YB_MIN_RAYDIST 0.00005f      // CMakeLists.txt:44
#define YB_SHADOW_BIAS 0.0005f       // CMakeLists.txt:48
#define YB_LOFFS_DELTA 4567u         // mcintegrator.cc:43

// scr_halton.h:47-69 (dim < 50 on this path: SURVEY.md §5)
YB_D uint32_t yb_umulhi(uint32_t a, uint32_t b)
{
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (uint32_t)(((uint64_t)a * (uint64_t)b) >> 32);
#endif
}
YB_D double scrHalton(int dim, uint32_t n)
{
    double value = 0.0;
    if (dim > 0 && dim < 50) {
        const unsigned char *sigma = yb_faure_flat + yb_faure_offs[dim];
        uint32_t base = (uint32_t)yb_prims[dim];
        double f, factor, dn = (double)n;
        f = factor = yb_invPrims[dim];
        const uint32_t magic = yb_primMagic[dim];
        while (n > 0) {
            // n % base without a runtime division: the mul-hi quotient is exact or one too large
            uint32_t r = n - yb_umulhi(n, magic) * base;
            if ((int32_t)r < 0) r += base;
            value += (double)YB_LDG_U8(sigma + r) * factor;
            dn *= f;
            n = (uint32_t)dn;
            factor *= f;
        }
    }
    return value < 1.0e-36 ? 1.0e-36 : (value > 1.0 ? 1.0 : value);
}

// one direct-lighting estimate still to do: task index in the current wave, light index | bin << 24 (bin = group key x 4 + slot of the
// material within the group: jobs are ordered by bin, a group's kernel runs over its four bins at once - materials of one class still
// take different branches (Blinn / anisotropic lobes, Lambert / Oren-Nayar), so a warp should see one material where it can)
struct LightJob { uint32_t task, lightKey; };
// group of a job: light type (LIGHT_POINT .. LIGHT_SPHERE; mesh lights run in the ext build, which is not staged) x material class
// (0..2 = shinydiffuse / glossy / coated_glossy, the materials with eval / pdf bodies; 3 = everything else) x (trajectory split or not:
// a sampled light takes light.samples / rayDivision samples, and a warp that mixes 8-sample and 1-sample jobs idles 3/4 of its lanes)
#define YB_LJOB_MCLASSES 4
YB_HD uint32_t light_job_key(int lightType, int matType, uint32_t rayDivision)
{
    const bool sampled = !(lightType == LIGHT_POINT || lightType == LIGHT_DIRECTIONAL);
    return (((uint32_t)lightType * YB_LJOB_MCLASSES + (uint32_t)(matType < 3 ? matType : 3)) << 1) | (sampled && rayDivision > 1 ? 1u : 0u);
}

struct RenderArgs {
    SceneDev sc;
    Task *tasks;            // chunk being processed
    Hit *hits;
    Task *out_tasks;        // children appended here
    ShadowTask *shadow;     // shadow rays appended here
    uint32_t *counters;     // [0] out_tasks count, [1] shadow count
    const uint32_t *nPtr;   // optional: the number of tasks of this wave lives in device memory (the launch's n is an upper bound)
    const uint32_t *order;  // shading order of this wave (0 = identity): tasks binned by shade_key
    uint32_t *orderOut; unsigned char *keys; uint32_t *bins;   // binning scratch: bins[0..7] counts, bins[8..15] cursors
    f4 *acc;                // per camera sample: rgb + alpha
    const uint32_t *pix;    // owned pixels: x | y << 16
    float *film;            // (x + w*y)*5: r,g,b,a,weight
    float *depthT;          // z_channel: per camera sample, c_ray.tmax after integrate() (NaN: no depth sample); 0 = off
    float *depthFilm;       // z_channel: (x + w*y)*2: sum(depth*w), sum(w)  (gray2DImage_t / pixelGray_t)
    uint32_t sampleBase;    // first global camera-sample index of this batch
    uint32_t outCap, shadowCap;
    // staged direct lighting (yb_direct.cu): (task, light) jobs queued by the staged shade kernel, then grouped by
    // (light type x material class); counters[YB_CNT_LJOBS] = number of jobs, counters[YB_CNT_LBINS + key] = jobs per group
    LightJob *ljobs, *ljobsSorted;
    uint32_t ljobCap;
};

// slots of the per-wave counter block (RenderArgs::counters)
enum { YB_CNT_TASKS = 0, YB_CNT_SHADOW = 1, YB_CNT_TRACE_A = 2, YB_CNT_TRACE_B = 3, YB_CNT_LJOBS = 4, YB_CNT_LBINS = 8, YB_LGROUPS = 64, YB_LBINS = 4 * YB_LGROUPS,
       YB_CNT_LCURSOR = YB_CNT_LBINS + YB_LBINS, YB_CNT_WORDS = YB_CNT_LCURSOR + YB_LBINS };

// per-camera-sample identity (integrator.cc:251-273)
struct SampleId { int x, y, s; uint32_t samplingOffs; int pixelSample; };
YB_D SampleId sample_id(const RenderArgs &a, uint32_t sid)
{
    SampleId r;
    const uint32_t gid = a.sampleBase + sid;
    const uint32_t ns = (uint32_t)a.sc.film.n_samples;
    const uint32_t p = a.pix[gid / ns];
    r.s = (int)(gid % ns);
    r.x = (int)(p & 0xffffu); r.y = (int)(p >> 16);
    r.samplingOffs = fnv_32a_buf((uint32_t)r.y * fnv_32a_buf((uint32_t)r.x));
    r.pixelSample = a.sc.film.pass_offs + r.s;
    return r;
}
// sub-pixel position of a sample (integrator.cc:238,277-286)
YB_D void sample_dxdy(const Film &f, const SampleId &id, float &dx, float &dy)
{
    dx = 0.5f; dy = 0.5f;
    if (f.AA_passes > 1) { dx = RI_vdC((uint32_t)id.pixelSample, id.samplingOffs); dy = RI_S((uint32_t)id.pixelSample, id.samplingOffs); }
    else if (f.n_samples > 1) {
        const float d1 = (float)(1.0 / (float)f.n_samples);
        dx = (float)((0.5 + (double)(float)id.s) * (double)d1);
        dy = RI_LP((uint32_t)id.s + id.samplingOffs);
    }
}

YB_D void acc_add(const RenderArgs &a, uint32_t sid, col3 c)
{
    YB_ATOMIC_ADD_F32(&a.acc[sid].x, c.x); YB_ATOMIC_ADD_F32(&a.acc[sid].y, c.y); YB_ATOMIC_ADD_F32(&a.acc[sid].z, c.z);
}
YB_D void acc_add_alpha(const RenderArgs &a, uint32_t sid, float v) { YB_ATOMIC_ADD_F32(&a.acc[sid].w, v); }
// ------------------------------------------------------------------------------------------------ queues
// the shadow queue as three words, so that out-of-line functions take it by value instead of a reference to the kernel's
// parameter block (a reference would make every thread copy the 2 KB RenderArgs to local memory)
struct ShadowQ { ShadowTask *shadow; uint32_t *counters; uint32_t shadowCap; };
YB_D ShadowQ shadow_q(const RenderArgs &a) { ShadowQ q; q.shadow = a.shadow; q.counters = a.counters; q.shadowCap = a.shadowCap; return q; }
YB_D void push_shadow(const ShadowQ q, const Ray &r, col3 c, uint32_t sid)
{
    if (isBlack(c)) return;                                   // contributes nothing either way
    const uint32_t i = YB_ATOMIC_ADD_U32(q.counters + 1, 1u);
    if (i >= q.shadowCap) return;                             // cannot happen: the host sizes chunks for the worst case
    ShadowTask st; st.ray = r; st.c = c; st.sid = sid;
    q.shadow[i] = st;
}
YB_D void push_shadow(const RenderArgs &a, const Ray &r, col3 c, uint32_t sid) { push_shadow(shadow_q(a), r, c, sid); }
YB_D void push_task(const RenderArgs &a, const Task &t)
{
    const uint32_t i = YB_ATOMIC_ADD_U32(a.counters + 0, 1u);
    if (i >= a.outCap) return;
    a.out_tasks[i] = t;
}

YB_D void put_light_job(const RenderArgs &a, uint32_t i, uint32_t task, uint32_t light, uint32_t key, const Material *mat)
{
    key = (key << 2) | ((uint32_t)(mat - a.sc.materials) & 3u);
    if (i >= a.ljobCap) return;                               // cannot happen: sized for (tasks of a wave) x (lights)
    LightJob j; j.task = task; j.lightKey = light | (key << 24);
    a.ljobs[i] = j;
}
YB_D void push_light_job(const RenderArgs &a, uint32_t task, uint32_t light, uint32_t key, const Material *mat)
{
    put_light_job(a, YB_ATOMIC_ADD_U32(a.counters + YB_CNT_LJOBS, 1u), task, light, key, mat);
}

// background_t::operator() of the supported backgrounds: constant (constant colour) and gradientBackground_t::eval (gradientback.cc:60-80)
YB_D col3 background_eval(const Integrator &I, v3 dir)
{
    if (!I.bgGradient) return I.bgColor;
    float blend = dir.z;
    col3 color;
    if (blend >= 0.f) color = blend * I.bgSZenith + (1.f - blend) * I.bgSHoriz;
    else { blend = -blend; color = blend * I.bgGZenith + (1.f - blend) * I.bgGHoriz; }
    if (fminf_(color.x, fminf_(color.y, color.z)) < 1e-6f) color = col3(1e-5f);
    return color;
}

// the light estimateOneDirectLight picks for path sample n (mcintegrator.cc:62-66)
YB_D int pick_one_light(int n, int lightNum)
{
    Halton2 hal2;
    hal2.setStart((uint32_t)(n - 1));
    int lnum = (int)(hal2.getNext() * (float)lightNum);
    if (lnum > lightNum - 1) lnum = lightNum - 1;
    return lnum;
}


// the three builds of k_shade (yb_shade.inc): base = the four core materials; ext = + rough_glass, blend_mat, mesh lights;
// nodes = + shader nodes / textures / bump
void launch_shade_base(uint32_t n, const RenderArgs &a, rt_stream s);
void launch_shade_ext(uint32_t n, const RenderArgs &a, rt_stream s);
void launch_shade_nodes(uint32_t n, const RenderArgs &a, rt_stream s);
// the base build for waves that hold one kind of task only (kind = TASK_INTEGRATE / TASK_PATH): half the kernel each
void launch_shade_base_kind(uint32_t kind, uint32_t n, const RenderArgs &a, rt_stream s);
// the same for scenes whose materials are all shinydiffuse (material dispatch folded)
void launch_shade_sd_kind(uint32_t kind, uint32_t n, const RenderArgs &a, rt_stream s);
// staged form of the base build: launch_shade_staged queues the direct-lighting estimates, launch_direct evaluates the n jobs
// of one (light type, material class) group that start at ljobsSorted[base]
void launch_shade_staged(uint32_t n, const RenderArgs &a, rt_stream s);
void launch_direct(uint32_t key, uint32_t base, uint32_t n, const RenderArgs &a, rt_stream s);

}  // namespace yb
