// yb_render.cu — the wavefront form of tiledIntegrator_t::renderTile + directLighting_t / pathIntegrator_t
// ::integrate (src/yafraycore/integrator.cc:229-339, src/integrators/directlight.cc:112-182,
// src/integrators/pathtracer.cc:134-333, src/yafraycore/mcintegrator.cc:45-195,421-683) and of
// imageFilm_t::addSample / flush (src/yafraycore/imagefilm.cc:453-511,383-443).
//
// The reference recursion integrate() -> recursiveRaytrace() -> integrate() is linear in the returned colour,
// so it is unrolled into a pool of independent tasks: every task carries the product of the weights above it
// (`thr`) and adds its own terms straight into the camera sample's accumulator.  One wave = closest-hit trace of
// a chunk of tasks, shading (which appends shadow rays and child tasks), any-hit trace of the shadow rays.
// All sample streams are replayed exactly as SURVEY.md §3.6 maps them.
#include "yb_shade_common.cuh"
#include "yb_trace_fast.cuh"
#include <chrono>
#include <vector>
#include <algorithm>

namespace yb {

// ------------------------------------------------------------------------------------------------ raygen
YB_KERNEL(k_raygen, RenderArgs)
{
    const SampleId id = sample_id(a, tid);
    float dx, dy;
    sample_dxdy(a.sc.film, id, dx, dy);
    float lens_u = 0.5f, lens_v = 0.5f;
    if (a.sc.cam.sampleLens) {          // integrator.cc:248-249,266-267,287-291
        Halton3 halU; Halton5 halV;
        halU.setStart((uint32_t)a.sc.film.pass_offs + id.samplingOffs);
        halV.setStart((uint32_t)a.sc.film.pass_offs + id.samplingOffs);
        for (int k = 0; k <= id.s; ++k) { lens_u = halU.getNext(); lens_v = halV.getNext(); }
    }
    float wt;
    Task t;
    t.ray = cam_shootRay(a.sc.cam, (float)id.x + dx, (float)id.y + dy, lens_u, lens_v, wt);
    t.thr = col3(1.f); t.aw = 1.f; t.sid = tid;
    // wt == 0 (angular camera outside its circle): the reference adds a black, alpha-0 sample without tracing
    // (integrator.cc:293-297); here the slot keeps a ray that cannot hit (tmin = tmax = 0) and carries zero weights
    if (wt == 0.f) { t.thr = col3(0.f); t.aw = 0.f; }
    if (a.depthT) a.depthT[tid] = wt == 0.f ? i2f(0x7fc00000) : t.ray.tmax;   // integrator.cc:293-297: no depth sample when wt == 0
    t.bits = task_bits(TASK_INTEGRATE, true, 0, 0);
    t.rayDivision = 1; t.rayOffset = 0; t.dc1 = 0.f; t.dc2 = 0.f; t.pathOffs = 0; t.pad = 0;
    a.tasks[tid] = t;
    f4 z; z.x = z.y = z.z = z.w = 0.f;
    a.acc[tid] = z;
}

// ------------------------------------------------------------------------------------------------ trace kernels
YB_KERNEL(k_trace_closest, RenderArgs)
{
    Hit h;
    scene_intersect(a.sc, a.tasks[tid].ray, h);
    a.hits[tid] = h;
}
YB_KERNEL(k_trace_shadow, RenderArgs)
{
    const ShadowTask st = a.shadow[tid];
    if (a.sc.integ.trShad) {
        col3 filt;
        if (!scene_isShadowedT(a.sc, st.ray, a.sc.integ.sDepth, filt)) acc_add(a, st.sid, mul(st.c, filt));
    } else if (!scene_isShadowed(a.sc, st.ray)) acc_add(a, st.sid, st.c);
}

// ---- shading order.  After the first wave the tasks of one warp take very different code paths (hit / miss, the
// light estimateOneDirectLight picks): measured 8 of 32 lanes active.  Tasks are binned by a small key so that warps
// of the shade kernel are homogeneous; the kernel then gathers its task through `order`.
#define YB_SHADE_BINS 256
// key: 0 = miss; 8..35 = integrate() task by material type of the hit (the BSDF code of the material classes shares nothing, and
// recursiveRaytrace children of a glossy / specular surface land on every material: 1080p materials scene, shade 14.4 s -> see
// profiles/) x material slot (index & 3: materials of one type still branch on their parameters - Blinn / anisotropic lobe,
// Lambert / Oren-Nayar, mirror / transparency components); 64..255 = path vertex by (light estimateOneDirectLight picks: 0, 1, 2+;
// not when direct lighting is staged) x (material type) x (material slot)
YB_D uint32_t shade_key(const RenderArgs &a, uint32_t t)
{
    const int tri = a.hits[t].tri;
    if (tri < 0) return 0u;
    const uint32_t bits = a.tasks[t].bits;
    const uint32_t mi = (uint32_t)f2i(a.sc.tri_v[3 * (size_t)tri + 1].w) & TRI_MAT_MASK;
    const uint32_t ty = (uint32_t)a.sc.materials[mi].type;
    if (task_kind(bits) == TASK_INTEGRATE) return 8u + 4u * (ty < 6u ? ty : 6u) + (mi & 3u);
    const int l = a.ljobs ? 0 : pick_one_light((int)a.tasks[t].pathOffs, a.sc.nLights);
    return 64u + 64u * (uint32_t)(l < 0 ? 0 : (l < 2 ? l : 2)) + 4u * (ty < 7u ? ty : 7u) + (mi & 3u);      // l = -1 without lights
}
#ifndef YB_HOSTEMU
__global__ void __launch_bounds__(256) k_bin_count_kernel(uint32_t n, const RenderArgs a)
{
    __shared__ uint32_t h[YB_SHADE_BINS];
    static_assert(YB_SHADE_BINS == 256, "one thread per bin");
    if (a.nPtr) n = min(n, *a.nPtr);
    h[threadIdx.x] = 0u;
    __syncthreads();
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
        const uint32_t k = shade_key(a, t);
        a.keys[t] = (unsigned char)k;
        atomicAdd(&h[k], 1u);
    }
    __syncthreads();
    if (h[threadIdx.x]) atomicAdd(&a.bins[threadIdx.x], h[threadIdx.x]);
}
// exclusive prefix over 256 per-thread values (one per bin), result in sh[threadIdx.x]
__device__ __forceinline__ void block_exclusive_scan_256(uint32_t own, uint32_t *sh)
{
    uint32_t v = own;
    sh[threadIdx.x] = v;
    __syncthreads();
    for (uint32_t off = 1; off < 256u; off <<= 1) {
        const uint32_t add = threadIdx.x >= off ? sh[threadIdx.x - off] : 0u;
        __syncthreads();
        v += add; sh[threadIdx.x] = v;
        __syncthreads();
    }
    sh[threadIdx.x] = v - own;
    __syncthreads();
}
// rank of this thread's item among the items of the same bin in the block, in thread order (a stable counting sort), and the bin
// sizes of the block in h[]; 256 threads, keys < 256
__device__ __forceinline__ uint32_t block_stable_rank_256(uint32_t k, bool valid, uint32_t (*wcnt)[256], uint32_t *h)
{
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
#pragma unroll
    for (int w = 0; w < 8; ++w) wcnt[w][threadIdx.x] = 0u;
    __syncthreads();
    const unsigned peers = __match_any_sync(0xffffffffu, valid ? k : 0x100u + lane);
    const uint32_t r = (uint32_t)__popc(peers & ((1u << lane) - 1u));
    if (valid && r == 0u) wcnt[warp][k] = (uint32_t)__popc(peers);
    __syncthreads();
    uint32_t run = 0u;
#pragma unroll
    for (int w = 0; w < 8; ++w) { const uint32_t c = wcnt[w][threadIdx.x]; wcnt[w][threadIdx.x] = run; run += c; }
    h[threadIdx.x] = run;
    __syncthreads();
    return valid ? wcnt[warp][k] + r : 0u;
}
__global__ void __launch_bounds__(256) k_bin_scatter_kernel(uint32_t n, const RenderArgs a)
{
    __shared__ uint32_t h[YB_SHADE_BINS], base[YB_SHADE_BINS], start[YB_SHADE_BINS];
    if (a.nPtr) n = min(n, *a.nPtr);
    block_exclusive_scan_256(a.bins[threadIdx.x], start);
    __shared__ uint32_t wcnt[8][YB_SHADE_BINS];
    for (uint32_t c0 = blockIdx.x * blockDim.x; c0 < n; c0 += gridDim.x * blockDim.x) {     // same trip count for the whole block
        const uint32_t t = c0 + threadIdx.x;
        uint32_t k = 0u;
        if (t < n) k = a.keys[t];
        const uint32_t rank = block_stable_rank_256(k, t < n, wcnt, h);      // tasks of a bin keep their pool order within the block
        base[threadIdx.x] = start[threadIdx.x] + (h[threadIdx.x] ? atomicAdd(&a.bins[YB_SHADE_BINS + threadIdx.x], h[threadIdx.x]) : 0u);
        __syncthreads();
        if (t < n) a.orderOut[base[k] + rank] = t;
        __syncthreads();
    }
}
#endif
static void bin_tasks(uint32_t n, const RenderArgs &a, rt_stream s)
{
    rt_memset(a.bins, 0, 2 * YB_SHADE_BINS * sizeof(uint32_t), s);
#ifndef YB_HOSTEMU
    uint32_t blocks = (n + 255u) / 256u;
    k_bin_count_kernel<<<blocks > 148u * 8u ? 148u * 8u : blocks, 256, 0, s>>>(n, a);
    k_bin_scatter_kernel<<<blocks > 148u * 8u ? 148u * 8u : blocks, 256, 0, s>>>(n, a);
    YB_CUDA_CHECK(cudaGetLastError());
#else
    if (a.nPtr && *a.nPtr < n) n = *a.nPtr;
    for (uint32_t t = 0; t < n; ++t) { a.keys[t] = (unsigned char)shade_key(a, t); a.bins[a.keys[t]]++; }
    uint32_t base[YB_SHADE_BINS], acc = 0;
    for (int b = 0; b < YB_SHADE_BINS; ++b) { base[b] = acc; acc += a.bins[b]; }
    for (uint32_t t = 0; t < n; ++t) a.orderOut[base[a.keys[t]]++] = t;
#endif
    g_kernel_launches += 2;
}

// ---- staged direct lighting: group the (task, light) jobs of a wave by (light type x material class) so that each group runs as
//      one launch of its own specialised kernel (yb_direct.cu).  The job count lives in device memory (counters[YB_CNT_LJOBS]); the
//      host reads the group sizes afterwards, together with the other counters of the wave.
#ifndef YB_HOSTEMU
__global__ void __launch_bounds__(256) k_ljob_count_kernel(const RenderArgs a)
{
    __shared__ uint32_t h[YB_LBINS];
    const uint32_t n = min(a.counters[YB_CNT_LJOBS], a.ljobCap);
    h[threadIdx.x] = 0u;
    __syncthreads();
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) atomicAdd(&h[a.ljobs[t].lightKey >> 24], 1u);
    __syncthreads();
    if (h[threadIdx.x]) atomicAdd(&a.counters[YB_CNT_LBINS + threadIdx.x], h[threadIdx.x]);
}
__global__ void __launch_bounds__(256) k_ljob_scatter_kernel(const RenderArgs a)
{
    __shared__ uint32_t h[YB_LBINS], base[YB_LBINS], start[YB_LBINS];
    const uint32_t n = min(a.counters[YB_CNT_LJOBS], a.ljobCap);
    static_assert(YB_LBINS == 256, "one thread per bin");
    block_exclusive_scan_256(a.counters[YB_CNT_LBINS + threadIdx.x], start);
    __shared__ uint32_t wcnt[8][YB_LBINS];
    for (uint32_t c0 = blockIdx.x * blockDim.x; c0 < n; c0 += gridDim.x * blockDim.x) {     // same trip count for the whole block
        const uint32_t t = c0 + threadIdx.x;
        LightJob j; j.task = 0u; j.lightKey = 0u;
        uint32_t k = 0u;
        if (t < n) { j = a.ljobs[t]; k = j.lightKey >> 24; }
        // stable rank of the job among the jobs of its bin in this block (order of arrival kept: neighbouring jobs stay neighbours, and
        // with them the surface points whose shadow rays one warp of the direct-lighting kernel queues side by side)
        const uint32_t rank = block_stable_rank_256(k, t < n, wcnt, h);
        base[threadIdx.x] = start[threadIdx.x] + (h[threadIdx.x] ? atomicAdd(&a.counters[YB_CNT_LCURSOR + threadIdx.x], h[threadIdx.x]) : 0u);
        __syncthreads();
        if (t < n) a.ljobsSorted[base[k] + rank] = j;
        __syncthreads();
    }
}
#endif
static void group_light_jobs(size_t maxJobs, const RenderArgs &a, rt_stream s)
{
#ifndef YB_HOSTEMU
    uint32_t blocks = (uint32_t)std::min<size_t>((maxJobs + 255u) / 256u, 148u * 8u);
    if (blocks < 1u) blocks = 1u;
    k_ljob_count_kernel<<<blocks, 256, 0, s>>>(a);
    k_ljob_scatter_kernel<<<blocks, 256, 0, s>>>(a);
    YB_CUDA_CHECK(cudaGetLastError());
#else
    (void)maxJobs;
    const uint32_t n = std::min(a.counters[YB_CNT_LJOBS], a.ljobCap);
    for (uint32_t t = 0; t < n; ++t) a.counters[YB_CNT_LBINS + (a.ljobs[t].lightKey >> 24)]++;
    uint32_t base[YB_LBINS], acc = 0;
    for (int b = 0; b < YB_LBINS; ++b) { base[b] = acc; acc += a.counters[YB_CNT_LBINS + b]; }
    for (uint32_t t = 0; t < n; ++t) a.ljobsSorted[base[a.ljobs[t].lightKey >> 24]++] = a.ljobs[t];
#endif
    g_kernel_launches += 2;
}

// ------------------------------------------------------------------------------------------------ film
// imageFilm_t::addSample (imagefilm.cc:453-511); film layout (x + w*y)*5 = r,g,b,a,weight (pixel_t)
YB_D void film_splat_sample(const RenderArgs &a, uint32_t tid)
{
    const Film &f = a.sc.film;
    const SampleId id = sample_id(a, tid);
    float dx, dy;
    sample_dxdy(f, id, dx, dy);
    f4 c = a.acc[tid];                                           // wt == 1 for the perspective camera
#ifdef YB_HOSTEMU      // emulator only: per-sample radiance dump for chasing a differing pixel against the reference
    if (getenv("YB200_DEBUG_ACC")) fprintf(stderr, "ACC %u %d %d %d %.9g %.9g %.9g %.9g\n", tid, id.x, id.y, id.s, c.x, c.y, c.z, c.w);
#endif
    if (f.clamp) {
        c.x = c.x < 0.f ? 0.f : (c.x > 1.f ? 1.f : c.x); c.y = c.y < 0.f ? 0.f : (c.y > 1.f ? 1.f : c.y); c.z = c.z < 0.f ? 0.f : (c.z > 1.f ? 1.f : c.z);
    }
    const int x = id.x, y = id.y;
    int dx0 = Round2Int((double)dx - f.filterw), dx1 = Round2Int((double)dx + f.filterw - 1.0);
    int dy0 = Round2Int((double)dy - f.filterw), dy1 = Round2Int((double)dy + f.filterw - 1.0);
    if (dx0 < f.cx0 - x) dx0 = f.cx0 - x;
    if (dx1 > f.cx1 - x - 1) dx1 = f.cx1 - x - 1;
    if (dy0 < f.cy0 - y) dy0 = f.cy0 - y;
    if (dy1 > f.cy1 - y - 1) dy1 = f.cy1 - y - 1;
    int xIndex[YB_MAX_FILTER_SIZE + 1], yIndex[YB_MAX_FILTER_SIZE + 1];
    const double x_offs = dx - 0.5, y_offs = dy - 0.5;
    for (int i = dx0, n = 0; i <= dx1 && n <= YB_MAX_FILTER_SIZE; ++i, ++n) xIndex[n] = Floor2Int(fabs(((double)i - x_offs) * f.tableScale));
    for (int i = dy0, n = 0; i <= dy1 && n <= YB_MAX_FILTER_SIZE; ++i, ++n) yIndex[n] = Floor2Int(fabs(((double)i - y_offs) * f.tableScale));
    const int x0 = x + dx0, x1 = x + dx1, y0 = y + dy0, y1 = y + dy1;
    for (int j = y0; j <= y1; ++j)
        for (int i = x0; i <= x1; ++i) {
            const int offset = yIndex[j - y0] * YB_FILTER_TABLE_SIZE + xIndex[i - x0];
            const float filterWt = f.filterTable[offset];
            if (f.premult) { c.x *= c.w; c.y *= c.w; c.z *= c.w; }      // sic: inside the footprint loop (imagefilm.cc:502)
            float *px = a.film + ((size_t)(i - f.cx0) + (size_t)f.w * (size_t)(j - f.cy0)) * 5;
            YB_ATOMIC_ADD_F32(px + 0, c.x * filterWt); YB_ATOMIC_ADD_F32(px + 1, c.y * filterWt);
            YB_ATOMIC_ADD_F32(px + 2, c.z * filterWt); YB_ATOMIC_ADD_F32(px + 3, c.w * filterWt);
            YB_ATOMIC_ADD_F32(px + 4, filterWt);
        }
    if (a.depthT) {                                              // integrator.cc:313-334 + imageFilm_t::addDepthSample (imagefilm.cc:513-564)
        const float tmax = a.depthT[tid];
        if (tmax == tmax) {
            float depth = 0.f;
            if (f.zNorm) { if (tmax > 0.f) depth = 1.f - (tmax - f.minDepth) * f.maxDepth; }
            else { depth = tmax; if (depth <= 0.f) depth = 99999997952.f; }
            for (int j = y0; j <= y1; ++j)
                for (int i = x0; i <= x1; ++i) {
                    const float filterWt = f.filterTable[yIndex[j - y0] * YB_FILTER_TABLE_SIZE + xIndex[i - x0]];
                    float *px = a.depthFilm + ((size_t)(i - f.cx0) + (size_t)f.w * (size_t)(j - f.cy0)) * 2;
                    YB_ATOMIC_ADD_F32(px + 0, depth * filterWt); YB_ATOMIC_ADD_F32(px + 1, filterWt);
                }
        }
    }
}
YB_KERNEL(k_film, RenderArgs) { film_splat_sample(a, tid); }
#ifndef YB_HOSTEMU
// Warp-aggregated form of the splat for the common case (no premult, footprint of every sample of the warp inside the 3x3
// neighbourhood of its pixel: filter half-width up to about 1.5 pixels - box and lanczos: AA_pixelwidth / 2, gauss: AA_pixelwidth,
// mitchell: 1.3 AA_pixelwidth; wider filters send the warp to the per-sample path): consecutive lanes hold consecutive samples of the same pixel, so the 3x3 x 5 partial sums
// are reduced over each run of equal pixels with shuffles and only the run's first lane touches the film - 45 atomics per pixel
// and warp instead of 20 per sample.  Anything else (wide filters, premult, the depth channel) takes k_film's per-sample path.
__global__ void __launch_bounds__(256) k_film_fast_kernel(uint32_t n, const RenderArgs a)
{
    const Film &f = a.sc.film;
    const unsigned lane = threadIdx.x & 31u;
    for (uint32_t base = blockIdx.x * blockDim.x; base < n; base += gridDim.x * blockDim.x) {
        const uint32_t tid = base + threadIdx.x;
        const bool valid = tid < n;
        bool fits = false;
        int x = 0, y = 0, dx0 = 0, dx1 = -1, dy0 = 0, dy1 = -1;
        int xIndex[3] = {0, 0, 0}, yIndex[3] = {0, 0, 0};
        f4 c; c.x = c.y = c.z = c.w = 0.f;
        uint32_t key = 0xffffffffu - lane;                        // invalid lanes: runs of their own
        if (valid) {
            const SampleId id = sample_id(a, tid);
            float dx, dy;
            sample_dxdy(f, id, dx, dy);
            c = a.acc[tid];
            if (f.clamp) {
                c.x = c.x < 0.f ? 0.f : (c.x > 1.f ? 1.f : c.x); c.y = c.y < 0.f ? 0.f : (c.y > 1.f ? 1.f : c.y); c.z = c.z < 0.f ? 0.f : (c.z > 1.f ? 1.f : c.z);
            }
            x = id.x; y = id.y;
            dx0 = Round2Int((double)dx - f.filterw); dx1 = Round2Int((double)dx + f.filterw - 1.0);
            dy0 = Round2Int((double)dy - f.filterw); dy1 = Round2Int((double)dy + f.filterw - 1.0);
            if (dx0 < f.cx0 - x) dx0 = f.cx0 - x;
            if (dx1 > f.cx1 - x - 1) dx1 = f.cx1 - x - 1;
            if (dy0 < f.cy0 - y) dy0 = f.cy0 - y;
            if (dy1 > f.cy1 - y - 1) dy1 = f.cy1 - y - 1;
            fits = dx0 >= -1 && dx1 <= 1 && dy0 >= -1 && dy1 <= 1;
            if (fits) {
                const double x_offs = dx - 0.5, y_offs = dy - 0.5;
                for (int i = dx0; i <= dx1; ++i) xIndex[i + 1] = Floor2Int(fabs(((double)i - x_offs) * f.tableScale));
                for (int i = dy0; i <= dy1; ++i) yIndex[i + 1] = Floor2Int(fabs(((double)i - y_offs) * f.tableScale));
            }
            key = (a.sampleBase + tid) / (uint32_t)f.n_samples;  // pixel slot: equal for the samples of one pixel
        }
        const bool fast = __all_sync(0xffffffffu, !valid || fits) && !f.premult;
        if (!fast) { if (valid) film_splat_sample(a, tid); continue; }
        float v[45];
        uint32_t cover = 0u;
#pragma unroll
        for (int j = -1; j <= 1; ++j)
#pragma unroll
            for (int i = -1; i <= 1; ++i) {
                const int cell = (j + 1) * 3 + (i + 1);
                const bool in = valid && i >= dx0 && i <= dx1 && j >= dy0 && j <= dy1;
                const float wgt = in ? f.filterTable[yIndex[j + 1] * YB_FILTER_TABLE_SIZE + xIndex[i + 1]] : 0.f;
                if (in) cover |= 1u << cell;
                v[5 * cell + 0] = c.x * wgt; v[5 * cell + 1] = c.y * wgt; v[5 * cell + 2] = c.z * wgt; v[5 * cell + 3] = c.w * wgt; v[5 * cell + 4] = wgt;
            }
        // segmented reduction over runs of equal keys (runs are contiguous: samples of a pixel are consecutive)
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t k2 = __shfl_down_sync(0xffffffffu, key, off);
            const uint32_t c2 = __shfl_down_sync(0xffffffffu, cover, off);
            const bool take = (lane + off < 32u) && k2 == key;
            if (take) cover |= c2;
#pragma unroll
            for (int q = 0; q < 45; ++q) { const float t = __shfl_down_sync(0xffffffffu, v[q], off); if (take) v[q] += t; }
        }
        const uint32_t kprev = __shfl_up_sync(0xffffffffu, key, 1);
        if (valid && (lane == 0u || kprev != key)) {
#pragma unroll
            for (int cell = 0; cell < 9; ++cell)
                if (cover & (1u << cell)) {
                    const int i = cell % 3 - 1, j = cell / 3 - 1;
                    float *px = a.film + ((size_t)(x + i - f.cx0) + (size_t)f.w * (size_t)(y + j - f.cy0)) * 5;
                    atomicAdd(px + 0, v[5 * cell + 0]); atomicAdd(px + 1, v[5 * cell + 1]); atomicAdd(px + 2, v[5 * cell + 2]);
                    atomicAdd(px + 3, v[5 * cell + 3]); atomicAdd(px + 4, v[5 * cell + 4]);
                }
        }
    }
}
static void k_film_fast(uint32_t n, const RenderArgs &a, rt_stream s)
{
    if (n == 0) return;
    uint32_t blocks = (n + 255u) / 256u;
    if (blocks > 148u * 32u) blocks = 148u * 32u;
    k_film_fast_kernel<<<blocks, 256, 0, s>>>(n, a);
    YB_CUDA_CHECK(cudaGetLastError());
    g_kernel_launches++;
}
#endif
// pixelGray_t::normalized (image_buffers.h:46-57)
struct DepthResolveArgs { const float *dfilm; float *z; };
YB_KERNEL(k_resolve_depth, DepthResolveArgs)
{
    const float v = a.dfilm[2 * (size_t)tid], w = a.dfilm[2 * (size_t)tid + 1];
    a.z[tid] = w > 0.f ? v / w : 0.f;
}

// imageFilm_t::flush (imagefilm.cc:383-443) + pixel_t::normalized (image_buffers.h:32-43), memoryIO_t layout
struct ResolveArgs { Film f; const float *film; float *rgba; };
YB_KERNEL(k_resolve, ResolveArgs)
{
    const float *px = a.film + (size_t)tid * 5;
    float r = 0.f, g = 0.f, b = 0.f, al = 0.f;
    const float wgt = px[4];
    if (wgt > 0.f) { const float inv = 1.f / wgt; r = px[0] * inv; g = px[1] * inv; b = px[2] * inv; al = px[3] * inv; }
    if (r < 0.f) r = 0.f; if (g < 0.f) g = 0.f; if (b < 0.f) b = 0.f;
    if (a.f.correctGamma) { r = fPow(r, a.f.gamma); g = fPow(g, a.f.gamma); b = fPow(b, a.f.gamma); }
    if (a.f.premult) { r *= al; g *= al; b *= al; }
    float *o = a.rgba + (size_t)tid * 4;
    o[0] = r; o[1] = g; o[2] = b; o[3] = al;
}

// ------------------------------------------------------------------------------------------------ parity / bench hook kernels
struct TraceArgs { SceneDev sc; const Ray *rays; void *out; int shadowDepth; };
struct ClosestOut { int32_t prim; float t, b1, b2; };
struct TShadowOut { int32_t s; float r, g, b; };
YB_KERNEL(k_batch_closest, TraceArgs)
{
    Hit h;
    ClosestOut o;
    if (scene_intersect(a.sc, a.rays[tid], h)) { o.prim = f2i(a.sc.tri_v[3 * h.tri].w); o.t = h.t; o.b1 = h.b1; o.b2 = h.b2; }
    else { o.prim = -1; o.t = 0.f; o.b1 = 0.f; o.b2 = 0.f; }
    ((ClosestOut *)a.out)[tid] = o;
}
YB_KERNEL(k_batch_shadow, TraceArgs) { ((int32_t *)a.out)[tid] = scene_isShadowed(a.sc, a.rays[tid]) ? 1 : 0; }
YB_KERNEL(k_batch_tshadow, TraceArgs)
{
    col3 filt;
    TShadowOut o;
    o.s = scene_isShadowedT(a.sc, a.rays[tid], a.shadowDepth, filt) ? 1 : 0;
    o.r = filt.x; o.g = filt.y; o.b = filt.z;
    ((TShadowOut *)a.out)[tid] = o;
}

// imageFilm_t::nextPass (imagefilm.cc:213-263): flag the pixels whose normalised brightness differs from a neighbour's
// by AA_threshold or more; both pixels of such a pair are resampled in the next pass.  film = raw sums of ALL ranks.
struct FlagArgs { const float *film; unsigned char *flags; int32_t w, h; float thresh; };
YB_D float film_bri(const FlagArgs &a, int x, int y, bool absv)
{
    const float *p = a.film + ((size_t)x + (size_t)a.w * (size_t)y) * 5;
    const float wgt = p[4];
    float r = 0.f, g = 0.f, b = 0.f;
    if (wgt > 0.f) { const float f = (float)(1.0 / (double)wgt); r = p[0] * f; g = p[1] * f; b = p[2] * f; }   // pixel_t::normalized, color.h:329-333
    if (absv) { r = fabsf(r); g = fabsf(g); b = fabsf(b); }
    return 0.2126f * r + 0.7152f * g + 0.0722f * b;                                                         // color.h:79-80
}
YB_KERNEL(k_aa_flags, FlagArgs)
{
    const int x = (int)(tid % (uint32_t)(a.w - 1)), y = (int)(tid / (uint32_t)(a.w - 1));
    const float c = film_bri(a, x, y, true);
    bool need = false;
    if (fabsf(c - film_bri(a, x + 1, y, false)) >= a.thresh) { need = true; a.flags[(size_t)(x + 1) + (size_t)a.w * y] = 1; }
    if (fabsf(c - film_bri(a, x, y + 1, false)) >= a.thresh) { need = true; a.flags[(size_t)x + (size_t)a.w * (y + 1)] = 1; }
    if (fabsf(c - film_bri(a, x + 1, y + 1, false)) >= a.thresh) { need = true; a.flags[(size_t)(x + 1) + (size_t)a.w * (y + 1)] = 1; }
    if (x > 0 && fabsf(c - film_bri(a, x - 1, y + 1, false)) >= a.thresh) { need = true; a.flags[(size_t)(x - 1) + (size_t)a.w * (y + 1)] = 1; }
    if (need) a.flags[(size_t)x + (size_t)a.w * y] = 1;
}

// ================================================================================================ host driver
struct RenderBuffers {
    DevBuf<Task> poolA, poolB;
    DevBuf<Hit> hits;
    DevBuf<ShadowTask> shadow;
    DevBuf<LightJob> ljobs, ljobsSorted;     // staged direct lighting
    DevBuf<uint32_t> counters, order, bins, waveCnt;
#ifndef YB_HOSTEMU
    std::vector<cudaEvent_t> events;     // per-wave stage timers of the device-counted wave loop, read once per batch
    // staged direct lighting: the groups of a wave are independent launches; side streams let the small ones (a few dozen CTAs that
    // would each hold the GPU for the latency of one job) run beside the large ones
    enum { kSideStreams = 4 };
    cudaStream_t side[kSideStreams] = {0, 0, 0, 0};
    cudaEvent_t sideDone[kSideStreams] = {0, 0, 0, 0}, sideFork = 0;
    void sideStreams()
    {
        if (sideFork) return;
        for (int i = 0; i < kSideStreams; ++i) {
            YB_CUDA_CHECK(cudaStreamCreateWithFlags(&side[i], cudaStreamNonBlocking));
            YB_CUDA_CHECK(cudaEventCreateWithFlags(&sideDone[i], cudaEventDisableTiming));
        }
        YB_CUDA_CHECK(cudaEventCreateWithFlags(&sideFork, cudaEventDisableTiming));
    }
    ~RenderBuffers()
    {
        for (size_t i = 0; i < events.size(); ++i) cudaEventDestroy(events[i]);
        for (int i = 0; i < kSideStreams; ++i) { if (side[i]) cudaStreamDestroy(side[i]); if (sideDone[i]) cudaEventDestroy(sideDone[i]); }
        if (sideFork) cudaEventDestroy(sideFork);
    }
    cudaEvent_t event(size_t i)
    {
        while (events.size() <= i) { cudaEvent_t e; YB_CUDA_CHECK(cudaEventCreate(&e)); events.push_back(e); }
        return events[i];
    }
#endif
    DevBuf<unsigned char> keys, aaFlags;
    DevBuf<f4> acc;
    DevBuf<uint32_t> pix;
    DevBuf<float> film, rgba;
    DevBuf<float> depthT, depthFilm, depthZ;   // z_channel
    DevBuf<char> io_in, io_out;     // staging for host-pointer ray batches
    // the pixels a rank owns depend only on the film window, the tile size and (rank, world): kept across frames, on the host
    // and - while `pix` still holds them - on the device (2 M push_backs + 8 MB of H2D per 1080p frame otherwise)
    std::vector<uint32_t> owned;
    int ownedKey[7] = {-1, -1, -1, -1, -1, -1, -1};
    bool ownedOnDevice = false;
};
void render_release(RenderBuffers *rb) { delete rb; }

struct StageTimer {
#ifndef YB_HOSTEMU
    cudaEvent_t e0, e1; rt_stream s;
    StageTimer(rt_stream s_) : s(s_) { cudaEventCreate(&e0); cudaEventCreate(&e1); }
    ~StageTimer() { cudaEventDestroy(e0); cudaEventDestroy(e1); }
    void start() { cudaEventRecord(e0, s); }
    void stop() { cudaEventRecord(e1, s); }
    double ms() { cudaEventSynchronize(e1); float m = 0; cudaEventElapsedTime(&m, e0, e1); return m; }
#else
    std::chrono::steady_clock::time_point t0, t1;
    StageTimer(rt_stream) {}
    void start() { t0 = std::chrono::steady_clock::now(); }
    void stop() { t1 = std::chrono::steady_clock::now(); }
    double ms() { return std::chrono::duration<double, std::milli>(t1 - t0).count(); }
#endif
};

static size_t env_size(const char *name, size_t dflt)
{
    const char *e = getenv(name);
    if (!e || !*e) return dflt;
    return (size_t)strtoull(e, 0, 10);
}

// tuning knob of the traversal kernels' postponed triangle phase (yb_trace_fast.cuh); 1 = never postpone
static uint32_t tri_threshold()
{
    static const uint32_t v = (uint32_t)env_size("YB200_TRI_THRESHOLD", 1);
    return v < 1u ? 1u : (v > 32u ? 32u : v);
}

static void trace_tuning(FastJob &j)
{
    j.surf.tri_v = 0; j.surf.tri_n = 0; j.surf.materials = 0; j.surf.prim_ng = 0; j.shadowDepth = 0;
    static const uint32_t burst = (uint32_t)env_size("YB200_BURST", 6), idle = (uint32_t)env_size("YB200_REFILL_IDLE", 8);
    j.triThreshold = tri_threshold(); j.oneBits = 0x3F800000u;
    j.burst = burst < 1u ? 1u : burst; j.refillIdle = idle < 1u ? 1u : (idle > 32u ? 32u : idle);
}

// tiledIntegrator_t::precalcDepths (integrator.cc:99-130): the range the normalised depth channel maps to [0, 1].  Without a
// far clip the reference shoots one ray per pixel - with row and column swapped, shootRay(i, j) for i < h, j < w, kept here.
static bool precalc_depths(Context &c, Film &f)
{
    const Camera &cam = c.sc.cam;
    float minDepth = 1e38f, maxDepth = 0.f;
    if (cam.farClip > -1) { minDepth = cam.nearClip; maxDepth = cam.farClip; }
    else {
        const int w = cam.resx, h = cam.resy;
        const size_t n = (size_t)w * h;
        std::vector<Ray> rays(n);
        for (int i = 0; i < h; ++i)
            for (int j = 0; j < w; ++j) { float wt; rays[(size_t)i * w + j] = cam_shootRay(cam, (float)i, (float)j, 0.5f, 0.5f, wt); }
        std::vector<ClosestOut> out(n);
        RenderBuffers &rb = *c.rb;
        rb.io_in.reserve(n * sizeof(Ray)); rb.io_out.reserve(n * sizeof(ClosestOut));
        rt_h2d(rb.io_in.p, rays.data(), n * sizeof(Ray), c.stream);
        trace_batch_device(c, (const float *)rb.io_in.p, n, TRACE_CLOSEST, 0, rb.io_out.p);
        rt_d2h(out.data(), rb.io_out.p, n * sizeof(ClosestOut), c.stream);
        for (size_t k = 0; k < n; ++k) {
            const float tmax = out[k].prim >= 0 ? out[k].t : rays[k].tmax;      // scene_t::intersect writes ray.tmax = Z on a hit
            if (tmax > maxDepth) maxDepth = tmax;
            if (tmax < minDepth && tmax >= 0.f) minDepth = tmax;
        }
    }
    if (maxDepth > 0.f) maxDepth = 1.f / (maxDepth - minDepth);
    f.minDepth = minDepth; f.maxDepth = maxDepth;
    return true;
}
float *depth_film_device(Context &c) { return (c.rb && c.sc.film.zChan) ? c.rb->depthFilm.p : 0; }
bool depth_resolve_device(Context &c, const float *d_depthFilm, float *z_host)
{
    if (!c.rb || !d_depthFilm) return c.fail("depth_resolve: no depth film (render with z_channel = true first)");
    const size_t npx = (size_t)c.sc.film.w * c.sc.film.h;
    c.rb->depthZ.reserve(npx + 1);
    DepthResolveArgs da; da.dfilm = d_depthFilm; da.z = c.rb->depthZ.p;
    k_resolve_depth((uint32_t)npx, da, c.stream);
    rt_d2h(z_host, c.rb->depthZ.p, npx * sizeof(float), c.stream);
    return true;
}

// transparent shadows in the persistent kernel: the tables material->getTransparency needs (materials without shader nodes only)
static void trace_tshadow(FastJob &j, const Context &c, int shadowDepth)
{
    j.surf.tri_v = c.sc.tri_v; j.surf.tri_n = c.sc.tri_n; j.surf.materials = c.sc.materials; j.surf.prim_ng = c.sc.prim_ng; j.shadowDepth = shadowDepth;
}

bool render_film_device(Context &c, float *d_film, const float *d_filmAll, int rank, int world, int onlyPass,
                        int (*should_abort)(void *), void (*progress)(void *, int, int), void *user)
{
    if (!c.sceneReady) return c.fail("render: scene not updated");
    if (world < 1 || rank < 0 || rank >= world) return c.fail("render_film: bad rank/world");
    try {
        if (!c.rb) c.rb = new RenderBuffers();
        RenderBuffers &rb = *c.rb;
        rt_stream s = c.stream;
        const Film &f = c.sc.film;
        const Integrator &I = c.sc.integ;
        auto tstart = std::chrono::steady_clock::now();
        // ---- owned pixels: tiles of imageSpliter_t (imagesplitter.cc:29-52) dealt round-robin to ranks
        std::vector<uint32_t> &owned = rb.owned;
        const int ownedKey[7] = {f.w, f.h, f.cx0, f.cy0, f.tileSize, rank, world};
        if (memcmp(ownedKey, rb.ownedKey, sizeof(ownedKey)) != 0) {
            if (f.cx1 > 65535 || f.cy1 > 65535) return c.fail("render: image coordinates above 65535 are not supported");
            owned.clear(); rb.ownedOnDevice = false;
            memcpy(rb.ownedKey, ownedKey, sizeof(ownedKey));
            const int bs = f.tileSize, nx = (f.w + bs - 1) / bs, ny = (f.h + bs - 1) / bs;
            int tile = 0;
            for (int j = 0; j < ny; ++j)
                for (int i = 0; i < nx; ++i, ++tile) {
                    if (tile % world != rank) continue;
                    const int x0 = f.cx0 + i * bs, y0 = f.cy0 + j * bs;
                    const int tw = std::min(bs, f.cx0 + f.w - x0), th = std::min(bs, f.cy0 + f.h - y0);
                    for (int y = y0; y < y0 + th; ++y)
                        for (int x = x0; x < x0 + tw; ++x) owned.push_back((uint32_t)x | ((uint32_t)y << 16));
                }
        }
        // ---- worst-case fan-out per task, from the integrator and the materials present
        bool anyGlossy = false, anyRoughGlass = false, extShade = false, nodeShade = false, allShinyDiffuse = !c.materials.empty(); int specChildren = 0;
        for (size_t i = 0; i < c.materials.size(); ++i) {
            const Material &m = c.materials[i];
            if (m.type != MAT_SHINYDIFFUSE) allShinyDiffuse = false;
            if (m.bsdfFlags & BSDF_GLOSSY) anyGlossy = true;
            if (m.type == MAT_ROUGH_GLASS) anyRoughGlass = true;     // two children per glossy sample
            if (m.type == MAT_ROUGH_GLASS || m.type == MAT_BLEND || m.hasBeer) extShade = true;     // absorbing glass: Beer's law lives in the ext build
            if (m.nNodes > 0) nodeShade = true;
            int sc_ = 0;
            if (m.type == MAT_GLASS || m.type == MAT_BLEND) sc_ = 2;
            else if (m.type == MAT_SHINYDIFFUSE) sc_ = (m.isMirror ? 1 : 0) + (m.isTransparent ? 1 : 0);
            else if (m.type == MAT_COATED_GLOSSY || m.type == MAT_MIRROR) sc_ = 1;
            specChildren = std::max(specChildren, sc_);
        }
        size_t F = (I.type == INTEG_PATH ? (size_t)std::max(1, I.nPaths) : 0) + (anyGlossy ? (anyRoughGlass ? 16 : 8) : 0) + (size_t)specChildren;
        if (F < 1) F = 1;
        size_t S = 0;
        for (size_t i = 0; i < c.lights.size(); ++i) if (c.lights[i].type == LIGHT_MESH) extShade = true;
        for (size_t i = 0; i < c.lights.size(); ++i) S += light_dirac(c.lights[i]) ? 1 : 2 * (size_t)std::max(1, c.lights[i].samples);
        if (I.type == INTEG_DIRECT && I.useAO) S += (size_t)std::max(1, I.aoSamples);
        if (S < 1) S = 1;
        // ---- depth channel: own film of this rank, cleared at the start of a frame; normalisation range first
        Film fz = f;
        if (f.zChan) {
            const size_t npx = (size_t)f.w * f.h;
            rb.depthFilm.reserve(2 * npx + 2);
            if (onlyPass <= 0) rt_memset(rb.depthFilm.p, 0, 2 * npx * sizeof(float), s);
            if (f.zNorm) { if (!precalc_depths(c, fz)) return false; }
        }
        // mcIntegrator_t::createCausticMap in preprocess() (directlight.cc:78-83, pathtracer.cc:90-93): once per frame
        if (onlyPass <= 0 && !build_caustic_map(c)) return false;
        const bool gatherCaustics = c.sc.photonGrid.nPhotons > 0;
        StageTimer tc(s), tsh(s), tsd(s), tf(s);
        uint64_t raysClosest = 0, raysShadow = 0, samplesDone = 0; double msClosest = 0, msShadow = 0, msShade = 0, msFilm = 0;
        const uint64_t launches0 = g_kernel_launches;
        // ---- passes (tiledIntegrator_t::render, integrator.cc:153-164): pass 0 = AA_minsamples on every owned pixel; pass i
        //      = AA_inc_samples on the owned pixels flagged by nextPass from the film of ALL ranks after pass i-1
        const int nPasses = std::max(1, f.AA_passes);
        if (onlyPass >= nPasses) return c.fail("render: pass index beyond AA_passes");
        if (onlyPass < 0 && world > 1 && nPasses > 1)
            return c.fail("render: adaptive passes on several ranks need the film of all ranks between passes - use yb200_render_film_pass");
        for (int pass = (onlyPass < 0 ? 0 : onlyPass); pass < (onlyPass < 0 ? nPasses : onlyPass + 1); ++pass) {
        if (c.abortFlag || (should_abort && should_abort(user))) break;
        std::vector<uint32_t> flagged;
        const std::vector<uint32_t> *pixv = &owned;
        const int passSamples = pass == 0 ? f.n_samples : std::max(1, f.AA_inc_samples);
        const int passOffs = pass == 0 ? 0 : f.n_samples + (pass - 1) * std::max(1, f.AA_inc_samples);
        if (pass > 0 && f.AA_threshold > 0.f) {
            const size_t npx = (size_t)f.w * f.h;
            rb.aaFlags.reserve(npx + 1);
            rt_memset(rb.aaFlags.p, 0, npx, s);
            if (f.w > 1 && f.h > 1) {
                FlagArgs fa; fa.film = d_filmAll ? d_filmAll : d_film; fa.flags = rb.aaFlags.p; fa.w = f.w; fa.h = f.h; fa.thresh = f.AA_threshold;
                k_aa_flags((uint32_t)((size_t)(f.w - 1) * (f.h - 1)), fa, s);
            }
            std::vector<unsigned char> hf(npx);
            rt_d2h(hf.data(), rb.aaFlags.p, npx, s);
            for (size_t i = 0; i < owned.size(); ++i) {
                const int x = (int)(owned[i] & 0xffffu) - f.cx0, y = (int)(owned[i] >> 16) - f.cy0;   // doMoreSamples (imagefilm.cc:445-448)
                if (hf[(size_t)x + (size_t)f.w * y]) flagged.push_back(owned[i]);
            }
            pixv = &flagged;
        }
        const std::vector<uint32_t> &pix = *pixv;
        const uint64_t totalSamples = (uint64_t)pix.size() * (uint64_t)passSamples;
        if (totalSamples >= 0xffffffffull) return c.fail("render: more than 2^32 camera samples per rank are not supported");
        if (totalSamples == 0) continue;
        if (!(pixv == &owned && rb.ownedOnDevice)) {
            rb.pix.reserve(pix.size() + 1);
            rt_h2d(rb.pix.p, pix.data(), pix.size() * sizeof(uint32_t), s);
            rt_sync(s);
            rb.ownedOnDevice = (pixv == &owned);
        }
        // staged direct lighting (yb_direct.cu): where tasks fan out (so that the host counts every wave anyway) and the base build shades
        // (YB200_STAGE_DIRECT = 0: never, 2: also without fan-out - the waves are then counted by the host)
        const size_t stageKnob = env_size("YB200_STAGE_DIRECT", 1);
        const bool stageDirect = (F > 1 || stageKnob == 2) && !extShade && !nodeShade && !c.lights.empty() && stageKnob != 0;
#ifdef YB_HOSTEMU
        const size_t budget = env_size("YB200_QUEUE_BYTES", (size_t)512 << 20), Bmax = env_size("YB200_BATCH_SAMPLES", (size_t)1 << 16);
#else
        // camera samples per batch: 4 M; 8 M with staged direct lighting, whose waves are cut into one launch per job group and want to
        // be longer (materials config, 1080p / 256 spp: 6.2 s per frame at 4 M, 5.9 s at 8 M, no change at 16 M)
        const size_t budget = env_size("YB200_QUEUE_BYTES", (size_t)32 << 30), Bmax = env_size("YB200_BATCH_SAMPLES", stageDirect ? (size_t)1 << 23 : (size_t)1 << 22);
#endif
        const size_t L = stageDirect ? c.lights.size() : 0;       // jobs per task: every light (integrate) or one (path vertex)
        const size_t perSample = (F + 1) * sizeof(Task) * 2 + S * sizeof(ShadowTask) + sizeof(Hit) + sizeof(f4) + 2 * L * sizeof(LightJob);
        size_t B = std::min<size_t>(Bmax, std::max<size_t>(1024, budget / perSample));
        B = std::min<size_t>(B, (size_t)totalSamples);
        const size_t poolCap = B * (F + 1), shadowCap = B * S;
        rb.poolA.reserve(poolCap); rb.poolB.reserve(poolCap);
        rb.hits.reserve(B); rb.shadow.reserve(shadowCap); rb.counters.reserve(YB_CNT_WORDS); rb.acc.reserve(B);
        if (stageDirect) { rb.ljobs.reserve(B * L); rb.ljobsSorted.reserve(B * L); }
        rb.order.reserve(B); rb.keys.reserve(B); rb.bins.reserve(2 * YB_SHADE_BINS);
        if (f.zChan) rb.depthT.reserve(B);
        RenderArgs a;
        memset((void *)&a, 0, sizeof(a));
        a.sc = c.sc; a.sc.film.minDepth = fz.minDepth; a.sc.film.maxDepth = fz.maxDepth; a.sc.film.n_samples = passSamples; a.sc.film.pass_offs = passOffs; a.hits = rb.hits.p; a.shadow = rb.shadow.p; a.counters = rb.counters.p; a.acc = rb.acc.p; a.pix = rb.pix.p;
        a.orderOut = rb.order.p; a.keys = rb.keys.p; a.bins = rb.bins.p;
        const bool binShade = env_size("YB200_BIN_SHADE", 1) != 0;
        const bool binFirst = env_size("YB200_BIN_FIRST", 0) != 0;       // also bin the camera wave (its pixel order is already coherent: off)
        const bool filmFast = env_size("YB200_FILM_FAST", 1) != 0;       // warp-aggregated film splat (k_film_fast_kernel)
        (void)filmFast;                                                  // (unused in the emulator build)
        const bool logWaves = env_size("YB200_LOG_WAVES", 0) != 0;
        const bool kindShade = env_size("YB200_KIND_SHADE", 1) != 0;           // device-counted waves: shade kernel compiled for the wave's kind of task
        const bool sideStreams = env_size("YB200_SIDE_STREAMS", 1) != 0;       // staged direct lighting: groups of a wave on several streams
        (void)sideStreams;
        // coherence sort of the ray waves: 0 off, 1 closest-hit waves after the camera wave, 2 + shadow waves after the first, 3 all
        const int sortRays = (int)env_size("YB200_SORT_RAYS", 0);
        a.depthT = f.zChan ? rb.depthT.p : 0; a.depthFilm = f.zChan ? rb.depthFilm.p : 0;
        a.film = d_film; a.shadowCap = (uint32_t)std::min<size_t>(shadowCap, 0xffffffffu);
        a.ljobs = stageDirect ? rb.ljobs.p : 0; a.ljobsSorted = stageDirect ? rb.ljobsSorted.p : 0; a.ljobCap = (uint32_t)std::min<size_t>(B * L, 0xffffffffu);
        Task *pool = rb.poolA.p, *alt = rb.poolB.p;
        const bool useFast = env_size("YB200_SIMPLE_TRACE", 0) == 0;
        // longest chain of tasks one camera sample can spawn when every task has at most one child: the camera task, one path
        // vertex per bounce, one specular child per recursion level (raylevel is capped at rDepth, mcintegrator.cc:428,588)
        const int maxChain = 1 + (I.type == INTEG_PATH ? std::max(0, I.maxBounces) : 0) + (specChildren ? std::max(0, I.rDepth) + 1 : 0);
        const bool tsFast = useFast && !nodeShade;      // IntersectTS in the persistent kernel (node materials evaluate their graph: per-ray kernel)
        const bool deviceWaves = !stageDirect && F <= 1 && useFast && (!I.trShad || tsFast) && !sortRays && !logWaves && env_size("YB200_DEVICE_WAVES", 1) != 0;
        const int nBatches = (int)((totalSamples + B - 1) / B);
        for (int b = 0; b < nBatches; ++b) {
            if (c.abortFlag || (should_abort && should_abort(user))) break;
            const uint32_t base = (uint32_t)((uint64_t)b * B);
            const uint32_t nb = (uint32_t)std::min<uint64_t>(B, totalSamples - base);
            a.sampleBase = base; a.tasks = pool;
            k_raygen(nb, a, s);
            size_t top = nb;
            bool firstWave = true;      // camera rays come in pixel order: already coherent
            bool firstShadowWave = true;
            // ---- device-counted waves.  With a fan-out of at most one child per task the pool never outgrows the batch, every
            //      wave consumes the whole pool and ping-pongs to the other one - so the host does not need the counts to plan
            //      the next wave: each kernel reads its task / ray count from the counter block the previous kernel filled, and
            //      the host reads all counters and stage timers ONCE per batch instead of blocking three times per wave
            //      (160 waves per frame of the benchmark scene).
            if (deviceWaves) {
                uint32_t pending = nb;
                bool firstWaveOfBatch = true;      // the chain of waves below starts with the camera wave only in its first round
                while (pending > 0) {
                    const size_t nSlots = (size_t)maxChain + 1;
                    rb.waveCnt.reserve(4 * nSlots);
                    rt_memset(rb.waveCnt.p, 0, 4 * nSlots * sizeof(uint32_t), s);
                    rt_h2d(rb.waveCnt.p, &pending, sizeof(uint32_t), s);          // slot 0: input count of the first wave
                    for (int w = 0; w < maxChain; ++w) {
                        a.tasks = pool; a.out_tasks = alt; a.outCap = (uint32_t)std::min<size_t>(poolCap, 0xffffffffu);
                        a.nPtr = rb.waveCnt.p + 4 * (size_t)w;
                        a.counters = rb.waveCnt.p + 4 * (size_t)(w + 1);
#ifndef YB_HOSTEMU
                        YB_CUDA_CHECK(cudaEventRecord(rb.event(4 * (size_t)w + 0), s));
#endif
                        {
                            FastJob j; j.nodes = c.sc.nodes; j.tri_v = c.sc.tri_v; j.rays = (const char *)a.tasks; j.stride = sizeof(Task);
                            j.n = pending; j.nPtr = a.nPtr; j.order = 0; j.counter = a.counters + 2; j.out = a.hits; j.acc = a.acc;
                            j.nNodes = c.sc.nNodes; trace_tuning(j); j.counterZeroed = true;
                            launch_trace_fast<false, OUT_HITS>(j, s);
                        }
#ifndef YB_HOSTEMU
                        YB_CUDA_CHECK(cudaEventRecord(rb.event(4 * (size_t)w + 1), s));
#endif
                        a.order = 0;
                        if (binShade && (!firstWave || binFirst)) { bin_tasks(pending, a, s); a.order = rb.order.p; }
                        // without fan-out a wave holds one kind of task: the camera wave integrate() calls, the later ones path vertices
                        // (pathtracing) or the integrate() calls of the one specular child (directlighting)
                        const uint32_t waveKind = (firstWaveOfBatch && w == 0) ? (uint32_t)TASK_INTEGRATE : (I.type == INTEG_PATH ? (uint32_t)TASK_PATH : (uint32_t)TASK_INTEGRATE);
                        if (nodeShade) launch_shade_nodes(pending, a, s);
                        else if (extShade) launch_shade_ext(pending, a, s);
                        else if (kindShade && allShinyDiffuse) launch_shade_sd_kind(waveKind, pending, a, s);
                        else if (kindShade) launch_shade_base_kind(waveKind, pending, a, s);
                        else launch_shade_base(pending, a, s);
                        if (gatherCaustics) launch_caustic_gather(pending, a, s);
#ifndef YB_HOSTEMU
                        YB_CUDA_CHECK(cudaEventRecord(rb.event(4 * (size_t)w + 2), s));
#endif
                        firstWave = false;
                        {
                            FastJob j; j.nodes = c.sc.nodes; j.tri_v = c.sc.tri_v; j.rays = (const char *)a.shadow; j.stride = sizeof(ShadowTask);
                            j.n = (uint32_t)std::min<size_t>((size_t)pending * S, 0xffffffffu); j.nPtr = a.counters + 1; j.order = 0;
                            j.counter = a.counters + 3; j.out = 0; j.acc = a.acc; j.nNodes = c.sc.nNodes; trace_tuning(j); j.counterZeroed = true;
                            if (I.trShad) { trace_tshadow(j, c, I.sDepth); launch_trace_fast<true, OUT_TSHADOW_ACC>(j, s); }
                            else launch_trace_fast<true, OUT_SHADOW_ACC>(j, s);
                        }
#ifndef YB_HOSTEMU
                        YB_CUDA_CHECK(cudaEventRecord(rb.event(4 * (size_t)w + 3), s));
#endif
                        std::swap(pool, alt);
                    }
                    std::vector<uint32_t> hc(4 * nSlots);
                    rt_d2h(hc.data(), rb.waveCnt.p, hc.size() * sizeof(uint32_t), s);      // waits for the waves above
                    for (int w = 0; w < maxChain; ++w) {
                        const uint32_t nIn = hc[4 * (size_t)w], nChildren = hc[4 * (size_t)(w + 1)], nShadow = hc[4 * (size_t)(w + 1) + 1];
                        if (nChildren > poolCap || nShadow > shadowCap || nChildren > nIn) throw RtError("render: internal queue overflow");
                        raysClosest += nIn; raysShadow += nShadow;
#ifndef YB_HOSTEMU
                        if (nIn) {
                            float m = 0.f;
                            cudaEventElapsedTime(&m, rb.events[4 * (size_t)w + 0], rb.events[4 * (size_t)w + 1]); msClosest += m;
                            cudaEventElapsedTime(&m, rb.events[4 * (size_t)w + 1], rb.events[4 * (size_t)w + 2]); msShade += m;
                            if (nShadow) { cudaEventElapsedTime(&m, rb.events[4 * (size_t)w + 2], rb.events[4 * (size_t)w + 3]); msShadow += m; }
                        }
#endif
                    }
                    pending = hc[4 * (size_t)maxChain];      // tasks still alive after maxChain waves (none unless the bound below is too low)
                    firstWaveOfBatch = false;
                }
                top = 0;
                a.nPtr = 0; a.counters = rb.counters.p;
            }
            while (top > 0) {
                size_t M = std::min(top, B);
                M = std::min(M, shadowCap / S);
                if (M * F > poolCap - top) M = (poolCap - top) / F;
                if (M == 0) throw RtError("render: task pool exhausted (raise YB200_QUEUE_BYTES)");
                const bool whole = (M == top);
                a.tasks = pool + (top - M);
                a.out_tasks = whole ? alt : pool + top;
                a.outCap = (uint32_t)std::min<size_t>(whole ? poolCap : poolCap - top, 0xffffffffu);
                rt_memset(rb.counters.p, 0, YB_CNT_WORDS * sizeof(uint32_t), s);
                tc.start();
                if (useFast) {
                    FastJob j; j.nodes = c.sc.nodes; j.tri_v = c.sc.tri_v; j.rays = (const char *)a.tasks; j.stride = sizeof(Task); j.n = (uint32_t)M; j.nPtr = 0; j.counterZeroed = false;
                    j.order = (sortRays && !firstWave) ? sort_rays((const char *)a.tasks, sizeof(Task), (uint32_t)M, c.sc.boundLo, c.sc.boundHi, c.raySort, s) : 0;
                    j.counter = rb.counters.p + 2; j.out = a.hits; j.acc = a.acc; j.nNodes = c.sc.nNodes; trace_tuning(j);
                    launch_trace_fast<false, OUT_HITS>(j, s);
                } else
                    k_trace_closest((uint32_t)M, a, s);
                tc.stop();
                tsd.start();
                a.order = 0;
                if (binShade && (!firstWave || binFirst)) { bin_tasks((uint32_t)M, a, s); a.order = rb.order.p; }
                if (nodeShade) launch_shade_nodes((uint32_t)M, a, s);
                else if (extShade) launch_shade_ext((uint32_t)M, a, s);
                else if (stageDirect) launch_shade_staged((uint32_t)M, a, s);
                else launch_shade_base((uint32_t)M, a, s);
                if (gatherCaustics) launch_caustic_gather((uint32_t)M, a, s);
                uint32_t cnt[YB_CNT_WORDS];
                if (stageDirect) {      // the queued estimates, one launch per (light type x material class) group present in this wave
                    group_light_jobs(M * L, a, s);
                    rt_d2h(cnt, rb.counters.p, sizeof(cnt), s);
                    if (cnt[YB_CNT_LJOBS] > a.ljobCap) throw RtError("render: internal queue overflow (light jobs)");
                    uint32_t gbase[YB_LGROUPS], gcount[YB_LGROUPS], base = 0;
                    for (uint32_t k = 0; k < YB_LGROUPS; ++k) {
                        gbase[k] = base; gcount[k] = 0;
                        for (uint32_t q = 0; q < 4; ++q) gcount[k] += cnt[YB_CNT_LBINS + 4 * k + q];
                        base += gcount[k];
                    }
#ifndef YB_HOSTEMU
                    if (sideStreams) {
                        // largest first (a sampled light costs about eight point lights per job), dealt round-robin over the main stream
                        // and the side streams; the shadow queue takes the pushes of all of them through one atomic counter
                        std::vector<std::pair<uint64_t, uint32_t>> groups;
                        for (uint32_t k = 0; k < YB_LGROUPS; ++k)
                            if (gcount[k]) {
                                const uint32_t lt = (k >> 1) / YB_LJOB_MCLASSES;
                                groups.push_back(std::make_pair((uint64_t)gcount[k] * ((lt == LIGHT_POINT || lt == LIGHT_DIRECTIONAL) ? 1u : 8u), k));
                            }
                        std::sort(groups.begin(), groups.end(), [](const std::pair<uint64_t, uint32_t> &x, const std::pair<uint64_t, uint32_t> &y) { return x.first > y.first; });
                        rb.sideStreams();
                        YB_CUDA_CHECK(cudaEventRecord(rb.sideFork, s));
                        bool used[RenderBuffers::kSideStreams] = {false, false, false, false};
                        for (size_t g = 0; g < groups.size(); ++g) {
                            const uint32_t k = groups[g].second;
                            const size_t lane = g % (RenderBuffers::kSideStreams + 1);
                            rt_stream st = s;
                            if (lane > 0) {
                                st = rb.side[lane - 1];
                                if (!used[lane - 1]) { YB_CUDA_CHECK(cudaStreamWaitEvent(st, rb.sideFork, 0)); used[lane - 1] = true; }
                            }
                            launch_direct(k, gbase[k], gcount[k], a, st);
                        }
                        for (int i = 0; i < RenderBuffers::kSideStreams; ++i)
                            if (used[i]) { YB_CUDA_CHECK(cudaEventRecord(rb.sideDone[i], rb.side[i])); YB_CUDA_CHECK(cudaStreamWaitEvent(s, rb.sideDone[i], 0)); }
                    } else
#endif
                    for (uint32_t k = 0; k < YB_LGROUPS; ++k) launch_direct(k, gbase[k], gcount[k], a, s);
                }
                tsd.stop();
                firstWave = false;
                rt_d2h(cnt, rb.counters.p, 4 * sizeof(uint32_t), s);
                if (cnt[0] > a.outCap || cnt[1] > a.shadowCap) throw RtError("render: internal queue overflow");
                tsh.start();
                if (useFast && !I.trShad) {
                    FastJob j; j.nodes = c.sc.nodes; j.tri_v = c.sc.tri_v; j.rays = (const char *)a.shadow; j.stride = sizeof(ShadowTask); j.n = cnt[1]; j.nPtr = 0; j.counterZeroed = false;
                    j.order = (sortRays >= 2 && (!firstShadowWave || sortRays >= 3)) ? sort_rays((const char *)a.shadow, sizeof(ShadowTask), cnt[1], c.sc.boundLo, c.sc.boundHi, c.raySort, s) : 0;
                    j.counter = rb.counters.p + 3; j.out = 0; j.acc = a.acc; j.nNodes = c.sc.nNodes; trace_tuning(j);
                    launch_trace_fast<true, OUT_SHADOW_ACC>(j, s);
                } else if (tsFast && I.trShad) {
                    FastJob j; j.nodes = c.sc.nodes; j.tri_v = c.sc.tri_v; j.rays = (const char *)a.shadow; j.stride = sizeof(ShadowTask); j.n = cnt[1]; j.nPtr = 0; j.counterZeroed = false;
                    j.order = 0; j.counter = rb.counters.p + 3; j.out = 0; j.acc = a.acc; j.nNodes = c.sc.nNodes; trace_tuning(j); trace_tshadow(j, c, I.sDepth);
                    launch_trace_fast<true, OUT_TSHADOW_ACC>(j, s);
                } else
                    k_trace_shadow(cnt[1], a, s);
                tsh.stop();
                firstShadowWave = false;
                raysClosest += M; raysShadow += cnt[1];
                if (logWaves) fprintf(stderr, "yb200 wave: batch %d closest %zu shadow %u children %u\n", b, M, cnt[1], cnt[0]);
                msClosest += tc.ms(); msShade += tsd.ms(); if (cnt[1]) msShadow += tsh.ms();
                if (whole) { std::swap(pool, alt); top = cnt[0]; }
                else {
                    const size_t k = std::min<size_t>(M, cnt[0]);
                    rt_d2d(pool + (top - M), pool + (top + cnt[0] - k), k * sizeof(Task), s);
                    top = top - M + cnt[0];
                }
            }
            a.tasks = pool;
            tf.start();
#ifndef YB_HOSTEMU
            if (filmFast && !a.depthT) k_film_fast(nb, a, s); else
#endif
            k_film(nb, a, s);
            tf.stop();
            rt_sync(s);
            msFilm += tf.ms();
            if (progress) progress(user, pass * nBatches + b + 1, nPasses * nBatches);
        }
        samplesDone += totalSamples;
        }   // passes
        rt_sync(s);
        Stats &st = c.stats;
        st.rays_closest = raysClosest; st.rays_shadow = raysShadow; st.camera_samples = samplesDone;
        st.kernel_launches = g_kernel_launches - launches0;
        st.trace_closest_ms = msClosest; st.trace_shadow_ms = msShadow; st.shade_ms = msShade; st.film_ms = msFilm;
        st.render_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tstart).count();
    } catch (const RtError &e) { return c.fail(std::string("render failed: ") + e.what()); }
    return true;
}

// ------------------------------------------------------------------------------------------------ several devices, one context
struct AddArgs { float *dst; const float *src; };
YB_KERNEL(k_film_add, AddArgs) { a.dst[tid] += a.src[tid]; }
void film_add_device(float *d_dst, const float *d_src, size_t n, rt_stream s)
{
    AddArgs a; a.dst = d_dst; a.src = d_src;
    for (size_t o = 0; o < n; o += (size_t)1 << 30) { AddArgs b = a; b.dst += o; b.src += o; k_film_add((uint32_t)std::min<size_t>(n - o, (size_t)1 << 30), b, s); }
}

}  // namespace yb
#ifndef YB_HOSTEMU
#include <thread>
#include <atomic>
#endif
namespace yb {

#ifndef YB_HOSTEMU
static int group_should_abort(void *flag) { return ((std::atomic<int> *)flag)->load() != 0; }
// dst (on the leader's device, leader's stream) += src (on device srcDev) through the staging buffer tmp
static void film_add_peer(Context &L, float *dst, const float *src, int srcDev, float *tmp, size_t n)
{
    YB_CUDA_CHECK(cudaMemcpyPeerAsync(tmp, L.device, src, srcDev, n * sizeof(float), L.stream));
    film_add_device(dst, tmp, n, L.stream);
}
#endif

// The tiles of the frame are dealt round-robin over the devices (rank = position in yb200_set_devices' list), every device runs
// the same wavefront driver on its own stream from its own host thread, and the films - raw sums, so the reduction is a plain
// addition - meet on the leader's device by peer copies.  Adaptive passes need the sum of all films between
// passes (imageFilm_t::nextPass looks at neighbouring pixels): it is reduced and handed back to every device after each pass.
bool render_group(Context &c, float *rgba_host, int (*should_abort)(void *), void (*progress)(void *, int, int), void *user)
{
#ifdef YB_HOSTEMU
    (void)rgba_host; (void)should_abort; (void)progress; (void)user;
    return c.fail("several devices are not part of the host emulator");
#else
    const int world = 1 + (int)c.helpers.size();
    if (!c.ensureDevice()) return false;
    const Film &f = c.sc.film;
    const size_t npx = (size_t)f.w * f.h, nf = npx * 5;
    const int nPasses = std::max(1, f.AA_passes);
    try {
        auto t0 = std::chrono::steady_clock::now();
        // ---- 1. the scene reaches the other devices (peer copies of the leader's buffers), all helpers at once
        {
            std::vector<std::thread> th;
            std::vector<int> ok(c.helpers.size(), 0);
            for (size_t i = 0; i < c.helpers.size(); ++i)
                th.emplace_back([&, i]() { ok[i] = c.helpers[i]->adoptScene(c) ? 1 : 0; });
            for (size_t i = 0; i < th.size(); ++i) th[i].join();
            for (size_t i = 0; i < ok.size(); ++i) if (!ok[i]) return c.fail("device " + std::to_string(c.helpers[i]->device) + ": " + c.helpers[i]->err);
        }
        const double adoptMs = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        YB_CUDA_CHECK(cudaSetDevice(c.device));
        for (size_t i = 0; i < c.helpers.size(); ++i) {      // the films come back over NVLink as well
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, c.device, c.helpers[i]->device) == cudaSuccess && can) cudaDeviceEnablePeerAccess(c.helpers[i]->device, 0);
            cudaGetLastError();
        }
        c.groupFilm.reserve(nf); c.groupFilmAll.reserve(nf); c.groupTmp.reserve(std::max(nf, 2 * npx));
        rt_memset(c.groupFilm.p, 0, nf * sizeof(float), c.stream);
        std::atomic<int> abortAll(0);
        Stats total; memset(&total, 0, sizeof(total));
        for (int pass = 0; pass < nPasses; ++pass) {
            if (c.abortFlag || (should_abort && should_abort(user))) abortAll = 1;
            if (abortAll.load()) break;
            // ---- 2. every device renders its tiles of this pass
            std::vector<std::thread> th;
            std::vector<int> ok(c.helpers.size(), 0);
            for (size_t i = 0; i < c.helpers.size(); ++i)
                th.emplace_back([&, i, pass]() {
                    Context &h = *c.helpers[i];
                    if (cudaSetDevice(h.device) != cudaSuccess) { h.fail("cudaSetDevice failed"); return; }
                    try {
                        h.groupFilm.reserve(nf);
                        if (nPasses > 1) h.groupFilmAll.reserve(nf);      // allocated here: this thread has the helper's device current
                        if (pass == 0) { rt_memset(h.groupFilm.p, 0, nf * sizeof(float), h.stream); h.abortFlag = 0; }
                        ok[i] = render_film_device(h, h.groupFilm.p, pass > 0 ? h.groupFilmAll.p : 0, (int)i + 1, world, nPasses > 1 ? pass : -1,
                                                   group_should_abort, 0, &abortAll) ? 1 : 0;
                    } catch (const std::exception &e) { h.fail(e.what()); }
                });
            struct LeaderCb { int (*fn)(void *); void *user; std::atomic<int> *all; };
            LeaderCb lcb = {should_abort, user, &abortAll};
            const bool okL = render_film_device(c, c.groupFilm.p, pass > 0 ? c.groupFilmAll.p : 0, 0, world, nPasses > 1 ? pass : -1,
                                                [](void *u) -> int {
                                                    LeaderCb *l = (LeaderCb *)u;
                                                    if (l->fn && l->fn(l->user)) l->all->store(1);
                                                    return l->all->load();
                                                }, progress ? progress : 0, &lcb);
            for (size_t i = 0; i < th.size(); ++i) th[i].join();
            if (!okL) return false;
            for (size_t i = 0; i < ok.size(); ++i) if (!ok[i]) return c.fail("device " + std::to_string(c.helpers[i]->device) + ": " + c.helpers[i]->err);
            YB_CUDA_CHECK(cudaSetDevice(c.device));
            // counters of this pass
            total.rays_closest += c.stats.rays_closest; total.rays_shadow += c.stats.rays_shadow; total.camera_samples += c.stats.camera_samples;
            total.kernel_launches += c.stats.kernel_launches;
            double ms[5] = {c.stats.trace_closest_ms, c.stats.trace_shadow_ms, c.stats.shade_ms, c.stats.film_ms, c.stats.render_ms};
            for (size_t i = 0; i < c.helpers.size(); ++i) {
                const Stats &hs = c.helpers[i]->stats;
                total.rays_closest += hs.rays_closest; total.rays_shadow += hs.rays_shadow; total.camera_samples += hs.camera_samples;
                total.kernel_launches += hs.kernel_launches;
                ms[0] = std::max(ms[0], hs.trace_closest_ms); ms[1] = std::max(ms[1], hs.trace_shadow_ms); ms[2] = std::max(ms[2], hs.shade_ms);
                ms[3] = std::max(ms[3], hs.film_ms); ms[4] = std::max(ms[4], hs.render_ms);
            }
            total.trace_closest_ms += ms[0]; total.trace_shadow_ms += ms[1]; total.shade_ms += ms[2]; total.film_ms += ms[3];
            // ---- 3. sum of the films on the leader's device
            rt_d2d(c.groupFilmAll.p, c.groupFilm.p, nf * sizeof(float), c.stream);
            for (size_t i = 0; i < c.helpers.size(); ++i) film_add_peer(c, c.groupFilmAll.p, c.helpers[i]->groupFilm.p, c.helpers[i]->device, c.groupTmp.p, nf);
            rt_sync(c.stream);
            if (pass + 1 < nPasses)       // every device needs the whole picture to flag its pixels for the next pass
                for (size_t i = 0; i < c.helpers.size(); ++i) {
                    Context &h = *c.helpers[i];
                    YB_CUDA_CHECK(cudaMemcpyPeerAsync(h.groupFilmAll.p, h.device, c.groupFilmAll.p, c.device, nf * sizeof(float), c.stream));
                }
            rt_sync(c.stream);
        }
        // depth channel: the raw depth films add up like the colour film
        if (f.zChan && depth_film_device(c))
            for (size_t i = 0; i < c.helpers.size(); ++i)
                if (depth_film_device(*c.helpers[i])) film_add_peer(c, depth_film_device(c), depth_film_device(*c.helpers[i]), c.helpers[i]->device, c.groupTmp.p, 2 * npx);
        // ---- 4. imageFilm_t::flush on the leader
        c.groupTmp.reserve(std::max(nf, 4 * npx));
        film_resolve_device(c, c.groupFilmAll.p, c.groupTmp.p);
        rt_d2h(rgba_host, c.groupTmp.p, npx * 4 * sizeof(float), c.stream);
        const double build_ms = c.stats.build_ms;
        const uint64_t nt = c.stats.n_triangles, nn = c.stats.n_bvh_nodes, bb = c.stats.bvh_bytes, tb = c.stats.tri_bytes;
        c.stats = total;
        c.stats.build_ms = build_ms + adoptMs; c.stats.n_triangles = nt; c.stats.n_bvh_nodes = nn; c.stats.bvh_bytes = bb; c.stats.tri_bytes = tb;
        c.stats.render_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    } catch (const RtError &e) { return c.fail(std::string("multi-device render failed: ") + e.what()); }
    return true;
#endif
}

void film_resolve_device(Context &c, const float *d_film, float *d_rgba)
{
    ResolveArgs ra; ra.f = c.sc.film; ra.film = d_film; ra.rgba = d_rgba;
    k_resolve((uint32_t)((size_t)c.sc.film.w * c.sc.film.h), ra, c.stream);
}

void trace_batch_device(Context &c, const float *d_rays, size_t n, int mode, int shadow_depth, void *d_out)
{
    TraceArgs ta; ta.sc = c.sc; ta.rays = (const Ray *)d_rays; ta.out = d_out; ta.shadowDepth = shadow_depth;
    const size_t step = (size_t)1 << 30;
    const bool useFast = env_size("YB200_SIMPLE_TRACE", 0) == 0;
    for (size_t o = 0; o < n; o += step) {
        const uint32_t m = (uint32_t)std::min(step, n - o);
        TraceArgs t2 = ta; t2.rays = ta.rays + o;
        bool anyNodes = false;
        for (size_t i = 0; i < c.materials.size(); ++i) if (c.materials[i].nNodes > 0) anyNodes = true;
        if (useFast && (mode != TRACE_TSHADOW || !anyNodes)) {
            c.d_traceCounter.reserve(4);
            FastJob j; j.nodes = c.sc.nodes; j.tri_v = c.sc.tri_v; j.rays = (const char *)(ta.rays + o); j.stride = sizeof(Ray); j.n = m;
            j.order = 0; j.nPtr = 0; j.counterZeroed = false; j.counter = c.d_traceCounter.p; j.acc = 0; j.nNodes = c.sc.nNodes; trace_tuning(j);
            if (mode == TRACE_CLOSEST) { j.out = (ClosestOut *)d_out + o; launch_trace_fast<false, OUT_CLOSEST_PRIM>(j, c.stream); }
            else if (mode == TRACE_SHADOW) { j.out = (int32_t *)d_out + o; launch_trace_fast<true, OUT_SHADOW_FLAG>(j, c.stream); }
            else { j.out = (TShadowOut *)d_out + o; trace_tshadow(j, c, shadow_depth); launch_trace_fast<true, OUT_TSHADOW_RES>(j, c.stream); }
            continue;
        }
        if (mode == TRACE_CLOSEST) { t2.out = (ClosestOut *)d_out + o; k_batch_closest(m, t2, c.stream); }
        else if (mode == TRACE_SHADOW) { t2.out = (int32_t *)d_out + o; k_batch_shadow(m, t2, c.stream); }
        else { t2.out = (TShadowOut *)d_out + o; k_batch_tshadow(m, t2, c.stream); }
    }
}

}  // namespace yb
