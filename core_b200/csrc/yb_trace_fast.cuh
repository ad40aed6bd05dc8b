// yb_trace_fast.cuh — the production traversal kernels (sm_100a only; the per-ray functions of yb_trace.cuh
// stay as the readable statement of the same algorithm for the host emulator and the transparent-shadow mode).
//
// Design (B200 has no RT cores; this path is instruction-issue and latency bound, see profiles/):
//  * persistent warps: grid = SMs x resident blocks; a warp pulls rays from a global counter with one
//    warp-aggregated atomic and REFILLS lanes whose ray has terminated once enough lanes are idle, so lanes
//    do not idle behind the longest ray of the warp (ballot/popc compaction of the idle lanes);
//  * 8-wide quantised nodes fetched with five 128-bit loads; child planes are turned into floats with one PRMT
//    each (byte -> mantissa of 1.xxx) and one FFMA; the 32-bit hit mask carries internal children (bits 24..31,
//    ordered by the ray octant) and triangles (bits 0..23) as in Ylitie et al. 2017;
//  * traversal stack: the first YB_SSTACK entries of every lane live in shared memory ([entry][thread], conflict
//    free), the rest spill to local memory;
//  * triangles: three 128-bit loads, the reference's Moeller-Trumbore operation order without FMA contraction.
#pragma once
#include "yb_trace.cuh"
#include <stdio.h>

#ifdef YB_HOSTEMU
// ---- emulator shims: the kernel body below runs as ONE lane that owns a whole (1-wide) warp, so that the very same
//      traversal code can be checked against the golden kd-tree results on a machine without a GPU (tests/emu).
#include <cmath>
namespace yb { namespace emu {
struct uint2_ { uint32_t x, y; };
struct uint4_ { uint32_t x, y, z, w; };
struct float4_ { float x, y, z, w; };
struct int4_ { int x, y, z, w; };
struct tidx_ { unsigned x; };
inline uint32_t byte_perm(uint32_t x, uint32_t y, uint32_t s)
{
    uint64_t src = ((uint64_t)y << 32) | x;
    uint32_t r = 0;
    for (int i = 0; i < 4; ++i) {
        uint32_t sel = (s >> (4 * i)) & 0xf;
        uint32_t b = (uint32_t)(src >> (8 * (sel & 7))) & 0xff;
        if (sel & 8) b = (b & 0x80) ? 0xff : 0x00;
        r |= b << (8 * i);
    }
    return r;
}
} }
#define uint2 yb::emu::uint2_
#define uint4 yb::emu::uint4_
#define float4 yb::emu::float4_
#define int4 yb::emu::int4_
#define make_uint2(a, b) (yb::emu::uint2_{(uint32_t)(a), (uint32_t)(b)})
#define make_float4(a, b, c, d) (yb::emu::float4_{(a), (b), (c), (d)})
#define make_int4(a, b, c, d) (yb::emu::int4_{(a), (b), (c), (d)})
#define __global__
#define __shared__ static
#define __device__
#define __forceinline__ inline
#define __launch_bounds__(a, b)
#define __syncwarp()
#define __ballot_sync(m, p) ((p) ? 1u : 0u)
#define __shfl_sync(m, v, l) (v)
#define __popc(x) __builtin_popcount(x)
#define __ffs(x) __builtin_ffs((int)(x))
#define __clz(x) ((x) ? __builtin_clz((unsigned)(x)) : 32)
#define __ldg(p) (*(p))
#define __byte_perm(x, y, s) yb::emu::byte_perm((x), (y), (s))
#define __uint_as_float(u) yb::i2f((int32_t)(u))
#define __float_as_uint(f) ((uint32_t)yb::f2i(f))
#define __int_as_float(i) yb::i2f(i)
#define __float_as_int(f) yb::f2i(f)
#define atomicAdd(p, v) yb::emu_fetch_add((p), (v))
#define YB_FULLMASK 1u
#define YB_TB 1
static yb::emu::tidx_ threadIdx = {0};
#else
#define YB_FULLMASK 0xffffffffu
#define YB_TB 128            // threads per block
#endif

namespace yb {

#ifndef YB_SSTACK
#define YB_SSTACK 8          // shared-memory stack entries per lane
#endif
#define YB_LSTACK (YB_STACK_SIZE - YB_SSTACK)
#define YB_REFILL_IDLE 8     // refill when at least this many lanes are idle (or all are)
#ifndef YB_TRACE_CTAS
#define YB_TRACE_CTAS 8      // resident CTAs per SM the kernel is compiled for (8 x 128 threads: 64 registers, no spills)
#endif

#if defined(YB_HOSTEMU) && defined(YB_TRACE_STATS)
static uint64_t g_stat_nodes = 0, g_stat_tris = 0, g_stat_rays = 0;
struct StatPrinter { ~StatPrinter() { if (g_stat_rays) fprintf(stderr, "TRACE STATS: rays %llu nodes/ray %.2f tris/ray %.2f\n", (unsigned long long)g_stat_rays, (double)g_stat_nodes / g_stat_rays, (double)g_stat_tris / g_stat_rays); } };
static StatPrinter g_stat_printer;
#endif
// OUT_TSHADOW_*: scene_t::isShadowed(state, ray, maxDepth, filt) -> IntersectTS (scene.cc:904-928, kdtree.cc:953-1108): an opaque
// hit shadows the ray; a transparent one multiplies filt by material->getTransparency, more than shadowDepth of them shadow it too.
// A BVH holds every triangle in exactly one leaf, so the std::set the kd-tree needs against double counting is not needed.
enum { OUT_HITS = 0, OUT_SHADOW_ACC = 1, OUT_SHADOW_FLAG = 2, OUT_CLOSEST_PRIM = 3, OUT_TSHADOW_ACC = 4, OUT_TSHADOW_RES = 5 };

struct FastJob {
    const BvhNode *nodes;
    const f4 *tri_v;
    const char *rays;        // ray records {from, tmin, dir, tmax} at offset 0 of each `stride`-byte record
    uint32_t stride, n;      // n: number of rays, or - with nPtr - an upper bound that only sizes the grid
    const uint32_t *nPtr;    // optional: the ray count lives in device memory (written by the kernel that filled the queue)
    const uint32_t *order;   // optional: fetch order of the rays (coherence sort); results still go to the ray's own slot
    uint32_t *counter;       // ray-fetch counter; cleared on the stream by the launcher unless counterZeroed
    bool counterZeroed;
    void *out;               // OUT_HITS: Hit[n]; OUT_SHADOW_FLAG: int32[n]; OUT_CLOSEST_PRIM: {prim,t,b1,b2}[n]
    f4 *acc;                 // OUT_SHADOW_ACC: ShadowTask{ray, c, sid}: acc[sid].rgb += c when unoccluded
    uint32_t nNodes;
    uint32_t oneBits;        // 0x3F800000 (see plane_f)
    uint32_t burst, refillIdle;   // traversal steps between refill checks; idle lanes that trigger a refill
    uint32_t triThreshold;   // triangle phase runs when this many lanes (or as many as still have node work) hold a triangle group
    // transparent shadows only (OUT_TSHADOW_*): what material->getTransparency needs at the hit
    SurfTables surf; int32_t shadowDepth;
};

// replicate the msb of every byte over the byte (prmt sign mode; __byte_perm only honours 3 selector bits per nibble)
__device__ __forceinline__ uint32_t sign_extend_s8x4(uint32_t x)
{
#ifdef YB_HOSTEMU
    return yb::emu::byte_perm(x, 0u, 0xba98u);
#else
    uint32_t r;
    asm("prmt.b32 %0, %1, 0, 0x0000ba98;" : "=r"(r) : "r"(x));
    return r;
#endif
}
// byte j of w into the top mantissa byte of 1.0f.  `one` (= 0x3F800000) arrives as a kernel parameter rather than a
// literal so that it lives in ONE register and the selector becomes the PRMT immediate; with the literal, ptxas
// encodes 0x3F800000 as the immediate and re-materialises the selector in a register before each of the 48 PRMTs.
__device__ __forceinline__ float plane_f(uint32_t w, uint32_t one, uint32_t sel) { return __uint_as_float(__byte_perm(w, one, sel)); }

// any-hit rays keep the old rule (no node step while triangles are pending): the next triangle may end the ray, a node visited
// meanwhile is wasted (10 M-triangle soup, AO rays: 128 -> 127 ms; benchmark scene: no difference)
#ifndef ANYHIT_WAITS
#define ANYHIT_WAITS 1
#endif
template <bool ANYHIT, int OUT, bool SORTED>
__global__ void __launch_bounds__(YB_TB, YB_TRACE_CTAS) k_trace_fast(const FastJob job)
{
    __shared__ uint2 sstack[YB_SSTACK][YB_TB];
    uint2 lstack[YB_LSTACK];
    const unsigned lane = threadIdx.x & 31u;   // 0 in the emulator
    const unsigned FULL = YB_FULLMASK;
    const uint32_t one = job.oneBits;
    const uint32_t nRays = job.nPtr ? __ldg(job.nPtr) : job.n;

    bool active = false, exhausted = false;
    uint32_t rayIdx = 0;
    float ox = 0, oy = 0, oz = 0, dx = 0, dy = 0, dz = 0, idx_ = 0, idy = 0, idz = 0, tlo = 0, thi = 0;
    uint32_t oct = 0;
    int sp = 0;
    uint2 ng = make_uint2(0, 0);      // node group: x = child base, y = ordered hit bits << 24 | imask
    uint2 tg = make_uint2(0, 0);      // triangle group: x = triangle base, y = hit bits (0..23)
    int hitTri = -1; float hitT = 0, hitU = 0, hitV = 0;
    constexpr bool TSHADOW = (OUT == OUT_TSHADOW_ACC || OUT == OUT_TSHADOW_RES);
    float filtR = 1.f, filtG = 1.f, filtB = 1.f; int tsDepth = 0;      // transparent shadows: filter colour so far, transparent hits so far

    for (;;) {
        // ------------------------------------------------------------ refill idle lanes (warp-converged here)
        __syncwarp();
        const unsigned idle = __ballot_sync(FULL, !active);
        if (idle == FULL || (!exhausted && __popc(idle) >= (int)job.refillIdle)) {
            if (!exhausted) {
                const int nIdle = __popc(idle);
                const int leader = __ffs(idle) - 1;
                uint32_t base = 0;
                if ((int)lane == leader) base = atomicAdd(job.counter, (uint32_t)nIdle);
                base = __shfl_sync(FULL, base, leader);
                if (!active) {
                    uint32_t i = base + (uint32_t)__popc(idle & ((1u << lane) - 1u));
                    if (i < nRays) {
                        if (SORTED) i = __ldg(job.order + i);          // coherence-sorted fetch order (opt-in, measured slower: DESIGN.md)
                        const float4 r0 = __ldg((const float4 *)(job.rays + (size_t)i * job.stride));
                        const float4 r1 = __ldg((const float4 *)(job.rays + (size_t)i * job.stride + 16));
                        rayIdx = i;
                        dx = r1.x; dy = r1.y; dz = r1.z;
                        if (ANYHIT) {          // scene_t::isShadowed: advance the origin by tmin, dist = tmax - 2 tmin (scene.cc:884-889)
                            ox = r0.x + dx * r0.w; oy = r0.y + dy * r0.w; oz = r0.z + dz * r0.w;
                            tlo = TSHADOW ? r0.w : 0.f; thi = r1.w < 0 ? INFINITY : r1.w - 2 * r0.w;     // IntersectTS accepts t >= tmin once more (kdtree.cc:1055)
                            filtR = filtG = filtB = 1.f; tsDepth = 0;
                        } else {               // scene_t::intersect (scene.cc:856-857)
                            ox = r0.x; oy = r0.y; oz = r0.z;
                            tlo = r0.w; thi = r1.w < 0 ? INFINITY : r1.w;
                        }
                        const float eps = 1e-30f;
                        idx_ = 1.0f / (fabsf(dx) > eps ? dx : (dx < 0 ? -eps : eps));
                        idy = 1.0f / (fabsf(dy) > eps ? dy : (dy < 0 ? -eps : eps));
                        idz = 1.0f / (fabsf(dz) > eps ? dz : (dz < 0 ? -eps : eps));
                        oct = (dx >= 0 ? 1u : 0u) | (dy >= 0 ? 2u : 0u) | (dz >= 0 ? 4u : 0u);
                        sp = 0;
                        ng = make_uint2(0u, ((1u << (0u ^ oct)) << 24) | 1u);   // virtual group holding the root
                        tg = make_uint2(0u, 0u);
                        hitTri = -1; hitT = 0.f; hitU = 0.f; hitV = 0.f;
                        active = job.nNodes != 0;
                        if (!active) {          // empty scene: miss
                            if (OUT == OUT_HITS || OUT == OUT_CLOSEST_PRIM) ((int4 *)job.out)[i] = make_int4(-1, 0, 0, 0);
                            else if (OUT == OUT_SHADOW_FLAG) ((int32_t *)job.out)[i] = 0;
                            else if (OUT == OUT_TSHADOW_RES) ((float4 *)job.out)[i] = make_float4(__int_as_float(0), 1.f, 1.f, 1.f);
                            else { const float4 c = __ldg((const float4 *)(job.rays + (size_t)i * job.stride + 32));
                                   const uint32_t sid = __float_as_uint(c.w);
                                   atomicAdd(&job.acc[sid].x, c.x); atomicAdd(&job.acc[sid].y, c.y); atomicAdd(&job.acc[sid].z, c.z); }
                        }
                    }
                }
                if (base + (uint32_t)nIdle >= nRays) exhausted = true;
            }
            if (__ballot_sync(FULL, active) == 0u) { if (exhausted) break; else continue; }
        }

        // ------------------------------------------------------------ a burst of traversal steps
#pragma unroll 1
        for (int burst = 0; burst < (int)job.burst; ++burst) {
            bool done = false;
            // ---- node phase: every lane that has a node to visit descends one node - also while it still holds triangles of an
            //      earlier leaf (those are tested in the triangle phase of the same round, so a lane advances on both fronts per
            //      round instead of sitting out the node step once per triangle; triangles found meanwhile are parked on the stack).
            //      Closest hit: a parked triangle is tested a little later than in strict order, i.e. against a `thi` that can only
            //      be smaller - the result is the same, a few more nodes are visited.
            bool progressed = false;
            if (active && !(ANYHIT_WAITS && ANYHIT && tg.y != 0u)) {
                // current node group exhausted: take the next one from the stack; parked triangle groups above it are picked up on the
                // way if this lane holds none (otherwise they stay and this round is a triangle step only)
                while (ng.y <= 0x00ffffffu && sp > 0) {
                    const uint2 e = (sp - 1 < YB_SSTACK) ? sstack[sp - 1][threadIdx.x] : lstack[sp - 1 - YB_SSTACK];
                    if (e.y > 0x00ffffffu) { ng = e; --sp; }
                    else if (tg.y == 0u) { tg = e; --sp; progressed = true; }
                    else break;
                }
                if (ng.y > 0x00ffffffu) {
                    progressed = true;
                    const int bit = 31 - __clz((int)ng.y);
                    ng.y &= ~(1u << bit);
                    if (ng.y > 0x00ffffffu) {               // siblings still pending: push the rest of the group
                        if (sp < YB_SSTACK) sstack[sp][threadIdx.x] = ng; else lstack[sp - YB_SSTACK] = ng;
                        ++sp;
                    }
                    const uint32_t slot = ((uint32_t)bit - 24u) ^ oct;
                    const uint32_t nodeIdx = ng.x + (uint32_t)__popc(ng.y & ((1u << slot) - 1u) & 0xffu);
#if defined(YB_HOSTEMU) && defined(YB_TRACE_STATS)
                    g_stat_nodes++;
#endif
                    const uint4 *np = (const uint4 *)(job.nodes + nodeIdx);
                    const uint4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
                    // t = (1 + q 2^-15) * A + B per plane; A = cell * 2^15 / dir, B = (p - o) / dir - A; planes widened by 1/64 cell
                    const float Ax = __uint_as_float(((n0.w & 0xffu) + 15u) << 23) * idx_;
                    const float Ay = __uint_as_float((((n0.w >> 8) & 0xffu) + 15u) << 23) * idy;
                    const float Az = __uint_as_float((((n0.w >> 16) & 0xffu) + 15u) << 23) * idz;
                    const float Bx = (__uint_as_float(n0.x) - ox) * idx_ - Ax;
                    const float By = (__uint_as_float(n0.y) - oy) * idy - Ay;
                    const float Bz = (__uint_as_float(n0.z) - oz) * idz - Az;
                    const float wx = fabsf(Ax) * 4.76837158e-7f, wy = fabsf(Ay) * 4.76837158e-7f, wz = fabsf(Az) * 4.76837158e-7f;   // 2^-21 |A| = cell/64
                    const float Bnx = Bx - wx, Bfx = Bx + wx, Bny = By - wy, Bfy = By + wy, Bnz = Bz - wz, Bfz = Bz + wz;
                    const bool nx = dx < 0, ny = dy < 0, nz = dz < 0;
                    // words: n2 = {lox[0..3], lox[4..7], loy[0..3], loy[4..7]}, n3 = {loz, loz, hix, hix}, n4 = {hiy, hiy, hiz, hiz}
                    const uint32_t nX0 = nx ? n3.z : n2.x, nX1 = nx ? n3.w : n2.y, fX0 = nx ? n2.x : n3.z, fX1 = nx ? n2.y : n3.w;
                    const uint32_t nY0 = ny ? n4.x : n2.z, nY1 = ny ? n4.y : n2.w, fY0 = ny ? n2.z : n4.x, fY1 = ny ? n2.w : n4.y;
                    const uint32_t nZ0 = nz ? n4.z : n3.x, nZ1 = nz ? n4.w : n3.y, fZ0 = nz ? n3.x : n4.z, fZ1 = nz ? n3.y : n4.w;
                    uint32_t hitmask = 0u;
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const uint32_t meta4 = h ? n1.w : n1.z;
                        const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
                        const uint32_t inner_mask4 = sign_extend_s8x4(is_inner4 << 3);             // 0xff where the child is internal
                        const uint32_t bit_index4 = (meta4 ^ ((oct * 0x01010101u) & inner_mask4)) & 0x1f1f1f1fu;
                        const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
                        const uint32_t wnx = h ? nX1 : nX0, wfx = h ? fX1 : fX0, wny = h ? nY1 : nY0, wfy = h ? fY1 : fY0, wnz = h ? nZ1 : nZ0, wfz = h ? fZ1 : fZ0;
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const uint32_t sel = 0x7604u | ((uint32_t)j << 4);
                            const float t0x = fmaf(plane_f(wnx, one, sel), Ax, Bnx), t1x = fmaf(plane_f(wfx, one, sel), Ax, Bfx);
                            const float t0y = fmaf(plane_f(wny, one, sel), Ay, Bny), t1y = fmaf(plane_f(wfy, one, sel), Ay, Bfy);
                            const float t0z = fmaf(plane_f(wnz, one, sel), Az, Bnz), t1z = fmaf(plane_f(wfz, one, sel), Az, Bfz);
                            const float tn = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, tlo));
                            const float tf = fminf(fminf(t1x, t1y), fminf(t1z, thi));
                            if (tn <= tf) {
                                const uint32_t cb = (child_bits4 >> (8 * j)) & 0xffu, bi = (bit_index4 >> (8 * j)) & 0xffu;
                                hitmask |= cb << bi;
                            }
                        }
                    }
                    ng = make_uint2(n1.x, (hitmask & 0xff000000u) | (n0.w >> 24));
                    const uint2 tnew = make_uint2(n1.y, hitmask & 0x00ffffffu);
                    if (tnew.y != 0u) {
                        if (tg.y == 0u) tg = tnew;
                        else { if (sp < YB_SSTACK) sstack[sp][threadIdx.x] = tnew; else lstack[sp - YB_SSTACK] = tnew; ++sp; }
                    }
                }
            }
            // ---- triangle phase (warp-synchronous): every lane holding a triangle group tests ONE triangle.  With a threshold above
            //      one (YB200_TRI_THRESHOLD) the phase waits until that many lanes hold a group or no lane could advance on its nodes.
            const bool hasTri = active && !done && tg.y != 0u;
            const unsigned triMask = __ballot_sync(FULL, hasTri);
            if (triMask != 0u) {
                bool run = true;
                if (job.triThreshold > 1u) run = __popc(triMask) >= (int)job.triThreshold || __ballot_sync(FULL, progressed && !hasTri) == 0u;
                if (run) {
                    if (hasTri) {
                        const int tb = 31 - __clz((int)tg.y);
                        tg.y &= ~(1u << tb);
                        const uint32_t ti = tg.x + (uint32_t)tb;
                        const float4 A = __ldg((const float4 *)(job.tri_v + 3 * (size_t)ti));
                        const float4 B = __ldg((const float4 *)(job.tri_v + 3 * (size_t)ti + 1));
                        const float4 C = __ldg((const float4 *)(job.tri_v + 3 * (size_t)ti + 2));
                        float t, u, v;
#if defined(YB_HOSTEMU) && defined(YB_TRACE_STATS)
                        g_stat_tris++;
#endif
                        if (tri_intersect(v3(A.x, A.y, A.z), v3(B.x, B.y, B.z), v3(C.x, C.y, C.z), v3(ox, oy, oz), v3(dx, dy, dz), t, u, v) && t < thi && t >= tlo) {
                            if (TSHADOW) {
                                const Material &m = job.surf.materials[(uint32_t)__float_as_int(B.w) & TRI_MAT_MASK];
                                if (!mat_isTransparent<1>(m) || tsDepth >= job.shadowDepth) { hitTri = (int)ti; done = true; }
                                else {
                                    Surf sps;
                                    tri_getSurface(job.surf, (int)ti, v3(ox, oy, oz) + t * v3(dx, dy, dz), u, v, sps);
                                    const col3 tr = mat_getTransparency<1>(m, sps, v3(dx, dy, dz));
                                    filtR *= tr.x; filtG *= tr.y; filtB *= tr.z;
                                    ++tsDepth;
                                }
                            } else if (ANYHIT) { hitTri = (int)ti; done = true; }
                            else { thi = t; hitTri = (int)ti; hitT = t; hitU = u; hitV = v; }
                        }
                    }
                }
            }
            // nothing left on any front: the ray ends in this round (not in the node phase of the next one)
            if (active && ng.y <= 0x00ffffffu && sp == 0 && tg.y == 0u) done = true;
            if (done) {
                active = false;
                if (OUT == OUT_HITS) ((float4 *)job.out)[rayIdx] = make_float4(__int_as_float(hitTri), hitT, hitU, hitV);
                else if (OUT == OUT_CLOSEST_PRIM) {
                    const int prim = hitTri >= 0 ? __float_as_int(__ldg(&job.tri_v[3 * (size_t)hitTri].w)) : -1;
                    ((float4 *)job.out)[rayIdx] = make_float4(__int_as_float(prim), hitT, hitU, hitV);
                } else if (OUT == OUT_SHADOW_FLAG) ((int32_t *)job.out)[rayIdx] = hitTri >= 0 ? 1 : 0;
                else if (OUT == OUT_TSHADOW_RES) ((float4 *)job.out)[rayIdx] = make_float4(__int_as_float(hitTri >= 0 ? 1 : 0), filtR, filtG, filtB);
                else if (hitTri < 0) {
                    float4 c = __ldg((const float4 *)(job.rays + (size_t)rayIdx * job.stride + 32));
                    const uint32_t sid = __float_as_uint(c.w);
                    if (TSHADOW) { c.x *= filtR; c.y *= filtG; c.z *= filtB; }
                    atomicAdd(&job.acc[sid].x, c.x); atomicAdd(&job.acc[sid].y, c.y); atomicAdd(&job.acc[sid].z, c.z);
                }
            }
        }
    }
}

template <bool ANYHIT, int OUT>
static void launch_trace_fast(const FastJob &job, rt_stream s)
{
    if (job.n == 0) return;
#ifdef YB_HOSTEMU
#ifdef YB_TRACE_STATS
    g_stat_rays += job.n;
#endif
    if (!job.counterZeroed) *job.counter = 0;
    if (job.order) k_trace_fast<ANYHIT, OUT, true>(job); else k_trace_fast<ANYHIT, OUT, false>(job);
    g_kernel_launches++;
    return;
#else
    if (!job.counterZeroed) YB_CUDA_CHECK(cudaMemsetAsync(job.counter, 0, sizeof(uint32_t), s));
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    uint32_t blocks = (uint32_t)sms * (uint32_t)YB_TRACE_CTAS;
    const uint32_t need = (job.n + YB_TB - 1) / YB_TB;
    if (blocks > need) blocks = need;
    if (job.order) k_trace_fast<ANYHIT, OUT, true><<<blocks, YB_TB, 0, s>>>(job);
    else k_trace_fast<ANYHIT, OUT, false><<<blocks, YB_TB, 0, s>>>(job);
    YB_CUDA_CHECK(cudaGetLastError());
    g_kernel_launches++;
#endif
}

}  // namespace yb
#ifdef YB_HOSTEMU
#undef uint2
#undef uint4
#undef float4
#undef int4
#undef atomicAdd
#undef __ldg
#undef __global__
#undef __shared__
#undef __device__
#undef __forceinline__
#endif
