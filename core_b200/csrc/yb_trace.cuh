// yb_trace.cuh — per-ray traversal of the 8-wide quantised BVH (replaces triKdTree_t::Intersect /
// IntersectS / IntersectTS, src/yafraycore/kdtree.cc:675-1108) and the reference triangle test
// (triangle_t::intersect, include/yafraycore/triangle_inline.h:27-64) restated with the same operation
// order and without FMA contraction so that t/u/v agree to rounding.
//
// Acceptance rules kept from the reference:
//   closest : t_hit <  Z   && t_hit >= ray.tmin   (kdtree.cc:772,791)    Z starts at `dist`
//   shadow  : t_hit <  dist && t_hit >= 0          (kdtree.cc:916)  on the pre-advanced origin (scene.cc:884-889)
//   tshadow : t_hit <  dist && t_hit >= ray.tmin   (kdtree.cc:1055) on the pre-advanced origin (scene.cc:906-910)
#pragma once
#include "yb_lights.cuh"

namespace yb {

#if defined(__CUDA_ARCH__)
#define YB_MUL(a, b) __fmul_rn((a), (b))
#define YB_ADD(a, b) __fadd_rn((a), (b))
#define YB_SUB(a, b) __fsub_rn((a), (b))
YB_D u4 ld16(const void *p) { uint4 v = __ldg((const uint4 *)p); u4 r; r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w; return r; }
YB_D f4 ld16f(const void *p) { float4 v = __ldg((const float4 *)p); f4 r; r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w; return r; }
YB_D int yb_clz(uint32_t x) { return __clz((int)x); }
YB_D int yb_popc(uint32_t x) { return __popc(x); }
#else
#define YB_MUL(a, b) ((a) * (b))
#define YB_ADD(a, b) ((a) + (b))
#define YB_SUB(a, b) ((a) - (b))
YB_HD u4 ld16(const void *p) { return *(const u4 *)p; }
YB_HD f4 ld16f(const void *p) { return *(const f4 *)p; }
YB_HD int yb_clz(uint32_t x) { return x ? __builtin_clz(x) : 32; }
YB_HD int yb_popc(uint32_t x) { return __builtin_popcount(x); }
#endif

YB_HD float dot_rn(v3 a, v3 b) { return YB_ADD(YB_ADD(YB_MUL(a.x, b.x), YB_MUL(a.y, b.y)), YB_MUL(a.z, b.z)); }
YB_HD v3 cross_rn(v3 a, v3 b)
{
    return v3(YB_SUB(YB_MUL(a.y, b.z), YB_MUL(a.z, b.y)), YB_SUB(YB_MUL(a.z, b.x), YB_MUL(a.x, b.z)), YB_SUB(YB_MUL(a.x, b.y), YB_MUL(a.y, b.x)));
}
YB_HD v3 sub_rn(v3 a, v3 b) { return v3(YB_SUB(a.x, b.x), YB_SUB(a.y, b.y), YB_SUB(a.z, b.z)); }

// triangle_inline.h:27-64
YB_HD bool tri_intersect(v3 a, v3 b, v3 c, v3 from, v3 dir, float &t, float &u, float &v)
{
    v3 edge1 = sub_rn(b, a), edge2 = sub_rn(c, a);
    v3 pvec = cross_rn(dir, edge2);
    float det = dot_rn(edge1, pvec);
    if (det == 0.0f) return false;
#if defined(__CUDA_ARCH__)
    float inv_det = __fdiv_rn(1.0f, det);
#else
    float inv_det = (float)(1.0 / det);
#endif
    v3 tvec = sub_rn(from, a);
    u = YB_MUL(dot_rn(tvec, pvec), inv_det);
    if (u < 0.0f || u > 1.0f) return false;
    v3 qvec = cross_rn(tvec, edge1);
    v = YB_MUL(dot_rn(dir, qvec), inv_det);
    if ((v < 0.0f) || (YB_ADD(u, v) > 1.0f)) return false;
    t = YB_MUL(dot_rn(edge2, qvec), inv_det);
    return true;
}

#define YB_STACK_SIZE 48
enum { TRACE_CLOSEST = 0, TRACE_SHADOW = 1, TRACE_TSHADOW = 2 };

// Result of a transparent-shadow query (scene_t::isShadowed(state, ray, maxDepth, filt))
struct TSResult { bool shadowed; col3 filt; };

// Generic traversal.  `from`/`dir` are the (already advanced, for shadow rays) origin and direction; hits
// are accepted for tlo <= t < thi.  MODE selects closest / any-hit / transparent-shadow behaviour.
template <int MODE>
YB_HD bool bvh_traverse(const SceneDev &sc, v3 from, v3 dir, float tlo, float thi, Hit &hit, int maxDepth, col3 &filt)
{
    hit.tri = -1; hit.t = 0.f; hit.b1 = 0.f; hit.b2 = 0.f;
    if (sc.nNodes == 0) return false;
    const float eps = 1e-30f;
    v3 idir(1.0f / (fabsf(dir.x) > eps ? dir.x : (dir.x < 0 ? -eps : eps)),
            1.0f / (fabsf(dir.y) > eps ? dir.y : (dir.y < 0 ? -eps : eps)),
            1.0f / (fabsf(dir.z) > eps ? dir.z : (dir.z < 0 ? -eps : eps)));
    // ray octant: bit set when the component is positive.  Children sit in slots by their position relative to
    // the node centre (bit set = positive side), so the slot entered first is ~oct and the visiting key is slot^oct.
    const uint32_t oct = (dir.x >= 0 ? 1u : 0u) | (dir.y >= 0 ? 2u : 0u) | (dir.z >= 0 ? 4u : 0u);
    uint32_t stackBase[YB_STACK_SIZE], stackBits[YB_STACK_SIZE];
    int sp = 0;
    int depth = 0;                      // TSHADOW: number of transparent surfaces crossed (kdtree.cc:1064)
    // start: a virtual group holding only the root (slot 0, imask bit 0)
    uint32_t gBase = 0, gBits = (1u << (0 ^ oct)) | (1u << 8);
    bool found = false;
    for (;;) {
        if ((gBits & 0xffu) == 0) {
            if (sp == 0) break;
            --sp; gBase = stackBase[sp]; gBits = stackBits[sp];
            continue;
        }
        const int k = 31 - yb_clz(gBits & 0xffu);
        gBits &= ~(1u << k);
        const uint32_t slot = (uint32_t)k ^ oct;
        const uint32_t imaskG = (gBits >> 8) & 0xffu;
        const uint32_t nodeIdx = gBase + (uint32_t)yb_popc(imaskG & ((1u << slot) - 1u));
        if (gBits & 0xffu) {
            if (sp < YB_STACK_SIZE) { stackBase[sp] = gBase; stackBits[sp] = gBits; ++sp; }
        }
        // ---- fetch + decode node
        const char *np = (const char *)(sc.nodes + nodeIdx);
        const u4 n0 = ld16(np), n1 = ld16(np + 16), n2 = ld16(np + 32), n3 = ld16(np + 48), n4 = ld16(np + 64);
        const float px = i2f((int32_t)n0.x), py = i2f((int32_t)n0.y), pz = i2f((int32_t)n0.z);
        const float cx = i2f((int32_t)((n0.w & 0xffu) << 23)), cy = i2f((int32_t)(((n0.w >> 8) & 0xffu) << 23)), cz = i2f((int32_t)(((n0.w >> 16) & 0xffu) << 23));
        const uint32_t imask = n0.w >> 24;
        const float ax = cx * idir.x, ay = cy * idir.y, az = cz * idir.z;
        const float bx = (px - from.x) * idir.x, by = (py - from.y) * idir.y, bz = (pz - from.z) * idir.z;
        // near/far byte planes per axis according to direction sign
        const uint32_t lox0 = n2.x, lox1 = n2.y, loy0 = n2.z, loy1 = n2.w, loz0 = n3.x, loz1 = n3.y;
        const uint32_t hix0 = n3.z, hix1 = n3.w, hiy0 = n4.x, hiy1 = n4.y, hiz0 = n4.z, hiz1 = n4.w;
        const bool nx = dir.x < 0, ny = dir.y < 0, nz = dir.z < 0;
        const uint32_t nearx0 = nx ? hix0 : lox0, nearx1 = nx ? hix1 : lox1, farx0 = nx ? lox0 : hix0, farx1 = nx ? lox1 : hix1;
        const uint32_t neary0 = ny ? hiy0 : loy0, neary1 = ny ? hiy1 : loy1, fary0 = ny ? loy0 : hiy0, fary1 = ny ? loy1 : hiy1;
        const uint32_t nearz0 = nz ? hiz0 : loz0, nearz1 = nz ? hiz1 : loz1, farz0 = nz ? loz0 : hiz0, farz1 = nz ? loz1 : hiz1;
        uint32_t childHits = 0;         // ordered bits of internal children that are hit
        uint32_t leafHits = 0;          // slot bits of leaf children that are hit
        for (int s = 0; s < 8; ++s) {
            const uint32_t meta = ((s < 4 ? n1.z : n1.w) >> (8 * (s & 3))) & 0xffu;
            if (meta == 0) continue;
            const int sh = 8 * (s & 3);
            const float qnx = (float)(((s < 4 ? nearx0 : nearx1) >> sh) & 0xffu), qfx = (float)(((s < 4 ? farx0 : farx1) >> sh) & 0xffu);
            const float qny = (float)(((s < 4 ? neary0 : neary1) >> sh) & 0xffu), qfy = (float)(((s < 4 ? fary0 : fary1) >> sh) & 0xffu);
            const float qnz = (float)(((s < 4 ? nearz0 : nearz1) >> sh) & 0xffu), qfz = (float)(((s < 4 ? farz0 : farz1) >> sh) & 0xffu);
            const float t0x = qnx * ax + bx, t1x = qfx * ax + bx;
            const float t0y = qny * ay + by, t1y = qfy * ay + by;
            const float t0z = qnz * az + bz, t1z = qfz * az + bz;
            const float tn = fmaxf_(fmaxf_(t0x, t0y), fmaxf_(t0z, tlo));
            const float tf = fminf_(fminf_(t1x, t1y), fminf_(t1z, thi)) * 1.0000005f;
            if (tn <= tf) {
                if ((meta & 0x1fu) >= 24u) childHits |= 1u << ((uint32_t)s ^ oct);
                else leafHits |= 1u << s;
            }
        }
        // ---- leaf children: test their triangles now (shrinks thi before descending)
        while (leafHits) {
            const int s = 31 - yb_clz(leafHits);
            leafHits &= ~(1u << s);
            const uint32_t meta = ((s < 4 ? n1.z : n1.w) >> (8 * (s & 3))) & 0xffu;
            const uint32_t first = n1.y + (meta & 0x1fu), cnt = (uint32_t)yb_popc(meta >> 5);
            for (uint32_t ti = first; ti < first + cnt; ++ti) {
                const f4 A = ld16f(sc.tri_v + 3 * ti), B = ld16f(sc.tri_v + 3 * ti + 1), C = ld16f(sc.tri_v + 3 * ti + 2);
                float t, u, v;
                if (!tri_intersect(v3(A.x, A.y, A.z), v3(B.x, B.y, B.z), v3(C.x, C.y, C.z), from, dir, t, u, v)) continue;
                if (!(t < thi && t >= tlo)) continue;
                if (MODE == TRACE_CLOSEST) {
                    // strict `<`: on exact t ties the first-tested triangle stays, as in the reference (kdtree.cc:772);
                    // test order differs between a kd-tree and a BVH, so ties are compared by t, not by id
                    thi = t; hit.tri = (int32_t)ti; hit.t = t; hit.b1 = u; hit.b2 = v; found = true;
                } else if (MODE == TRACE_SHADOW) {
                    hit.tri = (int32_t)ti; hit.t = t; hit.b1 = u; hit.b2 = v;
                    return true;
                } else {
                    const Material &m = sc.materials[(uint32_t)f2i(B.w) & TRI_MAT_MASK];
                    if (!mat_isTransparent<1>(m)) return true;
                    if (depth >= maxDepth) return true;
                    Surf sps;
                    tri_getSurface(sc, (int)ti, from + t * dir, u, v, sps);
                    Material rmat;                      // shinydiffuse.cc:435-455 evaluates the nodes itself for shadow rays
                    const Material *mp = &m;
                    if (m.nNodes > 0) { TexPoint tp; tri_getTexPoint(sc, (int)ti, sps.P, u, v, sps.Ng, tp); mp = &mat_resolve<2>(sc, m, sps.P, sps.Ng, tp, rmat, m.type == MAT_BLEND ? &sps : 0); }
                    filt *= mat_getTransparency<1>(*mp, sps, dir);
                    ++depth;
                }
            }
        }
        gBase = n1.x;
        gBits = childHits | (imask << 8);
    }
    return found;
}

// scene_t::intersect (scene.cc:852-879) up to the hit record
YB_HD bool scene_intersect(const SceneDev &sc, const Ray &ray, Hit &hit)
{
    float dis = ray.tmax < 0 ? INFINITY : ray.tmax;
    col3 dummy(1.f);
    return bvh_traverse<TRACE_CLOSEST>(sc, ray.from, ray.dir, ray.tmin, dis, hit, 0, dummy);
}
// scene_t::isShadowed(state, ray) (scene.cc:881-902)
YB_HD bool scene_isShadowed(const SceneDev &sc, const Ray &ray)
{
    v3 from = ray.from + ray.dir * ray.tmin;
    float dis = ray.tmax < 0 ? INFINITY : ray.tmax - 2 * ray.tmin;
    Hit h; col3 dummy(1.f);
    return bvh_traverse<TRACE_SHADOW>(sc, from, ray.dir, 0.f, dis, h, 0, dummy);
}
// scene_t::isShadowed(state, ray, maxDepth, filt) (scene.cc:904-928)
YB_HD bool scene_isShadowedT(const SceneDev &sc, const Ray &ray, int maxDepth, col3 &filt)
{
    v3 from = ray.from + ray.dir * ray.tmin;
    float dis = ray.tmax < 0 ? INFINITY : ray.tmax - 2 * ray.tmin;
    Hit h;
    filt = col3(1.f);
    return bvh_traverse<TRACE_TSHADOW>(sc, from, ray.dir, ray.tmin, dis, h, maxDepth, filt);
}

}  // namespace yb
