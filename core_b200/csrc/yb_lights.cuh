// yb_lights.cuh — device restatement of light_t for pointLight_t / sunLight_t / areaLight_t, the
// perspective camera and the triangle surface-point computation.
//   src/lights/pointlight.cc:54-91, src/lights/sunlight.cc:74-95, src/lights/arealight.cc:68-154
//   src/cameras/perspectiveCamera.cc:127-149, include/utilities/geometry.h:33-36
//   src/yafraycore/triangle.cc:12-108 (getSurface), include/yafraycore/triangle_inline.h:27-64
#pragma once
#include "yb_bsdf.cuh"

namespace yb {

struct LSample { float s1, s2; col3 col; float pdf; };   // lSample_t fields used on this path (light.h)

// light_t::diracLight / canIntersect of the six light types
// LT >= 0: the light type is known at compile time (the staged direct-lighting kernels, one per light type: every switch on
// the type folds); LT < 0: read it from the record
template <int LT = -1> YB_HD bool light_dirac(const Light &l) { const int ty = LT >= 0 ? LT : l.type; return ty == LIGHT_POINT || ty == LIGHT_DIRECTIONAL || (ty == LIGHT_SPOT && !l.softShadows); }
template <int LT = -1> YB_HD bool light_canIntersect(const Light &l) { const int ty = LT >= 0 ? LT : l.type; return ty == LIGHT_SUN || ty == LIGHT_AREA || ty == LIGHT_MESH || (ty == LIGHT_SPOT && l.softShadows); }
YB_HD bool tri_intersect(v3 a, v3 b, v3 c, v3 from, v3 dir, float &t, float &u, float &v);   // yb_trace.cuh

// spot cone falloff (spotlight.cc:117-126,163-172,263-272): smoothstep between cosEnd and cosStart
YB_HD float spot_falloff(const Light &l, float cosa)
{
    if (cosa >= l.cosStart) return 1.f;
    float v = (cosa - l.cosEnd) * l.icosDiff;
    return v * v * (3.f - 2.f * v);
}
// pointLight_t::illuminate (pointlight.cc:54-70), directionalLight_t::illuminate (directional.cc:77-97),
// spotLight_t::illuminate (spotlight.cc:103-132)
template <int LT = -1> YB_HD bool light_illuminate(const Light &l, v3 P, col3 &col, Ray &wi)
{
    const int ty = LT >= 0 ? LT : l.type;
    if (ty == LIGHT_DIRECTIONAL) {
        if (!l.infinite) {               // outside of the illuminated cylinder?
            v3 vec = l.position - P;
            float dist = length(cross(l.direction, vec));
            if (dist > l.radius) return false;
            wi.tmax = dot(vec, l.direction);
            if (wi.tmax <= 0.0f) return false;
        } else wi.tmax = -1.0f;
        wi.dir = l.direction;
        col = l.color;
        return true;
    }
    if (ty == LIGHT_SPOT) {
        if (l.photonOnly) return false;
        v3 ldir = l.position - P;
        float dist_sqr = dot(ldir, ldir);
        float dist = fSqrt(dist_sqr);
        if (dist == 0.0f) return false;
        float idist_sqr = 1.f / dist_sqr;
        ldir *= 1.f / dist;
        float cosa = dot(l.ndir, ldir);
        if (cosa < l.cosEnd) return false;
        if (cosa >= l.cosStart) col = l.color * idist_sqr;
        else {
            float v = (cosa - l.cosEnd) * l.icosDiff;
            v = v * v * (3.f - 2.f * v);
            col = l.color * (v * idist_sqr);
        }
        wi.tmax = dist;
        wi.dir = ldir;
        return true;
    }
    if (ty != LIGHT_POINT) return false;
    v3 ldir = l.position - P;
    float dist_sqr = ldir.x * ldir.x + ldir.y * ldir.y + ldir.z * ldir.z;
    float dist = fSqrt(dist_sqr);
    if (dist == 0.0f) return false;
    float idist_sqr = 1.f / dist_sqr;
    ldir *= 1.f / dist;
    wi.tmax = dist;
    wi.dir = ldir;
    col = l.color * idist_sqr;
    return true;
}
// spherelight.cc:86-98 (the 2.0 / 4.0 literals make eb, osc and the roots double expressions rounded to float)
YB_HD bool sphereIntersect(v3 from, v3 dir, v3 c, float R2, float &d1, float &d2)
{
    v3 vf = from - c;
    float ea = dot(dir, dir);
    float eb = (float)(2.0 * (double)dot(vf, dir));
    float ec = dot(vf, vf) - R2;
    float osc = (float)((double)(eb * eb) - 4.0 * (double)ea * (double)ec);
    if (osc < 0) { d1 = fSqrt(ec / ea); return false; }
    osc = fSqrt(osc);
    d1 = (float)((double)(-eb - osc) / (2.0 * (double)ea));
    d2 = (float)((double)(-eb + osc) / (2.0 * (double)ea));
    return true;
}
// sunLight_t::illumSample (sunlight.cc:74-85), areaLight_t::illumSample (arealight.cc:68-96)
// meshLight_t::sampleSurface (meshlight.cc:78-96) = pdf1D_t::DSample (sample_utils.h:141-157) + triangle_t::sample (triangle.cc:156-167)
YB_HD bool mesh_sampleSurface(const Light &l, v3 &p, v3 &n, float s1, float s2)
{
    int primNum = 0;
    if (s1 != 0.f) {
        int lo = 0, hi = l.meshCount + 1;                 // std::lower_bound(cdf, cdf + count + 1, s1)
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (l.meshCdf[mid] < s1) lo = mid + 1; else hi = mid; }
        primNum = lo - 1;
        if (primNum < 0) primNum = 0;
    }
    if (primNum >= l.meshCount) return false;             // "MeshLight: Sampling error!"
    float ss1, delta = l.meshCdf[primNum + 1];
    if (primNum > 0) { delta -= l.meshCdf[primNum]; ss1 = (s1 - l.meshCdf[primNum]) / delta; }
    else ss1 = s1 / delta;
    const f4 A = l.meshTris[3 * primNum], B = l.meshTris[3 * primNum + 1], C = l.meshTris[3 * primNum + 2];
    const float su1 = fSqrt(ss1), u = 1.f - su1, v = s2 * su1;
    p = u * v3(A.x, A.y, A.z) + v * v3(B.x, B.y, B.z) + (1.f - u - v) * v3(C.x, C.y, C.z);
    n = v3(A.w, B.w, C.w);
    return true;
}
template <int EXT, int LT = -1> YB_HD bool light_illumSample(const Light &l, v3 P, LSample &s, Ray &wi)
{
    const int ty = LT >= 0 ? LT : l.type;
    if (EXT && ty == LIGHT_MESH) {           // meshLight_t::illumSample (meshlight.cc:100-137)
        v3 n(0.f), p(0.f);
        mesh_sampleSurface(l, p, n, s.s1, s.s2);
        v3 ldir = p - P;
        float dist_sqr = dot(ldir, ldir);
        float dist = fSqrt(dist_sqr);
        if (dist <= 0.0f) return false;
        ldir *= 1.f / dist;
        float cos_angle = -dot(ldir, n);
        if (cos_angle <= 0) {
            if (l.doubleSided) cos_angle = -cos_angle;
            else return false;
        }
        wi.tmax = dist;
        wi.dir = ldir;
        s.col = l.color;
        float area_mul_cosangle = l.area * cos_angle;
        s.pdf = (float)((double)dist_sqr * YB_M_PI / (double)((area_mul_cosangle == 0.f) ? 1e-8f : area_mul_cosangle));
        return true;
    }
    if (ty == LIGHT_SUN) {
        wi.dir = sampleCone(l.direction, l.du, l.dv, (float)l.cosAngle, s.s1, s.s2);
        wi.tmax = -1.f;
        s.col = l.colPdf;
        s.pdf = l.pdf;
        return true;
    }
    if (ty == LIGHT_SPOT) {                  // spotLight_t::illumSample (spotlight.cc:134-177)
        if (l.photonOnly) return false;
        v3 ldir = l.position - P;
        float dist_sqr = dot(ldir, ldir);
        if (dist_sqr == 0.0f) return false;
        float dist = fSqrt(dist_sqr);
        ldir *= 1.f / dist;
        float cosa = dot(l.ndir, ldir);
        if (cosa < l.cosEnd) return false;
        wi.tmax = dist;
        wi.dir = sampleCone(ldir, l.du, l.dv, l.cosEnd, s.s1 * l.shadowFuzzy, s.s2 * l.shadowFuzzy);
        s.col = l.color * spot_falloff(l, cosa);
        s.pdf = dist_sqr;
        return true;
    }
    if (ty == LIGHT_SPHERE) {                // sphereLight_t::illumSample (spherelight.cc:100-132)
        v3 cdir = l.position - P;
        float dist_sqr = dot(cdir, cdir);
        if (dist_sqr <= l.squareRadius) return false;     // only emits on the outside
        float dist = fSqrt(dist_sqr);
        float idist_sqr = 1.f / dist_sqr;
        float cosAlpha = fSqrt(1.f - l.squareRadius * idist_sqr);
        cdir *= 1.f / dist;
        v3 du, dv;
        createCS(cdir, du, dv);
        wi.dir = sampleCone(cdir, du, dv, cosAlpha, s.s1, s.s2);
        float d1, d2;
        if (!sphereIntersect(wi.from, wi.dir, l.position, l.squareRadiusEps, d1, d2)) return false;
        wi.tmax = d1;
        s.pdf = 1.f / (2.f * (1.f - cosAlpha));
        s.col = l.color;
        return true;
    }
    if (ty == LIGHT_AREA) {
        v3 p = l.corner + s.s1 * l.toX + s.s2 * l.toY;
        v3 ldir = p - P;
        float dist_sqr = dot(ldir, ldir);
        float dist = fSqrt(dist_sqr);
        if (dist <= 0.0f) return false;
        ldir *= 1.f / dist;
        float cos_angle = dot(ldir, l.fnormal);
        if (cos_angle <= 0) return false;
        wi.tmax = dist;
        wi.dir = ldir;
        s.col = l.color;
        s.pdf = (float)(dist_sqr * YB_M_PI / (l.area * cos_angle));
        return true;
    }
    return false;
}
// arealight.cc:117-134
YB_HD bool light_triIntersect(v3 a, v3 b, v3 c, v3 from, v3 dir, float &t)
{
    v3 edge1 = b - a, edge2 = c - a;
    v3 pvec = cross(dir, edge2);
    float det = dot(edge1, pvec);
    if (det == 0.0f) return false;
    float inv_det = (float)(1.0 / det);
    v3 tvec = from - a;
    float u = dot(tvec, pvec) * inv_det;
    if (u < 0.0f || u > 1.0f) return false;
    v3 qvec = cross(tvec, edge1);
    float v = dot(dir, qvec) * inv_det;
    if ((v < 0.0f) || ((u + v) > 1.0f)) return false;
    t = dot(edge2, qvec) * inv_det;
    return true;
}
// sunLight_t::intersect (sunlight.cc:87-95), areaLight_t::intersect (arealight.cc:136-154)
template <int EXT, int LT = -1> YB_HD bool light_intersect(const Light &l, v3 from, v3 dir, float tmin, float &t, col3 &col, float &ipdf)
{
    const int ty = LT >= 0 ? LT : l.type;
    if (EXT && ty == LIGHT_MESH) {           // meshLight_t::intersect (meshlight.cc:181-205): closest hit on the emitter's own triangles
        // (the reference walks a private kd-tree; here the emitter list is tested in order, acceptance rule of kdtree.cc:772,791)
        float Z = t < 0 ? INFINITY : t;
        int hit = -1;
        for (int i = 0; i < l.meshCount; ++i) {
            const f4 A = l.meshTris[3 * i], B = l.meshTris[3 * i + 1], C = l.meshTris[3 * i + 2];
            float th, u, v;
            if (!tri_intersect(v3(A.x, A.y, A.z), v3(B.x, B.y, B.z), v3(C.x, C.y, C.z), from, dir, th, u, v)) continue;
            if (th < Z && th >= tmin) { Z = th; hit = i; }
        }
        if (hit < 0) return false;
        t = Z;
        const v3 n(l.meshTris[3 * hit].w, l.meshTris[3 * hit + 1].w, l.meshTris[3 * hit + 2].w);
        float cos_angle = dot(dir, -n);
        if (cos_angle <= 0) {
            if (l.doubleSided) cos_angle = fabsf(cos_angle);
            else return false;
        }
        float idist_sqr = 1.f / (t * t);
        ipdf = (float)((double)(idist_sqr * l.area * cos_angle) * (1.f / YB_M_PI));
        col = l.color;
        return true;
    }
    if (ty == LIGHT_SUN) {
        float cosine = dot(dir, l.direction);
        if ((double)cosine < l.cosAngle) return false;
        col = l.colPdf;
        t = -1;
        ipdf = l.invpdf;
        return true;
    }
    if (ty == LIGHT_SPOT) {                  // spotLight_t::intersect (spotlight.cc:243-280), quirks included
        float cosA = dot(l.direction, dir);
        if (cosA == 0.f) return false;
        t = dot(l.direction, l.position - from) / cosA;
        if (t < 0.f) return false;
        v3 P = from + t * dir;
        if (dot(l.direction, P - l.position) == 0) {
            if (dot(P, P) <= 1e-2f) {
                float cosa = dot(l.direction, dir);
                if (cosa < l.cosEnd) return false;
                col = l.color * spot_falloff(l, cosa);
                ipdf = 1.f / (t * t);
                return true;
            }
        }
        return false;
    }
    if (ty == LIGHT_AREA) {
        float cos_angle = dot(dir, l.fnormal);
        if (cos_angle <= 0) return false;
        if (!light_triIntersect(l.corner, l.c2, l.c3, from, dir, t)) {
            if (!light_triIntersect(l.corner, l.c3, l.c4, from, dir, t)) return false;
        }
        if (!(t > 1.0e-10f)) return false;
        col = l.color;
        ipdf = (float)(1.f / (t * t) * l.area * cos_angle * YB_M_1_PI);
        return true;
    }
    return false;
}

// perspectiveCam_t::biasDist / sampleTSD / getLensUV (perspectiveCamera.cc:73-122)
YB_HD void cam_biasDist(const Camera &c, float &r)
{
    if (c.bkhbias == 1) r = fSqrt(fSqrt(r) * r);            // BB_CENTER
    else if (c.bkhbias == 2) r = fSqrt(1.0f - r * r);       // BB_EDGE
    else r = fSqrt(r);
}
YB_HD void cam_getLensUV(const Camera &c, float r1, float r2, float &u, float &v)
{
    if (c.bkhtype >= 3 && c.bkhtype <= 6) {                 // BK_TRI .. BK_HEXA: sampleTSD
        const float fn = (float)c.bkhtype;
        int idx = (int)(r1 * fn);
        r1 = (r1 - ((float)idx) / fn) * fn;
        cam_biasDist(c, r1);
        const float b1 = r1 * r2, b0 = r1 - b1;
        idx <<= 1;
        u = c.LS[idx] * b0 + c.LS[idx + 2] * b1;
        v = c.LS[idx + 1] * b0 + c.LS[idx + 3] * b1;
    } else if (c.bkhtype == 1 || c.bkhtype == 7) {          // BK_DISK2, BK_RING
        const float w = (float)YB_M_2PI * r2;
        if (c.bkhtype == 7) r1 = fSqrt(0.707106781f + 0.292893218f);
        else cam_biasDist(c, r1);
        u = r1 * fCos(w); v = r1 * fSin(w);
    } else ShirleyDisk(r1, r2, u, v);                       // BK_DISK1
}
// perspectiveCam_t::shootRay (perspectiveCamera.cc:127-149)
YB_HD Ray cam_shootRay(const Camera &c, float px, float py, float lu, float lv, float &wt)
{
    Ray ray;
    wt = 1;
    if (c.camType == CAM_ORTHO) {                               // orthoCam_t::shootRay (orthographicCamera.cc:52-64)
        ray.from = c.opos + c.vright * px + c.vup * py;
        ray.dir = c.vto;
        float den = dot(ray.dir, c.camZ);
        ray.tmin = dot(c.camZ, c.near_p - ray.from) / den;
        ray.tmax = dot(c.camZ, c.far_p - ray.from) / den;
        return ray;
    }
    if (c.camType == CAM_ANGULAR) {                             // angularCam_t::shootRay (angularCamera.cc:50-71)
        ray.from = c.position;
        ray.dir = v3(0.f, 0.f, 1.f); ray.tmin = 0.f; ray.tmax = 0.f;
        float u = 1.f - 2.f * (px / (float)c.resx);
        float v = 2.f * (py / (float)c.resy) - 1.f;
        v *= c.aspectRatio;
        float radius = fSqrt(u * u + v * v);
        if (c.circular && radius > c.maxR) { wt = 0; return ray; }
        float theta = 0;
        if (!((u == 0) && (v == 0))) theta = atan2f(v, u);
        float phi = radius * c.horPhi;
        ray.dir = fSin(phi) * (fCos(theta) * c.vright + fSin(theta) * c.vup) + fCos(phi) * c.vto;
        float den = dot(ray.dir, c.camZ);
        ray.tmin = dot(c.camZ, c.near_p - ray.from) / den;
        ray.tmax = dot(c.camZ, c.far_p - ray.from) / den;
        return ray;
    }
    ray.from = c.position;
    ray.dir = normalize(c.vright * px + c.vup * py + c.vto);
    float denom = dot(ray.dir, c.camZ);
    ray.tmin = dot(c.camZ, c.near_p - ray.from) / denom;      // geometry.h:33-36
    ray.tmax = dot(c.camZ, c.far_p - ray.from) / denom;
    if (c.aperture != 0) {
        float u, v;
        cam_getLensUV(c, lu, lv, u, v);
        v3 LI = c.dof_rt * u + c.dof_up * v;
        ray.from = ray.from + LI;
        ray.dir = normalize((ray.dir * c.dof_distance) - LI);
    }
    return ray;
}

// triangle_t::getSurface (triangle.cc:12-108) without orco/uv (rejected at flatten time):
// sp.N from the vertex normals, createCS(N, NU, NV); Ng = normalize((b-a)^(c-a)) as recNormal (triangle_inline.h:100-107)
// the geometry tables it reads, so that the traversal kernel (which carries no SceneDev) can call it too
struct SurfTables { const f4 *tri_v; const v3 *tri_n; const Material *materials; const v3 *prim_ng; };
YB_HD void tri_getSurface(const SurfTables &sc, int tri, v3 hitP, float b1, float b2, Surf &sp)
{
    const f4 A = sc.tri_v[3 * tri], B = sc.tri_v[3 * tri + 1], C = sc.tri_v[3 * tri + 2];
    v3 a(A.x, A.y, A.z), b(B.x, B.y, B.z), c(C.x, C.y, C.z);
    uint32_t info = (uint32_t)f2i(B.w);
    sp.hasBlendVal = 0; sp.blendVal = 0.f;
    // instances keep the BASE triangle's normal pushed through objToWorld (not perpendicular under non-uniform scale)
    sp.Ng = (info & TRI_EXPLICIT_NG) ? sc.prim_ng[f2i(A.w)] : normalize(cross(b - a, c - a));
    float u = 1 - b1 - b2, v = b1, w = b2;        // data.b0, b1, b2
    if (info & TRI_SMOOTH) {
        v3 va = (info & TRI_HAS_NA) ? sc.tri_n[3 * tri] : sp.Ng;
        v3 vb = (info & TRI_HAS_NB) ? sc.tri_n[3 * tri + 1] : sp.Ng;
        v3 vc = (info & TRI_HAS_NC) ? sc.tri_n[3 * tri + 2] : sp.Ng;
        sp.N = normalize(u * va + v * vb + w * vc);
    } else sp.N = sp.Ng;
    sp.P = hitP;
    sp.mat = &sc.materials[info & TRI_MAT_MASK];
    createCS(sp.N, sp.NU, sp.NV);
}

YB_HD void tri_getSurface(const SceneDev &sc, int tri, v3 hitP, float b1, float b2, Surf &sp)
{
    SurfTables t; t.tri_v = sc.tri_v; t.tri_n = sc.tri_n; t.materials = sc.materials; t.prim_ng = sc.prim_ng;
    tri_getSurface(t, tri, hitP, b1, b2, sp);
}

// the texture-coordinate part of triangle_t::getSurface (triangle.cc:28-105): orco point / normal, interpolated UV and - when
// the shading frame is given (bump mapping from image textures) - dPdU / dPdV brought into shading space
YB_HD void tri_getTexPoint(const SceneDev &sc, int tri, v3 hitP, float b1, float b2, v3 Ng, TexPoint &tp, const Surf *sp = 0)
{
    const f4 A = sc.tri_v[3 * tri], B = sc.tri_v[3 * tri + 1], C = sc.tri_v[3 * tri + 2];
    const uint32_t info = (uint32_t)f2i(B.w);
    const int prim = f2i(A.w);
    const float u = 1 - b1 - b2, v = b1, w = b2;
    if ((info & TRI_HAS_ORCO) && sc.prim_orco) {
        const v3 p0 = sc.prim_orco[3 * (size_t)prim], p1 = sc.prim_orco[3 * (size_t)prim + 1], p2 = sc.prim_orco[3 * (size_t)prim + 2];
        tp.orcoP = u * p0 + v * p1 + w * p2;
        tp.orcoNg = normalize(cross(p1 - p0, p2 - p0));
    } else { tp.orcoP = hitP; tp.orcoNg = Ng; }
    tp.U = 0.f; tp.V = 0.f;                              // left uninitialised by the reference without UVs (triangle.cc:85-90)
    const v3 p0(A.x, A.y, A.z), p1(B.x, B.y, B.z), p2(C.x, C.y, C.z);
    v3 dPdU = p1 - p0, dPdV = p2 - p1;                   // implicit mapping (triangle.cc:78-90)
    if ((info & TRI_HAS_UV) && sc.prim_uv) {
        const float *q = sc.prim_uv + 6 * (size_t)prim;
        tp.U = u * q[0] + v * q[2] + w * q[4];
        tp.V = u * q[1] + v * q[3] + w * q[5];
        const float du1 = q[0] - q[4], du2 = q[2] - q[4], dv1 = q[1] - q[5], dv2 = q[3] - q[5];
        const float det = du1 * dv2 - dv1 * du2;
        if (fabsf(det) > 1e-30f) {
            const float invdet = 1.f / det;
            const v3 dp1 = p0 - p2, dp2 = p1 - p2;
            dPdU = (dv2 * dp1 - dv1 * dp2) * invdet;
            dPdV = (du1 * dp2 - du2 * dp1) * invdet;
        }
    }
    tp.dSdU = tp.dSdV = v3(0.f);
    if (sp) {
        dPdU = normalize(dPdU); dPdV = normalize(dPdV);
        tp.dSdU = v3(dot(sp->NU, dPdU), dot(sp->NV, dPdU), dot(sp->N, dPdU));
        tp.dSdV = v3(dot(sp->NU, dPdV), dot(sp->NV, dPdV), dot(sp->N, dPdV));
    }
}

}  // namespace yb
