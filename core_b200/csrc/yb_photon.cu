// yb_photon.cu — the caustic photon map of mcIntegrator_t:
//   createCausticMap       src/yafraycore/mcintegrator.cc:197-382   -> k_photon_shoot + build_photon_grid
//   estimateCausticPhotons src/yafraycore/mcintegrator.cc:384-419   -> k_caustic_gather
// over photonMap_t / pointKdTree (src/yafraycore/photon.cc, include/yafraycore/pkdtree.h), light_t::emitPhoton / totalEnergy of the
// light files and material_t::scatterPhoton (src/yafraycore/material.cc:31-48).
//
// B200 form.  Photon paths are independent and every sample value is a closed function of the path index (RI_vdC, scrHalton),
// so one thread walks one photon path with the per-ray BVH traversal and appends its caustic deposits to a flat array; the
// reference's serial loop and its `curr` counter become the thread index, setNumPaths(curr) becomes an atomic maximum.  The
// balanced kd-tree of photonMap_t is replaced by a uniform grid with cells of the gather radius (counting sort by cell): the
// gather visits the 27 cells around the shading point.  What the lookup RETURNS is restated exactly - the nCausSearch nearest
// photons inside the radius, and, once that many were found, the radius shrunk to the farthest of them (the kd-tree lookup
// lowers maxDistSquared by reference and estimateCausticPhotons builds its kernel from the lowered value).
#include "yb_shade_common.cuh"
#include <vector>
#include <algorithm>
#ifndef YB_HOSTEMU
#include <cub/cub.cuh>
#endif

namespace yb {

#define YB_MAX_CAUS_SEARCH 256

// ------------------------------------------------------------------------------------------------ light_t::emitPhoton
YB_HD v3 SampleSphere(float s1, float s2)                     // sample_utils.h:54-72
{
    v3 dir;
    dir.z = 1.0f - 2.0f * s1;
    float r = 1.0f - dir.z * dir.z;
    if (r > 0.0f) {
        r = fSqrt(r);
        const float a = (float)YB_M_2PI * s2;
        dir.x = fCos(a) * r; dir.y = fSin(a) * r;
    } else { dir.x = 0.f; dir.y = 0.f; }
    return dir;
}
YB_HD void minRot(v3 D, v3 U, v3 D2, v3 &U2, v3 &V2)          // sample_utils.h:168-176
{
    const float cosAlpha = dot(D, D2);
    const float sinAlpha = fSqrt(1 - cosAlpha * cosAlpha);
    const v3 v = cross(D, D2);
    U2 = cosAlpha * U + (1.f - cosAlpha) * dot(v, U) * v3(1.f, 1.f, 1.f) + sinAlpha * cross(v, U);
    V2 = cross(D2, U2);
}
// the smoothstep table of spotLight_t (spotlight.cc:74-84) through pdf1D_t::Sample (sample_utils.h:118-136)
YB_HD float spot_pdfSample(float u, float &pdf, float &integral)
{
    float func[65], cdf[66];
    for (int i = 0; i < 65; ++i) { const float v = (float)i / 64.f; func[i] = v * v * (3.f - 2.f * v); }
    double c = 0.0; const double delta = 1.0 / 65.0;
    cdf[0] = 0.f;
    for (int i = 1; i < 66; ++i) { c += (double)func[i - 1] * delta; cdf[i] = (float)c; }
    integral = (float)c;
    for (int i = 1; i < 66; ++i) cdf[i] /= integral;
    int lo = 0, hi = 66;                                     // std::lower_bound(cdf, cdf + count + 1, u)
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (cdf[mid] < u) lo = mid + 1; else hi = mid; }
    int index = lo - 1;
    if (index < 0) index = 0;
    const float d = (u - cdf[index]) / (cdf[index + 1] - cdf[index]);
    pdf = func[index] * (1.f / integral);
    return (float)index + d;
}

struct PhotonScene {        // what the lights' init(scene) derive from the scene bound (sunlight.cc:65-72, directional.cc:63-76)
    v3 worldCenter; float worldRadius;
};

// color = emitPhoton(s1, s2, s3, s4, ray, ipdf) of the light files
YB_HD col3 light_emitPhoton(const Light &l, const PhotonScene &ps, float s1, float s2, float s3, float s4, Ray &ray, float &ipdf)
{
    switch (l.type) {
    case LIGHT_POINT:                                        // pointlight.cc:91-97
        ray.from = l.position; ray.dir = SampleSphere(s1, s2); ipdf = 4.0f * (float)YB_M_PI;
        return l.color;
    case LIGHT_SUN: {                                        // sunlight.cc:97-112
        float u, v;
        ShirleyDisk(s3, s4, u, v);
        const v3 ldir = sampleCone(l.direction, l.du, l.dv, (float)l.cosAngle, s3, s4);
        v3 du2, dv2;
        minRot(l.direction, l.du, ldir, du2, dv2);
        ipdf = l.invpdf;
        ray.from = ps.worldCenter + ps.worldRadius * (u * du2 + v * dv2 + ldir);
        ray.dir = -ldir;
        const float ePdf = (float)YB_M_PI * ps.worldRadius * ps.worldRadius;
        return l.colPdf * ePdf;
    }
    case LIGHT_AREA:                                         // arealight.cc:98-104
        ipdf = l.area;
        ray.from = l.corner + s3 * l.toX + s4 * l.toY;
        ray.dir = SampleCosHemisphere(l.normal, l.du, l.dv, s1, s2);
        return l.color;
    case LIGHT_DIRECTIONAL: {                                // directional.cc:106-116 (init :63-76: an infinite light covers the scene)
        float u, v;
        ray.dir = -l.direction;
        ShirleyDisk(s1, s2, u, v);
        const v3 pos = l.infinite ? ps.worldCenter : l.position;
        const float radius = l.infinite ? ps.worldRadius : l.radius;
        ray.from = pos + radius * (u * l.du + v * l.dv);
        if (l.infinite) ray.from = ray.from + l.direction * ps.worldRadius;
        ipdf = (float)YB_M_PI * radius * radius;
        return l.color;
    }
    case LIGHT_SPOT: {                                       // spotlight.cc:177-198
        ray.from = l.position;
        double interv1 = 1.0 - (double)l.cosStart, interv2 = 0.5 * ((double)l.cosStart - (double)l.cosEnd);     // spotlight.cc:93-100
        float sum = fabsf((float)interv1) + fabsf((float)interv2);
        if (sum > 0.f) sum = (float)(1.0 / sum);
        const float i1 = (float)(interv1 * sum), i2 = (float)(interv2 * sum);
        if (s3 <= i1) {
            ray.dir = sampleCone(l.direction, l.du, l.dv, l.cosStart, s1, s2);
            ipdf = (float)YB_M_2PI * (1.f - l.cosStart) / i1;
            return l.color;
        }
        float spdf, integral;
        const float sm2 = spot_pdfSample(s2, spdf, integral) * (1.f / 65.f);
        ipdf = (float)YB_M_2PI * (l.cosStart - l.cosEnd) / (i2 * spdf);
        const double cosAng = l.cosEnd + (l.cosStart - l.cosEnd) * (double)sm2;
        const double sinAng = fSqrt((float)(1.0 - cosAng * cosAng));
        const float t1 = (float)YB_M_2PI * s1;
        ray.dir = (l.du * fCos(t1) + l.dv * fSin(t1)) * (float)sinAng + l.direction * (float)cosAng;
        return l.color * (spdf * integral);
    }
    case LIGHT_SPHERE: {                                     // spherelight.cc:169-178
        const v3 sdir = SampleSphere(s3, s4);
        ray.from = l.position + l.radius * sdir;
        v3 du, dv;
        createCS(sdir, du, dv);
        ray.dir = SampleCosHemisphere(sdir, du, dv, s1, s2);
        ipdf = l.area;
        return l.color;
    }
    default: {                                               // LIGHT_MESH: meshlight.cc:140-155
        v3 normal(0.f), du, dv;
        ipdf = l.area;
        mesh_sampleSurface(l, ray.from, normal, s3, s4);
        createCS(normal, du, dv);
        if (l.doubleSided) {
            ipdf *= 2.f;
            if (s1 > 0.5f) ray.dir = SampleCosHemisphere(-normal, du, dv, (s1 - 0.5f) * 2.f, s2);
            else ray.dir = SampleCosHemisphere(normal, du, dv, s1 * 2.f, s2);
        } else ray.dir = SampleCosHemisphere(normal, du, dv, s1, s2);
        return l.color;
    }
    }
}
// light_t::totalEnergy().energy() (color.h: energy = (R + G + B) * 0.333333)
static float light_totalEnergy(const Light &l, const PhotonScene &ps)
{
    col3 e = l.color;
    switch (l.type) {
    case LIGHT_POINT: e = l.color * (4.0f * (float)YB_M_PI); break;                                          // pointlight.cc:32
    case LIGHT_SUN: e = l.color * ((float)YB_M_PI * ps.worldRadius * ps.worldRadius); break;                  // sunlight.cc:32
    case LIGHT_AREA: e = l.color * l.area; break;                                                             // arealight.cc:66
    case LIGHT_DIRECTIONAL: { const float r = l.infinite ? ps.worldRadius : l.radius; e = l.color * (r * r * (float)YB_M_PI); break; }   // directional.cc:33
    case LIGHT_SPOT: e = l.color * ((float)YB_M_2PI * (1.f - 0.5f * (l.cosStart + l.cosEnd))); break;          // spotlight.cc:108-111
    case LIGHT_SPHERE: e = l.color * l.area; break;                                                           // spherelight.cc:86
    default: e = l.doubleSided ? (l.color * l.area) * 2.f : l.color * l.area; break;                          // meshlight.cc:100
    }
    return (e.x + e.y + e.z) * 0.333333f;
}

// ------------------------------------------------------------------------------------------------ createCausticMap
struct ShootArgs {
    SceneDev sc; PhotonScene ps;
    const float *lightFunc, *lightCdf;       // pdf1D_t over the lights' energies (func[n], cdf[n + 1])
    float lightInvIntegral; int32_t nLights;
    uint32_t nPhotons; int32_t causDepth;
    Photon *photons; uint32_t cap;
    uint32_t *counters;                      // [0] photons stored, [1] max path index that stored one
};

template <int EXT>
YB_D void photon_path(const ShootArgs &a, uint32_t curr)
{
    const float s1 = RI_vdC(curr), s2 = (float)scrHalton(2, curr), s3 = (float)scrHalton(3, curr), s4 = (float)scrHalton(4, curr);
    const float sL = (float)curr / (float)a.nPhotons;
    // pdf1D_t::DSample (sample_utils.h:141-157)
    int lightNum = 0;
    if (sL != 0.f) {
        int lo = 0, hi = a.nLights + 1;
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (a.lightCdf[mid] < sL) lo = mid + 1; else hi = mid; }
        lightNum = lo - 1;
        if (lightNum < 0) lightNum = 0;
    }
    if (lightNum >= a.nLights) return;
    const float lightNumPdf = a.lightFunc[lightNum] * a.lightInvIntegral;
    Ray ray; float lightPdf = 0.f;
    ray.tmax = -1.f;
    col3 pcol = light_emitPhoton(a.sc.lights[lightNum], a.ps, s1, s2, s3, s4, ray, lightPdf);
    ray.tmin = YB_MIN_RAYDIST; ray.tmax = -1.f;
    pcol *= (float)a.nLights * lightPdf / lightNumPdf;
    if (isBlack(pcol)) return;
    int nBounces = 0;
    bool causticPhoton = false, directPhoton = true;
    bool prevBeer = false; col3 prevSigma(0.f); v3 prevNg(0.f);           // the material the segment leaves: its inside absorbs (beer volume)
    Hit H;
    while (scene_intersect(a.sc, ray, H)) {
        if (pcol.x != pcol.x || pcol.y != pcol.y || pcol.z != pcol.z) break;
        // mcintegrator.cc:294-304: transmittance of the segment just traversed, by the volume handler of the material it left
        col3 transm(1.f);
        if (prevBeer && dot(prevNg, ray.dir) < 0.f) transm = beer_transmittance(prevSigma, true, H.t);
        Surf sp;
        tri_getSurface(a.sc, H.tri, ray.from + H.t * ray.dir, H.b1, H.b2, sp);
        Material rmat;
        if (EXT >= 2 && sp.mat->nNodes > 0) {
            TexPoint tp;
            tri_getTexPoint(a.sc, H.tri, sp.P, H.b1, H.b2, sp.Ng, tp, &sp);
            sp.mat = &mat_resolve<EXT>(a.sc, *sp.mat, sp.P, sp.Ng, tp, rmat, &sp);
        }
        const Material &mat = *sp.mat;
        MatDat md; uint32_t bsdfs;
        mat_initBSDF<EXT>(mat, md, bsdfs);
        mat_blendSurface<EXT>(mat, sp);
        const v3 wi = -ray.dir;
        if ((bsdfs & (BSDF_DIFFUSE | BSDF_GLOSSY)) && causticPhoton) {           // deposit (mcintegrator.cc:313-322)
            const uint32_t slot = YB_ATOMIC_ADD_U32(a.counters + 0, 1u);
            if (slot < a.cap) { Photon p; p.pos = sp.P; p.dir = wi; p.c = pcol; a.photons[slot] = p; }
            YB_ATOMIC_MAX_I32(a.counters + 1, (int)curr);                        // causticMap.setNumPaths(curr)
        }
        if (nBounces == a.causDepth) break;
        const int d5 = 3 * nBounces + 5;
        const float s5 = (float)scrHalton(d5, curr), s6 = (float)scrHalton(d5 + 1, curr), s7 = (float)scrHalton(d5 + 2, curr);
        // material_t::scatterPhoton (material.cc:31-48) with BSDF_ALL_SPECULAR | BSDF_GLOSSY | BSDF_FILTER | BSDF_DISPERSIVE
        BSample s(s5, s6, BSDF_SPECULAR | BSDF_REFLECT | BSDF_TRANSMIT | BSDF_GLOSSY | BSDF_FILTER | BSDF_DISPERSIVE);
        float W = 0.f; v3 wo(0.f);
        const col3 scol = mat_sample<EXT>(mat, md, sp, wi, wo, s, W);
        if (!(s.pdf > 1.0e-6f)) break;
        const col3 cnew = mul(mul(pcol, transm), scol) * W;                      // lcol * alpha * scol * W
        const float new_max = fmaxf_(cnew.x, fmaxf_(cnew.y, cnew.z)), old_max = fmaxf_(pcol.x, fmaxf_(pcol.y, pcol.z));
        const float prob = fminf_(1.f, new_max / old_max);
        if (!(s7 <= prob && prob > 1e-4f)) break;                                // absorbed
        pcol = col3(cnew.x / prob, cnew.y / prob, cnew.z / prob);               // color_t::operator/ (color.h:252-255)
        const uint32_t sf = s.sampledFlags;
        const bool nc = ((sf & (BSDF_GLOSSY | BSDF_SPECULAR | BSDF_DISPERSIVE)) && directPhoton) ||
                        ((sf & (BSDF_GLOSSY | BSDF_SPECULAR | BSDF_FILTER | BSDF_DISPERSIVE)) && causticPhoton);
        causticPhoton = nc;
        directPhoton = (sf & BSDF_FILTER) && directPhoton;
        if (!(causticPhoton || directPhoton)) break;
        prevBeer = (bsdfs & BSDF_VOLUMETRIC) && mat.hasBeer; prevSigma = mat.beerSigma; prevNg = sp.Ng;
        ray.from = sp.P; ray.dir = wo; ray.tmin = YB_MIN_RAYDIST; ray.tmax = -1.f;
        ++nBounces;
    }
}
YB_KERNEL(k_photon_shoot, ShootArgs) { photon_path<2>(a, tid); }

// ------------------------------------------------------------------------------------------------ grid
struct GridArgs { PhotonGrid g; const Photon *in; Photon *out; uint32_t *cellOf, *index; uint32_t n; };
YB_HD uint32_t grid_cell(const PhotonGrid &g, v3 p)
{
    int ix = (int)floorf((p.x - g.lo.x) * g.invCell), iy = (int)floorf((p.y - g.lo.y) * g.invCell), iz = (int)floorf((p.z - g.lo.z) * g.invCell);
    ix = ix < 0 ? 0 : (ix >= g.nx ? g.nx - 1 : ix); iy = iy < 0 ? 0 : (iy >= g.ny ? g.ny - 1 : iy); iz = iz < 0 ? 0 : (iz >= g.nz ? g.nz - 1 : iz);
    return (uint32_t)ix + (uint32_t)g.nx * ((uint32_t)iy + (uint32_t)g.ny * (uint32_t)iz);
}
YB_KERNEL(k_photon_cell, GridArgs) { a.cellOf[tid] = grid_cell(a.g, a.in[tid].pos); a.index[tid] = tid; }
YB_KERNEL(k_photon_reorder, GridArgs) { a.out[tid] = a.in[a.index[tid]]; }
// cellStart[c] = first sorted photon whose cell is >= c (cellOf is sorted): one thread per cell boundary
struct StartArgs { const uint32_t *cellOf; uint32_t *cellStart; uint32_t n, nCells; };
YB_KERNEL(k_photon_starts, StartArgs)
{
    // tid in [0, n]: photon tid starts the cells (prevCell, cell]
    const uint32_t prev = tid == 0 ? 0u : a.cellOf[tid - 1] + 1u;
    const uint32_t cur = tid == a.n ? a.nCells : a.cellOf[tid];
    for (uint32_t c = prev; c <= cur; ++c) a.cellStart[c] = tid;
}

// ------------------------------------------------------------------------------------------------ estimateCausticPhotons
YB_HD float photon_kernel(float r_photon2, float ir_gather2)    // sample_utils.h:27-31
{
    const float s = (1.f - r_photon2 * ir_gather2);
    return 3.f * ir_gather2 * (float)YB_M_1_PI * s * s;
}
struct GatherArgs { RenderArgs r; uint32_t nTasks; };
template <int EXT>
YB_D void caustic_gather(const RenderArgs &a, uint32_t t)
{
    const Task T = a.tasks[t];
    const Hit H = a.hits[t];
    if (task_kind(T.bits) != TASK_INTEGRATE || H.tri < 0) return;
    const PhotonGrid &g = a.sc.photonGrid;
    Surf sp;
    tri_getSurface(a.sc, H.tri, T.ray.from + H.t * T.ray.dir, H.b1, H.b2, sp);
    Material rmat;
    if (EXT >= 2 && sp.mat->nNodes > 0) {
        TexPoint tp;
        tri_getTexPoint(a.sc, H.tri, sp.P, H.b1, H.b2, sp.Ng, tp, &sp);
        sp.mat = &mat_resolve<EXT>(a.sc, *sp.mat, sp.P, sp.Ng, tp, rmat, &sp);
    }
    const Material &mat = *sp.mat;
    MatDat md; uint32_t bsdfs;
    mat_initBSDF<EXT>(mat, md, bsdfs);
    mat_blendSurface<EXT>(mat, sp);
    if (!(bsdfs & BSDF_DIFFUSE)) return;
    const v3 wo = -T.ray.dir;
    // photonMap_t::gather: the K nearest photons with dist2 < maxDistSquared; maxDistSquared drops to the farthest kept one
    // as soon as K are held (photonGather_t::operator(), photon.cc:46-68)
    float hd[YB_MAX_CAUS_SEARCH]; uint32_t hi[YB_MAX_CAUS_SEARCH];
    const int K = g.nSearch;
    int found = 0; float maxD2 = g.radius * g.radius; int worst = 0;
    const int cx = (int)floorf((sp.P.x - g.lo.x) * g.invCell), cy = (int)floorf((sp.P.y - g.lo.y) * g.invCell), cz = (int)floorf((sp.P.z - g.lo.z) * g.invCell);
    for (int dz = -1; dz <= 1; ++dz)
        for (int dy = -1; dy <= 1; ++dy) {
            const int iy = cy + dy, iz = cz + dz;
            if (iy < 0 || iy >= g.ny || iz < 0 || iz >= g.nz) continue;
            int x0 = cx - 1, x1 = cx + 1;
            if (x0 < 0) x0 = 0;
            if (x1 >= g.nx) x1 = g.nx - 1;
            if (x0 > x1) continue;
            const uint32_t row = (uint32_t)g.nx * ((uint32_t)iy + (uint32_t)g.ny * (uint32_t)iz);
            const uint32_t p0 = g.cellStart[row + (uint32_t)x0], p1 = g.cellStart[row + (uint32_t)x1 + 1u];     // the three cells of a row are contiguous
            for (uint32_t p = p0; p < p1; ++p) {
                const v3 v = g.photons[p].pos - sp.P;
                const float d2 = v.x * v.x + v.y * v.y + v.z * v.z;
                if (!(d2 < maxD2)) continue;
                if (found < K) {
                    hd[found] = d2; hi[found] = p; ++found;
                    if (found == K) { worst = 0; for (int q = 1; q < K; ++q) if (hd[q] > hd[worst]) worst = q; maxD2 = hd[worst]; }
                } else {
                    hd[worst] = d2; hi[worst] = p;
                    worst = 0; for (int q = 1; q < K; ++q) if (hd[q] > hd[worst]) worst = q;
                    maxD2 = hd[worst];
                }
            }
        }
    if (found == 0) return;
    const float ir2 = 1.f / maxD2;
    col3 sum(0.f);
    for (int i = 0; i < found; ++i) {
        const Photon ph = g.photons[hi[i]];
        const col3 surfCol = mat_eval<EXT>(mat, md, sp, wo, ph.dir, BSDF_ALL);
        const float k = photon_kernel(hd[i], ir2);
        sum += mul(surfCol * k, ph.c);
    }
    sum *= 1.f / (float)g.nPaths;
    acc_add(a, T.sid, mul(T.thr, sum));
}
YB_KERNEL(k_caustic_gather, RenderArgs)
{
    if (a.nPtr && tid >= *a.nPtr) return;
    caustic_gather<2>(a, tid);
}
void launch_caustic_gather(uint32_t n, const RenderArgs &a, rt_stream s) { k_caustic_gather(n, a, s); }

// ------------------------------------------------------------------------------------------------ host
struct PhotonBuffers {
    DevBuf<Photon> raw, sorted;
    DevBuf<uint32_t> counters, cellOf, cellOf2, index, index2, cellStart;
    DevBuf<float> lightFunc, lightCdf;
    DevBuf<char> sortTmp;
};
void photon_release(PhotonBuffers *p) { delete p; }

// scene_t::getSceneBound = the kd-tree's bound: the box of all triangles grown by 0.1 % per side (kdtree.cc:109-115)
static void scene_bound(const Context &c, v3 &lo, v3 &hi)
{
    lo = c.boundLo; hi = c.boundHi;
    float *a[3] = {&lo.x, &lo.y, &lo.z}, *g[3] = {&hi.x, &hi.y, &hi.z};
    for (int i = 0; i < 3; ++i) { const double foo = (double)(*g[i] - *a[i]) * 0.001; *a[i] = (float)((double)*a[i] - foo); *g[i] = (float)((double)*g[i] + foo); }
}

bool build_caustic_map(Context &c)
{
    const Integrator &I = c.sc.integ;
    c.sc.photonGrid.nPhotons = 0; c.sc.photonGrid.photons = 0;
    if (!I.usePhotonCaustics) return true;
    if (I.nCausSearch > YB_MAX_CAUS_SEARCH || I.nCausSearch < 1)
        return c.fail("caustic_mix (photons per estimate) must be between 1 and " + std::to_string(YB_MAX_CAUS_SEARCH));
    if (c.lights.empty() || I.nCausPhotons <= 0 || c.nTris == 0) return true;       // "No caustic source lights found"
    for (size_t i = 0; i < c.materials.size(); ++i) {
        if (c.materials[i].type == MAT_BLEND) return c.fail("caustic photon maps: blend_mat scatters photons through both sub-materials (blend.cc:359-380), which is not supported");
    }
    try {
        if (!c.photonBufs) c.photonBufs = new PhotonBuffers();
        PhotonBuffers &pb = *c.photonBufs;
        rt_stream s = c.stream;
        v3 lo, hi;
        scene_bound(c, lo, hi);
        PhotonScene ps;
        ps.worldRadius = 0.5f * length(hi - lo);
        ps.worldCenter = 0.5f * (lo + hi);
        // pdf1D_t over the lights' energies (mcintegrator.cc:218-224, sample_utils.h:85-116)
        const int nl = (int)c.lights.size();
        std::vector<float> func(nl), cdf(nl + 1);
        for (int i = 0; i < nl; ++i) func[i] = light_totalEnergy(c.lights[i], ps);
        double acc = 0.0; const double delta = 1.0 / (double)nl;
        cdf[0] = 0.f;
        for (int i = 1; i < nl + 1; ++i) { acc += (double)func[i - 1] * delta; cdf[i] = (float)acc; }
        const float integral = (float)acc;
        for (int i = 1; i < nl + 1; ++i) cdf[i] /= integral;
        pb.lightFunc.reserve(nl + 1); pb.lightCdf.reserve(nl + 2);
        rt_h2d(pb.lightFunc.p, func.data(), nl * sizeof(float), s);
        rt_h2d(pb.lightCdf.p, cdf.data(), (nl + 1) * sizeof(float), s);
        // a path deposits at most once per diffuse / glossy hit after a specular one; more than a few per path is rare
        const size_t cap = std::min<size_t>((size_t)I.nCausPhotons * 4 + 1024, (size_t)1 << 28);
        pb.raw.reserve(cap); pb.counters.reserve(4);
        rt_memset(pb.counters.p, 0, 4 * sizeof(uint32_t), s);
        ShootArgs sa;
        memset((void *)&sa, 0, sizeof(sa));
        sa.sc = c.sc; sa.ps = ps; sa.lightFunc = pb.lightFunc.p; sa.lightCdf = pb.lightCdf.p; sa.lightInvIntegral = 1.f / integral; sa.nLights = nl;
        sa.nPhotons = (uint32_t)I.nCausPhotons; sa.causDepth = I.causDepth; sa.photons = pb.raw.p; sa.cap = (uint32_t)cap; sa.counters = pb.counters.p;
        k_photon_shoot((uint32_t)I.nCausPhotons, sa, s);
        uint32_t cnt[4];
        rt_d2h(cnt, pb.counters.p, sizeof(cnt), s);
        if (cnt[0] > cap) return c.fail("caustic photon map: more photons stored than the buffer holds (" + std::to_string(cnt[0]) + ")");
        const uint32_t n = cnt[0];
        c.stats.caustic_photons = n;
        if (n == 0) return true;                                                   // nothing stored: causticMap.ready() stays false
        // ---- grid with cells of the gather radius over the stored photons
        std::vector<Photon> hp;
        PhotonGrid g;
        memset((void *)&g, 0, sizeof(g));
        g.radius = I.causRadius; g.nSearch = I.nCausSearch; g.nPaths = (int32_t)cnt[1]; g.nPhotons = n;
        g.lo = lo;
        float cell = I.causRadius > 0.f ? I.causRadius : 1e-3f;
        for (;;) {                                                                 // at most 2^24 cells: coarser cells are still >= the radius
            const double nx = floor((double)(hi.x - lo.x) / cell) + 1, ny = floor((double)(hi.y - lo.y) / cell) + 1, nz = floor((double)(hi.z - lo.z) / cell) + 1;
            if (nx * ny * nz <= (double)(1 << 24)) { g.nx = (int32_t)nx; g.ny = (int32_t)ny; g.nz = (int32_t)nz; break; }
            cell *= 1.26f;
        }
        g.invCell = 1.f / cell;
        const uint32_t nCells = (uint32_t)g.nx * (uint32_t)g.ny * (uint32_t)g.nz;
        pb.sorted.reserve(n); pb.cellOf.reserve(n + 1); pb.cellOf2.reserve(n + 1); pb.index.reserve(n + 1); pb.index2.reserve(n + 1); pb.cellStart.reserve((size_t)nCells + 2);
        GridArgs ga; ga.g = g; ga.in = pb.raw.p; ga.out = pb.sorted.p; ga.cellOf = pb.cellOf.p; ga.index = pb.index.p; ga.n = n;
        k_photon_cell(n, ga, s);
#ifndef YB_HOSTEMU
        size_t tmpBytes = 0;
        YB_CUDA_CHECK(cub::DeviceRadixSort::SortPairs(0, tmpBytes, pb.cellOf.p, pb.cellOf2.p, pb.index.p, pb.index2.p, (int)n, 0, 24, s));
        pb.sortTmp.reserve(tmpBytes + 16);
        YB_CUDA_CHECK(cub::DeviceRadixSort::SortPairs(pb.sortTmp.p, tmpBytes, pb.cellOf.p, pb.cellOf2.p, pb.index.p, pb.index2.p, (int)n, 0, 24, s));
#else
        {
            std::vector<uint32_t> order(n);
            for (uint32_t i = 0; i < n; ++i) order[i] = i;
            std::stable_sort(order.begin(), order.end(), [&](uint32_t x, uint32_t y) { return pb.cellOf.p[x] < pb.cellOf.p[y]; });
            for (uint32_t i = 0; i < n; ++i) { pb.cellOf2.p[i] = pb.cellOf.p[order[i]]; pb.index2.p[i] = order[i]; }
        }
#endif
        ga.index = pb.index2.p;
        k_photon_reorder(n, ga, s);
        StartArgs st; st.cellOf = pb.cellOf2.p; st.cellStart = pb.cellStart.p; st.n = n; st.nCells = nCells;
        k_photon_starts(n + 1, st, s);
        rt_sync(s);
        g.photons = pb.sorted.p; g.cellStart = pb.cellStart.p;
        c.sc.photonGrid = g;
    } catch (const RtError &e) { return c.fail(std::string("caustic photon map failed: ") + e.what()); }
    return true;
}

}  // namespace yb
