// yb_nodes.cuh — device restatement of the shader-node tree of nodeMaterial_t for the node and texture types listed in
// yb_types.h:
//   textureMapper_t / valueNode_t / mixNode_t family   src/textures/basicnodes.cc:96-160,312-600
//   layerNode_t                                        src/textures/layernode.cc:12-110, include/core_api/shader.h:112-225
//   rgbCube_t / textureBlend_t / textureClouds_t / textureMarble_t / textureWood_t   src/textures/basictex.cc:44-244,478-502
//   newPerlin_t + turbulence                           src/textures/noise.cc:161-183,757-769, include/textures/noise.h:46-53
// Every supported node is view independent, so a material's nodes are evaluated once per surface point (what initBSDF does
// with allViewindep, e.g. shinydiffuse.cc:159-161) and the results replace the constants of a LOCAL copy of the Material
// record (mat_resolve at the end of yb_bsdf.cuh); the BSDF code then runs unchanged.  Bump and normal mapping: node_evalDerivative + the frame update in mat_resolve.
#pragma once
#include "yb_types.h"

namespace yb {

#if defined(__CUDACC__) && !defined(YB_HOSTEMU)
#define YB_NTABLE static __device__ __constant__
#else
#define YB_NTABLE static
#endif
#include "yb_noise_tab.inc"
// texture / node evaluation is called from many places of the (fully inlined) shade kernel with small by-value arguments:
// out of line it keeps the EXT = 2 build at a compilable size (-DYB_INLINE_NODES restores inlining for A/B runs)
#if defined(__CUDACC__) && !defined(YB_HOSTEMU) && !defined(YB_INLINE_NODES)
#define YB_HD_NODE static __host__ __device__ __noinline__
#else
#define YB_HD_NODE YB_HD
#endif

struct NodeRes { float r, g, b, a, f; };          // nodeResult_t: colorA_t + scalar (shader.h:32-38)
struct TexPoint { float U, V; v3 orcoP, orcoNg, dSdU, dSdV; };   // the surfacePoint_t fields texture_mapper reads besides P / Ng / NU / NV

// ---- noise ------------------------------------------------------------------------------------------
YB_HD int nperm(int i) { return (int)yb_noise_perm[i & 255]; }     // hash[512] holds the permutation twice
YB_HD float np_fade(float t) { return t * t * t * (t * (t * 6 - 15) + 10); }
YB_HD float np_grad(int hash, float x, float y, float z)
{
    int h = hash & 15;
    float u = h < 8 ? x : y, v = h < 4 ? y : (h == 12 || h == 14) ? x : z;
    return ((h & 1) == 0 ? u : -u) + ((h & 2) == 0 ? v : -v);
}
YB_HD float np_lerp(float t, float a, float b) { return a + t * (b - a); }
YB_HD float newPerlin(v3 pt)
{
    float x = pt.x, y = pt.y, z = pt.z;
    float u = floorf(x), v = floorf(y), w = floorf(z);
    int X = ((int)u) & 255, Y = ((int)v) & 255, Z = ((int)w) & 255;
    x -= u; y -= v; z -= w;
    u = np_fade(x); v = np_fade(y); w = np_fade(z);
    int A = nperm(X) + Y, AA = nperm(A) + Z, AB = nperm(A + 1) + Z, B = nperm(X + 1) + Y, BA = nperm(B) + Z, BB = nperm(B + 1) + Z;
    float nv = np_lerp(w, np_lerp(v, np_lerp(u, np_grad(nperm(AA), x, y, z), np_grad(nperm(BA), x - 1, y, z)),
                                     np_lerp(u, np_grad(nperm(AB), x, y - 1, z), np_grad(nperm(BB), x - 1, y - 1, z))),
                          np_lerp(v, np_lerp(u, np_grad(nperm(AA + 1), x, y, z - 1), np_grad(nperm(BA + 1), x - 1, y, z - 1)),
                                     np_lerp(u, np_grad(nperm(AB + 1), x, y - 1, z - 1), np_grad(nperm(BB + 1), x - 1, y - 1, z - 1))));
    return (float)(0.5 + 0.5 * (double)nv);
}
// blenderNoise_t (noise.cc:373-447)
YB_HD float blenderNoise(v3 pt)
{
    float cn1, cn2, cn3, cn4, cn5, cn6, i;
    const float *h;
    float ox, oy, oz, jx, jy, jz;
    float n = 0.5f;
    int ix, iy, iz, b00, b01, b10, b11, b20, b21;
    const float x = pt.x, y = pt.y, z = pt.z;
    ox = (x - (float)(ix = (int)floorf(x)));
    oy = (y - (float)(iy = (int)floorf(y)));
    oz = (z - (float)(iz = (int)floorf(z)));
    jx = ox - 1; jy = oy - 1; jz = oz - 1;
    cn1 = ox * ox; cn2 = oy * oy; cn3 = oz * oz;
    cn4 = jx * jx; cn5 = jy * jy; cn6 = jz * jz;
    cn1 = (float)(1.0 - 3.0 * (double)cn1 + 2.0 * (double)cn1 * (double)ox);
    cn2 = (float)(1.0 - 3.0 * (double)cn2 + 2.0 * (double)cn2 * (double)oy);
    cn3 = (float)(1.0 - 3.0 * (double)cn3 + 2.0 * (double)cn3 * (double)oz);
    cn4 = (float)(1.0 - 3.0 * (double)cn4 - 2.0 * (double)cn4 * (double)jx);
    cn5 = (float)(1.0 - 3.0 * (double)cn5 - 2.0 * (double)cn5 * (double)jy);
    cn6 = (float)(1.0 - 3.0 * (double)cn6 - 2.0 * (double)cn6 * (double)jz);
    b00 = nperm(nperm(ix & 255) + (iy & 255));
    b10 = nperm(nperm((ix + 1) & 255) + (iy & 255));
    b01 = nperm(nperm(ix & 255) + ((iy + 1) & 255));
    b11 = nperm(nperm((ix + 1) & 255) + ((iy + 1) & 255));
    b20 = iz & 255; b21 = (iz + 1) & 255;
    i = (cn1 * cn2 * cn3); h = yb_noise_vect + 3 * nperm(b20 + b00); n += i * (h[0] * ox + h[1] * oy + h[2] * oz);
    i = (cn1 * cn2 * cn6); h = yb_noise_vect + 3 * nperm(b21 + b00); n += i * (h[0] * ox + h[1] * oy + h[2] * jz);
    i = (cn1 * cn5 * cn3); h = yb_noise_vect + 3 * nperm(b20 + b01); n += i * (h[0] * ox + h[1] * jy + h[2] * oz);
    i = (cn1 * cn5 * cn6); h = yb_noise_vect + 3 * nperm(b21 + b01); n += i * (h[0] * ox + h[1] * jy + h[2] * jz);
    i = cn4 * cn2 * cn3; h = yb_noise_vect + 3 * nperm(b20 + b10); n += i * (h[0] * jx + h[1] * oy + h[2] * oz);
    i = cn4 * cn2 * cn6; h = yb_noise_vect + 3 * nperm(b21 + b10); n += i * (h[0] * jx + h[1] * oy + h[2] * jz);
    i = cn4 * cn5 * cn3; h = yb_noise_vect + 3 * nperm(b20 + b11); n += i * (h[0] * jx + h[1] * jy + h[2] * oz);
    i = (cn4 * cn5 * cn6); h = yb_noise_vect + 3 * nperm(b21 + b11); n += i * (h[0] * jx + h[1] * jy + h[2] * jz);
    if (n < 0.0f) n = 0.0f; else if (n > 1.0f) n = 1.0f;
    return n;
}
// distance metrics of voronoi_t (noise.cc:335-370)
YB_HD float voronoi_dist(int dm, float x, float y, float z, float e)
{
    switch (dm) {
        case 1: return x * x + y * y + z * z;
        case 2: { x = fabsf(x); y = fabsf(y); z = fabsf(z); const float t = (x > y) ? x : y; return (z > t) ? z : t; }
        case 3: { const float d = fSqrt(fabsf(x)) + fSqrt(fabsf(y)) + fSqrt(fabsf(z)); return d * d; }
        case 4: { x *= x; y *= y; z *= z; return fSqrt(fSqrt(x * x + y * y + z * z)); }
        case 5: return fPow(fPow(fabsf(x), e) + fPow(fabsf(y), e) + fPow(fabsf(z), e), 1.0f / e);
        default: return fSqrt(x * x + y * y + z * z);
    }
}
// voronoi_t::getFeatures (noise.cc:498-541): the four nearest feature points and their distances
YB_HD void voronoi_features(v3 pt, int dm, float mex, float *da, v3 *pa)
{
    const float x = pt.x, y = pt.y, z = pt.z;
    const int xi = (int)floorf(x), yi = (int)floorf(y), zi = (int)floorf(z);
    da[0] = da[1] = da[2] = da[3] = 1e10f;
    pa[0] = pa[1] = pa[2] = pa[3] = v3(0.f);
    for (int xx = xi - 1; xx <= xi + 1; xx++)
        for (int yy = yi - 1; yy <= yi + 1; yy++)
            for (int zz = zi - 1; zz <= zi + 1; zz++) {
                const float *p = yb_noise_pnt + 3 * nperm(nperm(nperm(zz & 255) + yy) + xx);   // HASHPNT (noise.cc:6)
                const float xd = x - (p[0] + (float)xx), yd = y - (p[1] + (float)yy), zd = z - (p[2] + (float)zz);
                const float d = voronoi_dist(dm, xd, yd, zd, mex);
                const v3 q(p[0] + (float)xx, p[1] + (float)yy, p[2] + (float)zz);
                if (d < da[0]) { da[3] = da[2]; da[2] = da[1]; da[1] = da[0]; da[0] = d; pa[3] = pa[2]; pa[2] = pa[1]; pa[1] = pa[0]; pa[0] = q; }
                else if (d < da[1]) { da[3] = da[2]; da[2] = da[1]; da[1] = d; pa[3] = pa[2]; pa[2] = pa[1]; pa[1] = q; }
                else if (d < da[2]) { da[3] = da[2]; da[2] = d; pa[3] = pa[2]; pa[2] = q; }
                else if (d < da[3]) { da[3] = d; pa[3] = q; }
            }
}
// voronoi_t::operator() as a noise basis: real distance (noise.cc:543-565)
YB_HD float voronoiNoise(v3 pt, int vType)
{
    float da[4]; v3 pa[4];
    voronoi_features(pt, 0, 2.5f, da, pa);
    switch (vType) {
        case 1: return da[1];
        case 2: return da[2];
        case 3: return da[3];
        case 4: return da[1] - da[0];
        case 5: { const float t = (float)(10.0 * (double)(da[1] - da[0])); return (t > 1.0f) ? 1.0f : t; }
        default: return da[0];
    }
}
YB_HD float cellNoise(v3 pt)                             // noise.cc:568-576
{
    const int xi = (int)floorf(pt.x), yi = (int)floorf(pt.y), zi = (int)floorf(pt.z);
    unsigned int n = (unsigned int)(xi + yi * 1301 + zi * 314159);
    n ^= (n << 13);
    return (float)((double)(float)(n * (n * n * 15731u + 789221u) + 1376312589u) / 4294967296.0);
}
// stdPerlin_t (noise.cc:289-345)
YB_HD void stdp_setup(float v, int &b0, int &b1, float &r0, float &r1)
{
    const float t = (float)((double)v + 10000.0);
    b0 = ((int)t) & 255; b1 = (b0 + 1) & 255;
    r0 = t - (float)(int)t; r1 = (float)((double)r0 - 1.0);
}
YB_HD float stdp_surve(float t) { return (float)((double)t * (double)t * (3.0 - 2.0 * (double)t)); }
YB_HD float stdp_at(int gi, float rx, float ry, float rz) { const float *q = yb_stdp_g + 3 * gi; return rx * q[0] + ry * q[1] + rz * q[2]; }
YB_HD float stdPerlin(v3 pt)
{
    int bx0, bx1, by0, by1, bz0, bz1;
    float rx0, rx1, ry0, ry1, rz0, rz1;
    stdp_setup(pt.x, bx0, bx1, rx0, rx1); stdp_setup(pt.y, by0, by1, ry0, ry1); stdp_setup(pt.z, bz0, bz1, rz0, rz1);
    const int i = yb_stdp_p[bx0], j = yb_stdp_p[bx1];
    const int b00 = yb_stdp_p[i + by0], b10 = yb_stdp_p[j + by0], b01 = yb_stdp_p[i + by1], b11 = yb_stdp_p[j + by1];
    const float sx = stdp_surve(rx0), sy = stdp_surve(ry0), sz = stdp_surve(rz0);
    float u, v, a, b;
    u = stdp_at(b00 + bz0, rx0, ry0, rz0); v = stdp_at(b10 + bz0, rx1, ry0, rz0); a = np_lerp(sx, u, v);
    u = stdp_at(b01 + bz0, rx0, ry1, rz0); v = stdp_at(b11 + bz0, rx1, ry1, rz0); b = np_lerp(sx, u, v);
    const float c = np_lerp(sy, a, b);
    u = stdp_at(b00 + bz1, rx0, ry0, rz1); v = stdp_at(b10 + bz1, rx1, ry0, rz1); a = np_lerp(sx, u, v);
    u = stdp_at(b01 + bz1, rx0, ry1, rz1); v = stdp_at(b11 + bz1, rx1, ry1, rz1); b = np_lerp(sx, u, v);
    const float d = np_lerp(sy, a, b);
    return (float)(0.5 + 0.75 * (double)np_lerp(sz, c, d));
}
YB_HD float noise_gen(int noiseType, v3 p)
{
    if (noiseType == 0) return newPerlin(p);
    if (noiseType == 1) return blenderNoise(p);
    if (noiseType == 2) return cellNoise(p);
    if (noiseType == 9) return stdPerlin(p);
    return voronoiNoise(p, noiseType - 3);
}
YB_HD float turbulence(int noiseType, v3 pt, int oct, float size, bool hard)
{
    float val, amp = 1, sum = 0;
    v3 tp = (noiseType == 1 ? pt + v3(1.0f) : pt) * size;      // only blenderNoise_t::offset adds (1, 1, 1)
    for (int i = 0; i <= oct; i++, amp *= 0.5f, tp *= 2.0f) {
        val = noise_gen(noiseType, tp);
        if (hard) val = fabsf((float)(2.0 * (double)val - 1.0));
        sum += amp * val;
    }
    return sum * ((float)(1 << oct) / (float)((1 << (oct + 1)) - 1));
}
YB_HD float signedNoise(int nt, v3 p) { return 2.0f * noise_gen(nt, p) - 1.0f; }          // noise.h:259-262
// the five musgrave generators (noise.cc:594-745)
YB_HD float musgrave(const Texture &t, v3 pt)
{
    const int nt = t.noiseType;
    const float H = t.musH, lacunarity = t.musLacu, octaves = t.musOcts, offset = t.musOffs, gain = t.musGain;
    const float pwHL = fPow(lacunarity, -H);
    v3 tp = pt;
    const float rmd = octaves - floorf(octaves);
    if (t.musType == 1) {                       // mFractal_t
        float value = 1, pwr = 1;
        for (int i = 0; i < (int)octaves; i++) { value *= (pwr * signedNoise(nt, tp) + 1.0f); pwr *= pwHL; tp *= lacunarity; }
        if (rmd != 0.0f) value *= (rmd * signedNoise(nt, tp) * pwr + 1.0f);
        return value;
    }
    if (t.musType == 2) {                       // heteroTerrain_t
        float pwr = pwHL;
        float value = offset + signedNoise(nt, tp);
        tp *= lacunarity;
        for (int i = 1; i < (int)octaves; i++) { const float increment = (signedNoise(nt, tp) + offset) * pwr * value; value += increment; pwr *= pwHL; tp *= lacunarity; }
        if (rmd != 0.0f) { const float increment = (signedNoise(nt, tp) + offset) * pwr * value; value += rmd * increment; }
        return value;
    }
    if (t.musType == 3) {                       // hybridMFractal_t
        float pwr = pwHL;
        float result = signedNoise(nt, tp) + offset;
        float weight = gain * result;
        tp *= lacunarity;
        for (int i = 1; (weight > 0.001f) && (i < (int)octaves); i++) {
            if (weight > 1.0f) weight = 1.0f;
            const float signal = (signedNoise(nt, tp) + offset) * pwr;
            pwr *= pwHL;
            result += weight * signal;
            weight *= gain * signal;
            tp *= lacunarity;
        }
        if (rmd != 0.0f) result += rmd * ((signedNoise(nt, tp) + offset) * pwr);
        return result;
    }
    if (t.musType == 4) {                       // ridgedMFractal_t
        float pwr = pwHL;
        float signal = offset - fabsf(signedNoise(nt, tp));
        signal *= signal;
        float result = signal, weight = 1.0f;
        for (int i = 1; i < (int)octaves; i++) {
            tp *= lacunarity;
            weight = signal * gain;
            if (weight > 1.0f) weight = 1.0f; else if (weight < 0.0f) weight = 0.0f;
            signal = offset - fabsf(signedNoise(nt, tp));
            signal *= signal;
            signal *= weight;
            result += signal * pwr;
            pwr *= pwHL;
        }
        return result;
    }
    float value = 0, pwr = 1;                   // fBm_t
    for (int i = 0; i < (int)octaves; i++) { value += signedNoise(nt, tp) * pwr; pwr *= pwHL; tp *= lacunarity; }
    if (rmd != 0.f) value += rmd * signedNoise(nt, tp) * pwr;
    return value;
}
YB_HD float voronoi_intensity(const Texture &t, const float *da)
{
    return t.iscale * fabsf(t.vw[0] * da[0] + t.vw[1] * da[1] + t.vw[2] * da[2] + t.vw[3] * da[3]);
}
YB_HD float wave_shape(int wshape, float w)             // the shape switch of marble / wood (basictex.cc:112-124,180-192)
{
    if (wshape == 1) { w *= (float)(0.5 * YB_M_1_PI); w -= floorf(w); return w; }
    if (wshape == 2) { w *= (float)(0.5 * YB_M_1_PI); return fabsf(2.0f * (w - floorf(w)) - 1.0f); }
    return 0.5f + 0.5f * fSin(w);
}

// ---- image texture (imagetex.cc:44-107 interpolateImage, :152-218 doMapping, :109-126 getColor / getNoGammaColor) -------
struct C4 { float r, g, b, a; };
YB_HD C4 c4(f4 p) { C4 c; c.r = p.x; c.g = p.y; c.b = p.z; c.a = p.w; return c; }
YB_HD C4 operator+(C4 a, C4 b) { C4 c; c.r = a.r + b.r; c.g = a.g + b.g; c.b = a.b + b.b; c.a = a.a + b.a; return c; }
YB_HD C4 operator-(C4 a, C4 b) { C4 c; c.r = a.r - b.r; c.g = a.g - b.g; c.b = a.b - b.b; c.a = a.a - b.a; return c; }
YB_HD C4 operator*(C4 a, C4 b) { C4 c; c.r = a.r * b.r; c.g = a.g * b.g; c.b = a.b * b.b; c.a = a.a * b.a; return c; }
YB_HD C4 operator*(C4 a, float f) { C4 c; c.r = a.r * f; c.g = a.g * f; c.b = a.b * f; c.a = a.a * f; return c; }
YB_HD C4 operator*(float f, C4 a) { return a * f; }
YB_HD C4 img_px(const Texture &t, int x, int y) { return c4(t.pixels[(size_t)y * (size_t)t.resx + (size_t)x]); }
YB_HD C4 cubicInterpolate(C4 y0, C4 y1, C4 y2, C4 y3, double mu)          // interpolation.h:47-61 with T = colorA_t
{
    const float m = (float)mu, m2 = (float)(mu * mu);
    C4 mu2; mu2.r = mu2.g = mu2.b = mu2.a = m2;
    const C4 a0 = y3 - y2 - y0 + y1, a1 = y0 - y1 - a0, a2 = y2 - y0, a3 = y1;
    return (a0 * m * mu2 + a1 * mu2 + a2 * m + a3);
}
YB_HD C4 img_interpolate(const Texture &t, v3 p)
{
    const int resx = t.resx, resy = t.resy;
    float xf = ((float)resx * (p.x - floorf(p.x)));
    float yf = ((float)resy * (p.y - floorf(p.y)));
    if (t.intp != 0) { xf -= 0.5f; yf -= 0.5f; }
    int x = (int)xf; x = x > resx - 1 ? resx - 1 : x; x = x < 0 ? 0 : x;
    int y = (int)yf; y = y > resy - 1 ? resy - 1 : y; y = y < 0 ? 0 : y;
    const C4 c1 = img_px(t, x, y);
    if (t.intp == 0) return c1;
    const int x2 = x + 1 < resx - 1 ? x + 1 : resx - 1, y2 = y + 1 < resy - 1 ? y + 1 : resy - 1;
    const C4 c2 = img_px(t, x2, y), c3 = img_px(t, x, y2), c4_ = img_px(t, x2, y2);
    const float dx = xf - floorf(xf), dy = yf - floorf(yf);
    if (t.intp == 1) {
        const float w0 = (1 - dx) * (1 - dy), w1 = (1 - dx) * dy, w2 = dx * (1 - dy), w3 = dx * dy;
        return (w0 * c1) + (w1 * c3) + (w2 * c2) + (w3 * c4_);
    }
    const int x0 = x - 1 > 0 ? x - 1 : 0, x3 = x2 + 1 < resx - 1 ? x2 + 1 : resx - 1;
    const int y0 = y - 1 > 0 ? y - 1 : 0, y3 = y2 + 1 < resy - 1 ? y2 + 1 : resy - 1;
    C4 c0 = cubicInterpolate(img_px(t, x0, y0), img_px(t, x, y0), img_px(t, x2, y0), img_px(t, x3, y0), dx);
    C4 c8 = cubicInterpolate(img_px(t, x0, y), c1, c2, img_px(t, x3, y), dx);
    C4 cA = cubicInterpolate(img_px(t, x0, y2), c3, c4_, img_px(t, x3, y2), dx);
    C4 cC = cubicInterpolate(img_px(t, x0, y3), img_px(t, x, y3), img_px(t, x2, y3), img_px(t, x3, y3), dx);
    return cubicInterpolate(c0, c8, cA, cC, dy);
}
YB_HD bool img_doMapping(const Texture &t, v3 &texpt)
{
    bool outside = false;
    texpt = 0.5f * texpt + v3(0.5f);
    if (t.clipmode == TCL_REPEAT) {
        if (t.xrepeat > 1) { texpt.x *= (float)t.xrepeat; if (texpt.x > 1.0f) texpt.x -= (float)(int)texpt.x; else if (texpt.x < 0.0f) texpt.x += (float)(1 - (int)texpt.x); }
        if (t.yrepeat > 1) { texpt.y *= (float)t.yrepeat; if (texpt.y > 1.0f) texpt.y -= (float)(int)texpt.y; else if (texpt.y < 0.0f) texpt.y += (float)(1 - (int)texpt.y); }
    }
    if (t.cropx) texpt.x = t.cropminx + texpt.x * (t.cropmaxx - t.cropminx);
    if (t.cropy) texpt.y = t.cropminy + texpt.y * (t.cropmaxy - t.cropminy);
    if (t.rot90) { const float tmp = texpt.x; texpt.x = texpt.y; texpt.y = tmp; }
    int mode = t.clipmode;
    if (mode == TCL_CLIPCUBE) {
        if ((texpt.x < 0) || (texpt.x > 1) || (texpt.y < 0) || (texpt.y > 1) || (texpt.z < -1) || (texpt.z > 1)) outside = true;
        return outside;
    }
    if (mode == TCL_CHECKER) {
        const int xs = (int)floorf(texpt.x), ys = (int)floorf(texpt.y);
        texpt.x -= (float)xs; texpt.y -= (float)ys;
        if (!t.checkerOdd && !((xs + ys) & 1)) return true;
        if (!t.checkerEven && ((xs + ys) & 1)) return true;
        if (t.checkerDist < 1.0f) {
            texpt.x = (float)(((double)texpt.x - 0.5) / (1.0 - (double)t.checkerDist) + 0.5);
            texpt.y = (float)(((double)texpt.y - 0.5) / (1.0 - (double)t.checkerDist) + 0.5);
        }
        mode = TCL_CLIP;                                    // falls through
    }
    if (mode == TCL_CLIP) {
        if ((texpt.x < 0) || (texpt.x > 1) || (texpt.y < 0) || (texpt.y > 1)) outside = true;
        return outside;
    }
    if (mode == TCL_EXTEND) {
        if (texpt.x > 0.99999f) texpt.x = 0.99999f; else if (texpt.x < 0) texpt.x = 0;
        if (texpt.y > 0.99999f) texpt.y = 0.99999f; else if (texpt.y < 0) texpt.y = 0;
    }
    return false;
}
YB_HD_NODE C4 img_getNoGammaColor(const Texture &t, v3 p)
{
    v3 p1(p.x, -p.y, p.z);
    C4 ret; ret.r = ret.g = ret.b = ret.a = 0.f;
    if (img_doMapping(t, p1)) return ret;
    return img_interpolate(t, p1);
}

// ---- textures: texture_t::getFloat / getColor --------------------------------------------------------
YB_HD_NODE float tex_getFloat(const Texture &t, v3 p)
{
    switch (t.type) {
        case TEX_RGB_CUBE: {
            const float r = clamp01f(p.x), g = clamp01f(p.y), b = clamp01f(p.z);
            return (r + g + b) * 0.333333f;                                  // color_t::energy
        }
        case TEX_IMAGE: { const C4 c = img_getNoGammaColor(t, p); return 0.2126f * c.r + 0.7152f * c.g + 0.0722f * c.b; }   // texture.h:38: col2bri of the NO-gamma colour
        case TEX_VORONOI: { float da[4]; v3 pa[4]; voronoi_features(p * t.size, t.distMetric, t.mkExp, da, pa); return voronoi_intensity(t, da); }
        case TEX_MUSGRAVE: return t.iscale * musgrave(t, p * t.size);
        case TEX_DISTNOISE: {                                                // basictex.cc:427-435
            const v3 ofs(13.5f, 13.5f, 13.5f);
            const v3 tp = p * t.size;
            const v3 rv(signedNoise(t.noiseType, tp + ofs), signedNoise(t.noiseType, tp), signedNoise(t.noiseType, tp - ofs));
            return signedNoise(t.noiseType2, tp + rv * t.distort);
        }
        case TEX_BLEND: {
            float blend = (float)(0.5 * ((double)p.x + 1.));
            return fmaxf_(0.f, fminf_(blend, 1.f));
        }
        case TEX_CLOUDS: {
            float v = turbulence(t.noiseType, p, t.depth, t.size, t.hard != 0);
            if (t.bias) { v *= v; if (t.bias == 1) return -v; }
            return v;
        }
        case TEX_MARBLE: {
            float w = (float)((double)(p.x + p.y + p.z) * 5.0 + ((t.turb == 0.0f) ? 0.0 : (double)(t.turb * turbulence(t.noiseType, p, t.depth, t.size, t.hard != 0))));
            return fPow(wave_shape(t.wshape, w), t.sharpness);
        }
        default: {                                                           // TEX_WOOD
            float w;
            if (t.rings) w = (float)((double)fSqrt(p.x * p.x + p.y * p.y + p.z * p.z) * 20.0);
            else w = (float)((double)(p.x + p.y + p.z) * 10.0);
            w += (t.turb == 0.0f) ? 0.0f : t.turb * turbulence(t.noiseType, p, t.depth, t.size, t.hard != 0);
            return wave_shape(t.wshape, w);
        }
    }
}
YB_HD_NODE void tex_getColor(const Texture &t, v3 p, float *c)
{
    if (t.type == TEX_RGB_CUBE) { c[0] = clamp01f(p.x); c[1] = clamp01f(p.y); c[2] = clamp01f(p.z); c[3] = 1.f; return; }
    if (t.type == TEX_BLEND) { c[0] = c[1] = c[2] = c[3] = 1.f; return; }
    if (t.type == TEX_VORONOI) {                                             // basictex.cc:288-318
        float da[4]; v3 pa[4];
        voronoi_features(p * t.size, t.distMetric, t.mkExp, da, pa);
        const float inte = voronoi_intensity(t, da);
        if (!t.coltype) { c[0] = c[1] = c[2] = c[3] = inte; return; }
        float col[4] = {0.f, 0.f, 0.f, 0.f};
        for (int k = 0; k < 4; ++k) {                                        // aw_k * cellNoiseColor(pa[k]) (noise.cc:748-755), alpha 1
            const int xi = (int)floorf(pa[k].x), yi = (int)floorf(pa[k].y), zi = (int)floorf(pa[k].z);
            const float *cl = yb_noise_pnt + 3 * nperm(nperm(nperm(zi & 255) + yi) + xi);
            col[0] += t.vaw[k] * cl[0]; col[1] += t.vaw[k] * cl[1]; col[2] += t.vaw[k] * cl[2]; col[3] += t.vaw[k] * 1.0f;
        }
        float mulv = t.iscale;
        if (t.coltype >= 2) {
            float t1 = (float)((double)(da[1] - da[0]) * 10.0);
            if (t1 > 1) t1 = 1;
            if (t.coltype == 3) t1 *= inte; else t1 *= t.iscale;
            mulv = t1;
        }
        c[0] = col[0] * mulv; c[1] = col[1] * mulv; c[2] = col[2] * mulv; c[3] = col[3] * mulv;
        return;
    }
    if (t.type == TEX_IMAGE) {
        C4 r = img_getNoGammaColor(t, p);
        if (t.gamma != 1.f && !t.isHDR) { r.r = fPow(r.r, t.gamma); r.g = fPow(r.g, t.gamma); r.b = fPow(r.b, t.gamma); }
        c[0] = r.r; c[1] = r.g; c[2] = r.b; c[3] = r.a;
        return;
    }
    const float f = tex_getFloat(t, p);                                      // color1 + getFloat(p) * (color2 - color1), alpha 1
    c[0] = t.color1.x + f * (t.color2.x - t.color1.x); c[1] = t.color1.y + f * (t.color2.y - t.color1.y);
    c[2] = t.color1.z + f * (t.color2.z - t.color1.z); c[3] = 1.f;
}

// ---- nodes ---------------------------------------------------------------------------------------------
YB_HD float col2bri(const float *c) { return 0.2126f * c[0] + 0.7152f * c[1] + 0.0722f * c[2]; }   // color.h:79-80
// texture_rgb_blend / texture_value_blend (shader.h:112-225); blend types: mix_modes of shader.h:62
YB_HD void texture_rgb_blend(const float *tex, const float *out, float fact, float facg, int mode, float *res)
{
    for (int k = 0; k < 3; ++k) {
        const float t = tex[k], o = out[k];
        float f = fact, r;
        switch (mode) {
            case 2: f *= facg; r = ((1.f - facg) + f * t) * o; break;                                    // MN_MULT
            case 4: f *= facg; r = 1.f - ((1.f - facg) + f * (1.f - t)) * (1.f - o); break;              // MN_SCREEN
            case 3: f = -f; f *= facg; r = f * t + o; break;                                             // MN_SUB
            case 1: f *= facg; r = f * t + o; break;                                                     // MN_ADD
            case 5: { f *= facg; const float it = t == 0.f ? 0.f : 1.f / t; r = (1.f - f) * o + f * o * it; break; }   // MN_DIV (invertRGB, color.h)
            case 6: f *= facg; r = (1.f - f) * o + f * fabsf(t - o); break;                              // MN_DIFF
            case 7: { f *= facg; const float c = f * t; r = c < o ? c : o; break; }                      // MN_DARK (darkenRGB)
            case 8: { f *= facg; const float c = f * t; r = c > o ? c : o; break; }                      // MN_LIGHT (lightenRGB)
            default: f *= facg; r = f * t + (1.f - f) * o; break;                                        // MN_MIX
        }
        res[k] = r;
    }
}
YB_HD float texture_value_blend(float tex, float out, float fact, float facg, int mode, bool flip)
{
    fact *= facg;
    float facm = 1.f - fact;
    if (flip) { const float t = fact; fact = facm; facm = t; }
    switch (mode) {
        case 2: facm = 1.f - facg; return (facm + fact * tex) * out;
        case 4: facm = 1.f - facg; return 1.f - (facm + fact * (1.f - tex)) * (1.f - out);
        case 3: fact = -fact; return fact * tex + out;
        case 1: return fact * tex + out;
        case 5: if (tex == 0.f) return 0.f; return facm * out + fact * out / tex;
        case 6: return facm * out + fact * fabsf(tex - out);
        case 7: { const float c = fact * tex; return c < out ? c : out; }
        case 8: { const float c = fact * tex; return c > out ? c : out; }
        default: return fact * tex + facm * out;
    }
}

// textureMapper_t::getCoords (basicnodes.cc:121-139) and doMapping (:96-119)
YB_HD v3 texmap_coords(const ShaderNode &n, v3 P, const TexPoint &tp)
{
    switch (n.texCoords) {
        case TXC_UV: return v3(tp.U, tp.V, 0.f);
        case TXC_ORCO: return tp.orcoP;
        case TXC_TRAN: return v3(n.mtx[0] * P.x + n.mtx[1] * P.y + n.mtx[2] * P.z + n.mtx[3], n.mtx[4] * P.x + n.mtx[5] * P.y + n.mtx[6] * P.z + n.mtx[7],
                                 n.mtx[8] * P.x + n.mtx[9] * P.y + n.mtx[10] * P.z + n.mtx[11]);
        default: return P;
    }
}
// tubemap / spheremap / cubemap of basicnodes.cc:41-91 (discrete textures only, :270)
YB_HD v3 texmap_doMapping(const ShaderNode &n, v3 texpt, v3 N)
{
    if (n.texCoords == TXC_UV) texpt = v3(2.0f * texpt.x - 1.0f, 2.0f * texpt.y - 1.0f, texpt.z);
    const float texmap[4] = {0.f, texpt.x, texpt.y, texpt.z};
    texpt = v3(texmap[n.mapX], texmap[n.mapY], texmap[n.mapZ]);
    if (n.mapType == TXP_TUBE) {
        const v3 p = texpt; v3 res(0.f, p.z, 0.f);
        const float d = p.x * p.x + p.y * p.y;
        if (d > 0) { res.z = (float)(1.0 / (double)fSqrt(d)); res.x = (float)((double)(-atan2f(p.x, p.y)) * YB_M_1_PI); }
        texpt = res;
    } else if (n.mapType == TXP_SPHERE) {
        const v3 p = texpt; v3 res(0.f);
        const float d = p.x * p.x + p.y * p.y + p.z * p.z;
        if (d > 0) {
            res.z = fSqrt(d);
            if ((p.x != 0) && (p.y != 0)) res.x = (float)((double)(-atan2f(p.x, p.y)) * YB_M_1_PI);
            res.y = (float)(1.0 - 2.0 * ((double)fAcos(p.z / res.z) * YB_M_1_PI));
        }
        texpt = res;
    } else if (n.mapType == TXP_CUBE) {
        int axis;
        if (fabsf(N.z) >= fabsf(N.x) && fabsf(N.z) >= fabsf(N.y)) axis = 2;
        else if (fabsf(N.y) >= fabsf(N.x) && fabsf(N.y) >= fabsf(N.z)) axis = 1;
        else axis = 0;
        const v3 p = texpt;
        texpt = axis == 0 ? v3(p.y, p.z, p.x) : (axis == 1 ? v3(p.x, p.z, p.y) : p);
    }
    return mul(texpt, n.scale) + n.offset;
}
YB_HD v3 texmap_Ng(const ShaderNode &n, v3 Ng, const TexPoint &tp)        // the normal getCoords pairs with the coordinates
{
    if (n.texCoords == TXC_ORCO) return tp.orcoNg;
    if (n.texCoords == TXC_TRAN) return v3(n.mtx[0] * Ng.x + n.mtx[1] * Ng.y + n.mtx[2] * Ng.z, n.mtx[4] * Ng.x + n.mtx[5] * Ng.y + n.mtx[6] * Ng.z,
                                            n.mtx[8] * Ng.x + n.mtx[9] * Ng.y + n.mtx[10] * Ng.z);
    return Ng;
}
// shaderNode_t::evalDerivative for bump mapping: texture_mapper on a procedural (non-discrete) texture (basicnodes.cc:218-232),
// layer (layernode.cc:118-146); every other node type keeps the base-class zero (shader.h:83-84)
YB_HD_NODE void node_evalDerivative(const ShaderNode &n, const Texture *textures, v3 P, v3 Ng, v3 NU, v3 NV, const TexPoint &tp, NodeRes *st, int self)
{
    NodeRes r; r.r = r.g = r.b = r.a = r.f = 0.f;
    if (n.type == NODE_TEXMAP) {
        v3 texpt = texmap_coords(n, P, tp);
        const v3 mNg = texmap_Ng(n, Ng, tp);
        const Texture &t = textures[n.tex];
        float du = 0.f, dv = 0.f;
        if (t.type == TEX_IMAGE) {                     // discrete texture (basicnodes.cc:173-216)
            texpt = texmap_doMapping(n, texpt, mNg);
            v3 norm(0.f);
            if (t.normalmap) {
                const C4 color = img_getNoGammaColor(t, texpt);
                norm = 2.f * v3(color.r, color.g, color.b) - v3(1.f);
                du = dot(norm, tp.dSdU); dv = dot(norm, tp.dSdV);          // overwritten below, as in the reference
            } else {
                const v3 pDU(n.dU, 0.f, 0.f), pDV(0.f, n.dV, 0.f);
                const float dfdu = (tex_getFloat(t, texpt - pDU) - tex_getFloat(t, texpt + pDU)) / n.dU;
                const float dfdv = (tex_getFloat(t, texpt - pDV) - tex_getFloat(t, texpt + pDV)) / n.dV;
                v3 vecU = tp.dSdU, vecV = tp.dSdV;
                vecU.z = dfdu; vecV.z = dfdv;
                norm = cross(vecU, vecV);
            }
            norm = normalize(norm);
            if (fabsf(norm.z) > 1e-30f) {
                const float NF = (float)(1.0 / (double)norm.z * (double)n.bumpStr);
                du = norm.x * NF; dv = norm.y * NF;
            } else du = dv = 0.f;
        } else {
            const v3 i0 = texmap_doMapping(n, texpt - n.dU * NU, mNg), i1 = texmap_doMapping(n, texpt + n.dU * NU, mNg);
            const v3 j0 = texmap_doMapping(n, texpt - n.dV * NV, mNg), j1 = texmap_doMapping(n, texpt + n.dV * NV, mNg);
            du = (tex_getFloat(t, i0) - tex_getFloat(t, i1)) / n.dU;
            dv = (tex_getFloat(t, j0) - tex_getFloat(t, j1)) / n.dV;
            du *= n.bumpStr; dv *= n.bumpStr;
        }
        r.r = du; r.g = dv;
    } else if (n.type == NODE_LAYER) {
        float rdu = 0.f, rdv = 0.f, stencilTin = 1.f;
        if (n.upper >= 0) { rdu = st[n.upper].r; rdv = st[n.upper].g; stencilTin = st[n.upper].a; }
        float tdu = st[n.input].r, tdv = st[n.input].g;
        if (n.texflag & TXF_NEGATIVE) { tdu = -tdu; tdv = -tdv; }
        r.r = rdu + tdu; r.g = rdv + tdv; r.a = stencilTin;
    }
    st[self] = r;
}

YB_HD_NODE void node_eval(const ShaderNode &n, const Texture *textures, v3 P, v3 Ng, const TexPoint &tp, NodeRes *st, int self)
{
    NodeRes r; r.r = r.g = r.b = r.a = r.f = 0.f;
    if (n.type == NODE_TEXMAP) {                      // textureMapper_t::eval
        const v3 texpt = texmap_doMapping(n, texmap_coords(n, P, tp), texmap_Ng(n, Ng, tp));
        const Texture &t = textures[n.tex];
        float c[4];
        tex_getColor(t, texpt, c);
        r.r = c[0]; r.g = c[1]; r.b = c[2]; r.a = c[3];
        r.f = n.doScalar ? tex_getFloat(t, texpt) : 0.f;
    } else if (n.type == NODE_VALUE) {
        r.r = n.col1[0]; r.g = n.col1[1]; r.b = n.col1[2]; r.a = n.col1[3]; r.f = n.val1;
    } else if (n.type == NODE_MIX) {                  // mixNode_t::getInputs + the eval of the selected subclass
        const float f2 = n.inFactor >= 0 ? st[n.inFactor].f : n.cfactor, f1 = 1.f - f2;
        float c1[4], c2[4], fin1, fin2;
        if (n.in1 >= 0) { c1[0] = st[n.in1].r; c1[1] = st[n.in1].g; c1[2] = st[n.in1].b; c1[3] = st[n.in1].a; fin1 = st[n.in1].f; }
        else { c1[0] = n.col1[0]; c1[1] = n.col1[1]; c1[2] = n.col1[2]; c1[3] = n.col1[3]; fin1 = n.val1; }
        if (n.in2 >= 0) { c2[0] = st[n.in2].r; c2[1] = st[n.in2].g; c2[2] = st[n.in2].b; c2[3] = st[n.in2].a; fin2 = st[n.in2].f; }
        else { c2[0] = n.col2[0]; c2[1] = n.col2[1]; c2[2] = n.col2[2]; c2[3] = n.col2[3]; fin2 = n.val2; }
        float o[4], of;
        switch (n.mode) {
            case 1: for (int k = 0; k < 4; ++k) o[k] = c1[k] + f2 * c2[k]; of = fin1 + f2 * fin2; break;                    // add
            case 2: for (int k = 0; k < 4; ++k) o[k] = c1[k] * (f1 + f2 * c2[k]); of = fin1; break;                          // mult (sic: scalar = fin1)
            case 3: for (int k = 0; k < 4; ++k) o[k] = c1[k] - f2 * c2[k]; of = fin1 - f2 * fin2; break;                    // sub
            case 4: for (int k = 0; k < 4; ++k) o[k] = 1.f - (f1 + f2 * (1.f - c2[k])) * (1.f - c1[k]);
                    of = (float)(1.0 - (double)((f1 + f2 * (1.f - fin2)) * (1.f - fin1))); break;                           // screen
            case 6: for (int k = 0; k < 4; ++k) o[k] = f1 * c1[k] + f2 * fabsf(c1[k] - c2[k]); of = f1 * fin1 + f2 * fabsf(fin1 - fin2); break;   // diff
            case 7: for (int k = 0; k < 4; ++k) { const float x = c2[k] * f2; o[k] = x < c1[k] ? x : c1[k]; }
                    { const float x = fin2 * f2; of = x < fin1 ? x : fin1; } break;                                         // dark
            case 8: for (int k = 0; k < 4; ++k) { const float x = c2[k] * f2; o[k] = x > c1[k] ? x : c1[k]; }
                    { const float x = fin2 * f2; of = x > fin1 ? x : fin1; } break;                                         // light
            case 9: for (int k = 0; k < 4; ++k)
                        o[k] = (c1[k] < 0.5f) ? c1[k] * (f1 + 2.0f * f2 * c2[k]) : (float)(1.0 - ((double)f1 + (double)(2.0f * f2) * (1.0 - (double)c2[k])) * (1.0 - (double)c1[k]));
                    of = (fin1 < 0.5f) ? fin1 * (f1 + 2.0f * f2 * fin2) : (float)(1.0 - ((double)f1 + (double)(2.0f * f2) * (1.0 - (double)fin2)) * (1.0 - (double)fin1));
                    break;                                                                                                  // overlay
            default: for (int k = 0; k < 4; ++k) o[k] = f1 * c1[k] + f2 * c2[k]; of = f1 * fin1 + f2 * fin2; break;         // mix (also mode 5: no case in the factory)
        }
        r.r = o[0]; r.g = o[1]; r.b = o[2]; r.a = o[3]; r.f = of;
    } else {                                          // layerNode_t::eval (layernode.cc:12-110)
        float rcol[4], texcolor[4] = {0.f, 0.f, 0.f, 0.f};
        float rval, Tin = 0.f, Ta = 1.f, stencilTin;
        if (n.upper >= 0) { rcol[0] = st[n.upper].r; rcol[1] = st[n.upper].g; rcol[2] = st[n.upper].b; rcol[3] = st[n.upper].a; rval = st[n.upper].f; }
        else { rcol[0] = n.upperCol[0]; rcol[1] = n.upperCol[1]; rcol[2] = n.upperCol[2]; rcol[3] = n.upperCol[3]; rval = n.upperVal; }
        stencilTin = rcol[3];
        bool TEX_RGB = n.colorInput != 0;
        if (n.colorInput) { texcolor[0] = st[n.input].r; texcolor[1] = st[n.input].g; texcolor[2] = st[n.input].b; texcolor[3] = st[n.input].a; Ta = texcolor[3]; }
        else Tin = st[n.input].f;
        if (n.texflag & TXF_RGBTOINT) { Tin = col2bri(texcolor); TEX_RGB = false; }
        if (n.texflag & TXF_NEGATIVE) {
            if (TEX_RGB) for (int k = 0; k < 4; ++k) texcolor[k] = 1.f - texcolor[k];
            Tin = 1.f - Tin;
        }
        if (n.texflag & TXF_STENCIL) {
            float fact;
            if (TEX_RGB) { fact = Ta; Ta *= stencilTin; stencilTin *= fact; }
            else { fact = Tin; Tin *= stencilTin; stencilTin *= fact; }
        }
        if (n.doColor) {
            if (!TEX_RGB) { texcolor[0] = n.defCol.x; texcolor[1] = n.defCol.y; texcolor[2] = n.defCol.z; texcolor[3] = 1.f; }
            else Tin = Ta;
            Tin = Tin > 1.f ? 1.f : (Tin < 0.f ? 0.f : Tin);               // inRange(1, 0, Tin)
            float res[3];
            texture_rgb_blend(texcolor, rcol, Tin, stencilTin * n.colfac, n.mode, res);
            rcol[0] = fmaxf_(0.f, res[0]); rcol[1] = fmaxf_(0.f, res[1]); rcol[2] = fmaxf_(0.f, res[2]);   // clampRGB0; alpha of the blend is 1 (color_t -> colorA_t)
            rcol[3] = 1.f;
        }
        if (n.doScalarL) {
            if (TEX_RGB) {
                if (n.useAlpha) { Tin = Ta; if (n.texflag & TXF_NEGATIVE) Tin = 1.f - Tin; }
                else Tin = col2bri(texcolor);
            }
            rval = texture_value_blend(n.defVal, rval, Tin, stencilTin * n.valfac, n.mode, !n.doScalarL);
            if (rval < 0.f) rval = 0.f;
        }
        rcol[3] = stencilTin;
        r.r = rcol[0]; r.g = rcol[1]; r.b = rcol[2]; r.a = rcol[3]; r.f = rval;
    }
    st[self] = r;
}

}  // namespace yb
