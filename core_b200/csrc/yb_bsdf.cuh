// yb_bsdf.cuh — device restatement of the material_t interface for the supported materials:
//   shinyDiffuseMat_t  src/materials/shinydiffuse.cc:132-469
//   glossyMat_t        src/materials/glossy.cc:82-351       (+ include/materials/microfacet.h)
//   coatedGlossyMat_t  src/materials/coatedglossy.cc:104-418
//   glassMat_t/mirrorMat_t/nullMat_t  src/materials/glass.cc:84-253,336-390
//   lightMat_t         src/materials/simple.cc:54-75
// The BSDF code reads constants only: shader nodes are evaluated first and written over the constants of a local copy of the
// record (mat_resolve at the end of this file).  The reference's behaviours
// that look like bugs are kept on purpose: W = |wi.N|/(0.99 pdf + 0.01), no 1/pi, ...
#pragma once
#include "yb_types.h"
#include "yb_nodes.cuh"

namespace yb {

struct Surf {            // the fields of surfacePoint_t the supported materials read (triangle.cc:12-108)
    v3 P, N, Ng, NU, NV;
    const Material *mat;
    // blend_mat with a mask shader node (blend.cc:49-66): the blend value of THIS point, written by mat_resolve; the material record
    // itself stays where it is (its sub-materials are addressed relative to it)
    float blendVal; int32_t hasBlendVal;
};
struct MatDat { float c[4]; };   // SDDat_t::component[4]  |  MDat_t{mDiffuse,mGlossy,pDiffuse}
struct BSample {                 // sample_t (material.h:49-60) minus the reverse-path fields
    float s1, s2, pdf;
    uint32_t flags, sampledFlags;
    YB_HD BSample(float a, float b, uint32_t f) : s1(a), s2(b), pdf(0.f), flags(f), sampledFlags(BSDF_NONE) {}
};
struct RState { int raylevel; bool includeLights; };

#define YB_FACE_FORWARD(Ng, N, I) ((dot((Ng), (I)) < 0) ? -(N) : (N))

// ================================================================== microfacet.h
YB_HD float pdfDivisor(float c) { return 8.f * (float)YB_M_PI * (c * 0.99f + 0.04f); }
YB_HD float ASDivisor(float cos1, float cosI, float cosO) { return 8.f * (float)YB_M_PI * ((cos1 * fmaxf_(cosI, cosO)) * 0.99f + 0.04f); }
YB_HD void sample_quadrant_aniso(v3 &H, float s1, float s2, float e_u, float e_v)
{
    float phi = atanf(fSqrt((e_u + 1.f) / (e_v + 1.f)) * tanf((float)YB_M_PI_2 * s1));
    float cosPhi = fCos(phi), sinPhi = fSin(phi);
    float cosPhi2 = cosPhi * cosPhi, sinPhi2 = 1.f - cosPhi2;
    float cosTheta = fPow(1.f - s2, 1.f / (e_u * cosPhi2 + e_v * sinPhi2 + 1.f));
    float sinTheta = fSqrt(1.f - cosTheta * cosTheta);
    H = v3(sinTheta * cosPhi, sinTheta * sinPhi, cosTheta);
}
YB_HD float AS_Aniso_D(v3 h, float e_u, float e_v)
{
    if (h.z <= 0.f) return 0.f;
    float exponent = (e_u * h.x * h.x + e_v * h.y * h.y) / (1.00001f - h.z * h.z);
    return fSqrt((e_u + 1.f) * (e_v + 1.f)) * fPow(fmaxf_(0.f, h.z), exponent);
}
YB_HD float AS_Aniso_Pdf(v3 h, float cos_w_H, float e_u, float e_v) { return AS_Aniso_D(h, e_u, e_v) / pdfDivisor(cos_w_H); }
YB_HD void AS_Aniso_Sample(v3 &H, float s1, float s2, float e_u, float e_v)
{
    if (s1 < 0.25f) sample_quadrant_aniso(H, 4.f * s1, s2, e_u, e_v);
    else if (s1 < 0.5f) { sample_quadrant_aniso(H, 1.f - 4.f * (0.5f - s1), s2, e_u, e_v); H.x = -H.x; }
    else if (s1 < 0.75f) { sample_quadrant_aniso(H, 4.f * (s1 - 0.5f), s2, e_u, e_v); H.x = -H.x; H.y = -H.y; }
    else { sample_quadrant_aniso(H, 1.f - 4.f * (1.f - s1), s2, e_u, e_v); H.y = -H.y; }
}
YB_HD float Blinn_D(float cos_h, float e) { return (e + 1.f) * fPow(cos_h, e); }
YB_HD float Blinn_Pdf(float costheta, float cos_w_H, float e) { return Blinn_D(costheta, e) / pdfDivisor(cos_w_H); }
YB_HD void Blinn_Sample(v3 &H, float s1, float s2, float exponent)
{
    float cosTheta = fPow(1.f - s2, 1.f / (exponent + 1.f));
    float sinTheta = fSqrt(1.f - cosTheta * cosTheta);
    float phi = s1 * (float)YB_M_2PI;
    H = v3(sinTheta * fCos(phi), sinTheta * fSin(phi), cosTheta);
}
YB_HD float SchlickFresnel(float costheta, float R)
{
    float c1 = (1.f - costheta);
    float c2 = c1 * c1;
    return R + ((1.f - R) * c1 * c2 * c2);
}
YB_HD col3 diffuseReflect(float wiN, float woN, float mGlossy, float mDiffuse, col3 diff_base)
{
    float f_wi = (1.f - (0.5f * wiN));
    float temp = f_wi * f_wi;
    f_wi = temp * temp * f_wi;
    float f_wo = (1.f - (0.5f * woN));
    temp = f_wo * f_wo;
    f_wo = temp * temp * f_wo;
    return (float)(0.387507688 * mDiffuse * (1.f - mGlossy) * (1.f - f_wi) * (1.f - f_wo)) * diff_base;
}

// ================================================================== shinyDiffuseMat_t
YB_HD void sd_getFresnel(const Material &m, v3 wo, v3 n, float &Kr)      // shinydiffuse.cc:100-128
{
    if (m.hasFresnel) {
        v3 N = (dot(wo, n) < 0.f) ? -n : n;
        float c = dot(wo, N);
        float g = m.iorSquared + c * c - 1.f;
        if (g < 0.f) g = 0.f; else g = fSqrt(g);
        float aux = c * (g + c);
        Kr = ((0.5f * (g - c) * (g - c)) / ((g + c) * (g + c))) * (1.f + ((aux - 1) * (aux - 1)) / ((aux + 1) * (aux + 1)));
    } else Kr = 1.f;
}
YB_HD void sd_accumulate(const float *component, float *accum, float Kr)  // :133-143
{
    accum[0] = component[0] * Kr;
    float acc = 1.f - accum[0];
    accum[1] = component[1] * acc;
    acc *= 1.f - component[1];
    accum[2] = component[2] * acc;
    acc *= 1.f - component[2];
    accum[3] = component[3] * acc;
}
YB_HD float sd_OrenNayar(const Material &m, v3 wi, v3 wo, v3 N)            // :185-220
{
    float cos_ti = fmaxf_(1e-8f, fminf_(1.f, dot(N, wi)));
    float cos_to = fmaxf_(1e-8f, fminf_(1.f, dot(N, wo)));
    float maxcos_f = 0.f;
    if (cos_ti < 0.9999f && cos_to < 0.9999f) {
        v3 v1 = normalize(wi - N * cos_ti);
        v3 v2 = normalize(wo - N * cos_to);
        maxcos_f = fmaxf_(0.f, dot(v1, v2));
    }
    float sin_alpha, tan_beta;
    if (cos_to >= cos_ti) { sin_alpha = fSqrt(1.f - cos_ti * cos_ti); tan_beta = fSqrt(1.f - cos_to * cos_to) / cos_to; }
    else { sin_alpha = fSqrt(1.f - cos_to * cos_to); tan_beta = fSqrt(1.f - cos_ti * cos_ti) / cos_ti; }
    return m.orenA + m.orenB * maxcos_f * sin_alpha * tan_beta;
}
YB_HD void sd_initBSDF(const Material &m, MatDat &d)                       // :145-166 + getComponents :82-98
{
    d.c[0] = d.c[1] = d.c[2] = d.c[3] = 0.f;
    if (m.isMirror) d.c[0] = m.mirrorStrength;
    if (m.isTransparent) d.c[1] = m.transparencyStrength;
    if (m.isTranslucent) d.c[2] = m.translucencyStrength;
    if (m.isDiffuse) d.c[3] = m.diffuseStrength;
}
YB_HD col3 sd_eval(const Material &m, const MatDat &d, const Surf &sp, v3 wo, v3 wl, uint32_t bsdfs)   // :223-249
{
    float cos_Ng_wo = dot(sp.Ng, wo), cos_Ng_wl = dot(sp.Ng, wl);
    v3 N = YB_FACE_FORWARD(sp.Ng, sp.N, wo);
    if (!(bsdfs & m.bsdfFlags & BSDF_DIFFUSE)) return col3(0.f);
    float Kr;
    sd_getFresnel(m, wo, N, Kr);
    float mT = (1.f - Kr * d.c[0]) * (1.f - d.c[1]);
    bool transmit = (cos_Ng_wo * cos_Ng_wl) < 0.f;
    if (transmit) { if (m.isTranslucent) return (d.c[2] * mT) * m.diffuseColor; }
    if (dot(N, wl) < 0.0f) return col3(0.f);
    float mD = mT * (1.f - d.c[2]) * d.c[3];
    if (m.useOrenNayar) mD *= sd_OrenNayar(m, wo, wl, N);
    return mD * m.diffuseColor;
}
YB_HD col3 sd_sample(const Material &m, const MatDat &d, const Surf &sp, v3 wo, v3 &wi, BSample &s, float &W)   // :259-336
{
    float accumC[4];
    float cos_Ng_wo = dot(sp.Ng, wo), cos_Ng_wi, cos_N;
    v3 N = YB_FACE_FORWARD(sp.Ng, sp.N, wo);
    float Kr;
    sd_getFresnel(m, wo, N, Kr);
    sd_accumulate(d.c, accumC, Kr);
    // component list of shinyDiffuseMat_t::sample (shinydiffuse.cc:268-296), walked in its fixed order (mirror, transparency,
    // translucency, diffuse: cFlags / cIndex are filled in that order) with static indices: the same partial sums, the same
    // normalisation and the same pick as the arrays val[] / width[] / choice[] of the reference, kept in registers
    const uint32_t CF[4] = {BSDF_SPECULAR | BSDF_REFLECT, BSDF_TRANSMIT | BSDF_FILTER, BSDF_DIFFUSE | BSDF_TRANSMIT, BSDF_DIFFUSE | BSDF_REFLECT};
    const bool on[4] = {m.isMirror && (s.flags & CF[0]) == CF[0], m.isTransparent && (s.flags & CF[1]) == CF[1],
                        m.isTranslucent && (s.flags & CF[2]) == CF[2], m.isDiffuse && (s.flags & CF[3]) == CF[3]};
    float sum = 0.f, val[4] = {0.f, 0.f, 0.f, 0.f};
    int nMatch = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int c = 0; c < 4; ++c)
        if (on[c]) { sum += accumC[c]; val[c] = sum; ++nMatch; }
    if (!nMatch || sum < 0.00001) { s.sampledFlags = BSDF_NONE; s.pdf = 0.f; return col3(1.f); }
    float inv_sum = 1.f / sum;
    int pc = -1, lastOn = 0;
    float wPick = 0.f, prevPick = 0.f, lastW = 0.f, lastPrev = 0.f, runPrev = 0.f;
    bool hasPrevPick = false, lastHasPrev = false, anyPrev = false;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int c = 0; c < 4; ++c)
        if (on[c]) {
            const float vN = val[c] * inv_sum, wN = accumC[c] * inv_sum;
            if (pc < 0 && s.s1 <= vN) { pc = c; wPick = wN; prevPick = runPrev; hasPrevPick = anyPrev; }
            lastOn = c; lastW = wN; lastPrev = runPrev; lastHasPrev = anyPrev;
            runPrev = vN; anyPrev = true;
        }
    if (pc < 0) { pc = lastOn; wPick = lastW; prevPick = lastPrev; hasPrevPick = lastHasPrev; }
    float s1;
    if (hasPrevPick) s1 = (s.s1 - prevPick) / wPick;
    else s1 = s.s1 / wPick;
    col3 scolor(0.f);
    switch (pc) {
        case 0:         // BSDF_SPECULAR | BSDF_REFLECT
            wi = reflect_dir(N, wo);
            s.pdf = wPick;
            scolor = m.mirrorColor * accumC[0];
            scolor *= 1.f / fabsf(dot(sp.N, wi));
            break;
        case 1:         // BSDF_TRANSMIT | BSDF_FILTER
            wi = -wo;
            scolor = accumC[1] * (m.transmitFilterStrength * m.diffuseColor + col3(1.f - m.transmitFilterStrength));
            cos_N = fabsf(dot(wi, N));
            if (cos_N < 1e-6) s.pdf = 0.f; else s.pdf = wPick;
            break;
        case 2:         // BSDF_DIFFUSE | BSDF_TRANSMIT
            wi = SampleCosHemisphere(-N, sp.NU, sp.NV, s1, s.s2);
            cos_Ng_wi = dot(sp.Ng, wi);
            if (cos_Ng_wo * cos_Ng_wi < 0) scolor = accumC[2] * m.diffuseColor;
            s.pdf = fabsf(dot(wi, N)) * wPick;
            break;
        default:        // BSDF_DIFFUSE | BSDF_REFLECT
            wi = SampleCosHemisphere(N, sp.NU, sp.NV, s1, s.s2);
            cos_Ng_wi = dot(sp.Ng, wi);
            if (cos_Ng_wo * cos_Ng_wi > 0) scolor = accumC[3] * m.diffuseColor;
            if (m.useOrenNayar) scolor *= sd_OrenNayar(m, wo, wi, N);
            s.pdf = fabsf(dot(wi, N)) * wPick;
            break;
    }
    s.sampledFlags = pc == 0 ? CF[0] : (pc == 1 ? CF[1] : (pc == 2 ? CF[2] : CF[3]));
    W = fabsf(dot(wi, sp.N)) / (s.pdf * 0.99f + 0.01f);
    return scolor;
}
YB_HD float sd_pdf(const Material &m, const MatDat &d, const Surf &sp, v3 wo, v3 wi, uint32_t bsdfs)   // :338-377
{
    if (!(bsdfs & BSDF_DIFFUSE)) return 0.f;
    float pdf = 0.f, accumC[4];
    float cos_Ng_wo = dot(sp.Ng, wo), cos_Ng_wi;
    v3 N = YB_FACE_FORWARD(sp.Ng, sp.N, wo);
    float Kr;
    sd_getFresnel(m, wo, N, Kr);
    sd_accumulate(d.c, accumC, Kr);
    // the component list in its fixed order (see sd_sample), static indices
    float sum = 0.f, width;
    int nMatch = 0;
    if (m.isMirror && (bsdfs & (BSDF_SPECULAR | BSDF_REFLECT))) { sum += accumC[0]; ++nMatch; }
    if (m.isTransparent && (bsdfs & (BSDF_TRANSMIT | BSDF_FILTER))) { sum += accumC[1]; ++nMatch; }
    if (m.isTranslucent && (bsdfs & (BSDF_DIFFUSE | BSDF_TRANSMIT))) {
        width = accumC[2];
        sum += width;
        cos_Ng_wi = dot(sp.Ng, wi);
        if (cos_Ng_wo * cos_Ng_wi < 0) pdf += fabsf(dot(wi, N)) * width;
        ++nMatch;
    }
    if (m.isDiffuse && (bsdfs & (BSDF_DIFFUSE | BSDF_REFLECT))) {
        width = accumC[3];
        sum += width;
        pdf += fabsf(dot(wi, N)) * width;
        ++nMatch;
    }
    if (!nMatch || sum < 0.00001) return 0.f;
    return pdf / sum;
}
YB_HD void sd_getSpecular(const Material &m, const MatDat &d, const Surf &sp, v3 wo, bool &doReflect, bool &doRefract, v3 *wi, col3 *col)   // :391-433
{
    const bool backface = dot(wo, sp.Ng) < 0.f;
    const v3 N = backface ? -sp.N : sp.N;
    const v3 Ng = backface ? -sp.Ng : sp.Ng;
    float Kr;
    sd_getFresnel(m, wo, N, Kr);
    if (m.isTransparent) {
        doRefract = true;
        wi[1] = -wo;
        col3 tcol = m.transmitFilterStrength * m.diffuseColor + col3(1.f - m.transmitFilterStrength);
        col[1] = ((1.f - d.c[0] * Kr) * d.c[1]) * tcol;
    } else doRefract = false;
    if (m.isMirror) {
        doReflect = true;
        wi[0] = reflect_about(wo, N);
        float cos_wi_Ng = dot(wi[0], Ng);
        if (cos_wi_Ng < 0.01) {
            wi[0] += (float)(0.01 - cos_wi_Ng) * Ng;
            wi[0] = normalize(wi[0]);
        }
        col[0] = m.mirrorColor * (d.c[0] * Kr);
    } else doReflect = false;
}
YB_HD col3 sd_getTransparency(const Material &m, const Surf &sp, v3 wo)      // :435-455
{
    float accum = 1.f, Kr;
    v3 N = YB_FACE_FORWARD(sp.Ng, sp.N, wo);
    sd_getFresnel(m, wo, N, Kr);
    if (m.isMirror) accum = 1.f - Kr * m.mirrorStrength;
    if (m.isTransparent) accum *= m.transparencyStrength * accum;
    col3 tcol = m.transmitFilterStrength * m.diffuseColor + col3(1.f - m.transmitFilterStrength);
    return accum * tcol;
}
YB_HD float sd_getAlpha(const Material &m, const MatDat &d, const Surf &sp, v3 wo)   // :457-469
{
    if (m.isTransparent) {
        v3 N = YB_FACE_FORWARD(sp.Ng, sp.N, wo);
        float Kr;
        sd_getFresnel(m, wo, N, Kr);
        float refl = (1.f - d.c[0] * Kr) * d.c[1];
        return 1.f - refl;
    }
    return 1.f;
}

// ================================================================== glossyMat_t / coatedGlossyMat_t shared
YB_HD void gl_initBSDF(const Material &m, MatDat &d)                         // glossy.cc:82-95, coatedglossy.cc:104-117
{
    d.c[0] = m.mDiffuse;
    d.c[1] = m.reflectivity;
    d.c[2] = fminf_(0.6f, 1.f - (d.c[1] / (d.c[1] + (1.f - d.c[1]) * d.c[0])));
    d.c[3] = 0.f;
}
YB_HD float gl_OrenNayar(const Material &m, v3 wi, v3 wo, v3 N)               // glossy.cc:105-132
{
    float cos_ti = fmaxf_(-1.f, fminf_(1.f, dot(N, wi)));
    float cos_to = fmaxf_(-1.f, fminf_(1.f, dot(N, wo)));
    float maxcos_f = 0.f;
    if (cos_ti < 0.9999f && cos_to < 0.9999f) {
        v3 v1 = normalize(wi - N * cos_ti);
        v3 v2 = normalize(wo - N * cos_to);
        maxcos_f = fmaxf_(0.f, dot(v1, v2));
    }
    float sin_alpha, tan_beta;
    if (cos_to >= cos_ti) { sin_alpha = fSqrt(1.f - cos_ti * cos_ti); tan_beta = fSqrt(1.f - cos_to * cos_to) / ((cos_to == 0.f) ? 1e-8f : cos_to); }
    else { sin_alpha = fSqrt(1.f - cos_to * cos_to); tan_beta = fSqrt(1.f - cos_ti * cos_ti) / ((cos_ti == 0.f) ? 1e-8f : cos_ti); }
    return m.orenA + m.orenB * maxcos_f * sin_alpha * tan_beta;
}

// ================================================================== glossyMat_t
YB_HD col3 gl_eval(const Material &m, const MatDat &d, const Surf &sp, v3 wo, v3 wi, uint32_t bsdfs)    // glossy.cc:134-173
{
    if (!(bsdfs & BSDF_DIFFUSE) || (dot(sp.Ng, wi) * dot(sp.Ng, wo)) < 0.f) return col3(0.f);
    col3 col(0.f);
    bool diffuse_flag = bsdfs & BSDF_DIFFUSE;
    v3 N = YB_FACE_FORWARD(sp.Ng, sp.N, wo);
    float wiN = fabsf(dot(wi, N)), woN = fabsf(dot(wo, N));
    if ((m.asDiffuse && diffuse_flag) || (!m.asDiffuse && (bsdfs & BSDF_GLOSSY))) {
        v3 H = normalize(wo + wi);
        float cos_wi_H = fmaxf_(0.f, dot(wi, H));
        float glossy;
        if (m.anisotropic) {
            v3 Hs(dot(H, sp.NU), dot(H, sp.NV), dot(H, N));
            glossy = AS_Aniso_D(Hs, m.expU, m.expV) * SchlickFresnel(cos_wi_H, d.c[1]) / ASDivisor(cos_wi_H, woN, wiN);
        } else glossy = Blinn_D(dot(H, N), m.exponent) * SchlickFresnel(cos_wi_H, d.c[1]) / ASDivisor(cos_wi_H, woN, wiN);
        col = glossy * m.glossColor;
    }
    if (m.withDiffuse && diffuse_flag)
        col += (d.c[0] * (1.f - d.c[1])) * m.diffColor * (m.orenNayar ? gl_OrenNayar(m, wi, wo, N) : 1.f);
    return col;
}
YB_HD col3 gl_sample(const Material &m, const MatDat &d, const Surf &sp, v3 wo, v3 &wi, BSample &s, float &W)   // glossy.cc:176-305
{
    float cos_Ng_wo = dot(sp.Ng, wo), cos_Ng_wi;
    v3 N = YB_FACE_FORWARD(sp.Ng, sp.N, wo);
    v3 Hs;
    s.pdf = 0.f;
    float wiN = 0.f, woN = fabsf(dot(wo, N)), cos_wo_H = 0.f;
    col3 scolor(0.f);
    float s1 = s.s1;
    float cur_pDiffuse = d.c[2];
    bool use_glossy = m.asDiffuse ? (s.flags & BSDF_DIFFUSE) : (s.flags & BSDF_GLOSSY);
    bool use_diffuse = m.withDiffuse && (s.flags & BSDF_DIFFUSE);
    float glossy = 0.f;
    if (use_diffuse) {
        float s_pDiffuse = use_glossy ? cur_pDiffuse : 1.f;
        if (s1 < s_pDiffuse) {
            s1 /= s_pDiffuse;
            wi = SampleCosHemisphere(N, sp.NU, sp.NV, s1, s.s2);
            cos_Ng_wi = dot(sp.Ng, wi);
            if (cos_Ng_wi * cos_Ng_wo < 0.f) return scolor;
            wiN = fabsf(dot(wi, N));
            s.pdf = wiN;
            if (use_glossy) {
                v3 H = normalize(wi + wo);
                cos_wo_H = dot(wo, H);
                float cos_wi_H = fabsf(dot(wi, H));
                float cos_N_H = dot(N, H);
                if (m.anisotropic) {
                    v3 Hs2(dot(H, sp.NU), dot(H, sp.NV), cos_N_H);
                    s.pdf = s.pdf * cur_pDiffuse + AS_Aniso_Pdf(Hs2, cos_wo_H, m.expU, m.expV) * (1.f - cur_pDiffuse);
                    glossy = AS_Aniso_D(Hs2, m.expU, m.expV) * SchlickFresnel(cos_wi_H, d.c[1]) / ASDivisor(cos_wi_H, woN, wiN);
                } else {
                    s.pdf = s.pdf * cur_pDiffuse + Blinn_Pdf(cos_N_H, cos_wo_H, m.exponent) * (1.f - cur_pDiffuse);
                    glossy = Blinn_D(cos_N_H, m.exponent) * SchlickFresnel(cos_wi_H, d.c[1]) / ASDivisor(cos_wi_H, woN, wiN);
                }
            }
            s.sampledFlags = BSDF_DIFFUSE | BSDF_REFLECT;
            if (!(s.flags & BSDF_REFLECT)) return col3(0.f);
            scolor = glossy * m.glossColor;
            if (use_diffuse) scolor += diffuseReflect(wiN, woN, d.c[1], d.c[0], m.diffColor) * (m.orenNayar ? gl_OrenNayar(m, wi, wo, N) : 1.f);
            W = wiN / (s.pdf * 0.99f + 0.01f);
            return scolor;
        }
        s1 -= cur_pDiffuse;
        s1 /= (1.f - cur_pDiffuse);
    }
    if (use_glossy) {
        if (m.anisotropic) AS_Aniso_Sample(Hs, s1, s.s2, m.expU, m.expV);
        else Blinn_Sample(Hs, s1, s.s2, m.exponent);
        v3 H = Hs.x * sp.NU + Hs.y * sp.NV + Hs.z * N;
        cos_wo_H = dot(wo, H);
        if (cos_wo_H < 0.f) { H = reflect_about(H, N); cos_wo_H = dot(wo, H); }
        wi = reflect_dir(H, wo);
        cos_Ng_wi = dot(sp.Ng, wi);
        if (cos_Ng_wo * cos_Ng_wi < 0.f) return col3(0.f);
        wiN = fabsf(dot(wi, N));
        if (m.anisotropic) {
            s.pdf = AS_Aniso_Pdf(Hs, cos_wo_H, m.expU, m.expV);
            glossy = AS_Aniso_D(Hs, m.expU, m.expV) * SchlickFresnel(cos_wo_H, d.c[1]) / ASDivisor(cos_wo_H, woN, wiN);
        } else {
            float cosHN = dot(H, N);
            s.pdf = Blinn_Pdf(cosHN, cos_wo_H, m.exponent);
            glossy = Blinn_D(cosHN, m.exponent) * SchlickFresnel(cos_wo_H, d.c[1]) / ASDivisor(cos_wo_H, woN, wiN);
        }
        scolor = glossy * m.glossColor;
        s.sampledFlags = m.asDiffuse ? (BSDF_DIFFUSE | BSDF_REFLECT) : (BSDF_GLOSSY | BSDF_REFLECT);
    }
    if (use_diffuse) {
        s.pdf = wiN * cur_pDiffuse + s.pdf * (1.f - cur_pDiffuse);
        scolor += diffuseReflect(wiN, woN, d.c[1], d.c[0], m.diffColor) * (m.orenNayar ? gl_OrenNayar(m, wi, wo, N) : 1.f);
    }
    W = wiN / (s.pdf * 0.99f + 0.01f);
    return scolor;
}
YB_HD float gl_pdf(const Material &m, const MatDat &d, const Surf &sp, v3 wo, v3 wi, uint32_t flags)    // glossy.cc:307-351
{
    if (dot(sp.Ng, wo) * dot(sp.Ng, wi) < 0.f) return 0.f;
    v3 N = YB_FACE_FORWARD(sp.Ng, sp.N, wo);
    float pdf = 0.f, cos_wo_H, cos_N_H, cur_pDiffuse = d.c[2];
    bool use_glossy = m.asDiffuse ? (flags & BSDF_DIFFUSE) : (flags & BSDF_GLOSSY);
    bool use_diffuse = m.withDiffuse && (flags & BSDF_DIFFUSE);
    if (use_diffuse) {
        pdf = fabsf(dot(wi, N));
        if (use_glossy) {
            v3 H = normalize(wi + wo);
            cos_wo_H = dot(wo, H);
            cos_N_H = dot(N, H);
            if (m.anisotropic) {
                v3 Hs(dot(H, sp.NU), dot(H, sp.NV), cos_N_H);
                pdf = pdf * cur_pDiffuse + AS_Aniso_Pdf(Hs, cos_wo_H, m.expU, m.expV) * (1.f - cur_pDiffuse);
            } else pdf = pdf * cur_pDiffuse + Blinn_Pdf(cos_N_H, cos_wo_H, m.exponent) * (1.f - cur_pDiffuse);
        }
        return pdf;
    }
    if (use_glossy) {
        v3 H = normalize(wi + wo);
        cos_wo_H = dot(wo, H);
        cos_N_H = dot(N, H);
        if (m.anisotropic) {
            v3 Hs(dot(H, sp.NU), dot(H, sp.NV), cos_N_H);
            pdf = AS_Aniso_Pdf(Hs, cos_wo_H, m.expU, m.expV);
        } else pdf = Blinn_Pdf(cos_N_H, cos_wo_H, m.exponent);
    }
    return pdf;
}

// ================================================================== coatedGlossyMat_t
#define YB_C_SPECULAR 0
#define YB_C_GLOSSY 1
#define YB_C_DIFFUSE 2
YB_HD col3 cg_eval(const Material &m, const MatDat &d, const Surf &sp, v3 wo, v3 wi, uint32_t bsdfs)     // coatedglossy.cc:156-193
{
    col3 col(0.f);
    bool diffuse_flag = bsdfs & BSDF_DIFFUSE;
    if (!diffuse_flag || (dot(sp.Ng, wi) * dot(sp.Ng, wo)) < 0.f) return col;
    v3 N = YB_FACE_FORWARD(sp.Ng, sp.N, wo);
    float Kr, Kt;
    float wiN = fabsf(dot(wi, N)), woN = fabsf(dot(wo, N));
    fresnel(wo, N, m.ior, Kr, Kt);
    if ((m.asDiffuse && diffuse_flag) || (!m.asDiffuse && (bsdfs & BSDF_GLOSSY))) {
        v3 H = normalize(wo + wi);
        float cos_wi_H = dot(wi, H);
        float glossy;
        if (m.anisotropic) {
            v3 Hs(dot(H, sp.NU), dot(H, sp.NV), dot(H, N));
            glossy = Kt * AS_Aniso_D(Hs, m.expU, m.expV) * SchlickFresnel(cos_wi_H, d.c[1]) / ASDivisor(cos_wi_H, woN, wiN);
        } else glossy = Kt * Blinn_D(dot(H, N), m.exponent) * SchlickFresnel(cos_wi_H, d.c[1]) / ASDivisor(cos_wi_H, woN, wiN);
        col = glossy * m.glossColor;
    }
    if (m.withDiffuse && diffuse_flag)
        col += (d.c[0] * (1.f - d.c[1])) * m.diffColor * Kt * (m.orenNayar ? gl_OrenNayar(m, wi, wo, N) : 1.f);
    return col;
}
YB_HD col3 cg_sample(const Material &m, const MatDat &d, const Surf &sp, v3 wo, v3 &wi, BSample &s, float &W)   // :195-336
{
    float cos_Ng_wo = dot(sp.Ng, wo), cos_Ng_wi;
    v3 N = YB_FACE_FORWARD(sp.Ng, sp.N, wo);
    v3 Hs(0.f);
    s.pdf = 0.f;
    float Kr, Kt, wiN = 0.f, woN = 0.f;
    fresnel(wo, N, m.ior, Kr, Kt);
    // component list of coatedGlossyMat_t::sample (coatedglossy.cc:213-252) in its fixed order (specular, glossy, diffuse) with static
    // indices: the same partial sums, normalisation (none when a single component matches: width 1) and pick as the reference's
    // arrays val[] / width[] / cIndex[] / rcIndex[], kept in registers; wN[i] = normalised width of component i
    float accumC[3];
    accumC[0] = Kr;
    accumC[1] = Kt * (1.f - d.c[2]);
    accumC[2] = Kt * (d.c[2]);
    const bool use[3] = {m.nBSDF > 0 && (s.flags & m.cFlags[0]) == m.cFlags[0], m.nBSDF > 1 && (s.flags & m.cFlags[1]) == m.cFlags[1],
                         m.nBSDF > 2 && (s.flags & m.cFlags[2]) == m.cFlags[2]};
    float sum = 0.f, val[3] = {0.f, 0.f, 0.f}, wN[3] = {0.f, 0.f, 0.f};
    int nMatch = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int i = 0; i < 3; ++i)
        if (use[i]) { sum += accumC[i]; val[i] = sum; ++nMatch; }
    if (!nMatch || sum < 0.00001) return col3(0.f);
    int pc = -1, lastOn = 0;
    float wPick = 0.f, prevPick = 0.f, lastW = 0.f, lastPrev = 0.f, runPrev = 0.f;
    bool hasPrevPick = false, lastHasPrev = false, anyPrev = false;
    if (nMatch == 1) {
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int i = 0; i < 3; ++i)
            if (use[i]) { pc = i; wN[i] = 1.f; }
        wPick = 1.f;
    } else {
        float inv_sum = 1.f / sum;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int i = 0; i < 3; ++i)
            if (use[i]) {
                const float vN = val[i] * inv_sum;
                wN[i] = accumC[i] * inv_sum;
                if (pc < 0 && s.s1 <= vN) { pc = i; wPick = wN[i]; prevPick = runPrev; hasPrevPick = anyPrev; }
                lastOn = i; lastW = wN[i]; lastPrev = runPrev; lastHasPrev = anyPrev;
                runPrev = vN; anyPrev = true;
            }
        if (pc < 0) { pc = lastOn; wPick = lastW; prevPick = lastPrev; hasPrevPick = lastHasPrev; }
    }
    float s1;
    if (hasPrevPick) s1 = (s.s1 - prevPick) / wPick;
    else s1 = s.s1 / wPick;
    col3 scolor(0.f);
    wi = v3(0.f);   // reference reads an uninitialised wi for the glossy pick before overwriting wiN; value unused
    switch (pc) {
        case YB_C_SPECULAR:
            wi = reflect_dir(N, wo);
            scolor = m.mirrorColor * Kr;
            s.pdf = wPick;
            break;
        case YB_C_GLOSSY:
            if (m.anisotropic) AS_Aniso_Sample(Hs, s1, s.s2, m.expU, m.expV);
            else Blinn_Sample(Hs, s1, s.s2, m.exponent);
            break;
        case YB_C_DIFFUSE:
        default:
            wi = SampleCosHemisphere(N, sp.NU, sp.NV, s1, s.s2);
            cos_Ng_wi = dot(sp.Ng, wi);
            if (cos_Ng_wo * cos_Ng_wi < 0) return col3(0.f);
    }
    wiN = fabsf(dot(wi, N));
    woN = fabsf(dot(wo, N));
    if (pc != YB_C_SPECULAR) {
        if (use[YB_C_GLOSSY]) {
            float glossy, cos_wo_H;
            v3 H(0.f);
            if (pc != YB_C_GLOSSY) {
                H = normalize(wi + wo);
                Hs = v3(dot(H, sp.NU), dot(H, sp.NV), dot(H, N));
                cos_wo_H = dot(wo, H);
            } else {
                H = Hs.x * sp.NU + Hs.y * sp.NV + Hs.z * N;
                cos_wo_H = dot(wo, H);
                if (cos_wo_H < 0.f) { H = reflect_about(H, N); cos_wo_H = dot(wo, H); }
                wi = reflect_dir(H, wo);
                cos_Ng_wi = dot(sp.Ng, wi);
                if (cos_Ng_wo * cos_Ng_wi < 0) return col3(0.f);
            }
            wiN = fabsf(dot(wi, N));
            if (m.anisotropic) {
                s.pdf += AS_Aniso_Pdf(Hs, cos_wo_H, m.expU, m.expV) * wN[YB_C_GLOSSY];
                glossy = AS_Aniso_D(Hs, m.expU, m.expV) * SchlickFresnel(cos_wo_H, d.c[1]) / ASDivisor(cos_wo_H, woN, wiN);
            } else {
                float cosHN = dot(H, N);
                s.pdf += Blinn_Pdf(cosHN, cos_wo_H, m.exponent) * wN[YB_C_GLOSSY];
                glossy = Blinn_D(cosHN, m.exponent) * SchlickFresnel(cos_wo_H, d.c[1]) / ASDivisor(cos_wo_H, woN, wiN);
            }
            scolor = (glossy * Kt) * m.glossColor;
        }
        if (use[YB_C_DIFFUSE]) {
            scolor += Kt * diffuseReflect(wiN, woN, d.c[1], d.c[0], m.diffColor) * (m.orenNayar ? gl_OrenNayar(m, wi, wo, N) : 1.f);
            s.pdf += wiN * wN[YB_C_DIFFUSE];
        }
        W = wiN / (s.pdf * 0.99f + 0.01f);
    } else W = 1.f;
    s.sampledFlags = pc == 0 ? m.cFlags[0] : (pc == 1 ? m.cFlags[1] : m.cFlags[2]);
    return scolor;
}
YB_HD float cg_pdf(const Material &m, const MatDat &d, const Surf &sp, v3 wo, v3 wi, uint32_t flags)     // :338-382
{
    bool transmit = (dot(sp.Ng, wo) * dot(sp.Ng, wi)) < 0.f;
    if (transmit) return 0.f;
    v3 N = YB_FACE_FORWARD(sp.Ng, sp.N, wo);
    float pdf = 0.f, Kr, Kt;
    fresnel(wo, N, m.ior, Kr, Kt);
    float accumC[3], sum = 0.f, width;
    accumC[0] = Kr;
    accumC[1] = Kt * (1.f - d.c[2]);
    accumC[2] = Kt * (d.c[2]);
    int nMatch = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int i = 0; i < 3; ++i) {           // static indices (see cg_sample)
        if (i < m.nBSDF && (flags & m.cFlags[i]) == m.cFlags[i]) {
            width = accumC[i];
            sum += width;
            if (i == YB_C_GLOSSY) {
                v3 H = normalize(wi + wo);
                float cos_wo_H = dot(wo, H), cos_N_H = dot(N, H);
                if (m.anisotropic) {
                    v3 Hs(dot(H, sp.NU), dot(H, sp.NV), cos_N_H);
                    pdf += AS_Aniso_Pdf(Hs, cos_wo_H, m.expU, m.expV) * width;
                } else pdf += Blinn_Pdf(cos_N_H, cos_wo_H, m.exponent) * width;
            } else if (i == YB_C_DIFFUSE) pdf += fabsf(dot(wi, N)) * width;
            ++nMatch;
        }
    }
    if (!nMatch || sum < 0.00001) return 0.f;
    return pdf / sum;
}
// glass.cc:113-116 / coatedglossy.cc:387-398: the "bent" shading normal
YB_HD v3 bent_normal(const Surf &sp, v3 wo, bool outside)
{
    float cos_wo_N = dot(sp.N, wo);
    if (outside) return (cos_wo_N >= 0) ? sp.N : normalize(sp.N - (float)(1.00001 * cos_wo_N) * wo);
    return (cos_wo_N <= 0) ? sp.N : normalize(sp.N - (float)(1.00001 * cos_wo_N) * wo);
}
YB_HD void cg_getSpecular(const Material &m, const Surf &sp, v3 wo, int raylevel, bool &refl, bool &refr, v3 *dir, col3 *col)  // :384-418
{
    bool outside = dot(sp.Ng, wo) >= 0;
    v3 N = bent_normal(sp, wo, outside);
    v3 Ng = outside ? sp.Ng : -sp.Ng;
    float Kr, Kt;
    fresnel(wo, N, m.ior, Kr, Kt);
    refr = false;
    if (raylevel > 5) return;
    dir[0] = reflect_about(wo, N);
    col[0] = Kr * m.mirrorColor;
    float cos_wi_Ng = dot(dir[0], Ng);
    if (cos_wi_Ng < 0.01) {
        dir[0] += (float)(0.01 - cos_wi_Ng) * Ng;
        dir[0] = normalize(dir[0]);
    }
    refl = true;
}

// ================================================================== glassMat_t (no dispersion, no absorption: rejected at creation)
#define YB_MATCHES(bits, flags) (((bits) & (flags)) == (flags))
YB_HD col3 glass_sample(const Material &m, const Surf &sp, v3 wo, v3 &wi, BSample &s, float &W)          // glass.cc:84-188 (else-branch :143-186)
{
    if (!(s.flags & BSDF_SPECULAR)) { s.pdf = 0.f; return col3(0.f); }
    v3 refdir;
    bool outside = dot(sp.Ng, wo) > 0;
    v3 N = bent_normal(sp, wo, outside);
    s.pdf = 1.f;
    const uint32_t tmFlags = m.fakeShadow ? (BSDF_FILTER | BSDF_TRANSMIT) : (BSDF_SPECULAR | BSDF_TRANSMIT);
    if (refract(N, wo, refdir, m.ior)) {
        float Kr, Kt;
        fresnel(wo, N, m.ior, Kr, Kt);
        float pKr = (float)(0.01 + 0.99 * Kr), pKt = (float)(0.01 + 0.99 * Kt);
        if (s.s1 < pKt && YB_MATCHES(s.flags, tmFlags)) {
            wi = refdir;
            s.pdf = pKt;
            s.sampledFlags = tmFlags;
            W = 1.f;
            return m.filterCol;
        } else if (YB_MATCHES(s.flags, BSDF_SPECULAR | BSDF_REFLECT)) {
            wi = reflect_about(wo, N);
            s.pdf = pKr;
            s.sampledFlags = BSDF_SPECULAR | BSDF_REFLECT;
            W = 1.f;
            return m.specRefCol;
        }
    } else if (YB_MATCHES(s.flags, BSDF_SPECULAR | BSDF_REFLECT)) {   // total inner reflection
        wi = reflect_about(wo, N);
        s.sampledFlags = BSDF_SPECULAR | BSDF_REFLECT;
        W = 1.f;
        return col3(1.f);
    }
    s.pdf = 0.f;
    return col3(0.f);
}
YB_HD col3 glass_getTransparency(const Material &m, const Surf &sp, v3 wo)                                // :190-196
{
    v3 N = YB_FACE_FORWARD(sp.Ng, sp.N, wo);
    float Kr, Kt;
    fresnel(wo, N, m.ior, Kr, Kt);
    return Kt * m.filterCol;
}
YB_HD float col_energy(col3 c) { return (c.x + c.y + c.z) * 0.333333f; }                                  // color.h energy()
YB_HD void glass_getSpecular(const Material &m, const Surf &sp, v3 wo, int raylevel, bool &refl, bool &refr, v3 *dir, col3 *col)  // :204-247
{
    bool outside = dot(sp.Ng, wo) > 0;
    v3 N = bent_normal(sp, wo, outside);
    v3 refdir;
    if (refract(N, wo, refdir, m.ior)) {
        float Kr, Kt;
        fresnel(wo, N, m.ior, Kr, Kt);
        col[1] = Kt * m.filterCol;
        dir[1] = refdir;
        refr = true;
        if (outside || raylevel < 3) {
            dir[0] = reflect_about(wo, N);
            col[0] = m.specRefCol * Kr;
            refl = true;
        } else refl = false;
    } else {
        col[0] = col3(1.f);
        dir[0] = reflect_about(wo, N);
        refl = true;
        refr = false;
    }
}

// ================================================================== roughGlassMat_t (no dispersion, no absorption: rejected at creation)
// microfacet.h:118-193
YB_HD void GGX_Sample(v3 &H, float alpha2, float s1, float s2)
{
    float tanTheta2 = alpha2 * (s1 / (1.00001f - s1));
    float cosTheta = 1.f / fSqrt(1.f + tanTheta2);
    float sinTheta = fSqrt(1.00001f - (cosTheta * cosTheta));
    float phi = (float)(YB_M_2PI * (double)s2);
    H = v3(sinTheta * fCos(phi), sinTheta * fSin(phi), cosTheta);
}
YB_HD float GGX_D(float alpha2, float cosTheta2, float tanTheta2)
{
    float cosTheta4 = cosTheta2 * cosTheta2;
    float aTan = alpha2 + tanTheta2;
    float div = (float)(YB_M_PI * (double)cosTheta4 * (double)aTan * (double)aTan);
    return alpha2 / div;
}
YB_HD float GGX_G(float alpha2, float woN, float wiN)
{
    float woN2 = woN * woN, wiN2 = wiN * wiN;
    float sqrTerm1 = fSqrt(1.f + alpha2 * ((1.f - woN2) / woN2));
    float sqrTerm2 = fSqrt(1.f + alpha2 * ((1.f - wiN2) / wiN2));
    float G1wo = 2.f / (1.f + sqrTerm1), G1wi = 2.f / (1.f + sqrTerm2);
    return G1wo * G1wi;
}
YB_HD float GGX_Pdf(float D, float cosTheta, float Jacobian) { return D * cosTheta * Jacobian; }
YB_HD float microfacetFresnel(float woH, float ior)
{
    float c = fabsf(woH);
    float g = ior * ior - 1 + c * c;
    if (g > 0) {
        g = fSqrt(g);
        float A = (g - c) / (g + c);
        float B = (c * (g + c) - 1) / (c * (g - c) + 1);
        return 0.5f * A * A * (1 + B * B);
    }
    return 1.0f;
}
YB_HD bool refractMicrofacet(float eta, v3 wo, v3 &wi, v3 H, float woH, float woN, float &Kr, float &Kt)
{
    wi = v3(0.f);
    float c = -dot(wo, H);
    float sign = (c > 0.f) ? 1.f : -1.f;
    float t1 = 1 - (eta * eta * (1 - c * c));
    if (t1 < 0.f) return false;
    wi = eta * wo + (eta * c - sign * fSqrt(t1)) * H;
    wi = -wi;
    Kr = 0.f; Kt = 0.f;
    Kr = microfacetFresnel(woH, 1.f / eta);
    if (Kr == 1.f) return false;
    Kt = 1 - Kr;
    return true;
}
YB_HD void reflectMicrofacet(v3 wo, v3 &wi, v3 H, float woH)
{
    wi = wo + (2.f * dot(H, -wo) * H);
    wi = -wi;
}
// The two sample() overloads of roughGlassMat_t share everything up to the Fresnel split (roughglass.cc:56-90,146-181)
struct RGState { v3 N, H, wiT; bool outside, refr; float D, cosTheta, woH, woN, Kr, Kt; };
YB_HD void rg_setup(const Material &m, const Surf &sp, v3 wo, float s1, float s2, RGState &r)
{
    r.N = YB_FACE_FORWARD(sp.Ng, sp.N, wo);
    r.outside = dot(sp.Ng, wo) > 0.f;
    v3 H(0.f);
    GGX_Sample(H, m.rgAlpha2, s1, s2);
    H = H.x * sp.NU + H.y * sp.NV + H.z * r.N;
    r.H = normalize(H);
    r.D = 0.f;
    r.cosTheta = dot(r.H, r.N);
    float cosTheta2 = r.cosTheta * r.cosTheta;
    float tanTheta2 = (1.f - cosTheta2) / (cosTheta2 * 0.99f + 0.01f);
    if (r.cosTheta > 0.f) r.D = GGX_D(m.rgAlpha2, cosTheta2, tanTheta2);
    r.woH = dot(wo, r.H);
    r.woN = dot(wo, r.N);
    r.Kr = r.Kt = 0.f;
    r.refr = refractMicrofacet(r.outside ? 1.f / m.ior : m.ior, wo, r.wiT, r.H, r.woH, r.woN, r.Kr, r.Kt);
}
YB_HD col3 rg_transmit(const Material &m, const RGState &r, v3 wi, float &pdf, float &W)    // roughglass.cc:94-116
{
    float wiN = dot(wi, r.N), wiH = dot(wi, r.H), G = 0.f;
    if ((wiH * wiN) > 0.f && (r.woH * r.woN) > 0.f) G = GGX_G(m.rgAlpha2, wiN, r.woN);
    float IORwi = 1.f, IORwo = 1.f;
    if (r.outside) IORwi = m.ior; else IORwo = m.ior;
    float ht = IORwo * r.woH + IORwi * wiH;
    float Jacobian = (IORwi * IORwi) / fmaxf_(1.0e-8f, ht * ht);
    float glossy = fabsf((r.woH * wiH) / (wiN * r.woN)) * r.Kt * G * r.D * Jacobian;
    pdf = GGX_Pdf(r.D, r.cosTheta, Jacobian * fabsf(wiH));
    W = fabsf(wiN) / (pdf * 0.99f + 0.01f);
    return glossy * m.filterCol;
}
YB_HD col3 rg_reflect(const Material &m, const RGState &r, v3 wo, v3 &wi, float &pdf, float &W)   // roughglass.cc:118-136
{
    reflectMicrofacet(wo, wi, r.H, r.woH);
    float wiN = dot(wi, r.N), wiH = dot(wi, r.H);
    float G = GGX_G(m.rgAlpha2, wiN, r.woN);
    float Jacobian = 1.f / ((4.f * fabsf(wiH)) * 0.99f + 0.01f);
    float glossy = (r.Kr * G * r.D) / ((4.f * fabsf(r.woN * wiN)) * 0.99f + 0.01f);
    pdf = GGX_Pdf(r.D, r.cosTheta, Jacobian);
    W = fabsf(wiN) / (pdf * 0.99f + 0.01f);
    return glossy * m.specRefCol;
}
YB_HD col3 rg_sample(const Material &m, const Surf &sp, v3 wo, v3 &wi, BSample &s, float &W)      // roughglass.cc:56-139
{
    RGState r;
    s.pdf = 1.f;
    rg_setup(m, sp, wo, s.s1, s.s2, r);
    wi = r.wiT;
    col3 ret(0.f);
    if (r.refr) {
        if (s.s1 < r.Kt && (s.flags & BSDF_TRANSMIT)) {
            ret = rg_transmit(m, r, wi, s.pdf, W);
            s.sampledFlags = BSDF_GLOSSY | BSDF_TRANSMIT;
        } else if (s.flags & BSDF_REFLECT) {
            ret = rg_reflect(m, r, wo, wi, s.pdf, W);
            s.sampledFlags = BSDF_GLOSSY | BSDF_REFLECT;
        }
    } else {                                                            // total inner reflection
        wi = reflect_about(wo, r.H);
        s.sampledFlags = BSDF_GLOSSY | BSDF_REFLECT;
        ret = col3(1.f);
        W = 1.f;
    }
    return ret;
}
// the two-direction overload used by recursiveRaytrace (roughglass.cc:141-236): slot 0 = refraction (or the TIR mirror
// direction), slot 1 = reflection; the integrator reads slot 0 under BSDF_REFLECT and slot 1 under BSDF_TRANSMIT
// (mcintegrator.cc:551-575), so both slots are traced whenever refraction exists
YB_HD void rg_sample2(const Material &m, const Surf &sp, v3 wo, v3 *dir, col3 *col, float *W, BSample &s)
{
    RGState r;
    s.pdf = 1.f;
    rg_setup(m, sp, wo, s.s1, s.s2, r);
    s.sampledFlags = 0;
    col[0] = col[1] = col3(0.f); W[0] = W[1] = 0.f; dir[0] = dir[1] = v3(0.f);
    if (r.refr) {
        if (s.flags & BSDF_TRANSMIT) {
            col[0] = rg_transmit(m, r, r.wiT, s.pdf, W[0]);
            s.sampledFlags = BSDF_GLOSSY | BSDF_TRANSMIT;
            dir[0] = r.wiT;
        }
        if (s.flags & BSDF_REFLECT) {
            v3 wi;
            col[1] = rg_reflect(m, r, wo, wi, s.pdf, W[1]);
            s.sampledFlags |= BSDF_GLOSSY | BSDF_REFLECT;
            dir[1] = wi;
        }
    } else {
        dir[0] = reflect_about(wo, r.H);
        s.sampledFlags |= BSDF_GLOSSY | BSDF_REFLECT;
        col[0] = col3(1.f);
        W[0] = 1.f;
    }
}

// ================================================================== dispatch over the leaf materials: material_t virtuals
// FIX >= 0: the material type is known at compile time (staged direct lighting: one kernel per material type, the dispatch folds)
template <int FIX = -1> YB_HD void leaf_initBSDF(const Material &m, MatDat &d, uint32_t &bsdfs)
{
    bsdfs = m.bsdfFlags;
    switch (FIX >= 0 ? FIX : m.type) {
        case MAT_SHINYDIFFUSE: sd_initBSDF(m, d); break;
        case MAT_GLOSSY: case MAT_COATED_GLOSSY: gl_initBSDF(m, d); break;
        case MAT_NULL: bsdfs = BSDF_NONE;   // glass.cc:374
        default: d.c[0] = d.c[1] = d.c[2] = d.c[3] = 0.f; break;
    }
}
template <int FIX = -1> YB_HD col3 leaf_eval(const Material &m, const MatDat &d, const Surf &sp, v3 wo, v3 wl, uint32_t bsdfs)
{
    switch (FIX >= 0 ? FIX : m.type) {
        case MAT_SHINYDIFFUSE: return sd_eval(m, d, sp, wo, wl, bsdfs);
        case MAT_GLOSSY: return gl_eval(m, d, sp, wo, wl, bsdfs);
        case MAT_COATED_GLOSSY: return cg_eval(m, d, sp, wo, wl, bsdfs);
        default: return col3(0.f);
    }
}
template <int EXT, int FIX = -1> YB_HD col3 leaf_sample(const Material &m, const MatDat &d, const Surf &sp, v3 wo, v3 &wi, BSample &s, float &W)
{
    switch (FIX >= 0 ? FIX : m.type) {
        case MAT_SHINYDIFFUSE: return sd_sample(m, d, sp, wo, wi, s, W);
        case MAT_GLOSSY: return gl_sample(m, d, sp, wo, wi, s, W);
        case MAT_COATED_GLOSSY: return cg_sample(m, d, sp, wo, wi, s, W);
        case MAT_GLASS: return glass_sample(m, sp, wo, wi, s, W);
        case MAT_ROUGH_GLASS: if (EXT) return rg_sample(m, sp, wo, wi, s, W);   // falls through to the empty default in builds without it
        case MAT_MIRROR:                                                 // glass.cc:348-354
            wi = reflect_dir(sp.N, wo);
            s.sampledFlags = BSDF_SPECULAR | BSDF_REFLECT;
            W = 1.f;
            return m.specRefCol * (1.f / fabsf(dot(sp.N, wi)));
        default:                                                         // nullMat_t / lightMat_t
            s.pdf = 0.f; W = 0.f; wi = v3(0.f);
            return col3(0.f);
    }
}
template <int FIX = -1> YB_HD float leaf_pdf(const Material &m, const MatDat &d, const Surf &sp, v3 wo, v3 wi, uint32_t bsdfs)
{
    switch (FIX >= 0 ? FIX : m.type) {
        case MAT_SHINYDIFFUSE: return sd_pdf(m, d, sp, wo, wi, bsdfs);
        case MAT_GLOSSY: return gl_pdf(m, d, sp, wo, wi, bsdfs);
        case MAT_COATED_GLOSSY: return cg_pdf(m, d, sp, wo, wi, bsdfs);
        default: return 0.f;
    }
}
template <int FIX = -1> YB_HD col3 leaf_emit(const Material &m, const Surf &sp, v3 wo, bool includeLights)
{
    const int ty = FIX >= 0 ? FIX : m.type;
    if (ty == MAT_SHINYDIFFUSE) return m.emitColor;                     // shinydiffuse.cc:251-257
    if (ty == MAT_LIGHT) {                                          // simple.cc:63-70
        if (!includeLights) return col3(0.f);
        if (m.doubleSided) return m.lightCol;
        return (dot(wo, sp.N) > 0) ? m.lightCol : col3(0.f);
    }
    return col3(0.f);
}
template <int FIX = -1> YB_HD void leaf_getSpecular(const Material &m, const MatDat &d, const Surf &sp, v3 wo, int raylevel, bool &refl, bool &refr, v3 *dir, col3 *col)
{
    switch (FIX >= 0 ? FIX : m.type) {
        case MAT_SHINYDIFFUSE: sd_getSpecular(m, d, sp, wo, refl, refr, dir, col); break;
        case MAT_COATED_GLOSSY: cg_getSpecular(m, sp, wo, raylevel, refl, refr, dir, col); break;
        case MAT_GLASS: glass_getSpecular(m, sp, wo, raylevel, refl, refr, dir, col); break;
        case MAT_MIRROR: {                                               // glass.cc:356-364
            col[0] = m.specRefCol; col[1] = col3(1.f);
            v3 N = YB_FACE_FORWARD(sp.Ng, sp.N, wo);
            dir[0] = reflect_dir(N, wo);
            refl = true; refr = false;
            break;
        }
        default: refl = false; refr = false; break;                      // material.h:141-144
    }
}
YB_HD bool leaf_isTransparent(const Material &m)
{
    if (m.type == MAT_SHINYDIFFUSE) return m.isTransparent != 0;        // shinydiff.h isTransparent(){return mIsTransparent;}
    if (m.type == MAT_GLASS || m.type == MAT_ROUGH_GLASS) return m.fakeShadow != 0;
    return false;
}
YB_HD col3 leaf_getTransparency(const Material &m, const Surf &sp, v3 wo)
{
    if (m.type == MAT_SHINYDIFFUSE) return sd_getTransparency(m, sp, wo);
    if (m.type == MAT_GLASS || m.type == MAT_ROUGH_GLASS) return glass_getTransparency(m, sp, wo);   // roughglass.cc:238-244 is the same formula
    return col3(0.f);
}
template <int FIX = -1> YB_HD float leaf_getAlpha(const Material &m, const MatDat &d, const Surf &sp, v3 wo)
{
    const int ty = FIX >= 0 ? FIX : m.type;
    if (ty == MAT_SHINYDIFFUSE) return sd_getAlpha(m, d, sp, wo);
    if (ty == MAT_GLASS) {                                          // glass.cc:198-202
        float alpha = (float)(1.0 - col_energy(glass_getTransparency(m, sp, wo)));
        if (alpha < 0.0f) alpha = 0.0f;
        return alpha;
    }
    if (ty == MAT_ROUGH_GLASS)                                          // roughglass.cc:246-250
        return fmaxf_(0.f, fminf_(1.f, 1.f - col_energy(glass_getTransparency(m, sp, wo))));
    return 1.f;
}

// beer_t::transmittance (beer.cc:35-45) over a ray that ended at distance t (no hit: tmax stays -1 - black)
YB_HD col3 beer_transmittance(col3 sigma, bool hit, float t)
{
    if (!hit || t < 0.f || t > 1e30f) return col3(0.f);
    return col3(fExp(-t * sigma.x), fExp(-t * sigma.y), fExp(-t * sigma.z));
}

// ================================================================== blendMat_t (blend.cc:49-393) over two leaf materials, constant blend_value
// EXT: the shade kernel is compiled twice (yb_render.cu); scenes without rough_glass / blend_mat / mesh lights run the EXT = 0
// build, which carries none of this code (register pressure of the common path)
// userdata of the sub-materials (SDDat_t / MDat_t) depends only on the material constants here, so it is recomputed
// where the reference reads it back from state.userdata
YB_HD const Material &blend_m1(const Material &m) { return (&m)[m.blendOff1]; }
YB_HD const Material &blend_m2(const Material &m) { return (&m)[m.blendOff2]; }
YB_HD float blend_ival(float val) { return fminf_(1.f, fmaxf_(0.f, 1.f - val)); }                       // :60
YB_HD float blend_val(const Material &m, const Surf &sp) { return sp.hasBlendVal ? sp.blendVal : m.blendVal; }   // getBlendVal :49-66
YB_HD uint32_t blend_flags(const Material &sub) { return sub.type == MAT_NULL ? (uint32_t)BSDF_NONE : sub.bsdfFlags; }   // mat1Flags / mat2Flags

template <int EXT, int FIX = -1> YB_HD void mat_initBSDF(const Material &m, MatDat &d, uint32_t &bsdfs)
{
    if (!EXT || (FIX >= 0 ? FIX : m.type) != MAT_BLEND) { leaf_initBSDF<FIX>(m, d, bsdfs); return; }
    d.c[0] = d.c[1] = d.c[2] = d.c[3] = 0.f;
    bsdfs = blend_flags(blend_m1(m)) | blend_flags(blend_m2(m));
}
// blendMat_t::initBSDF replaces the surface point by blend_surface_points(sp, sp, ival) (blend.cc:72-82, surface.cc:82-96,
// interpolation.h:29-35): without bump mapping both inputs are equal, and v*a + v*(1-a) is kept as the reference computes it
template <int EXT> YB_HD void mat_blendSurface(const Material &m, Surf &sp)
{
    if (!EXT || m.type != MAT_BLEND) return;
    const double alpha = (double)blend_ival(blend_val(m, sp));
    const float a = (float)alpha, b = (float)(1.0 - alpha);
    sp.N = sp.N * a + sp.N * b;
    sp.NU = sp.NU * a + sp.NU * b;
    sp.NV = sp.NV * a + sp.NV * b;
}
template <int EXT, int FIX = -1> YB_HD col3 mat_eval(const Material &m, const MatDat &d, const Surf &sp, v3 wo, v3 wl, uint32_t bsdfs)
{
    if (!EXT || (FIX >= 0 ? FIX : m.type) != MAT_BLEND) return leaf_eval<FIX>(m, d, sp, wo, wl, bsdfs);
    const float val = blend_val(m, sp), ival = blend_ival(val);
    MatDat d1, d2; uint32_t f;
    leaf_initBSDF(blend_m1(m), d1, f); leaf_initBSDF(blend_m2(m), d2, f);
    col3 col1 = leaf_eval(blend_m1(m), d1, sp, wo, wl, bsdfs);
    col3 col2 = leaf_eval(blend_m2(m), d2, sp, wo, wl, bsdfs);
    return col1 * ival + col2 * val;
}
template <int EXT, int FIX = -1> YB_HD col3 mat_sample(const Material &m, const MatDat &d, const Surf &sp, v3 wo, v3 &wi, BSample &s, float &W)
{
    if (!EXT || (FIX >= 0 ? FIX : m.type) != MAT_BLEND) return leaf_sample<EXT, FIX>(m, d, sp, wo, wi, s, W);
    const float val = blend_val(m, sp), ival = blend_ival(val);
    const Material &m1 = blend_m1(m), &m2 = blend_m2(m);
    bool mat1Sampled = false, mat2Sampled = false;
    col3 col1(0.f), col2(0.f);
    BSample s1 = s, s2 = s;
    v3 wi1(0.f), wi2(0.f);
    float W1 = 0.f, W2 = 0.f;
    s2.pdf = s1.pdf = s.pdf = 0.f;
    MatDat dd; uint32_t f;
    if (s.flags & blend_flags(m1)) { leaf_initBSDF(m1, dd, f); col1 = leaf_sample<EXT>(m1, dd, sp, wo, wi1, s1, W1); mat1Sampled = true; }
    if (s.flags & blend_flags(m2)) { leaf_initBSDF(m2, dd, f); col2 = leaf_sample<EXT>(m2, dd, sp, wo, wi2, s2, W2); mat2Sampled = true; }
    if (mat1Sampled && mat2Sampled) {
        wi = normalize(wi2 + wi1);
        s.pdf = s1.pdf * ival + s2.pdf * val;
        s.sampledFlags = s1.sampledFlags | s2.sampledFlags;
        col1 = (col1 * W1) * ival + (col2 * W2) * val;
        W = 1.f;
    } else if (mat1Sampled) {
        wi = wi1; s.pdf = s1.pdf; s.sampledFlags = s1.sampledFlags; W = W1;
    } else if (mat2Sampled) {
        wi = wi2; s.pdf = s2.pdf; s.sampledFlags = s2.sampledFlags; col1 = col2; W = W2;
    }
    return col1;
}
template <int EXT, int FIX = -1> YB_HD float mat_pdf(const Material &m, const MatDat &d, const Surf &sp, v3 wo, v3 wi, uint32_t bsdfs)
{
    if (!EXT || (FIX >= 0 ? FIX : m.type) != MAT_BLEND) return leaf_pdf<FIX>(m, d, sp, wo, wi, bsdfs);
    const float val = blend_val(m, sp), ival = blend_ival(val);
    MatDat d1, d2; uint32_t f;
    leaf_initBSDF(blend_m1(m), d1, f); leaf_initBSDF(blend_m2(m), d2, f);
    float pdf1 = leaf_pdf(blend_m1(m), d1, sp, wo, wi, bsdfs), pdf2 = leaf_pdf(blend_m2(m), d2, sp, wo, wi, bsdfs);
    return pdf1 * ival + pdf2 * val;
}
template <int EXT, int FIX = -1> YB_HD col3 mat_emit(const Material &m, const Surf &sp, v3 wo, bool includeLights)
{
    if (!EXT || (FIX >= 0 ? FIX : m.type) != MAT_BLEND) return leaf_emit<FIX>(m, sp, wo, includeLights);
    const float val = blend_val(m, sp), ival = blend_ival(val);
    col3 col1 = leaf_emit(blend_m1(m), sp, wo, includeLights), col2 = leaf_emit(blend_m2(m), sp, wo, includeLights);
    return col1 * ival + col2 * val;
}
template <int EXT, int FIX = -1> YB_HD void mat_getSpecular(const Material &m, const MatDat &d, const Surf &sp, v3 wo, int raylevel, bool &refl, bool &refr, v3 *dir, col3 *col)
{
    if (!EXT || (FIX >= 0 ? FIX : m.type) != MAT_BLEND) { leaf_getSpecular<FIX>(m, d, sp, wo, raylevel, refl, refr, dir, col); return; }
    const float val = blend_val(m, sp), ival = blend_ival(val);
    refl = false; refr = false;
    bool m1_refl = false, m1_refr = false;
    v3 m1_dir[2]; col3 m1_col[2];
    m1_col[0] = m1_col[1] = col3(0.f); m1_dir[0] = m1_dir[1] = v3(0.f);
    MatDat dd; uint32_t f;
    leaf_initBSDF(blend_m1(m), dd, f);
    leaf_getSpecular(blend_m1(m), dd, sp, wo, raylevel, m1_refl, m1_refr, m1_dir, m1_col);
    leaf_initBSDF(blend_m2(m), dd, f);
    leaf_getSpecular(blend_m2(m), dd, sp, wo, raylevel, refl, refr, dir, col);
    if (refl && m1_refl) { col[0] = m1_col[0] * ival + col[0] * val; dir[0] = normalize(dir[0] + m1_dir[0]); }
    else if (m1_refl) { col[0] = m1_col[0] * ival; dir[0] = m1_dir[0]; }
    else col[0] = col[0] * val;
    if (refr && m1_refr) { col[1] = m1_col[1] * ival + col[1] * val; dir[1] = normalize(dir[1] + m1_dir[1]); }
    else if (m1_refr) { col[1] = m1_col[1] * ival; dir[1] = m1_dir[1]; }
    else col[1] = col[1] * val;
    refr = refr || m1_refr;
    refl = refl || m1_refl;
}
template <int EXT> YB_HD bool mat_isTransparent(const Material &m)
{
    if (!EXT || m.type != MAT_BLEND) return leaf_isTransparent(m);
    return leaf_isTransparent(blend_m1(m)) || leaf_isTransparent(blend_m2(m));
}
template <int EXT> YB_HD col3 mat_getTransparency(const Material &m, const Surf &sp, v3 wo)
{
    if (!EXT || m.type != MAT_BLEND) return leaf_getTransparency(m, sp, wo);
    const float val = blend_val(m, sp), ival = blend_ival(val);
    col3 col1 = leaf_getTransparency(blend_m1(m), sp, wo), col2 = leaf_getTransparency(blend_m2(m), sp, wo);
    return col1 * ival + col2 * val;
}
template <int EXT, int FIX = -1> YB_HD float mat_getAlpha(const Material &m, const MatDat &d, const Surf &sp, v3 wo)
{
    if (!EXT || (FIX >= 0 ? FIX : m.type) != MAT_BLEND) return leaf_getAlpha<FIX>(m, d, sp, wo);
    if (!mat_isTransparent<EXT>(m)) return 1.f;
    MatDat d1, d2; uint32_t f;
    leaf_initBSDF(blend_m1(m), d1, f); leaf_initBSDF(blend_m2(m), d2, f);
    float al1 = leaf_getAlpha(blend_m1(m), d1, sp, wo), al2 = leaf_getAlpha(blend_m2(m), d2, sp, wo);
    return fminf_(al1, al2);
}

// ================================================================== node materials: constants <- shader nodes
// Evaluates the material's nodes for this surface point and returns a Material whose constants are the node outputs
// (EXT >= 2 builds only; materials without nodes, and every material in the other builds, come back untouched).
// Slot use per material: shinydiffuse.cc:82-98,242-256,305-329,407,427,447-453; glossy.cc:93,164-169; coatedglossy.cc:115,186-190;
// glass.cc:123,168,240; roughglass.cc:132.
// sp != 0: the call stands for initBSDF, which first bumps the shading frame (nodeMaterial_t::evalBump, nodematerial.cc:132-139 +
// material_t::applyBump, material.cc:68-75); getTransparency for shadow rays (sp == 0) does not.
template <int EXT> YB_HD const Material &mat_resolve(const SceneDev &sc, const Material &m, v3 P, v3 Ng, const TexPoint &tp, Material &tmp, Surf *sp = 0)
{
    if (EXT < 2 || m.nNodes <= 0) return m;
    NodeRes st[YB_MAX_NODES];
    if (m.type == MAT_BLEND) {          // only the mask is a shader node here (blend.cc:406-414): its scalar is the blend value of this point
        for (int i = 0; i < m.nNodes && i < YB_MAX_NODES; ++i) node_eval(sc.shaderNodes[m.nodeBase + i], sc.textures, P, Ng, tp, st, i);
        if (sp && m.slot[SLOT_BLEND_MASK] >= 0) { sp->blendVal = st[m.slot[SLOT_BLEND_MASK]].f; sp->hasBlendVal = 1; }
        return m;
    }
    if (sp && m.slot[SLOT_BUMP] >= 0) {
        for (int i = 0; i < m.nNodes && i < YB_MAX_NODES; ++i)
            if (m.bumpMask & (1u << i)) node_evalDerivative(sc.shaderNodes[m.nodeBase + i], sc.textures, P, Ng, sp->NU, sp->NV, tp, st, i);
        const float du = st[m.slot[SLOT_BUMP]].r, dv = st[m.slot[SLOT_BUMP]].g;
        sp->NU += du * sp->N;
        sp->NV += dv * sp->N;
        sp->N = normalize(cross(sp->NU, sp->NV));
        sp->NU = normalize(sp->NU);
        sp->NV = normalize(cross(sp->N, sp->NU));
    }
    for (int i = 0; i < m.nNodes && i < YB_MAX_NODES; ++i) node_eval(sc.shaderNodes[m.nodeBase + i], sc.textures, P, Ng, tp, st, i);
    tmp = m;
    tmp.nNodes = 0;
    int s;
    if ((s = m.slot[SLOT_DIFFUSE]) >= 0) {
        const col3 c(st[s].r, st[s].g, st[s].b);
        if (m.type == MAT_SHINYDIFFUSE) { tmp.diffuseColor = c; tmp.emitColor = c * m.emitStrength; }
        else tmp.diffColor = c;
    }
    if ((s = m.slot[SLOT_MIRROR_COLOR]) >= 0) {
        const col3 c(st[s].r, st[s].g, st[s].b);
        if (m.type == MAT_SHINYDIFFUSE) tmp.mirrorColor = c; else tmp.specRefCol = c;       // glass / rough_glass: mirColS
    }
    if ((s = m.slot[SLOT_MIRROR]) >= 0) tmp.mirrorStrength = st[s].f;
    if ((s = m.slot[SLOT_TRANSPARENCY]) >= 0) tmp.transparencyStrength = st[s].f;
    if ((s = m.slot[SLOT_TRANSLUCENCY]) >= 0) tmp.translucencyStrength = st[s].f;
    if ((s = m.slot[SLOT_GLOSSY]) >= 0) tmp.glossColor = col3(st[s].r, st[s].g, st[s].b);
    if ((s = m.slot[SLOT_GLOSSY_REFLECT]) >= 0) tmp.reflectivity = st[s].f;
    return tmp;
}

}  // namespace yb
