/* oracle/yb_oracle.c — TEST INFRASTRUCTURE ONLY (never linked or called by the product, core_b200/).
 *
 * Plain-C CPU restatement of the reference's arithmetic for the ray/integrator hot path, used as the
 * portable checker next to the compiled reference itself (oracle/_ref, built by oracle/Makefile):
 *   - triangle_t::intersect                 include/yafraycore/triangle_inline.h:27-64
 *   - closest / any-hit acceptance rules    src/yafraycore/kdtree.cc:772,791,916 + src/yafraycore/scene.cc:852-902
 *     evaluated by BRUTE FORCE over all triangles (the kd-tree is an accelerator, not part of the result)
 *   - QMC: Halton, RI_vdC, RI_S, RI_LP, fnv_32a_buf, scrHalton   include/utilities/mcqmc.h:29-168,
 *     include/yafraycore/scr_halton.h:27-69 (Faure permutations generated, see tools/gen_faure.py)
 *   - fast math fExp2/fLog2/fPow/fSin/fCos  include/utilities/mathOptimizations.h:100-280
 * Parity pinning: the reference ships no golden vectors; this file is pinned by
 * tests/test_oracle.py against outputs of the compiled reference (yref_* hooks in oracle/ref_driver.cc)
 * and against tests/golden/ fixtures generated from it by tests/golden/make_golden.py.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>
#include <stddef.h>

/* ------------------------------------------------------------------ triangles */
static int tri_intersect(const float *a, const float *b, const float *c, const float *from, const float *dir,
                         float *t, float *u, float *v)
{
    float e1[3], e2[3], p[3], tv[3], q[3];
    int i;
    for (i = 0; i < 3; ++i) { e1[i] = b[i] - a[i]; e2[i] = c[i] - a[i]; }
    p[0] = dir[1] * e2[2] - dir[2] * e2[1]; p[1] = dir[2] * e2[0] - dir[0] * e2[2]; p[2] = dir[0] * e2[1] - dir[1] * e2[0];
    float det = e1[0] * p[0] + e1[1] * p[1] + e1[2] * p[2];
    if (det == 0.0f) return 0;
    float inv_det = (float)(1.0 / det);
    for (i = 0; i < 3; ++i) tv[i] = from[i] - a[i];
    *u = (tv[0] * p[0] + tv[1] * p[1] + tv[2] * p[2]) * inv_det;
    if (*u < 0.0f || *u > 1.0f) return 0;
    q[0] = tv[1] * e1[2] - tv[2] * e1[1]; q[1] = tv[2] * e1[0] - tv[0] * e1[2]; q[2] = tv[0] * e1[1] - tv[1] * e1[0];
    *v = (dir[0] * q[0] + dir[1] * q[1] + dir[2] * q[2]) * inv_det;
    if (*v < 0.0f || (*u + *v) > 1.0f) return 0;
    *t = (e2[0] * q[0] + e2[1] * q[1] + e2[2] * q[2]) * inv_det;
    return 1;
}

typedef struct { int32_t prim; float t, b1, b2; } yo_hit;

/* tris: n x 9 floats (a, b, c) in the global primitive order; rays: m x 8 floats {from, tmin, dir, tmax}.
 * Closest hit (scene_t::intersect): t < Z && t >= tmin, Z starts at tmax (inf if < 0); on exact t ties the
 * LOWEST primitive index is reported together with the number of tied primitives in ties[i] (may be NULL). */
void yo_closest(const float *tris, int n, const float *rays, int m, yo_hit *out, int32_t *ties)
{
    int i, k;
    for (i = 0; i < m; ++i) {
        const float *r = rays + 8 * i;
        float Z = r[7] < 0 ? INFINITY : r[7];
        yo_hit h; h.prim = -1; h.t = 0; h.b1 = 0; h.b2 = 0;
        int nt = 0;
        for (k = 0; k < n; ++k) {
            float t, u, v;
            if (!tri_intersect(tris + 9 * k, tris + 9 * k + 3, tris + 9 * k + 6, r, r + 4, &t, &u, &v)) continue;
            if (t >= r[3] && (t < Z || (h.prim >= 0 && t == h.t))) {
                if (h.prim >= 0 && t == h.t) { nt++; continue; }
                Z = t; h.prim = k; h.t = t; h.b1 = u; h.b2 = v; nt = 1;
            }
        }
        out[i] = h;
        if (ties) ties[i] = nt;
    }
}
/* scene_t::isShadowed(state, ray): origin advanced by dir*tmin, dist = tmax - 2 tmin, accept 0 <= t < dist */
void yo_shadow(const float *tris, int n, const float *rays, int m, int32_t *out)
{
    int i, k;
    for (i = 0; i < m; ++i) {
        const float *r = rays + 8 * i;
        float from[3] = { r[0] + r[4] * r[3], r[1] + r[5] * r[3], r[2] + r[6] * r[3] };
        float dist = r[7] < 0 ? INFINITY : r[7] - 2 * r[3];
        int s = 0;
        for (k = 0; k < n && !s; ++k) {
            float t, u, v;
            if (tri_intersect(tris + 9 * k, tris + 9 * k + 3, tris + 9 * k + 6, from, r + 4, &t, &u, &v) && t < dist && t >= 0.f) s = 1;
        }
        out[i] = s;
    }
}

/* ------------------------------------------------------------------ QMC */
double yo_halton(unsigned base, unsigned start, int nth)   /* setStart(start), then nth x getNext(); returns the last */
{
    double inv = 1.0 / (double)base, factor = inv, value = 0.0;
    unsigned i = start;
    float out = 0.f;
    int k;
    while (i > 0) { value += (double)(i % base) * factor; i /= base; factor *= inv; }
    for (k = 0; k < nth; ++k) {
        double r = 0.9999999999 - value;
        if (inv < r) value += inv;
        else { double hh = 0, h = inv; while (h >= r) { hh = h; h *= inv; } value += hh + h - 1.0; }
        out = (float)value; if (out > 1.f) out = 1.f; if (out < 0.f) out = 0.f;
    }
    return out;
}
static float clamp01(float v) { return v < 0.f ? 0.f : (v > 1.f ? 1.f : v); }
float yo_ri_vdc(unsigned bits, unsigned r)
{
    bits = (bits << 16) | (bits >> 16);
    bits = ((bits & 0x00ff00ff) << 8) | ((bits & 0xff00ff00) >> 8);
    bits = ((bits & 0x0f0f0f0f) << 4) | ((bits & 0xf0f0f0f0) >> 4);
    bits = ((bits & 0x33333333) << 2) | ((bits & 0xcccccccc) >> 2);
    bits = ((bits & 0x55555555) << 1) | ((bits & 0xaaaaaaaa) >> 1);
    return clamp01((float)((double)(bits ^ r) * 0.00000000023283064365386962890625));
}
float yo_ri_s(unsigned i, unsigned r)
{
    unsigned v;
    for (v = 1u << 31; i; i >>= 1, v ^= v >> 1) if (i & 1) r ^= v;
    return clamp01((float)((double)r * 0.00000000023283064365386962890625));
}
float yo_ri_lp(unsigned i, unsigned r)
{
    unsigned v;
    for (v = 1u << 31; i; i >>= 1, v |= v >> 1) if (i & 1) r ^= v;
    return clamp01((float)((double)r * 0.00000000023283064365386962890625));
}
unsigned yo_fnv(unsigned value)
{
    unsigned hash = 0x811c9dc5u; int i;
    for (i = 0; i < 4; ++i) { hash ^= (value >> (8 * i)) & 0xffu; hash *= 0x01000193u; }
    return hash;
}
/* Faure permutation of base b into out[b] (H. Faure 1992; tools/gen_faure.py holds the same recursion) */
static void faure_perm(int b, int *out)
{
    if (b == 1) { out[0] = 0; return; }
    if (b == 2) { out[0] = 0; out[1] = 1; return; }
    if (b % 2 == 0) {
        int h = b / 2, i;
        faure_perm(h, out);
        for (i = 0; i < h; ++i) { out[h + i] = 2 * out[i] + 1; }
        for (i = 0; i < h; ++i) out[i] = 2 * out[i];
    } else {
        int c = (b - 1) / 2, i, tmp[256];
        faure_perm(b - 1, tmp);
        for (i = 0; i < b - 1; ++i) if (tmp[i] >= c) tmp[i]++;
        for (i = 0; i < c; ++i) out[i] = tmp[i];
        out[c] = c;
        for (i = c; i < b - 1; ++i) out[i + 1] = tmp[i];
    }
}
static const int yo_prims[50] = {1, 2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37, 41, 43, 47, 53, 59, 61, 67, 71, 73, 79, 83, 89, 97,
    101, 103, 107, 109, 113, 127, 131, 137, 139, 149, 151, 157, 163, 167, 173, 179, 181, 191, 193, 197, 199, 211, 223, 227};
int yo_faure(int dim, int *out) { faure_perm(yo_prims[dim], out); return yo_prims[dim]; }
double yo_scr_halton(int dim, unsigned n)
{
    int sigma[256];
    double value = 0.0, f, factor, dn = (double)n;
    unsigned base = (unsigned)yo_prims[dim];
    faure_perm((int)base, sigma);
    f = factor = floor(1e9 / (double)base + 0.5) / 1e9;       /* invPrims: 1/p rounded to 9 decimals (scr_halton.h:33-45) */
    while (n > 0) { value += (double)sigma[n % base] * factor; dn *= f; n = (unsigned)dn; factor *= f; }
    return value < 1.0e-36 ? 1.0e-36 : (value > 1.0 ? 1.0 : value);
}

/* ------------------------------------------------------------------ fast math */
static float i2f(int32_t i) { float f; memcpy(&f, &i, 4); return f; }
static int32_t f2i(float f) { int32_t i; memcpy(&i, &f, 4); return i; }
float yo_fexp2(float x)
{
    if (x > 129.00000f) x = 129.00000f;
    if (x < -126.99999f) x = -126.99999f;
    int32_t ip = (int32_t)(x - 0.5f);
    float fp = x - (float)ip;
    float ex = i2f((ip + 127) << 23);
    return ex * (float)(fp * (fp * (fp * (fp * (fp * 1.8775767e-3f + 8.9893397e-3f) + 5.5826318e-2f) + 2.4015361e-1f) + 6.9315308e-1f) + 9.9999994e-1f);
}
float yo_flog2(float x)
{
    int32_t ii = f2i(x);
    float e = (float)(((ii & 0x7F800000) >> 23) - 127);
    float m = i2f((ii & 0x7FFFFF) | f2i(1.0f));
    return (float)(m * (m * (m * (m * (m * -3.4436006e-2f + 3.1821337e-1f) + -1.2315303f) + 2.5988452) + -3.3241990f) + 3.1157899f) * (m - 1.0f) + e;
}
float yo_fpow(float a, float b) { return yo_fexp2(yo_flog2(a) * b); }
float yo_fsin(float x)
{
    if (x > 6.28318530717958647692 || x < -6.28318530717958647692) x -= ((int)(x * (float)0.15915494309189533577)) * (float)6.28318530717958647692;
    if (x < -3.14159265358979323846) x += (float)6.28318530717958647692;
    else if (x > 3.14159265358979323846) x -= (float)6.28318530717958647692;
    x = ((float)1.27323954473516268615 * x) - ((float)0.40528473456935108578 * x * fabsf(x));
    float result = 0.225f * (x * fabsf(x) - x) + x;
    if (result <= -1.0f) return -1.0f;
    if (result >= 1.0f) return 1.0f;
    return result;
}
float yo_fcos(float x) { return yo_fsin(x + (float)1.57079632679489661923); }
