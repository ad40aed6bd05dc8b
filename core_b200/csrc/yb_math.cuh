// yb_math.cuh — device math library of the B200 ray/integrator hot path.
//
// Restates, operation for operation, the inline math the reference's hot path is built on:
//   vectors / colours       include/core_api/vector3d.h:180-334, src/yafraycore/vector3d.cc:86-182
//   fast-math approximations include/utilities/mathOptimizations.h:100-296 (FAST_MATH + FAST_TRIG on)
//   QMC + hash              include/utilities/mcqmc.h:29-168, include/yafraycore/scr_halton.h:27-69
//   sampling helpers        include/utilities/sample_utils.h:41-88,180-184
//   Round2Int/Floor2Int     include/utilities/math_utils.h:60-86 (Y_FAST_INT == 0 branch)
// Everything is YB_HD (host + device) so that the host-side factories (light/camera constants) use the
// very same approximations as the kernels, and so that tests/emu can single-step the kernels on a CPU.
#pragma once
#include <stdint.h>
#include <math.h>
#include <string.h>

#if defined(__CUDACC__) && !defined(YB_HOSTEMU)
#define YB_HD __host__ __device__ __forceinline__
#define YB_D __device__ __forceinline__
// large device functions shared by several call sites of the shade kernel.  Out-of-line (-DYB_OUTLINE) shrinks k_shade from
// 70 k to 44 k SASS instructions but was measured 5x SLOWER on B200 (shade 108 -> 590 ms per frame: the call ABI pins the
// surface point, material data and sample state in local memory), so everything stays inlined.
#ifdef YB_OUTLINE
#define YB_D_BIG static __device__ __noinline__
#else
#define YB_D_BIG __device__ __forceinline__
#endif
#else
#define YB_HD inline
#define YB_D inline
#define YB_D_BIG inline
#endif

namespace yb {

// ------------------------------------------------------------------ bit casts
YB_HD float i2f(int32_t i) { float f; memcpy(&f, &i, 4); return f; }
YB_HD int32_t f2i(float f) { int32_t i; memcpy(&i, &f, 4); return i; }
YB_HD float fminf_(float a, float b) { return a < b ? a : b; }
YB_HD float fmaxf_(float a, float b) { return a > b ? a : b; }

// ------------------------------------------------------------------ fast math (mathOptimizations.h)
#define YB_M_2PI 6.28318530717958647692
#define YB_M_1_2PI 0.15915494309189533577
#define YB_M_4_PI 1.27323954473516268615
#define YB_M_4_PI2 0.40528473456935108578
#define YB_M_PI 3.14159265358979323846
#define YB_M_PI_2 1.57079632679489661923
#define YB_M_PI_4 0.78539816339744830962
#define YB_M_1_PI 0.31830988618379067154
#define YB_M_LN2 0.69314718055994530942
#define YB_M_LOG2E 1.4426950408889634074

// mathOptimizations.h:87 POLYEXP / :88 POLYLOG (the 2.5988452 literal is a double there too)
YB_HD float polyexp(float x)
{
    return (float)(x * (x * (x * (x * (x * 1.8775767e-3f + 8.9893397e-3f) + 5.5826318e-2f) + 2.4015361e-1f) + 6.9315308e-1f) + 9.9999994e-1f);
}
YB_HD float polylog(float x)
{
    return (float)(x * (x * (x * (x * (x * -3.4436006e-2f + 3.1821337e-1f) + -1.2315303f) + 2.5988452) + -3.3241990f) + 3.1157899f);
}
// mathOptimizations.h:105-119
YB_HD float fExp2(float x)
{
    x = fminf_(x, 129.00000f);
    x = fmaxf_(x, -126.99999f);
    int32_t ip = (int32_t)(x - 0.5f);
    float fp = x - (float)ip;
    float ex = i2f((ip + 127) << 23);
    return ex * polyexp(fp);
}
// mathOptimizations.h:121-131
YB_HD float fLog2(float x)
{
    int32_t ii = f2i(x);
    float e = (float)(((ii & 0x7F800000) >> 23) - 127);
    float m = i2f((ii & 0x7FFFFF) | f2i(1.0f));
    return polylog(m) * (m - 1.0f) + e;
}
YB_HD float fPow(float a, float b) { return fExp2(fLog2(a) * b); }        // :200-207
YB_HD float fLog(float a) { return fLog2(a) * (float)YB_M_LN2; }          // :209-216
YB_HD float fExp(float a) { return fExp2((float)YB_M_LOG2E * a); }        // :218-225
// :227-237 — x87 fsqrt of a float then rounded to float == correctly rounded float sqrt
YB_HD float fSqrt(float a)
{
#if defined(__CUDA_ARCH__)
    return __fsqrt_rn(a);
#else
    return sqrtf(a);
#endif
}
// :249-271
YB_HD float fSin(float x)
{
    if (x > YB_M_2PI || x < -YB_M_2PI) x -= ((int)(x * (float)YB_M_1_2PI)) * (float)YB_M_2PI;
    if (x < -YB_M_PI) x += (float)YB_M_2PI;
    else if (x > YB_M_PI) x -= (float)YB_M_2PI;
    x = ((float)YB_M_4_PI * x) - ((float)YB_M_4_PI2 * x * fabsf(x));
    float result = 0.225f * (x * fabsf(x) - x) + x;
    if (result <= -1.0f) return -1.0f;
    else if (result >= 1.0f) return 1.0f;
    return result;
}
YB_HD float fCos(float x) { return fSin(x + (float)YB_M_PI_2); }          // :273-280
YB_HD float fAcos(float x)                                                // :282-288
{
    if (x <= -1.0f) return (float)YB_M_PI;
    else if (x >= 1.0f) return 0.0f;
    return acosf(x);
}
YB_HD float fAsin(float x)                                                // :290-296
{
    if (x <= -1.0f) return (float)-YB_M_PI_2;
    else if (x >= 1.0f) return (float)YB_M_PI_2;
    return asinf(x);
}
// math_utils.h:60-86, non-x86-32 branch
YB_HD int Round2Int(double v) { return (int)(v + (0.5 - 1.4e-11)); }
YB_HD int Floor2Int(double v) { return (int)floor(v); }

// ------------------------------------------------------------------ vectors (vector3d.h)
struct v3 {
    float x, y, z;
    YB_HD v3() {}
    YB_HD v3(float a, float b, float c) : x(a), y(b), z(c) {}
    YB_HD explicit v3(float a) : x(a), y(a), z(a) {}
    YB_HD float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
YB_HD v3 operator+(v3 a, v3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
YB_HD v3 operator-(v3 a, v3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
YB_HD v3 operator-(v3 a) { return v3(-a.x, -a.y, -a.z); }
YB_HD v3 operator*(v3 a, float f) { return v3(a.x * f, a.y * f, a.z * f); }
YB_HD v3 operator*(float f, v3 a) { return v3(a.x * f, a.y * f, a.z * f); }
YB_HD v3 mul(v3 a, v3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
YB_HD float dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
YB_HD v3 cross(v3 a, v3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
YB_HD v3 &operator+=(v3 &a, v3 b) { a.x += b.x; a.y += b.y; a.z += b.z; return a; }
YB_HD v3 &operator*=(v3 &a, float f) { a.x *= f; a.y *= f; a.z *= f; return a; }
YB_HD v3 &operator*=(v3 &a, v3 b) { a.x *= b.x; a.y *= b.y; a.z *= b.z; return a; }
YB_HD float length(v3 a) { return fSqrt(a.x * a.x + a.y * a.y + a.z * a.z); }     // vector3d.h:245-248
YB_HD v3 normalize(v3 a)                                                            // vector3d.h:250-261
{
    float len = a.x * a.x + a.y * a.y + a.z * a.z;
    if (len != 0) { len = 1.0f / fSqrt(len); a.x *= len; a.y *= len; a.z *= len; }
    return a;
}
YB_HD float normLen(v3 &a)                                                          // vector3d.h:67-76
{
    float vl = a.x * a.x + a.y * a.y + a.z * a.z;
    if (vl != 0.0f) { vl = fSqrt(vl); const float d = 1.0f / vl; a.x *= d; a.y *= d; a.z *= d; }
    return vl;
}
YB_HD bool isBlack(v3 c) { return c.x == 0 && c.y == 0 && c.z == 0; }
typedef v3 col3;   // color_t

// vector3d.h:285-292  vector3d_t::reflect
YB_HD v3 reflect_about(v3 v, v3 n)
{
    const float vn = 2.0f * (v.x * n.x + v.y * n.y + v.z * n.z);
    return v3(vn * n.x - v.x, vn * n.y - v.y, vn * n.z - v.z);
}
// vector3d.h:294-299  reflect_dir
YB_HD v3 reflect_dir(v3 n, v3 v)
{
    const float vn = dot(v, n);
    if (vn < 0) return -v;
    return 2 * vn * n - v;
}
// vector3d.h:316-334
YB_HD void createCS(v3 N, v3 &u, v3 &v)
{
    if ((N.x == 0) && (N.y == 0)) {
        if (N.z < 0) u = v3(-1, 0, 0); else u = v3(1, 0, 0);
        v = v3(0, 1, 0);
    } else {
        const float d = 1.0f / fSqrt(N.y * N.y + N.x * N.x);
        u = v3(N.y * d, -N.x * d, 0);
        v = cross(N, u);
    }
}
// vector3d.cc:86-108
YB_HD bool refract(v3 n, v3 wi, v3 &wo, float IOR)
{
    v3 N = n, I = -wi;
    float eta = IOR;
    float cos_v_n = dot(wi, n);
    if (cos_v_n < 0) { N = -n; cos_v_n = -cos_v_n; }
    else eta = 1.0f / IOR;
    float k = 1 - eta * eta * (1 - cos_v_n * cos_v_n);
    if (k <= 0.f) return false;
    wo = normalize(eta * I + (eta * cos_v_n - fSqrt(k)) * N);
    return true;
}
// vector3d.cc:110-141
YB_HD void fresnel(v3 I, v3 n, float IOR, float &Kr, float &Kt)
{
    float eta = IOR;
    v3 N = (dot(I, n) < 0) ? -n : n;
    float c = dot(I, N);
    float g = eta * eta + c * c - 1;
    if (g <= 0) g = 0; else g = fSqrt(g);
    float aux = c * (g + c);
    Kr = ((0.5f * (g - c) * (g - c)) / ((g + c) * (g + c))) * (1 + ((aux - 1) * (aux - 1)) / ((aux + 1) * (aux + 1)));
    if (Kr < 1.0f) Kt = 1 - Kr; else Kt = 0;
}
// vector3d.cc:155-182
YB_HD void ShirleyDisk(float r1, float r2, float &u, float &v)
{
    float phi = 0, r = 0, a = 2 * r1 - 1, b = 2 * r2 - 1;
    if (a > -b) {
        if (a > b) { r = a; phi = (float)YB_M_PI_4 * (b / a); }
        else { r = b; phi = (float)YB_M_PI_4 * (2 - a / b); }
    } else {
        if (a < b) { r = -a; phi = (float)YB_M_PI_4 * (4 + b / a); }
        else { r = -b; if (b != 0) phi = (float)YB_M_PI_4 * (6 - a / b); else phi = 0; }
    }
    u = r * fCos(phi);
    v = r * fSin(phi);
}

// ------------------------------------------------------------------ sampling helpers (sample_utils.h)
YB_HD v3 SampleCosHemisphere(v3 N, v3 Ru, v3 Rv, float s1, float s2)        // :41-50
{
    if (s1 >= 1.0f) return N;
    float z1 = s1;
    float z2 = s2 * (float)YB_M_2PI;
    return (Ru * fCos(z2) + Rv * fSin(z2)) * fSqrt(1.0f - z1) + N * fSqrt(z1);
}
YB_HD v3 sampleCone(v3 D, v3 U, v3 V, float maxCosAng, float s1, float s2)  // :78-84
{
    float cosAng = 1.f - (1.f - maxCosAng) * s2;
    float sinAng = fSqrt(1.f - cosAng * cosAng);
    float t1 = (float)YB_M_2PI * s1;
    return (U * fCos(t1) + V * fSin(t1)) * sinAng + D * cosAng;
}
YB_HD float addMod1(float a, float b) { float s = a + b; return s > 1 ? s - 1.f : s; }   // :180-184

// ------------------------------------------------------------------ QMC (mcqmc.h)
#define YB_MULT_RATIO (0.00000000023283064365386962890625)
YB_HD float clamp01f(float v) { return fmaxf_(0.f, fminf_(1.f, v)); }
YB_HD float RI_vdC(uint32_t bits, uint32_t r = 0)                              // :100-108
{
    bits = (bits << 16) | (bits >> 16);
    bits = ((bits & 0x00ff00ff) << 8) | ((bits & 0xff00ff00) >> 8);
    bits = ((bits & 0x0f0f0f0f) << 4) | ((bits & 0xf0f0f0f0) >> 4);
    bits = ((bits & 0x33333333) << 2) | ((bits & 0xcccccccc) >> 2);
    bits = ((bits & 0x55555555) << 1) | ((bits & 0xaaaaaaaa) >> 1);
    return clamp01f((float)((double)(bits ^ r) * YB_MULT_RATIO));
}
YB_HD float RI_S(uint32_t i, uint32_t r = 0)                                   // :110-115
{
    for (uint32_t v = 1u << 31; i; i >>= 1, v ^= v >> 1) if (i & 1) r ^= v;
    return clamp01f((float)((double)r * YB_MULT_RATIO));
}
YB_HD float RI_LP(uint32_t i, uint32_t r = 0)                                  // :117-122
{
    for (uint32_t v = 1u << 31; i; i >>= 1, v |= v >> 1) if (i & 1) r ^= v;
    return clamp01f((float)((double)r * YB_MULT_RATIO));
}
YB_HD uint32_t fnv_32a_buf(uint32_t value)                                     // :155-168 (little endian bytes)
{
    uint32_t hash = 0x811c9dc5u;
    for (int i = 0; i < 4; i++) { hash ^= (value >> (8 * i)) & 0xffu; hash *= 0x01000193u; }
    return hash;
}
// mcqmc.h:29-94 — incremental Halton; FP64 state, getNext() advances THEN returns.
// The base is a template parameter so the digit loop divides by a constant (mul-hi instead of a 30-instruction
// runtime division); the FP64 operation sequence — and so every bit of the stream — is the reference's.
// factor of digit k in HaltonT::setStart: invBase multiplied k times by invBase, in the same FP64 operations as the loop (a
// compile-time constant wherever k is one)
constexpr double halton_digit_factor(uint32_t base, int k)
{
    const double inv = 1.0 / (double)base;
    double f = inv;
    for (int j = 0; j < k; ++j) f *= inv;
    return f;
}
template <uint32_t BASE> struct HaltonT {
    double mValue;
    YB_HD HaltonT() : mValue(0) {}
    YB_HD static double invBase() { return 1.0 / (double)BASE; }
    YB_HD void setStart(uint32_t i)
    {
        if (BASE == 2u) {
            // all partial sums are exact in binary: the loop result is the bit-reversed index scaled by 2^-32
            uint32_t r = i;
            r = ((r >> 1) & 0x55555555u) | ((r & 0x55555555u) << 1);
            r = ((r >> 2) & 0x33333333u) | ((r & 0x33333333u) << 2);
            r = ((r >> 4) & 0x0f0f0f0fu) | ((r & 0x0f0f0f0fu) << 4);
            r = ((r >> 8) & 0x00ff00ffu) | ((r & 0x00ff00ffu) << 8);
            r = (r >> 16) | (r << 16);
            mValue = (double)r * 2.3283064365386963e-10;   // 2^-32
            return;
        }
        if (BASE == 3u) {
            // same partial sums as the loop below, digit by digit: the factors are constants, digit * factor is exact (0, f, 2 f),
            // adding a zero digit changes nothing - what is left per digit is one mul-hi division and one FP64 add
            mValue = 0.0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
            for (int k = 0; k < 21; ++k) {                     // 3^21 > 2^32
                if (i == 0u) break;
                const uint32_t q = i / 3u, d = i - 3u * q;
                const double f = halton_digit_factor(3u, k);
                if (d) mValue += (d == 1u ? f : 2.0 * f);
                i = q;
            }
            return;
        }
        const double mInvBase = invBase();
        double factor = mInvBase;
        mValue = 0.0;
        while (i > 0) { mValue += (double)(i % BASE) * factor; i /= BASE; factor *= mInvBase; }
    }
    YB_HD float getNext()
    {
        const double mInvBase = invBase();
        double r = 0.9999999999 - mValue;
        if (mInvBase < r) mValue += mInvBase;
        else {
            double hh = 0, h = mInvBase;
            while (h >= r) { hh = h; h *= mInvBase; }
            mValue += hh + h - 1.0;
        }
        return fmaxf_(0.f, fminf_(1.f, (float)mValue));
    }
};
// Base 2: every quantity of the incremental scheme is a dyadic rational with <= 33 bits, so the FP64 arithmetic of
// setStart / getNext is exact and the stream is simply the bit-reversed counter; 0.9999999999 only decides comparisons
// that are >= 1e-10 away from a tie.  The counter wrap at 2^32 (where the reference's value is NOT Phi2(0)) falls back
// to the generic incremental code.
struct Halton2 {
    uint32_t idx; bool slow; HaltonT<2> gen;
    YB_HD Halton2() : idx(0), slow(false) {}
    YB_HD static uint32_t brev(uint32_t r)
    {
#if defined(__CUDA_ARCH__)
        return __brev(r);
#else
        r = ((r >> 1) & 0x55555555u) | ((r & 0x55555555u) << 1);
        r = ((r >> 2) & 0x33333333u) | ((r & 0x33333333u) << 2);
        r = ((r >> 4) & 0x0f0f0f0fu) | ((r & 0x0f0f0f0fu) << 4);
        r = ((r >> 8) & 0x00ff00ffu) | ((r & 0x00ff00ffu) << 8);
        return (r >> 16) | (r << 16);
#endif
    }
    YB_HD void setStart(uint32_t i) { idx = i; slow = false; }
    YB_HD float getNext()
    {
        if (!slow && idx != 0xffffffffu) {
            ++idx;
            return fmaxf_(0.f, fminf_(1.f, (float)((double)brev(idx) * 2.3283064365386963e-10)));
        }
        if (!slow) { gen.setStart(idx); slow = true; }
        return gen.getNext();
    }
};
typedef HaltonT<3> Halton3;
typedef HaltonT<5> Halton5;

}  // namespace yb
