// rt_hd.h — host/device helpers: float3 algebra and the *exact* binary32 operations the parity contract needs.
//
// Everything that feeds a hit/miss decision, a hit distance or a colour is evaluated in strict IEEE binary32 in the
// reference's operation order (Unity.Mathematics 1.2.6 semantics).  The CUDA translation units are
// compiled with -fmad=false so `a*b+c` is never contracted; where an FMA is wanted (conservative BVH slab tests only)
// it is requested explicitly through rt_fma().
#pragma once
#include <stdint.h>
#include <math.h>
#include <float.h>
#include <string.h>

#if defined(__CUDACC__)
#define RT_HD __host__ __device__ __forceinline__
#else
#define RT_HD inline
#endif

struct f3 { float x, y, z; };
RT_HD f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
RT_HD f3 bc3(float s) { return mk3(s, s, s); }
RT_HD f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
RT_HD f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
RT_HD f3 operator*(f3 a, f3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
RT_HD f3 operator*(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
RT_HD f3 operator*(float s, f3 a) { return mk3(s * a.x, s * a.y, s * a.z); }

RT_HD float rt_sqrt(float x) {
#if defined(__CUDA_ARCH__)
    return __fsqrt_rn(x);          // correctly rounded == (float)System.Math.Sqrt((double)x)
#else
    return sqrtf(x);
#endif
}
RT_HD float rt_div(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fdiv_rn(a, b);
#else
    return a / b;
#endif
}
RT_HD float rt_fma(float a, float b, float c) {
#if defined(__CUDA_ARCH__)
    return __fmaf_rn(a, b, c);
#else
    return fmaf(a, b, c);
#endif
}
RT_HD f3 operator/(f3 a, float s) { return mk3(rt_div(a.x, s), rt_div(a.y, s), rt_div(a.z, s)); }

// ---- Unity.Mathematics semantics -----------------------------------------------------------------------------
RT_HD float um_min(float x, float y) { return (y != y || x < y) ? x : y; }   // math.min: first arg when the second is NaN
RT_HD float um_max(float x, float y) { return (y != y || x > y) ? x : y; }
RT_HD float um_dot(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
RT_HD f3 um_cross(f3 a, f3 b) {                                              // (x*y.yzx - x.yzx*y).yzx
    float rx = a.x * b.y - a.y * b.x;
    float ry = a.y * b.z - a.z * b.y;
    float rz = a.z * b.x - a.x * b.z;
    return mk3(ry, rz, rx);
}
RT_HD float um_rsqrt(float x) { return rt_div(1.0f, rt_sqrt(x)); }
RT_HD f3 um_normalize(f3 a) { return um_rsqrt(um_dot(a, a)) * a; }
RT_HD float um_length(f3 a) { return rt_sqrt(um_dot(a, a)); }
// (float)System.Math.Pow((double)x, (double)y) (RayTracingSetup.cs:399).  CUDA's double pow() is ~170 instructions per lane (ncu:
// 2.6 % of the fused shadow kernel).  PhongExponent is a small whole number in every shipped material, so on the device x^n
// is evaluated by squaring in double (<= 12 DMUL, relative error < 64 double ulps for n <= 64) and rounded to float ONLY when
// that rounding is certain: the double lies more than 256 ulps from every float rounding boundary and in float's normal range
// (or is so small that the float is 0).  In that case the result IS the correctly rounded true power, i.e. what any pow()
// with < 256 ulps of error returns; otherwise (1e-6 of the calls, and every non-integer or large exponent) the library pow runs.
RT_HD float um_pow(float x, float y) {
#if defined(__CUDA_ARCH__)
    const int n = (int)y;
    if ((float)n == y && n >= 0 && n <= 64) {
        double b = (double)x, r = 1.0;
        for (int e = n; e; ) { if (e & 1) r *= b; e >>= 1; if (e) b *= b; }
        const long long bits = __double_as_longlong(r);
        const int low = (int)((unsigned)bits & 0x1fffffffu);                // the 29 mantissa bits a float does not have
        const int off = low - 0x10000000;                                  // distance from the round-to-nearest midpoint
        const double a = fabs(r);
        const bool mid_ok = off > 256 || off < -256;
        if ((a >= 1.1754943508222875e-38 && a < 3.0e38 && mid_ok) || a < 3.0e-46) return (float)r;
    }
#endif
    return (float)pow((double)x, (double)y);
}

#define RT_EPS 0.00001f             /* RMath.cs:9 */
#define RT_SHADOW_EPS 0.0001f       /* RayTracingSetup.cs:42 */
// Largest float x with  degrees((float)acos((double)x)) > 90f  (RayTracingSetup.cs:384-391); the predicate is monotone,
// so `angle > 90` == `x <= RT_SPEC_CUTOFF` bit for bit.  Pinned by tests/test_oracle.py against the oracle's acos.
#define RT_SPEC_CUTOFF (-0x1.bbbd3p-24f)

// float <-> int bit casts
RT_HD float i2f(int32_t i) {
#if defined(__CUDA_ARCH__)
    return __int_as_float(i);
#else
    float f; memcpy(&f, &i, 4); return f;
#endif
}
RT_HD int32_t f2i(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_int(f);
#else
    int32_t i; memcpy(&i, &f, 4); return i;
#endif
}

#if !defined(__CUDACC__)
struct float4 { float x, y, z, w; };
static inline float4 make_float4(float x, float y, float z, float w) { float4 r{x, y, z, w}; return r; }
#endif
