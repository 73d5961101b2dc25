// vecmath.cuh -- small FP32 vector vocabulary for the device code.
//
// Every helper spells out its evaluation order: sums run left to right and nothing is
// re-associated, so the GPRT_MATH_PARITY build (compiled with --fmad=false, IEEE div/sqrt)
// produces the same bits as a left-to-right scalar evaluation.  The GPRT_MATH_FAST build compiles
// the same source with --use_fast_math (FMA contraction, approximate div/sqrt/transcendentals).
#pragma once

#ifndef GPRT_HD
#if defined(__CUDACC__)
#define GPRT_HD __device__ __forceinline__
#else
#define GPRT_HD inline
#endif
#endif

#include <math.h>
#include <stdint.h>

namespace gprt {

struct V2 { float x, y; };
struct V3 { float x, y, z; };
struct V4 { float x, y, z, w; };

GPRT_HD V2 v2(float x, float y) { V2 r; r.x = x; r.y = y; return r; }
GPRT_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
GPRT_HD V3 v3s(float s) { return v3(s, s, s); }
GPRT_HD V4 v4(float x, float y, float z, float w) { V4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
GPRT_HD V4 v4s(float s) { return v4(s, s, s, s); }

GPRT_HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
GPRT_HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
GPRT_HD V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
GPRT_HD V3 operator/(V3 a, V3 b) { return v3(a.x / b.x, a.y / b.y, a.z / b.z); }
GPRT_HD V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
GPRT_HD V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
GPRT_HD V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
GPRT_HD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
GPRT_HD V2 operator+(V2 a, V2 b) { return v2(a.x + b.x, a.y + b.y); }
GPRT_HD V2 operator-(V2 a, V2 b) { return v2(a.x - b.x, a.y - b.y); }
GPRT_HD V2 operator*(V2 a, V2 b) { return v2(a.x * b.x, a.y * b.y); }
GPRT_HD V2 operator*(float s, V2 a) { return v2(s * a.x, s * a.y); }
GPRT_HD V4 operator+(V4 a, V4 b) { return v4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
GPRT_HD V4 operator-(V4 a, V4 b) { return v4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
GPRT_HD V4 operator*(V4 a, V4 b) { return v4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
GPRT_HD V4 operator*(float s, V4 a) { return v4(s * a.x, s * a.y, s * a.z, s * a.w); }

GPRT_HD float dot(V2 a, V2 b) { return a.x * b.x + a.y * b.y; }
GPRT_HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
GPRT_HD float length(V2 a) { return sqrtf(dot(a, a)); }
GPRT_HD float length(V3 a) { return sqrtf(dot(a, a)); }
// HLSL normalize lowers to v * rsqrt(dot(v, v))
GPRT_HD float rsqrt_ieee(float x) {
#if defined(GPRT_FAST_MATH) && defined(__CUDA_ARCH__)
    return rsqrtf(x);
#else
    return 1.0f / sqrtf(x);
#endif
}
GPRT_HD V3 normalize(V3 a) { const float r = rsqrt_ieee(dot(a, a)); return v3(a.x * r, a.y * r, a.z * r); }
GPRT_HD V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
GPRT_HD V3 reflect(V3 i, V3 n) { const float s = 2.0f * dot(i, n); return v3(i.x - s * n.x, i.y - s * n.y, i.z - s * n.z); }
GPRT_HD V3 vabs(V3 a) { return v3(fabsf(a.x), fabsf(a.y), fabsf(a.z)); }
GPRT_HD V3 vmax(V3 a, float s) { return v3(fmaxf(a.x, s), fmaxf(a.y, s), fmaxf(a.z, s)); }
GPRT_HD V2 vmax(V2 a, float s) { return v2(fmaxf(a.x, s), fmaxf(a.y, s)); }
GPRT_HD V2 vabs(V2 a) { return v2(fabsf(a.x), fabsf(a.y)); }

GPRT_HD float saturate(float x) { return fminf(fmaxf(x, 0.0f), 1.0f); }
GPRT_HD float lerp(float a, float b, float s) { return a + s * (b - a); }
GPRT_HD V3 lerp(V3 a, V3 b, float s) { return v3(a.x + s * (b.x - a.x), a.y + s * (b.y - a.y), a.z + s * (b.z - a.z)); }
GPRT_HD V4 lerp(V4 a, V4 b, float s) { return v4(a.x + s * (b.x - a.x), a.y + s * (b.y - a.y), a.z + s * (b.z - a.z), a.w + s * (b.w - a.w)); }
GPRT_HD float frac(float x) { return x - floorf(x); }
// HLSL pow(x, y) lowers to exp2(log2(x) * y); exp(x) to exp2(x * log2 e)
GPRT_HD float hlsl_pow(float x, float y) { return exp2f(log2f(x) * y); }
GPRT_HD float hlsl_exp(float x) { return exp2f(x * 1.44269504088896340736f); }
// HLSL fmod(x, y) lowers to q = x / y; r = frac(|q|); (q >= 0 ? r : -r) * y
GPRT_HD float hlsl_fmod(float x, float y) {
    const float q = x / y;
    const float aq = fabsf(q);
    const float r = aq - floorf(aq);
    return (q >= 0.0f ? r : -r) * y;
}
GPRT_HD float hlsl_smoothstep01(float x) { const float s = saturate(x / 1.0f); return s * s * (3.0f - 2.0f * s); }

// Row-vector * 4x4 (row-major), point (w = 1) and vector forms; matches mul(float4(v,1), M).xyz
// and mul(v, (float3x3)M) of the shaders.
GPRT_HD V3 mul_point_rowvec(V3 v, const float* m) {
    return v3(v.x * m[0] + v.y * m[4] + v.z * m[8] + 1.0f * m[12],
              v.x * m[1] + v.y * m[5] + v.z * m[9] + 1.0f * m[13],
              v.x * m[2] + v.y * m[6] + v.z * m[10] + 1.0f * m[14]);
}
GPRT_HD V3 mul_vector_rowvec(V3 v, const float* m) {
    return v3(v.x * m[0] + v.y * m[4] + v.z * m[8], v.x * m[1] + v.y * m[5] + v.z * m[9], v.x * m[2] + v.y * m[6] + v.z * m[10]);
}
// 3x4 (column-vector form, D3D12 instance transform): M * [p; 1] and M3x3 * v
GPRT_HD V3 xform_point_3x4(const float* m, V3 p) {
    return v3(m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3], m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7], m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]);
}
GPRT_HD V3 xform_vector_3x4(const float* m, V3 v) {
    return v3(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[4] * v.x + m[5] * v.y + m[6] * v.z, m[8] * v.x + m[9] * v.y + m[10] * v.z);
}

} // namespace gprt
