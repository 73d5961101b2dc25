// hlsl_shim.h -- TEST INFRASTRUCTURE (oracle).  Not part of the product; nothing under
// gpuraytracer_b200/ may include this.
//
// A small HLSL-in-C++ vocabulary, just large enough that the reference's shader sources
// (src/Raytracing.hlsl and the five .hlsli files, after the mechanical rewrites done by
// oracle/build_ref.py) compile as C++17.  The reference's own src/HlslCompat.h is only five
// typedefs (XMFLOAT3 -> float3 ...) and supplies none of this.
//
// Numeric conventions are pinned to how dxc lowers the intrinsics in the reference's
// precompiled blob (res/compiled/Raytracing.hlsl.h):
//   normalize(v) = v * (1/sqrt(dot(v,v)))      length(v) = sqrt(x*x + y*y + z*z)
//   pow(x,y)     = exp2(log2(x) * y)            exp(x)    = exp2(x * log2(e))
//   fmod(x,y)    = q = x/y; r = frac(|q|); (q >= 0 ? r : -r) * y
//   smoothstep(0,1,x) = s*s*(3-2s), s = saturate(x)
// Sums are evaluated left to right; the build uses -ffp-contract=off so no FMA is formed.
// Everything lives in namespace hlsl so unqualified sqrt/pow/abs/... resolve here and never
// to the double-precision C library functions.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

namespace hlsl {

typedef unsigned int uint;

struct float2; struct float3; struct float4; struct uint2; struct uint3;

// ---- swizzle proxies (members of anonymous unions; must stay trivial) ----------------------
template <int N, int A, int B> struct Swz2 {
    float d[N];
    operator float2() const;
    Swz2& operator=(const float2& v);
};
template <int N, int A, int B, int C> struct Swz3 {
    float d[N];
    operator float3() const;
    Swz3& operator=(const float3& v);
    Swz3& operator/=(float s) { d[A] /= s; d[B] /= s; d[C] /= s; return *this; }
    Swz3& operator*=(float s) { d[A] *= s; d[B] *= s; d[C] *= s; return *this; }
};
template <int N, int A, int B> struct USwz2 {
    uint d[N];
    operator uint2() const;
    operator float2() const;
};

struct float2 {
    union {
        struct { float x, y; };
        Swz2<2, 0, 1> xy; Swz2<2, 1, 0> yx; Swz2<2, 0, 0> xx; Swz2<2, 1, 1> yy;
        Swz3<2, 0, 1, 1> xyy; Swz3<2, 1, 1, 0> yyx; Swz3<2, 1, 0, 1> yxy; Swz3<2, 0, 0, 0> xxx;
    };
    float2() : x(0), y(0) {}
    float2(float s) : x(s), y(s) {}
    float2(float x_, float y_) : x(x_), y(y_) {}
    float2(const float2& o) : x(o.x), y(o.y) {}
    float2(const uint2& u);
    float2& operator=(const float2& o) { x = o.x; y = o.y; return *this; }
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
};

struct float3 {
    union {
        struct { float x, y, z; };
        Swz2<3, 0, 1> xy; Swz2<3, 0, 2> xz; Swz2<3, 1, 2> yz; Swz2<3, 1, 0> yx; Swz2<3, 2, 0> zx;
        Swz3<3, 0, 1, 2> xyz; Swz3<3, 0, 2, 1> xzy;
    };
    float3() : x(0), y(0), z(0) {}
    float3(float s) : x(s), y(s), z(s) {}
    float3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    float3(const float2& a, float z_) : x(a.x), y(a.y), z(z_) {}
    float3(const float3& o) : x(o.x), y(o.y), z(o.z) {}
    float3& operator=(const float3& o) { x = o.x; y = o.y; z = o.z; return *this; }
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
};

struct float4 {
    union {
        struct { float x, y, z, w; };
        Swz2<4, 0, 1> xy; Swz2<4, 0, 2> xz;
        Swz3<4, 0, 1, 2> xyz;
    };
    float4() : x(0), y(0), z(0), w(0) {}
    float4(float s) : x(s), y(s), z(s), w(s) {}
    float4(float x_, float y_, float z_, float w_) : x(x_), y(y_), z(z_), w(w_) {}
    float4(const float3& a, float w_) : x(a.x), y(a.y), z(a.z), w(w_) {}
    float4(const float2& a, float z_, float w_) : x(a.x), y(a.y), z(z_), w(w_) {}
    float4(const float4& o) : x(o.x), y(o.y), z(o.z), w(o.w) {}
    float4& operator=(const float4& o) { x = o.x; y = o.y; z = o.z; w = o.w; return *this; }
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
};

struct uint2 {
    uint x, y;
    uint2() : x(0), y(0) {}
    uint2(uint x_, uint y_) : x(x_), y(y_) {}
};
struct uint3 {
    union {
        struct { uint x, y, z; };
        USwz2<3, 0, 1> xy;
    };
    uint3() : x(0), y(0), z(0) {}
    uint3(uint x_, uint y_, uint z_) : x(x_), y(y_), z(z_) {}
    uint& operator[](int i) { return (&x)[i]; }
    uint operator[](int i) const { return (&x)[i]; }
};
struct int3 {
    int x, y, z;
    int3() : x(0), y(0), z(0) {}
    int3(int x_, int y_, int z_) : x(x_), y(y_), z(z_) {}
};
struct bool3 {
    bool x, y, z;
    operator int3() const { return int3(x ? 1 : 0, y ? 1 : 0, z ? 1 : 0); }
};

inline float2::float2(const uint2& u) : x((float)u.x), y((float)u.y) {}
template <int N, int A, int B> Swz2<N, A, B>::operator float2() const { return float2(d[A], d[B]); }
template <int N, int A, int B> Swz2<N, A, B>& Swz2<N, A, B>::operator=(const float2& v) { d[A] = v.x; d[B] = v.y; return *this; }
template <int N, int A, int B, int C> Swz3<N, A, B, C>::operator float3() const { return float3(d[A], d[B], d[C]); }
template <int N, int A, int B, int C> Swz3<N, A, B, C>& Swz3<N, A, B, C>::operator=(const float3& v) { d[A] = v.x; d[B] = v.y; d[C] = v.z; return *this; }
template <int N, int A, int B> USwz2<N, A, B>::operator uint2() const { return uint2(d[A], d[B]); }
template <int N, int A, int B> USwz2<N, A, B>::operator float2() const { return float2((float)d[A], (float)d[B]); }

inline uint2 operator+(const uint2& a, const uint2& b) { return uint2(a.x + b.x, a.y + b.y); }

// ---- component-wise arithmetic ---------------------------------------------------------------
#define HLSL_BINOP(OP)                                                                              \
    inline float2 operator OP(const float2& a, const float2& b) { return float2(a.x OP b.x, a.y OP b.y); } \
    inline float3 operator OP(const float3& a, const float3& b) { return float3(a.x OP b.x, a.y OP b.y, a.z OP b.z); } \
    inline float4 operator OP(const float4& a, const float4& b) { return float4(a.x OP b.x, a.y OP b.y, a.z OP b.z, a.w OP b.w); }
HLSL_BINOP(+) HLSL_BINOP(-) HLSL_BINOP(*) HLSL_BINOP(/)
#undef HLSL_BINOP
#define HLSL_ASSIGNOP(OP)                                                                           \
    inline float2& operator OP##=(float2& a, const float2& b) { a = a OP b; return a; }              \
    inline float3& operator OP##=(float3& a, const float3& b) { a = a OP b; return a; }              \
    inline float4& operator OP##=(float4& a, const float4& b) { a = a OP b; return a; }
HLSL_ASSIGNOP(+) HLSL_ASSIGNOP(-) HLSL_ASSIGNOP(*) HLSL_ASSIGNOP(/)
#undef HLSL_ASSIGNOP
inline float2 operator-(const float2& a) { return float2(-a.x, -a.y); }
inline float3 operator-(const float3& a) { return float3(-a.x, -a.y, -a.z); }
inline float4 operator-(const float4& a) { return float4(-a.x, -a.y, -a.z, -a.w); }

inline bool3 operator>(const float3& a, const float3& b) { return bool3{a.x > b.x, a.y > b.y, a.z > b.z}; }
inline bool3 operator!=(const float3& a, const float3& b) { return bool3{a.x != b.x, a.y != b.y, a.z != b.z}; }
inline float3 select(const bool3& c, const float3& a, const float3& b) {
    return float3(c.x ? a.x : b.x, c.y ? a.y : b.y, c.z ? a.z : b.z);
}

// ---- scalar intrinsics -----------------------------------------------------------------------
inline float sqrt(float x) { return ::sqrtf(x); }
inline float rsqrt(float x) { return 1.0f / ::sqrtf(x); }
inline float abs(float x) { return ::fabsf(x); }
inline float min(float a, float b) { return ::fminf(a, b); }
inline float max(float a, float b) { return ::fmaxf(a, b); }
inline float floor(float x) { return ::floorf(x); }
inline float frac(float x) { return x - ::floorf(x); }
inline float saturate(float x) { return ::fminf(::fmaxf(x, 0.0f), 1.0f); }
inline float clamp(float x, float lo, float hi) { return ::fminf(::fmaxf(x, lo), hi); }
inline float lerp(float a, float b, float s) { return a + s * (b - a); }
inline float sign(float x) { return (float)((x > 0.0f) - (x < 0.0f)); }
inline float log2(float x) { return ::log2f(x); }
inline float exp2(float x) { return ::exp2f(x); }
inline float pow(float x, float y) { return ::exp2f(::log2f(x) * y); }
inline float exp(float x) { return ::exp2f(x * 1.44269504088896340736f); }
inline float sin(float x) { return ::sinf(x); }
inline float cos(float x) { return ::cosf(x); }
inline float atan2(float y, float x) { return ::atan2f(y, x); }
inline float fmod(float x, float y) {
    float q = x / y;
    float aq = ::fabsf(q);
    float r = aq - ::floorf(aq);
    return (q >= 0.0f ? r : -r) * y;
}
inline float smoothstep(float a, float b, float x) {
    float s = saturate((x - a) / (b - a));
    return s * s * (3.0f - 2.0f * s);
}

// ---- vector intrinsics -----------------------------------------------------------------------
#define HLSL_MAP1(F)                                                                     \
    inline float2 F(const float2& a) { return float2(F(a.x), F(a.y)); }                   \
    inline float3 F(const float3& a) { return float3(F(a.x), F(a.y), F(a.z)); }           \
    inline float4 F(const float4& a) { return float4(F(a.x), F(a.y), F(a.z), F(a.w)); }
HLSL_MAP1(abs) HLSL_MAP1(floor) HLSL_MAP1(frac) HLSL_MAP1(saturate)
#undef HLSL_MAP1
#define HLSL_MAP2(F)                                                                                     \
    inline float2 F(const float2& a, const float2& b) { return float2(F(a.x, b.x), F(a.y, b.y)); }        \
    inline float3 F(const float3& a, const float3& b) { return float3(F(a.x, b.x), F(a.y, b.y), F(a.z, b.z)); } \
    inline float4 F(const float4& a, const float4& b) { return float4(F(a.x, b.x), F(a.y, b.y), F(a.z, b.z), F(a.w, b.w)); }
HLSL_MAP2(min) HLSL_MAP2(max) HLSL_MAP2(fmod)
#undef HLSL_MAP2
inline float2 lerp(const float2& a, const float2& b, float s) { return a + float2(s) * (b - a); }
inline float3 lerp(const float3& a, const float3& b, float s) { return a + float3(s) * (b - a); }
inline float4 lerp(const float4& a, const float4& b, float s) { return a + float4(s) * (b - a); }

inline float dot(const float2& a, const float2& b) { return a.x * b.x + a.y * b.y; }
inline float dot(const float3& a, const float3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float dot(const float4& a, const float4& b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
inline float length(const float2& a) { return ::sqrtf(dot(a, a)); }
inline float length(const float3& a) { return ::sqrtf(dot(a, a)); }
inline float2 normalize(const float2& a) { return a * float2(rsqrt(dot(a, a))); }
inline float3 normalize(const float3& a) { return a * float3(rsqrt(dot(a, a))); }
inline float3 reflect(const float3& i, const float3& n) { return i - float3(2.0f * dot(i, n)) * n; }
inline float3 cross(const float3& a, const float3& b) {
    return float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}

// ---- matrices (row-major storage, as the reference compiles with /Zpr) ---------------------
struct float3x4 { float m[3][4]; };
struct float4x4 { float m[4][4]; };
struct float3x3 {
    float m[3][3];
    float3x3() { std::memset(m, 0, sizeof m); }
    explicit float3x3(const float4x4& a) { for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) m[i][j] = a.m[i][j]; }
    explicit float3x3(const float3x4& a) { for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) m[i][j] = a.m[i][j]; }
};
struct float2x2 {
    float m[2][2];
    float2x2(float a, float b, float c, float d) { m[0][0] = a; m[0][1] = b; m[1][0] = c; m[1][1] = d; }
};
// row vector * matrix
inline float3 mul(const float3& v, const float3x3& a) {
    return float3(v.x * a.m[0][0] + v.y * a.m[1][0] + v.z * a.m[2][0],
                  v.x * a.m[0][1] + v.y * a.m[1][1] + v.z * a.m[2][1],
                  v.x * a.m[0][2] + v.y * a.m[1][2] + v.z * a.m[2][2]);
}
inline float4 mul(const float4& v, const float4x4& a) {
    float4 r;
    for (int j = 0; j < 4; j++) r[j] = v.x * a.m[0][j] + v.y * a.m[1][j] + v.z * a.m[2][j] + v.w * a.m[3][j];
    return r;
}
// matrix * column vector
inline float3 mul(const float3x3& a, const float3& v) {
    return float3(a.m[0][0] * v.x + a.m[0][1] * v.y + a.m[0][2] * v.z,
                  a.m[1][0] * v.x + a.m[1][1] * v.y + a.m[1][2] * v.z,
                  a.m[2][0] * v.x + a.m[2][1] * v.y + a.m[2][2] * v.z);
}
inline float2 mul(const float2x2& a, const float2& v) {
    return float2(a.m[0][0] * v.x + a.m[0][1] * v.y, a.m[1][0] * v.x + a.m[1][1] * v.y);
}

// ---- resources -------------------------------------------------------------------------------
struct ByteAddressBuffer {
    const unsigned char* base = nullptr;
    uint2 Load2(uint byte_offset) const {
        uint v[2];
        std::memcpy(v, base + byte_offset, 8);
        return uint2(v[0], v[1]);
    }
};
template <class T> struct StructuredBuffer {
    const T* base = nullptr;
    const T& operator[](uint i) const { return base[i]; }
};
template <class T> struct ConstantBuffer : public T {
    ConstantBuffer& operator=(const T& t) { static_cast<T&>(*this) = t; return *this; }
};
template <class T> struct RWTexture2D {
    T* base = nullptr;
    uint pitch = 0; // elements per row
    uint origin_x = 0, origin_y = 0; // launch index of base[0]
    T& operator[](const uint2& p) { return base[(size_t)(p.y - origin_y) * pitch + (p.x - origin_x)]; }
};
struct RaytracingAccelerationStructure {};
struct RayDesc {
    float3 Origin;
    float TMin;
    float3 Direction;
    float TMax;
};
struct BuiltInTriangleIntersectionAttributes { float2 barycentrics; };

enum : uint {
    RAY_FLAG_NONE = 0x00,
    RAY_FLAG_FORCE_OPAQUE = 0x01,
    RAY_FLAG_FORCE_NON_OPAQUE = 0x02,
    RAY_FLAG_ACCEPT_FIRST_HIT_AND_END_SEARCH = 0x04,
    RAY_FLAG_SKIP_CLOSEST_HIT_SHADER = 0x08,
    RAY_FLAG_CULL_BACK_FACING_TRIANGLES = 0x10,
    RAY_FLAG_CULL_FRONT_FACING_TRIANGLES = 0x20,
    RAY_FLAG_CULL_OPAQUE = 0x40,
    RAY_FLAG_CULL_NON_OPAQUE = 0x80,
};

} // namespace hlsl
