// prims.cuh -- ray vs procedural primitive tests in the primitive's LOCAL space.
//
// Device-side counterparts of the reference's intersection-shader bodies:
//   analytic    RayAnalyticGeometryIntersectionTest  src/ProceduralPrimitivesLibrary.hlsli:32-46,
//               src/AnalyticPrimitives.hlsli:28-219
//   volumetric  RayMetaballsIntersectionTest         src/VolumetricPrimitives.hlsli:36-202
//   SDF         RaySignedDistancePrimitiveTest       src/SignedDistancePrimitives.hlsli:55-319,
//               src/ProceduralPrimitivesLibrary.hlsli:63-98, src/SignedDistanceFractals.hlsli:34-63
// Written from the algorithm, organised for the wavefront kernels (one primitive type per
// instantiation, march state explicit so a lane can be refilled).  Arithmetic order follows the
// HLSL expressions so the parity build agrees with the CPU oracle to the last bit wherever only
// + - * / sqrt are involved.
//
// Both TraceRay call sites of the reference pass RAY_FLAG_CULL_BACK_FACING_TRIANGLES and never
// CULL_FRONT (src/Raytracing.hlsl:104,136), so IsCulled reduces to dot(dir, normal) > 0.
#pragma once
#include "vecmath.cuh"

namespace gprt {

struct LocalRay {
    V3 o, d;
    float tmin, tcur; // RayTMin(), RayTCurrent()
};

// IsAValidHit, src/RaytracingShaderHelper.hlsli:136-139 (IsInRange + IsCulled)
GPRT_HD bool is_valid_hit(const LocalRay& r, float thit, V3 n) {
    const bool in_range = (thit >= r.tmin) && (thit <= r.tcur);
    return in_range && !(dot(r.d, n) > 0.0f);
}

// ---------------------------------------------------------------------------------------------
// Analytic primitives
// ---------------------------------------------------------------------------------------------

// SolveQuadraticEqn, src/AnalyticPrimitives.hlsli:28-43
GPRT_HD bool solve_quadratic(float a, float b, float c, float& x0, float& x1) {
    const float discr = b * b - 4.0f * a * c;
    if (discr < 0.0f) return false;
    if (discr == 0.0f) {
        x0 = x1 = -0.5f * b / a;
    } else {
        const float sq = sqrtf(discr);
        const float q = (b > 0.0f) ? -0.5f * (b + sq) : -0.5f * (b - sq);
        x0 = q / a;
        x1 = c / q;
    }
    if (x0 > x1) { const float t = x0; x0 = x1; x1 = t; }
    return true;
}

// SolveRaySphereIntersectionEquation, src/AnalyticPrimitives.hlsli:54-61
GPRT_HD bool ray_sphere_roots(const LocalRay& r, V3 center, float radius, float& t0, float& t1) {
    const V3 L = r.o - center;
    const float a = dot(r.d, r.d);
    const float b = 2.0f * dot(r.d, L);
    const float c = dot(L, L) - radius * radius;
    return solve_quadratic(a, b, c, t0, t1);
}

GPRT_HD V3 sphere_normal(const LocalRay& r, float thit, V3 center) { // CalculateNormalForARaySphereHit, :46-50
    const V3 p = r.o + thit * r.d;
    return normalize(p - center);
}

// RaySphereIntersectionTest (hollow sphere), src/AnalyticPrimitives.hlsli:64-100
GPRT_HD bool ray_hollow_sphere(const LocalRay& r, V3 center, float radius, float& thit, V3& n) {
    float t0, t1;
    if (!ray_sphere_roots(r, center, radius, t0, t1)) return false;
    if (t0 < r.tmin) {
        if (t1 < r.tmin) return false;
        n = sphere_normal(r, t1, center);
        if (is_valid_hit(r, t1, n)) { thit = t1; return true; }
    } else {
        n = sphere_normal(r, t0, center);
        if (is_valid_hit(r, t0, n)) { thit = t0; return true; }
        n = sphere_normal(r, t1, center);
        if (is_valid_hit(r, t1, n)) { thit = t1; return true; }
    }
    return false;
}

// RaySpheresIntersectionTest, src/AnalyticPrimitives.hlsli:119-153
GPRT_HD bool ray_spheres(const LocalRay& r, float& thit, V3& n) {
    const float cx[3] = {-0.3f, 0.1f, 0.35f}, cy[3] = {-0.3f, 0.1f, 0.35f}, cz[3] = {-0.3f, 0.4f, 0.0f};
    const float radii[3] = {0.6f, 0.3f, 0.15f};
    bool found = false;
    thit = r.tcur;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        float t;
        V3 nn = v3s(0.0f);
        if (ray_hollow_sphere(r, v3(cx[i], cy[i], cz[i]), radii[i], t, nn)) {
            if (t < thit) { thit = t; n = nn; found = true; }
        }
    }
    return found;
}

// RayAABBIntersectionTest (both overloads), src/AnalyticPrimitives.hlsli:158-218, box [-1,1]^3
GPRT_HD bool ray_unit_box(const LocalRay& r, float& thit, V3& n) {
    const float kInf = __builtin_inff();
    // select(d != 0, 1/d, select(d > 0, +inf, -inf))
    const float ix = (r.d.x != 0.0f) ? 1.0f / r.d.x : -kInf;
    const float iy = (r.d.y != 0.0f) ? 1.0f / r.d.y : -kInf;
    const float iz = (r.d.z != 0.0f) ? 1.0f / r.d.z : -kInf;
    // aabb[1 - sign] is the near plane, aabb[sign] the far plane (sign = d > 0)
    const float nx = r.d.x > 0.0f ? -1.0f : 1.0f, fx = r.d.x > 0.0f ? 1.0f : -1.0f;
    const float ny = r.d.y > 0.0f ? -1.0f : 1.0f, fy = r.d.y > 0.0f ? 1.0f : -1.0f;
    const float nz = r.d.z > 0.0f ? -1.0f : 1.0f, fz = r.d.z > 0.0f ? 1.0f : -1.0f;
    const float tmin = fmaxf(fmaxf((nx - r.o.x) * ix, (ny - r.o.y) * iy), (nz - r.o.z) * iz);
    const float tmax = fminf(fminf((fx - r.o.x) * ix, (fy - r.o.y) * iy), (fz - r.o.z) * iz);
    if (!(tmax > tmin && tmax >= r.tmin && tmin <= r.tcur)) return false;
    if (tmin < r.tmin || tmin > r.tcur) return false; // only hits entering from outside
    thit = tmin;
    const V3 p = r.o + thit * r.d;
    const V3 d0 = vabs(v3s(-1.0f) - p), d1 = vabs(v3s(1.0f) - p);
    const float eps = 0.0001f;
    n = v3s(0.0f); // attr is zero-initialised by the intersection shader (src/Raytracing.hlsl:296)
    if (d0.x < eps) n = v3(-1, 0, 0);
    else if (d0.y < eps) n = v3(0, -1, 0);
    else if (d0.z < eps) n = v3(0, 0, -1);
    else if (d1.x < eps) n = v3(1, 0, 0);
    else if (d1.y < eps) n = v3(0, 1, 0);
    else if (d1.z < eps) n = v3(0, 0, 1);
    return is_valid_hit(r, thit, n);
}

// ---------------------------------------------------------------------------------------------
// Metaballs
// ---------------------------------------------------------------------------------------------
template <int NB> struct Metaballs {
    float cx[NB], cy[NB], cz[NB], radius[NB];
    int n_active; // nActiveMetaballs (only meaningful for NB >= 5: LIMIT_TO_ACTIVE_METABALLS)
};

// CalculateAnimationInterpolant, src/RaytracingShaderHelper.hlsli:36-41
GPRT_HD float animation_interpolant(float elapsed, float cycle) {
    float c = hlsl_fmod(elapsed, cycle) / cycle;
    c = (c <= 0.5f) ? 2.0f * c : 1.0f - 2.0f * (c - 0.5f);
    return hlsl_smoothstep01(c);
}

// InitializeAnimatedMetaballs, src/VolumetricPrimitives.hlsli:88-120
template <int NB> GPRT_HD void init_metaballs(Metaballs<NB>& mb, float elapsed_time) {
    const float ta = animation_interpolant(elapsed_time, 12.0f);
    if (NB == 5) {
        const float k0[5][3] = {{-0.7f, 0, 0}, {0.7f, 0, 0}, {0, -0.7f, 0}, {0, 0.7f, 0}, {0, 0, 0}};
        const float k1[5][3] = {{0.7f, 0, 0}, {-0.7f, 0, 0}, {0, 0.7f, 0}, {0, -0.7f, 0}, {0, 0, 0}};
        const float rr[5] = {0.35f, 0.35f, 0.35f, 0.35f, 0.25f};
#pragma unroll
        for (int j = 0; j < NB; j++) {
            mb.cx[j] = lerp(k0[j][0], k1[j][0], ta); mb.cy[j] = lerp(k0[j][1], k1[j][1], ta); mb.cz[j] = lerp(k0[j][2], k1[j][2], ta);
            mb.radius[j] = rr[j];
        }
    } else {
        const float k0[3][3] = {{-0.3f, -0.3f, -0.4f}, {0.0f, -0.2f, 0.5f}, {0.4f, 0.4f, 0.4f}};
        const float k1[3][3] = {{0.3f, -0.3f, -0.0f}, {0.0f, 0.4f, 0.5f}, {-0.4f, 0.2f, -0.4f}};
        const float rr[3] = {0.45f, 0.55f, 0.45f};
#pragma unroll
        for (int j = 0; j < NB; j++) {
            mb.cx[j] = lerp(k0[j][0], k1[j][0], ta); mb.cy[j] = lerp(k0[j][1], k1[j][1], ta); mb.cz[j] = lerp(k0[j][2], k1[j][2], ta);
            mb.radius[j] = rr[j];
        }
    }
    mb.n_active = NB;
}

// CalculateMetaballPotential, src/VolumetricPrimitives.hlsli:36-57
GPRT_HD float metaball_potential(V3 p, float cx, float cy, float cz, float r) {
    const float distance = length(p - v3(cx, cy, cz));
    if (distance <= r) {
        const float d = r - distance;
        return 6.0f * (d * d * d * d * d) / (r * r * r * r * r) - 15.0f * (d * d * d * d) / (r * r * r * r) + 10.0f * (d * d * d) / (r * r * r);
    }
    return 0.0f;
}

// CalculateMetaballsPotential, :60-73.  NB >= 5 builds loop over the active balls only.
template <int NB> GPRT_HD float metaballs_potential(V3 p, const Metaballs<NB>& mb) {
    float sum = 0.0f;
#pragma unroll
    for (int j = 0; j < NB; j++) {
        if (NB >= 5 && j >= mb.n_active) break;
        sum += metaball_potential(p, mb.cx[j], mb.cy[j], mb.cz[j], mb.radius[j]);
    }
    return sum;
}

// CalculateMetaballsNormal (central differences), :76-86
template <int NB> GPRT_HD V3 metaballs_normal(V3 p, const Metaballs<NB>& mb) {
    const float e = 0.5773f * 0.00001f;
    return normalize(v3(metaballs_potential(p + v3(-e, 0, 0), mb) - metaballs_potential(p + v3(e, 0, 0), mb),
                        metaballs_potential(p + v3(0, -e, 0), mb) - metaballs_potential(p + v3(0, e, 0), mb),
                        metaballs_potential(p + v3(0, 0, -e), mb) - metaballs_potential(p + v3(0, 0, e), mb)));
}

// FindIntersectingMetaballs, :124-147 (RaySolidSphereIntersectionTest, AnalyticPrimitives :104-116)
template <int NB> GPRT_HD void find_intersecting_metaballs(const LocalRay& r, Metaballs<NB>& mb, float& tmin, float& tmax) {
    const float kInf = __builtin_inff();
    tmin = kInf;
    tmax = -kInf;
    int n_active = 0;
#pragma unroll
    for (int i = 0; i < NB; i++) {
        float t0, t1;
        if (ray_sphere_roots(r, v3(mb.cx[i], mb.cy[i], mb.cz[i]), mb.radius[i], t0, t1)) {
            const float th = fmaxf(t0, r.tmin), tm = fminf(t1, r.tcur);
            tmin = fminf(th, tmin);
            tmax = fmaxf(tm, tmax);
            if (NB >= 5) {
                // blobs[nActive++] = blobs[i]  (compaction, n_active <= i)
#pragma unroll
                for (int k = 0; k < NB; k++)
                    if (k == n_active) { mb.cx[k] = mb.cx[i]; mb.cy[k] = mb.cy[i]; mb.cz[k] = mb.cz[i]; mb.radius[k] = mb.radius[i]; }
                n_active++;
            } else {
                n_active = NB;
            }
        }
    }
    mb.n_active = n_active;
    tmin = fmaxf(tmin, r.tmin);
    tmax = fminf(tmax, r.tcur);
}

// RayMetaballsIntersectionTest, :151-202.  max_steps = MAX_STEPS (128).
template <int NB> GPRT_HD bool ray_metaballs(const LocalRay& r, float elapsed_time, uint32_t max_steps, float& thit, V3& n, uint32_t& work) {
    Metaballs<NB> mb;
    init_metaballs<NB>(mb, elapsed_time);
    float tmin, tmax;
    find_intersecting_metaballs<NB>(r, mb, tmin, tmax);
    // No bounding sphere crossed: tmin = +inf, every sample is at infinity and the field is 0
    // there, so the reference's 128 iterations cannot report a hit.
    if (tmin == __builtin_inff()) return false;
    float t = tmin;
    const float min_t_step = (tmax - tmin) / (float)(max_steps / 1u);
    for (uint32_t i = 0; i < max_steps; i++) {
        const V3 p = r.o + t * r.d;
        const float sum = metaballs_potential<NB>(p, mb);
        work++;
        if (sum >= 0.25f) {
            const V3 nn = metaballs_normal<NB>(p, mb);
            if (is_valid_hit(r, t, nn)) { thit = t; n = nn; return true; }
        }
        t += min_t_step;
    }
    return false;
}

// ---------------------------------------------------------------------------------------------
// Signed distance primitives
// ---------------------------------------------------------------------------------------------
GPRT_HD float sd_sphere(V3 p, float s) { return length(p) - s; }
GPRT_HD float sd_box(V3 p, V3 b) {
    const V3 d = vabs(p) - b;
    return fminf(fmaxf(d.x, fmaxf(d.y, d.z)), 0.0f) + length(vmax(d, 0.0f));
}
GPRT_HD float ud_round_box(V3 p, V3 b, float r) { return length(vmax(vabs(p) - b, 0.0f)) - r; }
GPRT_HD V3 op_rep(V3 p, V3 c) {
    return v3(hlsl_fmod(p.x, c.x) - 0.5f * c.x, hlsl_fmod(p.y, c.y) - 0.5f * c.y, hlsl_fmod(p.z, c.z) - 0.5f * c.z);
}
GPRT_HD float sd_torus(V3 p, V2 t) {
    const V2 q = v2(length(v2(p.x, p.z)) - t.x, p.y);
    return length(q) - t.y;
}
GPRT_HD float length8(V2 p) { // length_toPowNegative8
    p = p * p; p = p * p; p = p * p;
    return hlsl_pow(p.x + p.y, 1.0f / 8.0f);
}
GPRT_HD float sd_torus82(V3 p, V2 t) {
    const V2 q = v2(length(v2(p.x, p.z)) - t.x, p.y);
    return length8(q) - t.y;
}
GPRT_HD float sd_cylinder(V3 p, V2 h) {
    const V2 d = vabs(v2(length(v2(p.x, p.z)), p.y)) - h;
    return fminf(fmaxf(d.x, d.y), 0.0f) + length(vmax(d, 0.0f));
}
GPRT_HD V3 op_twist(V3 p) {
    const float c = cosf(3.0f * p.y), s = sinf(3.0f * p.y);
    // float2x2(c, -s, s, c) * p.xz
    return v3(c * p.x + (-s) * p.z, s * p.x + c * p.z, p.y);
}
GPRT_HD float sd_pyramid(V3 p, V3 h) { // sdOctahedron + opS(octa, p.y)
    const float d = fmaxf(fabsf(p.x), fabsf(p.z)) * h.x + fabsf(p.y) * h.y;
    const float octa = d - h.y * h.z;
    return fmaxf(octa, -p.y);
}
// sdFractalPyramid, src/SignedDistanceFractals.hlsli:34-63; Scale = 2.  One fold towards the nearest of the five
// vertices: position = Scale * position - v * (Scale - 1).
GPRT_HD V3 fractal_fold(V3 p, float a, float hz) {
    const V3 v1 = v3(0, hz, 0), v2_ = v3(-a, 0, a), v3_ = v3(a, 0, -a), v4_ = v3(a, 0, a), v5_ = v3(-a, 0, -a);
    V3 v = v1;
    V3 q = p - v1;
    float dist = dot(q, q), d;
    q = p - v2_; d = dot(q, q); if (d < dist) { v = v2_; dist = d; }
    q = p - v3_; d = dot(q, q); if (d < dist) { v = v3_; dist = d; }
    q = p - v4_; d = dot(q, q); if (d < dist) { v = v4_; dist = d; }
    q = p - v5_; d = dot(q, q); if (d < dist) { v = v5_; dist = d; }
    return v3s(2.0f) * p - v * v3s(2.0f - 1.0f);
}
// `iters` = FRACTAL_ITERATIONS_COUNT (src/ConstantBuffers.h:27; 4 in the reference, a run-time knob here, <= 8).
GPRT_HD float sd_fractal_pyramid(V3 p, V3 h, uint32_t iters) {
    const float a = h.z * h.y / h.x;
    // Not unrolled: the folds are one dependent chain anyway, and the march kernels hold this body once per SDF type --
    // ncu showed their warps starved for instructions (stall_no_instruction 6 per issue) with 215 KB of unrolled code.
#pragma unroll 1
    for (uint32_t n = 0; n < iters; n++) p = fractal_fold(p, a, h.z);
    // pow(Scale, -n) = exp2(log2(2) * -n) = 2^-n exactly
#if defined(__CUDA_ARCH__)
    const float inv_scale = __uint_as_float((127u - iters) << 23);
#else
    const float inv_scale = ldexpf(1.0f, -(int)iters);
#endif
    return sd_pyramid(p, h) * inv_scale;
}

// GetDistanceFromSignedDistancePrimitive, src/ProceduralPrimitivesLibrary.hlsli:63-98
template <int TYPE> GPRT_HD float sdf_distance(V3 p, uint32_t fractal_iters = 4u) {
    if (TYPE == 0) { // MiniSpheres
        return fmaxf(sd_sphere(op_rep(p + v3s(1.0f), v3s(2.0f) / v3s(4.0f)), 0.65f / 4.0f), sd_box(p, v3s(1.0f)));
    } else if (TYPE == 1) { // IntersectedRoundCube: opS(opS(roundbox, sphere 1.20), -sphere 1.32)
        const float a = fmaxf(ud_round_box(p, v3s(0.75f), 0.2f), -sd_sphere(p, 1.20f));
        return fmaxf(a, -(-sd_sphere(p, 1.32f)));
    } else if (TYPE == 2) { // SquareTorus
        return sd_torus82(p, v2(0.75f, 0.15f));
    } else if (TYPE == 3) { // TwistedTorus
        return sd_torus(op_twist(p), v2(0.6f, 0.2f));
    } else if (TYPE == 4) { // Cog
        const V3 q = v3(atan2f(p.z, p.x) / 6.2831f, 1.0f, 0.015f + 0.25f * length(p)) + v3s(1.0f);
        return fmaxf(sd_torus82(p, v2(0.60f, 0.3f)), -sd_cylinder(op_rep(q, v3(0.05f, 1.0f, 0.075f)), v2(0.02f, 0.8f)));
    } else if (TYPE == 5) { // Cylinder
        return fmaxf(sd_cylinder(op_rep(p + v3(1, 1, 1), v3(1, 2, 1)), v2(0.3f, 2.0f)), sd_box(p + v3(1, 1, 1), v3(2, 2, 2)));
    } else if (TYPE == 6) { // FractalPyramid
        return sd_fractal_pyramid(p + v3(0, 1, 0), v3(0.894f, 0.447f, 2.0f), fractal_iters);
    }
    return 0.0f;
}

// sdCalculateNormal (tetrahedral taps), src/SignedDistancePrimitives.hlsli:275-284
template <int TYPE> GPRT_HD V3 sdf_normal(V3 pos, uint32_t fractal_iters = 4u) {
    const float ex = 1.0f * 0.5773f * 0.0001f, ey = -1.0f * 0.5773f * 0.0001f;
    const V3 xyy = v3(ex, ey, ey), yyx = v3(ey, ey, ex), yxy = v3(ey, ex, ey), xxx = v3(ex, ex, ex);
    const V3 s = ((xyy * v3s(sdf_distance<TYPE>(pos + xyy, fractal_iters)) + yyx * v3s(sdf_distance<TYPE>(pos + yyx, fractal_iters))) +
                  yxy * v3s(sdf_distance<TYPE>(pos + yxy, fractal_iters))) +
                 xxx * v3s(sdf_distance<TYPE>(pos + xxx, fractal_iters));
    return normalize(s);
}

// RaySignedDistancePrimitiveTest (sphere tracing), src/SignedDistancePrimitives.hlsli:287-319
template <int TYPE> GPRT_HD bool ray_sdf(const LocalRay& r, float step_scale, uint32_t max_steps, float& thit, V3& n, uint32_t& work, uint32_t fractal_iters = 4u) {
    const float threshold = 0.0001f;
    float t = r.tmin;
    for (uint32_t i = 0; i < max_steps && t <= r.tcur; i++) {
        const V3 p = r.o + t * r.d;
        const float distance = sdf_distance<TYPE>(p, fractal_iters);
        work++;
        if (distance <= threshold * t) {
            const V3 nn = sdf_normal<TYPE>(p, fractal_iters);
            work += 4u;
            if (is_valid_hit(r, t, nn)) { thit = t; n = nn; return true; }
        }
        t += step_scale * distance;
    }
    return false;
}

} // namespace gprt
