// ref_harness.cpp -- TEST INFRASTRUCTURE (oracle, kind "reference").
//
// This translation unit compiles the reference's OWN shader sources as C++:
//   src/ConstantBuffers.h, src/HlslCompat.h, src/RaytracingShaderHelper.hlsli,
//   src/ProceduralPrimitivesLibrary.hlsli, src/AnalyticPrimitives.hlsli,
//   src/VolumetricPrimitives.hlsli, src/SignedDistancePrimitives.hlsli,
//   src/SignedDistanceFractals.hlsli, src/Raytracing.hlsl
// read from /root/reference/src by oracle/build_ref.py, which applies a handful of mechanical
// rewrites (in/out/inout -> C++ references, 1.#INF, the resource declarations) into a scratch
// directory outside the repository and compiles this file against them.  No reference source
// is copied into the repository; only the resulting shared object lands in oracle/_ref/.
//
// What is written here (nothing of it exists in the reference, which relies on the D3D12
// driver): the DXR execution model the shaders assume -- TraceRay, ReportHit, the system-value
// intrinsics, traversal over {plane triangles, procedural AABBs}, shader-table indexing.
// The semantics implemented are the "canonical order" of DESIGN.md section 3:
//   (i)   the triangle instance is tested first (Moller-Trumbore in object space, back faces
//         culled when the ray flag says so, hit needs TMin < t < TCurrent);
//   (ii)  every procedural AABB whose slab test passes on [TMin, min(TMax, t_triangle)] runs its
//         intersection shader with RayTCurrent() = min(TMax, t_triangle), independently;
//   (iii) the smallest reported t wins, ties go to the lowest geometry index, and a procedural
//         hit at exactly the triangle's t replaces it (ReportHit's interval is closed);
//   (iv)  RAY_FLAG_ACCEPT_FIRST_HIT_AND_END_SEARCH stops at the first reported hit.
// With several TLAS instances of a BLAS (two-level instancing) the same order runs
// over all instances of a kind in TLAS order: (i) over every triangle instance (the closest
// t wins, the first instance on a tie), (ii) over every AABB instance with the ray taken into that
// instance's object space and the hit-group record found through ITS
// InstanceContributionToHitGroupIndex; ties go to the lowest (instance, geometry index).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load the library built from this file.  The product (libgprt.so) never does.
#include "hlsl_shim.h"

#include <algorithm>
#include <cstdio>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/gprt.h" // input struct layouts only (the scene ABI)

#ifndef ORACLE_COUNTERS
#define ORACLE_COUNTERS 1
#endif

namespace hlsl {

// ---------------------------------------------------------------------------------------------
// Run-time knobs that replace compile-time constants of the reference (build_ref.py rewrites
// `#define MAX_RAY_RECURSION_DEPTH 3`, `const UINT MaxSteps = 512;`, `UINT MAX_STEPS = 128;`).
// ---------------------------------------------------------------------------------------------
static uint g_max_recursion_depth = 3;
static uint g_sdf_max_steps = 512;
static uint g_metaball_max_steps = 128;
static int g_last_threads = 1;
static uint g_fractal_iterations = 4; // FRACTAL_ITERATIONS_COUNT, src/ConstantBuffers.h:27

// ---------------------------------------------------------------------------------------------
// Work counters (thread-local, summed after a render).
// ---------------------------------------------------------------------------------------------
struct Counters {
    uint64_t radiance_rays = 0, shadow_rays = 0;
    uint64_t isect[3] = {0, 0, 0};
    uint64_t sdf_evals[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    uint64_t metaball_steps = 0;
    uint64_t aabb_tests = 0;
    void add(const Counters& o) {
        radiance_rays += o.radiance_rays; shadow_rays += o.shadow_rays;
        for (int i = 0; i < 3; i++) isect[i] += o.isect[i];
        for (int i = 0; i < 8; i++) sdf_evals[i] += o.sdf_evals[i];
        metaball_steps += o.metaball_steps; aabb_tests += o.aabb_tests;
    }
};
static thread_local Counters tl_counters;
#if ORACLE_COUNTERS
#define ORACLE_COUNT_SDF(type) (tl_counters.sdf_evals[(unsigned)(type) & 7u]++)
#define ORACLE_COUNT_MB_STEP() (tl_counters.metaball_steps++)
#else
#define ORACLE_COUNT_SDF(type) ((void)0)
#define ORACLE_COUNT_MB_STEP() ((void)0)
#endif

// ---------------------------------------------------------------------------------------------
// DXR system values, served from a thread-local ray context.
// ---------------------------------------------------------------------------------------------
struct DxrContext {
    float3 world_origin, world_direction;
    float3 object_origin, object_direction;
    float tmin = 0, tcurrent = 0;
    uint flags = 0;
    uint primitive_index = 0;
    float3x4 object_to_world;
};
static thread_local DxrContext tl_ray;
static thread_local uint3 tl_launch_index;
static uint3 g_launch_dim;

inline float3 WorldRayOrigin() { return tl_ray.world_origin; }
inline float3 WorldRayDirection() { return tl_ray.world_direction; }
inline float3 ObjectRayOrigin() { return tl_ray.object_origin; }
inline float3 ObjectRayDirection() { return tl_ray.object_direction; }
inline float RayTMin() { return tl_ray.tmin; }
inline float RayTCurrent() { return tl_ray.tcurrent; }
inline uint RayFlags() { return tl_ray.flags; }
inline uint PrimitiveIndex() { return tl_ray.primitive_index; }
inline float3x4 ObjectToWorld3x4() { return tl_ray.object_to_world; }
inline uint3 DispatchRaysIndex() { return tl_launch_index; }
inline uint3 DispatchRaysDimensions() { return g_launch_dim; }

} // namespace hlsl

// The reference's shared C++/HLSL ABI header, compiled in its HLSL flavour.
namespace hlsl {
#define HLSL
#include "ConstantBuffers.h"
} // namespace hlsl

namespace hlsl {

// ---------------------------------------------------------------------------------------------
// Resources the shaders bind through root signatures (src/Raytracing.hlsl:16-27; those lines
// are removed by the rewrite step and re-declared here).
// ---------------------------------------------------------------------------------------------
static RaytracingAccelerationStructure g_scene;
static thread_local RWTexture2D<float4> g_renderTarget;
static ConstantBuffer<SceneConstantBuffer> g_sceneCB;
static ByteAddressBuffer g_indices;
static StructuredBuffer<Vertex> g_vertices;
static StructuredBuffer<PrimitiveInstancePerFrameBuffer> g_AABBPrimitiveAttributes;
static thread_local ConstantBuffer<PrimitiveConstantBuffer> l_materialCB;
static thread_local ConstantBuffer<PrimitiveInstanceConstantBuffer> l_aabbCB;

void TraceRay(const RaytracingAccelerationStructure&, uint flags, uint mask, uint hit_group_offset,
              uint geometry_stride, uint miss_index, const RayDesc& ray, RayPayload& payload);
void TraceRay(const RaytracingAccelerationStructure&, uint flags, uint mask, uint hit_group_offset,
              uint geometry_stride, uint miss_index, const RayDesc& ray, ShadowRayPayload& payload);
bool ReportHit(float thit, uint hit_kind, const ProceduralPrimitiveAttributes& attr);

// The reference's shader library.
#include "Raytracing.hlsl"

// ---------------------------------------------------------------------------------------------
// Scene as handed over by the caller (same bytes the product receives through gprt.h).
// ---------------------------------------------------------------------------------------------
struct HostBvhNode {
    float lo[3], hi[3];
    int left, right; // children; leaf when left < 0: prim = ~left
};
struct Scene {
    std::vector<gprt_aabb> aabbs;
    std::vector<gprt_hit_group_record> records;
    std::vector<gprt_vertex> vertices;
    std::vector<uint16_t> indices; // padded by one dword so Load2 never reads past the end
    uint32_t n_indices = 0;
    struct Inst { float3x4 o2w, w2o; uint contrib = 0; };
    std::vector<Inst> tri_insts, aabb_insts; // TLAS instances per BLAS kind, in TLAS order
    bool has_tri = false, has_aabb = false;
    std::vector<gprt_prim_per_frame> per_frame;
    std::vector<HostBvhNode> bvh;
    bool use_bvh = false;
};
static Scene g_sc;

static float3x4 invert_affine(const float3x4& a) {
    // inverse of [R|t] computed in double, rounded once
    double r[3][3], t[3];
    for (int i = 0; i < 3; i++) { for (int j = 0; j < 3; j++) r[i][j] = a.m[i][j]; t[i] = a.m[i][3]; }
    double det = r[0][0] * (r[1][1] * r[2][2] - r[1][2] * r[2][1]) - r[0][1] * (r[1][0] * r[2][2] - r[1][2] * r[2][0]) +
                 r[0][2] * (r[1][0] * r[2][1] - r[1][1] * r[2][0]);
    double inv[3][3];
    inv[0][0] = (r[1][1] * r[2][2] - r[1][2] * r[2][1]) / det;
    inv[0][1] = (r[0][2] * r[2][1] - r[0][1] * r[2][2]) / det;
    inv[0][2] = (r[0][1] * r[1][2] - r[0][2] * r[1][1]) / det;
    inv[1][0] = (r[1][2] * r[2][0] - r[1][0] * r[2][2]) / det;
    inv[1][1] = (r[0][0] * r[2][2] - r[0][2] * r[2][0]) / det;
    inv[1][2] = (r[0][2] * r[1][0] - r[0][0] * r[1][2]) / det;
    inv[2][0] = (r[1][0] * r[2][1] - r[1][1] * r[2][0]) / det;
    inv[2][1] = (r[0][1] * r[2][0] - r[0][0] * r[2][1]) / det;
    inv[2][2] = (r[0][0] * r[1][1] - r[0][1] * r[1][0]) / det;
    float3x4 o;
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) o.m[i][j] = (float)inv[i][j];
        o.m[i][3] = (float)(-(inv[i][0] * t[0] + inv[i][1] * t[1] + inv[i][2] * t[2]));
    }
    return o;
}

static inline float3 xform_point(const float3x4& a, const float3& p) {
    return float3(a.m[0][0] * p.x + a.m[0][1] * p.y + a.m[0][2] * p.z + a.m[0][3],
                  a.m[1][0] * p.x + a.m[1][1] * p.y + a.m[1][2] * p.z + a.m[1][3],
                  a.m[2][0] * p.x + a.m[2][1] * p.y + a.m[2][2] * p.z + a.m[2][3]);
}
static inline float3 xform_vector(const float3x4& a, const float3& v) {
    return float3(a.m[0][0] * v.x + a.m[0][1] * v.y + a.m[0][2] * v.z,
                  a.m[1][0] * v.x + a.m[1][1] * v.y + a.m[1][2] * v.z,
                  a.m[2][0] * v.x + a.m[2][1] * v.y + a.m[2][2] * v.z);
}

// Canonical ray/AABB predicate: closed interval, IEEE 1/d, fminf/fmaxf.
static inline bool slab_test(const float3& o, const float3& inv_d, const float* lo, const float* hi,
                             float tmin, float tmax) {
    float t0x = (lo[0] - o.x) * inv_d.x, t1x = (hi[0] - o.x) * inv_d.x;
    float t0y = (lo[1] - o.y) * inv_d.y, t1y = (hi[1] - o.y) * inv_d.y;
    float t0z = (lo[2] - o.z) * inv_d.z, t1z = (hi[2] - o.z) * inv_d.z;
    float tnear = ::fmaxf(::fmaxf(::fminf(t0x, t1x), ::fminf(t0y, t1y)), ::fmaxf(::fminf(t0z, t1z), tmin));
    float tfar = ::fminf(::fminf(::fmaxf(t0x, t1x), ::fmaxf(t0y, t1y)), ::fminf(::fmaxf(t0z, t1z), tmax));
    return tnear <= tfar;
}

// Host BVH over the AABBs (median split); only used to find the same candidate set faster
// when there are many primitives.  The slab predicate is monotone under box inclusion, so
// the candidate set equals the brute-force one.
static int build_bvh_rec(std::vector<HostBvhNode>& nodes, std::vector<uint32_t>& ids, int lo, int hi) {
    HostBvhNode n;
    for (int k = 0; k < 3; k++) { n.lo[k] = INFINITY; n.hi[k] = -INFINITY; }
    for (int i = lo; i < hi; i++) {
        const float* b = &g_sc.aabbs[ids[i]].min_x;
        for (int k = 0; k < 3; k++) { n.lo[k] = ::fminf(n.lo[k], b[k]); n.hi[k] = ::fmaxf(n.hi[k], b[3 + k]); }
    }
    int self = (int)nodes.size();
    nodes.push_back(n);
    if (hi - lo == 1) { nodes[self].left = ~(int)ids[lo]; nodes[self].right = -1; return self; }
    int axis = 0;
    float ext = -1;
    for (int k = 0; k < 3; k++) if (n.hi[k] - n.lo[k] > ext) { ext = n.hi[k] - n.lo[k]; axis = k; }
    int mid = (lo + hi) / 2;
    std::nth_element(ids.begin() + lo, ids.begin() + mid, ids.begin() + hi, [&](uint32_t a, uint32_t b) {
        const float* pa = &g_sc.aabbs[a].min_x; const float* pb = &g_sc.aabbs[b].min_x;
        return pa[axis] + pa[axis + 3] < pb[axis] + pb[axis + 3];
    });
    int l = build_bvh_rec(nodes, ids, lo, mid);
    int r = build_bvh_rec(nodes, ids, mid, hi);
    nodes[self].left = l; nodes[self].right = r;
    return self;
}

// ---------------------------------------------------------------------------------------------
// TraceRay core.
// ---------------------------------------------------------------------------------------------
struct ReportedHit { bool any = false; float t = 0; ProceduralPrimitiveAttributes attr; };
static thread_local ReportedHit tl_reported;
static thread_local int tl_trace_depth = 0;
static thread_local uint32_t tl_primary_hit_id = GPRT_HIT_MISS;

bool ReportHit(float thit, uint /*hit_kind*/, const ProceduralPrimitiveAttributes& attr) {
    if (!(thit >= tl_ray.tmin && thit <= tl_ray.tcurrent)) return false;
    tl_reported.any = true;
    tl_reported.t = thit;
    tl_reported.attr = attr;
    return true;
}

struct CommittedHit {
    int kind = 0; // 0 miss, 1 triangle, 2 procedural
    float t = 0;
    uint index = 0; // triangle index or geometry index
    uint inst = 0;  // ordinal of the TLAS instance among those of its kind
    ProceduralPrimitiveAttributes attr;
    float2 bary;
};

static const gprt_hit_group_record& hit_record(uint idx) {
    if (idx >= g_sc.records.size()) { std::fprintf(stderr, "oracle: hit group index %u out of range\n", idx); std::abort(); }
    return g_sc.records[idx];
}
static void bind_local_root(const gprt_hit_group_record& rec) {
    std::memcpy(static_cast<PrimitiveConstantBuffer*>(&l_materialCB), &rec.material_cb, sizeof(PrimitiveConstantBuffer));
    std::memcpy(static_cast<PrimitiveInstanceConstantBuffer*>(&l_aabbCB), &rec.aabb_cb, sizeof(PrimitiveInstanceConstantBuffer));
}

static void run_candidate(uint inst, uint geom, uint hit_group_offset, uint stride, bool first_hit, CommittedHit& best, bool& done) {
    const gprt_hit_group_record& rec = hit_record(hit_group_offset + stride * geom + g_sc.aabb_insts[inst].contrib);
    bind_local_root(rec);
    tl_reported.any = false;
    switch (rec.shader_id[0]) {
    case GPRT_HG_AABB_ANALYTIC: tl_counters.isect[0]++; MyIntersectionShader_AnalyticPrimitive(); break;
    case GPRT_HG_AABB_VOLUMETRIC: tl_counters.isect[1]++; MyIntersectionShader_VolumetricPrimitive(); break;
    case GPRT_HG_AABB_SDF: tl_counters.isect[2]++; MyIntersectionShader_SignedDistancePrimitive(); break;
    default: std::fprintf(stderr, "oracle: bad hit group kind %u\n", rec.shader_id[0]); std::abort();
    }
    if (tl_reported.any) {
        if (best.kind != 2 || tl_reported.t < best.t || (tl_reported.t == best.t && (inst < best.inst || (inst == best.inst && geom < best.index)))) {
            best.kind = 2; best.t = tl_reported.t; best.index = geom; best.inst = inst; best.attr = tl_reported.attr;
        }
        if (first_hit) done = true;
    }
}

template <class Payload> struct PayloadOps;
template <> struct PayloadOps<RayPayload> {
    static void miss(RayPayload& p) { MyMissShader(p); }
    static void closest_triangle(RayPayload& p, const BuiltInTriangleIntersectionAttributes& a) { MyClosestHitShader_Triangle(p, a); }
    static void closest_aabb(RayPayload& p, const ProceduralPrimitiveAttributes& a) { MyClosestHitShader_AABB(p, a); }
    static constexpr bool radiance = true;
};
template <> struct PayloadOps<ShadowRayPayload> {
    static void miss(ShadowRayPayload& p) { MyMissShader_ShadowRay(p); }
    static void closest_triangle(ShadowRayPayload&, const BuiltInTriangleIntersectionAttributes&) {}
    static void closest_aabb(ShadowRayPayload&, const ProceduralPrimitiveAttributes&) {}
    static constexpr bool radiance = false;
};

template <class Payload>
static void trace_core(uint flags, uint hit_group_offset, uint stride, const RayDesc& ray, Payload& payload) {
    const DxrContext saved = tl_ray;
    const ConstantBuffer<PrimitiveConstantBuffer> saved_mat = l_materialCB;
    const ConstantBuffer<PrimitiveInstanceConstantBuffer> saved_aabb = l_aabbCB;
    const bool is_primary = (tl_trace_depth == 0);
    tl_trace_depth++;
    if (PayloadOps<Payload>::radiance) tl_counters.radiance_rays++; else tl_counters.shadow_rays++;

    tl_ray.world_origin = ray.Origin;
    tl_ray.world_direction = ray.Direction;
    tl_ray.tmin = ray.TMin;
    tl_ray.tcurrent = ray.TMax;
    tl_ray.flags = flags;

    const bool first_hit = (flags & RAY_FLAG_ACCEPT_FIRST_HIT_AND_END_SEARCH) != 0;
    const bool cull_back = (flags & RAY_FLAG_CULL_BACK_FACING_TRIANGLES) != 0;
    const bool cull_front = (flags & RAY_FLAG_CULL_FRONT_FACING_TRIANGLES) != 0;
    CommittedHit tri_hit, proc_hit;
    bool done = false;

    // (i) triangle instances
    float tbest = ray.TMax;
    for (uint ti = 0; g_sc.has_tri && ti < (uint)g_sc.tri_insts.size() && !done; ti++) {
        const float3 o = xform_point(g_sc.tri_insts[ti].w2o, ray.Origin);
        const float3 d = xform_vector(g_sc.tri_insts[ti].w2o, ray.Direction);
        for (uint tri = 0; tri * 3 + 2 < g_sc.n_indices && !done; tri++) {
            const gprt_vertex& a = g_sc.vertices[g_sc.indices[tri * 3 + 0]];
            const gprt_vertex& b = g_sc.vertices[g_sc.indices[tri * 3 + 1]];
            const gprt_vertex& c = g_sc.vertices[g_sc.indices[tri * 3 + 2]];
            const float3 v0(a.position[0], a.position[1], a.position[2]);
            const float3 e1 = float3(b.position[0], b.position[1], b.position[2]) - v0;
            const float3 e2 = float3(c.position[0], c.position[1], c.position[2]) - v0;
            const float3 pvec = cross(d, e2);
            const float det = dot(e1, pvec);
            // clockwise (left-handed) = front face  <=>  det > 0
            if (det == 0.0f) continue;
            if (cull_back && !(det > 0.0f)) continue;
            if (cull_front && !(det < 0.0f)) continue;
            const float inv_det = 1.0f / det;
            const float3 tvec = o - v0;
            const float u = dot(tvec, pvec) * inv_det;
            if (u < 0.0f || u > 1.0f) continue;
            const float3 qvec = cross(tvec, e1);
            const float v = dot(d, qvec) * inv_det;
            if (v < 0.0f || u + v > 1.0f) continue;
            const float t = dot(e2, qvec) * inv_det;
            if (t > ray.TMin && t < tbest) {
                tbest = t;
                tri_hit.kind = 1; tri_hit.t = t; tri_hit.index = tri; tri_hit.inst = ti; tri_hit.bary = float2(u, v);
                if (first_hit) done = true;
            }
        }
    }

    // (ii) procedural AABB instances, all candidates see TCurrent = min(TMax, t_triangle)
    for (uint k = 0; g_sc.has_aabb && k < (uint)g_sc.aabb_insts.size() && !done; k++) {
        const float tcur = tri_hit.kind ? tri_hit.t : ray.TMax;
        tl_ray.tcurrent = tcur;
        tl_ray.object_origin = xform_point(g_sc.aabb_insts[k].w2o, ray.Origin);
        tl_ray.object_direction = xform_vector(g_sc.aabb_insts[k].w2o, ray.Direction);
        tl_ray.object_to_world = g_sc.aabb_insts[k].o2w;
        const float3 o = tl_ray.object_origin;
        const float3 inv_d(1.0f / tl_ray.object_direction.x, 1.0f / tl_ray.object_direction.y, 1.0f / tl_ray.object_direction.z);
        if (!g_sc.use_bvh) {
            for (uint i = 0; i < (uint)g_sc.aabbs.size() && !done; i++) {
                tl_counters.aabb_tests++;
                if (slab_test(o, inv_d, &g_sc.aabbs[i].min_x, &g_sc.aabbs[i].max_x, ray.TMin, tcur))
                    run_candidate(k, i, hit_group_offset, stride, first_hit, proc_hit, done);
            }
        } else {
            uint32_t cand[256];
            std::vector<uint32_t> spill;
            uint32_t ncand = 0;
            int stack[128];
            int sp = 0;
            stack[sp++] = 0;
            while (sp) {
                const HostBvhNode& n = g_sc.bvh[stack[--sp]];
                tl_counters.aabb_tests++;
                if (!slab_test(o, inv_d, n.lo, n.hi, ray.TMin, tcur)) continue;
                if (n.left < 0) {
                    if (ncand < 256) cand[ncand++] = (uint32_t)~n.left; else spill.push_back((uint32_t)~n.left);
                } else { stack[sp++] = n.right; stack[sp++] = n.left; }
            }
            if (!spill.empty()) { spill.insert(spill.end(), cand, cand + ncand); std::sort(spill.begin(), spill.end()); for (uint32_t g : spill) { if (done) break; run_candidate(k, g, hit_group_offset, stride, first_hit, proc_hit, done); } }
            else { std::sort(cand, cand + ncand); for (uint32_t c = 0; c < ncand && !done; c++) run_candidate(k, cand[c], hit_group_offset, stride, first_hit, proc_hit, done); }
        }
    }

    // (iii) commit
    CommittedHit hit = tri_hit;
    if (proc_hit.kind && (!tri_hit.kind || proc_hit.t <= tri_hit.t)) hit = proc_hit;
    if (is_primary && PayloadOps<Payload>::radiance)
        tl_primary_hit_id = hit.kind == 0 ? GPRT_HIT_MISS
                          : hit.kind == 1 ? (GPRT_HIT_TRIANGLE | (hit.inst * (g_sc.n_indices / 3u) + hit.index))
                                          : (GPRT_HIT_AABB | (hit.inst * (uint)g_sc.aabbs.size() + hit.index));

    if (hit.kind == 0) {
        PayloadOps<Payload>::miss(payload);
    } else if (!(flags & RAY_FLAG_SKIP_CLOSEST_HIT_SHADER)) {
        tl_ray.tcurrent = hit.t;
        tl_ray.primitive_index = hit.kind == 1 ? hit.index : 0;
        if (hit.kind == 1) {
            tl_ray.object_to_world = g_sc.tri_insts[hit.inst].o2w;
            bind_local_root(hit_record(hit_group_offset + stride * 0 + g_sc.tri_insts[hit.inst].contrib));
            BuiltInTriangleIntersectionAttributes a; a.barycentrics = hit.bary;
            PayloadOps<Payload>::closest_triangle(payload, a);
        } else {
            tl_ray.object_to_world = g_sc.aabb_insts[hit.inst].o2w;
            bind_local_root(hit_record(hit_group_offset + stride * hit.index + g_sc.aabb_insts[hit.inst].contrib));
            PayloadOps<Payload>::closest_aabb(payload, hit.attr);
        }
    }

    tl_trace_depth--;
    tl_ray = saved;
    l_materialCB = saved_mat;
    l_aabbCB = saved_aabb;
}

void TraceRay(const RaytracingAccelerationStructure&, uint flags, uint, uint hit_group_offset, uint geometry_stride, uint,
              const RayDesc& ray, RayPayload& payload) { trace_core(flags, hit_group_offset, geometry_stride, ray, payload); }
void TraceRay(const RaytracingAccelerationStructure&, uint flags, uint, uint hit_group_offset, uint geometry_stride, uint,
              const RayDesc& ray, ShadowRayPayload& payload) { trace_core(flags, hit_group_offset, geometry_stride, ray, payload); }

static inline unsigned char to_unorm8(float c) {
    if (!(c == c)) return 0;
    c = c < 0.0f ? 0.0f : (c > 1.0f ? 1.0f : c);
    return (unsigned char)::floorf(c * 255.0f + 0.5f);
}

} // namespace hlsl

// =============================================================================================
// C interface (ctypes / dlopen from tests and bench.py)
// =============================================================================================
extern "C" {

struct oracle_config {
    uint32_t max_recursion_depth, sdf_max_steps, metaball_max_steps, force_brute_force;
    uint32_t fractal_iterations; // 0 = the reference's 4
};
struct oracle_counters {
    uint64_t radiance_rays, shadow_rays, isect[3], sdf_evals[8], metaball_steps, aabb_tests;
};

int oracle_metaballs_count(void) { return METABALLS_COUNT; }
const char* oracle_kind(void) { return "reference"; }

int oracle_set_scene(const gprt_aabb* aabbs, uint32_t n_aabbs, const gprt_vertex* vertices, uint32_t n_vertices,
                     const uint16_t* indices, uint32_t n_indices, const gprt_instance_desc* instances, uint32_t n_instances,
                     const void* records, uint32_t n_records, uint32_t stride_bytes, const oracle_config* cfg) {
    using namespace hlsl;
    g_sc = Scene();
    g_sc.aabbs.assign(aabbs, aabbs + n_aabbs);
    g_sc.vertices.assign(vertices, vertices + n_vertices);
    g_sc.indices.assign(indices, indices + n_indices);
    g_sc.indices.resize(n_indices + 4, 0);
    g_sc.n_indices = n_indices;
    g_sc.records.resize(n_records);
    for (uint32_t i = 0; i < n_records; i++) std::memcpy(&g_sc.records[i], (const char*)records + (size_t)i * stride_bytes, sizeof(gprt_hit_group_record));
    for (uint32_t i = 0; i < n_instances; i++) {
        float3x4 o2w;
        std::memcpy(o2w.m, instances[i].transform, sizeof o2w.m);
        uint contrib = instances[i].hit_group_contrib_and_flags & 0xFFFFFFu;
        Scene::Inst inst;
        inst.o2w = o2w; inst.w2o = invert_affine(o2w); inst.contrib = contrib;
        if (instances[i].acceleration_structure == GPRT_BLAS_TRIANGLES) {
            g_sc.has_tri = n_indices >= 3; g_sc.tri_insts.push_back(inst);
        } else if (instances[i].acceleration_structure == GPRT_BLAS_AABBS) {
            g_sc.has_aabb = n_aabbs > 0; g_sc.aabb_insts.push_back(inst);
        } else return -1;
    }
    if (cfg) {
        g_max_recursion_depth = cfg->max_recursion_depth;
        g_sdf_max_steps = cfg->sdf_max_steps;
        g_metaball_max_steps = cfg->metaball_max_steps;
        g_fractal_iterations = cfg->fractal_iterations ? cfg->fractal_iterations : 4u;
    }
    g_sc.use_bvh = n_aabbs > 64 && !(cfg && cfg->force_brute_force);
    if (g_sc.use_bvh) {
        std::vector<uint32_t> ids(n_aabbs);
        for (uint32_t i = 0; i < n_aabbs; i++) ids[i] = i;
        g_sc.bvh.reserve(2 * (size_t)n_aabbs);
        build_bvh_rec(g_sc.bvh, ids, 0, (int)n_aabbs);
    }
    g_indices.base = reinterpret_cast<const unsigned char*>(g_sc.indices.data());
    g_vertices.base = reinterpret_cast<const Vertex*>(g_sc.vertices.data());
    static_assert(sizeof(Vertex) == sizeof(gprt_vertex), "Vertex layout");
    static_assert(sizeof(PrimitiveInstancePerFrameBuffer) == sizeof(gprt_prim_per_frame), "per-frame layout");
    static_assert(sizeof(PrimitiveConstantBuffer) == sizeof(gprt_material), "material layout");
    static_assert(sizeof(SceneConstantBuffer) == 136, "scene cb layout (HLSL packing)");
    return 0;
}

// Render `tile` of a width x height dispatch.  Any of the outputs may be null.
int oracle_render(const gprt_scene_cb* scene, const gprt_prim_per_frame* per_frame, uint32_t n_per_frame,
                  uint32_t width, uint32_t height, gprt_rect tile, uint8_t* rgba8, float* rgba_f32, uint32_t* hit_ids,
                  oracle_counters* counters, int n_threads) {
    using namespace hlsl;
    std::memcpy(static_cast<SceneConstantBuffer*>(&g_sceneCB), scene, sizeof(SceneConstantBuffer));
    g_sc.per_frame.assign(per_frame, per_frame + n_per_frame);
    g_AABBPrimitiveAttributes.base = reinterpret_cast<const PrimitiveInstancePerFrameBuffer*>(g_sc.per_frame.data());
    g_launch_dim = uint3(width, height, 1);
    Counters total;
#ifdef _OPENMP
    if (n_threads <= 0) n_threads = omp_get_max_threads();
#else
    n_threads = 1;
#endif
#pragma omp parallel num_threads(n_threads)
    {
#ifdef _OPENMP
#pragma omp single
        g_last_threads = omp_get_num_threads();
#endif
        tl_counters = Counters();
        std::vector<float4> row(tile.w);
#pragma omp for schedule(dynamic, 1)
        for (int64_t ty = 0; ty < (int64_t)tile.h; ty++) {
            g_renderTarget.base = row.data();
            g_renderTarget.pitch = 0;
            g_renderTarget.origin_x = tile.x;
            g_renderTarget.origin_y = tile.y + (uint32_t)ty;
            for (uint32_t tx = 0; tx < tile.w; tx++) {
                tl_launch_index = uint3(tile.x + tx, tile.y + (uint32_t)ty, 0);
                tl_trace_depth = 0;
                tl_primary_hit_id = GPRT_HIT_MISS;
                MyRaygenShader();
                const size_t o = (size_t)ty * tile.w + tx;
                if (hit_ids) hit_ids[o] = tl_primary_hit_id;
                const float4 c = row[tx];
                if (rgba_f32) { rgba_f32[4 * o + 0] = c.x; rgba_f32[4 * o + 1] = c.y; rgba_f32[4 * o + 2] = c.z; rgba_f32[4 * o + 3] = c.w; }
                if (rgba8) { rgba8[4 * o + 0] = to_unorm8(c.x); rgba8[4 * o + 1] = to_unorm8(c.y); rgba8[4 * o + 2] = to_unorm8(c.z); rgba8[4 * o + 3] = to_unorm8(c.w); }
            }
        }
#pragma omp critical
        total.add(tl_counters);
    }
    if (counters) {
        counters->radiance_rays = total.radiance_rays; counters->shadow_rays = total.shadow_rays;
        for (int i = 0; i < 3; i++) counters->isect[i] = total.isect[i];
        for (int i = 0; i < 8; i++) counters->sdf_evals[i] = total.sdf_evals[i];
        counters->metaball_steps = total.metaball_steps; counters->aabb_tests = total.aabb_tests;
    }
    return 0;
}

// threads the last oracle_render really ran on (omp_get_num_threads inside its parallel region)
int oracle_last_threads(void) { return hlsl::g_last_threads; }

int oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

// ---- function-level probes (known-answer pins of tests/test_oracle_functions.py) -------------
// One intersection-shader body in the primitive's LOCAL space (no transforms), as
// RayAnalyticGeometryIntersectionTest / RayVolumetricGeometryIntersectionTest /
// RaySignedDistancePrimitiveTest see it.
int oracle_intersect_local(uint32_t hit_group_kind, uint32_t primitive_type, const float* origin, const float* direction,
                           float tmin, float tcurrent, uint32_t ray_flags, float elapsed_time, float step_scale,
                           float* thit_out, float* normal_out) {
    using namespace hlsl;
    const DxrContext saved = tl_ray;
    tl_ray.tmin = tmin; tl_ray.tcurrent = tcurrent; tl_ray.flags = ray_flags;
    Ray ray = {float3(origin[0], origin[1], origin[2]), float3(direction[0], direction[1], direction[2])};
    float thit = 0;
    ProceduralPrimitiveAttributes attr = ProceduralPrimitiveAttributes{};
    bool hit = false;
    switch (hit_group_kind) {
    case GPRT_HG_AABB_ANALYTIC: hit = RayAnalyticGeometryIntersectionTest(ray, (AnalyticPrimitive::Enum)primitive_type, thit, attr); break;
    case GPRT_HG_AABB_VOLUMETRIC: hit = RayVolumetricGeometryIntersectionTest(ray, (VolumetricPrimitive::Enum)primitive_type, thit, attr, elapsed_time); break;
    case GPRT_HG_AABB_SDF: hit = RaySignedDistancePrimitiveTest(ray, (SignedDistancePrimitive::Enum)primitive_type, thit, attr, step_scale); break;
    default: break;
    }
    tl_ray = saved;
    if (hit) { *thit_out = thit; normal_out[0] = attr.normal.x; normal_out[1] = attr.normal.y; normal_out[2] = attr.normal.z; }
    return hit ? 1 : 0;
}
// FRACTAL_ITERATIONS_COUNT for the function-level probes below (oracle_set_scene sets it for renders)
void oracle_set_fractal_iterations(uint32_t n) { hlsl::g_fractal_iterations = n ? n : 4u; }

float oracle_sdf_distance(uint32_t primitive_type, const float* p) {
    return hlsl::GetDistanceFromSignedDistancePrimitive(hlsl::float3(p[0], p[1], p[2]), (hlsl::SignedDistancePrimitive::Enum)primitive_type);
}
float oracle_checkers_box_filter(const float* uv, const float* dpdx, const float* dpdy, uint32_t ratio) {
    return hlsl::CheckersTextureBoxFilter(hlsl::float2(uv[0], uv[1]), hlsl::float2(dpdx[0], dpdx[1]), hlsl::float2(dpdy[0], dpdy[1]), ratio);
}
void oracle_fresnel_schlick(const float* i, const float* n, const float* f0, float* out) {
    hlsl::float3 r = hlsl::FresnelReflectanceSchlick(hlsl::float3(i[0], i[1], i[2]), hlsl::float3(n[0], n[1], n[2]), hlsl::float3(f0[0], f0[1], f0[2]));
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
float oracle_animation_interpolant(float elapsed, float cycle) { return hlsl::CalculateAnimationInterpolant(elapsed, cycle); }
void oracle_camera_ray(const gprt_scene_cb* scene, uint32_t width, uint32_t height, uint32_t x, uint32_t y, float* origin, float* direction) {
    using namespace hlsl;
    std::memcpy(static_cast<SceneConstantBuffer*>(&g_sceneCB), scene, sizeof(SceneConstantBuffer));
    g_launch_dim = uint3(width, height, 1);
    Ray r = GenerateCameraRay(uint2(x, y), g_sceneCB.camera_position.xyz, g_sceneCB.projection_to_world);
    origin[0] = r.origin.x; origin[1] = r.origin.y; origin[2] = r.origin.z;
    direction[0] = r.direction.x; direction[1] = r.direction.y; direction[2] = r.direction.z;
}

} // extern "C"
