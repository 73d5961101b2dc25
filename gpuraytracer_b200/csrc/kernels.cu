// kernels.cu -- the wavefront kernels.  Compiled twice into libgprt.so:
//   -DGPRT_KERNELS_PARITY   --fmad=false                  (IEEE, oracle-exact arithmetic)
//   -DGPRT_KERNELS_FAST     --use_fast_math -DGPRT_FAST_MATH
//   -DGPRT_KERNELS_FMA      --fmad=true, IEEE div/sqrt, no flush-to-zero (march kernels only)
// See wavefront.cuh for the pipeline; prims.cuh / shade.cuh for the math and the reference
// file:line each piece follows.
#include <cuda_runtime.h>

#include <cstdlib>

#include "prims.cuh"
#include "wavefront.cuh"

namespace gprt {
namespace {

constexpr int kBlock = 256;
constexpr unsigned kFull = 0xFFFFFFFFu;
constexpr int kRefillMin = 8; // idle lanes of a warp that trigger a refill from the queue (measured: 1, 8, 16, 24 within 3 %)

__device__ __forceinline__ unsigned long long pack_key(float t, uint32_t code) {
    // t >= 0 here, so the IEEE bit pattern orders like the value; fabsf folds -0 into +0
    return ((unsigned long long)__float_as_uint(fabsf(t)) << 32) | (unsigned long long)code;
}

__device__ __forceinline__ V3 ld3(const float4 v) { return v3(v.x, v.y, v.z); }

// Canonical ray/AABB predicate: closed interval, IEEE 1/d, fminf/fmaxf.
__device__ __forceinline__ bool slab_test(V3 o, V3 inv_d, const float* lo, const float* hi, float tmin, float tmax) {
    const float t0x = (lo[0] - o.x) * inv_d.x, t1x = (hi[0] - o.x) * inv_d.x;
    const float t0y = (lo[1] - o.y) * inv_d.y, t1y = (hi[1] - o.y) * inv_d.y;
    const float t0z = (lo[2] - o.z) * inv_d.z, t1z = (hi[2] - o.z) * inv_d.z;
    const float tnear = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fmaxf(fminf(t0z, t1z), tmin));
    const float tfar = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fminf(fmaxf(t0z, t1z), tmax));
    return tnear <= tfar;
}

// ---------------------------------------------------------------------------------------------
// traverse: plane triangles (Moller-Trumbore, back faces culled) then the LBVH.
// A ray's candidates are collected in registers while it walks the tree and appended to the
// per-type queues afterwards, at a point where the whole warp is convergent: per-warp counts come
// from shared-memory atomics, then ten lanes issue the (at most ten) global reservations at once.
// One global-atomic latency per warp and batch of 32 rays, instead of one per leaf hit in the
// middle of the walk (measured: the walk spent 18 % of its time waiting for those returns and for
// the primitive's type, which now travels in the leaf reference).
// ---------------------------------------------------------------------------------------------
constexpr int kMaxPending = 6;         // candidates kept in registers per ray; more are appended at once (rare)
constexpr uint32_t kLeafTypeShift = 27; // leaf ref = ~(primitive | type id << 27)
constexpr uint32_t kLeafPrimMask = (1u << kLeafTypeShift) - 1u;

__device__ __noinline__ void emit_candidate_now(const PassParams& pp, const WaveBuffers& wb, WaveCtrl* ctrl, uint32_t ray, uint32_t prim,
                                                   uint32_t type) {
    // warp-aggregated append from divergent code: one atomic per (warp, type)
    const unsigned peers = __match_any_sync(__activemask(), type);
    const int leader = __ffs(peers) - 1;
    const int lane = threadIdx.x & 31;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(&ctrl->cand_count[type], (uint32_t)__popc(peers));
    base = __shfl_sync(peers, base, leader);
    const uint32_t idx = base + (uint32_t)__popc(peers & ((1u << lane) - 1u));
    if (idx < pp.cand_capacity) wb.cand[type][idx] = make_uint2(ray, prim);
    else ctrl->overflow = 1u;
}

struct Pending {
    uint32_t e[kMaxPending]; // primitive | type << 27
    int n;
};

__device__ __forceinline__ void add_candidate(const PassParams& pp, const WaveBuffers& wb, WaveCtrl* ctrl, uint32_t ray, uint32_t prim, uint32_t type,
                                              Pending& pend) {
    if (pend.n < kMaxPending) {
#pragma unroll
        for (int k = 0; k < kMaxPending; k++)
            if (k == pend.n) pend.e[k] = prim | (type << kLeafTypeShift);
        pend.n++;
    } else {
        emit_candidate_now(pp, wb, ctrl, ray, prim, type);
    }
}

// A leaf the ray's slab test passed: `tagged` = primitive | type id << 27.  The tag is the type in the FIRST AABB
// instance's hit groups; if another instance maps a primitive to another intersection shader, the record decides.
// (MULTI = false: the reference's TLAS shape, one instance per BLAS; the kernels are compiled for both.)
template <bool MULTI>
__device__ __forceinline__ void add_leaf(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, WaveCtrl* ctrl, uint32_t r, uint32_t flat0,
                                         uint32_t tagged, Pending& pend) {
    const uint32_t flat = (MULTI ? flat0 : 0u) + (tagged & kLeafPrimMask);
    const uint32_t type = (!MULTI || sc.leaf_types_valid) ? (tagged >> kLeafTypeShift) : sc.prims[flat].type_id;
    add_candidate(pp, wb, ctrl, r, flat, type, pend);
}

// The triangle instances of the TLAS, in TLAS order (the reference has one: the ground quad): Moller-Trumbore, back faces
// culled, open interval; a shadow ray stops at its first hit.  Updates tcur / key.
template <bool MULTI>
__device__ __forceinline__ void plane_triangles(const SceneDev& sc, V3 o, V3 d, bool shadow, float& tcur, unsigned long long& key) {
    const float tmin = 0.0f;
    const uint32_t n_ti = MULTI ? sc.n_tri_insts : min(sc.n_tri_insts, 1u);
    for (uint32_t ti = 0; ti < n_ti; ti++) {
        const float* w2o = (MULTI && ti) ? sc.tri_insts[ti].w2o : sc.tri_w2o;
        const V3 oo = xform_point_3x4(w2o, o), dd = xform_vector_3x4(w2o, d);
        for (uint32_t tri = 0; tri < sc.n_tris; tri++) {
            const float* p = sc.tri_pos + 9 * tri;
            const V3 v0 = v3(p[0], p[1], p[2]);
            const V3 e1 = v3(p[3], p[4], p[5]) - v0, e2 = v3(p[6], p[7], p[8]) - v0;
            const V3 pvec = cross(dd, e2);
            const float det = dot(e1, pvec);
            if (!(det > 0.0f)) continue; // clockwise = front; back faces culled
            const float inv_det = 1.0f / det;
            const V3 tvec = oo - v0;
            const float u = dot(tvec, pvec) * inv_det;
            if (u < 0.0f || u > 1.0f) continue;
            const V3 qvec = cross(tvec, e1);
            const float v = dot(dd, qvec) * inv_det;
            if (v < 0.0f || u + v > 1.0f) continue;
            const float t = dot(e2, qvec) * inv_det;
            if (t > tmin && t < tcur) {
                tcur = t;
                key = pack_key(t, kCodePlane | ((MULTI ? ti * sc.n_tris : 0u) + tri));
                if (shadow) return;
            }
        }
    }
}

// One ray without the tree: plane, then EVERY AABB of every instance (GPRT_FLAG_BRUTE_FORCE_TRAVERSAL, and scenes of a
// single primitive, which have no tree).  Writes hitkey / ray_tcur, collects candidates in `pend`.  The LBVH walk is in
// walk.cuh.
template <bool PRIMARY, bool MULTI>
__device__ __forceinline__ void traverse_ray(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, WaveCtrl* ctrl, int pq, uint32_t r,
                                             Pending& pend) {
    V3 o, d;
    uint32_t meta;
    if (PRIMARY) {
        const uint32_t px = pp.tile_x + r % pp.tile_w, py = pass_global_row(pp, r / pp.tile_w);
        generate_camera_ray(px, py, pp.fc, o, d);
        meta = make_meta(r, 1u, 0u);
        wb.ray_o[pq][r] = make_float4(o.x, o.y, o.z, 0.0f);
        wb.ray_d[pq][r] = make_float4(d.x, d.y, d.z, __uint_as_float(meta));
    } else {
        const float4 ro = wb.ray_o[pq][r], rd = wb.ray_d[pq][r];
        o = ld3(ro); d = ld3(rd);
        meta = __float_as_uint(rd.w);
    }
    const bool shadow = meta_shadow(meta) != 0u;
    const float tmin = 0.0f;
    float tcur = kRayTMax;
    unsigned long long key = kKeyMiss;
    plane_triangles<MULTI>(sc, o, d, shadow, tcur, key);
    wb.hitkey[r] = key;
    wb.ray_tcur[r] = tcur;
    if (shadow && key != kKeyMiss) return; // ACCEPT_FIRST_HIT_AND_END_SEARCH
    if (!sc.has_aabb_instance || sc.n_prims == 0) return;
    const uint32_t n_ai = MULTI ? sc.n_aabb_insts : 1u;
    for (uint32_t k = 0; k < n_ai; k++) {
        const float* w2o = (MULTI && k) ? sc.aabb_insts[k].w2o : sc.aabb_w2o;
        const V3 oo = xform_point_3x4(w2o, o), dd = xform_vector_3x4(w2o, d);
        const V3 inv_d = v3(1.0f / dd.x, 1.0f / dd.y, 1.0f / dd.z);
        for (uint32_t i = 0; i < sc.n_prims; i++) {
            const float* b = sc.aabbs + 6 * i;
            const uint32_t flat = (MULTI ? k * sc.n_prims : 0u) + i;
            if (slab_test(oo, inv_d, b, b + 3, tmin, tcur)) add_candidate(pp, wb, ctrl, r, flat, sc.prims[flat].type_id, pend);
        }
    }
}

// Convergent append of a warp's pending candidates (all 32 lanes call this).
__device__ __forceinline__ void flush_pending(const PassParams& pp, const WaveBuffers& wb, WaveCtrl* ctrl, uint32_t r, const Pending& pend,
                                              uint32_t (*s_cnt)[kNumTypes], uint32_t (*s_base)[kNumTypes]) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (!__any_sync(kFull, pend.n > 0)) return;
    if (lane < kNumTypes) s_cnt[warp][lane] = 0u;
    __syncwarp();
    uint32_t rank[kMaxPending];
#pragma unroll
    for (int k = 0; k < kMaxPending; k++)
        rank[k] = (k < pend.n) ? atomicAdd(&s_cnt[warp][pend.e[k] >> kLeafTypeShift], 1u) : 0u;
    __syncwarp();
    if (lane < kNumTypes) {
        const uint32_t c = s_cnt[warp][lane];
        s_base[warp][lane] = c ? atomicAdd(&ctrl->cand_count[lane], c) : 0u;
    }
    __syncwarp();
#pragma unroll
    for (int k = 0; k < kMaxPending; k++)
        if (k < pend.n) {
            const uint32_t type = pend.e[k] >> kLeafTypeShift;
            const uint32_t idx = s_base[warp][type] + rank[k];
            if (idx < pp.cand_capacity) wb.cand[type][idx] = make_uint2(r, pend.e[k] & kLeafPrimMask);
            else ctrl->overflow = 1u;
        }
    __syncwarp();
}

// traversal statistics (gprt_get_wave_stats): kept per lane, two atomics per warp when the kernel ends
// (same-address atomics retire at about one per two clocks: per batch of 32 rays they cost 0.2 ms at 1080p)
struct VisitStats {
    uint32_t sum = 0, most = 0;
    __device__ __forceinline__ void add(uint32_t visits) { sum += visits; most = max(most, visits); }
    __device__ __forceinline__ void commit(WaveCtrl* ctrl) const {
        const uint32_t vsum = __reduce_add_sync(kFull, sum), vmax = __reduce_max_sync(kFull, most);
        if ((threadIdx.x & 31) == 0 && vsum) { atomicAdd(&ctrl->node_visits, vsum); atomicMax(&ctrl->max_node_visits, vmax); }
    }
};

template <bool PRIMARY, bool MULTI>
__global__ void __launch_bounds__(kBlock, 6) k_traverse_brute(const __grid_constant__ SceneDev sc, const __grid_constant__ PassParams pp,
                                                              const __grid_constant__ WaveBuffers wb, int wave) {
    __shared__ uint32_t s_cnt[kBlock / 32][kNumTypes], s_base[kBlock / 32][kNumTypes];
    WaveCtrl* ctrl = &wb.ctrl[wave];
    const int pq = wave & 1;
    const uint32_t n_rays = PRIMARY ? pp.n_pixels : min(ctrl->n_rays, pp.ray_capacity);
    const uint32_t n_round = (n_rays + 31u) & ~31u;
    for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n_round; r += gridDim.x * blockDim.x) {
        Pending pend;
        pend.n = 0;
#pragma unroll
        for (int k = 0; k < kMaxPending; k++) pend.e[k] = 0u;
        if (r < n_rays) traverse_ray<PRIMARY, MULTI>(sc, pp, wb, ctrl, pq, r, pend);
        flush_pending(pp, wb, ctrl, r, pend, s_cnt, s_base);
    }
}

// ---------------------------------------------------------------------------------------------
// intersect<TYPE>: the intersection-shader body of one primitive type over its candidate queue
// (MyIntersectionShader_* + GetRayInAABBPrimitiveLocalSpace, src/Raytracing.hlsl:277-341).
// Persistent blocks pull chunks of the queue; its length is only known on the device.
// ---------------------------------------------------------------------------------------------
// The TLAS instance a flattened candidate / hit id belongs to: its world <-> object transforms.
struct InstXform { const float* w2o; const float* o2w; };
__device__ __forceinline__ InstXform inst_xform(const SceneDev& sc, uint32_t flat) {
    InstXform x;
    x.w2o = sc.aabb_w2o; x.o2w = sc.aabb_o2w;
    if (sc.n_aabb_insts > 1u) {
        // flat / n_prims without the division sequence (inv == 0: one primitive, flat is the instance)
        const uint32_t k = sc.inv_n_prims ? (uint32_t)__umul64hi((unsigned long long)flat, sc.inv_n_prims) : flat;
        if (k) { x.w2o = sc.aabb_insts[k].w2o; x.o2w = sc.aabb_insts[k].o2w; }
    }
    return x;
}

template <int TYPE, int NB>
__device__ __forceinline__ bool run_test(const LocalRay& lr, const Material& mat, const PassParams& pp, float& thit, V3& n, uint32_t& work) {
    if (TYPE == 0) { work++; return ray_unit_box(lr, thit, n); }
    if (TYPE == 1) { work++; return ray_spheres(lr, thit, n); }
    if (TYPE == 2) return ray_metaballs<NB>(lr, pp.fc.elapsed_time, pp.metaball_max_steps, thit, n, work);
    return ray_sdf<(TYPE >= 3 ? TYPE - 3 : 0)>(lr, mat.step_scale, pp.sdf_max_steps, thit, n, work, pp.fractal_iterations);
}

template <int TYPE, int NB>
__global__ void __launch_bounds__(kBlock) k_intersect(const __grid_constant__ SceneDev sc, const __grid_constant__ PassParams pp,
                                                      const __grid_constant__ WaveBuffers wb, int wave) {
    WaveCtrl* ctrl = &wb.ctrl[wave];
    const int pq = wave & 1;
    const uint32_t count = min(ctrl->cand_count[TYPE], pp.cand_capacity);
    __shared__ uint32_t s_base;
    uint32_t work = 0;
    for (;;) {
        if (threadIdx.x == 0) s_base = atomicAdd(&ctrl->cursor[TYPE], (uint32_t)kBlock);
        __syncthreads();
        const uint32_t base = s_base;
        __syncthreads();
        if (base >= count) break;
        const uint32_t c = base + threadIdx.x;
        if (c >= count) continue;
        const uint2 cand = wb.cand[TYPE][c];
        const uint32_t r = cand.x, prim = cand.y;
        const float4 rd = wb.ray_d[pq][r];
        const uint32_t meta = __float_as_uint(rd.w);
        float4 res = make_float4(0.0f, 0.0f, 0.0f, -1.0f);
        if (!(meta_shadow(meta) && wb.hitkey[r] != kKeyMiss)) { // shadow ray already occluded: skip
            const float4 ro = wb.ray_o[pq][r];
            const PrimRecord pr = sc.prims[prim];
            const float* pf = sc.per_frame + 32 * (size_t)pr.instance_index;
            const InstXform ix = inst_xform(sc, prim);
            const V3 oo = xform_point_3x4(ix.w2o, ld3(ro)), dd = xform_vector_3x4(ix.w2o, ld3(rd));
            LocalRay lr;
            lr.o = mul_point_rowvec(oo, pf + 16);
            lr.d = mul_vector_rowvec(dd, pf + 16);
            lr.tmin = 0.0f;
            lr.tcur = wb.ray_tcur[r];
            float thit = 0.0f;
            V3 n = v3s(0.0f);
            const uint32_t work0 = work;
            const bool hit = run_test<TYPE, NB>(lr, pr.mat, pp, thit, n, work);
            { // per-candidate cost statistics (this one-candidate-per-thread kernel is the profiling path)
                const uint32_t dw = work - work0;
                atomicMax(&ctrl->max_evals[TYPE], dw);
                atomicAdd(&ctrl->evals_hist[TYPE][dw ? min(11, 32 - __clz(dw)) : 0], 1u);
            }
            if (hit && thit >= lr.tmin && thit <= lr.tcur) {
                // attr.normal = mul(n, (float3x3)local_to_blas); normalize(mul((float3x3)ObjectToWorld3x4(), attr.normal))
                const V3 nb = mul_vector_rowvec(n, pf);
                const V3 nw = normalize(xform_vector_3x4(ix.o2w, nb));
                atomicMin(&wb.hitkey[r], pack_key(thit, prim));
                res = make_float4(nw.x, nw.y, nw.z, fabsf(thit));
            }
        }
        wb.cand_res[TYPE][c] = res;
    }
    // work statistics: one atomic per warp
    for (int off = 16; off; off >>= 1) work += __shfl_xor_sync(kFull, work, off);
    if ((threadIdx.x & 31) == 0 && work) atomicAdd(&ctrl->work[TYPE], work);
}

// ---------------------------------------------------------------------------------------------
// resolve: the candidate whose (t, primitive) equals the ray's winning key publishes its normal.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) k_resolve(const __grid_constant__ PassParams pp, const __grid_constant__ WaveBuffers wb, int wave,
                                                    uint32_t type_mask) {
    const WaveCtrl* ctrl = &wb.ctrl[wave];
    { // one slice of the grid (blockIdx.y) per primitive type: the ten queues are resolved side by side
        const int type = (int)blockIdx.y;
        if (!((type_mask >> type) & 1u)) return;
        const uint32_t count = min(ctrl->cand_count[type], pp.cand_capacity);
        // a chain of three dependent loads per candidate and nothing to compute: kResolveIlp candidates per thread and trip,
        // each stage of the chain issued for all of them before the next one is waited for
        constexpr int kResolveIlp = 4;
        const uint32_t stride = gridDim.x * blockDim.x;
        for (uint32_t c0 = blockIdx.x * blockDim.x + threadIdx.x; c0 < count; c0 += kResolveIlp * stride) {
            float4 res[kResolveIlp];
            uint2 cand[kResolveIlp];
            unsigned long long key[kResolveIlp];
#pragma unroll
            for (int i = 0; i < kResolveIlp; i++) {
                const uint32_t c = c0 + (uint32_t)i * stride;
                res[i] = c < count ? wb.cand_res[type][c] : make_float4(0.0f, 0.0f, 0.0f, -1.0f);
            }
#pragma unroll
            for (int i = 0; i < kResolveIlp; i++)
                if (!(res[i].w < 0.0f)) cand[i] = wb.cand[type][c0 + (uint32_t)i * stride];
#pragma unroll
            for (int i = 0; i < kResolveIlp; i++)
                if (!(res[i].w < 0.0f)) key[i] = wb.hitkey[cand[i].x];
#pragma unroll
            for (int i = 0; i < kResolveIlp; i++)
                if (!(res[i].w < 0.0f) && key[i] == pack_key(res[i].w, cand[i].y)) wb.hit_n[cand[i].x] = res[i];
        }
    }
}

// ---------------------------------------------------------------------------------------------
// shade_spawn: per ray of wave `wave`: record the hit for the final fold and emit the next
// wave's rays -- the TraceRay calls inside MyClosestHitShader_* (src/Raytracing.hlsl:189-207,
// 226-242) under the depth rule of TraceRadianceRay / TraceShadowRayAndReportIfHit (:89,:117).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) k_shade_spawn(const __grid_constant__ SceneDev sc, const __grid_constant__ PassParams pp,
                                                        const __grid_constant__ WaveBuffers wb, int wave) {
    WaveCtrl* ctrl = &wb.ctrl[wave];
    WaveCtrl* next = &wb.ctrl[wave + 1];
    const int pq = wave & 1, nq = pq ^ 1;
    const uint32_t n_rays = (wave == 1) ? pp.n_pixels : min(ctrl->n_rays, pp.ray_capacity);
    const uint32_t n_round = (n_rays + (uint32_t)kBlock - 1u) & ~((uint32_t)kBlock - 1u); // whole blocks: the loop has block-wide barriers
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __shared__ uint32_t s_total[2][kBlock / 32], s_base[2]; // double-buffered by trip parity: two barriers a trip are enough
    uint32_t cnt_radiance = 0, cnt_shadow = 0, trip = 0;
    for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n_round; r += gridDim.x * blockDim.x) {
        bool spawn_shadow = false, spawn_refl = false;
        V3 hit_pos = v3s(0.0f), sdir = v3s(0.0f), rdir = v3s(0.0f);
        uint32_t pixel = 0, level = 0;
        if (r < n_rays) {
            const float4 rd = wb.ray_d[pq][r];
            const uint32_t meta = __float_as_uint(rd.w);
            pixel = meta_pixel(meta);
            level = meta_level(meta);
            const unsigned long long key = wb.hitkey[r];
            const size_t li = (size_t)(level - 1u) * pp.n_pixels + pixel;
            if (meta_shadow(meta)) {
                cnt_shadow++;
                wb.lvl_shadow[li] = (key != kKeyMiss) ? 1 : 0;
            } else {
                cnt_radiance++;
                const uint32_t code = (uint32_t)(key & 0xFFFFFFFFull);
                wb.lvl_code[li] = (key == kKeyMiss) ? kCodeMiss : code;
                if (key != kKeyMiss) {
                    const float t = __uint_as_float((uint32_t)(key >> 32));
                    const float4 ro = wb.ray_o[pq][r];
                    const V3 o = ld3(ro), d = ld3(rd);
                    V3 n;
                    float refl;
                    if (code & kCodePlane) {
                        const uint32_t ft = code & ~kCodePlane; // triangle instance * n_tris + triangle
                        const float* tn = sc.tri_nrm + 3 * (sc.n_tri_insts > 1u ? ft % sc.n_tris : ft);
                        n = v3(tn[0], tn[1], tn[2]); // the vertex normal as it is: the shader applies no object-to-world (src/Raytracing.hlsl:184)
                        refl = (sc.n_tri_insts > 1u ? sc.plane_mats[ft / sc.n_tris] : sc.plane_mat).reflectance_coefficient;
                    } else {
                        n = ld3(wb.hit_n[r]);
                        refl = sc.prims[code].mat.reflectance_coefficient;
                    }
                    wb.lvl_o_t[li] = make_float4(o.x, o.y, o.z, t);
                    wb.lvl_d[li] = make_float4(d.x, d.y, d.z, 0.0f);
                    wb.lvl_n[li] = make_float4(n.x, n.y, n.z, 0.0f);
                    if (level < pp.max_depth) {
                        hit_pos = o + t * d; // HitWorldPosition()
                        const V3 lp = v3(pp.fc.light_position[0], pp.fc.light_position[1], pp.fc.light_position[2]);
                        sdir = normalize(lp - hit_pos);
                        spawn_shadow = true;
                        if (refl > 0.001f) { rdir = reflect(d, n); spawn_refl = true; }
                    }
                }
            }
        }
        // ballot/popc compaction into the next wave's queue.  ONE reservation per block and trip: the counter is a single
        // address, and one atomic per warp (a million of them for an 8K frame, serialised in L2 at ~2 ns each) was what this
        // kernel spent most of its time on.
        const unsigned bs = __ballot_sync(kFull, spawn_shadow), br = __ballot_sync(kFull, spawn_refl);
        const uint32_t total = (uint32_t)(__popc(bs) + __popc(br));
        const uint32_t pb = trip++ & 1u;
        if (lane == 0) s_total[pb][warp] = total;
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t sum = 0;
#pragma unroll
            for (int i = 0; i < kBlock / 32; i++) sum += s_total[pb][i];
            s_base[pb] = sum ? atomicAdd(&next->n_rays, sum) : 0u;
        }
        __syncthreads();
        if (total) {
            uint32_t base = s_base[pb];
            for (int i = 0; i < warp; i++) base += s_total[pb][i];
            const unsigned lt = (1u << lane) - 1u;
            if (spawn_shadow) {
                const uint32_t s = base + (uint32_t)__popc(bs & lt);
                if (s < pp.ray_capacity) {
                    wb.ray_o[nq][s] = make_float4(hit_pos.x, hit_pos.y, hit_pos.z, 0.0f);
                    wb.ray_d[nq][s] = make_float4(sdir.x, sdir.y, sdir.z, __uint_as_float(make_meta(pixel, level, 1u)));
                }
            }
            if (spawn_refl) {
                const uint32_t s = base + (uint32_t)__popc(bs) + (uint32_t)__popc(br & lt);
                if (s < pp.ray_capacity) {
                    wb.ray_o[nq][s] = make_float4(hit_pos.x, hit_pos.y, hit_pos.z, 0.0f);
                    wb.ray_d[nq][s] = make_float4(rdir.x, rdir.y, rdir.z, __uint_as_float(make_meta(pixel, level + 1u, 0u)));
                }
            }
        }
    }
    // ray statistics (one atomic per warp)
    for (int off = 16; off; off >>= 1) {
        cnt_radiance += __shfl_xor_sync(kFull, cnt_radiance, off);
        cnt_shadow += __shfl_xor_sync(kFull, cnt_shadow, off);
    }
    if (lane == 0) {
        if (cnt_radiance) atomicAdd(&ctrl->n_radiance, cnt_radiance);
        if (cnt_shadow) atomicAdd(&ctrl->n_shadow, cnt_shadow);
    }
}

// ---------------------------------------------------------------------------------------------
// final_color: fold the per-level hit records back to front -- the order in which the recursive
// closest-hit / miss shaders return -- and store R8G8B8A8_UNORM (src/Raytracing.hlsl:164).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock, 5) k_final_color(const __grid_constant__ SceneDev sc, const __grid_constant__ PassParams pp,
                                                        const __grid_constant__ WaveBuffers wb, const __grid_constant__ OutputBuffers ob) {
    for (uint32_t pixel = blockIdx.x * blockDim.x + threadIdx.x; pixel < pp.n_pixels; pixel += gridDim.x * blockDim.x) {
        const uint32_t tx = pixel % pp.tile_w, ty = pixel / pp.tile_w;
        const uint32_t px = pp.tile_x + tx, py = pass_global_row(pp, ty);
        V4 color = v4s(0.0f); // TraceRadianceRay beyond the depth limit returns 0 (src/Raytracing.hlsl:89-92)
        for (int level = (int)pp.max_depth; level >= 1; level--) {
            const size_t li = (size_t)(level - 1) * pp.n_pixels + pixel;
            const uint32_t code = wb.lvl_code[li];
            if (code == kCodeNone) continue;
            if (code == kCodeMiss) { color = background_color(); continue; } // MyMissShader
            const float4 ot = wb.lvl_o_t[li], dd = wb.lvl_d[li], nn = wb.lvl_n[li];
            const bool is_tri = (code & kCodePlane) != 0u;
            const Material mat = is_tri ? (sc.n_tri_insts > 1u ? sc.plane_mats[(code & ~kCodePlane) / sc.n_tris] : sc.plane_mat) : sc.prims[code].mat;
            const bool in_shadow = ((uint32_t)level < pp.max_depth) ? (wb.lvl_shadow[li] != 0) : false;
            color = closest_hit_color(pp.fc, ld3(ot), ld3(dd), ot.w, ld3(nn), mat, in_shadow, color, is_tri, px, py);
        }
        const uint32_t ox = ob.out_x0 + tx, oy = ob.out_y0 + ty;
        *reinterpret_cast<uint32_t*>(ob.rgba8 + (size_t)oy * ob.pitch + (size_t)ox * 4u) = pack_unorm8(color);
        const size_t oi = (size_t)oy * ob.out_w + ox;
        if (ob.rgba_f32) ob.rgba_f32[oi] = make_float4(color.x, color.y, color.z, color.w);
        if (ob.hit_ids) {
            const uint32_t code = wb.lvl_code[pixel];
            ob.hit_ids[oi] = (code == kCodeMiss || code == kCodeNone) ? 0xFFFFFFFFu : (code & kCodePlane) ? (0x00000000u | (code & ~kCodePlane)) : (0x10000000u | code);
        }
    }
}

} // namespace
} // namespace gprt

#include "walk.cuh"  // k_walk*: the LBVH walk (4-wide nodes, top levels and stack in shared memory, per-lane refill)
#include "march.cuh" // queue plumbing, analytic tests (same anonymous namespace)
#include "march_coop.cuh" // k_intersect_coop: warp-cooperative sphere tracer (the production path)
#include "megakernel.cuh" // k_megakernel: one thread per pixel, no queues (the comparison point of SURVEY 7.3, GPRT_FLAG_MEGAKERNEL)
#include "march_lane.cuh" // k_intersect_lane: per-lane sphere tracer (measured alternative, GPRT_ISECT_VARIANT=2)

namespace gprt {
namespace {

// ---------------------------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------------------------
int g_sm_count = 0;
int sm_count() {
    if (!g_sm_count) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (g_sm_count <= 0) g_sm_count = 148;
    }
    return g_sm_count;
}
int grid_for(uint32_t n_items, int blocks_per_sm) {
    const long long need = ((long long)n_items + kBlock - 1) / kBlock;
    const long long cap = (long long)sm_count() * blocks_per_sm;
    return (int)(need < 1 ? 1 : (need < cap ? need : cap));
}

// Blocks of `kernel` one SM holds at once: grid-stride and cursor-fed kernels are launched as exactly one resident
// wave of blocks (a partial second wave of a fixed-stride grid left SMs idle 40 % of the walk kernel's time).
int resident_blocks(const void* kernel) {
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kBlock, 0) != cudaSuccess || n < 1) n = 4;
    return n;
}

// Dynamic shared memory of a walk block: the staged top array + its threads' stacks.
size_t walk_smem(const SceneDev& sc, int threads, int depth) { return (size_t)sc.n_top * 8 * sizeof(float4) + (size_t)threads * depth * sizeof(uint32_t); }
template <class K> int walk_grid(K kernel, int threads, size_t smem) {
    // one resident wave of blocks (the kernels are persistent); cached per kernel for the last shared-memory size asked
    thread_local K cached_kernel = nullptr;
    thread_local size_t cached_smem = 0;
    thread_local int cached_grid = 0;
    if (cached_kernel == kernel && cached_smem == smem && cached_grid) return cached_grid;
    cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem);
    if (per_sm < 1) per_sm = 1;
    cached_kernel = kernel; cached_smem = smem; cached_grid = sm_count() * per_sm;
    return cached_grid;
}

template <bool MULTI>
int launch_traverse_t(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, int wave, cudaStream_t s) {
    const uint32_t upper = (wave == 1) ? pp.n_pixels : pp.ray_capacity;
    if (pp.brute_force || sc.root_ref < 0) { // no tree to walk: every AABB (test flag; scenes of one primitive)
        static const int per_sm_primary = resident_blocks((const void*)k_traverse_brute<true, MULTI>), per_sm_secondary = resident_blocks((const void*)k_traverse_brute<false, MULTI>);
        if (wave == 1) k_traverse_brute<true, MULTI><<<grid_for(upper, per_sm_primary), kBlock, 0, s>>>(sc, pp, wb, wave);
        else k_traverse_brute<false, MULTI><<<grid_for(upper, per_sm_secondary), kBlock, 0, s>>>(sc, pp, wb, wave);
        return 1;
    }
    if (!pp.spill_budget) {
        // small scenes: no hand-over rounds (three empty launches per wave would cost more than the walk); the deep-stack
        // kernel walks every ray to its end
        const size_t smem = walk_smem(sc, kWalkBlockLast, kStackDepthLast);
        if (wave == 1) { const int g = walk_grid(k_walk_deep<true, MULTI>, kWalkBlockLast, smem); k_walk_deep<true, MULTI><<<g, kWalkBlockLast, smem, s>>>(sc, pp, wb, wave); }
        else { const int g = walk_grid(k_walk_deep<false, MULTI>, kWalkBlockLast, smem); k_walk_deep<false, MULTI><<<g, kWalkBlockLast, smem, s>>>(sc, pp, wb, wave); }
        return 1;
    }
    const size_t smem = walk_smem(sc, kWalkBlock, kStackDepth);
    if (wave == 1) { const int g = walk_grid(k_walk<true, MULTI>, kWalkBlock, smem); k_walk<true, MULTI><<<g, kWalkBlock, smem, s>>>(sc, pp, wb, wave); }
    else { const int g = walk_grid(k_walk<false, MULTI>, kWalkBlock, smem); k_walk<false, MULTI><<<g, kWalkBlock, smem, s>>>(sc, pp, wb, wave); }
    const int g_round = walk_grid(k_walk_round<MULTI>, kWalkBlock, smem);
    for (int round = 1; round < kSpillRounds; round++) k_walk_round<MULTI><<<g_round, kWalkBlock, smem, s>>>(sc, pp, wb, wave, round);
    const size_t smem_last = walk_smem(sc, kWalkBlockLast, kStackDepthLast);
    k_walk_last<MULTI><<<walk_grid(k_walk_last<MULTI>, kWalkBlockLast, smem_last), kWalkBlockLast, smem_last, s>>>(sc, pp, wb, wave, kSpillRounds);
    return 1 + kSpillRounds;
}
int launch_traverse(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, int wave, void* stream) {
    // the reference's TLAS (one instance per BLAS) runs the kernels compiled without the instance loops
    if (sc.n_tri_insts > 1u || sc.n_aabb_insts > 1u) return launch_traverse_t<true>(sc, pp, wb, wave, (cudaStream_t)stream);
    return launch_traverse_t<false>(sc, pp, wb, wave, (cudaStream_t)stream);
}

// GPRT_ISECT_VARIANT: 0 (default) the warp-cooperative k_intersect_coop; 2 the per-lane k_intersect_lane (A/B, march_lane.cuh);
// -1 (profiling aid, with GPRT_PER_TYPE_LAUNCHES=1) the one-candidate-per-thread kernels, which also fill the per-candidate
// cost statistics of gprt_get_wave_stats.
int isect_variant() {
    static int v = -1000;
    if (v == -1000) { const char* e = getenv("GPRT_ISECT_VARIANT"); v = e ? atoi(e) : 0; }
    return v;
}
bool stats_kernels() { return isect_variant() == -1; }
// GPRT_ISECT_BLOCKS (tuning aid): resident blocks per SM of the persistent intersection kernel (default: what fits, 4)
int isect_blocks_limit() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("GPRT_ISECT_BLOCKS"); v = e ? atoi(e) : 0; }
    return v;
}

template <class K> int persistent_grid(K kernel, int block = kCoopBlock) {
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, block, 0);
    if (per_sm < 1) per_sm = 1;
    return sm_count() * per_sm;
}

// one launch per primitive type (GPRT_PER_TYPE_LAUNCHES=1: per-type device times)
template <int TYPE, int NB> void launch_intersect_t(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, int wave, cudaStream_t s) {
    if constexpr (TYPE >= 3) {
        if (!stats_kernels()) {
            static const int grid = persistent_grid(k_coop_sdf<TYPE - 3, kRefillMin>);
            k_coop_sdf<TYPE - 3, kRefillMin><<<grid, kCoopBlock, 0, s>>>(sc, pp, wb, wave);
            return;
        }
    }
    if constexpr (TYPE == 2) {
        if (!stats_kernels()) {
            static const int grid = persistent_grid(k_scan_metaballs<NB>);
            k_scan_metaballs<NB><<<grid, kCoopBlock, 0, s>>>(sc, pp, wb, wave);
            return;
        }
    }
    static const int grid = persistent_grid(k_intersect<TYPE, NB>, kBlock);
    k_intersect<TYPE, NB><<<grid, kBlock, 0, s>>>(sc, pp, wb, wave);
}

int launch_intersect(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, int wave, int type_id, int metaballs, void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    switch (type_id) {
    case 0: launch_intersect_t<0, 3>(sc, pp, wb, wave, s); break;
    case 1: launch_intersect_t<1, 3>(sc, pp, wb, wave, s); break;
    case 2:
        if (metaballs == 5) launch_intersect_t<2, 5>(sc, pp, wb, wave, s);
        else launch_intersect_t<2, 3>(sc, pp, wb, wave, s);
        break;
    case 3: launch_intersect_t<3, 3>(sc, pp, wb, wave, s); break;
    case 4: launch_intersect_t<4, 3>(sc, pp, wb, wave, s); break;
    case 5: launch_intersect_t<5, 3>(sc, pp, wb, wave, s); break;
    case 6: launch_intersect_t<6, 3>(sc, pp, wb, wave, s); break;
    case 7: launch_intersect_t<7, 3>(sc, pp, wb, wave, s); break;
    case 8: launch_intersect_t<8, 3>(sc, pp, wb, wave, s); break;
    case 9: launch_intersect_t<9, 3>(sc, pp, wb, wave, s); break;
    default: return 0;
    }
    return 1;
}

// the production path: every primitive type in one persistent launch per wave
template <int NB> void launch_coop_t(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, int wave, uint32_t type_mask, cudaStream_t s, int blocks_per_sm) {
    static const int full = persistent_grid(k_intersect_coop<NB, kRefillMin>);
    int lim = isect_blocks_limit(); // GPRT_ISECT_BLOCKS (tuning aid)
    if (blocks_per_sm > 0 && (lim <= 0 || blocks_per_sm < lim)) lim = blocks_per_sm;
    const int grid = (lim > 0 && lim * sm_count() < full) ? lim * sm_count() : full;
    k_intersect_coop<NB, kRefillMin><<<grid, kCoopBlock, 0, s>>>(sc, pp, wb, wave, type_mask);
}
template <int NB> void launch_lane_t(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, int wave, uint32_t type_mask, cudaStream_t s) {
    static const int per_sm = [] {
        int n = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_intersect_lane<NB>, kCoopBlock, 0);
        if (n < 1) n = 1;
        const int lim = isect_blocks_limit();
        return (lim > 0 && lim < n) ? lim : n;
    }();
    k_intersect_lane<NB><<<sm_count() * per_sm, kCoopBlock, 0, s>>>(sc, pp, wb, wave, type_mask);
}
int launch_intersect_all(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, int wave, uint32_t type_mask, int metaballs, void* stream, int blocks_per_sm) {
    if (isect_variant() == 2) {
        if (metaballs == 5) launch_lane_t<5>(sc, pp, wb, wave, type_mask, (cudaStream_t)stream);
        else launch_lane_t<3>(sc, pp, wb, wave, type_mask, (cudaStream_t)stream);
        return 1;
    }
    if (metaballs == 5) launch_coop_t<5>(sc, pp, wb, wave, type_mask, (cudaStream_t)stream, blocks_per_sm);
    else launch_coop_t<3>(sc, pp, wb, wave, type_mask, (cudaStream_t)stream, blocks_per_sm);
    return 1;
}
int launch_resolve(const SceneDev&, const PassParams& pp, const WaveBuffers& wb, int wave, uint32_t type_mask, void* stream) {
    k_resolve<<<dim3((unsigned)sm_count() * 4u, (unsigned)kNumTypes), kBlock, 0, (cudaStream_t)stream>>>(pp, wb, wave, type_mask);
    return 1;
}

int launch_shade_spawn(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, int wave, void* stream) {
    const uint32_t upper = (wave == 1) ? pp.n_pixels : pp.ray_capacity;
    static const int per_sm = resident_blocks((const void*)k_shade_spawn);
    k_shade_spawn<<<grid_for(upper, per_sm), kBlock, 0, (cudaStream_t)stream>>>(sc, pp, wb, wave);
    return 1;
}

int launch_megakernel(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, const OutputBuffers& ob, int metaballs, void* stream) {
    const int grid = (int)((pp.n_pixels + 127u) / 128u);
    if (metaballs == 5) k_megakernel<5><<<grid, 128, 0, (cudaStream_t)stream>>>(sc, pp, wb, ob);
    else k_megakernel<3><<<grid, 128, 0, (cudaStream_t)stream>>>(sc, pp, wb, ob);
    return 1;
}

int launch_final_color(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, const OutputBuffers& ob, void* stream) {
    static const int per_sm = resident_blocks((const void*)k_final_color);
    k_final_color<<<grid_for(pp.n_pixels, per_sm), kBlock, 0, (cudaStream_t)stream>>>(sc, pp, wb, ob);
    return 1;
}

} // namespace

#if defined(GPRT_KERNELS_FMA)
const KernelTable* kernels_fma() {
    static const KernelTable t = {launch_traverse, launch_intersect, launch_intersect_all, launch_resolve, launch_shade_spawn, launch_final_color, launch_megakernel, "fma"};
    return &t;
}
#elif defined(GPRT_KERNELS_FAST)
const KernelTable* kernels_fast() {
    static const KernelTable t = {launch_traverse, launch_intersect, launch_intersect_all, launch_resolve, launch_shade_spawn, launch_final_color, launch_megakernel, "fast"};
    return &t;
}
#else
const KernelTable* kernels_parity() {
    static const KernelTable t = {launch_traverse, launch_intersect, launch_intersect_all, launch_resolve, launch_shade_spawn, launch_final_color, launch_megakernel, "parity"};
    return &t;
}
#endif

} // namespace gprt
