// march.cuh -- warp-coherent ray marchers with lane refill (included by kernels.cu).
//
// The reference's sphere tracer (RaySignedDistancePrimitiveTest, src/SignedDistancePrimitives.hlsli:
// 287-319) and metaball marcher (RayMetaballsIntersectionTest, src/VolumetricPrimitives.hlsli:
// 151-202) take anything between 1 and 512 (128) iterations per ray.  One-candidate-per-thread
// leaves most lanes of a warp idle while the slowest lane finishes.  Here every lane is a small
// state machine and the expensive field evaluation is a single, convergent call site per loop
// trip:
//   mode 0        march:   p = o + t d, evaluate the field, test the hit criterion
//   mode 1..TAPS  normal:  evaluate the field at the k-th finite-difference tap (the reference
//                 computes the normal inside the loop because IsAValidHit culls on it)
// A lane that finishes (hit, cull-and-continue exhausted, or ray end) publishes its result and is
// refilled from a warp-private chunk of the type's candidate queue; chunks come from one global
// atomic per kChunk candidates.  Arithmetic (and so the parity build's bits) is unchanged.
#pragma once

namespace gprt {
namespace {

constexpr uint32_t kNoCand = 0xFFFFFFFFu;
constexpr uint32_t kChunk = 32; // one candidate per lane per grab: the queues are short relative to the machine

struct WarpChunk {
    uint32_t next = 0, end = 0;
    bool exhausted = false;
};

// Hands one queue index to every lane with need == true while the queue lasts (warp-uniform control flow).
__device__ __forceinline__ uint32_t warp_take(WarpChunk& q, bool need, uint32_t* cursor, uint32_t count, int lane) {
    uint32_t idx = kNoCand;
    unsigned m = __ballot_sync(kFull, need);
    while (m) {
        if (q.next >= q.end) {
            if (q.exhausted) break;
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(cursor, kChunk);
            base = __shfl_sync(kFull, base, 0);
            if (base >= count) { q.exhausted = true; break; }
            q.next = base;
            q.end = min(base + kChunk, count);
        }
        const uint32_t avail = q.end - q.next;
        const uint32_t rank = (uint32_t)__popc(m & ((1u << lane) - 1u));
        if (need && idx == kNoCand && rank < avail) idx = q.next + rank;
        q.next += min(avail, (uint32_t)__popc(m));
        m = __ballot_sync(kFull, need && idx == kNoCand);
    }
    return idx;
}

// Refill policy: fetching a candidate is a chain of dependent global loads that stalls the whole
// warp, so idle lanes are only refilled once at least RMIN of them wait (or none is active).
template <int RMIN> __device__ __forceinline__ bool refill_now(bool active) {
    const int idle = __popc(__ballot_sync(kFull, !active));
    return idle >= RMIN || idle == 32;
}

struct CandState {
    uint32_t c, r, prim;
    const float* pf; // per-frame transforms of the primitive: local->BLAS (16 floats), BLAS->local (16 floats)
    LocalRay lr;
    float step_scale;
};

// GetRayInAABBPrimitiveLocalSpace (src/Raytracing.hlsl:277-287).  Returns false (after writing a
// "no hit" result) when the candidate belongs to a shadow ray that is already occluded.
template <int TYPE>
__device__ __forceinline__ bool fetch_candidate(const SceneDev& sc, const WaveBuffers& wb, int pq, uint32_t c, CandState& s) {
    const uint2 cand = wb.cand[TYPE][c];
    s.c = c; s.r = cand.x; s.prim = cand.y;
    const float4 rd = wb.ray_d[pq][s.r];
    const uint32_t meta = __float_as_uint(rd.w);
    if (meta_shadow(meta) && wb.hitkey[s.r] != kKeyMiss) {
        wb.cand_res[TYPE][c] = make_float4(0.0f, 0.0f, 0.0f, -1.0f);
        return false;
    }
    const float4 ro = wb.ray_o[pq][s.r];
    const PrimRecord* pr = sc.prims + s.prim;
    s.step_scale = pr->mat.step_scale;
    s.pf = sc.per_frame + 32 * (size_t)pr->instance_index;
    const V3 oo = xform_point_3x4(sc.aabb_w2o, ld3(ro)), dd = xform_vector_3x4(sc.aabb_w2o, ld3(rd));
    s.lr.o = mul_point_rowvec(oo, s.pf + 16);
    s.lr.d = mul_vector_rowvec(dd, s.pf + 16);
    s.lr.tmin = 0.0f;
    s.lr.tcur = wb.ray_tcur[s.r];
    return true;
}

// ReportHit + the normal transform of the intersection shaders (src/Raytracing.hlsl:297-304)
template <int TYPE>
__device__ __forceinline__ void publish_hit(const SceneDev& sc, const WaveBuffers& wb, const CandState& s, float thit, V3 n_local) {
    if (thit >= s.lr.tmin && thit <= s.lr.tcur) {
        const V3 nb = mul_vector_rowvec(n_local, s.pf);
        const V3 nw = normalize(xform_vector_3x4(sc.aabb_o2w, nb));
        atomicMin(&wb.hitkey[s.r], pack_key(thit, s.prim));
        wb.cand_res[TYPE][s.c] = make_float4(nw.x, nw.y, nw.z, fabsf(thit));
    } else {
        wb.cand_res[TYPE][s.c] = make_float4(0.0f, 0.0f, 0.0f, -1.0f);
    }
}
template <int TYPE> __device__ __forceinline__ void publish_miss(const WaveBuffers& wb, const CandState& s) {
    wb.cand_res[TYPE][s.c] = make_float4(0.0f, 0.0f, 0.0f, -1.0f);
}

// ---------------------------------------------------------------------------------------------
// Sphere tracer for SDF type SDF (0..6); queue / kernel type id is SDF + 3.
// ---------------------------------------------------------------------------------------------
template <int SDF, int RMIN>
__device__ __forceinline__ void march_sdf_queue(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, WaveCtrl* ctrl, int pq, int lane) {
    constexpr int TYPE = SDF + 3;
    constexpr int kTrips = (SDF == 6) ? 4 : 8; // loop trips between refill checks
    const uint32_t count = min(ctrl->cand_count[TYPE], pp.cand_capacity);
    const uint32_t max_steps = pp.sdf_max_steps;
    const float threshold = 0.0001f;
    // tetrahedral taps of sdCalculateNormal: e.xyy, e.yyx, e.yxy, e.xxx
    const float ex = 1.0f * 0.5773f * 0.0001f, ey = -1.0f * 0.5773f * 0.0001f;

    WarpChunk q;
    CandState s;
    bool active = false;
    int mode = 0;
    uint32_t step = 0, work = 0;
    float t = 0.0f, dist_hit = 0.0f;
    V3 pos = v3s(0.0f), acc = v3s(0.0f);

    for (;;) {
        const uint32_t idx = warp_take(q, refill_now<RMIN>(active) && !active, &ctrl->cursor[TYPE], count, lane);
        if (idx != kNoCand && fetch_candidate<TYPE>(sc, wb, pq, idx, s)) {
            active = true; mode = 0; step = 0; t = s.lr.tmin;
        }
        if (!__any_sync(kFull, active)) {
            if (q.exhausted && q.next >= q.end) break;
            continue;
        }
#pragma unroll 1
        for (int trip = 0; trip < kTrips; trip++) {
            // while (i++ < MaxSteps && t <= RayTCurrent())
            if (active && mode == 0 && !(step < max_steps && t <= s.lr.tcur)) { publish_miss<TYPE>(wb, s); active = false; }
            if (active) {
                V3 tap = v3s(0.0f);
                if (mode == 1) tap = v3(ex, ey, ey);
                else if (mode == 2) tap = v3(ey, ey, ex);
                else if (mode == 3) tap = v3(ey, ex, ey);
                else if (mode == 4) tap = v3(ex, ex, ex);
                const V3 p = (mode == 0) ? s.lr.o + t * s.lr.d : pos + tap;
                const float dist = sdf_distance<SDF>(p); // the one convergent evaluation site
                work++;
                if (mode == 0) {
                    step++;
                    if (dist <= threshold * t) { mode = 1; pos = p; dist_hit = dist; }
                    else t += s.step_scale * dist;
                } else {
                    const V3 term = tap * v3s(dist);
                    acc = (mode == 1) ? term : acc + term;
                    if (mode == 4) {
                        const V3 n = normalize(acc);
                        if (is_valid_hit(s.lr, t, n)) { publish_hit<TYPE>(sc, wb, s, t, n); active = false; }
                        else { t += s.step_scale * dist_hit; mode = 0; }
                    } else {
                        mode++;
                    }
                }
            }
        }
    }
    for (int off = 16; off; off >>= 1) work += __shfl_xor_sync(kFull, work, off);
    if (lane == 0 && work) atomicAdd(&ctrl->work[TYPE], work);
}

// ---------------------------------------------------------------------------------------------
// Metaball marcher (type id 2): fixed step, 6 central-difference taps for the normal.
// ---------------------------------------------------------------------------------------------
template <int NB, int RMIN>
__device__ __forceinline__ void march_metaballs_queue(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, WaveCtrl* ctrl, int pq, int lane) {
    constexpr int TYPE = 2;
    constexpr int kTrips = 8;
    const uint32_t count = min(ctrl->cand_count[TYPE], pp.cand_capacity);
    const uint32_t max_steps = pp.metaball_max_steps;
    const float e = 0.5773f * 0.00001f;

    WarpChunk q;
    CandState s;
    Metaballs<NB> mb;
    bool active = false;
    int mode = 0;
    uint32_t step = 0, work = 0;
    float t = 0.0f, t_step = 0.0f, prev = 0.0f;
    V3 pos = v3s(0.0f), nrm = v3s(0.0f);

    for (;;) {
        const uint32_t idx = warp_take(q, refill_now<RMIN>(active) && !active, &ctrl->cursor[TYPE], count, lane);
        if (idx != kNoCand && fetch_candidate<TYPE>(sc, wb, pq, idx, s)) {
            init_metaballs<NB>(mb, pp.fc.elapsed_time);
            float tmin, tmax;
            find_intersecting_metaballs<NB>(s.lr, mb, tmin, tmax);
            if (tmin == __builtin_inff()) {
                publish_miss<TYPE>(wb, s); // no bounding sphere crossed: the field is 0 at every sample
            } else {
                active = true; mode = 0; step = 0; t = tmin;
                t_step = (tmax - tmin) / (float)(max_steps / 1u);
            }
        }
        if (!__any_sync(kFull, active)) {
            if (q.exhausted && q.next >= q.end) break;
            continue;
        }
#pragma unroll 1
        for (int trip = 0; trip < kTrips; trip++) {
            if (active && mode == 0 && !(step < max_steps)) { publish_miss<TYPE>(wb, s); active = false; }
            if (active) {
                V3 tap = v3s(0.0f);
                if (mode == 1) tap = v3(-e, 0, 0);
                else if (mode == 2) tap = v3(e, 0, 0);
                else if (mode == 3) tap = v3(0, -e, 0);
                else if (mode == 4) tap = v3(0, e, 0);
                else if (mode == 5) tap = v3(0, 0, -e);
                else if (mode == 6) tap = v3(0, 0, e);
                const V3 p = (mode == 0) ? s.lr.o + t * s.lr.d : pos + tap;
                const float sum = metaballs_potential<NB>(p, mb); // the one convergent evaluation site
                if (mode == 0) {
                    step++;
                    work++;
                    if (sum >= 0.25f) { mode = 1; pos = p; }
                    else t += t_step;
                } else {
                    if (mode & 1) prev = sum;
                    else if (mode == 2) nrm.x = prev - sum;
                    else if (mode == 4) nrm.y = prev - sum;
                    else nrm.z = prev - sum;
                    if (mode == 6) {
                        const V3 n = normalize(nrm);
                        if (is_valid_hit(s.lr, t, n)) { publish_hit<TYPE>(sc, wb, s, t, n); active = false; }
                        else { t += t_step; mode = 0; }
                    } else {
                        mode++;
                    }
                }
            }
        }
    }
    for (int off = 16; off; off >>= 1) work += __shfl_xor_sync(kFull, work, off);
    if (lane == 0 && work) atomicAdd(&ctrl->work[TYPE], work);
}

// Analytic types (0 box, 1 spheres): one cheap closed-form test per candidate, no state machine.
template <int TYPE>
__device__ __forceinline__ void analytic_queue(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, WaveCtrl* ctrl, int pq, int lane) {
    const uint32_t count = min(ctrl->cand_count[TYPE], pp.cand_capacity);
    WarpChunk q;
    uint32_t work = 0;
    for (;;) {
        const uint32_t idx = warp_take(q, true, &ctrl->cursor[TYPE], count, lane);
        if (idx != kNoCand) {
            CandState s;
            if (fetch_candidate<TYPE>(sc, wb, pq, idx, s)) {
                float thit = 0.0f;
                V3 n = v3s(0.0f);
                work++;
                const bool hit = (TYPE == 0) ? ray_unit_box(s.lr, thit, n) : ray_spheres(s.lr, thit, n);
                if (hit) publish_hit<TYPE>(sc, wb, s, thit, n); else publish_miss<TYPE>(wb, s);
            }
        }
        if (q.exhausted && q.next >= q.end) break;
    }
    for (int off = 16; off; off >>= 1) work += __shfl_xor_sync(kFull, work, off);
    if (lane == 0 && work) atomicAdd(&ctrl->work[TYPE], work);
}

// ---- one kernel per type (profiling mode: per-type device times) -------------------------------
template <int SDF>
__global__ void __launch_bounds__(kBlock) k_march_sdf(const __grid_constant__ SceneDev sc, const __grid_constant__ PassParams pp,
                                                      const __grid_constant__ WaveBuffers wb, int wave) {
    march_sdf_queue<SDF, 1>(sc, pp, wb, &wb.ctrl[wave], wave & 1, threadIdx.x & 31);
}
template <int NB>
__global__ void __launch_bounds__(kBlock) k_march_metaballs(const __grid_constant__ SceneDev sc, const __grid_constant__ PassParams pp,
                                                            const __grid_constant__ WaveBuffers wb, int wave) {
    march_metaballs_queue<NB, 1>(sc, pp, wb, &wb.ctrl[wave], wave & 1, threadIdx.x & 31);
}

// One candidate per lane, run to completion, then the warp takes the next 32 (no state machine).
template <int TYPE, int NB>
__device__ __forceinline__ void simple_queue(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, WaveCtrl* ctrl, int pq, int lane) {
    const uint32_t count = min(ctrl->cand_count[TYPE], pp.cand_capacity);
    WarpChunk q;
    uint32_t work = 0;
    for (;;) {
        const uint32_t idx = warp_take(q, true, &ctrl->cursor[TYPE], count, lane);
        if (idx != kNoCand) {
            CandState s;
            if (fetch_candidate<TYPE>(sc, wb, pq, idx, s)) {
                float thit = 0.0f;
                V3 n = v3s(0.0f);
                Material m;
                m.step_scale = s.step_scale;
                if (run_test<TYPE, NB>(s.lr, m, pp, thit, n, work)) publish_hit<TYPE>(sc, wb, s, thit, n); else publish_miss<TYPE>(wb, s);
            }
        }
        if (q.exhausted && q.next >= q.end) break;
    }
    for (int off = 16; off; off >>= 1) work += __shfl_xor_sync(kFull, work, off);
    if (lane == 0 && work) atomicAdd(&ctrl->work[TYPE], work);
}

// ---- all types in one persistent launch ---------------------------------------------------------
// Every warp works through the type queues in a fixed order, most expensive first (longest
// processing time first): while the last long pyramid rays finish, the other warps are already
// draining the cheaper queues, so no SM idles in a per-type tail and a wave costs one launch
// instead of ten.  RMIN: refill threshold of the marchers; RMIN == 0 selects simple_queue.
template <int SDF, int NB, int RMIN>
__device__ __forceinline__ void sdf_queue(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, WaveCtrl* ctrl, int pq, int lane) {
    if (RMIN == 0) simple_queue<SDF + 3, NB>(sc, pp, wb, ctrl, pq, lane);
    else march_sdf_queue<SDF, (RMIN ? RMIN : 1)>(sc, pp, wb, ctrl, pq, lane);
}
template <int NB, int RMIN>
__global__ void __launch_bounds__(kBlock) k_intersect_all(const __grid_constant__ SceneDev sc, const __grid_constant__ PassParams pp,
                                                          const __grid_constant__ WaveBuffers wb, int wave, uint32_t type_mask) {
    WaveCtrl* ctrl = &wb.ctrl[wave];
    const int pq = wave & 1;
    const int lane = threadIdx.x & 31;
    // the closed-form tests cost next to nothing and may occlude a shadow ray before its marches start
    if (type_mask & (1u << 1)) analytic_queue<1>(sc, pp, wb, ctrl, pq, lane);
    if (type_mask & (1u << 0)) analytic_queue<0>(sc, pp, wb, ctrl, pq, lane);
    if (type_mask & (1u << 9)) sdf_queue<6, NB, RMIN>(sc, pp, wb, ctrl, pq, lane); // fractal pyramid
    if (type_mask & (1u << 7)) sdf_queue<4, NB, RMIN>(sc, pp, wb, ctrl, pq, lane); // cog
    if (type_mask & (1u << 2)) {
        if (RMIN == 0) simple_queue<2, NB>(sc, pp, wb, ctrl, pq, lane);
        else march_metaballs_queue<NB, (RMIN ? RMIN : 1)>(sc, pp, wb, ctrl, pq, lane);
    }
    if (type_mask & (1u << 8)) sdf_queue<5, NB, RMIN>(sc, pp, wb, ctrl, pq, lane); // cylinders
    if (type_mask & (1u << 3)) sdf_queue<0, NB, RMIN>(sc, pp, wb, ctrl, pq, lane); // mini spheres
    if (type_mask & (1u << 4)) sdf_queue<1, NB, RMIN>(sc, pp, wb, ctrl, pq, lane); // intersected round cube
    if (type_mask & (1u << 6)) sdf_queue<3, NB, RMIN>(sc, pp, wb, ctrl, pq, lane); // twisted torus
    if (type_mask & (1u << 5)) sdf_queue<2, NB, RMIN>(sc, pp, wb, ctrl, pq, lane); // square torus
}

} // namespace
} // namespace gprt
