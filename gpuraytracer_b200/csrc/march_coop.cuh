// march_coop.cuh -- warp-cooperative sphere tracer (included by kernels.cu after march.cuh; same
// anonymous namespace).
//
// Measured on the reference scene (tools/wave_stats.py): the mean SDF candidate needs 15-80 field
// evaluations, but in the shadow / reflection waves a handful of rays per type need ~2500.  Those
// rays start ON a surface (TMin = 0), stay inside its 1e-4*t hit shell and leave it back-facing:
// every one of their 512 march steps finds a hit candidate, computes the 4-tap normal, is culled by
// IsAValidHit and advances by the tiny hit distance (RaySignedDistancePrimitiveTest,
// src/SignedDistancePrimitives.hlsli:287-319).  One lane doing that serially is 5 x 512 dependent
// field evaluations -- about 1 ms -- and a wave cannot end before its slowest ray, so the whole
// frame was bound by the critical path of a few dozen rays, not by throughput.
//
// The only true dependency of the sphere tracer is t(i+1) = t(i) + s * f(p(t(i))).  The 4 normal
// taps of a hit candidate and the NEXT march sample (which needs only the hit distance, not the
// normal) are independent of each other.  So each trip of the loop below
//   1. every active lane posts the points it needs into a warp-private shared-memory list:
//      1 point (marching) or 5 points (hit candidate: 4 taps + the speculative next sample),
//   2. all 32 lanes evaluate the posted points, one point per lane per pass (a single convergent
//      call site of the field function, dense lanes whatever the rays' states are),
//   3. the owners read their distances back and advance their state machines.
// A stuck ray now costs 512 trips of ONE evaluation latency instead of 2560, idle lanes of a
// sparse warp do useful work, and the arithmetic per point -- so every bit of the parity build's
// result -- is unchanged.  Lanes are refilled from the type's candidate queue as in march.cuh.
#pragma once

namespace gprt {
namespace {

#ifndef GPRT_COOP_BLOCK
#define GPRT_COOP_BLOCK 256
#endif
constexpr int kCoopBlock = GPRT_COOP_BLOCK; // threads per block of the persistent march kernels
constexpr int kWarpsPerBlock = kCoopBlock / 32;
constexpr int kCoopPts = 32 * 5; // worst case: every lane posts 5 points

struct CoopSmem { // warp-private scratch
    float px[kWarpsPerBlock][kCoopPts], py[kWarpsPerBlock][kCoopPts], pz[kWarpsPerBlock][kCoopPts];
    float dist[kWarpsPerBlock][kCoopPts];
};

// GetDistanceFromSignedDistancePrimitive with a run-time (warp-uniform) primitive type: one copy of
// the march loop serves all seven SDF types; the branch is uniform, so it costs a jump.
__device__ __forceinline__ float sdf_distance_dyn(int sdf, V3 p, uint32_t fractal_iters) {
    switch (sdf) {
    case 0: return sdf_distance<0>(p);
    case 1: return sdf_distance<1>(p);
    case 2: return sdf_distance<2>(p);
    case 3: return sdf_distance<3>(p);
    case 4: return sdf_distance<4>(p);
    case 5: return sdf_distance<5>(p);
    default: return sdf_distance<6>(p, fractal_iters);
    }
}

// One lane's march state.  The candidate's identity (ray slot, primitive, transforms) is re-read
// from the queue entry `c` when the lane publishes a hit, so the state stays small enough to be
// moved between lanes (pool compaction below).
struct MarchLane {
    V3 o, d;        // ray in the primitive's local space
    float tcur;     // RayTCurrent() of the intersection shader (range test of ReportHit / IsAValidHit)
    float t_stop;   // loop bound: min(tcur, closest t another candidate of the ray has committed)
    float t, scale, dist_hit;
    uint32_t c, r;  // queue entry, ray slot
    uint32_t step;
    int mode;       // 0: needs the field at p(t); 1: hit candidate at p(t), needs the 4 taps (+ next sample)
    bool active, shadow;
};

// GetRayInAABBPrimitiveLocalSpace (src/Raytracing.hlsl:277-287) for queue entry c of type id `type`.
// Returns false (after writing a "no hit" result) for a shadow ray that is already occluded.
__device__ __forceinline__ bool fetch_lane(const SceneDev& sc, const WaveBuffers& wb, int pq, int type, uint32_t c, MarchLane& m) {
    const uint2 cand = wb.cand[type][c];
    m.c = c; m.r = cand.x;
    const float4 rd = wb.ray_d[pq][m.r];
    m.shadow = meta_shadow(__float_as_uint(rd.w)) != 0u;
    // the result starts as "no hit": only publish_lane_hit writes again, the retire paths of the march loops write nothing
    wb.cand_res[type][c] = make_float4(0.0f, 0.0f, 0.0f, -1.0f);
    if (m.shadow && wb.hitkey[m.r] != kKeyMiss) return false;
    const float4 ro = wb.ray_o[pq][m.r];
    const PrimRecord* pr = sc.prims + cand.y;
    m.scale = pr->mat.step_scale;
    const float* pf = sc.per_frame + 32 * (size_t)pr->instance_index;
    const InstXform ix = inst_xform(sc, cand.y);
    const V3 oo = xform_point_3x4(ix.w2o, ld3(ro)), dd = xform_vector_3x4(ix.w2o, ld3(rd));
    m.o = mul_point_rowvec(oo, pf + 16);
    m.d = mul_vector_rowvec(dd, pf + 16);
    m.tcur = wb.ray_tcur[m.r];
    m.t_stop = m.tcur;
    m.t = 0.0f; // RayTMin()
    m.step = 0; m.mode = 0; m.dist_hit = 0.0f;
    m.active = true;
    return true;
}

// ReportHit + the normal transform of the intersection shaders (src/Raytracing.hlsl:297-304)
__device__ __forceinline__ void publish_lane_hit(const SceneDev& sc, const WaveBuffers& wb, int type, const MarchLane& m, float thit, V3 n_local) {
    if (thit >= 0.0f && thit <= m.tcur) {
        const uint32_t prim = wb.cand[type][m.c].y;
        const float* pf = sc.per_frame + 32 * (size_t)sc.prims[prim].instance_index;
        const V3 nb = mul_vector_rowvec(n_local, pf);
        const V3 nw = normalize(xform_vector_3x4(inst_xform(sc, prim).o2w, nb));
        atomicMin(&wb.hitkey[m.r], pack_key(thit, prim));
        wb.cand_res[type][m.c] = make_float4(nw.x, nw.y, nw.z, fabsf(thit));
    }
}

// Early out against what other candidates of the same ray have committed meanwhile (the hit key
// lives in L2, hence __ldcg).  A shadow ray needs any hit (ACCEPT_FIRST_HIT_AND_END_SEARCH); a
// radiance ray keeps the closest, so samples beyond the committed t cannot win the atomicMin --
// the same shrinking RayTCurrent() a DXR traversal would show this shader.  The image is unchanged.
__device__ __forceinline__ void poll_committed(const WaveBuffers& wb, int type, MarchLane& m) {
    if (!m.active) return;
    const unsigned long long key = __ldcg(&wb.hitkey[m.r]);
    if (key == kKeyMiss) return;
    if (m.shadow) m.active = false;
    else m.t_stop = fminf(m.t_stop, __uint_as_float((uint32_t)(key >> 32)));
}

// Up to `trips` cooperative trips of the warp (see the header comment).  Returns the number of
// field evaluations the reference's loop would have made for this lane's candidate.
__device__ __forceinline__ uint32_t coop_trips(const int SDF, MarchLane& m, int trips, const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb,
                                               int lane, CoopSmem& sm) {
    const int TYPE = SDF + 3;
    const int warp = threadIdx.x >> 5;
    float* const spx = sm.px[warp];
    float* const spy = sm.py[warp];
    float* const spz = sm.pz[warp];
    float* const sdist = sm.dist[warp];
    const uint32_t max_steps = pp.sdf_max_steps;
    const float threshold = 0.0001f;
    // tetrahedral taps of sdCalculateNormal: e.xyy, e.yyx, e.yxy, e.xxx
    const float ex = 1.0f * 0.5773f * 0.0001f, ey = -1.0f * 0.5773f * 0.0001f;
    const unsigned lt = (1u << lane) - 1u;
    uint32_t evals = 0;
#pragma unroll 1
    for (int trip = 0; trip < trips; trip++) {
        // while (i++ < MaxSteps && t <= RayTCurrent())
        if (m.active && m.mode == 0 && !(m.step < max_steps && m.t <= m.t_stop)) m.active = false;
        const unsigned b1 = __ballot_sync(kFull, m.active && m.mode == 0), b5 = __ballot_sync(kFull, m.active && m.mode == 1);
        if (!(b1 | b5)) break;
        const uint32_t my_off = (uint32_t)__popc(b1 & lt) + 5u * (uint32_t)__popc(b5 & lt);
        const uint32_t total = (uint32_t)__popc(b1) + 5u * (uint32_t)__popc(b5);
        float t_next = 0.0f;
        V3 pos = v3s(0.0f);
        if (m.active) {
            pos = m.o + m.t * m.d;
            if (m.mode == 0) {
                spx[my_off] = pos.x; spy[my_off] = pos.y; spz[my_off] = pos.z;
            } else {
                spx[my_off + 0] = pos.x + ex; spy[my_off + 0] = pos.y + ey; spz[my_off + 0] = pos.z + ey;
                spx[my_off + 1] = pos.x + ey; spy[my_off + 1] = pos.y + ey; spz[my_off + 1] = pos.z + ex;
                spx[my_off + 2] = pos.x + ey; spy[my_off + 2] = pos.y + ex; spz[my_off + 2] = pos.z + ey;
                spx[my_off + 3] = pos.x + ex; spy[my_off + 3] = pos.y + ex; spz[my_off + 3] = pos.z + ex;
                // the sample the loop takes next if this candidate is culled: t += stepScale * distance
                t_next = m.t + m.scale * m.dist_hit;
                const V3 p = m.o + t_next * m.d;
                spx[my_off + 4] = p.x; spy[my_off + 4] = p.y; spz[my_off + 4] = p.z;
            }
        }
        __syncwarp();
        for (uint32_t j = lane; j < total; j += 32u) sdist[j] = sdf_distance_dyn(SDF, v3(spx[j], spy[j], spz[j]), pp.fractal_iterations); // the one evaluation site
        __syncwarp();
        if (m.active) {
            bool have_sample = false;
            float dist = 0.0f;
            if (m.mode == 0) {
                dist = sdist[my_off];
                have_sample = true;
            } else {
                const float d0 = sdist[my_off + 0], d1 = sdist[my_off + 1], d2 = sdist[my_off + 2], d3 = sdist[my_off + 3];
                evals += 4u;
                const V3 acc = ((v3(ex, ey, ey) * v3s(d0) + v3(ey, ey, ex) * v3s(d1)) + v3(ey, ex, ey) * v3s(d2)) + v3(ex, ex, ex) * v3s(d3);
                const V3 n = normalize(acc);
                // IsAValidHit: in range and not back-facing
                if (m.t >= 0.0f && m.t <= m.tcur && !(dot(m.d, n) > 0.0f)) {
                    publish_lane_hit(sc, wb, TYPE, m, m.t, n);
                    m.active = false;
                } else {
                    m.t = t_next;
                    m.mode = 0;
                    if (!(m.step < max_steps && m.t <= m.t_stop)) m.active = false;
                    else { dist = sdist[my_off + 4]; have_sample = true; }
                }
            }
            if (have_sample) {
                evals++;
                m.step++;
                // a hit candidate needs its normal only if IsAValidHit could accept it: t outside [RayTMin(), RayTCurrent()]
                // is rejected whatever the normal is, and the step that follows is the ordinary one (see march_lane.cuh)
                if (dist <= threshold * m.t && m.t >= 0.0f && m.t <= m.tcur) { m.mode = 1; m.dist_hit = dist; }
                else m.t += m.scale * dist;
            }
        }
        __syncwarp(); // the point list is rewritten by the next trip
    }
    return evals;
}

// ---- warp-private driver: lanes are refilled from the queue once RMIN of them idle -----------------
template <int RMIN>
__device__ __forceinline__ void coop_march_sdf(const int SDF, const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, WaveCtrl* ctrl, int pq,
                                               int lane, CoopSmem& sm) {
    const int TYPE = SDF + 3;
    constexpr int kTrips = 8; // loop trips between refill checks
    const uint32_t count = min(ctrl->cand_count[TYPE], pp.cand_capacity);
    if (count == 0) return;
    WarpChunk q;
    MarchLane m;
    m.active = false; m.mode = 0; m.step = 0; m.c = 0; m.r = 0; m.shadow = false;
    m.o = v3s(0.0f); m.d = v3s(0.0f); m.t = m.tcur = m.t_stop = m.scale = m.dist_hit = 0.0f;
    uint32_t work = 0;
    for (;;) {
        poll_committed(wb, TYPE, m);
        const uint32_t idx = warp_take(q, refill_now<RMIN>(m.active) && !m.active, &ctrl->cursor[TYPE], count, lane);
        if (idx != kNoCand) fetch_lane(sc, wb, pq, TYPE, idx, m);
        if (!__any_sync(kFull, m.active)) {
            if (q.exhausted && q.next >= q.end) break;
            continue;
        }
        work += coop_trips(SDF, m, kTrips, sc, pp, wb, lane, sm);
    }
    for (int off = 16; off; off >>= 1) work += __shfl_xor_sync(kFull, work, off);
    if (lane == 0 && work) atomicAdd(&ctrl->work[TYPE], work);
}

// ---------------------------------------------------------------------------------------------
// Metaballs: step-parallel scan.
// RayMetaballsIntersectionTest (src/VolumetricPrimitives.hlsli:151-202) is not a sphere tracer: it
// samples the field at 128 FIXED steps t(i+1) = t(i) + dt, so all samples of a ray are known up
// front.  One candidate per lane leaves most lanes idle (measured 10-13 of 32: three quarters of
// the candidates cross no bounding sphere at all, the rest run 16..128 steps).  Here
//   phase A  every lane takes a candidate, transforms the ray and runs FindIntersectingMetaballs;
//            the survivors are appended to a warp-private list in shared memory,
//   phase B  the warp walks the list one ray at a time, lane k sampling step (base + k); the first
//            lane whose potential reaches the threshold gets its 6-tap normal from lanes 0..5, and
//            IsAValidHit decides between ReportHit and going on with the next crossing.
// Each lane accumulates t by the same repeated addition as the reference loop (k adds on top of
// the pass's base), so every sample position -- and the parity build's result -- is bit-identical.
// ---------------------------------------------------------------------------------------------
template <int NB> __device__ __forceinline__ float metaballs_potential_masked(V3 p, const Metaballs<NB>& mb, uint32_t mask) {
    float sum = 0.0f; // NB >= 5 builds sum over the balls whose bounding sphere the ray crosses, in index order
#pragma unroll
    for (int j = 0; j < NB; j++)
        if (NB < 5 || ((mask >> j) & 1u)) sum += metaball_potential(p, mb.cx[j], mb.cy[j], mb.cz[j], mb.radius[j]);
    return sum;
}

// FindIntersectingMetaballs (:124-147) without the in-place compaction: span + mask of crossed bounding spheres
template <int NB>
__device__ __forceinline__ uint32_t find_metaball_span(V3 o, V3 d, float ray_tmin, float ray_tcur, const Metaballs<NB>& mb, float& tmin, float& tmax) {
    const float kInf = __builtin_inff();
    LocalRay r;
    r.o = o; r.d = d; r.tmin = ray_tmin; r.tcur = ray_tcur;
    tmin = kInf;
    tmax = -kInf;
    uint32_t mask = 0;
#pragma unroll
    for (int i = 0; i < NB; i++) {
        float t0, t1;
        if (ray_sphere_roots(r, v3(mb.cx[i], mb.cy[i], mb.cz[i]), mb.radius[i], t0, t1)) {
            const float th = fmaxf(t0, r.tmin), tm = fminf(t1, r.tcur);
            tmin = fminf(th, tmin);
            tmax = fmaxf(tm, tmax);
            mask |= 1u << i;
        }
    }
    tmin = fmaxf(tmin, r.tmin);
    tmax = fminf(tmax, r.tcur);
    return mask;
}

template <int NB>
__device__ __forceinline__ void scan_metaballs_queue(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, WaveCtrl* ctrl, int pq, int lane,
                                                     CoopSmem& sm) {
    constexpr int TYPE = 2;
    const uint32_t count = min(ctrl->cand_count[TYPE], pp.cand_capacity);
    if (count == 0) return;
    const int warp = threadIdx.x >> 5;
    // warp-private ray list, 12 words per entry, aliased onto the sphere tracer's point list (640 words per warp)
    float* const L = sm.px[warp]; // px, py, pz, dist of one warp are NOT contiguous: use px/py/pz (3 x 160 words) as three 4-word-per-entry banks
    float* const L1 = sm.py[warp];
    float* const L2 = sm.pz[warp];
    const uint32_t max_steps = pp.metaball_max_steps;
    const float e = 0.5773f * 0.00001f;
    Metaballs<NB> mb;
    init_metaballs<NB>(mb, pp.fc.elapsed_time);
    const unsigned lt = (1u << lane) - 1u;
    WarpChunk q;
    uint32_t work = 0;
    for (;;) {
        // ---- phase A: one candidate per lane ----
        const uint32_t idx = warp_take(q, true, &ctrl->cursor[TYPE], count, lane);
        bool keep = false;
        MarchLane m;
        float tmin = 0.0f, tmax = 0.0f;
        uint32_t mask = 0;
        if (idx != kNoCand && fetch_lane(sc, wb, pq, TYPE, idx, m)) {
            mask = find_metaball_span<NB>(m.o, m.d, 0.0f, m.tcur, mb, tmin, tmax);
            keep = tmin != __builtin_inff(); // otherwise the field is 0 at every sample
        }
        const unsigned kb = __ballot_sync(kFull, keep);
        const int n_list = __popc(kb);
        if (keep) {
            const int slot = __popc(kb & lt);
            L[4 * slot + 0] = m.o.x; L[4 * slot + 1] = m.o.y; L[4 * slot + 2] = m.o.z; L[4 * slot + 3] = tmin;
            L1[4 * slot + 0] = m.d.x; L1[4 * slot + 1] = m.d.y; L1[4 * slot + 2] = m.d.z;
            L1[4 * slot + 3] = (tmax - tmin) / (float)(max_steps / 1u);
            L2[4 * slot + 0] = m.tcur; L2[4 * slot + 1] = __uint_as_float(mask); L2[4 * slot + 2] = __uint_as_float(m.c);
            L2[4 * slot + 3] = __uint_as_float(m.r);
        }
        __syncwarp();
        // ---- phase B: one ray at a time, 32 steps per pass ----
#pragma unroll 1
        for (int ei = 0; ei < n_list; ei++) {
            MarchLane r;
            r.o = v3(L[4 * ei + 0], L[4 * ei + 1], L[4 * ei + 2]);
            r.d = v3(L1[4 * ei + 0], L1[4 * ei + 1], L1[4 * ei + 2]);
            float t_base = L[4 * ei + 3];
            const float dt = L1[4 * ei + 3];
            r.tcur = L2[4 * ei + 0];
            const uint32_t bmask = __float_as_uint(L2[4 * ei + 1]);
            r.c = __float_as_uint(L2[4 * ei + 2]);
            r.r = __float_as_uint(L2[4 * ei + 3]);
            bool done = false;
            uint32_t steps_done = max_steps;
#pragma unroll 1
            for (uint32_t base = 0; base < max_steps && !done; base += 32u) {
                // t of step (base + lane): the reference's `t += minTStep`, lane times
                float t = t_base;
                for (int k = 0; k < 31; k++) if (k < lane) t += dt;
                const V3 p = r.o + t * r.d;
                const float sum = metaballs_potential_masked<NB>(p, mb, bmask);
                unsigned cross = __ballot_sync(kFull, sum >= 0.25f && base + (uint32_t)lane < max_steps);
                while (cross) {
                    const int k = __ffs(cross) - 1;
                    cross &= cross - 1u;
                    // CalculateMetaballsNormal at lane k's sample: taps -x +x -y +y -z +z on lanes 0..5
                    const V3 pk = v3(__shfl_sync(kFull, p.x, k), __shfl_sync(kFull, p.y, k), __shfl_sync(kFull, p.z, k));
                    const float tk = __shfl_sync(kFull, t, k);
                    float f = 0.0f;
                    if (lane < 6) {
                        const float s = (lane & 1) ? e : -e;
                        const V3 tap = v3(lane < 2 ? s : 0.0f, (lane >> 1) == 1 ? s : 0.0f, lane >= 4 ? s : 0.0f);
                        f = metaballs_potential_masked<NB>(pk + tap, mb, bmask);
                    }
                    const float f0 = __shfl_sync(kFull, f, 0), f1 = __shfl_sync(kFull, f, 1), f2 = __shfl_sync(kFull, f, 2);
                    const float f3 = __shfl_sync(kFull, f, 3), f4 = __shfl_sync(kFull, f, 4), f5 = __shfl_sync(kFull, f, 5);
                    const V3 n = normalize(v3(f0 - f1, f2 - f3, f4 - f5));
                    if (tk >= 0.0f && tk <= r.tcur && !(dot(r.d, n) > 0.0f)) { // IsAValidHit
                        if (lane == 0) publish_lane_hit(sc, wb, TYPE, r, tk, n);
                        steps_done = base + (uint32_t)k + 1u;
                        done = true;
                        break;
                    }
                }
                // next pass starts one step after lane 31's sample
                t_base = __shfl_sync(kFull, t, 31) + dt;
            }
            work += steps_done;
        }
        __syncwarp();
        if (q.exhausted && q.next >= q.end) break;
    }
    if (lane == 0 && work) atomicAdd(&ctrl->work[TYPE], work);
}

// ---- all types in one persistent launch ---------------------------------------------------------
// The closed-form tests run first (they may occlude a shadow ray before its marches start).  The
// marching types are then visited most expensive first -- pyramid, cog, metaballs, cylinders, mini
// spheres, round cube, twisted torus, square torus -- but every SM starts at its own position of
// that list (%smid) and wraps: all queues drain concurrently from the start of the launch, while
// the warps resident on one SM run the same code (measured: with a per-warp rotation
// stall_no_instruction became the top stall).
template <int NB, int RMIN>
__global__ void __launch_bounds__(kCoopBlock, 1024 / kCoopBlock) k_intersect_coop(const __grid_constant__ SceneDev sc, const __grid_constant__ PassParams pp,
                                                              const __grid_constant__ WaveBuffers wb, int wave, uint32_t type_mask) {
    __shared__ CoopSmem sm;
    WaveCtrl* ctrl = &wb.ctrl[wave];
    const int pq = wave & 1;
    const int lane = threadIdx.x & 31;
    if (type_mask & (1u << 1)) analytic_queue<1>(sc, pp, wb, ctrl, pq, lane);
    if (type_mask & (1u << 0)) analytic_queue<0>(sc, pp, wb, ctrl, pq, lane);
    uint32_t rot;
    asm("mov.u32 %0, %%smid;" : "=r"(rot));
#pragma unroll 1
    for (uint32_t i = 0; i < 8u; i++) {
        const int type = (int)((0x56438279u >> (4u * ((i + rot) & 7u))) & 15u);
        if (!((type_mask >> type) & 1u)) continue;
        if (type == 2) scan_metaballs_queue<NB>(sc, pp, wb, ctrl, pq, lane, sm);
        else coop_march_sdf<RMIN>(type - 3, sc, pp, wb, ctrl, pq, lane, sm);
    }
}

// one SDF type per launch (profiling mode)
template <int SDF, int RMIN>
__global__ void __launch_bounds__(kCoopBlock) k_coop_sdf(const __grid_constant__ SceneDev sc, const __grid_constant__ PassParams pp,
                                                     const __grid_constant__ WaveBuffers wb, int wave) {
    __shared__ CoopSmem sm;
    coop_march_sdf<RMIN>(SDF, sc, pp, wb, &wb.ctrl[wave], wave & 1, threadIdx.x & 31, sm);
}

template <int NB>
__global__ void __launch_bounds__(kCoopBlock) k_scan_metaballs(const __grid_constant__ SceneDev sc, const __grid_constant__ PassParams pp,
                                                           const __grid_constant__ WaveBuffers wb, int wave) {
    __shared__ CoopSmem sm;
    scan_metaballs_queue<NB>(sc, pp, wb, &wb.ctrl[wave], wave & 1, threadIdx.x & 31, sm);
}

} // namespace
} // namespace gprt
