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

constexpr int kWarpsPerBlock = kBlock / 32;
constexpr int kCoopPts = 32 * 5; // worst case: every lane posts 5 points

struct CoopSmem { // warp-private scratch
    float px[kWarpsPerBlock][kCoopPts], py[kWarpsPerBlock][kCoopPts], pz[kWarpsPerBlock][kCoopPts];
    float dist[kWarpsPerBlock][kCoopPts];
};

// GetDistanceFromSignedDistancePrimitive with a run-time (warp-uniform) primitive type: one copy of
// the march loop serves all seven SDF types; the branch is uniform, so it costs a jump.
__device__ __forceinline__ float sdf_distance_dyn(int sdf, V3 p) {
    switch (sdf) {
    case 0: return sdf_distance<0>(p);
    case 1: return sdf_distance<1>(p);
    case 2: return sdf_distance<2>(p);
    case 3: return sdf_distance<3>(p);
    case 4: return sdf_distance<4>(p);
    case 5: return sdf_distance<5>(p);
    default: return sdf_distance<6>(p);
    }
}

// fetch_candidate / publish for a run-time SDF type (the queues are indexed by type id = sdf + 3)
__device__ __forceinline__ bool fetch_candidate_dyn(const SceneDev& sc, const WaveBuffers& wb, int pq, int type, uint32_t c, CandState& s, bool& shadow) {
    const uint2 cand = wb.cand[type][c];
    s.c = c; s.r = cand.x; s.prim = cand.y;
    const float4 rd = wb.ray_d[pq][s.r];
    const uint32_t meta = __float_as_uint(rd.w);
    shadow = meta_shadow(meta) != 0u;
    if (shadow && wb.hitkey[s.r] != kKeyMiss) {
        wb.cand_res[type][c] = make_float4(0.0f, 0.0f, 0.0f, -1.0f);
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
__device__ __forceinline__ void publish_hit_dyn(const SceneDev& sc, const WaveBuffers& wb, int type, const CandState& s, float thit, V3 n_local) {
    if (thit >= s.lr.tmin && thit <= s.lr.tcur) {
        const V3 nb = mul_vector_rowvec(n_local, s.pf);
        const V3 nw = normalize(xform_vector_3x4(sc.aabb_o2w, nb));
        atomicMin(&wb.hitkey[s.r], pack_key(thit, s.prim));
        wb.cand_res[type][s.c] = make_float4(nw.x, nw.y, nw.z, fabsf(thit));
    } else {
        wb.cand_res[type][s.c] = make_float4(0.0f, 0.0f, 0.0f, -1.0f);
    }
}

template <int RMIN>
__device__ __forceinline__ void coop_march_sdf(const int SDF, const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, WaveCtrl* ctrl, int pq,
                                               int lane, CoopSmem& sm) {
    const int TYPE = SDF + 3;
    constexpr int kTrips = 8; // loop trips between refill checks
    const uint32_t count = min(ctrl->cand_count[TYPE], pp.cand_capacity);
    if (count == 0) return;
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

    WarpChunk q;
    CandState s;
    bool active = false;
    int mode = 0; // 0: needs the field at p(t); 1: hit candidate at pos, needs the 4 taps (+ next sample)
    uint32_t step = 0, work = 0;
    float t = 0.0f, dist_hit = 0.0f, t_stop = 0.0f;
    bool shadow = false;
    V3 pos = v3s(0.0f);

    for (;;) {
        // Early out against what other candidates of the same ray have committed meanwhile (the hit key
        // lives in L2, hence __ldcg).  A shadow ray needs any hit (ACCEPT_FIRST_HIT_AND_END_SEARCH); a
        // radiance ray keeps the closest, so samples beyond the committed t cannot win the atomicMin --
        // the same shrinking RayTCurrent() a DXR traversal would show this shader.  The image is unchanged.
        if (active) {
            const unsigned long long key = __ldcg(&wb.hitkey[s.r]);
            if (key != kKeyMiss) {
                if (shadow) { wb.cand_res[TYPE][s.c] = make_float4(0.0f, 0.0f, 0.0f, -1.0f); active = false; }
                else t_stop = fminf(t_stop, __uint_as_float((uint32_t)(key >> 32)));
            }
        }
        const uint32_t idx = warp_take(q, refill_now<RMIN>(active) && !active, &ctrl->cursor[TYPE], count, lane);
        if (idx != kNoCand && fetch_candidate_dyn(sc, wb, pq, TYPE, idx, s, shadow)) {
            active = true; mode = 0; step = 0; t = s.lr.tmin; t_stop = s.lr.tcur;
        }
        if (!__any_sync(kFull, active)) {
            if (q.exhausted && q.next >= q.end) break;
            continue;
        }
#pragma unroll 1
        for (int trip = 0; trip < kTrips; trip++) {
            // while (i++ < MaxSteps && t <= RayTCurrent())
            if (active && mode == 0 && !(step < max_steps && t <= t_stop)) { wb.cand_res[TYPE][s.c] = make_float4(0.0f, 0.0f, 0.0f, -1.0f); active = false; }
            const unsigned b1 = __ballot_sync(kFull, active && mode == 0), b5 = __ballot_sync(kFull, active && mode == 1);
            if (!(b1 | b5)) break;
            const uint32_t my_off = (uint32_t)__popc(b1 & lt) + 5u * (uint32_t)__popc(b5 & lt);
            const uint32_t total = (uint32_t)__popc(b1) + 5u * (uint32_t)__popc(b5);
            float t_next = 0.0f;
            if (active) {
                if (mode == 0) {
                    const V3 p = s.lr.o + t * s.lr.d;
                    spx[my_off] = p.x; spy[my_off] = p.y; spz[my_off] = p.z;
                } else {
                    spx[my_off + 0] = pos.x + ex; spy[my_off + 0] = pos.y + ey; spz[my_off + 0] = pos.z + ey;
                    spx[my_off + 1] = pos.x + ey; spy[my_off + 1] = pos.y + ey; spz[my_off + 1] = pos.z + ex;
                    spx[my_off + 2] = pos.x + ey; spy[my_off + 2] = pos.y + ex; spz[my_off + 2] = pos.z + ey;
                    spx[my_off + 3] = pos.x + ex; spy[my_off + 3] = pos.y + ex; spz[my_off + 3] = pos.z + ex;
                    // the sample the loop takes next if this candidate is culled: t += stepScale * distance
                    t_next = t + s.step_scale * dist_hit;
                    const V3 p = s.lr.o + t_next * s.lr.d;
                    spx[my_off + 4] = p.x; spy[my_off + 4] = p.y; spz[my_off + 4] = p.z;
                }
            }
            __syncwarp();
            for (uint32_t j = lane; j < total; j += 32u) sdist[j] = sdf_distance_dyn(SDF, v3(spx[j], spy[j], spz[j])); // the one evaluation site
            __syncwarp();
            if (active) {
                bool have_sample = false;
                float dist = 0.0f;
                if (mode == 0) {
                    dist = sdist[my_off];
                    have_sample = true;
                } else {
                    const float d0 = sdist[my_off + 0], d1 = sdist[my_off + 1], d2 = sdist[my_off + 2], d3 = sdist[my_off + 3];
                    work += 4u;
                    const V3 acc = ((v3(ex, ey, ey) * v3s(d0) + v3(ey, ey, ex) * v3s(d1)) + v3(ey, ex, ey) * v3s(d2)) + v3(ex, ex, ex) * v3s(d3);
                    const V3 n = normalize(acc);
                    if (is_valid_hit(s.lr, t, n)) {
                        publish_hit_dyn(sc, wb, TYPE, s, t, n);
                        active = false;
                    } else {
                        t = t_next;
                        mode = 0;
                        if (!(step < max_steps && t <= t_stop)) { wb.cand_res[TYPE][s.c] = make_float4(0.0f, 0.0f, 0.0f, -1.0f); active = false; }
                        else { dist = sdist[my_off + 4]; have_sample = true; }
                    }
                }
                if (have_sample) {
                    work++;
                    step++;
                    if (dist <= threshold * t) { mode = 1; pos = s.lr.o + t * s.lr.d; dist_hit = dist; }
                    else t += s.step_scale * dist;
                }
            }
            __syncwarp(); // the point list is rewritten by the next trip
        }
    }
    for (int off = 16; off; off >>= 1) work += __shfl_xor_sync(kFull, work, off);
    if (lane == 0 && work) atomicAdd(&ctrl->work[TYPE], work);
}

// ---- all types in one persistent launch ---------------------------------------------------------
// The closed-form tests run first (they may occlude a shadow ray before its marches start).  The
// marching types are then visited in an order ROTATED by the warp's global index, so all type
// queues are drained concurrently from the start of the launch: the longest rays of different
// types overlap instead of queueing behind one another.
template <int NB, int RMIN>
__global__ void __launch_bounds__(kBlock, 3) k_intersect_coop(const __grid_constant__ SceneDev sc, const __grid_constant__ PassParams pp,
                                                           const __grid_constant__ WaveBuffers wb, int wave, uint32_t type_mask) {
    __shared__ CoopSmem sm;
    WaveCtrl* ctrl = &wb.ctrl[wave];
    const int pq = wave & 1;
    const int lane = threadIdx.x & 31;
    if (type_mask & (1u << 1)) analytic_queue<1>(sc, pp, wb, ctrl, pq, lane);
    if (type_mask & (1u << 0)) analytic_queue<0>(sc, pp, wb, ctrl, pq, lane);
    // marching types, most expensive first: pyramid, cog, metaballs, cylinders, mini spheres, round cube,
    // twisted torus, square torus (type ids).  Every SM starts at its own position of this list and
    // wraps: all queues drain concurrently from the start of the launch, while the warps resident on
    // one SM run the same code (the kernel is ~150 KB of SASS; mixed types on one SM thrash its
    // instruction cache -- measured: stall_no_instruction was the top stall with a per-warp rotation).
    uint32_t rot;
    asm("mov.u32 %0, %%smid;" : "=r"(rot));
#pragma unroll 1
    for (uint32_t i = 0; i < 8u; i++) {
        const int type = (int)((0x56438279u >> (4u * ((i + rot) & 7u))) & 15u);
        if (!((type_mask >> type) & 1u)) continue;
        if (type == 2) simple_queue<2, NB>(sc, pp, wb, ctrl, pq, lane); // fixed-step scan: one candidate per lane measured fastest
        else coop_march_sdf<RMIN>(type - 3, sc, pp, wb, ctrl, pq, lane, sm);
    }
}

// one SDF type per launch (profiling mode)
template <int SDF, int RMIN>
__global__ void __launch_bounds__(kBlock) k_coop_sdf(const __grid_constant__ SceneDev sc, const __grid_constant__ PassParams pp,
                                                     const __grid_constant__ WaveBuffers wb, int wave) {
    __shared__ CoopSmem sm;
    coop_march_sdf<RMIN>(SDF, sc, pp, wb, &wb.ctrl[wave], wave & 1, threadIdx.x & 31, sm);
}

} // namespace
} // namespace gprt
