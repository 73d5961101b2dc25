// march_lane.cuh -- the per-lane sphere tracer (included by kernels.cu after march_coop.cuh; same anonymous namespace):
// RaySignedDistancePrimitiveTest (src/SignedDistancePrimitives.hlsli:287-319) over the per-type candidate queues.
// GPRT_ISECT_VARIANT=2 selects it; the shipped kernel is k_intersect_coop (march_coop.cuh).
//
// One candidate per lane, the march state in registers, one field evaluation per lane and trip with no point list and no
// shared-memory round trip; only HIT CANDIDATES (distance <= 1e-4 * t, t in range) go through the warp-cooperative path:
// their four normal taps and the sample the loop takes next if IsAValidHit culls them are posted to the warp's list once six
// candidates wait (or nobody marches, or every 8 trips) and evaluated by all 32 lanes.  One copy of the loop serves the seven
// SDF types (warp-uniform switch to a __noinline__ field function per type).
//
// Measured against k_intersect_coop on a B200 (round 2): within +-4 % on every config -- the same
// instruction count (312 M vs 329 M warp instructions for wave 1 of config 2; the field evaluations are ~80 % of either) and
// the same serial floor (the longest march: 512 dependent steps of ~330 instructions at ~4.5 cycles each, whatever feeds
// them).  A first version with the loop specialised per type was 13 456 instructions (215 KB) and starved for instructions
// (stall_no_instruction 6.0 per issue, 39 % issue slots busy); a continuation kernel that marched the exported long rays with
// five lanes each (taps + next sample by shuffles) ran the same chains no faster (1 400-1 600 cycles per step on an idle
// scheduler: the chain is bound by dependent-issue latency, not by contention) and was removed.
#pragma once

namespace gprt {
namespace {

constexpr float kSdfThreshold = 0.0001f;                                       // src/SignedDistancePrimitives.hlsli:289
constexpr float kTapX = 1.0f * 0.5773f * 0.0001f, kTapY = -1.0f * 0.5773f * 0.0001f; // sdCalculateNormal: e.xyy, e.yyx, e.yxy, e.xxx

// "Has the ray intersected the primitive?" (distance <= threshold * t) AND could IsAValidHit accept it.  The reference
// computes the 4-tap normal of every such sample and only then finds t outside [RayTMin(), RayTCurrent()]
// (src/SignedDistancePrimitives.hlsli:302-311, src/RaytracingShaderHelper.hlsli:136-139); the range test needs no normal,
// and a rejected sample advances by t += stepScale * distance exactly like an ordinary step.  The rays that use all 512
// steps start INSIDE a surface (distance <= 0 at t = 0), step backwards to t < 0 and stay there: a hit candidate at every
// step that can never be in range.
__device__ __forceinline__ bool hit_candidate(float dist, const MarchLane& m) {
    return dist <= kSdfThreshold * m.t && m.t >= 0.0f && m.t <= m.tcur;
}

// One copy of each field function per kernel (the march step and the cooperative pass both call it), reached through a
// warp-uniform switch: code size, not arithmetic, was what starved the schedulers of the per-type version.
template <int SDF> __device__ __noinline__ float sdf_eval(float x, float y, float z, uint32_t iters) { return sdf_distance<SDF>(v3(x, y, z), iters); }
__device__ __forceinline__ float sdf_eval_dyn(int sdf, float x, float y, float z, uint32_t iters) {
    switch (sdf) {
    case 0: return sdf_eval<0>(x, y, z, iters);
    case 1: return sdf_eval<1>(x, y, z, iters);
    case 2: return sdf_eval<2>(x, y, z, iters);
    case 3: return sdf_eval<3>(x, y, z, iters);
    case 4: return sdf_eval<4>(x, y, z, iters);
    case 5: return sdf_eval<5>(x, y, z, iters);
    default: return sdf_eval<6>(x, y, z, iters);
    }
}

// The hit candidates of the warp (lanes of b5): normal taps + speculative next sample through the point list.
__device__ __forceinline__ uint32_t hit_candidates(const int SDF, MarchLane& m, unsigned b5, const SceneDev& sc, const PassParams& pp,
                                                   const WaveBuffers& wb, int lane, CoopSmem& sm) {
    const int TYPE = SDF + 3;
    const int warp = threadIdx.x >> 5;
    float* const spx = sm.px[warp];
    float* const spy = sm.py[warp];
    float* const spz = sm.pz[warp];
    float* const sdist = sm.dist[warp];
    const bool mine = (b5 >> lane) & 1u;
    const uint32_t off = 5u * (uint32_t)__popc(b5 & ((1u << lane) - 1u));
    const uint32_t total = 5u * (uint32_t)__popc(b5);
    float t_next = 0.0f;
    if (mine) {
        const V3 pos = m.o + m.t * m.d;
        spx[off + 0] = pos.x + kTapX; spy[off + 0] = pos.y + kTapY; spz[off + 0] = pos.z + kTapY;
        spx[off + 1] = pos.x + kTapY; spy[off + 1] = pos.y + kTapY; spz[off + 1] = pos.z + kTapX;
        spx[off + 2] = pos.x + kTapY; spy[off + 2] = pos.y + kTapX; spz[off + 2] = pos.z + kTapY;
        spx[off + 3] = pos.x + kTapX; spy[off + 3] = pos.y + kTapX; spz[off + 3] = pos.z + kTapX;
        t_next = m.t + m.scale * m.dist_hit; // the sample the loop takes next if this candidate is culled: t += stepScale * distance
        const V3 p = m.o + t_next * m.d;
        spx[off + 4] = p.x; spy[off + 4] = p.y; spz[off + 4] = p.z;
    }
    __syncwarp();
    for (uint32_t j = lane; j < total; j += 32u) sdist[j] = sdf_eval_dyn(SDF, spx[j], spy[j], spz[j], pp.fractal_iterations);
    __syncwarp();
    uint32_t evals = 0;
    if (mine) {
        const float d0 = sdist[off + 0], d1 = sdist[off + 1], d2 = sdist[off + 2], d3 = sdist[off + 3];
        evals = 4u;
        const V3 acc = ((v3(kTapX, kTapY, kTapY) * v3s(d0) + v3(kTapY, kTapY, kTapX) * v3s(d1)) + v3(kTapY, kTapX, kTapY) * v3s(d2)) + v3(kTapX, kTapX, kTapX) * v3s(d3);
        const V3 n = normalize(acc);
        if (m.t >= 0.0f && m.t <= m.tcur && !(dot(m.d, n) > 0.0f)) { // IsAValidHit: in range and not back-facing
            publish_lane_hit(sc, wb, TYPE, m, m.t, n);
            m.active = false;
        } else {
            m.t = t_next;
            m.mode = 0;
            if (!(m.step < pp.sdf_max_steps && m.t <= m.t_stop)) m.active = false; // while (i++ < MaxSteps && t <= RayTCurrent())
            else {
                const float dist = sdist[off + 4];
                evals++;
                m.step++;
                if (hit_candidate(dist, m)) { m.mode = 1; m.dist_hit = dist; }
                else m.t += m.scale * dist;
            }
        }
    }
    __syncwarp(); // the point list is rewritten by the next trip
    return evals;
}

__device__ __noinline__ void march_sdf_queue(const int SDF, const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, WaveCtrl* ctrl, int pq, CoopSmem& sm) {
    const int TYPE = SDF + 3;
    constexpr int kTrips = 8; // march steps between refill / export / poll checks
    const int lane = threadIdx.x & 31;
    const uint32_t count = min(ctrl->cand_count[TYPE], pp.cand_capacity);
    if (count == 0) return;
    const uint32_t max_steps = pp.sdf_max_steps, iters = pp.fractal_iterations;
    WarpChunk q;
    MarchLane m;
    m.active = false; m.mode = 0; m.step = 0; m.c = 0; m.r = 0; m.shadow = false;
    m.o = v3s(0.0f); m.d = v3s(0.0f); m.t = m.tcur = m.t_stop = m.scale = m.dist_hit = 0.0f;
    uint32_t work = 0;
    for (;;) {
        poll_committed(wb, TYPE, m);
        const uint32_t idx = warp_take(q, refill_now<kRefillMin>(m.active) && !m.active, &ctrl->cursor[TYPE], count, lane);
        if (idx != kNoCand) fetch_lane(sc, wb, pq, TYPE, idx, m);
        if (!__any_sync(kFull, m.active)) {
            if (q.exhausted && q.next >= q.end) break;
            continue;
        }
#pragma unroll 1
        for (int trip = 0; trip < kTrips; trip++) {
            if (m.active && m.mode == 0) {
                if (!(m.step < max_steps && m.t <= m.t_stop)) m.active = false; // while (i++ < MaxSteps && t <= RayTCurrent())
                else {
                    const V3 p = m.o + m.t * m.d;
                    const float dist = sdf_eval_dyn(SDF, p.x, p.y, p.z, iters);
                    m.step++;
                    work++;
                    if (hit_candidate(dist, m)) { m.mode = 1; m.dist_hit = dist; }
                    else m.t += m.scale * dist;
                }
            }
            // Hit candidates wait until six of them fill an evaluation pass (5 points each), nobody is left marching, or the
            // block of trips ends: one cooperative pass per ~8 trips instead of a 5-point pass in almost every trip.
            const unsigned b5 = __ballot_sync(kFull, m.active && m.mode == 1);
            const unsigned b1 = __ballot_sync(kFull, m.active && m.mode == 0);
            if (b5 && (__popc(b5) >= 6 || !b1 || trip == kTrips - 1)) work += hit_candidates(SDF, m, b5, sc, pp, wb, lane, sm);
            else if (!(b1 | b5)) break;
        }
    }
    for (int off = 16; off; off >>= 1) work += __shfl_xor_sync(kFull, work, off);
    if (lane == 0 && work) atomicAdd(&ctrl->work[TYPE], work);
}

// ---- all types in one persistent launch (see k_intersect_coop for the order and the %smid rotation) ----
template <int NB>
__global__ void __launch_bounds__(kCoopBlock, 1024 / kCoopBlock) k_intersect_lane(const __grid_constant__ SceneDev sc, const __grid_constant__ PassParams pp,
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
        else march_sdf_queue(type - 3, sc, pp, wb, ctrl, pq, sm);
    }
}

} // namespace
} // namespace gprt
