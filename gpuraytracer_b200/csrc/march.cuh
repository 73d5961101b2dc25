// march.cuh -- candidate-queue plumbing shared by the intersection kernels (included by kernels.cu):
// warp-aggregated queue reservation (warp_take), GetRayInAABBPrimitiveLocalSpace for one queue entry
// (fetch_candidate), ReportHit (publish_hit) and the closed-form analytic tests over their queues.
// The marchers are in march_coop.cuh.
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
    uint32_t c, r, prim; // prim: flattened instance * n_prims + primitive
    const float* o2w;    // ObjectToWorld3x4() of the candidate's TLAS instance
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
    const InstXform ix = inst_xform(sc, s.prim);
    s.o2w = ix.o2w;
    const V3 oo = xform_point_3x4(ix.w2o, ld3(ro)), dd = xform_vector_3x4(ix.w2o, ld3(rd));
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
        const V3 nw = normalize(xform_vector_3x4(s.o2w, nb));
        atomicMin(&wb.hitkey[s.r], pack_key(thit, s.prim));
        wb.cand_res[TYPE][s.c] = make_float4(nw.x, nw.y, nw.z, fabsf(thit));
    } else {
        wb.cand_res[TYPE][s.c] = make_float4(0.0f, 0.0f, 0.0f, -1.0f);
    }
}
template <int TYPE> __device__ __forceinline__ void publish_miss(const WaveBuffers& wb, const CandState& s) {
    wb.cand_res[TYPE][s.c] = make_float4(0.0f, 0.0f, 0.0f, -1.0f);
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

} // namespace
} // namespace gprt
