// walk.cuh -- the LBVH walk of the traversal stage (included by kernels.cu; same anonymous namespace).
//
// What DXR's TraceRay does between the raygen / closest-hit shaders and the intersection shaders
// (src/Raytracing.hlsl:103,135; configured at src/Renderer.cpp:567-835, PREFER_FAST_TRACE): find every procedural AABB
// the ray's interval [TMin, TCurrent] crosses.  B200 has no RT cores; this is the software form, shaped by three
// measurements of round 1's walk (one thread per ray, binary 64-byte nodes fetched with __ldg, a 64-entry stack in local
// memory; profiles/r01_ncu_full_k_traverse_config5.json): 2.2 G node fetches per 8K frame = 141 GB through L2
// (~11.5 TB/s: L2-bandwidth bound), 272-336 bytes of local-memory stack per thread, lane efficiency 0.35-0.42 (a warp of
// 32 rays ran until its longest ray was done).
//
//   * 4-wide nodes (lbvh.cu, k_collapse4): two binary levels per fetch, four independent slab tests per node.
//   * The top of the tree lives in SHARED MEMORY: every block copies the breadth-first top array (k_build_top; 320 nodes
//     = the first four levels and most of the fifth, 40 KB) once and every ray starts there, so the levels all rays cross
//     cost no L2 traffic at all.
//   * The traversal stack is in shared memory too: kStackDepth entries per thread, [entry][thread] so a warp's accesses
//     never conflict.  A push onto a full stack goes to the hand-over queue instead (never dropped).
//   * Persistent warps take batches of 32 consecutive rays from their chunk (a run of consecutive rays claimed from the
//     wave's cursor, one atomic per chunk, the next chunk requested before the current one runs out; guided chunk sizes so
//     the warps finish together).
//   * A node budget per ray (pp.spill_budget fetches): what is left of a long walk is handed to the next of kSpillRounds
//     rounds as (ray, subtree) entries that other lanes walk in parallel.  The two budgeted rounds expand their entries
//     BREADTH FIRST for kRoundBudget fetches (a 16-entry ring in the same shared-memory column), so a big subtree leaves
//     them as up to 16 pieces of similar size; the last round walks the pieces to the end depth first on a 48-entry stack.
//     Lanes of the round kernels are refilled one by one (entries are cheap to start and of every size).  Measured on the
//     8K frame: the longest walk of any item is 32 fetches (900 without the rounds), the three round launches of a wave
//     take 0.25 ms instead of 0.6 ms, and a 1/8 stripe of the frame -- one rank of eight -- loses 0.3 ms less per wave.
//     The candidate set does not depend on the visiting order, so images and hit ids are those of the brute-force loop.
// Candidates are collected in registers (kMaxPending per ray) and appended to the per-type queues after each batch, where
// the warp is convergent (flush_pending).
#pragma once

namespace gprt {
namespace {

constexpr int kStackDepth = 16;      // shared-memory stack entries per thread (main walk and budgeted rounds)
constexpr int kStackDepthLast = 48;  // ... of the last round, which must finish whatever it is handed (128 threads per block)
constexpr int kWalkBlock = 256;
constexpr int kWalkBlockLast = 128;
constexpr uint32_t kChunkMax = 128;  // rays a warp claims from the wave's cursor at once (512: a 1/8 stripe of an 8K frame is two chunks per warp, too coarse)
constexpr uint32_t kRoundBudget = 12; // node fetches per entry in the budgeted hand-over rounds (breadth first: see walk_trips)
constexpr uint32_t kNoItem = 0xFFFFFFFFu;

// The slab test of four child boxes at once: the canonical predicate of slab_test() per child (closed interval, IEEE 1/d,
// fminf/fmaxf), nothing shared between the children, so every child sees exactly the arithmetic of the binary walk.
__device__ __forceinline__ unsigned slab_test4(V3 o, V3 inv_d, const float4 lx, const float4 ly, const float4 lz, const float4 hx, const float4 hy,
                                               const float4 hz, float tmin, float tmax) {
    const float l0[3] = {lx.x, ly.x, lz.x}, h0[3] = {hx.x, hy.x, hz.x};
    const float l1[3] = {lx.y, ly.y, lz.y}, h1[3] = {hx.y, hy.y, hz.y};
    const float l2[3] = {lx.z, ly.z, lz.z}, h2[3] = {hx.z, hy.z, hz.z};
    const float l3[3] = {lx.w, ly.w, lz.w}, h3[3] = {hx.w, hy.w, hz.w};
    unsigned m = 0;
    if (slab_test(o, inv_d, l0, h0, tmin, tmax)) m |= 1u;
    if (slab_test(o, inv_d, l1, h1, tmin, tmax)) m |= 2u;
    if (slab_test(o, inv_d, l2, h2, tmin, tmax)) m |= 4u;
    if (slab_test(o, inv_d, l3, h3, tmin, tmax)) m |= 8u;
    return m;
}

// A warp's claim on the wave's work list: [next, end) is the chunk it is handing out to its lanes, [pre, pre_end) the
// one it asked for in advance.
struct WarpWork {
    uint32_t next = 0, end = 0, pre = 0, pre_end = 0;
    bool drained = false; // the cursor has passed the end of the list
};

__device__ __forceinline__ uint32_t chunk_size(uint32_t n_items, uint32_t base, uint32_t warps) {
    // guided: half of an even share of what is left, so the last chunks are small and the warps finish together
    const uint32_t left = n_items > base ? n_items - base : 0u;
    uint32_t c = left / (2u * warps);
    c = c < 32u ? 32u : (c > kChunkMax ? kChunkMax : c);
    return (c + 31u) & ~31u;
}

__device__ __forceinline__ void claim_chunk(WarpWork& w, uint32_t* cursor, uint32_t n_items, uint32_t warps, int lane) {
    if (w.drained) { w.pre = w.pre_end = 0; return; }
    uint32_t base = 0, size = chunk_size(n_items, w.end, warps);
    if (lane == 0) base = atomicAdd(cursor, size);
    base = __shfl_sync(kFull, base, 0);
    if (base >= n_items) { w.drained = true; w.pre = w.pre_end = 0; return; }
    w.pre = base;
    w.pre_end = min(base + size, n_items);
}

// Hands one item index to every lane with need == true while the list lasts (warp-uniform control flow).
__device__ __forceinline__ uint32_t take_items(WarpWork& w, bool need, uint32_t* cursor, uint32_t n_items, uint32_t warps, int lane) {
    uint32_t idx = kNoItem;
    unsigned m = __ballot_sync(kFull, need);
    while (m) {
        if (w.next >= w.end) {
            if (w.pre >= w.pre_end) claim_chunk(w, cursor, n_items, warps, lane);
            if (w.pre >= w.pre_end) break; // nothing left
            w.next = w.pre; w.end = w.pre_end;
            claim_chunk(w, cursor, n_items, warps, lane); // ask for the one after: its latency hides behind this chunk's walks
        }
        const uint32_t avail = w.end - w.next;
        const uint32_t rank = (uint32_t)__popc(m & ((1u << lane) - 1u));
        if (need && idx == kNoItem && rank < avail) idx = w.next + rank;
        w.next += min(avail, (uint32_t)__popc(m));
        m = __ballot_sync(kFull, need && idx == kNoItem);
    }
    return idx;
}
__device__ __forceinline__ bool work_left(const WarpWork& w) { return w.next < w.end || w.pre < w.pre_end || !w.drained; }

// One lane's walk state.
struct WalkLane {
    V3 oo, inv_d;     // ray in the BLAS object space of instance `inst`
    float tcur;
    uint32_t r;       // ray slot
    uint32_t inst;    // TLAS instance being walked (MULTI)
    uint32_t sp;      // stack entries
    uint32_t head;    // FIFO rounds: ring position of the oldest entry (0 in the depth-first kernels)
    uint32_t visits;  // node fetches of this item
    bool busy;
    bool unbounded;   // the hand-over queue was full: this item is walked to its end here
};

template <bool MULTI>
__device__ __forceinline__ void enter_instance(const SceneDev& sc, WalkLane& w, V3 o, V3 d, uint32_t k) {
    const float* w2o = (MULTI && k) ? sc.aabb_insts[k].w2o : sc.aabb_w2o;
    w.oo = xform_point_3x4(w2o, o);
    const V3 dd = xform_vector_3x4(w2o, d);
    w.inv_d = v3(1.0f / dd.x, 1.0f / dd.y, 1.0f / dd.z);
    w.inst = k;
}

// (ray, subtree) entries handed to the next round: the `n` stack entries (newest first) and / or one extra ref.  Returns how
// many were taken (the queue may be full: slots below its capacity are always written by whoever reserved them, what does
// not fit -- the oldest entries -- stays on the caller's stack).  MULTI packs the instance above the ray slot (kSpillRayBits).
template <bool MULTI>
__device__ __noinline__ uint32_t hand_over(const WaveBuffers& wb, uint2* q, uint32_t* q_count, const WalkLane& w, const uint32_t* stack, uint32_t n,
                                              uint32_t extra_ref, bool has_extra, uint32_t depth) {
    const uint32_t total = n + (has_extra ? 1u : 0u);
    if (!total || !q) return 0u;
    const uint32_t at = atomicAdd(q_count, total);
    const uint32_t fit = at >= wb.spill_capacity ? 0u : min(total, wb.spill_capacity - at);
    const uint32_t tag = MULTI ? (w.r | (w.inst << kSpillRayBits)) : w.r;
    for (uint32_t i = 0; i < fit; i++) q[at + i] = make_uint2(tag, i < n ? stack[(size_t)((w.head + (n - 1u - i)) % depth) * blockDim.x] : extra_ref); // newest first: what does not fit is the oldest, left in place
    return fit;
}

// This lane's walk of the subtrees on its stack, to the end or to the node budget.  stack: this thread's column of the
// shared-memory stack.
// FIFO (the budgeted hand-over rounds): the OLDEST entry is expanded first, so a big subtree is opened breadth first and what
// is handed on after kRoundBudget fetches are up to DEPTH pieces of similar size -- the longest chain shrinks by that factor
// per round, where handing over a depth-first stack (whose bottom entries are the big far subtrees) only peels it.
template <bool MULTI, int DEPTH, bool FIFO = false>
__device__ __forceinline__ void walk_trips(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, WaveCtrl* ctrl, const float4* s_top, uint32_t* stack,
                                           WalkLane& w, Pending& pend, uint32_t budget, uint2* q, uint32_t* q_count, int max_trips = 0x7FFFFFFF) {
    const uint32_t flat0 = MULTI ? w.inst * sc.n_prims : 0u;
#pragma unroll 1
    for (int trip = 0; trip < max_trips && w.sp != 0u; trip++) {
        if (w.visits >= budget && !w.unbounded) { // the rest of this walk goes to the next round
            const uint32_t took = hand_over<MULTI>(wb, q, q_count, w, stack, w.sp, 0u, false, FIFO ? DEPTH : 0x40000000u);
            if (took == w.sp) { w.sp = 0u; break; }
            // queue full: keep what did not fit (the oldest entries) and finish this item here, whatever it takes
            w.sp -= took;
            w.unbounded = true;
        }
        uint32_t ref;
        if (FIFO) { ref = stack[(size_t)w.head * blockDim.x]; w.head = (w.head + 1u) % DEPTH; w.sp--; }
        else ref = stack[(size_t)(--w.sp) * blockDim.x]; // depth first: head stays 0
        w.visits++;
        float4 lx, ly, lz, hx, hy, hz, rf;
        if (ref & kWideTopFlag) {
            const float4* nd = s_top + 8u * (ref & ~kWideTopFlag);
            lx = nd[0]; ly = nd[1]; lz = nd[2]; hx = nd[3]; hy = nd[4]; hz = nd[5]; rf = nd[6];
        } else {
            const float4* nd = sc.wide_nodes + 8 * (size_t)ref;
            lx = __ldg(nd); ly = __ldg(nd + 1); lz = __ldg(nd + 2); hx = __ldg(nd + 3); hy = __ldg(nd + 4); hz = __ldg(nd + 5); rf = __ldg(nd + 6);
        }
        const unsigned hit = slab_test4(w.oo, w.inv_d, lx, ly, lz, hx, hy, hz, 0.0f, w.tcur);
        const int32_t refs[4] = {__float_as_int(rf.x), __float_as_int(rf.y), __float_as_int(rf.z), __float_as_int(rf.w)};
#pragma unroll
        for (int c = 0; c < 4; c++) {
            if (!((hit >> c) & 1u) || refs[c] == kWideInvalid) continue;
            if (refs[c] < 0) add_leaf<MULTI>(sc, pp, wb, ctrl, w.r, flat0, (uint32_t)~refs[c], pend);
            else if (w.sp < (uint32_t)DEPTH) stack[(size_t)(FIFO ? (w.head + w.sp++) % DEPTH : w.sp++) * blockDim.x] = (uint32_t)refs[c];
            else if (!hand_over<MULTI>(wb, q, q_count, w, stack, 0u, (uint32_t)refs[c], true, FIFO ? DEPTH : 0x40000000u)) ctrl->overflow = 1u; // full stack: this subtree goes to the
                                                                                                                     // next round; nowhere to put it is an error, never a silent drop
        }
    }
}

// Ray setup of the main kernel: camera ray (primary) or the queued ray; the plane triangles; returns false when there is
// nothing to walk (a shadow ray the plane already occludes, no AABB instance).  o, d: the world-space ray.
template <bool PRIMARY, bool MULTI>
__device__ __forceinline__ bool begin_ray(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, int pq, uint32_t r, V3& o, V3& d) {
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
    float tcur = kRayTMax;
    unsigned long long key = kKeyMiss;
    plane_triangles<MULTI>(sc, o, d, shadow, tcur, key);
    wb.hitkey[r] = key;
    wb.ray_tcur[r] = tcur;
    if (shadow && key != kKeyMiss) return false; // ACCEPT_FIRST_HIT_AND_END_SEARCH
    return sc.has_aabb_instance && sc.n_prims != 0u;
}

// ---------------------------------------------------------------------------------------------
// k_walk: the main traversal kernel of a wave (rays) -- ROUND == 0 -- and its hand-over rounds (entries) -- ROUND > 0.
// Dynamic shared memory: [n_top * 8 float4 top array][DEPTH * blockDim.x stack words].
// ---------------------------------------------------------------------------------------------
template <bool PRIMARY, bool MULTI, bool ENTRIES, int DEPTH, bool FIFO_ROUND = false>
__device__ __forceinline__ void walk_body(const SceneDev& sc, const PassParams& pp, const WaveBuffers& wb, int wave, int round, bool last) {
    extern __shared__ float4 s_dyn[];
    __shared__ uint32_t s_cnt[kWalkBlock / 32][kNumTypes], s_base[kWalkBlock / 32][kNumTypes];
    float4* s_top = s_dyn;
    uint32_t* stack = reinterpret_cast<uint32_t*>(s_dyn + 8u * sc.n_top) + threadIdx.x;
    for (uint32_t i = threadIdx.x; i < 8u * sc.n_top; i += blockDim.x) s_top[i] = __ldg(sc.top_nodes + i);
    __syncthreads();

    WaveCtrl* ctrl = &wb.ctrl[wave];
    const int pq = wave & 1;
    const int lane = threadIdx.x & 31;
    const uint32_t warps = gridDim.x * (blockDim.x >> 5);
    uint32_t n_items, *cursor;
    const uint2* in = nullptr;
    uint2* out = nullptr;
    uint32_t* out_count = nullptr;
    if (!ENTRIES) {
        n_items = PRIMARY ? pp.n_pixels : min(ctrl->n_rays, pp.ray_capacity);
        cursor = &ctrl->trav_cursor;
        if (pp.spill_budget) { out = wb.spill[0]; out_count = &ctrl->spill_count[0]; }
    } else {
        in = wb.spill[(round - 1) & 1];
        n_items = min(ctrl->spill_count[round - 1], wb.spill_capacity);
        cursor = &ctrl->spill_cursor[round - 1];
        if (!last) { out = wb.spill[round & 1]; out_count = &ctrl->spill_count[round]; }
    }
    const uint32_t budget = (out && pp.spill_budget) ? pp.spill_budget : 0xFFFFFFFFu;
    const uint32_t n_ai = MULTI ? sc.n_aabb_insts : 1u;

    WarpWork work;
    VisitStats stats;
    if constexpr (ENTRIES) {
        // Hand-over entries are subtrees of every size (one node to hundreds), and an entry costs nothing to start (its ray
        // is two loads): here lanes ARE refilled one by one as their entry ends -- kRefillLanes idle lanes trigger a refill,
        // kEntryTrips fetches run between two checks.  (Measured on the 8K frame, wave 2: batches of 32 entries ran at a
        // lane efficiency of 0.26 / 0.28 / 0.17 in the three rounds.)
        constexpr int kRefillLanes = 8, kEntryTrips = 4;
        WalkLane w;
        w.busy = false; w.unbounded = false; w.sp = 0; w.head = 0; w.visits = 0; w.r = 0; w.inst = 0; w.tcur = 0.0f;
        w.oo = v3s(0.0f); w.inv_d = v3s(0.0f);
        Pending pend;
        pend.n = 0;
#pragma unroll
        for (int k = 0; k < kMaxPending; k++) pend.e[k] = 0u;
        for (;;) {
            if (w.busy && w.sp == 0u) { stats.add(w.visits); w.busy = false; }
            if (__popc(__ballot_sync(kFull, !w.busy)) >= kRefillLanes) {
                flush_pending(pp, wb, ctrl, w.r, pend, s_cnt, s_base); // convergent; every lane's candidates so far
                pend.n = 0;
                const uint32_t idx = take_items(work, !w.busy, cursor, n_items, warps, lane);
                if (idx != kNoItem) {
                    const uint2 e = in[idx];
                    w.r = MULTI ? (e.x & ((1u << kSpillRayBits) - 1u)) : e.x;
                    const V3 o = ld3(wb.ray_o[pq][w.r]), d = ld3(wb.ray_d[pq][w.r]);
                    enter_instance<MULTI>(sc, w, o, d, MULTI ? (e.x >> kSpillRayBits) : 0u);
                    w.tcur = wb.ray_tcur[w.r];
                    w.visits = 0; w.unbounded = false; w.busy = true;
                    stack[0] = e.y;
                    w.sp = 1u; w.head = 0u;
                }
                if (!__any_sync(kFull, w.busy)) {
                    if (!work_left(work)) break;
                    continue;
                }
            }
            walk_trips<MULTI, DEPTH, FIFO_ROUND>(sc, pp, wb, ctrl, s_top, stack, w, pend, FIFO_ROUND ? min(budget, kRoundBudget) : budget, out, out_count, kEntryTrips);
        }
    } else {
    // A batch of 32 consecutive rays per warp: neighbouring rays walk nearly the same nodes and take nearly the same
    // number of fetches (measured lane efficiency 31/32 on primary rays; refilling single lanes as they finish was tried
    // here too and lost to the divergent ray set-up it needs: 24/32), and the node budget bounds what the longest ray of a
    // batch can cost the other 31.
    for (;;) {
        const uint32_t idx = take_items(work, true, cursor, n_items, warps, lane);
        if (!__any_sync(kFull, idx != kNoItem)) break;
        WalkLane w;
        w.busy = false; w.unbounded = false; w.sp = 0; w.head = 0; w.visits = 0; w.r = 0; w.inst = 0; w.tcur = 0.0f;
        w.oo = v3s(0.0f); w.inv_d = v3s(0.0f);
        Pending pend;
        pend.n = 0;
#pragma unroll
        for (int k = 0; k < kMaxPending; k++) pend.e[k] = 0u;
        if (idx != kNoItem) {
            V3 o, d;
            w.r = idx;
            if (begin_ray<PRIMARY, MULTI>(sc, pp, wb, pq, idx, o, d)) {
                w.tcur = wb.ray_tcur[idx];
                w.busy = true;
                for (uint32_t k = 0; k < n_ai; k++) { // one walk of the shared BLAS per TLAS instance, in that instance's object space
                    enter_instance<MULTI>(sc, w, o, d, k);
                    stack[0] = kWideTopFlag; // the root is top[0]
                    w.sp = 1u; w.head = 0u;
                    walk_trips<MULTI, DEPTH>(sc, pp, wb, ctrl, s_top, stack, w, pend, budget, out, out_count);
                }
            }
        }
        flush_pending(pp, wb, ctrl, w.r, pend, s_cnt, s_base); // convergent
        stats.add(w.visits);
    }
    }
    stats.commit(ctrl);
}

template <bool PRIMARY, bool MULTI>
__global__ void __launch_bounds__(kWalkBlock, 3) k_walk(const __grid_constant__ SceneDev sc, const __grid_constant__ PassParams pp,
                                                        const __grid_constant__ WaveBuffers wb, int wave) {
    walk_body<PRIMARY, MULTI, false, kStackDepth>(sc, pp, wb, wave, 0, false);
}
// small scenes (no hand-over rounds): every ray walked to its end on the deep stack
template <bool PRIMARY, bool MULTI>
__global__ void __launch_bounds__(kWalkBlockLast, 6) k_walk_deep(const __grid_constant__ SceneDev sc, const __grid_constant__ PassParams pp,
                                                                 const __grid_constant__ WaveBuffers wb, int wave) {
    walk_body<PRIMARY, MULTI, false, kStackDepthLast>(sc, pp, wb, wave, 0, false);
}
template <bool MULTI>
__global__ void __launch_bounds__(kWalkBlock, 2) k_walk_round(const __grid_constant__ SceneDev sc, const __grid_constant__ PassParams pp,
                                                              const __grid_constant__ WaveBuffers wb, int wave, int round) {
    walk_body<false, MULTI, true, kStackDepth, true>(sc, pp, wb, wave, round, false);
}
template <bool MULTI>
__global__ void __launch_bounds__(kWalkBlockLast, 4) k_walk_last(const __grid_constant__ SceneDev sc, const __grid_constant__ PassParams pp,
                                                                 const __grid_constant__ WaveBuffers wb, int wave, int round) {
    walk_body<false, MULTI, true, kStackDepthLast>(sc, pp, wb, wave, round, true);
}

} // namespace
} // namespace gprt
