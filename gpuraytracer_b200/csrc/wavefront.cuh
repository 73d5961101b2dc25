// wavefront.cuh -- data layout shared by the wavefront kernels (kernels.cu, compiled once per
// math mode) and the host-side context (gprt_api.cu).
//
// The reference's DispatchRays runs one recursive shader invocation tree per pixel
// (src/Raytracing.hlsl:154-165 and the TraceRay calls under it).  Here the tree is flattened into
// `max_recursion_depth` WAVES.  Wave k holds the level-k radiance rays plus the level-(k-1)
// shadow rays; for each wave:
//   traverse      ray vs plane triangles, then the LBVH walk (per TLAS instance) -> (ray, primitive)
//                 candidates appended to one queue per primitive type; long walks are handed on to
//                 spill rounds after a node budget
//   intersect     one persistent launch: the type's intersection test on every candidate of
//                 every queue, 64-bit atomicMin of (t, primitive) per ray
//   resolve       the winning candidate of each ray publishes its normal
//   shade_spawn   stores the level's hit record per pixel, emits the next wave's shadow and
//                 reflection rays (ballot/popc compaction)
// and one final kernel folds the per-level hit records back to front into the pixel colour
// (exactly the order in which the recursive closest-hit shaders return).
#pragma once
#include <stdint.h>

#include "lbvh.h"
#include "shade.cuh"

namespace gprt {

constexpr int kNumTypes = 10; // 0 analytic AABB, 1 analytic spheres, 2 metaballs, 3..9 SDF types 0..6
constexpr int kMaxDepth = 8;  // waves; the reference uses 3
constexpr float kRayTMax = 10000.0f; // src/Raytracing.hlsl:99,126
constexpr uint32_t kCodeMiss = 0xFFFFFFFFu;
constexpr uint32_t kCodeNone = 0xFEFEFEFEu; // no ray traced at this level (memset pattern 0xFE)
constexpr uint32_t kCodePlane = 0x80000000u; // | triangle index
constexpr unsigned long long kKeyMiss = 0xFFFFFFFFFFFFFFFFull;

// ray meta word: pixel (26 bits) | level (4 bits, 1-based) | kind (1 bit: 1 = shadow)
constexpr uint32_t kMetaPixelMask = (1u << 26) - 1u;
__host__ __device__ inline uint32_t make_meta(uint32_t pixel, uint32_t level, uint32_t shadow) { return pixel | (level << 26) | (shadow << 30); }
__host__ __device__ inline uint32_t meta_pixel(uint32_t m) { return m & kMetaPixelMask; }
__host__ __device__ inline uint32_t meta_level(uint32_t m) { return (m >> 26) & 15u; }
__host__ __device__ inline uint32_t meta_shadow(uint32_t m) { return (m >> 30) & 1u; }

struct PrimRecord { // per geometry index; 48 bytes
    Material mat;            // 36 B
    uint32_t type_id;        // 0..kNumTypes-1
    uint32_t instance_index; // into the per-frame transform buffer
    uint32_t _pad;
};

// One TLAS instance (two-level instancing): the extra instances of a BLAS beyond the first of its kind.
struct InstDev {
    float w2o[12]; // world -> BLAS object space (3x4), inverted on the host in double
    float o2w[12]; // ObjectToWorld3x4()
};
constexpr uint32_t kMaxInstancesPerKind = 256; // 8 bits of a spill entry carry the AABB instance
constexpr uint32_t kSpillRayBits = 23;         // multi-instance scenes: ray slots per pass < 2^23 (gprt_api.cu: kMaxPassPixelsMulti = 2^22)

struct SceneDev {
    const float4* bvh_nodes; // binary LBVH, 4 x float4 per internal node: child0 min|max, child1 min|max, refs (build / refit only)
    const float4* wide_nodes; // what the walk reads: 8 x float4 per node, four children (lbvh.cu, k_collapse4)
    const float4* top_nodes;  // the first n_top wide nodes breadth first, refs into itself: staged in shared memory by every walk block
    uint32_t n_top;
    int32_t root_ref;        // >= 0 internal node, < 0 leaf (~primitive), INT32_MIN: no primitives
    uint32_t n_prims;
    const float* aabbs;      // 6 floats per primitive (BLAS object space)
    const PrimRecord* prims;
    const float* per_frame;  // 32 floats per instance_index: local->BLAS 4x4, BLAS->local 4x4
    const float* tri_pos;    // 9 floats per triangle: v0, v1, v2 (object space)
    const float* tri_nrm;    // 3 floats per triangle: normal of the vertex indices[0] names
    uint32_t n_tris;
    uint32_t has_aabb_instance;
    float tri_w2o[12];       // world -> triangle BLAS object space (3x4)
    float aabb_w2o[12];      // world -> AABB BLAS object space
    float aabb_o2w[12];      // ObjectToWorld3x4() of the AABB instance
    Material plane_mat;
    // two-level instancing: n_*_insts TLAS instances of each BLAS.  Instance 0 of each kind is the
    // set of fields above; candidates, hit codes and `prims` are flattened as instance * n_prims + primitive
    // (triangles: instance * n_tris + triangle), so everything downstream of the traversal indexes one table.
    uint32_t n_tri_insts, n_aabb_insts;
    const InstDev* tri_insts;    // [n_tri_insts]
    const InstDev* aabb_insts;   // [n_aabb_insts]
    const Material* plane_mats;  // [n_tri_insts]: the triangle hit group of each instance's contribution
    uint32_t leaf_types_valid;   // the type tag in the leaf refs holds for every AABB instance
    unsigned long long inv_n_prims; // floor(2^64 / n_prims) + 1: flat / n_prims == __umul64hi(flat, inv_n_prims) for flat < 2^27
};

constexpr int kSpillRounds = 3; // traversal hand-over rounds after the main walk (the last one has no node budget)

struct WaveCtrl { // one per wave, zeroed at the start of a pass
    uint32_t n_rays;
    uint32_t n_radiance, n_shadow;
    uint32_t cand_count[kNumTypes];
    uint32_t cursor[kNumTypes];
    uint32_t resolve_cursor;
    uint32_t overflow;
    uint32_t work[kNumTypes];   // per type: analytic tests / metaball march steps / SDF distance evaluations
    uint32_t max_evals[kNumTypes];      // profiling: most field evaluations spent on one candidate
    uint32_t evals_hist[kNumTypes][12]; // profiling: candidates by floor(log2(evaluations)) (bin 0: none)
    uint32_t trav_cursor;       // next unclaimed ray of the wave's traversal (warps claim batches)
    uint32_t node_visits, max_node_visits; // LBVH nodes fetched by the wave; by its longest ray
    uint32_t spill_count[kSpillRounds];    // (ray, subtree) entries handed to spill round i + 1
    uint32_t spill_cursor[kSpillRounds];   // next unclaimed entry of that round
};

struct PassParams {
    FrameConstants fc;
    uint32_t tile_x, tile_y, tile_w, tile_h; // this pass's pixel rectangle (global coordinates)
    uint32_t n_pixels;                       // tile_w * tile_h
    uint32_t max_depth, metaball_max_steps, sdf_max_steps;
    uint32_t fractal_iterations;             // FRACTAL_ITERATIONS_COUNT (src/ConstantBuffers.h:27)
    uint32_t ray_capacity;                   // 2 * n_pixels
    uint32_t cand_capacity;                  // per type
    uint32_t brute_force;
    uint32_t spill_budget;                   // node fetches per thread before a walk is handed on; 0 = never (small scenes)
    // multi-GPU row striping: with stripe_world > 1 the pass covers LOCAL rows tile_y .. tile_y+tile_h of
    // rank stripe_rank, local row l being global row ((l / stripe_rows) * stripe_world + stripe_rank) *
    // stripe_rows + l % stripe_rows.  DispatchRaysIndex / Dimensions stay global either way.
    uint32_t stripe_rows, stripe_rank, stripe_world;
};

__host__ __device__ inline uint32_t pass_global_row(const PassParams& pp, uint32_t ty) {
    const uint32_t l = pp.tile_y + ty;
    if (pp.stripe_world <= 1u) return l;
    return ((l / pp.stripe_rows) * pp.stripe_world + pp.stripe_rank) * pp.stripe_rows + l % pp.stripe_rows;
}

struct WaveBuffers {
    float4* ray_o[2];  // ping-pong by wave parity: xyz origin
    float4* ray_d[2];  // xyz direction, w = meta bits
    float* ray_tcur;   // RayTCurrent() seen by the intersection tests (min(TMax, plane t))
    unsigned long long* hitkey; // (float_bits(t) << 32) | code, atomicMin
    float4* hit_n;     // world normal of the winning procedural hit
    uint2* cand[kNumTypes];      // (ray slot, primitive)
    float4* cand_res[kNumTypes]; // world normal xyz, w = t (or -1: no hit)
    WaveCtrl* ctrl;    // [kMaxDepth + 1]
    uint2* spill[2];   // traversal hand-over queues (ray slot, subtree root), ping-pong by round
    uint32_t spill_capacity;

    // per-level, per-pixel hit records: index (level - 1) * n_pixels + pixel
    float4* lvl_o_t;   // ray origin xyz, committed t
    float4* lvl_d;     // ray direction xyz
    float4* lvl_n;     // world normal xyz
    uint32_t* lvl_code;
    uint8_t* lvl_shadow;
};

struct OutputBuffers {
    uint8_t* rgba8;    // tile row y at rgba8 + y * pitch
    size_t pitch;
    uint32_t* hit_ids; // optional, tile-packed
    float4* rgba_f32;  // optional, tile-packed
    uint32_t out_x0, out_y0; // position of this pass inside the caller's tile
    uint32_t out_w;          // width of the caller's tile (for the packed optional outputs)
};

// Launcher table exported by each compilation of kernels.cu.
struct KernelTable {
    // all launchers enqueue on `stream` and return the number of kernels launched
    int (*traverse)(const SceneDev&, const PassParams&, const WaveBuffers&, int wave, void* stream);
    int (*intersect)(const SceneDev&, const PassParams&, const WaveBuffers&, int wave, int type_id, int metaballs, void* stream);
    // every type of `type_mask` in one persistent launch (the production path; `intersect` per type is the profiling path)
    // blocks_per_sm: 0 = a full resident wave of blocks; k = at most k per SM (two launches sharing the machine)
    int (*intersect_all)(const SceneDev&, const PassParams&, const WaveBuffers&, int wave, uint32_t type_mask, int metaballs, void* stream, int blocks_per_sm);
    int (*resolve)(const SceneDev&, const PassParams&, const WaveBuffers&, int wave, uint32_t type_mask, void* stream);
    int (*shade_spawn)(const SceneDev&, const PassParams&, const WaveBuffers&, int wave, void* stream);
    int (*final_color)(const SceneDev&, const PassParams&, const WaveBuffers&, const OutputBuffers&, void* stream);
    // the whole frame in one kernel, one thread per pixel (GPRT_FLAG_MEGAKERNEL: the measured comparison point)
    int (*megakernel)(const SceneDev&, const PassParams&, const WaveBuffers&, const OutputBuffers&, int metaballs, void* stream);
    const char* name;
};

const KernelTable* kernels_parity();
const KernelTable* kernels_fast();
const KernelTable* kernels_fma(); // FMA contraction, IEEE div/sqrt, exact transcendentals

} // namespace gprt
