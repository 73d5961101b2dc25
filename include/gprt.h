/*
 * gprt.h -- C ABI of the B200-native replacement for GPURaytracer's DXR hot path.
 *
 * The reference has no plugin/FFI interface: the hot path sits behind two private
 * members of `Renderer` (all citations relative to the reference tree):
 *   - Renderer::build_acceleration_structures()   src/Renderer.cpp:788-835  (once, blocking)
 *   - Renderer::do_raytracing()                   src/Renderer.cpp:401-452  (per frame)
 * and their inputs are `Renderer` members.  This header restates those inputs as
 * plain C structs with the SAME byte layout the reference uploads to the GPU, so a
 * scene built by the reference drops in unchanged.
 *
 * No torch types, no C++ types, no exceptions: every entry point returns 0 or a
 * negative gprt_status.  One gprt_ctx per GPU; a context is not thread-safe; all
 * work is ordered on the context's CUDA stream.
 */
#ifndef GPRT_H
#define GPRT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GPRT_ABI_VERSION 2

/* ------------------------------------------------------------------------------------------
 * Scene / shader ABI (byte-for-byte the structs of src/ConstantBuffers.h; matrices are
 * row-major and used with the row-vector convention mul(v, M), as the HLSL is built /Zpr).
 * ---------------------------------------------------------------------------------------- */

/* SceneConstantBuffer, src/ConstantBuffers.h:49-58.  136 bytes used, C++ sizeof 144. */
typedef struct gprt_scene_cb {
    float projection_to_world[16]; /* @0   inverse(view*proj), row-major            */
    float camera_position[4];      /* @64                                           */
    float light_position[4];       /* @80                                           */
    float light_ambient_color[4];  /* @96                                           */
    float light_diffuse_color[4];  /* @112                                          */
    float reflectance;             /* @128 never read by the shaders                */
    float elapsed_time;            /* @132 drives the metaball animation            */
    float _pad[2];                 /* @136 XMMATRIX alignment tail                  */
} gprt_scene_cb;

/* PrimitiveConstantBuffer (material), src/ConstantBuffers.h:61-71.  48 bytes. */
typedef struct gprt_material {
    float albedo[4];
    float reflectance_coefficient;
    float diffuse_coefficient;
    float specular_coefficient;
    float specular_power;
    float step_scale; /* sphere-tracing step scale of SDF primitives */
    float padding[3];
} gprt_material;

/* PrimitiveInstanceConstantBuffer, src/ConstantBuffers.h:74-78.  8 bytes. */
typedef struct gprt_prim_instance_cb {
    uint32_t instance_index; /* index into the per-frame transform buffer */
    uint32_t primitive_type; /* enum inside the primitive's class         */
} gprt_prim_instance_cb;

/* PrimitiveInstancePerFrameBuffer, src/ConstantBuffers.h:81-85.  128 bytes. */
typedef struct gprt_prim_per_frame {
    float local_space_to_bottom_level_as[16];
    float bottom_level_as_to_local_space[16];
} gprt_prim_per_frame;

/* Vertex, src/ConstantBuffers.h:87-91.  24 bytes. */
typedef struct gprt_vertex {
    float position[3];
    float normal[3];
} gprt_vertex;

/* D3D12_RAYTRACING_AABB as filled at src/Renderer.cpp:499-503.  24 bytes, BLAS object space. */
typedef struct gprt_aabb {
    float min_x, min_y, min_z, max_x, max_y, max_z;
} gprt_aabb;

/* D3D12_RAYTRACING_INSTANCE_DESC as filled at src/Renderer.cpp:732-786.  64 bytes. */
typedef struct gprt_instance_desc {
    float    transform[3][4];            /* object -> world, 3x4 (column-vector form)   */
    uint32_t instance_id_and_mask;       /* InstanceID:24 | InstanceMask:8              */
    uint32_t hit_group_contrib_and_flags;/* InstanceContributionToHitGroupIndex:24 | Flags:8 */
    uint64_t acceleration_structure;     /* BLAS handle: GPRT_BLAS_TRIANGLES / GPRT_BLAS_AABBS */
} gprt_instance_desc;

/* What the 32-byte D3D12 shader identifier of a hit-group record stands for.  The reference
 * gets the identifier from ID3D12StateObjectProperties::GetShaderIdentifier
 * (src/Renderer.cpp:852-876); here the first two words name the hit group directly. */
enum gprt_hit_group_kind {
    GPRT_HG_TRIANGLE        = 0, /* MyHitGroup_Triangle[_ShadowRay]                   */
    GPRT_HG_AABB_ANALYTIC   = 1, /* MyHitGroup_AABB_AnalyticPrimitive[_ShadowRay]     */
    GPRT_HG_AABB_VOLUMETRIC = 2, /* MyHitGroup_AABB_VolumetricPrimitive[_ShadowRay]   */
    GPRT_HG_AABB_SDF        = 3  /* MyHitGroup_AABB_SignedDistancePrimitive[_ShadowRay] */
};

/* One hit-group shader record, laid out like the reference's (src/Renderer.cpp:938-979):
 * 32 B identifier + LocalRootSignature::AABB::RootArguments (48 + 8) -> 88, stride 96.
 * Record index = rayType + 2*geometryIndex + InstanceContributionToHitGroupIndex. */
typedef struct gprt_hit_group_record {
    uint32_t              shader_id[8]; /* [0] gprt_hit_group_kind, [1] ray type, rest 0 */
    gprt_material         material_cb;
    gprt_prim_instance_cb aabb_cb;      /* unused for the triangle records */
    uint32_t              _pad[2];
} gprt_hit_group_record;

#define GPRT_BLAS_TRIANGLES 0ull
#define GPRT_BLAS_AABBS     1ull

/* Enumerations of src/ConstantBuffers.h:137-176 and src/RaytracingSceneDefines.h:111-117. */
enum { GPRT_ANALYTIC_AABB = 0, GPRT_ANALYTIC_SPHERES = 1, GPRT_ANALYTIC_COUNT = 2 };
enum { GPRT_VOLUMETRIC_METABALLS = 0, GPRT_VOLUMETRIC_COUNT = 1 };
enum {
    GPRT_SDF_MINI_SPHERES = 0, GPRT_SDF_INTERSECTED_ROUND_CUBE, GPRT_SDF_SQUARE_TORUS,
    GPRT_SDF_TWISTED_TORUS, GPRT_SDF_COG, GPRT_SDF_CYLINDER, GPRT_SDF_FRACTAL_PYRAMID,
    GPRT_SDF_COUNT
};

/* ------------------------------------------------------------------------------------------
 * Library configuration: the reference's compile-time knobs as run-time parameters.
 * ---------------------------------------------------------------------------------------- */
typedef struct gprt_config {
    uint32_t abi_version;          /* GPRT_ABI_VERSION                                            */
    uint32_t max_recursion_depth;  /* MAX_RAY_RECURSION_DEPTH, src/ConstantBuffers.h:31 (3)       */
    uint32_t metaballs_count;      /* METABALLS_COUNT, src/ConstantBuffers.h:12 (3 or 5)          */
    uint32_t metaball_max_steps;   /* MAX_STEPS, src/VolumetricPrimitives.hlsli:160 (128)         */
    uint32_t sdf_max_steps;        /* MaxSteps, src/SignedDistancePrimitives.hlsli:291 (512)      */
    uint32_t math_mode;            /* GPRT_MATH_PARITY, GPRT_MATH_FMA or GPRT_MATH_FAST           */
    uint32_t cand_per_ray;         /* candidate-queue capacity per ray slot (0 = default 2)       */
    uint32_t flags;                /* GPRT_FLAG_*                                                 */
    uint32_t fractal_iterations;   /* FRACTAL_ITERATIONS_COUNT, src/ConstantBuffers.h:27 (4; 0 = 4, at most 8) */
} gprt_config;

enum { GPRT_MATH_PARITY = 0, /* IEEE mul/add/div/sqrt, no FMA contraction */
       GPRT_MATH_FAST   = 1, /* FMA contraction + fast intrinsics (what DXIL `fast` allows) */
       GPRT_MATH_FMA    = 2  /* FMA contraction only: IEEE div/sqrt, no flush-to-zero, exact transcendentals (march kernels) */ };
enum { GPRT_FLAG_BRUTE_FORCE_TRAVERSAL = 1u, /* test every AABB instead of walking the LBVH */
       GPRT_FLAG_PROFILE_KERNELS = 2u,       /* CUDA events around every kernel (gprt_get_kernel_profile) */
       GPRT_FLAG_MEGAKERNEL = 4u             /* the whole frame in ONE kernel, one thread per pixel, no queues: the comparison point
                                                beside the wavefront form (same image bit for bit; scenes of <= 64 primitives with one
                                                instance per BLAS, GPRT_E_UNSUPPORTED otherwise) */ };

typedef struct gprt_rect { uint32_t x, y, w, h; } gprt_rect; /* pixel tile, (0,0) = top-left */

/* Per-dispatch work counters (device-side, read back by gprt_get_counters). */
typedef struct gprt_counters {
    uint64_t radiance_rays;
    uint64_t shadow_rays;
    uint64_t candidates[3];  /* intersection-shader invocations: analytic, volumetric, SDF */
    uint64_t bvh_node_visits;
    uint64_t queue_overflow; /* non-zero: a queue overflowed and the image is incomplete   */
} gprt_counters;

/* Per-kernel device time of the last dispatch (only with GPRT_FLAG_PROFILE_KERNELS) and the work
 * each intersection kernel did (always counted). */
enum { GPRT_K_TRAVERSE = 0, GPRT_K_INTERSECT0 = 1 /* + type id 0..9: one launch per type, only with GPRT_PER_TYPE_LAUNCHES=1 */,
       GPRT_K_RESOLVE = 11, GPRT_K_SHADE_SPAWN = 12, GPRT_K_FINAL = 13,
       GPRT_K_INTERSECT_ALL = 14 /* the production intersection kernel: all types in one persistent launch per wave */,
       GPRT_K_COUNT = 15 };
typedef struct gprt_kernel_profile {
    float    ms[GPRT_K_COUNT];
    uint32_t launches[GPRT_K_COUNT];
    uint64_t work[10];       /* per type id: analytic tests / metaball march steps / SDF distance evaluations */
    uint64_t candidates[10]; /* per type id: intersection-shader invocations */
} gprt_kernel_profile;

/* Per-wave statistics of the last dispatch (wave k = level-k radiance rays + level-(k-1) shadow
 * rays).  max_evals / evals_hist are filled by the one-candidate-per-thread intersection kernels
 * only (GPRT_FLAG_PROFILE_KERNELS with GPRT_ISECT_VARIANT=-1). */
typedef struct gprt_wave_stats {
    uint32_t n_radiance, n_shadow;
    uint32_t candidates[10];
    uint32_t work[10];
    uint32_t max_evals[10];
    uint32_t evals_hist[10][12]; /* candidates by floor(log2(field evaluations)) + 1; bin 0 = none */
    uint32_t node_visits;        /* LBVH nodes fetched by the wave's traversal (first band)            */
    uint32_t max_node_visits;    /* ... by its single longest ray: the traversal's serial critical path */
} gprt_wave_stats;

typedef struct gprt_timing {
    float dispatch_ms;    /* CUDA events around the last gprt_dispatch only (== DX::GPUTimer) */
    float build_accel_ms; /* last gprt_build_accel                                            */
} gprt_timing;

typedef enum gprt_status {
    GPRT_OK                = 0,
    GPRT_E_INVALID_ARG     = -1,
    GPRT_E_CUDA            = -2, /* a CUDA runtime call failed; gprt_last_cuda_error has the code */
    GPRT_E_NO_DEVICE       = -3, /* no usable sm_100 device: there is NO CPU fallback             */
    GPRT_E_OUT_OF_MEMORY   = -4,
    GPRT_E_NOT_BUILT       = -5, /* dispatch before gprt_build_accel                              */
    GPRT_E_QUEUE_OVERFLOW  = -6, /* raise gprt_config.cand_per_ray or render in tiles            */
    GPRT_E_UNSUPPORTED     = -7,
    GPRT_E_ABI_MISMATCH    = -8,
    GPRT_E_TOO_LARGE       = -9  /* beyond the fixed scratch of a no-allocation entry point (GPRT_NOALLOC_MAX_PRIMS) */
} gprt_status;

typedef struct gprt_ctx gprt_ctx;

/* ------------------------------------------------------------------------------------------
 * Entry points
 * ---------------------------------------------------------------------------------------- */

/* Fill cfg with the reference's compile-time values (depth 3, 3 metaballs, 128/512 steps). */
void gprt_default_config(gprt_config* cfg);

/* Replaces DeviceResources device creation for a headless CUDA context
 * (src/Renderer.cpp:60-71).  Fails with GPRT_E_NO_DEVICE when there is no GPU. */
int gprt_create(gprt_ctx** out, int device_ordinal, const gprt_config* cfg);
void gprt_destroy(gprt_ctx* ctx);

/* Replaces Renderer::build_geometry + the geometry/instance descs of
 * build_geometry_descs_for_bottom_level_as / build_bottom_level_as_instance_descs
 * (src/Renderer.cpp:478-565, 567-612, 732-786).  Host arrays are copied during the call.
 * instances[i].acceleration_structure selects GPRT_BLAS_TRIANGLES or GPRT_BLAS_AABBS.  The
 * reference's TLAS is exactly {triangle BLAS instance, AABB BLAS instance}; any number of instances
 * of either BLAS (up to 256 of a kind) is accepted, each with its own 3x4 transform and
 * InstanceContributionToHitGroupIndex (two-level instancing).  Instances of a kind are numbered in
 * array order; instance_count * primitive_count must stay below 2^27. */
int gprt_set_geometry(gprt_ctx* ctx,
                      const gprt_aabb* aabbs, uint32_t n_aabbs,
                      const gprt_vertex* vertices, uint32_t n_vertices,
                      const uint16_t* indices, uint32_t n_indices,
                      const gprt_instance_desc* instances, uint32_t n_instances);

/* Replaces Renderer::build_shader_tables (src/Renderer.cpp:839-985): the hit-group table,
 * `n_records` records of `stride_bytes` (>= sizeof(gprt_hit_group_record)). */
int gprt_set_hit_group_table(gprt_ctx* ctx, const void* records, uint32_t n_records,
                             uint32_t stride_bytes);

/* Replaces Renderer::build_acceleration_structures (src/Renderer.cpp:788-835): builds the
 * LBVH over the AABBs on the GPU (Morton codes, radix sort, Karras tree, refit).  Blocking. */
int gprt_build_accel(gprt_ctx* ctx);

/* Refit node boxes after the AABBs moved (the reference never does this). */
int gprt_refit_accel(gprt_ctx* ctx, const gprt_aabb* aabbs, uint32_t n_aabbs);

/* Replaces Renderer::do_raytracing (src/Renderer.cpp:401-452): uploads the scene constant
 * buffer and the n per-frame transforms, then runs DispatchRays(width, height, 1) restricted to
 * `tile` (DispatchRaysIndex/Dimensions stay global).  Asynchronous on the context stream.
 *   d_rgba8      device pointer, R8G8B8A8_UNORM, row `y` of the tile at d_rgba8 + y*pitch_bytes
 *   d_hit_ids    optional device pointer, tile.w*tile.h uint32: primary hit id
 *                (GPRT_HIT_MISS, GPRT_HIT_TRIANGLE|tri, GPRT_HIT_AABB|geometryIndex)
 *   d_rgba_f32   optional device pointer, tile.w*tile.h float4: colour before UNORM conversion */
int gprt_dispatch(gprt_ctx* ctx, const gprt_scene_cb* scene,
                  const gprt_prim_per_frame* per_frame, uint32_t n_per_frame,
                  uint32_t width, uint32_t height, gprt_rect tile,
                  void* d_rgba8, size_t pitch_bytes,
                  uint32_t* d_hit_ids, float* d_rgba_f32);

/* Same, with HOST output: the library renders into its own device framebuffer and copies the
 * tile into h_rgba8 (pinned or pageable) on the context stream; returns after the copy
 * completed.  This is the end-to-end call (host in, host out). */
int gprt_dispatch_host(gprt_ctx* ctx, const gprt_scene_cb* scene,
                       const gprt_prim_per_frame* per_frame, uint32_t n_per_frame,
                       uint32_t width, uint32_t height, gprt_rect tile,
                       void* h_rgba8, size_t pitch_bytes, uint32_t* h_hit_ids);

/* Static inputs of Renderer::update_aabb_primitive_attributes (src/Renderer.cpp:302-356) for primitive i
 * (i == its instance_index): transform(t) = scaling(scale) * (rotates ? rotationY(-2 t) : I) *
 * translation(centre of AABB i).  16 bytes. */
typedef struct gprt_prim_anim {
    float    scale[3];
    uint32_t rotates;
} gprt_prim_anim;

/* Keep the per-frame transforms on the device: upload the animation inputs once ... */
int gprt_set_animation(gprt_ctx* ctx, const gprt_prim_anim* anim, uint32_t n);
/* ... then, per frame, compute all n PrimitiveInstancePerFrameBuffer entries on the GPU (same arithmetic
 * as the host path, bit-identical matrices) instead of uploading 128 bytes per primitive.  A following
 * gprt_dispatch* call with per_frame == NULL uses them.  Asynchronous on the context stream. */
int gprt_update_transforms(gprt_ctx* ctx, float animation_time);
/* Read the device-resident transforms back (tests). */
int gprt_get_transforms(gprt_ctx* ctx, gprt_prim_per_frame* out, uint32_t n);

/* Multi-GPU sharding (no counterpart in the single-GPU reference): the frame is cut into stripes of
 * `stripe_rows` rows dealt round-robin to `world` ranks; rank r renders stripes r, r + world, ...
 * into a PACKED buffer of gprt_stripe_local_rows() rows (local row l at d_rgba8 + l * pitch_bytes,
 * local row l = global row ((l / stripe_rows) * world + r) * stripe_rows + l % stripe_rows).
 * DispatchRaysIndex / DispatchRaysDimensions stay global, so the union of all ranks' rows is
 * bit-identical to one full-frame gprt_dispatch.  Asynchronous on the context stream. */
int gprt_dispatch_stripes(gprt_ctx* ctx, const gprt_scene_cb* scene,
                          const gprt_prim_per_frame* per_frame, uint32_t n_per_frame,
                          uint32_t width, uint32_t height,
                          uint32_t stripe_rows, uint32_t rank, uint32_t world,
                          void* d_rgba8, size_t pitch_bytes, uint32_t* d_hit_ids);
uint32_t gprt_stripe_local_rows(uint32_t height, uint32_t stripe_rows, uint32_t rank, uint32_t world);

/* gprt_dispatch_host without the final wait: the copy to h_rgba8 (pinned memory for a truly
 * asynchronous copy) is enqueued on the context stream and the call returns; gprt_sync() completes
 * it.  With one context per frame in flight this is the reference's frame pipelining
 * (frame_count = 3, src/Renderer.h:92; DeviceResources::move_to_next_frame). */
int gprt_dispatch_host_async(gprt_ctx* ctx, const gprt_scene_cb* scene,
                             const gprt_prim_per_frame* per_frame, uint32_t n_per_frame,
                             uint32_t width, uint32_t height, gprt_rect tile,
                             void* h_rgba8, size_t pitch_bytes, uint32_t* h_hit_ids);
/* Multi-GPU with a shared full-frame output: renders this rank's stripes (gprt_dispatch_stripes) and copies every row to
 * its GLOBAL row of the caller's frame h_frame_rgba8 (row y at h_frame_rgba8 + y * frame_pitch_bytes), asynchronously on
 * the context stream.  The frame may be page-locked host memory shared by all ranks -- gprt_alloc_host in one process,
 * shared memory registered with cudaHostRegister in each process: every GPU then delivers its rows over its own PCIe link
 * and nothing is gathered through one device -- or device memory of ANY GPU of the box (unified addressing; a peer copy
 * over NVLink after gprt_enable_peer_access).  gprt_sync() completes it. */
int gprt_dispatch_stripes_host_async(gprt_ctx* ctx, const gprt_scene_cb* scene,
                                     const gprt_prim_per_frame* per_frame, uint32_t n_per_frame,
                                     uint32_t width, uint32_t height,
                                     uint32_t stripe_rows, uint32_t rank, uint32_t world,
                                     void* h_frame_rgba8, size_t frame_pitch_bytes);
/* Let this context's device read / write memory of peer_device directly (cudaDeviceEnablePeerAccess; already enabled is
 * not an error).  GPRT_E_UNSUPPORTED when the two devices cannot reach each other: copies then go through the host. */
int gprt_enable_peer_access(gprt_ctx* ctx, int peer_device);
/* The context's CUDA stream (cudaStream_t as void*). */
void* gprt_get_stream(gprt_ctx* ctx);

#define GPRT_HIT_MISS     0xFFFFFFFFu
#define GPRT_HIT_TRIANGLE 0x00000000u /* | triangle instance ordinal * triangle count + PrimitiveIndex() */
#define GPRT_HIT_AABB     0x10000000u /* | AABB instance ordinal * AABB count + GeometryIndex            */

/* Waits for the context stream; returns GPRT_E_QUEUE_OVERFLOW if a queue of the last dispatch overflowed (the image is
 * then incomplete). */
int gprt_sync(gprt_ctx* ctx);
/* The same check without waiting, for callers of the asynchronous device-output gprt_dispatch that synchronise their own
 * stream: GPRT_OK while the dispatch is still running or after it finished cleanly, GPRT_E_QUEUE_OVERFLOW once it has
 * finished with an overflow.  An unseen overflow is also returned (once) by the next gprt_dispatch* on the context. */
int gprt_poll_status(gprt_ctx* ctx);
int gprt_get_counters(gprt_ctx* ctx, gprt_counters* out); /* of the last dispatch; syncs */
int gprt_last_timing(gprt_ctx* ctx, gprt_timing* out);    /* syncs                        */
int gprt_get_kernel_profile(gprt_ctx* ctx, gprt_kernel_profile* out); /* syncs */
int gprt_get_wave_stats(gprt_ctx* ctx, uint32_t wave, gprt_wave_stats* out); /* wave 1..depth; first band; syncs */
int gprt_last_cuda_error(gprt_ctx* ctx);
/* FP32 FMA issue-rate microbenchmark (TFLOP/s, 2 flop per FMA) on the context's device: the
 * measured denominator of the FP32 roofline.  Blocking; a few milliseconds. */
int gprt_measure_fp32_peak(gprt_ctx* ctx, float* tflops_out);
/* Use an existing CUDA stream (cudaStream_t as void*) for all work of this context. */
int gprt_set_stream(gprt_ctx* ctx, void* cuda_stream);
/* Number of kernels launched by the last dispatch (for bench.py's gpu_launches). */
int gprt_last_launch_count(gprt_ctx* ctx);

/* Page-locked host memory for the frames of gprt_dispatch_host_async / the *_async renderer calls (cudaHostAlloc /
 * cudaFreeHost), so that a host without the CUDA toolkit can keep copies asynchronous.  NULL on failure. */
void* gprt_alloc_host(size_t bytes);
void gprt_free_host(void* p);

/* One process per GPU: a DEVICE frame of one rank as peer memory of the others, so that every rank's rows go straight to
 * their place in rank 0's frame over NVLink (gprt_dispatch_stripes_host_async with the mapped pointer as h_frame_rgba8)
 * and nothing is gathered or composed afterwards.  The owner allocates with gprt_ipc_alloc and hands the 64 opaque bytes
 * to the other processes of the box by any means; they map it with gprt_ipc_open (peer access is enabled on first use)
 * and unmap it with gprt_ipc_close before the owner calls gprt_ipc_free.  Ordering between the ranks is the caller's:
 * the rows are in place once every rank's context stream has passed its copy (a barrier among the ranks' streams).
 * GPRT_E_UNSUPPORTED where the driver has no inter-process memory handles (cudaIpcGetMemHandle). */
typedef struct gprt_ipc_handle { unsigned char bytes[64]; } gprt_ipc_handle;
int gprt_ipc_alloc(int device_ordinal, size_t bytes, void** d_ptr, gprt_ipc_handle* handle);
int gprt_ipc_open(int device_ordinal, const gprt_ipc_handle* handle, void** d_ptr);
int gprt_ipc_close(int device_ordinal, void* d_ptr);
int gprt_ipc_free(int device_ordinal, void* d_ptr);

/* ---- LBVH build that allocates nothing --------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same builder as gprt_build_accel
 * (Morton codes, radix sort, Karras tree, box refit -- gprt::lbvh_build), over CALLER-owned device arrays, with its
 * scratch (sort keys / values, radix histograms, child / parent links, refit flags, node boxes) in a fixed array in the
 * library's device image instead of the context's cudaMalloc'd block.
 *   d_aabbs         in:  n_aabbs x 6 floats (gprt_aabb), device memory
 *   d_prim_tags     in:  optional (NULL: none) one byte per primitive, its low 4 bits stored in bits 27..30 of the
 *                        primitive's leaf reference (the context passes the primitive type id here)
 *   d_nodes         out: (n_aabbs - 1) x 64 bytes, the binary nodes (4 float4 each: two child boxes + two child refs)
 *   d_sorted_prims  out: n_aabbs x uint32, the primitive index of every Morton-sorted leaf
 *   root_ref        out (host): 0 = node 0 is the root, ~0 = a single leaf, INT32_MIN = empty
 * Asynchronous on `cuda_stream` (cudaStream_t as void*, NULL = the default stream); the call makes `device_ordinal`
 * current.  It calls no cudaMalloc / cudaFree or any allocating routine; the first call on a device looks the scratch
 * array's address up once (cudaGetSymbolAddress), so a caller that captures graphs makes one call before capturing.
 * Limits: n_aabbs <= GPRT_NOALLOC_MAX_PRIMS, else GPRT_E_TOO_LARGE and nothing is launched.  The scratch is
 * GPRT_NOALLOC_SCRATCH_BYTES of device memory on EVERY GPU that loads the library, held until the process exits.
 * ALL calls on a device share that one array and it keeps the previous call's contents (the builder clears the words that
 * must start at zero itself): two calls that may overlap on the GPU -- different streams, or two host threads -- RACE;
 * order them on one stream or with events. */
#define GPRT_NOALLOC_MAX_PRIMS     65536u
#define GPRT_NOALLOC_SCRATCH_BYTES (5u << 20)
int gprt_lbvh_build_noalloc(int device_ordinal, const gprt_aabb* d_aabbs, uint32_t n_aabbs, const uint8_t* d_prim_tags,
                            void* d_nodes, uint32_t* d_sorted_prims, int32_t* root_ref, void* cuda_stream);
/* Reads back what gprt_build_accel / gprt_refit_accel built in `ctx`, for comparing the two builders (tests; blocking):
 * the binary nodes (h_nodes: (n - 1) x 64 bytes), the sorted primitive indices (n x uint32), the tag bytes the context
 * passed to its builder (n bytes) and the root reference; any of the pointers may be NULL.  `n` must be the context's AABB
 * count.  Also no counterpart in the reference. */
int gprt_get_accel(gprt_ctx* ctx, uint32_t n, void* h_nodes, uint32_t* h_sorted_prims, uint8_t* h_prim_tags, int32_t* root_ref);

const char* gprt_error_string(int status);
uint32_t gprt_abi_version(void);

#ifdef __cplusplus
}
#endif

/* Layout checks: sizes and offsets the reference's shaders rely on. */
#if defined(__cplusplus)
static_assert(sizeof(gprt_scene_cb) == 144, "SceneConstantBuffer");
static_assert(offsetof(gprt_scene_cb, camera_position) == 64, "camera_position");
static_assert(offsetof(gprt_scene_cb, light_position) == 80, "light_position");
static_assert(offsetof(gprt_scene_cb, light_ambient_color) == 96, "light_ambient_color");
static_assert(offsetof(gprt_scene_cb, light_diffuse_color) == 112, "light_diffuse_color");
static_assert(offsetof(gprt_scene_cb, elapsed_time) == 132, "elapsed_time");
static_assert(sizeof(gprt_material) == 48, "PrimitiveConstantBuffer");
static_assert(offsetof(gprt_material, step_scale) == 32, "step_scale");
static_assert(sizeof(gprt_prim_instance_cb) == 8, "PrimitiveInstanceConstantBuffer");
static_assert(sizeof(gprt_prim_per_frame) == 128, "PrimitiveInstancePerFrameBuffer");
static_assert(sizeof(gprt_vertex) == 24, "Vertex");
static_assert(sizeof(gprt_aabb) == 24, "D3D12_RAYTRACING_AABB");
static_assert(sizeof(gprt_instance_desc) == 64, "D3D12_RAYTRACING_INSTANCE_DESC");
static_assert(sizeof(gprt_hit_group_record) == 96, "hit-group record stride");
static_assert(offsetof(gprt_hit_group_record, material_cb) == 32, "root arguments");
static_assert(offsetof(gprt_hit_group_record, aabb_cb) == 80, "aabb_cb");
#endif

#endif /* GPRT_H */
