// lbvh.h -- interface of the LBVH builder (lbvh.cu).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace gprt {

// Bytes of device scratch lbvh_build / lbvh_refit need for n primitives (kept by the context so
// gprt_refit_accel can reuse the topology).
size_t lbvh_scratch_bytes(uint32_t n);

// Builds the tree over d_aabbs (6 floats per primitive).  d_nodes: 4 float4 per internal node
// (n - 1 nodes).  d_sorted_prims: n uint32 (primitive index per Morton-sorted leaf).
// root_ref: 0 (internal root), ~0 (single leaf) or INT32_MIN (empty).  Returns kernels launched.
int lbvh_build(const float* d_aabbs, uint32_t n, void* d_scratch, float4* d_nodes, uint32_t* d_sorted_prims, int32_t* root_ref, void* stream,
               const uint8_t* d_prim_tags = nullptr); // optional 4-bit tag per primitive, stored in bits 27..30 of its leaf ref

// Recomputes the node boxes bottom-up for moved AABBs, topology unchanged.
int lbvh_refit(const float* d_aabbs, uint32_t n, void* d_scratch, float4* d_nodes, const uint32_t* d_sorted_prims, void* stream,
               const uint8_t* d_prim_tags = nullptr);

// The 4-wide form the traversal walks (lbvh.cu, "4-wide collapse"): d_wide holds 8 float4 per binary internal node
// (n - 1 nodes), d_top the first `top_capacity` nodes breadth first with refs into itself (staged in shared memory by the
// walk kernels), *d_n_top how many of them exist.  Call after lbvh_build / lbvh_refit.  Returns kernels launched.
constexpr int32_t kWideInvalid = 0x7FFFFFFF;   // empty child slot
constexpr uint32_t kWideTopFlag = 0x40000000u; // ref names a node of the staged top array
int lbvh_collapse(const float4* d_nodes, uint32_t n, float4* d_wide, float4* d_top, uint32_t top_capacity, uint32_t* d_n_top, void* stream);

// FP32 FMA issue-rate probe (fp32_probe.cu): best of a few launches, TFLOP/s.  Returns a cudaError_t.
int fp32_peak_probe(void* stream, double* tflops);

// transforms.cu: writes n PrimitiveInstancePerFrameBuffer entries from the static animation inputs; returns kernels launched.
int update_transforms(const void* d_anim, const float* d_aabbs, uint32_t n, float animation_time, void* d_per_frame, void* stream);

} // namespace gprt
