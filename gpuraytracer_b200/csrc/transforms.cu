// transforms.cu -- per-frame primitive transforms computed on the GPU.
//
// Renderer::update_aabb_primitive_attributes (src/Renderer.cpp:302-356) rebuilds, every frame and
// for every primitive, transform = scale * rotationY(-2 t) * translate(AABB centre) and its inverse,
// and do_raytracing uploads the 128 bytes per primitive (:438-443).  At a million primitives that
// is ~85 ms of host arithmetic and 128 MB over PCIe per frame -- more than the frame itself.  Here
// the static inputs (scale, "does it rotate", the AABB) stay on the device and one kernel writes
// the PrimitiveInstancePerFrameBuffer array in place.  It calls the SAME functions as the host path
// (host/mathx.h, double arithmetic, --fmad=false), so the matrices are bit-identical to the host's.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/gprt.h"
#include "../host/mathx.h"
#include "lbvh.h"

namespace gprt {
namespace {

__global__ void k_update_transforms(const gprt_prim_anim* __restrict__ anim, const float* __restrict__ aabbs, uint32_t n, gprt_host::Mat4 rotation,
                                    gprt_prim_per_frame* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    using namespace gprt_host;
    const gprt_prim_anim a = anim[i];
    const float* b = aabbs + 6 * (size_t)i;
    // v_translation = 0.5f * (min + max), float arithmetic as in the reference
    const float tx = 0.5f * (b[0] + b[3]), ty = 0.5f * (b[1] + b[4]), tz = 0.5f * (b[2] + b[5]);
    const Mat4 transform = mat_mul(mat_mul(mat_scaling(a.scale[0], a.scale[1], a.scale[2]), a.rotates ? rotation : mat_identity()), mat_translation(tx, ty, tz));
    mat_store(transform, out[i].local_space_to_bottom_level_as);
    mat_store(mat_inverse(transform), out[i].bottom_level_as_to_local_space);
}

} // namespace

int update_transforms(const void* d_anim, const float* d_aabbs, uint32_t n, float animation_time, void* d_per_frame, void* stream) {
    if (n == 0) return 0;
    const gprt_host::Mat4 rotation = gprt_host::mat_rotation_y(-2.0 * (double)animation_time); // on the host: same libm as the host path
    k_update_transforms<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>((const gprt_prim_anim*)d_anim, d_aabbs, n, rotation, (gprt_prim_per_frame*)d_per_frame);
    return 1;
}

} // namespace gprt
