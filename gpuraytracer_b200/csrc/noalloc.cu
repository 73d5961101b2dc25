// noalloc.cu -- gprt_lbvh_build_noalloc (include/gprt.h): the LBVH builder of lbvh.cu run with its scratch in a fixed array
// of the library's device image, for callers that may not allocate.  No counterpart in the reference.
// Nothing here calls cudaMalloc / cudaFree or an allocating library routine.
#include <cuda_runtime.h>

#include <climits>
#include <mutex>

#include "../../include/gprt.h"
#include "lbvh.h"

namespace {

// The builder's scratch: one copy per GPU that loads the library, held until the process exits.  Every call on a device
// shares it and it keeps the previous call's contents; lbvh_build clears what must start at zero (the scene bounds and the
// refit flags) and overwrites the rest before reading it.
static __device__ __align__(256) unsigned char g_lbvh_scratch[GPRT_NOALLOC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_mutex;
void* g_scratch_ptr[kMaxDevices] = {}; // device address of g_lbvh_scratch on each device, looked up once

// In host code `g_lbvh_scratch` names a host-side stand-in, not device memory: the kernels need the address
// cudaGetSymbolAddress returns, per device, with that device current.
int scratch_for(int device, void** out) {
    std::lock_guard<std::mutex> lock(g_mutex);
    if (!g_scratch_ptr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_lbvh_scratch);
        if (e != cudaSuccess || !p) { cudaGetLastError(); return GPRT_E_CUDA; }
        g_scratch_ptr[device] = p;
    }
    *out = g_scratch_ptr[device];
    return GPRT_OK;
}

} // namespace

extern "C" int gprt_lbvh_build_noalloc(int device_ordinal, const gprt_aabb* d_aabbs, uint32_t n_aabbs, const uint8_t* d_prim_tags, void* d_nodes,
                                       uint32_t* d_sorted_prims, int32_t* root_ref, void* cuda_stream) {
    if (!root_ref || device_ordinal < 0 || device_ordinal >= kMaxDevices) return GPRT_E_INVALID_ARG;
    *root_ref = INT_MIN;
    if (n_aabbs > GPRT_NOALLOC_MAX_PRIMS || gprt::lbvh_scratch_bytes(n_aabbs ? n_aabbs : 1) > (size_t)GPRT_NOALLOC_SCRATCH_BYTES) return GPRT_E_TOO_LARGE;
    if (n_aabbs == 0) return GPRT_OK;
    if (!d_aabbs || !d_sorted_prims || (n_aabbs > 1 && !d_nodes)) return GPRT_E_INVALID_ARG;
    if (cudaSetDevice(device_ordinal) != cudaSuccess) { cudaGetLastError(); return GPRT_E_NO_DEVICE; }
    void* scratch = nullptr;
    const int rc = scratch_for(device_ordinal, &scratch);
    if (rc != GPRT_OK) return rc;
    gprt::lbvh_build(reinterpret_cast<const float*>(d_aabbs), n_aabbs, scratch, static_cast<float4*>(d_nodes), d_sorted_prims, root_ref, cuda_stream, d_prim_tags);
    if (cudaGetLastError() != cudaSuccess) return GPRT_E_CUDA;
    return GPRT_OK;
}
