// gprt_common.cpp -- the parts of the C ABI that need no device.
#include "../../include/gprt.h"

extern "C" {

void gprt_default_config(gprt_config* cfg) {
    if (!cfg) return;
    cfg->abi_version = GPRT_ABI_VERSION;
    cfg->max_recursion_depth = 3; // MAX_RAY_RECURSION_DEPTH, src/ConstantBuffers.h:31
    cfg->metaballs_count = 3;     // METABALLS_COUNT, src/ConstantBuffers.h:12
    cfg->metaball_max_steps = 128; // src/VolumetricPrimitives.hlsli:160
    cfg->sdf_max_steps = 512;     // src/SignedDistancePrimitives.hlsli:291
    cfg->math_mode = GPRT_MATH_PARITY;
    cfg->cand_per_ray = 0;
    cfg->flags = 0;
    cfg->fractal_iterations = 4;  // FRACTAL_ITERATIONS_COUNT, src/ConstantBuffers.h:27
}

uint32_t gprt_abi_version(void) { return GPRT_ABI_VERSION; }

const char* gprt_error_string(int status) {
    switch (status) {
    case GPRT_OK: return "ok";
    case GPRT_E_INVALID_ARG: return "invalid argument";
    case GPRT_E_CUDA: return "CUDA runtime error (see gprt_last_cuda_error)";
    case GPRT_E_NO_DEVICE: return "no usable CUDA device (this library has no CPU fallback)";
    case GPRT_E_OUT_OF_MEMORY: return "out of memory";
    case GPRT_E_NOT_BUILT: return "acceleration structure not built";
    case GPRT_E_QUEUE_OVERFLOW: return "candidate queue overflow (raise gprt_config.cand_per_ray or render in tiles)";
    case GPRT_E_UNSUPPORTED: return "unsupported configuration";
    case GPRT_E_ABI_MISMATCH: return "ABI version mismatch";
    case GPRT_E_TOO_LARGE: return "size beyond the fixed scratch of this no-allocation entry point (GPRT_NOALLOC_MAX_PRIMS)";
    default: return "unknown gprt status";
    }
}

} // extern "C"
