// Development aid: lone-warp latency of the bare sphere-tracing chain (p = o + t d; dist = f(p); t += s dist) per SDF type, IEEE flags.
// nvcc -gencode arch=compute_100a,code=sm_100a --fmad=false -prec-div=true -prec-sqrt=true -ftz=false -o /tmp/chain tools/chain_latency.cu && /tmp/chain
// B200, round 2: 244 / 186 / 268 / 328 / 829 / 290 / 519 cycles per step for MiniSpheres / RoundCube / SquareTorus / TwistedTorus / Cog / Cylinder / FractalPyramid.
#include <cuda_runtime.h>
#include <stdint.h>
#include <cstdio>
#include "../gpuraytracer_b200/csrc/prims.cuh"
using namespace gprt;
template <int SDF> __global__ void k(float* out, long long* cyc, float3 o, float3 d, uint32_t iters, int steps) {
    float t = 0.0f;
    const long long c0 = clock64();
#pragma unroll 1
    for (int i = 0; i < steps; i++) {
        const V3 p = v3(o.x + t * d.x, o.y + t * d.y, o.z + t * d.z);
        const float dist = sdf_distance<SDF>(p, iters);
        if (dist <= 0.0001f * t && t >= 0.0f && t <= 10000.0f) break;
        t += 0.001f * dist;
    }
    const long long c1 = clock64();
    out[threadIdx.x] = t;
    if (threadIdx.x == 0) *cyc = c1 - c0;
}
int main() {
    float* out; long long* cyc;
    cudaMalloc(&out, 4096); cudaMallocManaged(&cyc, 8);
    const float3 o = make_float3(0.3f, 0.2f, -5.0f), d = make_float3(0.01f, 0.02f, 1.0f);
#define RUN(S) for (int w = 0; w < 2; w++) { k<S><<<1, 32>>>(out, cyc, o, d, 4, 512); cudaDeviceSynchronize(); } printf("sdf %d: %.1f cycles per step (one warp, 512 steps)\n", S, *cyc / 512.0);
    RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5) RUN(6)
    return 0;
}
