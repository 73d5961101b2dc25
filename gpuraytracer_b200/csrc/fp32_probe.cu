// fp32_probe.cu -- measured denominator of the FP32 roofline: dependent-chain-free FFMA issue
// rate over all SMs (MEASURED_PEAKS.json holds HBM and bf16 tensor peaks only).
#include <cuda_runtime.h>

#include "lbvh.h"

namespace gprt {
namespace {

constexpr int kIters = 4096;
constexpr int kChains = 16;

__global__ void __launch_bounds__(256) k_fma_probe(float* sink, float a, float b) {
    float x[kChains];
#pragma unroll
    for (int i = 0; i < kChains; i++) x[i] = (float)(threadIdx.x + i);
    for (int it = 0; it < kIters; it++) {
#pragma unroll
        for (int i = 0; i < kChains; i++) x[i] = fmaf(x[i], a, b);
    }
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < kChains; i++) s += x[i];
    if (s == 12345.678f) sink[0] = s; // never true; keeps the chain alive
}

} // namespace

int fp32_peak_probe(void* stream_, double* tflops) {
    cudaStream_t s = (cudaStream_t)stream_;
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    float* sink = nullptr;
    cudaError_t e = cudaMalloc((void**)&sink, 256);
    if (e != cudaSuccess) return (int)e;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int blocks = sms * 8;
    double best = 0.0;
    for (int rep = 0; rep < 6; rep++) {
        cudaEventRecord(e0, s);
        k_fma_probe<<<blocks, 256, 0, s>>>(sink, 0.999f, 0.001f);
        cudaEventRecord(e1, s);
        cudaEventSynchronize(e1);
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flops = 2.0 * (double)blocks * 256.0 * kIters * kChains;
        if (ms > 0.0f && rep > 0) { const double tf = flops / (ms * 1e-3) / 1e12; if (tf > best) best = tf; }
    }
    e = cudaGetLastError();
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    *tflops = best;
    return (int)e;
}

} // namespace gprt
