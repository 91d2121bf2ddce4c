// gc_peak.cuh -- FP32 (non-tensor) peak microbenchmark: the denominator of the
// FP32 roofline fraction bench.py reports (SURVEY.md 6.3 asks for a measured
// figure instead of 2 * 128 * n_SM * f_clk).
#pragma once

#include <cuda_runtime.h>

namespace gc {

constexpr int PEAK_ITERS = 4096;
constexpr int PEAK_CHAINS = 8;

__global__ void __launch_bounds__(256) k_fp32_peak(float *out, float a, float b)
{
    float acc[PEAK_CHAINS];
#pragma unroll
    for (int k = 0; k < PEAK_CHAINS; k++) acc[k] = (float) (threadIdx.x + k);
    for (int it = 0; it < PEAK_ITERS; it++) {
#pragma unroll
        for (int k = 0; k < PEAK_CHAINS; k++) acc[k] = fmaf(acc[k], a, b);
    }
    float s = 0.0f;
#pragma unroll
    for (int k = 0; k < PEAK_CHAINS; k++) s += acc[k];
    if (s == 123.456f) out[0] = s;   // never true; keeps the chains alive
}

// The same with Blackwell's packed fma.rn.f32x2 (two FMAs per issued instruction).
__global__ void __launch_bounds__(256) k_fp32x2_peak(float *out, float a, float b)
{
    float2 acc[PEAK_CHAINS / 2];
#pragma unroll
    for (int k = 0; k < PEAK_CHAINS / 2; k++) acc[k] = make_float2((float) (threadIdx.x + k), (float) k);
    const float2 a2 = make_float2(a, a), b2 = make_float2(b, b);
    for (int it = 0; it < PEAK_ITERS; it++) {
#pragma unroll
        for (int k = 0; k < PEAK_CHAINS / 2; k++) acc[k] = __ffma2_rn(acc[k], a2, b2);
    }
    float s = 0.0f;
#pragma unroll
    for (int k = 0; k < PEAK_CHAINS / 2; k++) s += acc[k].x + acc[k].y;
    if (s == 123.456f) out[0] = s;
}

// Returns TFLOP/s (2 flop per FMA), best of `reps`; packed = 1 times the f32x2 variant.
inline cudaError_t measure_fp32_peak(int device, double *tflops, int packed = 0)
{
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return e;
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    float *d = nullptr;
    if ((e = cudaMalloc(&d, 4)) != cudaSuccess) return e;
    cudaEvent_t t0, t1;
    cudaEventCreate(&t0);
    cudaEventCreate(&t1);
    const int blocks = sms * 8 * 4, threads = 256;
    double best = 0.0;
    for (int rep = 0; rep < 6; rep++) {
        cudaEventRecord(t0);
        if (packed) k_fp32x2_peak<<<blocks, threads>>>(d, 0.999f, 0.001f);
        else k_fp32_peak<<<blocks, threads>>>(d, 0.999f, 0.001f);
        cudaEventRecord(t1);
        if ((e = cudaEventSynchronize(t1)) != cudaSuccess) break;
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, t0, t1);
        double flop = 2.0 * PEAK_ITERS * PEAK_CHAINS * (double) blocks * threads;
        if (rep > 0 && ms > 0.0f) best = flop / (ms * 1.0e-3) / 1.0e12 > best ? flop / (ms * 1.0e-3) / 1.0e12 : best;
    }
    cudaEventDestroy(t0);
    cudaEventDestroy(t1);
    cudaFree(d);
    *tflops = best;
    return e;
}

} // namespace gc
