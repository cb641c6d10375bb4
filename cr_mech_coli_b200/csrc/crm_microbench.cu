// crm_microbench.cu — measures the two non-tensor peaks the force kernel is bounded by
// (neither is in MEASURED_PEAKS.json): dense FP32 FMA throughput and MUFU.EX2
// throughput.  Used by bench.py for the roofline denominator.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/crm_b200.h"

namespace {

__global__ void __launch_bounds__(256) k_fma_peak(float* out, int iters, float a, float b) {
    float x0 = threadIdx.x, x1 = x0 + 1.f, x2 = x0 + 2.f, x3 = x0 + 3.f, x4 = x0 + 4.f, x5 = x0 + 5.f, x6 = x0 + 6.f,
          x7 = x0 + 7.f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
            x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
        }
    }
    float s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
    if (s == 123.456f) out[0] = s;
}

__global__ void __launch_bounds__(256) k_mufu_peak(float* out, int iters) {
    float x0 = threadIdx.x * 1e-3f, x1 = x0 + .1f, x2 = x0 + .2f, x3 = x0 + .3f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x0));
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x1));
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x2));
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x3));
        }
    }
    float s = x0 + x1 + x2 + x3;
    if (s == 123.456f) out[0] = s;
}

}  // namespace

extern "C" int crm_microbench(int device, double* fp32_tflops, double* mufu_tops, double* sm_mhz_est) {
    if (cudaSetDevice(device) != cudaSuccess) return CRM_ERR_CUDA;
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device);
    float* out = nullptr;
    cudaMalloc(&out, 16);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int blocks = prop.multiProcessorCount * 8, threads = 256;
    double best_fma = 0, best_mufu = 0;
    for (int rep = 0; rep < 5; ++rep) {
        const int iters = 2000;
        cudaEventRecord(e0);
        k_fma_peak<<<blocks, threads>>>(out, iters, 0.999f, 0.001f);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        double flops = 2.0 * 8 * 16 * (double)iters * blocks * threads;
        if (rep > 0 && flops / (ms * 1e-3) > best_fma) best_fma = flops / (ms * 1e-3);
        cudaEventRecord(e0);
        k_mufu_peak<<<blocks, threads>>>(out, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        double ops = 4.0 * 16 * (double)iters * blocks * threads;
        if (rep > 0 && ops / (ms * 1e-3) > best_mufu) best_mufu = ops / (ms * 1e-3);
    }
    cudaFree(out);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (cudaGetLastError() != cudaSuccess) return CRM_ERR_CUDA;
    if (fp32_tflops) *fp32_tflops = best_fma * 1e-12;
    if (mufu_tops) *mufu_tops = best_mufu * 1e-12;
    // implied clock if every SM retires 128 FMA lanes per cycle
    if (sm_mhz_est) *sm_mhz_est = best_fma / (2.0 * 128 * prop.multiProcessorCount) * 1e-6;
    return 0;
}
