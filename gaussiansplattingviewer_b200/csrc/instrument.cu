// Measurement helpers (not on the render path): micro-benchmarks for the roofline denominators the
// blend kernel is judged against.  SURVEY.md 8(d): K5 is bound by FP32 issue and MUFU (ex2) throughput,
// whose peaks are "to be measured by a micro-benchmark on the box"; MEASURED_PEAKS.json only carries the
// HBM copy and bf16 GEMM figures.
#include "gsr_internal.h"

namespace {

// 8 independent FMA chains per thread, 4096 iterations: 2 * 8 * 4096 flop per thread.
__global__ void __launch_bounds__(256) k_peak_ffma(float* out, float a, float b, int iters) {
    float x0 = threadIdx.x, x1 = x0 + 1.f, x2 = x0 + 2.f, x3 = x0 + 3.f, x4 = x0 + 4.f, x5 = x0 + 5.f, x6 = x0 + 6.f,
          x7 = x0 + 7.f;
    for (int i = 0; i < iters; ++i) {
        x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
        x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
    }
    const float s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (s == 123.456f) out[0] = s;  // never true; keeps the chains alive
}

// 8 independent ex2.approx chains per thread.
__global__ void __launch_bounds__(256) k_peak_ex2(float* out, float a, int iters) {
    float x0 = threadIdx.x * 1e-3f, x1 = x0 + .1f, x2 = x0 + .2f, x3 = x0 + .3f, x4 = x0 + .4f, x5 = x0 + .5f,
          x6 = x0 + .6f, x7 = x0 + .7f;
    for (int i = 0; i < iters; ++i) {
        x0 = exp2f(x0 * a); x1 = exp2f(x1 * a); x2 = exp2f(x2 * a); x3 = exp2f(x3 * a);
        x4 = exp2f(x4 * a); x5 = exp2f(x5 * a); x6 = exp2f(x6 * a); x7 = exp2f(x7 * a);
    }
    const float s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (s == 123.456f) out[0] = s;
}

__global__ void __launch_bounds__(256) k_peak_copy(const float4* __restrict__ src, float4* __restrict__ dst, size_t n) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = src[i];
}

}  // namespace

// Returns 0 on success.  fp32_tflops: FFMA throughput (2 flop per FMA); ex2_tops: 1e12 ex2/s (each
// iteration also issues one FMUL, as the blend kernel's exp does); copy_gbs: read+write bytes of a
// device copy of 2 x 1 GiB.
int measure_peaks_impl(float* fp32_tflops, float* ex2_tops, float* copy_gbs) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    float* d_out = nullptr;
    if (cudaMalloc(&d_out, 256) != cudaSuccess) return -1;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int iters = 4096, grid = sms * 16, reps = 10;
    float best;
    // FFMA
    best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        cudaEventRecord(e0);
        k_peak_ffma<<<grid, 256>>>(d_out, 1.0000001f, 1e-7f, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    *fp32_tflops = (float)((double)grid * 256 * 8 * 2 * iters / (best * 1e-3) / 1e12);
    // EX2
    best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        cudaEventRecord(e0);
        k_peak_ex2<<<grid, 256>>>(d_out, -0.5f, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    *ex2_tops = (float)((double)grid * 256 * 8 * iters / (best * 1e-3) / 1e12);
    // copy
    *copy_gbs = 0.f;
    const size_t bytes = (size_t)1 << 30;
    float4 *a = nullptr, *b = nullptr;
    if (cudaMalloc(&a, bytes) == cudaSuccess && cudaMalloc(&b, bytes) == cudaSuccess) {
        cudaMemset(a, 1, bytes);
        best = 1e30f;
        for (int r = 0; r < reps; ++r) {
            cudaEventRecord(e0);
            k_peak_copy<<<sms * 16, 256>>>(a, b, bytes / 16);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms = 0;
            cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
        }
        *copy_gbs = (float)(2.0 * bytes / (best * 1e-3) / 1e9);
    }
    if (a) cudaFree(a);
    if (b) cudaFree(b);
    cudaFree(d_out);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}
