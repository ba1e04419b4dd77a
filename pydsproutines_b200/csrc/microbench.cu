// microbench.cu -- sustained FP32 FMA rate of the device (register-resident FFMA chains), the
// denominator of the FFT-flop roofline the bench reports next to the HBM one.  Not a reference
// function; MEASURED_PEAKS.json only carries HBM bandwidth and BF16 tensor throughput.
#include "internal.cuh"

namespace xc {

__global__ void __launch_bounds__(256) ffma_kernel(float* out, int iters, float a, float b) {
    float x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = threadIdx.x * 1e-3f + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int i = 0; i < 16; ++i) x[i] = fmaf(x[i], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += x[i];
    if (s == 123.456f) out[0] = s;   // never true; keeps the chains alive
}

}  // namespace xc

using namespace xc;

extern "C" int32_t xc_measure_fp32_tflops(double ms, double* tflops) {
    XC_API_GUARD;
    if (!tflops) return fail(XC_ERR_ARG, "tflops is NULL");
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail(XC_ERR_CUDA, "no CUDA device available");
    }
    float* d_out = reinterpret_cast<float*>(workspace(SLOT_TMP3, 256));
    if (!d_out) return XC_ERR_NOMEM;
    const int blocks = sm_count() * 8, threads = 256, iters = 4096;
    const double flop_per_launch = 2.0 * 16 * 8 * (double)iters * threads * blocks;
    cudaEvent_t e0, e1;
    XC_CUDA_TRY(cudaEventCreate(&e0));
    XC_CUDA_TRY(cudaEventCreate(&e1));
    ffma_kernel<<<blocks, threads>>>(d_out, iters, 0.999f, 0.001f);   // warm-up
    XC_LAUNCH_CHECK();
    XC_CUDA_TRY(cudaDeviceSynchronize());
    double best = 0.0, spent = 0.0;
    while (spent < ms) {
        XC_CUDA_TRY(cudaEventRecord(e0));
        ffma_kernel<<<blocks, threads>>>(d_out, iters, 0.999f, 0.001f);
        XC_LAUNCH_CHECK();
        XC_CUDA_TRY(cudaEventRecord(e1));
        XC_CUDA_TRY(cudaEventSynchronize(e1));
        float t = 0.f;
        XC_CUDA_TRY(cudaEventElapsedTime(&t, e0, e1));
        spent += t;
        const double tf = flop_per_launch / (t * 1e-3) / 1e12;
        // report the LAST (sustained) value rather than the burst: keep a running value of the tail
        best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *tflops = best;
    return XC_OK;
}
