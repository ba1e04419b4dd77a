// energy.cu -- sliding-window energy (the normalisation term of every correlator), the moving
// average/sum and |x|^2.  Replaces complex_magn.cu:8-20 and filter.cu:291-347 of the reference.
//
// One kernel serves all: the logical input is p[j] = 0 for j < front_pad, f(x[j - front_pad]) after,
// with f = |.|^2 for complex64 input or identity for float32; out[s] = scale * sum_{n<L} p[s+n].
// A CTA produces a tile of 2048 outputs: the first window is reduced directly in double, the rest by
// a block-wide scan (double) of d[j] = p[s0+j+L] - p[s0+j], so no long-range cancellation ever
// happens in float32 (the reference accumulates in double too: filter.cu:324-336).
#include "common.cuh"

namespace xc {

constexpr int SE_THREADS = 256;
constexpr int SE_PER_THREAD = 8;
constexpr int SE_TILE = SE_THREADS * SE_PER_THREAD;

template <bool COMPLEX_IN>
__device__ __forceinline__ float load_p(const void* x, int64_t j, int64_t front_pad, int64_t n_in) {
    const int64_t k = j - front_pad;
    if (k < 0 || k >= n_in) return 0.f;
    if (COMPLEX_IN) {
        const float2 v = __ldg(reinterpret_cast<const float2*>(x) + k);
        return fmaf(v.x, v.x, v.y * v.y);
    }
    return __ldg(reinterpret_cast<const float*>(x) + k);
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <bool COMPLEX_IN>
__global__ void __launch_bounds__(SE_THREADS) sliding_sum_kernel(const void* __restrict__ x, int64_t n_in,
                                                                 int64_t front_pad, int L, int64_t n_out,
                                                                 float* __restrict__ out, int average) {
    __shared__ double s_part[SE_THREADS / 32];
    __shared__ double s_w0;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t s0 = (int64_t)blockIdx.x * SE_TILE;

    // (1) first window of the tile
    double acc = 0.0;
    for (int n = tid; n < L; n += SE_THREADS) acc += (double)load_p<COMPLEX_IN>(x, s0 + n, front_pad, n_in);
    acc = warp_sum(acc);
    if (lane == 0) s_part[warp] = acc;
    __syncthreads();
    if (tid == 0) {
        double w = 0.0;
        for (int i = 0; i < SE_THREADS / 32; ++i) w += s_part[i];
        s_w0 = w;
    }
    __syncthreads();
    const double w0 = s_w0;
    __syncthreads();

    // (2) differences, thread-local inclusive prefix
    double d[SE_PER_THREAD];
    double run = 0.0;
    const int64_t jb = s0 + (int64_t)tid * SE_PER_THREAD;
#pragma unroll
    for (int r = 0; r < SE_PER_THREAD; ++r) {
        const int64_t j = jb + r;
        const float add = load_p<COMPLEX_IN>(x, j + L, front_pad, n_in);
        const float sub = load_p<COMPLEX_IN>(x, j, front_pad, n_in);
        run += (double)add - (double)sub;
        d[r] = run;
    }
    // (3) exclusive scan of the thread totals across the block
    double incl = run;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const double y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
    }
    if (lane == 31) s_part[warp] = incl;
    __syncthreads();
    double base = 0.0;
    for (int i = 0; i < warp; ++i) base += s_part[i];
    const double excl = base + incl - run;
    // out[s0 + k] = w0 + sum_{j<k} d_raw[j]
    const double inv = average ? 1.0 / (double)L : 1.0;
#pragma unroll
    for (int r = 0; r < SE_PER_THREAD; ++r) {
        const int64_t s = jb + r;
        if (s < n_out) {
            const double val = w0 + excl + (r == 0 ? 0.0 : d[r - 1]);
            out[s] = (float)(val * inv);
        }
    }
}

int32_t sliding_sum_device(const void* x, bool complex_in, int64_t n_in, int64_t front_pad, int L, int64_t n_out,
                           float* out, int average, cudaStream_t stream) {
    if (n_out <= 0) return XC_OK;
    const int64_t blocks = (n_out + SE_TILE - 1) / SE_TILE;
    if (complex_in)
        sliding_sum_kernel<true><<<(unsigned)blocks, SE_THREADS, 0, stream>>>(x, n_in, front_pad, L, n_out, out, average);
    else
        sliding_sum_kernel<false><<<(unsigned)blocks, SE_THREADS, 0, stream>>>(x, n_in, front_pad, L, n_out, out, average);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

// energy of arbitrary windows: one warp per shift-list entry (general step / explicit shift lists)
__global__ void window_energy_kernel(const float2* __restrict__ x, int64_t xlen, int L, int64_t shift0, int64_t step,
                                     const int64_t* __restrict__ shifts, int64_t K, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t i = warp; i < K; i += nwarps) {
        const int64_t s = shifts ? shifts[i] : shift0 + i * step;
        double acc = 0.0;
        if (s >= 0 && s + L <= xlen) {
            float part = 0.f;
            int cnt = 0;
            for (int n = lane; n < L; n += 32) {
                const float2 v = __ldg(x + s + n);
                part += fmaf(v.x, v.x, v.y * v.y);
                if (++cnt == 64) { acc += (double)part; part = 0.f; cnt = 0; }
            }
            acc += (double)part;
        }
        acc = warp_sum(acc);
        if (lane == 0) out[i] = (float)acc;
    }
}

int32_t window_energy_device(const float2* x, int64_t xlen, int L, int64_t shift0, int64_t step, const int64_t* shifts,
                             int64_t K, float* out, cudaStream_t stream) {
    if (K <= 0) return XC_OK;
    int64_t blocks = (K * 32 + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    window_energy_kernel<<<(unsigned)blocks, 256, 0, stream>>>(x, xlen, L, shift0, step, shifts, K, out);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

// sum |x|^2 of a short vector (cutout / templates) into a device float, double accumulation
__global__ void normsq_kernel(const float2* __restrict__ x, int64_t n, int64_t row_stride, float* __restrict__ out) {
    __shared__ double s_part[8];
    const float2* row = x + (int64_t)blockIdx.x * row_stride;
    double acc = 0.0;
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
        const float2 v = row[i];
        acc += (double)v.x * v.x + (double)v.y * v.y;
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double w = 0.0;
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) w += s_part[i];
        out[blockIdx.x] = (float)w;
    }
}

int32_t normsq_device(const float2* x, int64_t n, int64_t rows, int64_t row_stride, float* out, cudaStream_t stream) {
    normsq_kernel<<<(unsigned)rows, 256, 0, stream>>>(x, n, row_stride, out);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

template <typename OUT>
__global__ void magnsq_kernel(const float2* __restrict__ x, int64_t n, OUT* __restrict__ out) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float2 v = __ldg(x + i);
        const OUT re = (OUT)v.x, im = (OUT)v.y;
        out[i] = re * re + im * im;
    }
}

int32_t magnsq_device(const float2* x, int64_t n, void* out, int out_f64, cudaStream_t stream) {
    if (n <= 0) return XC_OK;
    int64_t blocks = (n + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (out_f64)
        magnsq_kernel<double><<<(unsigned)blocks, 256, 0, stream>>>(x, n, reinterpret_cast<double*>(out));
    else
        magnsq_kernel<float><<<(unsigned)blocks, 256, 0, stream>>>(x, n, reinterpret_cast<float*>(out));
    XC_LAUNCH_CHECK();
    return XC_OK;
}

}  // namespace xc
