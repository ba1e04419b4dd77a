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

// Very long windows (the reference's benchmark correlates a 1e6-sample cutout at 128 shifts): the first window of a tile is
// not summed sample by sample by one CTA, but taken from cumulative sums at chunk boundaries (cum[j] = sum of p[0 .. j SE_CHUNK),
// double) plus the two ragged edges.
constexpr int SE_CHUNK = 4096;
constexpr int SE_LONG_L = 4 * SE_CHUNK;    // windows from this length on use the chunk sums

template <bool COMPLEX_IN>
__global__ void __launch_bounds__(256) chunk_sum_kernel(const void* __restrict__ x, int64_t n_in, double* __restrict__ part) {
    __shared__ double s_part[8];
    const int64_t j0 = (int64_t)blockIdx.x * SE_CHUNK;
    double acc = 0.0;
    for (int i = threadIdx.x; i < SE_CHUNK; i += 256) acc += (double)load_p<COMPLEX_IN>(x, j0 + i, 0, n_in);
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double w = 0.0;
        for (int i = 0; i < 8; ++i) w += s_part[i];
        part[blockIdx.x] = w;
    }
}
// cum[0] = 0, cum[j + 1] = cum[j] + part[j], in place over part shifted by one (one block; n <= a few 10^4 chunks)
__global__ void __launch_bounds__(1024) chunk_scan_kernel(const double* __restrict__ part, int64_t n, double* __restrict__ cum) {
    __shared__ double s_tot[32];
    __shared__ double s_carry;
    if (threadIdx.x == 0) {
        s_carry = 0.0;
        cum[0] = 0.0;
    }
    __syncthreads();
    for (int64_t base = 0; base < n; base += 1024) {
        const int64_t i = base + threadIdx.x;
        double v = i < n ? part[i] : 0.0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const double y = __shfl_up_sync(0xffffffffu, v, o);
            if ((threadIdx.x & 31) >= o) v += y;
        }
        if ((threadIdx.x & 31) == 31) s_tot[threadIdx.x >> 5] = v;
        __syncthreads();
        double off = s_carry;
        for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) off += s_tot[w];
        if (i < n) cum[i + 1] = off + v;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = off + v;
        __syncthreads();
    }
}

template <bool COMPLEX_IN>
__global__ void __launch_bounds__(SE_THREADS) sliding_sum_kernel(const void* __restrict__ x, int64_t n_in,
                                                                 int64_t front_pad, int L, int64_t n_out,
                                                                 float* __restrict__ out, int average,
                                                                 const double* __restrict__ cum) {
    __shared__ double s_part[SE_THREADS / 32];
    __shared__ double s_w0;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t s0 = (int64_t)blockIdx.x * SE_TILE;

    // (1) first window of the tile
    double acc = 0.0;
    if (cum) {
        // whole chunks [c0, c1) from the cumulative sums, the ragged edges [s0, c0 CHUNK) and [c1 CHUNK, s0 + L) directly
        const int64_t c0 = (s0 + SE_CHUNK - 1) / SE_CHUNK, c1 = (s0 + L) / SE_CHUNK;
        const int64_t e0 = c0 * SE_CHUNK, e1 = c1 * SE_CHUNK;
        for (int64_t n = s0 + tid; n < e0; n += SE_THREADS) acc += (double)load_p<COMPLEX_IN>(x, n, 0, n_in);
        for (int64_t n = e1 + tid; n < s0 + L; n += SE_THREADS) acc += (double)load_p<COMPLEX_IN>(x, n, 0, n_in);
        if (tid == 0) acc += cum[c1] - cum[c0];
    } else {
        for (int n = tid; n < L; n += SE_THREADS) acc += (double)load_p<COMPLEX_IN>(x, s0 + n, front_pad, n_in);
    }
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
                           float* out, int average, cudaStream_t stream, bool scratch) {
    if (n_out <= 0) return XC_OK;
    const int64_t blocks = (n_out + SE_TILE - 1) / SE_TILE;
    const double* cum = nullptr;
    if (scratch && L >= SE_LONG_L && front_pad == 0 && n_out - 1 + (int64_t)L <= n_in) {
        // chunk sums of everything the windows cover, then their running totals
        int64_t cover = n_out - 1 + L;
        if (cover > n_in) cover = n_in;
        const int64_t nch = cover / SE_CHUNK;          // whole chunks only (c1 <= (s0 + L) / CHUNK <= cover / CHUNK)
        double* d = reinterpret_cast<double*>(workspace(SLOT_SCAN, sizeof(double) * (size_t)(2 * nch + 2)));
        if (!d) return XC_ERR_NOMEM;
        if (nch > 0) {
            if (complex_in) chunk_sum_kernel<true><<<(unsigned)nch, 256, 0, stream>>>(x, n_in, d + nch + 1);
            else chunk_sum_kernel<false><<<(unsigned)nch, 256, 0, stream>>>(x, n_in, d + nch + 1);
            XC_LAUNCH_CHECK();
        }
        chunk_scan_kernel<<<1, 1024, 0, stream>>>(d + nch + 1, nch, d);
        XC_LAUNCH_CHECK();
        cum = d;
    }
    if (complex_in)
        sliding_sum_kernel<true><<<(unsigned)blocks, SE_THREADS, 0, stream>>>(x, n_in, front_pad, L, n_out, out, average, cum);
    else
        sliding_sum_kernel<false><<<(unsigned)blocks, SE_THREADS, 0, stream>>>(x, n_in, front_pad, L, n_out, out, average, cum);
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

// one long vector: chunk sums by many CTAs, then their total in a fixed order
__global__ void __launch_bounds__(256) normsq_final_kernel(const double* __restrict__ part, int64_t n, const float2* __restrict__ tail,
                                                           int64_t ntail, float* __restrict__ out) {
    __shared__ double s_part[8];
    double acc = 0.0;
    for (int64_t i = threadIdx.x; i < n; i += 256) acc += part[i];
    for (int64_t i = threadIdx.x; i < ntail; i += 256) {
        const float2 v = tail[i];
        acc += (double)v.x * v.x + (double)v.y * v.y;
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double w = 0.0;
        for (int i = 0; i < 8; ++i) w += s_part[i];
        out[0] = (float)w;
    }
}

int32_t normsq_device(const float2* x, int64_t n, int64_t rows, int64_t row_stride, float* out, cudaStream_t stream, bool scratch) {
    if (scratch && rows == 1 && n >= 16 * SE_CHUNK) {
        const int64_t nch = n / SE_CHUNK;
        double* d = reinterpret_cast<double*>(workspace(SLOT_SCAN, sizeof(double) * (size_t)(2 * nch + 2)));
        if (!d) return XC_ERR_NOMEM;
        chunk_sum_kernel<true><<<(unsigned)nch, 256, 0, stream>>>(x, n, d);
        XC_LAUNCH_CHECK();
        normsq_final_kernel<<<1, 256, 0, stream>>>(d, nch, x + nch * SE_CHUNK, n - nch * SE_CHUNK, out);
        XC_LAUNCH_CHECK();
        return XC_OK;
    }
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
