// misc.cu -- small reductions around the correlators: the per-call peak (largest QF^2, lowest index
// on ties), row arg max (argmax.cu:93-152) and local-maxima peak finding (peakfinding.cu:14-58).
#include "internal.cuh"

namespace xc {

__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long v) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
        const unsigned long long y = __shfl_xor_sync(0xffffffffu, v, o);
        v = y > v ? y : v;
    }
    return v;
}

// ---- per-call peak -----------------------------------------------------------------------------
__global__ void peak_reduce_kernel(const float* __restrict__ qf2, int64_t K, unsigned long long* __restrict__ packed) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    unsigned long long best = 0ull;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < K; i += stride) {
        float v = qf2[i];
        if (!(v >= 0.f)) v = 0.f;   // NaN (zero-energy windows) never wins
        const unsigned long long p = pack_peak(v, 0xffffffffu - (uint32_t)i);
        best = p > best ? p : best;
    }
    best = warp_max_u64(best);
    if ((threadIdx.x & 31) == 0) atomicMax(packed, best);
}

// record[0] = packed (value bits, ~index); record[1] = {fidx, aux}; record[2] = shifts[index]: one 24-byte copy-out
__global__ void peak_fetch_kernel(unsigned long long* __restrict__ rec, const uint32_t* __restrict__ fidx,
                                  const int32_t* __restrict__ aux, const int64_t* __restrict__ shifts) {
    const unsigned long long i = 0xffffffffull - (rec[0] & 0xffffffffull);
    const unsigned long long f = fidx ? (unsigned long long)fidx[i] : 0ull;
    const unsigned long long a = aux ? (unsigned long long)(uint32_t)aux[i] : 0ull;
    rec[1] = f | (a << 32);
    rec[2] = shifts ? (unsigned long long)shifts[i] : 0ull;
}

int32_t peak_reduce_device(const float* qf2, const uint32_t* fidx, const int32_t* aux, int64_t K, unsigned long long* d_packed,
                           cudaStream_t stream, const int64_t* shifts) {
    if (K >= (1ll << 32)) return fail(XC_ERR_UNSUPPORTED, "peak reduction supports fewer than 2^32 shifts");
    XC_CUDA_TRY(cudaMemsetAsync(d_packed, 0, sizeof(unsigned long long), stream));
    if (K <= 0) return XC_OK;
    int64_t blocks = (K + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    peak_reduce_kernel<<<(unsigned)blocks, 256, 0, stream>>>(qf2, K, d_packed);
    XC_LAUNCH_CHECK();
    peak_fetch_kernel<<<1, 1, 0, stream>>>(d_packed, fidx, aux, shifts);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

// ---- row arg max ---------------------------------------------------------------------------------
template <bool NORMSQ>
__global__ void __launch_bounds__(256) row_argmax_kernel(const float2* __restrict__ x, int64_t rows, int64_t cols,
                                                         uint32_t* __restrict__ idx, float* __restrict__ val) {
    __shared__ unsigned long long s_best[8];
    for (int64_t r = blockIdx.x; r < rows; r += gridDim.x) {
        const float2* row = x + r * cols;
        unsigned long long best = 0ull;
        for (int64_t c = threadIdx.x; c < cols; c += blockDim.x) {
            const float2 z = __ldg(row + c);
            float m = NORMSQ ? fmaf(z.x, z.x, z.y * z.y) : hypotf(z.x, z.y);
            if (!(m >= 0.f)) m = 0.f;
            const unsigned long long p = pack_peak(m, 0xffffffffu - (uint32_t)c);
            best = p > best ? p : best;
        }
        best = warp_max_u64(best);
        __syncthreads();
        if ((threadIdx.x & 31) == 0) s_best[threadIdx.x >> 5] = best;
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned long long b = s_best[0];
            for (int w = 1; w < 8; ++w) b = s_best[w] > b ? s_best[w] : b;
            idx[r] = cols > 0 ? 0xffffffffu - (uint32_t)(b & 0xffffffffull) : 0u;
            if (val) val[r] = __uint_as_float((uint32_t)(b >> 32));
        }
    }
}

int32_t row_argmax_device(const float2* x, int64_t rows, int64_t cols, int use_normsq, uint32_t* idx, float* val,
                          cudaStream_t stream) {
    if (rows <= 0) return XC_OK;
    if (cols >= (1ll << 32)) return fail(XC_ERR_UNSUPPORTED, "row argmax supports fewer than 2^32 columns");
    int64_t blocks = rows;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    if (use_normsq)
        row_argmax_kernel<true><<<(unsigned)blocks, 256, 0, stream>>>(x, rows, cols, idx, val);
    else
        row_argmax_kernel<false><<<(unsigned)blocks, 256, 0, stream>>>(x, rows, cols, idx, val);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

// ---- local maxima: count per block, scan, ordered compaction ------------------------------------------
constexpr int LM_THREADS = 256;
constexpr int LM_PER_THREAD = 4;
constexpr int LM_TILE = LM_THREADS * LM_PER_THREAD;

__device__ __forceinline__ bool is_peak(const float* __restrict__ x, int64_t n, int64_t i, float h) {
    if (i >= n) return false;
    const float c = x[i];
    const float l = i > 0 ? x[i - 1] : 0.f;
    const float r = i + 1 < n ? x[i + 1] : 0.f;
    return c > h && c > l && c > r;
}

__global__ void __launch_bounds__(LM_THREADS) lm_count_kernel(const float* __restrict__ x, int64_t n, float h,
                                                              int32_t* __restrict__ block_counts) {
    __shared__ int s_cnt;
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * LM_TILE + (int64_t)threadIdx.x * LM_PER_THREAD;
    int c = 0;
#pragma unroll
    for (int r = 0; r < LM_PER_THREAD; ++r) c += is_peak(x, n, base + r, h) ? 1 : 0;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&s_cnt, c);
    __syncthreads();
    if (threadIdx.x == 0) block_counts[blockIdx.x] = s_cnt;
}

// single CTA: exclusive scan of the block counts in place, total written to *count
__global__ void __launch_bounds__(1024) lm_scan_kernel(int32_t* __restrict__ block_counts, int64_t nblocks,
                                                       int32_t* __restrict__ count) {
    __shared__ int s_warp[32];
    __shared__ int s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (int64_t b0 = 0; b0 < nblocks; b0 += 1024) {
        const int64_t b = b0 + threadIdx.x;
        const int v = b < nblocks ? block_counts[b] : 0;
        int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, incl, o);
            if ((threadIdx.x & 31) >= o) incl += y;
        }
        if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = incl;
        __syncthreads();
        int wbase = 0;
        for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) wbase += s_warp[w];
        const int carry = s_carry;
        if (b < nblocks) block_counts[b] = carry + wbase + incl - v;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = carry + wbase + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) *count = s_carry;
}

__global__ void __launch_bounds__(LM_THREADS) lm_write_kernel(const float* __restrict__ x, int64_t n, float h,
                                                              const int32_t* __restrict__ block_offsets,
                                                              int32_t* __restrict__ idx, int32_t cap) {
    __shared__ int s_warp[LM_THREADS / 32];
    const int64_t base = (int64_t)blockIdx.x * LM_TILE + (int64_t)threadIdx.x * LM_PER_THREAD;
    bool f[LM_PER_THREAD];
    int c = 0;
#pragma unroll
    for (int r = 0; r < LM_PER_THREAD; ++r) {
        f[r] = is_peak(x, n, base + r, h);
        c += f[r] ? 1 : 0;
    }
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, incl, o);
        if ((threadIdx.x & 31) >= o) incl += y;
    }
    if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = incl;
    __syncthreads();
    int pos = block_offsets[blockIdx.x] + incl - c;
    for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) pos += s_warp[w];
#pragma unroll
    for (int r = 0; r < LM_PER_THREAD; ++r)
        if (f[r]) {
            if (pos < cap) idx[pos] = (int32_t)(base + r);
            ++pos;
        }
}

int32_t local_maxima_device(const float* x, int64_t n, float min_height, int32_t* idx, int32_t cap, int32_t* count,
                            cudaStream_t stream) {
    if (n >= (1ll << 31)) return fail(XC_ERR_UNSUPPORTED, "peak finding supports fewer than 2^31 samples");
    if (n <= 0) {
        XC_CUDA_TRY(cudaMemsetAsync(count, 0, sizeof(int32_t), stream));
        return XC_OK;
    }
    const int64_t nblocks = (n + LM_TILE - 1) / LM_TILE;
    int32_t* d_counts = reinterpret_cast<int32_t*>(workspace(SLOT_TMP2, sizeof(int32_t) * (size_t)nblocks));
    if (!d_counts) return XC_ERR_NOMEM;
    lm_count_kernel<<<(unsigned)nblocks, LM_THREADS, 0, stream>>>(x, n, min_height, d_counts);
    XC_LAUNCH_CHECK();
    lm_scan_kernel<<<1, 1024, 0, stream>>>(d_counts, nblocks, count);
    XC_LAUNCH_CHECK();
    lm_write_kernel<<<(unsigned)nblocks, LM_THREADS, 0, stream>>>(x, n, min_height, d_counts, idx, cap);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

}  // namespace xc
