// ovs.cuh -- fastXcorr WITHOUT frequency search over a contiguous shift range as a fast correlation (overlap-save):
//     z[s] = sum_n rx[s + n] conj(cutout[n])      (np.vdot(rx[s:s+N], cutout) = conj(z[s]), xcorrRoutines.py:483-509)
// costs 8 N flops per shift as a sliding dot product, and 2 x 5 M log2 M / (M - N + 1) per shift through M-point blocks:
//     Z = FFT_M(rx[b S : b S + M]) conj(FFT_M(cutout, zero padded)),  z[b S + s] = IFFT_M(Z)[s],  s < S = M - N + 1.
// One thread group per block, on the block engine of fused_fft_kernel.cuh: forward transform, times the table H (conj of the
// cutout's spectrum, in the register order the transform leaves), conjugate, reorder through the exchange buffer, second
// forward transform (IFFT(y) = conj(FFT(conj(y))) / M), and the epilogue normalises the S valid outputs by ||cutout||^2 and
// the window energies.  N = 4096, 2^20 shifts: 256 blocks of 8192 points instead of 3.4e10 flops.
#pragma once
#include "bluestein.cuh"
#include "plans.cuh"

namespace xc {

// H[slot TG + t] = conj(FFT_M(cutout padded)[output_bin(t, slot)])  (one group, one launch per call)
template <class PL>
__global__ void __launch_bounds__(PL::TG) ovs_table_kernel(const float2* __restrict__ cutout, const int N, float2* __restrict__ H) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int TG = PL::TG, V = PL::V;
    const int t = threadIdx.x;
    float2* const tab = reinterpret_cast<float2*>(smem_raw);
    build_fft_tables<PL>(tab, t, TG);
    float2* const buf = reinterpret_cast<float2*>(smem_raw + PL::TABLE_BYTES);
    __syncthreads();
    FftThread<PL> th;
    th.init(t);
    const GroupBar<TG, 1> bar{1};
    float2 v[V];
#pragma unroll
    for (int n1 = 0; n1 < PL::R1; ++n1)
#pragma unroll
        for (int q = 0; q < V / PL::R1; ++q) {
            const int n = input_index<PL>(t, q, n1);
            v[q * PL::R1 + n1] = n < N ? cutout[n] : make_float2(0.f, 0.f);
        }
    block_fft<PL>(v, th, tab, buf, t, bar);
#pragma unroll
    for (int sl = 0; sl < V; ++sl) H[(size_t)sl * TG + t] = make_float2(v[sl].x, -v[sl].y);
}

struct OvsArgs {
    const float2* rx;
    int64_t rxlen;
    int64_t shift0;        // first shift; block b covers shifts shift0 + b S ... + S - 1
    int64_t K;
    int S;                 // valid outputs per block = M - N + 1
    const float2* H;
    const float* energy;   // [K]
    const float* refnormsq;
    float* qf2;            // [K] or NULL
    float2* cplx;          // [K] or NULL: conj(z) / (||cutout|| ||window||)
    float inv_m;           // 1 / M
};

template <class PL, int B>
__global__ void __launch_bounds__(PL::TG* B) ovs_kernel(const OvsArgs a, const int64_t nblocks) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int TG = PL::TG, V = PL::V, M = PL::N;
    const int g = threadIdx.x / TG, t = threadIdx.x % TG;
    float2* const tab = reinterpret_cast<float2*>(smem_raw);
    build_fft_tables<PL>(tab, threadIdx.x, TG * B);
    float2* const buf = reinterpret_cast<float2*>(smem_raw + PL::TABLE_BYTES) + (size_t)g * PL::BUF;
    __syncthreads();
    FftThread<PL> th;
    th.init(t);
    const GroupBar<TG, B> bar{1 + g};
    const float refn = *a.refnormsq;
    const unsigned int bbase = (unsigned int)output_bin_base<PL>(t);
    float2 v[V];
    // every group of the grid walks the blocks with the same trip count (the barriers are per group, the loop is uniform)
    const int64_t ngroups = (int64_t)gridDim.x * B;
    const int64_t trips = (nblocks + ngroups - 1) / ngroups;
    for (int64_t it = 0; it < trips; ++it) {
        const int64_t blk = it * ngroups + (int64_t)blockIdx.x * B + g;
        const bool act = blk < nblocks;
        const int64_t s0 = a.shift0 + blk * a.S;                   // rx index of the block's first sample
        constexpr int NQ = V / PL::R1;
#pragma unroll
        for (int n1 = 0; n1 < PL::R1; ++n1)
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                const int64_t p = s0 + input_index<PL>(t, q, n1);
                v[q * PL::R1 + n1] = (act && p >= 0 && p < a.rxlen) ? __ldg(a.rx + p) : make_float2(0.f, 0.f);
            }
        block_fft<PL>(v, th, tab, buf, t, bar);
#pragma unroll
        for (int sl = 0; sl < V; ++sl) {
            const float2 p = pcmul(v[sl], __ldg(a.H + (size_t)sl * TG + t));
            v[sl] = make_float2(p.x, -p.y);
        }
        bar();    // last-pass loads of the first transform are done: the buffer is free
#pragma unroll
        for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
            for (int j = 0; j < PL::RL; ++j) buf[blue_cell(output_bin<PL>(t, q, j))] = v[q * PL::RL + j];
        bar();
#pragma unroll
        for (int n1 = 0; n1 < PL::R1; ++n1)
#pragma unroll
            for (int q = 0; q < NQ; ++q) v[q * PL::R1 + n1] = buf[blue_cell(input_index<PL>(t, q, n1))];
        bar();
        block_fft<PL>(v, th, tab, buf, t, bar);
        // v[slot] = M conj(z[s]) at s = output_bin: conj(z) is np.vdot(rx[s:s+N], cutout)
        if (act) {
#pragma unroll
            for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                for (int j = 0; j < PL::RL; ++j) {
                    const unsigned int s = bbase + (unsigned int)output_bin_off<PL>(q, j);
                    const int64_t i = blk * a.S + s;
                    if (s < (unsigned int)a.S && i < a.K) {
                        const float2 d = make_float2(v[q * PL::RL + j].x * a.inv_m, v[q * PL::RL + j].y * a.inv_m);
                        const float e = a.energy[i];
                        if (a.qf2) a.qf2[i] = cnorm2(d) / refn / e;
                        if (a.cplx) {
                            const float sc = 1.f / (sqrtf(refn) * sqrtf(e));
                            a.cplx[i] = make_float2(d.x * sc, d.y * sc);
                        }
                    }
                }
        }
        bar();    // last-pass loads done before the next block's pass-1 stores
    }
}

template <int LOG2M>
int32_t ovs_run(const float2* cutout, int N, const float2* rx, int64_t rxlen, int64_t shift0, int64_t K, const float* energy,
                const float* refnormsq, float* qf2, float2* cplx, cudaStream_t stream) {
    using PF = PlanFor<LOG2M>;
    using PL = typename PF::PL;
    constexpr int B = PF::B, M = PL::N;
    float2* H = reinterpret_cast<float2*>(workspace(SLOT_TMP2, sizeof(float2) * (size_t)M));
    if (!H) return XC_ERR_NOMEM;
    {
        auto kern = ovs_table_kernel<PL>;
        constexpr size_t smem = PL::TABLE_BYTES + sizeof(float2) * (size_t)PL::BUF;
        static KernelOnce once;
        XC_TRY(once.configure(kern, max_smem_optin()));
        kern<<<1, PL::TG, smem, stream>>>(cutout, N, H);
        XC_LAUNCH_CHECK();
    }
    auto kern = ovs_kernel<PL, B>;
    constexpr size_t smem = PL::TABLE_BYTES + sizeof(float2) * (size_t)PL::BUF * B;
    if ((int)smem > max_smem_optin()) return fail(XC_ERR_UNSUPPORTED, "overlap-save kernel needs %zu B of shared memory", smem);
    static KernelOnce once;
    XC_TRY(once.configure(kern, max_smem_optin()));
    const int S = M - N + 1;
    const int64_t nblocks = (K + S - 1) / S;
    int64_t ctas = (nblocks + B - 1) / B;
    if (ctas > (int64_t)sm_count()) ctas = sm_count();
    OvsArgs a{rx, rxlen, shift0, K, S, H, energy, refnormsq, qf2, cplx, (float)(1.0 / (double)M)};
    profile_begin(stream);
    kern<<<(unsigned)ctas, PL::TG * B, smem, stream>>>(a, nblocks);
    profile_end(stream);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

}  // namespace xc
