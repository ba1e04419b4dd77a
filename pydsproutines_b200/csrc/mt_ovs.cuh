// mt_ovs.cuh -- multi-template sliding correlator (returnMax outputs) as a fast correlation: for templates of 512 ... 4096
// samples over long records the 8 L flops per (template, shift) of the dense contraction (tensor cores, mt_tc.cu) lose to
// overlap-save through 8192-point blocks: ONE forward transform of the x block serves all T templates, each template costs a
// spectrum product (table H_t made at first use) and one more transform; a thread keeps the running (best, template) of its
// 32 output positions in registers across the template loop.  C4 (256 x 1024 vs 2^24 samples): 2.1e6 flops per shift as a
// contraction, 1.9e4 through blocks of 7169 shifts.
#pragma once
#include "ovs.cuh"

namespace xc {

// H[t][slot TG + tid] = conj(FFT_M(conj(tmpl_t) padded)[output_bin(tid, slot)]): d_t[s] = sum_n tmpl_t[n] x[s + n] = IFFT(X H_t)[s]
template <class PL>
__global__ void __launch_bounds__(PL::TG) mt_ovs_table_kernel(const float2* __restrict__ templates, const int L, const int L8,
                                                              float2* __restrict__ H) {
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
    const float2* const tm = templates + (size_t)blockIdx.x * L8;
    float2 v[V];
#pragma unroll
    for (int n1 = 0; n1 < PL::R1; ++n1)
#pragma unroll
        for (int q = 0; q < V / PL::R1; ++q) {
            const int n = input_index<PL>(t, q, n1);
            const float2 c = n < L ? tm[n] : make_float2(0.f, 0.f);
            v[q * PL::R1 + n1] = make_float2(c.x, -c.y);
        }
    block_fft<PL>(v, th, tab, buf, t, bar);
    float2* const Ht = H + (size_t)blockIdx.x * PL::N;
#pragma unroll
    for (int sl = 0; sl < V; ++sl) Ht[(size_t)sl * TG + t] = make_float2(v[sl].x, -v[sl].y);
}

struct MtOvsArgs {
    const float2* x;
    int64_t xlen, start, K;
    int S, T;
    const float2* H;            // [T][M]
    const float* tenergy;       // [T]
    const float* xenergy;       // [K]
    float* best;                // [K] or NULL
    int32_t* tidx;              // [K] or NULL
    int sqrt_out;
    float inv_m2;
};

template <class PL>
__global__ void __launch_bounds__(PL::TG) mt_ovs_kernel(const MtOvsArgs a, const int64_t nblocks) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int TG = PL::TG, V = PL::V, M = PL::N;
    const int t = threadIdx.x;
    float2* const tab = reinterpret_cast<float2*>(smem_raw);
    build_fft_tables<PL>(tab, t, TG);
    float2* const buf = reinterpret_cast<float2*>(smem_raw + PL::TABLE_BYTES);
    float2* const rbuf = buf + PL::BUF;                 // M + M / 8 cells
    __syncthreads();
    FftThread<PL> th;
    th.init(t);
    const GroupBar<TG, 1> bar{1};
    const unsigned int bbase = (unsigned int)output_bin_base<PL>(t);
    float2 v[V], X[V];
    float best[V];
    int bidx[V];
    for (int64_t blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
        const int64_t s0 = a.start + blk * a.S;
        constexpr int NQ = V / PL::R1;
#pragma unroll
        for (int n1 = 0; n1 < PL::R1; ++n1)
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                const int64_t p = s0 + input_index<PL>(t, q, n1);
                v[q * PL::R1 + n1] = (p >= 0 && p < a.xlen) ? __ldg(a.x + p) : make_float2(0.f, 0.f);
            }
        block_fft<PL>(v, th, tab, buf, t, bar);
#pragma unroll
        for (int sl = 0; sl < V; ++sl) {
            X[sl] = v[sl];
            best[sl] = 0.f;
            bidx[sl] = 0;
        }
#pragma unroll 1
        for (int tt = 0; tt < a.T; ++tt) {
            const float2* const Ht = a.H + (size_t)tt * M + t;
#pragma unroll
            for (int sl = 0; sl < V; ++sl) {
                const float2 p = pcmul(X[sl], __ldg(Ht + (size_t)sl * TG));
                v[sl] = make_float2(p.x, -p.y);
            }
            // reorder from output bins to input order through a buffer of its own, one pad cell per 8 cells: the stores (bins at
            // a stride of R1 = 8 cells across the lanes) and the loads (consecutive cells) are both conflict free, and the
            // transform's exchange buffer needs no barrier of its own (the previous reads of rbuf are behind the barrier
            // inside the previous transform; the barrier below also says that everyone is done with the last pass of it)
#pragma unroll
            for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                for (int j = 0; j < PL::RL; ++j) {
                    const int i = output_bin<PL>(t, q, j);
                    rbuf[i + (i >> 3)] = v[q * PL::RL + j];
                }
            bar();
#pragma unroll
            for (int n1 = 0; n1 < PL::R1; ++n1)
#pragma unroll
                for (int q = 0; q < NQ; ++q) {
                    const int i = input_index<PL>(t, q, n1);
                    v[q * PL::R1 + n1] = rbuf[i + (i >> 3)];
                }
            block_fft<PL>(v, th, tab, buf, t, bar);
            const float ie = 1.f / __ldg(a.tenergy + tt);
#pragma unroll
            for (int sl = 0; sl < V; ++sl) {
                const float val = cnorm2(v[sl]) * ie;
                if (val > best[sl]) {         // strict: the lowest template wins ties, index 0 when nothing is > 0
                    best[sl] = val;
                    bidx[sl] = tt;
                }
            }
        }
#pragma unroll
        for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
            for (int j = 0; j < PL::RL; ++j) {
                const unsigned int s = bbase + (unsigned int)output_bin_off<PL>(q, j);
                const int64_t i = blk * a.S + s;
                if (s < (unsigned int)a.S && i < a.K) {
                    const float b = best[q * PL::RL + j] * a.inv_m2 / a.xenergy[i];
                    if (a.best) a.best[i] = a.sqrt_out ? sqrtf(b) : b;
                    if (a.tidx) a.tidx[i] = bidx[q * PL::RL + j];
                }
            }
        bar();
    }
}

}  // namespace xc
