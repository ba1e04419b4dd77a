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

// ---------------------------------------------------------------------------------------------------------------------
// Two thread groups per SM: the block spectrum X and the running (best, template) move from registers into TENSOR MEMORY
// (128 columns per warp: X 64, best 32, template 32; four warps per lane quarter = 512 columns), so that a thread needs
// 128 registers instead of 255 and sixteen warps fit an SM.
// ---------------------------------------------------------------------------------------------------------------------
#include "fused_fft_pp.cuh"

namespace xc {

template <class PL, int B>
__global__ void __launch_bounds__(PL::TG* B, 1) mt_ovs2_kernel(const MtOvsArgs a, const int64_t nblocks) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int TG = PL::TG, V = PL::V, M = PL::N;
    static_assert(V == 32 && TG * B == 512, "sixteen warps, 32 points per thread");
    const int g = threadIdx.x / TG, t = threadIdx.x % TG;
    const int warp_cta = threadIdx.x >> 5;
    float2* const tab = reinterpret_cast<float2*>(smem_raw);
    build_fft_tables<PL>(tab, threadIdx.x, TG * B);
    float2* const buf = reinterpret_cast<float2*>(smem_raw + PL::TABLE_BYTES) + (size_t)g * PL::BUF;
    uint32_t* const taddr_slot = reinterpret_cast<uint32_t*>(smem_raw + PL::TABLE_BYTES + sizeof(float2) * (size_t)PL::BUF * B);
    if (warp_cta == 0) tc::tmem_alloc(tc::smem_u32(taddr_slot), 512);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem0 = *taddr_slot;
    const uint32_t tb = tmem0 + (((uint32_t)(warp_cta & 3) * 32u) << 16) + (uint32_t)((warp_cta >> 2) * 128);
    constexpr int COL_X = 0, COL_BEST = 64, COL_IDX = 96;
    FftThread<PL> th;
    th.init(t);
    const GroupBar<TG, B> bar{1 + g};
    const unsigned int bbase = (unsigned int)output_bin_base<PL>(t);
    float2 v[V];
    const int64_t ngroups = (int64_t)gridDim.x * B;
    const int64_t trips = (nblocks + ngroups - 1) / ngroups;
    for (int64_t it = 0; it < trips; ++it) {
        const int64_t blk = it * ngroups + (int64_t)blockIdx.x * B + g;
        const bool act = blk < nblocks;
        const int64_t s0 = a.start + blk * a.S;
        constexpr int NQ = V / PL::R1;
#pragma unroll
        for (int n1 = 0; n1 < PL::R1; ++n1)
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                const int64_t p = s0 + input_index<PL>(t, q, n1);
                v[q * PL::R1 + n1] = (act && p >= 0 && p < a.xlen) ? __ldg(a.x + p) : make_float2(0.f, 0.f);
            }
        block_fft<PL>(v, th, tab, buf, t, bar);
        {
            float r[16];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    r[2 * i] = v[8 * c + i].x;
                    r[2 * i + 1] = v[8 * c + i].y;
                }
                tc::tmem_st16(tb + COL_X + 16 * c, r);
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) r[i] = 0.f;
            tc::tmem_st16(tb + COL_BEST, r);
            tc::tmem_st16(tb + COL_BEST + 16, r);
            tc::tmem_st16(tb + COL_IDX, r);
            tc::tmem_st16(tb + COL_IDX + 16, r);
            tc::tmem_st_wait();
        }
#pragma unroll 1
        for (int tt = 0; tt < a.T; ++tt) {
            const float2* const Ht = a.H + (size_t)tt * M + t;
            uint32_t xu[2][16];
            tc::tmem_ld16_issue(tb + COL_X, xu[0]);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                tc::tmem_ld_wait16(xu[c & 1]);
                if (c + 1 < 4) tc::tmem_ld16_issue(tb + COL_X + 16 * (c + 1), xu[(c + 1) & 1]);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int sl = 8 * c + i;
                    const float2 p = pcmul(make_float2(__uint_as_float(xu[c & 1][2 * i]), __uint_as_float(xu[c & 1][2 * i + 1])),
                                           __ldg(Ht + (size_t)sl * TG));
                    v[sl] = make_float2(p.x, -p.y);
                }
            }
            bar();    // last-pass loads of the previous transform are done: the buffer is free
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
            const float ie = 1.f / __ldg(a.tenergy + tt);
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                uint32_t bu[16], iu[16];
                tc::tmem_ld16_issue(tb + COL_BEST + 16 * c, bu);
                tc::tmem_ld16_issue(tb + COL_IDX + 16 * c, iu);
                tc::tmem_ld_wait16(bu);
                tc::tmem_ld_wait16(iu);
                float rb[16], ri[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const float val = cnorm2(v[16 * c + i]) * ie;
                    const float cur = __uint_as_float(bu[i]);
                    const bool up = val > cur;          // strict: the lowest template wins ties, index 0 when nothing is > 0
                    rb[i] = up ? val : cur;
                    ri[i] = up ? __int_as_float(tt) : __uint_as_float(iu[i]);
                }
                tc::tmem_st16(tb + COL_BEST + 16 * c, rb);
                tc::tmem_st16(tb + COL_IDX + 16 * c, ri);
            }
            tc::tmem_st_wait();
        }
        {
            uint32_t bu[2][16], iu[2][16];
            tc::tmem_ld16_issue(tb + COL_BEST, bu[0]);
            tc::tmem_ld16_issue(tb + COL_BEST + 16, bu[1]);
            tc::tmem_ld16_issue(tb + COL_IDX, iu[0]);
            tc::tmem_ld16_issue(tb + COL_IDX + 16, iu[1]);
            tc::tmem_ld_wait16(bu[0]);
            tc::tmem_ld_wait16(bu[1]);
            tc::tmem_ld_wait16(iu[0]);
            tc::tmem_ld_wait16(iu[1]);
            if (act) {
#pragma unroll
                for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                    for (int j = 0; j < PL::RL; ++j) {
                        const int sl = q * PL::RL + j;
                        const unsigned int s = bbase + (unsigned int)output_bin_off<PL>(q, j);
                        const int64_t i = blk * a.S + s;
                        if (s < (unsigned int)a.S && i < a.K) {
                            const float b = __uint_as_float(bu[sl >> 4][sl & 15]) * a.inv_m2 / a.xenergy[i];
                            if (a.best) a.best[i] = a.sqrt_out ? sqrtf(b) : b;
                            if (a.tidx) a.tidx[i] = (int32_t)iu[sl >> 4][sl & 15];
                        }
                    }
            }
        }
        bar();
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp_cta == 0) tc::tmem_dealloc(tmem0, 512);
}

}  // namespace xc
