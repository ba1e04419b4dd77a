// fused_blue_pp.cuh -- frequency search for cutout lengths that are NOT a power of two (Bluestein, see bluestein.cuh) in the
// ping-pong layout of fused_fft_pp.cuh: B free-running thread groups per SM over one staged rx window, every per-thread
// constant in tensor memory -- the chirped cutout conj(cutout[n] b[n]) (zero beyond the true length, so the padded part of
// the window needs no special loads), both folded twiddle sets, and FFT_M(chirp) in the register order the first transform
// leaves.  Per shift: FFT #1 (pp_block_fft_fold), times FFT_M(chirp), conjugate, reorder through the exchange buffer,
// FFT #2 from registers, |.|^2 of the bins below `nbins`, group maximum and lowest arg max.  (max, arg max) outputs only.
#pragma once
#include "fused_fft_pp.cuh"

namespace xc {

template <class PL, int B>
struct BluePPLayout {
    using L = PPLayout<PL, B, true>;
    static constexpr int COL_BH = L::CW;                                  // FFT_M(chirp)[output_bin(t, slot)], slot < V
    static constexpr int CW = L::CW + 2 * PL::V;
    static_assert(CW * L::COPIES <= 512, "tensor memory holds 512 columns");
    static constexpr int TMEM_COLS = (CW * L::COPIES <= 256) ? 256 : 512;
};

// second transform: the inputs are already in the registers in input_index order (no product), raw stores
template <class PL, int B, class Bar>
__device__ __forceinline__ void blue_block_fft_regs(float2* v, const uint32_t tb, float2* buf, const int t, Bar bar) {
    using L = PPLayout<PL, B, true>;
    constexpr int RL = PL::RL, ROW = PL::ROW, R1 = PL::R1, R2 = PL::R2, NQ1 = L::NQ1, NQ2 = L::NQ2;
    constexpr int K1STEP = (PL::M1 / RL) * ROW;
    constexpr int QSTEP1 = (PL::TG / RL) * ROW;
    const int o1 = (t / RL) * ROW + (t % RL);
    uint32_t cu[2][16];
#pragma unroll
    for (int q = 0; q < NQ1; ++q) {
        codelet_fft<R1>(v + q * R1);
#pragma unroll
        for (int j = 0; j < R1; ++j) buf[o1 + bitrev(j, ilog2(R1)) * K1STEP + q * QSTEP1] = v[q * R1 + j];
    }
    tc::tmem_ld16_issue(tb + L::COL_TW1, cu[0]);
    bar();
    constexpr int QSTEP2 = (PL::TG / PL::M2) * K1STEP;
    const int o2 = (t / PL::M2) * K1STEP + (t % PL::M2);
#pragma unroll
    for (int q = 0; q < NQ2; ++q)
#pragma unroll
        for (int n2 = 0; n2 < R2; ++n2) v[q * R2 + n2] = buf[o2 + q * QSTEP2 + n2 * ROW];
#pragma unroll
    for (int c = 0; c < 4 * NQ2; ++c) {
        tc::tmem_ld_wait16(cu[c & 1]);
        if (c + 1 < 4 * NQ2) tc::tmem_ld16_issue(tb + L::COL_TW1 + 16 * (c + 1), cu[(c + 1) & 1]);
        folded_stage_chunk(v + (c / 4) * R2, c % 4, cu[c & 1]);
    }
#pragma unroll
    for (int q = 0; q < NQ2; ++q) {
        codelet_fft<R2, 2>(v + q * R2);
#pragma unroll
        for (int j = 0; j < R2; ++j) buf[o2 + q * QSTEP2 + bitrev(j, ilog2(R2)) * ROW] = v[q * R2 + j];
    }
    tc::tmem_ld16_issue(tb + L::COL_TW2, cu[0]);
    __syncwarp();
    const float4* const b3 = reinterpret_cast<const float4*>(buf + t * ROW);
#pragma unroll
    for (int q = 0; q < NQ2; ++q)
#pragma unroll
        for (int d2 = 0; d2 < RL / 2; ++d2) {
            const float4 y = b3[(q * PL::TG * ROW) / 2 + d2];
            v[q * RL + 2 * d2] = make_float2(y.x, y.y);
            v[q * RL + 2 * d2 + 1] = make_float2(y.z, y.w);
        }
#pragma unroll
    for (int m = 0; m < 4; ++m) {
        tc::tmem_ld_wait16(cu[m & 1]);
        if (m + 1 < 4) tc::tmem_ld16_issue(tb + L::COL_TW2 + 16 * (m + 1), cu[(m + 1) & 1]);
#pragma unroll
        for (int q = 0; q < NQ2; ++q) folded_stage_chunk(v + q * RL, m, cu[m & 1]);
    }
#pragma unroll
    for (int q = 0; q < NQ2; ++q) codelet_fft<RL, 2>(v + q * RL);
}

struct BlueArgs {
    const float2* bhat;   // [M] FFT_M(chirp), natural bin order
    int nvalid;           // true cutout length N (samples of the window that must lie inside rx)
    int nbins;            // bins k < nbins count
    float inv_m2;         // 1 / M^2
    const float2* post;   // [nbins] post-chirp exp(+i 2 pi w k^2 / 2): X[k] = conj(post[k] v[k]) / M (complex QF only)
};

// CPLX: also the complex QF at the arg max (cztXcorr's default output, xcorrRoutines.py:413-457): candidate (bin, value) kept in
// registers until the winning bin is known one FFT later, alternating bin slots (as fused_fft_pp_kernel)
template <class PL, int B, bool CPLX>
__global__ void __launch_bounds__(PL::TG* B, 1) fused_blue_pp_kernel(const StagedArgs a, const BlueArgs ba, const FusedOut out,
                                                                     const int64_t K, const int chunk) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    using L = PPLayout<PL, B, true>;
    using BL = BluePPLayout<PL, B>;
    constexpr int TG = PL::TG, V = PL::V, W = L::W;
    const int g = threadIdx.x / TG, t = threadIdx.x % TG;
    const int warp_cta = threadIdx.x >> 5;
    float2* const buf = reinterpret_cast<float2*>(smem_raw + L::OFF_BUF) + (size_t)g * PL::BUF;
    float2* const stage = reinterpret_cast<float2*>(smem_raw + L::OFF_STAGE);
    unsigned int* const gmax = reinterpret_cast<unsigned int*>(smem_raw + L::OFF_GMAX) + 2 * g;
    unsigned int* const binmin = reinterpret_cast<unsigned int*>(smem_raw + L::OFF_BIN) + 2 * g;
    uint32_t* const taddr_slot = reinterpret_cast<uint32_t*>(smem_raw + L::OFF_TADDR);
    if (warp_cta == 0) tc::tmem_alloc(tc::smem_u32(taddr_slot), BL::TMEM_COLS);
    if (t == 0) {
        binmin[0] = 0xffffffffu;
        binmin[1] = 0xffffffffu;
        gmax[0] = 0u;
        gmax[1] = 0u;
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem0 = *taddr_slot;
    // lane quarter of this warp, column block of the warps of a group that share a quarter
    const uint32_t tb = tmem0 + (((uint32_t)(warp_cta & 3) * 32u) << 16) + (uint32_t)((t >> 7) * BL::CW);
    if (g < L::WRITERS) {
        float r[16];
#pragma unroll
        for (int c = 0; c < V / 8; ++c) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int j = 8 * c + i;
                const float2 cv = a.cutout[input_index<PL>(t, j / PL::R1, j % PL::R1)];
                r[2 * i] = cv.x;
                r[2 * i + 1] = -cv.y;
            }
            tc::tmem_st16(tb + L::COL_CUT + 16 * c, r);
        }
#pragma unroll 1
        for (int c = 0; c < 4 * L::NQ2; ++c) {
            const int e = t + TG * (c / 4), k1 = e / PL::M2, d = e % PL::M2;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int n2 = 4 * (c % 4) + (i >> 1) + 16 * (i & 1);
                const float2 w = unit_root(((n2 * PL::M2 + d) * k1) % PL::N, PL::N);
                r[2 * i] = w.x;
                r[2 * i + 1] = w.y;
            }
            tc::tmem_st16(tb + L::COL_TW1 + 16 * c, r);
        }
        const int k2 = t % PL::R2;
#pragma unroll 1
        for (int m = 0; m < 4; ++m) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int d = 4 * m + (i >> 1) + 16 * (i & 1);
                const float2 w = unit_root((d * k2) % PL::M1, PL::M1);
                r[2 * i] = w.x;
                r[2 * i + 1] = w.y;
            }
            tc::tmem_st16(tb + L::COL_TW2 + 16 * m, r);
        }
#pragma unroll
        for (int c = 0; c < V / 8; ++c) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int sl = 8 * c + i;
                const float2 w = ba.bhat[output_bin<PL>(t, sl / PL::RL, sl % PL::RL)];
                r[2 * i] = w.x;
                r[2 * i + 1] = w.y;
            }
            tc::tmem_st16(tb + BL::COL_BH + 16 * c, r);
        }
        tc::tmem_st_wait();
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();

    const int64_t i0 = (int64_t)blockIdx.x * B * (int64_t)chunk;
    int64_t n_left = K - i0;
    if (n_left < 0) n_left = 0;
    if (n_left > B * (int64_t)chunk) n_left = B * (int64_t)chunk;
    const int n_act = (int)n_left;
    const int step = a.step;
    const int64_t s_first = a.shift0 + i0 * step;
    const GroupBar<TG, B> bar{1 + g};
    const unsigned int bbase = (unsigned int)output_bin_base<PL>(t);
    float2 v[V];
    float pend = 0.f;
    float2 cand = make_float2(0.f, 0.f);
    unsigned int cand_bin = 0xffffffffu;
    int o = W + 1;

    for (int it = 0; it < chunk; ++it) {
        const int e = B * it + g;
        if (o + (B - 1) * step > W) {
            __syncthreads();
            const int64_t base = s_first + (int64_t)B * it * step;
            for (int j = threadIdx.x; j < PL::N + W; j += TG * B) {
                const int64_t gi = base + j;
                stage[j] = (gi >= 0 && gi < a.rxlen) ? __ldg(a.rx + gi) : make_float2(0.f, 0.f);
            }
            o = 0;
            __syncthreads();
        }
        const int64_t s = s_first + (int64_t)e * step;
        const bool ok = e < n_act && s >= 0 && s + ba.nvalid <= a.rxlen;
        const float2* const xw = stage + o + g * step + t;
        o += B * step;
        // ---- FFT #1 of the chirped product (the barrier inside also orders the previous shift's bin atomics) ----
        pp_block_fft_fold<PL, B>(v, xw, tb, buf, t, bar);
        unsigned int* const bslot = binmin + (CPLX ? (it & 1) : 0);
        if (t == 0 && it > 0) {
            const int p = e - B;
            if (p < n_act) {
                const unsigned int b = CPLX ? binmin[(it & 1) ^ 1] : *binmin;
                if (out.qf2) out.qf2[i0 + p] = pend;
                if (out.fidx) out.fidx[i0 + p] = (b == 0xffffffffu) ? 0u : b;
            }
            if (!CPLX) *binmin = 0xffffffffu;
        }
        if constexpr (CPLX) {
            if (cand_bin != 0xffffffffu && cand_bin == binmin[(it & 1) ^ 1]) {
                const float sc = sqrtf(ba.inv_m2) / (sqrtf(*out.refnormsq) * sqrtf(out.energy[i0 + e - B]));
                const float2 pz = cmul(__ldg(ba.post + cand_bin), cand);
                out.cplx[i0 + e - B] = make_float2(pz.x * sc, -pz.y * sc);
            }
            cand_bin = 0xffffffffu;
        }
        // ---- times FFT_M(chirp), conjugated for the inverse transform ----
        {
            uint32_t bh[2][16];
            tc::tmem_ld16_issue(tb + BL::COL_BH, bh[0]);
#pragma unroll
            for (int c = 0; c < V / 8; ++c) {
                tc::tmem_ld_wait16(bh[c & 1]);
                if (c + 1 < V / 8) tc::tmem_ld16_issue(tb + BL::COL_BH + 16 * (c + 1), bh[(c + 1) & 1]);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float2 p = pcmul(v[8 * c + i], make_float2(__uint_as_float(bh[c & 1][2 * i]), __uint_as_float(bh[c & 1][2 * i + 1])));
                    v[8 * c + i] = make_float2(p.x, -p.y);
                }
            }
        }
        bar();    // last-pass loads of FFT #1 done in every thread: the buffer is free
#pragma unroll
        for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
            for (int j = 0; j < PL::RL; ++j) {
                const int i = output_bin<PL>(t, q, j);
                buf[i + ((i >> 5) << 1)] = v[q * PL::RL + j];
            }
        bar();
#pragma unroll
        for (int n1 = 0; n1 < PL::R1; ++n1)
#pragma unroll
            for (int q = 0; q < V / PL::R1; ++q) {
                const int i = input_index<PL>(t, q, n1);
                v[q * PL::R1 + n1] = buf[i + ((i >> 5) << 1)];
            }
        bar();    // all reads done before the pass-1 stores of FFT #2
        blue_block_fft_regs<PL, B>(v, tb, buf, t, bar);
        // ---- |X[k]|^2 for k < nbins, group maximum, lowest arg max ----
        float nv[V], m8[V / 8];
#pragma unroll
        for (int j = 0; j < V; ++j)
            nv[j] = (bbase + (unsigned int)output_bin_off<PL>(j / PL::RL, j % PL::RL) < (unsigned int)ba.nbins) ? cnorm2(v[j]) : 0.f;
        float tmax = 0.f;
#pragma unroll
        for (int c = 0; c < V / 8; ++c) {
            m8[c] = fmaxf(fmaxf(fmaxf(nv[8 * c], nv[8 * c + 1]), fmaxf(nv[8 * c + 2], nv[8 * c + 3])),
                          fmaxf(fmaxf(nv[8 * c + 4], nv[8 * c + 5]), fmaxf(nv[8 * c + 6], nv[8 * c + 7])));
            tmax = fmaxf(tmax, m8[c]);
        }
        const unsigned int u = __reduce_max_sync(0xffffffffu, __float_as_uint(tmax));
        if ((t & 31) == 0) atomicMax(gmax + (it & 1), u);
        bar();
        const float m = __uint_as_float(gmax[it & 1]);
        if (t == 0) gmax[(it & 1) ^ 1] = 0u;
        if (CPLX && t == 0) binmin[(it & 1) ^ 1] = 0xffffffffu;       // read for the last time before this barrier
        if (ok && tmax == m) {
            unsigned int mybin = 0xffffffffu;
#pragma unroll
            for (int c = 0; c < V / 8; ++c) {
                if (m8[c] == m) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int j = 8 * c + i;
                        const unsigned int bj = bbase + (unsigned int)output_bin_off<PL>(j / PL::RL, j % PL::RL);
                        if (nv[j] == m && bj < (unsigned int)ba.nbins && bj < mybin) {
                            mybin = bj;
                            if (CPLX) cand = v[j];
                        }
                    }
                }
            }
            atomicMin(bslot, mybin);
            if (CPLX) cand_bin = mybin;
        }
        if (t == 0) pend = ok ? m * ba.inv_m2 / *out.refnormsq / out.energy[i0 + e] : 0.f;
        if (CPLX && t == 0 && e < n_act && !ok) out.cplx[i0 + e] = make_float2(0.f, 0.f);
    }
    bar();
    if (chunk > 0) {
        const int p = B * (chunk - 1) + g;
        const unsigned int b = binmin[CPLX ? ((chunk - 1) & 1) : 0];
        if (t == 0 && p < n_act) {
            if (out.qf2) out.qf2[i0 + p] = pend;
            if (out.fidx) out.fidx[i0 + p] = (b == 0xffffffffu) ? 0u : b;
        }
        if constexpr (CPLX) {
            if (cand_bin != 0xffffffffu && cand_bin == b) {
                const float sc = sqrtf(ba.inv_m2) / (sqrtf(*out.refnormsq) * sqrtf(out.energy[i0 + p]));
                const float2 pz = cmul(__ldg(ba.post + cand_bin), cand);
                out.cplx[i0 + p] = make_float2(pz.x * sc, -pz.y * sc);
            }
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp_cta == 0) tc::tmem_dealloc(tmem0, BL::TMEM_COLS);
}

template <class PL, int B>
int32_t launch_blue_pp(const StagedArgs& a, const BlueArgs& ba, const FusedOut& out, int64_t K, cudaStream_t stream) {
    auto kern = out.cplx ? fused_blue_pp_kernel<PL, B, true> : fused_blue_pp_kernel<PL, B, false>;
    constexpr size_t smem = PPLayout<PL, B, true>::BYTES;
    if ((int)smem > max_smem_optin()) return fail(XC_ERR_UNSUPPORTED, "ping-pong Bluestein kernel needs %zu B of shared memory", smem);
    static KernelOnce once[2];
    XC_TRY(once[out.cplx ? 1 : 0].configure(kern, max_smem_optin()));
    const int64_t per = (K + B - 1) / B;
    int64_t ctas = (int64_t)sm_count();
    if (ctas > per) ctas = per;
    if (ctas < 1) ctas = 1;
    const int64_t chunk = (per + ctas - 1) / ctas;
    if (chunk >= (1ll << 29)) return fail(XC_ERR_UNSUPPORTED, "too many shifts per thread group");
    profile_begin(stream);
    kern<<<(unsigned)ctas, PL::TG * B, smem, stream>>>(a, ba, out, K, (int)chunk);
    profile_end(stream);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

}  // namespace xc
