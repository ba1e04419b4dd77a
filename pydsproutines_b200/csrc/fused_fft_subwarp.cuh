// fused_fft_subwarp.cuh -- frequency search over a shift range at N = 128 ... 512: TG = N / 32 lanes per FFT, 32 / TG
// FFTs (consecutive shifts) per warp in lockstep.
//
// The layout of fused_fft_warp.cuh one size down: every lane still holds 32 points, N = 32 TG.
//   n = TG n1 + c, k = k1 + 32 k2:   pass 1: lane c (of its sub-group), DFT_32 over n1, the conj(cutout) product folded
//   into the first stage; exchange cell (k1, c) at k1 ROWP + c of the sub-group's buffer; pass 2: lane l takes the rows
//   k1 = l + TG j (j < 32 / TG), each a DFT_TG over c with the twiddles W_N^(c k1) folded into its first stage.
// Per-lane constants (conj(cutout) of the lane's column, the 32 pass-2 twiddles) in tensor memory, one staged rx window per
// CTA, no barrier wider than a warp; the epilogue reduces over the TG lanes of a sub-group (REDUX with a lane mask).
#pragma once
#include "fused_fft_warp.cuh"

namespace xc {

template <int LTG, int B>
struct SubwarpLayout {
    static_assert(LTG >= 1 && LTG <= 4, "64 ... 512 points (64 measured slower than the round-1 layout: not instantiated)");
    static constexpr int TG = 1 << LTG, N = 32 * TG, FPW = 32 / TG;
    static_assert(B % 4 == 0 && B >= 4 && B <= 32, "warps 0..3 write the constants of the four lane quarters");
    static constexpr int W = 2048;
    static constexpr int ROWP = TG + 2;                          // cells per exchange row (16-byte aligned, odd 16-byte pitch)
    // consecutive sub-groups of a warp sit side by side in bank space (stride = TG cells modulo 16): a warp's pass-1 store
    // covers every bank exactly twice
    static constexpr int SUBSTRIDE = 32 * ROWP + (TG < 16 ? TG : 0);
    static_assert(SUBSTRIDE % 2 == 0, "16-byte aligned sub-group buffers");
    static constexpr int WARPBUF = FPW * SUBSTRIDE;
    static constexpr size_t OFF_BUF = 0;
    static constexpr size_t OFF_STAGE = sizeof(float2) * (size_t)WARPBUF * B;
    static constexpr size_t OFF_TADDR = OFF_STAGE + sizeof(float2) * (size_t)(N + W);
    static constexpr size_t BYTES = OFF_TADDR + 16;
    static constexpr int COL_CUT = 0, COL_T2 = 64, TMEM_COLS = 128;
};

template <int LTG, int B, bool CPLX>
__global__ void __launch_bounds__(32 * B, 1) fused_fft_subwarp_kernel(const StagedArgs a, const FusedOut out, const int64_t K,
                                                                      const int chunk) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    using L = SubwarpLayout<LTG, B>;
    constexpr int TG = L::TG, N = L::N, FPW = L::FPW, ROWP = L::ROWP, W = L::W, HP = TG / 2;
    const int g = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int sub = lane >> LTG, c = lane & (TG - 1);
    float2* const buf = reinterpret_cast<float2*>(smem_raw + L::OFF_BUF) + (size_t)g * L::WARPBUF + sub * L::SUBSTRIDE;
    float2* const stage = reinterpret_cast<float2*>(smem_raw + L::OFF_STAGE);
    uint32_t* const taddr_slot = reinterpret_cast<uint32_t*>(smem_raw + L::OFF_TADDR);
    if (g == 0) tc::tmem_alloc(tc::smem_u32(taddr_slot), L::TMEM_COLS);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem0 = *taddr_slot;
    const uint32_t tb = tmem0 + (((uint32_t)(g & 3) * 32u) << 16);
    if (g < 4) {
        float r[16];
#pragma unroll 1
        for (int m = 0; m < 4; ++m) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int n1 = 4 * m + (i >> 1) + 16 * (i & 1);
                const float2 cv = a.cutout[TG * n1 + c];
                r[2 * i] = cv.x;
                r[2 * i + 1] = -cv.y;
            }
            tc::tmem_st16(tb + L::COL_CUT + 16 * m, r);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                // pair p = (row j, column i0) and its partner i0 + TG / 2 of the radix-TG codelet of row k1 = c + TG j
                const int p = 4 * m + (i >> 1), j = p / HP, cc = p % HP + HP * (i & 1);
                const float2 w = unit_root((cc * (c + TG * j)) % N, N);
                r[2 * i] = w.x;
                r[2 * i + 1] = w.y;
            }
            tc::tmem_st16(tb + L::COL_T2 + 16 * m, r);
        }
        tc::tmem_st_wait();
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();

    // this CTA owns shift-list entries [i0, i0 + B FPW chunk); warp g, sub-group s take i0 + (B it + g) FPW + s
    constexpr int SPI = B * FPW;                       // shifts per CTA iteration
    const int64_t i0 = (int64_t)blockIdx.x * SPI * (int64_t)chunk;
    int64_t n_left = K - i0;
    if (n_left < 0) n_left = 0;
    if (n_left > SPI * (int64_t)chunk) n_left = SPI * (int64_t)chunk;
    const int n_act = (int)n_left;
    const int step = a.step;
    const int64_t s_first = a.shift0 + i0 * step;
    const float refn = *out.refnormsq;
    const unsigned int submask = (TG == 32) ? 0xffffffffu : (((1u << TG) - 1u) << (sub * TG));
    float2 v[32];
    uint32_t cu[2][16];
    int o = W + 1;

    for (int it = 0; it < chunk; ++it) {
        const int e = (B * it + g) * FPW + sub;
        if (o + (SPI - 1) * step > W) {
            __syncthreads();
            const int64_t base = s_first + (int64_t)SPI * it * step;
            for (int j = threadIdx.x; j < N + W; j += 32 * B) {
                const int64_t gi = base + j;
                stage[j] = (gi >= 0 && gi < a.rxlen) ? __ldg(a.rx + gi) : make_float2(0.f, 0.f);
            }
            o = 0;
            __syncthreads();
        }
        const int64_t s = s_first + (int64_t)e * step;
        const bool ok = e < n_act && s >= 0 && s + N <= a.rxlen;
        const float2* const x = stage + o + (g * FPW + sub) * step + c;
        o += SPI * step;
        float en = 1.f;
        if (c == 0 && ok) en = __ldg(out.energy + i0 + e);
        // ---- pass 1: column c, product folded into the first stage ----
        tc::tmem_ld16_issue(tb + L::COL_CUT, cu[0]);
#pragma unroll
        for (int n1 = 0; n1 < 32; ++n1) v[n1] = x[TG * n1];
#pragma unroll
        for (int m = 0; m < 4; ++m) {
            tc::tmem_ld_wait16(cu[m & 1]);
            if (m + 1 < 4) tc::tmem_ld16_issue(tb + L::COL_CUT + 16 * (m + 1), cu[(m + 1) & 1]);
            folded_stage_chunk(v, m, cu[m & 1]);
        }
        tc::tmem_ld16_issue(tb + L::COL_T2, cu[0]);
        codelet_fft<32, 2>(v);
        __syncwarp();                      // the previous FFTs' row loads are done in every lane
#pragma unroll
        for (int j = 0; j < 32; ++j) buf[bitrev(j, 5) * ROWP + c] = v[j];
        __syncwarp();
        // ---- pass 2: rows k1 = c + TG j ----
#pragma unroll
        for (int j = 0; j < FPW; ++j) {
            const float4* const b3 = reinterpret_cast<const float4*>(buf + (c + TG * j) * ROWP);
#pragma unroll
            for (int d2 = 0; d2 < HP; ++d2) {
                const float4 y = b3[d2];
                v[j * TG + 2 * d2] = make_float2(y.x, y.y);
                v[j * TG + 2 * d2 + 1] = make_float2(y.z, y.w);
            }
        }
#pragma unroll
        for (int m = 0; m < 4; ++m) {
            tc::tmem_ld_wait16(cu[m & 1]);
            if (m + 1 < 4) tc::tmem_ld16_issue(tb + L::COL_T2 + 16 * (m + 1), cu[(m + 1) & 1]);
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int p = 4 * m + r, j = p / HP, i = p % HP;
                folded_butterfly(v[j * TG + i], v[j * TG + i + HP],
                                 make_float2(__uint_as_float(cu[m & 1][4 * r]), __uint_as_float(cu[m & 1][4 * r + 1])),
                                 make_float2(__uint_as_float(cu[m & 1][4 * r + 2]), __uint_as_float(cu[m & 1][4 * r + 3])));
            }
        }
#pragma unroll
        for (int j = 0; j < FPW; ++j) codelet_fft<TG, 2>(v + j * TG);
        // ---- |X|^2, maximum over the sub-group, lowest bin that reaches it; slot j TG + i holds bin c + TG j + 32 bitrev(i) ----
        float nv[32], m8[4];
#pragma unroll
        for (int j = 0; j < 32; ++j) nv[j] = cnorm2(v[j]);
        float tmax = 0.f;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            m8[q] = fmaxf(fmaxf(fmaxf(nv[8 * q], nv[8 * q + 1]), fmaxf(nv[8 * q + 2], nv[8 * q + 3])),
                          fmaxf(fmaxf(nv[8 * q + 4], nv[8 * q + 5]), fmaxf(nv[8 * q + 6], nv[8 * q + 7])));
            tmax = fmaxf(tmax, m8[q]);
        }
        const float mx = __uint_as_float(__reduce_max_sync(submask, __float_as_uint(tmax)));
        unsigned int mybin = 0xffffffffu;
        if (tmax == mx) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (m8[q] == mx) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int sl = 8 * q + i;
                        if (nv[sl] == mx) mybin = min(mybin, (unsigned int)(c + TG * (sl / TG) + 32 * bitrev(sl % TG, LTG)));
                    }
                }
            }
        }
        const unsigned int bin = __reduce_min_sync(submask, mybin);
        if (c == 0 && e < n_act) {
            if (out.qf2) out.qf2[i0 + e] = ok ? mx / refn / en : 0.f;
            if (out.fidx) out.fidx[i0 + e] = (ok && bin != 0xffffffffu) ? bin : 0u;
        }
        if constexpr (CPLX) {
            const float enb = __shfl_sync(0xffffffffu, en, sub * TG);
            if (e < n_act && c == (int)(ok ? (bin & (unsigned int)(TG - 1)) : 0u)) {
                float2 z = make_float2(0.f, 0.f);
                if (ok) {
#pragma unroll
                    for (int sl = 0; sl < 32; ++sl)
                        if ((unsigned int)(c + TG * (sl / TG) + 32 * bitrev(sl % TG, LTG)) == bin) z = v[sl];
                    const float sc = 1.f / (sqrtf(refn) * sqrtf(enb));
                    z = make_float2(z.x * sc, z.y * sc);
                }
                out.cplx[i0 + e] = z;
            }
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (g == 0) tc::tmem_dealloc(tmem0, L::TMEM_COLS);
}

template <int LTG, int B>
int32_t launch_fused_subwarp(const StagedArgs& a, const FusedOut& out, int64_t K, cudaStream_t stream) {
    using L = SubwarpLayout<LTG, B>;
    auto kern = out.cplx ? fused_fft_subwarp_kernel<LTG, B, true> : fused_fft_subwarp_kernel<LTG, B, false>;
    constexpr size_t smem = L::BYTES;
    if ((int)smem > max_smem_optin()) return fail(XC_ERR_UNSUPPORTED, "sub-warp kernel needs %zu B of shared memory", smem);
    static KernelOnce once[2];
    XC_TRY(once[out.cplx ? 1 : 0].configure(kern, max_smem_optin()));
    constexpr int SPI = B * L::FPW;
    const int64_t per = (K + SPI - 1) / SPI;
    int64_t ctas = (int64_t)sm_count();
    if (ctas > per) ctas = per;
    if (ctas < 1) ctas = 1;
    const int64_t chunk = (per + ctas - 1) / ctas;
    if (chunk >= (1ll << 24)) return fail(XC_ERR_UNSUPPORTED, "too many shifts per warp");
    profile_begin(stream);
    kern<<<(unsigned)ctas, 32 * B, smem, stream>>>(a, out, K, (int)chunk);
    profile_end(stream);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

}  // namespace xc
