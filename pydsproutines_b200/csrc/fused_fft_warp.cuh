// fused_fft_warp.cuh -- frequency search over a CONTIGUOUS shift range at N = 1024: ONE WARP per FFT.
//
// 1024 = 32 x 32: a warp holds the whole transform (32 points per lane, two radix-32 passes), so the one exchange is warp
// local and the kernel has no group barrier at all: B free-running warps per SM (four per scheduler, each in its own
// phase) over one staged rx window, which is what lets the FMA pipe and the shared-memory pipe overlap.  As in
// fused_fft_pp.cuh the per-lane constants live in TENSOR MEMORY (conj(cutout) of the lane's column and the twiddles of
// its row, 128 columns; written once by warps 0..3, one per lane quarter) and both ride on the first butterfly stage of
// the pass that consumes them (folded_butterfly()).
//   n = 32 n1 + c, k = k1 + 32 k2:  X[k] = sum_c W_32^(c k2) W_1024^(c k1) sum_n1 W_32^(n1 k1) x[32 n1 + c] conj(cutout[32 n1 + c])
//   pass 1: lane c, DFT_32 over n1 (product folded in); exchange cell (k1, c) at k1 ROW + c;
//   pass 2: lane k1, DFT_32 over c (twiddle W_1024^(c k1) folded in); bin k1 + 32 k2 in register bitrev(k2).
// Epilogue: REDUX max over the warp, lowest bin among the lanes that hold it by a second REDUX, lane 0 writes.
#pragma once
#include "fused_fft_pp.cuh"

namespace xc {

template <int B>
struct WarpLayout {
    using PL = FftPlan<10, 32, 32, 32, 1>;
    static_assert(B % 4 == 0 && B >= 4 && B <= 32, "warps 0..3 write the constants of the four lane quarters");
    static constexpr int W = 2048;                                        // extra samples behind the window
    static_assert(W % B == 0, "refills happen at the same iteration in every warp");
    static constexpr size_t OFF_BUF = 0;
    static constexpr size_t OFF_STAGE = sizeof(float2) * (size_t)PL::BUF * B;
    static constexpr size_t OFF_TADDR = OFF_STAGE + sizeof(float2) * (size_t)(PL::N + W);
    static constexpr size_t BYTES = OFF_TADDR + 16;
    static constexpr int COL_CUT = 0, COL_T2 = 64, TMEM_COLS = 128;
};

// CPLX: also writes the complex QF X[bin] / (||cutout|| ||window||) of the absResult = False variants (xcorrRoutines.py:533-548)
template <int B, bool CPLX>
__global__ void __launch_bounds__(32 * B, 1) fused_fft_warp_kernel(const StagedArgs a, const FusedOut out, const int64_t K,
                                                                   const int chunk) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    using L = WarpLayout<B>;
    using PL = typename L::PL;
    constexpr int ROW = PL::ROW, W = L::W, N = PL::N;
    const int g = threadIdx.x >> 5;       // warp = FFT group
    const int t = threadIdx.x & 31;
    float2* const buf = reinterpret_cast<float2*>(smem_raw + L::OFF_BUF) + (size_t)g * PL::BUF;
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
                const float2 cv = a.cutout[32 * n1 + t];
                r[2 * i] = cv.x;
                r[2 * i + 1] = -cv.y;
            }
            tc::tmem_st16(tb + L::COL_CUT + 16 * m, r);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int c = 4 * m + (i >> 1) + 16 * (i & 1);
                const float2 w = unit_root((c * t) % N, N);
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

    // this CTA owns shift-list entries [i0, i0 + B * chunk); warp g takes i0 + B * it + g
    const int64_t i0 = (int64_t)blockIdx.x * B * (int64_t)chunk;
    int64_t n_left = K - i0;
    if (n_left < 0) n_left = 0;
    if (n_left > B * (int64_t)chunk) n_left = B * (int64_t)chunk;
    const int n_act = (int)n_left;
    const int step = a.step;
    const int64_t s_first = a.shift0 + i0 * step;
    const float refn = *out.refnormsq;
    float2 v[32];
    uint32_t cu[2][16];
    int o = W + 1;

    for (int it = 0; it < chunk; ++it) {
        const int e = B * it + g;
        if (o + (B - 1) * step > W) {
            __syncthreads();               // no warp still reads the old window
            const int64_t base = s_first + (int64_t)B * it * step;
            for (int j = threadIdx.x; j < N + W; j += 32 * B) {
                const int64_t gi = base + j;
                stage[j] = (gi >= 0 && gi < a.rxlen) ? __ldg(a.rx + gi) : make_float2(0.f, 0.f);
            }
            o = 0;
            __syncthreads();
        }
        const int64_t s = s_first + (int64_t)e * step;
        const bool ok = e < n_act && s >= 0 && s + N <= a.rxlen;
        const float2* const x = stage + o + g * step + t;
        o += B * step;
        float en = 1.f;
        if (t == 0 && ok) en = __ldg(out.energy + i0 + e);     // in flight under the FFT
        // ---- pass 1: column c = t, product folded into the first stage ----
        tc::tmem_ld16_issue(tb + L::COL_CUT, cu[0]);
#pragma unroll
        for (int n1 = 0; n1 < 32; ++n1) v[n1] = x[32 * n1];
#pragma unroll
        for (int m = 0; m < 4; ++m) {
            tc::tmem_ld_wait16(cu[m & 1]);
            if (m + 1 < 4) tc::tmem_ld16_issue(tb + L::COL_CUT + 16 * (m + 1), cu[(m + 1) & 1]);
            folded_stage_chunk(v, m, cu[m & 1]);
        }
        tc::tmem_ld16_issue(tb + L::COL_T2, cu[0]);
        codelet_fft<32, 2>(v);
        __syncwarp();                      // the previous FFT's row loads are done in every lane
#pragma unroll
        for (int j = 0; j < 32; ++j) buf[bitrev(j, 5) * ROW + t] = v[j];
        __syncwarp();
        // ---- pass 2: row k1 = t ----
        const float4* const b3 = reinterpret_cast<const float4*>(buf + t * ROW);
#pragma unroll
        for (int d2 = 0; d2 < 16; ++d2) {
            const float4 y = b3[d2];
            v[2 * d2] = make_float2(y.x, y.y);
            v[2 * d2 + 1] = make_float2(y.z, y.w);
        }
#pragma unroll
        for (int m = 0; m < 4; ++m) {
            tc::tmem_ld_wait16(cu[m & 1]);
            if (m + 1 < 4) tc::tmem_ld16_issue(tb + L::COL_T2 + 16 * (m + 1), cu[(m + 1) & 1]);
            folded_stage_chunk(v, m, cu[m & 1]);
        }
        codelet_fft<32, 2>(v);
        // ---- |X|^2, warp maximum, lowest bin that reaches it ----
        float nv[32], m8[4];
#pragma unroll
        for (int j = 0; j < 32; ++j) nv[j] = cnorm2(v[j]);
        float tmax = 0.f;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            m8[c] = fmaxf(fmaxf(fmaxf(nv[8 * c], nv[8 * c + 1]), fmaxf(nv[8 * c + 2], nv[8 * c + 3])),
                          fmaxf(fmaxf(nv[8 * c + 4], nv[8 * c + 5]), fmaxf(nv[8 * c + 6], nv[8 * c + 7])));
            tmax = fmaxf(tmax, m8[c]);
        }
        const float mx = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(tmax)));
        unsigned int mybin = 0xffffffffu;
        if (tmax == mx) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                if (m8[c] == mx) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int j = 8 * c + i;
                        if (nv[j] == mx) mybin = min(mybin, (unsigned int)(t + 32 * bitrev(j, 5)));
                    }
                }
            }
        }
        const unsigned int bin = __reduce_min_sync(0xffffffffu, mybin);
        if (t == 0 && e < n_act) {
            if (out.qf2) out.qf2[i0 + e] = ok ? mx / refn / en : 0.f;
            if (out.fidx) out.fidx[i0 + e] = (ok && bin != 0xffffffffu) ? bin : 0u;
        }
        if constexpr (CPLX) {
            const float enb = __shfl_sync(0xffffffffu, en, 0);
            if (e < n_act && t == (int)(ok ? (bin & 31u) : 0u)) {        // the lane that holds the winning bin
                float2 z = make_float2(0.f, 0.f);
                if (ok) {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if ((unsigned int)(t + 32 * bitrev(j, 5)) == bin) z = v[j];
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

template <int B>
int32_t launch_fused_warp(const StagedArgs& a, const FusedOut& out, int64_t K, cudaStream_t stream) {
    auto kern = out.cplx ? fused_fft_warp_kernel<B, true> : fused_fft_warp_kernel<B, false>;
    constexpr size_t smem = WarpLayout<B>::BYTES;
    if ((int)smem > max_smem_optin()) return fail(XC_ERR_UNSUPPORTED, "warp-per-FFT kernel needs %zu B of shared memory", smem);
    static KernelOnce once[2];
    XC_TRY(once[out.cplx ? 1 : 0].configure(kern, max_smem_optin()));
    const int64_t per = (K + B - 1) / B;
    int64_t ctas = (int64_t)sm_count();
    if (ctas > per) ctas = per;
    if (ctas < 1) ctas = 1;
    const int64_t chunk = (per + ctas - 1) / ctas;
    if (chunk >= (1ll << 26)) return fail(XC_ERR_UNSUPPORTED, "too many shifts per warp");
    profile_begin(stream);
    kern<<<(unsigned)ctas, 32 * B, smem, stream>>>(a, out, K, (int)chunk);
    profile_end(stream);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

}  // namespace xc
