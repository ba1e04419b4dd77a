// group_warp.cuh -- GroupXcorrFFT at fftlen 1024 over a contiguous shift range: one warp per shift.
//
// The rotated-row fold of group_pp.cuh (ROT) on the warp layout of fused_fft_warp.cuh: rows are 32 cells, every group is cut
// into sub-groups of 32 samples, lane t holds the weight of sample (t - r) mod 32 of each sub-group in tensor memory and
// accumulates row A (t >= r) or the carry into row A + 1; the 32 row sums are exactly the lane's pass-1 inputs
// (cell 32 n1 + t, n1 = row).  Then the two radix-32 passes of the warp kernel, no barrier wider than a warp.
#pragma once
#include "fused_fft_warp.cuh"
#include "group_pp.cuh"

namespace xc {

template <int B>
struct GroupWarpLayout {
    using PL = FftPlan<10, 32, 32, 32, 1>;
    static_assert(B % 4 == 0 && B >= 4 && B <= 32, "warps 0..3 write the constants of the four lane quarters");
    static constexpr size_t OFF_BUF = 0;
    static constexpr size_t OFF_TADDR = sizeof(float2) * (size_t)PL::BUF * B;
    static constexpr size_t OFF_STAGE = OFF_TADDR + 16;                               // G x (L + W) samples + 32 zero cells
    __host__ __device__ static constexpr size_t bytes(int G, int L, int W) { return OFF_STAGE + sizeof(float2) * ((size_t)G * (size_t)(L + W) + 32); }
    static constexpr int COL_W = 0, COL_T2 = 2 * GR_MAX_E, TMEM_COLS = 512;
    static_assert(COL_T2 + 64 <= 512, "tensor memory holds 512 columns");
};

template <int B>
__global__ void __launch_bounds__(32 * B, 1) group_warp_kernel(const GroupPPArgs a, const __grid_constant__ GroupRotPlan rp,
                                                               const FusedOut out, const int64_t K, const int chunk) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    using LY = GroupWarpLayout<B>;
    using PL = typename LY::PL;
    constexpr int ROW = PL::ROW, N = PL::N;
    const int W = a.W;
    const int g = threadIdx.x >> 5, t = threadIdx.x & 31;
    float2* const buf = reinterpret_cast<float2*>(smem_raw + LY::OFF_BUF) + (size_t)g * PL::BUF;
    float2* const stage = reinterpret_cast<float2*>(smem_raw + LY::OFF_STAGE);
    uint32_t* const taddr_slot = reinterpret_cast<uint32_t*>(smem_raw + LY::OFF_TADDR);
    if (g == 0) tc::tmem_alloc(tc::smem_u32(taddr_slot), LY::TMEM_COLS);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem0 = *taddr_slot;
    const uint32_t tb = tmem0 + (((uint32_t)(g & 3) * 32u) << 16);
    if (g < 4) {
        float r[16];
        for (int e0 = 0; e0 < a.NC; e0 += GP_CH) {
#pragma unroll
            for (int i = 0; i < GP_CH; ++i) {
                const float2 w = a.w[(size_t)(e0 + i) * 32 + t];
                r[2 * i] = w.x;
                r[2 * i + 1] = w.y;
            }
            tc::tmem_st16(tb + LY::COL_W + 2 * e0, r);
        }
#pragma unroll 1
        for (int m = 0; m < 4; ++m) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int c = 4 * m + (i >> 1) + 16 * (i & 1);
                const float2 w = unit_root((c * t) % N, N);
                r[2 * i] = w.x;
                r[2 * i + 1] = w.y;
            }
            tc::tmem_st16(tb + LY::COL_T2 + 16 * m, r);
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
    const int64_t s_first = a.shift0 + i0;
    const float refn = *out.refnormsq;
    const int SW = a.L + W;
    const int nstage = a.G * SW;
    for (int i = threadIdx.x; i < 32; i += 32 * B) stage[nstage + i] = make_float2(0.f, 0.f);
    const uint32_t stage_s = tc::smem_u32(stage);
    float2 v[32];
    uint32_t cu[2][16];
    int o = W;

    for (int it = 0; it < chunk; ++it) {
        const int e = B * it + g;
        if (o == W) {
            __syncthreads();
            const int64_t base = s_first + (int64_t)B * it;
            for (int i = threadIdx.x; i < nstage; i += 32 * B) {
                const int gi = i / SW, j = i - gi * SW;
                const int64_t p = base + __ldg(a.starts + gi) + j;
                stage[i] = (p >= 0 && p < a.rxlen) ? __ldg(a.rx + p) : make_float2(0.f, 0.f);
            }
            o = 0;
            __syncthreads();
        }
        const int64_t s = s_first + e;
        const bool ok = e < n_act && s + a.lo >= 0 && s + a.hi <= a.rxlen;
        const uint32_t xs = stage_s + 8u * (uint32_t)(o + g);
        o += B;
        float en = 1.f;
        if (t == 0 && ok) en = __ldg(out.energy + i0 + e);
        // ---- rotated gather (see group_pp.cuh): row rho = pass-1 input n1 = rho of lane t ----
        {
            uint32_t wn[2];
            tc::tmem_ld2_issue(tb + LY::COL_W, wn);
            uint32_t cw = tb + LY::COL_W;
            int ee = 0;
            uint32_t d_n = rp.desc[0], d_nn = rp.desc[1];
            float2 x_n;
            asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(x_n.x), "=f"(x_n.y) : "r"(xs + 8u * ((d_n >> 7) + (uint32_t)((t - (int)(d_n & 127u)) & 31))));
            float2 carry = make_float2(0.f, 0.f);
#pragma unroll
            for (int rho = 0; rho < 32; ++rho) {
                float2 acc = carry;
                carry = make_float2(0.f, 0.f);
                const int n = rp.cnt[rho];
#pragma unroll 1
                for (int i = 0; i < n; ++i) {
                    const int r = (int)(d_n & 127u);
                    tc::tmem_ld_wait2(wn);
                    const float2 y = make_float2(__uint_as_float(wn[0]), __uint_as_float(wn[1]));
                    const float2 x = x_n;
                    d_n = d_nn;
                    ee += 1;
                    d_nn = rp.desc[ee + 1];
                    cw += 2;
                    tc::tmem_ld2_issue(cw, wn);
                    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(x_n.x), "=f"(x_n.y) : "r"(xs + 8u * ((d_n >> 7) + (uint32_t)((t - (int)(d_n & 127u)) & 31))));
                    const float2 p = pcmul(x, y);
                    if (t >= r) acc = padd(acc, p);
                    else carry = padd(carry, p);
                }
                v[rho] = acc;
            }
            v[0] = padd(v[0], carry);
            tc::tmem_ld_wait2(wn);
        }
        // ---- pass 1: column c = t ----
        tc::tmem_ld16_issue(tb + LY::COL_T2, cu[0]);
        codelet_fft<32>(v);
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 32; ++j) buf[bitrev(j, 5) * ROW + t] = v[j];
        __syncwarp();
        // ---- pass 2: row k1 = t, twiddles W_1024^(c k1) folded into the first stage ----
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
            if (m + 1 < 4) tc::tmem_ld16_issue(tb + LY::COL_T2 + 16 * (m + 1), cu[(m + 1) & 1]);
            folded_stage_chunk(v, m, cu[m & 1]);
        }
        codelet_fft<32, 2>(v);
        // ---- |X|^2, warp maximum, lowest bin ----
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
        // a transform twice (2^bs times) the requested fftlen holds the requested bins at the multiples of 2^bs (the DFT of
        // the products does not care where the fold wraps): lane t owns bins t + 32 k2
        const bool lane_on = (t & ((1 << a.bin_shift) - 1)) == 0;
        if (!lane_on) tmax = 0.f;
        const float mx = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(tmax)));
        unsigned int mybin = 0xffffffffu;
        if (lane_on && tmax == mx) {
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
            if (out.fidx) out.fidx[i0 + e] = (ok && bin != 0xffffffffu) ? (bin >> a.bin_shift) : 0u;
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (g == 0) tc::tmem_dealloc(tmem0, LY::TMEM_COLS);
}

}  // namespace xc
