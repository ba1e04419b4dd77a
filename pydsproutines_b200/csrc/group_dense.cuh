// group_dense.cuh -- GroupXcorrFFT at SHORT transform lengths (fftlen 64 / 128 / 256, e.g. the reference's default
// fftlen = group length, xcorrRoutines.py:1065-1068) over a contiguous shift range.
//
// There the F-point FFT is small change and the fold is the work: u_s[m] = sum_g w_g[m] x[s + s_g + l_g(m)] is G F complex
// multiply-adds per shift (w_g[m] = conj-applied sample l_g(m) = (m - s_g) mod F of group g, zero beyond the group length).
// A CTA takes a tile of ST consecutive shifts:
//   * fold phase: a thread owns two adjacent cells and eight consecutive shifts (16 accumulators); per group it reads its two
//     weights from tensor memory (one tcgen05.ld.x4) and a ten-sample window of the group's staged rx window with five
//     aligned LDS.128 (adjacent cells and shifts slide over the same samples: 0.31 loads per multiply-add), 32 packed FMAs;
//     a block whose two cells straddle the group's wrap-around point reads its two windows separately (rare);
//   * the tile u[ST][F] goes to shared memory, and half of the warps transform it with the sub-warp FFT of
//     fused_fft_subwarp.cuh (F / 32 lanes per shift, in place), |X|^2, maximum and lowest arg max per shift.
#pragma once
#include "fused_fft_subwarp.cuh"
#include "group_pp.cuh"

namespace xc {

#ifndef XC_GD_UNROLL
#define XC_GD_UNROLL 4
#endif
constexpr int GD_UNROLL = XC_GD_UNROLL;   // groups per trip of the fold loop
constexpr int GD_MAX_G = 112;          // groups (4 tensor-memory columns each beside the 64 twiddle columns)

namespace tc {
__device__ __forceinline__ void tmem_ld4_issue(uint32_t taddr, uint32_t* u) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]) : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait4(uint32_t* u) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(u[0]), "+r"(u[1]), "+r"(u[2]), "+r"(u[3]));
}
}  // namespace tc

template <int LF>
struct GroupDenseLayout {
    static_assert(LF >= 6 && LF <= 8, "fftlen 64, 128 or 256");
    static constexpr int F = 1 << LF, CB = F / 2, DUP = 128 / CB, ST = 32 * DUP, NSB = ST / 8;
    static constexpr int THREADS = 512;
    static_assert(CB * NSB == THREADS, "a thread per (cell pair, block of eight shifts)");
    static constexpr int TG = F / 32, LTG = LF - 5, FPW = 32 / TG;
    static constexpr int UP = F + 4;                                   // row pitch of the tile u (cells)
    static constexpr int COL_T2 = 0, COL_W = 64, TMEM_COLS = 512;
    static constexpr size_t OFF_U = 0;
    static constexpr size_t OFF_TADDR = sizeof(float2) * (size_t)ST * UP;
    static constexpr size_t OFF_DESC = OFF_TADDR + 16;                 // [G4][CB] uint32
    __host__ __device__ static constexpr size_t off_starts(int G4) { return (OFF_DESC + sizeof(uint32_t) * (size_t)G4 * CB + 15) / 16 * 16; }   // int64 [G4]
    __host__ __device__ static constexpr size_t off_stage(int G4) { return off_starts(G4) + sizeof(int64_t) * (size_t)G4; }
    __host__ __device__ static constexpr int sw(int L) { return (L + ST + 16 + 1) / 2 * 2; }    // staged samples per group
    __host__ __device__ static constexpr size_t bytes(int G4, int L) { return off_stage(G4) + sizeof(float2) * (size_t)G4 * sw(L); }
};

struct GroupDenseArgs {
    const float2* rx;
    int64_t rxlen;
    int64_t shift0;                // the shift of entry i is shift0 + i
    int64_t lo, hi;                // valid shifts satisfy s + lo >= 0 and s + hi <= rxlen
    const int64_t* starts;         // [G] device
    const float4* w;               // [G4][CB]: weights of cells 2 cb, 2 cb + 1 (conj applied), device
    const uint32_t* desc;          // [G4][CB]: l0 | l1 << 14 | mode << 28 (0 = one window at l0, even; 1 = odd; 2 = two windows)
    int G, G4, L;
};

__device__ __forceinline__ float2 gd_mac(float2 acc, float2 x, float wr, float wi) {
    return pfma(make_float2(x.y, x.x), make_float2(-wi, wi), pfma(x, make_float2(wr, wr), acc));
}

// stage[g SW + j] = rx[s0 + s_g + j], j < L + ST (zero outside the record, behind the window and for the padding groups);
// sst = the group starts in shared memory; eight independent loads in flight per thread
template <int LF>
__device__ __forceinline__ void gd_stage(const GroupDenseArgs& a, float2* stage, const int64_t* sst, const int SW, const int nstage,
                                         const int64_t s0, const int tid, const int nthreads) {
    constexpr int ST = GroupDenseLayout<LF>::ST;
    for (int i0 = tid; i0 < nstage; i0 += 8 * nthreads) {
        float2 val[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int i = i0 + k * nthreads;
            val[k] = make_float2(0.f, 0.f);
            if (i < nstage) {
                const int gi = i / SW, j = i - gi * SW;
                if (gi < a.G && j < a.L + ST) {
                    const int64_t p = s0 + sst[gi] + j;
                    if (p >= 0 && p < a.rxlen) val[k] = __ldg(a.rx + p);
                }
            }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int i = i0 + k * nthreads;
            if (i < nstage) stage[i] = val[k];
        }
    }
}

template <int LF>
__global__ void __launch_bounds__(512, 1) group_dense_kernel(const GroupDenseArgs a, const FusedOut out, const int64_t K) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    using LY = GroupDenseLayout<LF>;
    constexpr int F = LY::F, CB = LY::CB, ST = LY::ST, TG = LY::TG, LTG = LY::LTG, FPW = LY::FPW, UP = LY::UP, HP = TG / 2;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float2* const u = reinterpret_cast<float2*>(smem_raw + LY::OFF_U);
    uint32_t* const taddr_slot = reinterpret_cast<uint32_t*>(smem_raw + LY::OFF_TADDR);
    uint32_t* const desc = reinterpret_cast<uint32_t*>(smem_raw + LY::OFF_DESC);
    float2* const stage = reinterpret_cast<float2*>(smem_raw + LY::off_stage(a.G4));
    const int SW = LY::sw(a.L);
    if (warp == 0) tc::tmem_alloc(tc::smem_u32(taddr_slot), LY::TMEM_COLS);
    int64_t* const sst = reinterpret_cast<int64_t*>(smem_raw + LY::off_starts(a.G4));
    for (int i = threadIdx.x; i < a.G4 * CB; i += 512) desc[i] = a.desc[i];
    for (int i = threadIdx.x; i < a.G4; i += 512) sst[i] = i < a.G ? a.starts[i] : 0;
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem0 = *taddr_slot;
    const uint32_t tb = tmem0 + (((uint32_t)(warp & 3) * 32u) << 16);
    // fold phase: tensor-memory lane Lq = 32 (warp % 4) + lane holds the weights of cell pair cb = Lq % CB
    const int Lq = 32 * (warp & 3) + lane;
    const int cb = Lq % CB;
    const int sb = (warp >> 2) * LY::DUP + Lq / CB;          // block of eight shifts inside the tile
    // transform phase: lane c of sub-group `sub` of warp `warp` (warps < ST / FPW)
    const int sub = lane >> LTG, c = lane & (TG - 1);
    if (warp < 4) {
        float r[16];
#pragma unroll 1
        for (int m = 0; m < 4; ++m) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int p = 4 * m + (i >> 1), j = p / HP, cc = p % HP + HP * (i & 1);
                const float2 w = unit_root((cc * (c + TG * j)) % F, F);
                r[2 * i] = w.x;
                r[2 * i + 1] = w.y;
            }
            tc::tmem_st16(tb + LY::COL_T2 + 16 * m, r);
        }
        for (int g0 = 0; g0 < a.G4; g0 += 4) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float4 w = a.w[(size_t)(g0 + i) * CB + cb];
                r[4 * i] = w.x;
                r[4 * i + 1] = w.y;
                r[4 * i + 2] = w.z;
                r[4 * i + 3] = w.w;
            }
            tc::tmem_st16(tb + LY::COL_W + 4 * g0, r);
        }
        tc::tmem_st_wait();
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();

    const float refn = *out.refnormsq;
    const unsigned int submask = ((1u << TG) - 1u) << (sub * TG);
    const int64_t ntiles = (K + ST - 1) / ST;
    const int nstage = a.G4 * SW;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t i0 = tile * ST;                          // first shift-list entry of the tile
        const int64_t s0 = a.shift0 + i0;
        // (staging the next tile by the eight warps that do not transform measured slower: it outlasts the transform)
        gd_stage<LF>(a, stage, sst, SW, nstage, s0, threadIdx.x, 512);
        __syncthreads();
        // ---- fold: two cells x eight shifts per thread ----
        float2 acc0[8], acc1[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc0[j] = acc1[j] = make_float2(0.f, 0.f);
        uint32_t wn[4];
        tc::tmem_ld4_issue(tb + LY::COL_W, wn);
#pragma unroll GD_UNROLL
        for (int g = 0; g < a.G4; ++g) {
            const uint32_t d = desc[g * CB + cb];
            const int l0 = (int)(d & 0x3fffu), l1 = (int)((d >> 14) & 0x3fffu), mode = (int)(d >> 28);
            const float2* const xg = stage + g * SW + 8 * sb;
            tc::tmem_ld_wait4(wn);
            const float w0r = __uint_as_float(wn[0]), w0i = __uint_as_float(wn[1]);
            const float w1r = __uint_as_float(wn[2]), w1i = __uint_as_float(wn[3]);
            tc::tmem_ld4_issue(tb + LY::COL_W + 4 * (g + 1), wn);       // one group of padding follows the last
            if (mode == 0) {
                const float4* const x4 = reinterpret_cast<const float4*>(xg + l0);
                float2 x[10];
#pragma unroll
                for (int k = 0; k < 5; ++k) {
                    const float4 y = x4[k];
                    x[2 * k] = make_float2(y.x, y.y);
                    x[2 * k + 1] = make_float2(y.z, y.w);
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    acc0[j] = gd_mac(acc0[j], x[j], w0r, w0i);
                    acc1[j] = gd_mac(acc1[j], x[j + 1], w1r, w1i);
                }
            } else if (mode == 1) {
                const float4* const x4 = reinterpret_cast<const float4*>(xg + l0 - 1);
                float2 x[10];
#pragma unroll
                for (int k = 0; k < 5; ++k) {
                    const float4 y = x4[k];
                    x[2 * k] = make_float2(y.x, y.y);
                    x[2 * k + 1] = make_float2(y.z, y.w);
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    acc0[j] = gd_mac(acc0[j], x[j + 1], w0r, w0i);
                    acc1[j] = gd_mac(acc1[j], x[j + 2], w1r, w1i);
                }
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    acc0[j] = gd_mac(acc0[j], xg[l0 + j], w0r, w0i);
                    acc1[j] = gd_mac(acc1[j], xg[l1 + j], w1r, w1i);
                }
            }
        }
        tc::tmem_ld_wait4(wn);
        // ---- the tile u[shift][cell] ----
#pragma unroll
        for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(u + (8 * sb + j) * UP + 2 * cb) = make_float4(acc0[j].x, acc0[j].y, acc1[j].x, acc1[j].y);
        __syncthreads();
        // ---- transform: F / 32 lanes per shift, in place in the row of the shift ----
        if (warp < ST / FPW) {
            const int sh = warp * FPW + sub;
            float2* const row = u + sh * UP;
            const int64_t e = i0 + sh;
            const int64_t s = s0 + sh;
            const bool ok = e < K && s + a.lo >= 0 && s + a.hi <= a.rxlen;
            float en = 1.f;
            if (c == 0 && ok) en = __ldg(out.energy + e);
            float2 v[32];
            uint32_t cu[2][16];
            tc::tmem_ld16_issue(tb + LY::COL_T2, cu[0]);
#pragma unroll
            for (int n1 = 0; n1 < 32; ++n1) v[n1] = row[TG * n1 + c];
            codelet_fft<32>(v);
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 32; ++j) row[TG * bitrev(j, 5) + c] = v[j];
            __syncwarp();
#pragma unroll
            for (int j = 0; j < FPW; ++j) {
                const float4* const b3 = reinterpret_cast<const float4*>(row + (c + TG * j) * TG);
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
                if (m + 1 < 4) tc::tmem_ld16_issue(tb + LY::COL_T2 + 16 * (m + 1), cu[(m + 1) & 1]);
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
            float tmax = 0.f;
            float nv[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                nv[j] = cnorm2(v[j]);
                tmax = fmaxf(tmax, nv[j]);
            }
            const float mx = __uint_as_float(__reduce_max_sync(submask, __float_as_uint(tmax)));
            unsigned int mybin = 0xffffffffu;
            if (tmax == mx) {
#pragma unroll
                for (int sl = 0; sl < 32; ++sl)
                    if (nv[sl] == mx) mybin = min(mybin, (unsigned int)(c + TG * (sl / TG) + 32 * bitrev(sl % TG, LTG)));
            }
            const unsigned int bin = __reduce_min_sync(submask, mybin);
            if (c == 0 && e < K) {
                if (out.qf2) out.qf2[e] = ok ? mx / refn / en : 0.f;
                if (out.fidx) out.fidx[e] = (ok && bin != 0xffffffffu) ? bin : 0u;
            }
        }
        __syncthreads();     // u and the stage are free for the next tile
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem0, LY::TMEM_COLS);
}

}  // namespace xc
