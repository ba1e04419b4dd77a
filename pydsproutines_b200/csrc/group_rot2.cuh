// group_rot2.cuh -- GroupXcorrFFT at fftlen 8192 over a contiguous shift range: the rotated-row fold of group_pp.cuh (ROT) in
// front of the register-input block transform of fused_blue_pp.cuh (two groups of 128 threads, 64 points per thread, 64 rows of
// 128 cells).  Sub-groups of 128 samples sorted by row; thread t holds one weight per sub-group in tensor memory behind the two
// folded twiddle sets; software-pipelined gather (descriptor two entries ahead, sample and weight one ahead).
#pragma once
#include "fused_blue_pp.cuh"
#include "group_pp.cuh"

namespace xc {

constexpr int GR2_MAX_E = 96;          // sub-groups (2 tensor-memory columns each behind the 320 twiddle columns)
struct GroupRot2Plan {
    unsigned char cnt[64];             // sub-groups whose first cell lies in row rho (sorted by row)
    uint32_t desc[GR2_MAX_E];          // r | (first stage cell of the sub-group << 7)
};

template <class PL, int B>
struct GroupRot2Layout {
    using L = PPLayout<PL, B, true>;
    static_assert(PL::TG == 128 && L::COPIES == 1, "rows of 128 cells, one warp of a group per lane quarter");
    static constexpr int ROWS = PL::N / PL::TG;
    static_assert(ROWS == PL::V && ROWS <= 64, "one register slot per row");
    static constexpr int COL_W = L::CW;                                  // weights behind conj(cutout) (unused), T1, T2
    static_assert(COL_W + 2 * GR2_MAX_E <= 512, "tensor memory holds 512 columns");
    static constexpr size_t OFF_BUF = 0;
    static constexpr size_t OFF_GMAX = sizeof(float2) * (size_t)PL::BUF * B;
    static constexpr size_t OFF_BIN = OFF_GMAX + sizeof(unsigned int) * 2 * B;
    static constexpr size_t OFF_TADDR = OFF_BIN + sizeof(unsigned int) * B;
    static constexpr size_t OFF_STAGE = (OFF_TADDR + 16 + 15) / 16 * 16;              // G x (L + W) samples + TG zero cells
    __host__ __device__ static constexpr size_t bytes(int G, int L_, int W) {
        return OFF_STAGE + sizeof(float2) * ((size_t)G * (size_t)(L_ + W) + PL::TG);
    }
};

template <class PL, int B>
__global__ void __launch_bounds__(PL::TG* B, 1) group_rot2_kernel(const GroupPPArgs a, const __grid_constant__ GroupRot2Plan rp,
                                                                  const FusedOut out, const int64_t K, const int chunk) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    using L = PPLayout<PL, B, true>;
    using LY = GroupRot2Layout<PL, B>;
    constexpr int TG = PL::TG, V = PL::V, R1 = PL::R1;
    const int W = a.W;
    const int g = threadIdx.x / TG, t = threadIdx.x % TG;
    const int warp_cta = threadIdx.x >> 5;
    float2* const buf = reinterpret_cast<float2*>(smem_raw + LY::OFF_BUF) + (size_t)g * PL::BUF;
    float2* const stage = reinterpret_cast<float2*>(smem_raw + LY::OFF_STAGE);
    unsigned int* const gmax = reinterpret_cast<unsigned int*>(smem_raw + LY::OFF_GMAX) + 2 * g;
    unsigned int* const binmin = reinterpret_cast<unsigned int*>(smem_raw + LY::OFF_BIN) + g;
    uint32_t* const taddr_slot = reinterpret_cast<uint32_t*>(smem_raw + LY::OFF_TADDR);
    if (warp_cta == 0) tc::tmem_alloc(tc::smem_u32(taddr_slot), 512);
    if (t == 0) {
        *binmin = 0xffffffffu;
        gmax[0] = 0u;
        gmax[1] = 0u;
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem0 = *taddr_slot;
    const uint32_t tb = tmem0 + (((uint32_t)(warp_cta & 3) * 32u) << 16);
    if (g == 0) {
        float r[16];
        for (int e0 = 0; e0 < a.NC; e0 += GP_CH) {
#pragma unroll
            for (int i = 0; i < GP_CH; ++i) {
                const float2 w = a.w[(size_t)(e0 + i) * TG + t];
                r[2 * i] = w.x;
                r[2 * i + 1] = w.y;
            }
            tc::tmem_st16(tb + LY::COL_W + 2 * e0, r);
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
    const GroupBar<TG, B> bar{1 + g};
    const int SW = a.L + W;
    const int nstage = a.G * SW;
    for (int i = threadIdx.x; i < TG; i += TG * B) stage[nstage + i] = make_float2(0.f, 0.f);
    const uint32_t stage_s = tc::smem_u32(stage);
    float2 v[V];
    float pend = 0.f;
    int o = W;

    for (int it = 0; it < chunk; ++it) {
        const int e = B * it + g;
        if (o == W) {
            __syncthreads();
            const int64_t base = s_first + (int64_t)B * it;
            for (int i = threadIdx.x; i < nstage; i += TG * B) {
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
        // ---- rotated gather (group_pp.cuh, ROT): row rho = cells [128 rho, 128 rho + 128) = register slot (rho % QPR) R1 + rho / QPR ----
        {
            uint32_t wn[2];
            tc::tmem_ld2_issue(tb + LY::COL_W, wn);
            uint32_t cw = tb + LY::COL_W;
            int ee = 0;
            uint32_t d_n = rp.desc[0], d_nn = rp.desc[1];
            float2 x_n;
            asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(x_n.x), "=f"(x_n.y) : "r"(xs + 8u * ((d_n >> 7) + (uint32_t)((t - (int)(d_n & 127u)) & (TG - 1)))));
            float2 carry = make_float2(0.f, 0.f);
#pragma unroll
            for (int rho = 0; rho < LY::ROWS; ++rho) {
                constexpr int QPR = PL::M1 / TG;
                const int c = (rho % QPR) * R1 + rho / QPR;
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
                    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(x_n.x), "=f"(x_n.y) : "r"(xs + 8u * ((d_n >> 7) + (uint32_t)((t - (int)(d_n & 127u)) & (TG - 1)))));
                    const float2 p = pcmul(x, y);
                    if (t >= r) acc = padd(acc, p);
                    else carry = padd(carry, p);
                }
                v[c] = acc;
            }
            v[0] = padd(v[0], carry);
            tc::tmem_ld_wait2(wn);
        }
        // ---- transform from registers (one group barrier inside; it also orders the previous shift's bin atomics) ----
        blue_block_fft_regs<PL, B>(v, tb, buf, t, bar);
        if (t == 0 && it > 0) {
            const int p = e - B;
            if (p < n_act) {
                const unsigned int b = *binmin;
                if (out.qf2) out.qf2[i0 + p] = pend;
                if (out.fidx) out.fidx[i0 + p] = (b == 0xffffffffu) ? 0u : b;
            }
            *binmin = 0xffffffffu;
        }
        float nv[V], m8[V / 8];
#pragma unroll
        for (int j = 0; j < V; ++j) nv[j] = cnorm2(v[j]);
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
        if (ok && tmax == m) {
            unsigned int mybin = 0xffffffffu;
#pragma unroll
            for (int c = 0; c < V / 8; ++c) {
                if (m8[c] == m) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int j = 8 * c + i;
                        if (nv[j] == m) mybin = min(mybin, (unsigned int)output_bin<PL>(t, j / PL::RL, j % PL::RL));
                    }
                }
            }
            atomicMin(binmin, mybin);
        }
        if (t == 0) pend = ok ? m / *out.refnormsq / out.energy[i0 + e] : 0.f;
    }
    bar();
    if (t == 0 && chunk > 0) {
        const int p = B * (chunk - 1) + g;
        if (p < n_act) {
            const unsigned int b = *binmin;
            if (out.qf2) out.qf2[i0 + p] = pend;
            if (out.fidx) out.fidx[i0 + p] = (b == 0xffffffffu) ? 0u : b;
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp_cta == 0) tc::tmem_dealloc(tmem0, 512);
}

}  // namespace xc
