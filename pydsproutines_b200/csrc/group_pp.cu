// group_pp.cu -- plan builder and launcher of the ping-pong GroupXcorrFFT kernel (group_pp.cuh).
#include <algorithm>
#include <vector>

#include "group_pp.cuh"
#include "group_warp.cuh"
#include "group_dense.cuh"
#include "group_rot2.cuh"
#include "internal.cuh"

namespace xc {

using GPPL = FftPlan<12, 128, 4, 32, 32>;

#ifndef XC_GP_B
#define XC_GP_B 3
#endif
constexpr int GP_B = XC_GP_B;          // thread groups per SM

void group_pp_release(GroupPPPlan* p) {
    if (p->d_w) cudaFree(p->d_w);
    if (p->d_off) cudaFree(p->d_off);
    memset(p, 0, sizeof(*p));
}

using GPPL2K = FftPlan<11, 64, 2, 32, 32>;     // fftlen 2048: six groups of 64 threads, rows of 64 cells (ROT variant only)
constexpr int GP_B2K = 6;

// Rotated-row plan (group_pp.cuh, ROT): every group is cut into sub-groups of TG samples, sorted by the row their first
// cell lies in; thread t holds the weight of sample (t - r) mod TG of each sub-group.
using GPPL1K = FftPlan<10, 32, 32, 32, 1>;     // fftlen 1024: sixteen warps, one shift per warp (group_warp.cuh)
constexpr int GP_B1K = 16;

template <class PL, int B, class LY>
static int rot_window_slack(int G, int L) {
    // the stage keeps TG zero cells behind the last window
    const int64_t room = (int64_t)max_smem_optin() - (int64_t)LY::OFF_STAGE - (int64_t)sizeof(float2) * PL::TG;
    int W = ((int)(room / (int64_t)(sizeof(float2) * (size_t)G)) - L) / B * B;
    if (W > 252 / B * B) W = 252 / B * B;
    return W;
}
using GPPL8K = FftPlan<13, 128, 8, 32, 32>;    // fftlen 8192: two groups of 128 threads, 64 rows of 128 cells (group_rot2.cuh)
constexpr int GP_B8K = 2;

template <class PL, int B, class LY = GroupPPLayout<PL, B, true>, int MAXE = GR_MAX_E>
static int32_t group_rot_build(const float2* y, const int64_t* starts, int G, int L, GroupPPPlan* p) {
    constexpr int TG = PL::TG, F = PL::N;
    static_assert(GR_MAX_E == sizeof(p->rdesc) / sizeof(p->rdesc[0]) && MAXE <= GR_MAX_E, "descriptor table size");
    static_assert(F / TG <= (int)sizeof(p->rcnt), "row counters");
    const int per = (L + TG - 1) / TG;
    const int E = G * per;
    if (E + 2 > MAXE) return XC_OK;
    const int W = rot_window_slack<PL, B, LY>(G, L);
    if (W < 4 * B) return XC_OK;
    const int SW = L + W;
    struct Sub { int row, r, g, i; };
    std::vector<Sub> subs;
    for (int g = 0; g < G; ++g)
        for (int i = 0; i < per; ++i) {
            const int64_t a = (((starts[g] + (int64_t)TG * i) % F) + F) % F;
            subs.push_back(Sub{(int)(a / TG), (int)(a % TG), g, i});
        }
    std::stable_sort(subs.begin(), subs.end(), [](const Sub& x, const Sub& z) { return x.row < z.row; });
    const int NC = (E + 2 + GP_CH - 1) / GP_CH * GP_CH;   // padding entries (the gather prefetches ahead), whole st16 chunks
    std::vector<float2> hw((size_t)NC * TG, make_float2(0.f, 0.f));
    memset(p->rcnt, 0, sizeof(p->rcnt));
    for (int e = 0; e < E; ++e) {
        const Sub& sb = subs[(size_t)e];
        if (p->rcnt[sb.row] == 255) return XC_OK;
        ++p->rcnt[sb.row];
        p->rdesc[e] = (uint32_t)sb.r | ((uint32_t)(sb.g * SW + TG * sb.i) << 7);
        for (int t = 0; t < TG; ++t) {
            const int n = TG * sb.i + ((t - sb.r) & (TG - 1));
            if (n < L) hw[(size_t)e * TG + t] = y[(size_t)sb.g * L + n];
        }
    }
    for (int e = E; e < GR_MAX_E; ++e) p->rdesc[e] = 0;
    XC_CUDA_TRY(cudaMalloc(&p->d_w, sizeof(float2) * hw.size()));
    XC_CUDA_TRY(cudaMemcpy(p->d_w, hw.data(), sizeof(float2) * hw.size(), cudaMemcpyHostToDevice));
    p->G = G;
    p->L = L;
    p->W = W;
    p->NC = NC;
    p->nq_live = PL::V / PL::R1;
    p->rot = 1;
    p->B = B;
    p->F = F;
    p->usable = 1;
    return XC_OK;
}

// Dense-fold plan for short transforms (group_dense.cuh): per (group, cell pair) the two weights and the sample offsets
template <int LF>
static int32_t group_dense_build(const float2* y, const int64_t* starts, int G, int L, GroupPPPlan* p) {
    using LY = GroupDenseLayout<LF>;
    constexpr int F = LY::F, CB = LY::CB;
    const int G4 = (G + 3) / 4 * 4;
    if (G4 + 1 > GD_MAX_G - 1 || L > F) return XC_OK;
    if (LY::bytes(G4, L) > (size_t)max_smem_optin()) return XC_OK;
    std::vector<float4> hw((size_t)G4 * CB, make_float4(0.f, 0.f, 0.f, 0.f));
    std::vector<uint32_t> hd((size_t)G4 * CB, 0u | (1u << 14));
    for (int g = 0; g < G; ++g) {
        const int r = (int)(((starts[g] % F) + F) % F);
        for (int cb = 0; cb < CB; ++cb) {
            int l0 = ((2 * cb - r) % F + F) % F, l1 = ((2 * cb + 1 - r) % F + F) % F;
            const bool v0 = l0 < L, v1 = l1 < L;
            float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
            if (v0) { w.x = y[(size_t)g * L + l0].x; w.y = y[(size_t)g * L + l0].y; }
            if (v1) { w.z = y[(size_t)g * L + l1].x; w.w = y[(size_t)g * L + l1].y; }
            uint32_t mode;
            if (!v0 && !v1) {
                l0 = r & 1;                    // any window of the right parity
                l1 = l0 + 1;
                mode = (uint32_t)(l0 & 1);
            } else if (v0 && (l1 == l0 + 1)) {
                mode = (uint32_t)(l0 & 1);     // one sliding window (cell 1 may lie behind the group: weight zero)
            } else {
                if (!v0) l0 = 0;               // cell 1 opens the group (l1 = 0) or cell 0 closes it at the wrap-around point
                if (!v1) l1 = 0;
                mode = 2u;
            }
            hw[(size_t)g * CB + cb] = w;
            hd[(size_t)g * CB + cb] = (uint32_t)l0 | ((uint32_t)l1 << 14) | (mode << 28);
        }
    }
    XC_CUDA_TRY(cudaMalloc(&p->d_w, sizeof(float4) * hw.size()));
    XC_CUDA_TRY(cudaMalloc(&p->d_off, sizeof(uint32_t) * hd.size()));
    XC_CUDA_TRY(cudaMemcpy(p->d_w, hw.data(), sizeof(float4) * hw.size(), cudaMemcpyHostToDevice));
    XC_CUDA_TRY(cudaMemcpy(p->d_off, hd.data(), sizeof(uint32_t) * hd.size(), cudaMemcpyHostToDevice));
    p->G = G;
    p->G4 = G4;
    p->L = L;
    p->F = F;
    p->dense = 1;
    p->usable = 1;
    return XC_OK;
}

template <int LF>
static int32_t group_dense_launch(const GroupPPPlan& p, const int64_t* d_starts, const float2* rx, int64_t rxlen, int64_t shift0,
                                  int64_t K, int64_t lo, int64_t hi, const FusedOut& out, cudaStream_t stream) {
    using LY = GroupDenseLayout<LF>;
    auto kern = group_dense_kernel<LF>;
    const size_t smem = LY::bytes(p.G4, p.L);
    if ((int)smem > max_smem_optin()) return fail(XC_ERR_UNSUPPORTED, "group search kernel needs %zu B of shared memory", smem);
    static KernelOnce once;
    XC_TRY(once.configure(kern, max_smem_optin()));
    GroupDenseArgs a{rx, rxlen, shift0, lo, hi, d_starts, reinterpret_cast<const float4*>(p.d_w), p.d_off, p.G, p.G4, p.L};
    const int64_t ntiles = (K + LY::ST - 1) / LY::ST;
    int64_t ctas = (int64_t)sm_count();
    if (ctas > ntiles) ctas = ntiles;
    if (ctas < 1) ctas = 1;
    profile_begin(stream);
    kern<<<(unsigned)ctas, 512, smem, stream>>>(a, out, K);
    profile_end(stream);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

// y: host weights [G][L] (conjugation already applied); the plan is usable when the fold fits the kernel's limits
int32_t group_pp_build(const float2* y, const int64_t* starts, int G, int L, int F, GroupPPPlan* p) {
    memset(p, 0, sizeof(*p));
    if (F == 256) return group_dense_build<8>(y, starts, G, L, p);
    if (F == 128) return group_dense_build<7>(y, starts, G, L, p);
    if (F == 64) return group_dense_build<6>(y, starts, G, L, p);
    if (F == GPPL8K::N) return group_rot_build<GPPL8K, GP_B8K, GroupRot2Layout<GPPL8K, GP_B8K>, GR2_MAX_E>(y, starts, G, L, p);
    if (F == GPPL2K::N) return group_rot_build<GPPL2K, GP_B2K>(y, starts, G, L, p);
    if (F == GPPL1K::N) return group_rot_build<GPPL1K, GP_B1K, GroupWarpLayout<GP_B1K>>(y, starts, G, L, p);
    if (F == 512 || F == 32) {
        // X_F[k] = X_1024[k 1024 / F]: the 1024-point kernel, every (1024 / F)-th bin
        XC_TRY((group_rot_build<GPPL1K, GP_B1K, GroupWarpLayout<GP_B1K>>(y, starts, G, L, p)));
        if (p->usable) p->bin_shift = F == 512 ? 1 : 5;
        return XC_OK;
    }
    if (F != GPPL::N) return XC_OK;
    using LY = GroupPPLayout<GPPL, GP_B>;
    // window slack: the largest multiple of B (at most 252) that fits shared memory
    int W = ((int)(((int64_t)max_smem_optin() - (int64_t)LY::OFF_STAGE) / (int64_t)(sizeof(float2) * (size_t)G)) - L) / GP_B * GP_B;
    if (W > 252 / GP_B * GP_B) W = 252 / GP_B * GP_B;
    if (W < 4 * GP_B) return XC_OK;
    constexpr int TG = GPPL::TG, V = GPPL::V, R1 = GPPL::R1;
    struct Contrib { uint32_t off; float2 w; };
    std::vector<std::vector<Contrib>> cells(F);
    const int SW = L + W;
    for (int g = 0; g < G; ++g)
        for (int n = 0; n < L; ++n) {
            const int64_t a = starts[g] + n;
            const int64_t m = ((a % F) + F) % F;
            cells[(size_t)m].push_back(Contrib{(uint32_t)(((int64_t)g * SW + n) * 8), y[(size_t)g * L + n]});
        }
    int base[V + 1];
    base[0] = 0;
    int qtop = 0;
    for (int c = 0; c < V; ++c) {
        const int q = c / R1, n1 = c % R1;
        size_t len = 0;
        for (int t = 0; t < TG; ++t) {
            const size_t m = (size_t)(n1 * GPPL::M1 + t + TG * q);
            if (cells[m].size() > len) len = cells[m].size();
        }
        len = (len + 2 * GP_CH - 1) / (2 * GP_CH) * (2 * GP_CH);
        if (len > 248) {
            memset(p, 0, sizeof(*p));
            return group_rot_build<GPPL, GP_B>(y, starts, G, L, p);
        }
        if (len > 0) qtop = q;
        p->lens[c] = (unsigned char)len;
        base[c + 1] = base[c] + (int)len;
    }
    const int NC = base[V] + GP_CH;                       // one chunk of padding: the gather prefetches one chunk ahead
    if (base[V] == 0 || NC > GP_MAX_NC) {
        memset(p, 0, sizeof(*p));
        return group_rot_build<GPPL, GP_B>(y, starts, G, L, p);
    }
    int nq = 1;
    while (nq < qtop + 1) nq <<= 1;
    std::vector<float2> hw((size_t)NC * TG, make_float2(0.f, 0.f));
    std::vector<uint32_t> ho((size_t)NC * TG, 0u);
    for (int c = 0; c < V; ++c) {
        const int q = c / R1, n1 = c % R1;
        for (int t = 0; t < TG; ++t) {
            const size_t m = (size_t)(n1 * GPPL::M1 + t + TG * q);
            for (size_t j = 0; j < cells[m].size(); ++j) {
                hw[((size_t)base[c] + j) * TG + t] = cells[m][j].w;
                ho[((size_t)base[c] + j) * TG + t] = cells[m][j].off;
            }
        }
    }
    XC_CUDA_TRY(cudaMalloc(&p->d_w, sizeof(float2) * hw.size()));
    XC_CUDA_TRY(cudaMalloc(&p->d_off, sizeof(uint32_t) * ho.size()));
    XC_CUDA_TRY(cudaMemcpy(p->d_w, hw.data(), sizeof(float2) * hw.size(), cudaMemcpyHostToDevice));
    XC_CUDA_TRY(cudaMemcpy(p->d_off, ho.data(), sizeof(uint32_t) * ho.size(), cudaMemcpyHostToDevice));
    p->G = G;
    p->L = L;
    p->W = W;
    p->NC = NC;
    p->nq_live = nq;
    p->B = GP_B;
    p->F = F;
    p->usable = 1;
    return XC_OK;
}

template <class PL, int B, bool ROT_ONLY>
static int32_t group_pp_launch(const GroupPPPlan& p, const int64_t* d_starts, const float2* rx, int64_t rxlen, int64_t shift0,
                               int64_t K, int64_t lo, int64_t hi, const FusedOut& out, cudaStream_t stream) {
    using LY = GroupPPLayout<PL, B, ROT_ONLY>;
    constexpr int NQA = PL::V / PL::R1;
    void (*kern)(const GroupPPArgs, const GroupPPLens, const GroupRotPlan, const FusedOut, const int64_t, const int) =
        group_pp_kernel<PL, B, NQA, true>;
    int slot = 4;
    if constexpr (!ROT_ONLY) {
        if (!p.rot) {
            kern = p.nq_live == 1 ? group_pp_kernel<PL, B, 1, false> : p.nq_live == 2 ? group_pp_kernel<PL, B, 2, false>
                 : p.nq_live == 4 ? group_pp_kernel<PL, B, 4, false> : group_pp_kernel<PL, B, 8, false>;
            slot = p.nq_live == 1 ? 0 : p.nq_live == 2 ? 1 : p.nq_live == 4 ? 2 : 3;
        }
    } else if (!p.rot) {
        return fail(XC_ERR_UNSUPPORTED, "no per-cell list variant for fftlen %d", PL::N);
    }
    const size_t smem = LY::bytes(p.G, p.L, p.W) + (p.rot ? sizeof(float2) * PL::TG : 0);
    if ((int)smem > max_smem_optin()) return fail(XC_ERR_UNSUPPORTED, "group search kernel needs %zu B of shared memory", smem);
    static KernelOnce once[5];
    XC_TRY(once[slot].configure(kern, max_smem_optin()));
    GroupPPArgs a{rx, rxlen, shift0, lo, hi, d_starts, p.d_w, p.d_off, p.G, p.L, p.NC, p.nq_live, p.W};
    GroupPPLens lens;
    memcpy(lens.len, p.lens, sizeof(lens.len));
    GroupRotPlan rp;
    memcpy(rp.cnt, p.rcnt, sizeof(rp.cnt));
    memcpy(rp.desc, p.rdesc, sizeof(rp.desc));
    const int64_t per = (K + B - 1) / B;
    int64_t ctas = (int64_t)sm_count();
    if (ctas > per) ctas = per;
    if (ctas < 1) ctas = 1;
    const int64_t chunk = (per + ctas - 1) / ctas;
    if (chunk >= (1ll << 29)) return fail(XC_ERR_UNSUPPORTED, "too many shifts per thread group");
    profile_begin(stream);
    kern<<<(unsigned)ctas, PL::TG * B, smem, stream>>>(a, lens, rp, out, K, (int)chunk);
    profile_end(stream);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

static int32_t group_warp_launch(const GroupPPPlan& p, const int64_t* d_starts, const float2* rx, int64_t rxlen, int64_t shift0,
                                 int64_t K, int64_t lo, int64_t hi, const FusedOut& out, cudaStream_t stream) {
    constexpr int B = GP_B1K;
    using LY = GroupWarpLayout<B>;
    auto kern = group_warp_kernel<B>;
    const size_t smem = LY::bytes(p.G, p.L, p.W);
    if ((int)smem > max_smem_optin()) return fail(XC_ERR_UNSUPPORTED, "group search kernel needs %zu B of shared memory", smem);
    static KernelOnce once;
    XC_TRY(once.configure(kern, max_smem_optin()));
    GroupPPArgs a{rx, rxlen, shift0, lo, hi, d_starts, p.d_w, p.d_off, p.G, p.L, p.NC, p.nq_live, p.W, p.bin_shift};
    GroupRotPlan rp;
    memcpy(rp.cnt, p.rcnt, sizeof(rp.cnt));
    memcpy(rp.desc, p.rdesc, sizeof(rp.desc));
    const int64_t per = (K + B - 1) / B;
    int64_t ctas = (int64_t)sm_count();
    if (ctas > per) ctas = per;
    if (ctas < 1) ctas = 1;
    const int64_t chunk = (per + ctas - 1) / ctas;
    if (chunk >= (1ll << 26)) return fail(XC_ERR_UNSUPPORTED, "too many shifts per warp");
    profile_begin(stream);
    kern<<<(unsigned)ctas, 32 * B, smem, stream>>>(a, rp, out, K, (int)chunk);
    profile_end(stream);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

static int32_t group_rot2_launch(const GroupPPPlan& p, const int64_t* d_starts, const float2* rx, int64_t rxlen, int64_t shift0,
                                 int64_t K, int64_t lo, int64_t hi, const FusedOut& out, cudaStream_t stream) {
    constexpr int B = GP_B8K;
    using LY = GroupRot2Layout<GPPL8K, B>;
    auto kern = group_rot2_kernel<GPPL8K, B>;
    const size_t smem = LY::bytes(p.G, p.L, p.W);
    if ((int)smem > max_smem_optin()) return fail(XC_ERR_UNSUPPORTED, "group search kernel needs %zu B of shared memory", smem);
    static KernelOnce once;
    XC_TRY(once.configure(kern, max_smem_optin()));
    GroupPPArgs a{rx, rxlen, shift0, lo, hi, d_starts, p.d_w, p.d_off, p.G, p.L, p.NC, p.nq_live, p.W};
    GroupRot2Plan rp;
    memcpy(rp.cnt, p.rcnt, sizeof(rp.cnt));
    memcpy(rp.desc, p.rdesc, sizeof(rp.desc));
    const int64_t per = (K + B - 1) / B;
    int64_t ctas = (int64_t)sm_count();
    if (ctas > per) ctas = per;
    if (ctas < 1) ctas = 1;
    const int64_t chunk = (per + ctas - 1) / ctas;
    if (chunk >= (1ll << 29)) return fail(XC_ERR_UNSUPPORTED, "too many shifts per thread group");
    profile_begin(stream);
    kern<<<(unsigned)ctas, GPPL8K::TG * B, smem, stream>>>(a, rp, out, K, (int)chunk);
    profile_end(stream);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

int32_t group_pp_run(const GroupPPPlan& p, const int64_t* d_starts, const float2* rx, int64_t rxlen, int64_t shift0, int64_t K,
                     int64_t lo, int64_t hi, const FusedOut& out, cudaStream_t stream) {
    if (p.F == GPPL8K::N) return group_rot2_launch(p, d_starts, rx, rxlen, shift0, K, lo, hi, out, stream);
    if (p.dense && p.F == 256) return group_dense_launch<8>(p, d_starts, rx, rxlen, shift0, K, lo, hi, out, stream);
    if (p.dense && p.F == 128) return group_dense_launch<7>(p, d_starts, rx, rxlen, shift0, K, lo, hi, out, stream);
    if (p.dense && p.F == 64) return group_dense_launch<6>(p, d_starts, rx, rxlen, shift0, K, lo, hi, out, stream);
    if (p.F == GPPL1K::N) return group_warp_launch(p, d_starts, rx, rxlen, shift0, K, lo, hi, out, stream);
    if (p.F == GPPL2K::N) return group_pp_launch<GPPL2K, GP_B2K, true>(p, d_starts, rx, rxlen, shift0, K, lo, hi, out, stream);
    return group_pp_launch<GPPL, GP_B, false>(p, d_starts, rx, rxlen, shift0, K, lo, hi, out, stream);
}

}  // namespace xc
