// group_pp.cuh -- GroupXcorrFFT over a CONTIGUOUS shift range in the ping-pong layout of fused_fft_pp.cuh
// (B free-running thread groups per SM, per-thread constants in tensor memory), for fftlen 4096.
//
// The folded formulation of group.cu (one F-point FFT of the products folded into F cells) with three changes:
//   * the per-group rx windows rx[s + s_g, s + s_g + L + W) are staged in shared memory (G x (L + W) samples,
//     refilled every W / B iterations), so the gather of the G L products per shift reads LDS instead of L2
//     (profiles/r1_c3_group_ncu.md: 33 % of the samples sat in that loop with long-scoreboard stalls);
//   * the contributor lists of a thread (weight, stage offset) live in tensor memory next to the twiddle tables:
//     nothing per thread is re-read from shared or global memory;
//   * zero padding is not transformed: when the folded groups only reach the first 128 nq cells of each 1024-cell
//     quarter (nq in {1, 2, 4}; C3: 64 groups x 128 at starts 1024 g -> nq = 1), pass 1 runs nq of its 8 columns and
//     pass 2 is a radix-32 codelet with 4 nq non-zero inputs (copy stages cost nothing).
// Folds that do not fit (more than 96 padded contributors per thread, windows larger than the shared-memory budget,
// Bluestein lengths, shift lists with a step) stay on fused_fft_kernel<GroupLoader>.
#pragma once
#include "fused_fft_pp.cuh"

namespace xc {

constexpr int GP_MAX_NC = 96;      // padded contributors per thread incl. one chunk of padding (tensor-memory columns: 3 each)
constexpr int GP_CH = 8;           // contributors per tcgen05.ld pair (lists are padded to 2 GP_CH)

// ROT variant (group starts that are not aligned to the 128-cell rows of the transform: ragged spacings, starts 1000 g, ...).
// The per-cell lists above are padded to the longest list of a register slot, which a ragged fold blows up (2 contributors
// per cell on average become 16 per slot).  Here every group is cut into sub-groups of TG = 128 samples; sub-group (a = first
// cell mod F, r = a mod 128, A = a / 128) lands on rows A (cells r .. 127) and A + 1 (cells 0 .. r - 1) with the SAME thread
// rotation: thread t multiplies sample l = (t - r) mod 128 of the sub-group by its weight (one tensor-memory column pair per
// sub-group and thread, no duplicates) and adds the product to row A when t >= r, else to a carry that opens row A + 1.
// Sub-groups are sorted by row; the row loop is unrolled (static register index), the sub-group loop of a row is a run-time
// loop whose trip count is the same for every thread.  Capacity: GR_MAX_E sub-groups (2 tensor-memory columns each).
constexpr int GR_MAX_E = 192;
struct GroupRotPlan {
    unsigned char cnt[32];         // sub-groups whose first cell lies in row rho (sorted by row)
    uint32_t desc[GR_MAX_E];       // r | (first stage cell of the sub-group << 7)
};

struct GroupPPLens {
    unsigned char len[32];         // padded list length of register slot c (multiple of GP_CH), uniform over the threads
};

struct GroupPPArgs {
    const float2* rx;
    int64_t rxlen;
    int64_t shift0;                // the shift of entry i is shift0 + i
    int64_t lo, hi;                // valid shifts satisfy s + lo >= 0 and s + hi <= rxlen
    const int64_t* starts;         // [G] device
    const float2* w;               // [NC][TG] weights (conj applied), device
    const uint32_t* off;           // [NC][TG] byte offsets into the stage at iteration offset 0
    int G, L, NC, nq_live, W;
    int bin_shift = 0;             // group_warp_kernel: report bin >> bin_shift of the bins that are multiples of 2^bin_shift
};

namespace tc {
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t* r) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "r"(r[0]), "r"(r[1]),
                 "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
                 : "memory");
}
__device__ __forceinline__ void tmem_ld8_issue(uint32_t taddr, uint32_t* u) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7])
                 : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld2_issue(uint32_t taddr, uint32_t* u) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(u[0]), "=r"(u[1]) : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait2(uint32_t* u) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(u[0]), "+r"(u[1]));
}
__device__ __forceinline__ void tmem_ld_wait8(uint32_t* u) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(u[0]), "+r"(u[1]), "+r"(u[2]), "+r"(u[3]), "+r"(u[4]), "+r"(u[5]), "+r"(u[6]), "+r"(u[7]));
}
}  // namespace tc

// stages h >= H0 of codelet_fft<R> when only the first NZ inputs are non-zero: a stage whose partner distance
// R / (2 h) is >= NZ adds a zero, i.e. it copies (u + w 0, u - w 0) = (u, u)
template <int R, int NZ, int H0>
__device__ __forceinline__ void codelet_fft_pruned(float2* a) {
    constexpr int LG = ilog2(R);
#pragma unroll
    for (int h = H0; h < R; h <<= 1) {
#pragma unroll
        for (int blk = 0; blk < R; blk += 2 * h) {
#pragma unroll
            for (int j = 0; j < h; ++j) {
                if (R / (2 * h) >= NZ) {
                    a[bitrev(blk + j + h, LG)] = a[bitrev(blk + j, LG)];
                } else {
                    dit_butterfly(a[bitrev(blk + j, LG)], a[bitrev(blk + j + h, LG)], j * (16 / h));
                }
            }
        }
    }
}

template <class PL, int B, bool ROT = false>
struct GroupPPLayout {
    static_assert(PL::P == 3 && PL::R2 == 32 && PL::R3 == 32 && PL::V == 32 && (PL::TG == 128 || (ROT && PL::TG == 64)),
                  "4096-point plan (128 threads x 32 points); 2048 points (64 threads) in the rotated-row variant");
    static constexpr size_t OFF_BUF = 0;
    static constexpr size_t OFF_GMAX = sizeof(float2) * (size_t)PL::BUF * B;          // [B][2]
    static constexpr size_t OFF_BIN = OFF_GMAX + sizeof(unsigned int) * 2 * B;        // [B]
    static constexpr size_t OFF_TADDR = OFF_BIN + sizeof(unsigned int) * B;
    static constexpr size_t OFF_STAGE = (OFF_TADDR + 16 + 15) / 16 * 16;              // G x (L + W) samples
    __host__ __device__ static constexpr size_t bytes(int G, int L, int W) { return OFF_STAGE + sizeof(float2) * (size_t)G * (size_t)(L + W); }
    // tensor-memory columns of one thread
    static constexpr int COL_W = 0;                       // weights, 2 columns per contributor
    static constexpr int COL_O = 2 * GP_MAX_NC;           // stage byte offsets
    static constexpr int COL_T1 = ROT ? 2 * GR_MAX_E : 3 * GP_MAX_NC;   // pass-2 input twiddles, chunk layout of fused_fft_pp.cuh
    static constexpr int COL_T2 = COL_T1 + 64;            // pass-3 input twiddles
    static constexpr int TMEM_COLS = 512;
    static_assert(COL_T2 + 64 <= 512, "tensor memory holds 512 columns");
};

// pass 2 with NZ = 4 nq non-zero inputs (the columns c < 128 nq of the first exchange): loads, folded first stage,
// pruned tail, in-place stores
template <class PL, int NZ>
__device__ __forceinline__ void gp_pass2(float2* v, float2* buf, const int o2, const uint32_t t1) {
    constexpr int ROW = PL::ROW, R2 = PL::R2;
    uint32_t cu[2][16];
    tc::tmem_ld16_issue(t1, cu[0]);
#pragma unroll
    for (int n2 = 0; n2 < R2; ++n2) v[n2] = (n2 < NZ) ? buf[o2 + n2 * ROW] : make_float2(0.f, 0.f);
    if (NZ == R2) {
#pragma unroll
        for (int m = 0; m < 4; ++m) {
            tc::tmem_ld_wait16(cu[m & 1]);
            if (m + 1 < 4) tc::tmem_ld16_issue(t1 + 16 * (m + 1), cu[(m + 1) & 1]);
            folded_stage_chunk(v, m, cu[m & 1]);
        }
        codelet_fft<R2, 2>(v);
    } else {
        // a[i + 16] = 0: the first stage is a[i] = a[i + 16] = T1[i] a[i]
#pragma unroll
        for (int m = 0; m < NZ / 4; ++m) {
            tc::tmem_ld_wait16(cu[m & 1]);
            if (m + 1 < NZ / 4) tc::tmem_ld16_issue(t1 + 16 * (m + 1), cu[(m + 1) & 1]);
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int i = 4 * m + r;
                v[i] = pcmul(v[i], make_float2(__uint_as_float(cu[m & 1][4 * r]), __uint_as_float(cu[m & 1][4 * r + 1])));
                v[i + 16] = v[i];
            }
        }
        codelet_fft_pruned<R2, NZ, 2>(v);
    }
#pragma unroll
    for (int j = 0; j < R2; ++j) buf[o2 + bitrev(j, ilog2(R2)) * ROW] = v[j];
}

// NQ = number of live pass-1 columns (1, 2, 4 or 8 = all): a template parameter, so that the gather is only unrolled
// over the 4 NQ register slots that can have contributors and the code of one variant stays inside the instruction cache
template <class PL, int B, int NQ, bool ROT>
__global__ void __launch_bounds__(PL::TG* B, 1) group_pp_kernel(const GroupPPArgs a, const GroupPPLens lens,
                                                                const __grid_constant__ GroupRotPlan rp, const FusedOut out,
                                                                const int64_t K, const int chunk) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    static_assert(!ROT || NQ == PL::V / PL::R1, "the rotated fold reaches every row");
    using LY = GroupPPLayout<PL, B, ROT>;
    const int W = a.W;                    // extra samples behind every group window (a multiple of B)
    constexpr int TG = PL::TG, V = PL::V, RL = PL::RL, ROW = PL::ROW, R1 = PL::R1;
    const int g = threadIdx.x / TG;       // thread group within the CTA
    const int t = threadIdx.x % TG;
    const int warp_cta = threadIdx.x >> 5;
    float2* const buf = reinterpret_cast<float2*>(smem_raw + LY::OFF_BUF) + (size_t)g * PL::BUF;
    float2* const stage = reinterpret_cast<float2*>(smem_raw + LY::OFF_STAGE);
    unsigned int* const gmax = reinterpret_cast<unsigned int*>(smem_raw + LY::OFF_GMAX) + 2 * g;
    unsigned int* const binmin = reinterpret_cast<unsigned int*>(smem_raw + LY::OFF_BIN) + g;
    uint32_t* const taddr_slot = reinterpret_cast<uint32_t*>(smem_raw + LY::OFF_TADDR);
    if (warp_cta == 0) tc::tmem_alloc(tc::smem_u32(taddr_slot), LY::TMEM_COLS);
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
    // groups whose warps cover the four lane quarters between them write (one group of 128 threads, two of 64)
    if (g < (TG >= 128 ? 1 : 128 / TG)) {
        // ---- per-thread constants -> tensor memory (read by the same in-group thread of every group) ----
        for (int e0 = 0; e0 < a.NC; e0 += GP_CH) {
            float r[16];
            uint32_t o[8];
#pragma unroll
            for (int i = 0; i < GP_CH; ++i) {
                const float2 w = a.w[(size_t)(e0 + i) * TG + t];
                r[2 * i] = w.x;
                r[2 * i + 1] = w.y;
                if (!ROT) o[i] = a.off[(size_t)(e0 + i) * TG + t];
            }
            tc::tmem_st16(tb + LY::COL_W + 2 * e0, r);
            if (!ROT) tc::tmem_st8(tb + LY::COL_O + e0, o);
        }
        float r[16];
        const int k1 = t / PL::M2, d = t % PL::M2;
#pragma unroll 1
        for (int m = 0; m < 4; ++m) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int n2 = 4 * m + (i >> 1) + 16 * (i & 1);
                const float2 w = unit_root(((n2 * PL::M2 + d) * k1) % PL::N, PL::N);
                r[2 * i] = w.x;
                r[2 * i + 1] = w.y;
            }
            tc::tmem_st16(tb + LY::COL_T1 + 16 * m, r);
        }
        const int k2 = t % PL::R2;
#pragma unroll 1
        for (int m = 0; m < 4; ++m) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int dd = 4 * m + (i >> 1) + 16 * (i & 1);
                const float2 w = unit_root((dd * k2) % PL::M1, PL::M1);
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

    // this CTA owns shift-list entries [i0, i0 + B * chunk); group g takes i0 + B * it + g
    const int64_t i0 = (int64_t)blockIdx.x * B * (int64_t)chunk;
    int64_t n_left = K - i0;
    if (n_left < 0) n_left = 0;
    if (n_left > B * (int64_t)chunk) n_left = B * (int64_t)chunk;
    const int n_act = (int)n_left;
    const int64_t s_first = a.shift0 + i0;
    const GroupBar<TG, B> bar{1 + g};
    const int SW = a.L + W;
    const int nstage = a.G * SW;
    if (ROT) {   // sub-groups of a length that is not a multiple of TG read (with weight zero) up to TG cells behind the last window
        for (int i = threadIdx.x; i < TG; i += TG * B) stage[nstage + i] = make_float2(0.f, 0.f);
    }
    const uint32_t stage_s = tc::smem_u32(stage);
    constexpr int K1STEP = (PL::M1 / RL) * ROW;
    constexpr int QSTEP1 = (TG / RL) * ROW;
    const int o1 = (t / RL) * ROW + (t % RL);
    const int o2 = (t / PL::M2) * K1STEP + (t % PL::M2);
    float2 v[V];
    float pend = 0.f;
    int o = W;

    for (int it = 0; it < chunk; ++it) {
        const int e = B * it + g;
        if (o == W) {
            // refill: stage[gi (L + W) + j] = rx[s + s_gi + j] (zero outside the record)
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
        if constexpr (ROT) {
            // ---- rotated gather: sub-groups sorted by row; thread t takes sample (t - r) mod TG of each ----
            // software pipeline over the sorted sub-group list, independent of the row structure: descriptor two entries
            // ahead, sample and weight one entry ahead (the loop body would otherwise be one serial chain
            // LDC -> address -> LDS -> FMA per sub-group; two zero entries follow the last sub-group)
            uint32_t wn[2];
            tc::tmem_ld2_issue(tb + LY::COL_W, wn);
            uint32_t cw = tb + LY::COL_W;
            int e = 0;
            uint32_t d_n = rp.desc[0], d_nn = rp.desc[1];
            float2 x_n;
            asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(x_n.x), "=f"(x_n.y) : "r"(xs + 8u * ((d_n >> 7) + (uint32_t)((t - (int)(d_n & 127u)) & (TG - 1)))));
            float2 carry = make_float2(0.f, 0.f);
#pragma unroll
            for (int rho = 0; rho < V; ++rho) {
                // row rho = cells [128 rho, 128 rho + 128) = (n1, q) with 128 rho = n1 M1 + TG q
                constexpr int QPR = PL::M1 / TG;                 // rows per n1 block
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
                    // next entry on its way
                    d_n = d_nn;
                    e += 1;
                    d_nn = rp.desc[e + 1];
                    cw += 2;
                    tc::tmem_ld2_issue(cw, wn);
                    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(x_n.x), "=f"(x_n.y) : "r"(xs + 8u * ((d_n >> 7) + (uint32_t)((t - (int)(d_n & 127u)) & (TG - 1)))));
                    const float2 p = pcmul(x, y);
                    if (t >= r) acc = padd(acc, p);
                    else carry = padd(carry, p);
                }
                v[c] = acc;
            }
            v[0] = padd(v[0], carry);                            // row 31 wraps to row 0 (cells mod F)
            tc::tmem_ld_wait2(wn);
        } else {
        // ---- gather: v[c] = sum_j w[c][j] x[off[c][j]] over the contributor list of register slot c; the chunk of 8
            // (weights, offsets) after the current one is already on its way from tensor memory, four partial sums ----
            {
                uint32_t cw = tb + LY::COL_W, co = tb + LY::COL_O;
                uint32_t wv[2][16], ov[2][8];
                tc::tmem_ld16_issue(cw, wv[0]);
                tc::tmem_ld8_issue(co, ov[0]);
                int par = 0;                               // runtime parity of the chunk ring: kept uniform by two code copies
#pragma unroll
                for (int c = 0; c < R1 * NQ; ++c) {
                    float2 acc[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) acc[i] = make_float2(0.f, 0.f);
                    const int n_c = lens.len[c];
#pragma unroll 1
                    for (int j = 0; j < n_c; j += 2 * GP_CH) {
                        // lists are padded to a multiple of 16 contributors: two chunks per trip, static register parity
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            tc::tmem_ld_wait8(ov[h]);
                            float2 x[GP_CH];
#pragma unroll
                            for (int u = 0; u < GP_CH; ++u)
                                asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(x[u].x), "=f"(x[u].y) : "r"(xs + ov[h][u]));
                            tc::tmem_ld_wait16(wv[h]);
                            cw += 2 * GP_CH;
                            co += GP_CH;
                            tc::tmem_ld16_issue(cw, wv[h ^ 1]);      // the next chunk (one chunk of padding follows the last list)
                            tc::tmem_ld8_issue(co, ov[h ^ 1]);
#pragma unroll
                            for (int u = 0; u < GP_CH; ++u) {
                                const float2 y = make_float2(__uint_as_float(wv[h][2 * u]), __uint_as_float(wv[h][2 * u + 1]));
                                acc[u & 3] = pfma(make_float2(x[u].y, x[u].x), make_float2(-y.y, y.y), pfma(x[u], make_float2(y.x, y.x), acc[u & 3]));
                            }
                        }
                    }
                    v[c] = padd(padd(acc[0], acc[1]), padd(acc[2], acc[3]));
                }
                (void)par;
                tc::tmem_ld_wait8(ov[0]);                  // drain the last prefetch
#pragma unroll
                for (int c = R1 * NQ; c < V; ++c) v[c] = make_float2(0.f, 0.f);
            }
        }
        // ---- pass 1: radix-4 columns q < nq_live (the others are all zero), raw stores ----
#pragma unroll
        for (int q = 0; q < V / R1; ++q) {
            if (q < NQ) {
                codelet_fft<R1>(v + q * R1);
#pragma unroll
                for (int j = 0; j < R1; ++j) buf[o1 + bitrev(j, ilog2(R1)) * K1STEP + q * QSTEP1] = v[q * R1 + j];
            }
        }
        bar();
        // ---- pass 2 (in place, warp local from here on) ----
        gp_pass2<PL, ROT ? PL::R2 : 4 * NQ>(v, buf, o2, tb + LY::COL_T1);
        __syncwarp();
        // ---- pass 3 ----
        {
            uint32_t cu[2][16];
            tc::tmem_ld16_issue(tb + LY::COL_T2, cu[0]);
            const float4* const b3 = reinterpret_cast<const float4*>(buf + t * ROW);
#pragma unroll
            for (int d2 = 0; d2 < RL / 2; ++d2) {
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
            codelet_fft<RL, 2>(v);
        }
        if (t == 0 && it > 0) {                        // results of the previous FFT of this group
            const int p = e - B;
            if (p < n_act) {
                const unsigned int b = *binmin;
                if (out.qf2) out.qf2[i0 + p] = pend;
                if (out.fidx) out.fidx[i0 + p] = (b == 0xffffffffu) ? 0u : b;
            }
            *binmin = 0xffffffffu;
        }
        // ---- |X|^2, group maximum, lowest arg max (as fused_fft_pp_kernel) ----
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
        bar();   // all last-pass loads done (buffer free) + group maximum complete
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
                        if (nv[j] == m) mybin = min(mybin, (unsigned int)output_bin<PL>(t, j / RL, j % RL));
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
    if (warp_cta == 0) tc::tmem_dealloc(tmem0, LY::TMEM_COLS);
}

}  // namespace xc
