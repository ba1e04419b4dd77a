// fused_fft_pp.cuh -- frequency search over a CONTIGUOUS shift range: B independent thread groups per SM that
// share one staged rx window, with every per-thread constant (conj(cutout), both twiddle sets) held in TENSOR
// MEMORY instead of shared memory.
//
// Why: the staged layout (fused_fft_staged.cuh) keeps one group of 8 warps per SM in lockstep (CTA-wide
// barriers), so its FMA phases and its shared-memory phases overlap only through the instruction-level
// parallelism of the two FFTs a thread carries (profiles/r1_c2_staged_ncu.md: FMA pipe 66 %, shared-memory data
// pipe 67 %).  Two free-running groups give every scheduler two warps that are in DIFFERENT phases of their
// FFTs, but two 68 KB exchange buffers plus the 80 KB window only fit beside each other when the 64 KB
// conj(cutout) and the 16.5 KB twiddle tables leave shared memory.  They are per-thread constants (a thread
// always multiplies by the same V cutout values and the same twiddle row), which is exactly what a TMEM lane
// holds: tcgen05.st once per kernel, tcgen05.ld per FFT (TMEM read port, not the LSU / shared-memory pipe).
// No tcgen05.mma is issued: TMEM is used as a 256 KB per-SM register-file extension.
//
//   * group g of a CTA handles shift-list entries i0 + B * it + g; all groups read their windows from the one
//     staged copy rx[base, base + N + W) (refilled every W / B iterations behind a CTA-wide barrier);
//   * one exchange buffer per group, two group barriers (bar.sync id, TG) per FFT, the second exchange is
//     warp local and in place;
//   * TG = 128, V = 64 points per thread for N = 8192 (a warp owns two 1024-point sub-FFTs after pass 1).
// Results are bit-identical to fused_fft_kernel / fused_fft_staged_kernel (same codelets, same twiddle values,
// same operation order per point).  (max, arg max) outputs only.
#pragma once
#include "fused_fft_kernel.cuh"
#include "mt_tc.cuh"

namespace xc {

namespace tc {
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float* r) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
        "r"(__float_as_uint(r[0])), "r"(__float_as_uint(r[1])), "r"(__float_as_uint(r[2])), "r"(__float_as_uint(r[3])),
        "r"(__float_as_uint(r[4])), "r"(__float_as_uint(r[5])), "r"(__float_as_uint(r[6])), "r"(__float_as_uint(r[7])),
        "r"(__float_as_uint(r[8])), "r"(__float_as_uint(r[9])), "r"(__float_as_uint(r[10])), "r"(__float_as_uint(r[11])),
        "r"(__float_as_uint(r[12])), "r"(__float_as_uint(r[13])), "r"(__float_as_uint(r[14])), "r"(__float_as_uint(r[15]))
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// asynchronous load of 16 columns; the registers are valid after tmem_ld_wait16() on the same array
__device__ __forceinline__ void tmem_ld16_issue(uint32_t taddr, uint32_t* u) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]), "=r"(u[8]),
          "=r"(u[9]), "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15])
        : "r"(taddr));
}
// waits for every outstanding tcgen05.ld of this thread; the "+r" operands tie the consumers of `u` behind it.
// No "memory" clobber: the tables in tensor memory are written once per kernel (behind tcgen05.wait::st and a
// CTA barrier), and a clobber would pin every shared-memory load / store of the FFT to its side of the wait.
__device__ __forceinline__ void tmem_ld_wait16(uint32_t* u) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(u[0]), "+r"(u[1]), "+r"(u[2]), "+r"(u[3]), "+r"(u[4]), "+r"(u[5]), "+r"(u[6]), "+r"(u[7]),
                   "+r"(u[8]), "+r"(u[9]), "+r"(u[10]), "+r"(u[11]), "+r"(u[12]), "+r"(u[13]), "+r"(u[14]), "+r"(u[15]));
}
}  // namespace tc

// FOLD = true: no twiddle is applied behind a pass; the values are stored raw and the twiddles ride on the FIRST butterfly
// stage of the next pass (folded_butterfly(), five instead of six packed instructions per pair, and no power chain behind
// pass 1), read from per-thread tensor-memory tables: T1[q][n2] = W_N^((n2 M2 + d) k1) for the pass-2 inputs and
// T2[d] = W_M1^(d k2) for the pass-3 inputs, stored as x16 chunks {T[4m + r], T[4m + r + 16]}, r < 4, that feed four
// butterflies each.  9.7 % fewer FMA-pipe instructions than the table / power-chain form; the twiddles are exact
// (evaluated in double), so results differ from the B = 2 layout in the last bits.
template <class PL, int B, bool FOLD>
struct PPLayout {
    static_assert(PL::P == 3 && PL::R2 == 32 && PL::R3 == 32 && PL::V % 32 == 0, "three passes, radix-32 tail");
    static_assert(PL::TG == 64 || PL::TG % 128 == 0, "groups of 64 threads pair up over the four TMEM lane quarters");
    static_assert(PL::R1 == 8 || PL::R1 == 4 || PL::R1 == 2, "one tcgen05.ld.x16 = the 8 conj(cutout) values of 8 / R1 pass-1 columns");
    static_assert(FOLD || PL::R1 == 8, "the power-chain variant exists for the 8192-point plan only");
    // groups whose warps cover all four lane quarters between them write the constants (1 group of >= 128 threads,
    // 2 groups of 64); every other group reads the lanes of the writer with the same warp position
    static constexpr int WRITERS = (PL::TG >= 128) ? 1 : 128 / PL::TG;
    static constexpr int W = 2048;                                        // extra samples behind the window
    static_assert(W % B == 0, "refills happen at the same iteration in every group");
    static constexpr size_t OFF_BUF = 0;
    static constexpr size_t OFF_STAGE = sizeof(float2) * (size_t)PL::BUF * B;
    static constexpr size_t OFF_GMAX = OFF_STAGE + sizeof(float2) * (size_t)(PL::N + W);   // [B][2] bit patterns
    static constexpr size_t OFF_BIN = OFF_GMAX + sizeof(unsigned int) * 2 * B;             // [B]
    static constexpr size_t OFF_TADDR = OFF_BIN + sizeof(unsigned int) * 2 * B;            // OFF_BIN: [B][2]
    static constexpr size_t BYTES = OFF_TADDR + 16;
    // tensor-memory columns of one thread (32-bit each)
    static constexpr int NQ1 = PL::V / PL::R1;                            // pass-1 columns per thread
    static constexpr int NQ2 = PL::V / PL::R2;                            // pass-2 / pass-3 columns per thread
    static_assert(FOLD || NQ1 <= 8, "the pass-1 base twiddles fit one x16 load");
    static constexpr int COL_CUT = 0;                                     // conj(cutout): slot j = q * R1 + n1
    static constexpr int COL_TW1 = 2 * PL::V;                             // !FOLD: W_N^(t + TG q), q < NQ1; FOLD: T1[q][.]
    static constexpr int COL_TW2 = COL_TW1 + (FOLD ? 2 * PL::R2 * NQ2 : 16);   // !FOLD: W_M1^(d k2), k2 < 32; FOLD: T2[.]
    static constexpr int CW = COL_TW2 + 2 * PL::R2;
    static constexpr int COPIES = (PL::TG >= 128) ? PL::TG / 128 : 1;     // warps of a group per lane quarter
    static constexpr int TMEM_COLS = (CW * COPIES <= 32) ? 32 : (CW * COPIES <= 64) ? 64 : (CW * COPIES <= 128) ? 128 : (CW * COPIES <= 256) ? 256 : 512;
    static_assert(CW * COPIES <= 512, "tensor memory holds 512 columns");
};

// first butterfly stage of a radix-32 column with the pending twiddles of chunk m (pairs i = 4m .. 4m + 3)
__device__ __forceinline__ void folded_stage_chunk(float2* a, const int m, const uint32_t* u) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int i = 4 * m + r;
        folded_butterfly(a[i], a[i + 16], make_float2(__uint_as_float(u[4 * r]), __uint_as_float(u[4 * r + 1])),
                         make_float2(__uint_as_float(u[4 * r + 2]), __uint_as_float(u[4 * r + 3])));
    }
}

// One FFT of the group: products with conj(cutout) (TMEM) of the window x (shared), three passes.
// On return the registers hold the bins of output_bin().  One group barrier inside.
template <class PL, int B, class Bar>
__device__ __forceinline__ void pp_block_fft(float2* v, const float2* x, const uint32_t tb, float2* buf, const int t, Bar bar) {
    using L = PPLayout<PL, B, false>;
    constexpr int RL = PL::RL, ROW = PL::ROW, R1 = PL::R1, R2 = PL::R2, V = PL::V, NQ1 = L::NQ1;
    constexpr int K1STEP = (PL::M1 / RL) * ROW;
    constexpr int QSTEP1 = (PL::TG / RL) * ROW;
    const int o1 = (t / RL) * ROW + (t % RL);
    uint32_t cu[2][16], tw[16];
    tc::tmem_ld16_issue(tb + L::COL_CUT, cu[0]);
    tc::tmem_ld16_issue(tb + L::COL_TW1, tw);
    tc::tmem_ld_wait16(cu[0]);
    tc::tmem_ld_wait16(tw);
    // ---- products + pass 1 + first-exchange stores, column by column ----
#pragma unroll
    for (int q = 0; q < NQ1; ++q) {
        if (q + 1 < NQ1) tc::tmem_ld16_issue(tb + L::COL_CUT + 16 * (q + 1), cu[(q + 1) & 1]);
#pragma unroll
        for (int n1 = 0; n1 < R1; ++n1) {
            const float2 cv = make_float2(__uint_as_float(cu[q & 1][2 * n1]), __uint_as_float(cu[q & 1][2 * n1 + 1]));
            v[q * R1 + n1] = pcmul(x[input_index<PL>(0, q, n1)], cv);
        }
        codelet_fft<R1>(v + q * R1);
        twiddle_powers<R1>(v + q * R1, make_float2(__uint_as_float(tw[2 * q]), __uint_as_float(tw[2 * q + 1])));
#pragma unroll
        for (int j = 0; j < R1; ++j) buf[o1 + bitrev(j, ilog2(R1)) * K1STEP + q * QSTEP1] = v[q * R1 + j];
        if (q + 1 < NQ1) tc::tmem_ld_wait16(cu[(q + 1) & 1]);
    }
    tc::tmem_ld16_issue(tb + L::COL_TW2, cu[0]);      // first 8 pass-2 twiddles, in flight across the barrier
    bar();
    // ---- pass 2 (in place; from here on a warp only touches its own rows of the buffer) ----
    constexpr int QSTEP2 = (PL::TG / PL::M2) * K1STEP;
    const int o2 = (t / PL::M2) * K1STEP + (t % PL::M2);
#pragma unroll
    for (int q = 0; q < V / R2; ++q)
#pragma unroll
        for (int n2 = 0; n2 < R2; ++n2) v[q * R2 + n2] = buf[o2 + q * QSTEP2 + n2 * ROW];
#pragma unroll
    for (int q = 0; q < V / R2; ++q) codelet_fft<R2>(v + q * R2);
#pragma unroll
    for (int c = 0; c < R2 / 8; ++c) {
        tc::tmem_ld_wait16(cu[c & 1]);
        if (c + 1 < R2 / 8) tc::tmem_ld16_issue(tb + L::COL_TW2 + 16 * (c + 1), cu[(c + 1) & 1]);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int k2 = 8 * c + i;
            if (k2 == 0) continue;
            const float2 w = make_float2(__uint_as_float(cu[c & 1][2 * i]), __uint_as_float(cu[c & 1][2 * i + 1]));
            const int j = bitrev(k2, ilog2(R2));
#pragma unroll
            for (int q = 0; q < V / R2; ++q) v[q * R2 + j] = pcmul(v[q * R2 + j], w);
        }
    }
#pragma unroll
    for (int q = 0; q < V / R2; ++q)
#pragma unroll
        for (int j = 0; j < R2; ++j) buf[o2 + q * QSTEP2 + bitrev(j, ilog2(R2)) * ROW] = v[q * R2 + j];
    __syncwarp();   // the rows a warp wrote are exactly the rows it reads next
    // ---- last pass ----
    const float4* const b3 = reinterpret_cast<const float4*>(buf + t * ROW);
#pragma unroll
    for (int q = 0; q < V / RL; ++q)
#pragma unroll
        for (int d2 = 0; d2 < RL / 2; ++d2) {
            const float4 y = b3[(q * PL::TG * ROW) / 2 + d2];
            v[q * RL + 2 * d2] = make_float2(y.x, y.y);
            v[q * RL + 2 * d2 + 1] = make_float2(y.z, y.w);
        }
#pragma unroll
    for (int q = 0; q < V / RL; ++q) codelet_fft<RL>(v + q * RL);
}


// The same with folded twiddles (PPLayout<.., true>).
template <class PL, int B, class Bar>
__device__ __forceinline__ void pp_block_fft_fold(float2* v, const float2* x, const uint32_t tb, float2* buf, const int t, Bar bar) {
    using L = PPLayout<PL, B, true>;
    constexpr int RL = PL::RL, ROW = PL::ROW, R1 = PL::R1, R2 = PL::R2, NQ1 = L::NQ1, NQ2 = L::NQ2;
    constexpr int K1STEP = (PL::M1 / RL) * ROW;
    constexpr int QSTEP1 = (PL::TG / RL) * ROW;
    const int o1 = (t / RL) * ROW + (t % RL);
    uint32_t cu[2][16];
    tc::tmem_ld16_issue(tb + L::COL_CUT, cu[0]);
    // A chunk = 8 conj(cutout) values (one tcgen05.ld.x16) = CPC = 8 / R1 pass-1 columns.  Explicit software pipeline: the
    // window samples and the conj(cutout) chunk after the current one are requested before the arithmetic of the current
    // chunk (volatile loads keep their order; the empty asm ties the arithmetic behind them).
    constexpr int CPC = 8 / R1, NCH = PL::V / 8;
    float2 xs[2][8];
    const uint32_t xa = tc::smem_u32(x);
#pragma unroll
    for (int i = 0; i < 8; ++i) xs[0][i] = x[input_index<PL>(0, i / R1, i % R1)];
    tc::tmem_ld_wait16(cu[0]);
    // ---- products + pass 1 + first-exchange stores (raw), chunk by chunk ----
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
        if (c + 1 < NCH) {
            tc::tmem_ld16_issue(tb + L::COL_CUT + 16 * (c + 1), cu[(c + 1) & 1]);
#pragma unroll
            for (int i = 0; i < 8; ++i)
                asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];"
                             : "=f"(xs[(c + 1) & 1][i].x), "=f"(xs[(c + 1) & 1][i].y)
                             : "r"(xa + 8u * (uint32_t)((i % R1) * PL::M1 + PL::TG * ((c + 1) * CPC + i / R1))));
            asm volatile("" : "+r"(cu[c & 1][0]));
        }
#pragma unroll
        for (int qq = 0; qq < CPC; ++qq) {
            const int q = c * CPC + qq;
            // the products ride on the first butterfly stage of the radix-R1 codelet (pairs n1, n1 + R1 / 2)
#pragma unroll
            for (int n1 = 0; n1 < R1 / 2; ++n1) {
                const int i0 = qq * R1 + n1, i1 = i0 + R1 / 2;
                v[q * R1 + n1] = xs[c & 1][i0];
                v[q * R1 + n1 + R1 / 2] = xs[c & 1][i1];
                folded_butterfly(v[q * R1 + n1], v[q * R1 + n1 + R1 / 2],
                                 make_float2(__uint_as_float(cu[c & 1][2 * i0]), __uint_as_float(cu[c & 1][2 * i0 + 1])),
                                 make_float2(__uint_as_float(cu[c & 1][2 * i1]), __uint_as_float(cu[c & 1][2 * i1 + 1])));
            }
            codelet_fft<R1, 2>(v + q * R1);
#pragma unroll
            for (int j = 0; j < R1; ++j) buf[o1 + bitrev(j, ilog2(R1)) * K1STEP + q * QSTEP1] = v[q * R1 + j];
        }
        if (c + 1 < NCH) tc::tmem_ld_wait16(cu[(c + 1) & 1]);
    }
    tc::tmem_ld16_issue(tb + L::COL_TW1, cu[0]);      // first chunk of T1, in flight across the barrier
    bar();
    // ---- pass 2 (in place; from here on a warp only touches its own rows of the buffer) ----
    constexpr int QSTEP2 = (PL::TG / PL::M2) * K1STEP;
    const int o2 = (t / PL::M2) * K1STEP + (t % PL::M2);
#pragma unroll
    for (int q = 0; q < NQ2; ++q)
#pragma unroll
        for (int n2 = 0; n2 < R2; ++n2) v[q * R2 + n2] = buf[o2 + q * QSTEP2 + n2 * ROW];
#pragma unroll
    for (int c = 0; c < 4 * NQ2; ++c) {               // chunk c = 4 q + m
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
    tc::tmem_ld16_issue(tb + L::COL_TW2, cu[0]);      // first chunk of T2
    __syncwarp();   // the rows a warp wrote are exactly the rows it reads next
    // ---- last pass ----
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

// CPLX: also writes the complex QF X[bin] / (||cutout|| ||window||) of the absResult = False variants
// (xcorrRoutines.py:533-548): a thread that reaches the group maximum keeps its candidate (bin, value) in registers until the
// winning bin is known (one FFT later); the bin slots alternate so that the reset never races with those reads.
template <class PL, int B, bool FOLD, bool CPLX = false>
__global__ void __launch_bounds__(PL::TG* B, 1) fused_fft_pp_kernel(const StagedArgs a, const FusedOut out, const int64_t K,
                                                                    const int chunk) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    using L = PPLayout<PL, B, FOLD>;
    constexpr int TG = PL::TG, V = PL::V, W = L::W;
    const int g = threadIdx.x / TG;       // group within the CTA
    const int t = threadIdx.x % TG;       // thread within the group
    const int warp_cta = threadIdx.x >> 5;
    float2* const buf = reinterpret_cast<float2*>(smem_raw + L::OFF_BUF) + (size_t)g * PL::BUF;
    float2* const stage = reinterpret_cast<float2*>(smem_raw + L::OFF_STAGE);
    unsigned int* const gmax = reinterpret_cast<unsigned int*>(smem_raw + L::OFF_GMAX) + 2 * g;
    unsigned int* const binmin = reinterpret_cast<unsigned int*>(smem_raw + L::OFF_BIN) + 2 * g;
    uint32_t* const taddr_slot = reinterpret_cast<uint32_t*>(smem_raw + L::OFF_TADDR);
    if (warp_cta == 0) tc::tmem_alloc(tc::smem_u32(taddr_slot), L::TMEM_COLS);
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
    const uint32_t tb = tmem0 + (((uint32_t)(warp_cta & 3) * 32u) << 16) + (uint32_t)((t >> 7) * L::CW);
    if (g < L::WRITERS) {
        // ---- per-thread constants -> tensor memory (read by the same in-group thread of every group) ----
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
        if constexpr (!FOLD) {
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const float2 w = (q < L::NQ1) ? unit_root(t + TG * q, PL::N) : make_float2(0.f, 0.f);
                r[2 * q] = w.x;
                r[2 * q + 1] = w.y;
            }
            tc::tmem_st16(tb + L::COL_TW1, r);
            const int d = t % PL::M2;
#pragma unroll
            for (int c = 0; c < PL::R2 / 8; ++c) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float2 w = unit_root((d * (8 * c + i)) % PL::M1, PL::M1);
                    r[2 * i] = w.x;
                    r[2 * i + 1] = w.y;
                }
                tc::tmem_st16(tb + L::COL_TW2 + 16 * c, r);
            }
        } else {
            // T1: pass-2 cell e = t + TG q = k1 M2 + d, input n2 carries W_N^((n2 M2 + d) k1)
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
            // T2: pass-3 row f = t + TG q = k1 R2 + k2, input d carries W_M1^(d k2), k2 = t % R2 for every q
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
    const int step = a.step;
    const int64_t s_first = a.shift0 + i0 * step;
    const GroupBar<TG, B> bar{1 + g};
    float2 v[V];
    float pend = 0.f;                                  // thread 0 of the group: QF^2 of the previous FFT
    float2 cand = make_float2(0.f, 0.f);               // CPLX: this thread's candidate for the previous FFT
    unsigned int cand_bin = 0xffffffffu;
    int o = W + 1;                                     // offset of the current B shifts inside the stage

    for (int it = 0; it < chunk; ++it) {
        const int e = B * it + g;
        if (o + (B - 1) * step > W) {
            // refill: stage[j] = rx[s + j], j < N + W (zero outside the record).  The first barrier makes sure no
            // group still reads the old window.
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
        // windows that leave the record still run (the stage is zero filled there); their results are dropped
        const bool ok = e < n_act && s >= 0 && s + PL::N <= a.rxlen;
        const float2* const xw = stage + o + g * step + t;
        o += B * step;
        // the barrier inside also orders the previous FFT's bin-minimum atomics before thread 0 reads them
        if constexpr (FOLD)
            pp_block_fft_fold<PL, B>(v, xw, tb, buf, t, bar);
        else
            pp_block_fft<PL, B>(v, xw, tb, buf, t, bar);
        unsigned int* const bslot = binmin + (CPLX ? (it & 1) : 0);          // this FFT's bin slot
        if (t == 0 && it > 0) {                        // results of the previous FFT of this group
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
                const float sc = 1.f / (sqrtf(*out.refnormsq) * sqrtf(out.energy[i0 + e - B]));
                out.cplx[i0 + e - B] = make_float2(cand.x * sc, cand.y * sc);
            }
            cand_bin = 0xffffffffu;
        }
        // ---- |X|^2 and the group maximum: |X|^2 >= +0, so the float order is the order of the bit patterns:
        // one REDUX per warp, one shared atomic per warp into the slot of this iteration's parity ----
#ifdef XC_PP_FLATSEARCH
        float tmax = 0.f;
#pragma unroll
        for (int j = 0; j < V; ++j) tmax = fmaxf(tmax, cnorm2(v[j]));
        const unsigned int u = __reduce_max_sync(0xffffffffu, __float_as_uint(tmax));
        if ((t & 31) == 0) atomicMax(gmax + (it & 1), u);
        bar();   // all last-pass loads done (buffer free) + group maximum complete
        const float m = __uint_as_float(gmax[it & 1]);
        if (t == 0) gmax[(it & 1) ^ 1] = 0u;
        if (ok && tmax == m) {
            unsigned int mybin = 0xffffffffu;
#pragma unroll
            for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                for (int j = 0; j < PL::RL; ++j)
                    if (cnorm2(v[q * PL::RL + j]) == m) mybin = min(mybin, (unsigned int)output_bin<PL>(t, q, j));
            atomicMin(bslot, mybin);
        }
#else
        // maxima of chunks of 8 registers first: the one thread that holds the group maximum then only searches the
        // chunk(s) that reach it (about 40 instead of 160 instructions on the critical path of its warp)
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
        if (CPLX && t == 0) binmin[(it & 1) ^ 1] = 0xffffffffu;       // read for the last time before this barrier
        if (ok && tmax == m) {
            unsigned int mybin = 0xffffffffu;
#pragma unroll
            for (int c = 0; c < V / 8; ++c) {
                if (m8[c] == m) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int j = 8 * c + i;
                        if (nv[j] == m) {
                            const unsigned int bj = (unsigned int)output_bin<PL>(t, j / PL::RL, j % PL::RL);
                            if (bj < mybin) {
                                mybin = bj;
                                if (CPLX) cand = v[j];
                            }
                        }
                    }
                }
            }
            atomicMin(bslot, mybin);
            if (CPLX) cand_bin = mybin;
        }
#endif
        if (t == 0) pend = (ok) ? m / *out.refnormsq / out.energy[i0 + e] : 0.f;
        if (CPLX && t == 0 && e < n_act && !ok) out.cplx[i0 + e] = make_float2(0.f, 0.f);
    }
    bar();
    if (chunk > 0) {                                   // the last FFT of this group
        const int p = B * (chunk - 1) + g;
        const unsigned int b = binmin[CPLX ? ((chunk - 1) & 1) : 0];
        if (t == 0 && p < n_act) {
            if (out.qf2) out.qf2[i0 + p] = pend;
            if (out.fidx) out.fidx[i0 + p] = (b == 0xffffffffu) ? 0u : b;
        }
        if constexpr (CPLX) {
            if (cand_bin != 0xffffffffu && cand_bin == b) {
                const float sc = 1.f / (sqrtf(*out.refnormsq) * sqrtf(out.energy[i0 + p]));
                out.cplx[i0 + p] = make_float2(cand.x * sc, cand.y * sc);
            }
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp_cta == 0) tc::tmem_dealloc(tmem0, L::TMEM_COLS);
}

template <class PL, int B, bool FOLD>
int32_t launch_fused_pp(const StagedArgs& a, const FusedOut& out, int64_t K, cudaStream_t stream) {
    auto kern = (FOLD && out.cplx) ? fused_fft_pp_kernel<PL, B, FOLD, FOLD> : fused_fft_pp_kernel<PL, B, FOLD, false>;
    constexpr size_t smem = PPLayout<PL, B, FOLD>::BYTES;
    if ((int)smem > max_smem_optin()) return fail(XC_ERR_UNSUPPORTED, "ping-pong fused kernel needs %zu B of shared memory", smem);
    static KernelOnce once[2];
    XC_TRY(once[(FOLD && out.cplx) ? 1 : 0].configure(kern, max_smem_optin()));
    const int64_t per = (K + B - 1) / B;               // iterations if one CTA did everything
    int64_t ctas = (int64_t)sm_count();
    if (ctas > per) ctas = per;
    if (ctas < 1) ctas = 1;
    const int64_t chunk = (per + ctas - 1) / ctas;
    if (chunk >= (1ll << 29)) return fail(XC_ERR_UNSUPPORTED, "too many shifts per thread group");
    profile_begin(stream);
    kern<<<(unsigned)ctas, PL::TG * B, smem, stream>>>(a, out, K, (int)chunk);
    profile_end(stream);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

}  // namespace xc
