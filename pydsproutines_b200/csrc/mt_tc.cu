// mt_tc.cu -- multi-template sliding correlator on the 5th-generation tensor cores (tcgen05 / TMEM).
//     d[s, t] = sum_n x[s + n] * tmpl_t[n],   QF^2[s] = max_t |d[s,t]|^2 / E_t / E_x[s],   arg max over t
// Replaces multiTemplateSlidingDotProduct (multiplySlices.cu:251-399, cupyExtensions.py:563-642) and the
// returnMax path of TemplateCrossCorrelator.correlate (xcorrRoutines.py:339-371).  This is the one place of
// the detection path where the work is a dense complex contraction (8 L flops per (template, shift) cell).
//
// Formulation (one CTA = one SM, persistent over tiles of 128 consecutive shifts):
//   * GEMM roles: M = 128 shifts (TMEM lanes), N = up to 256 templates (TMEM columns), K = template samples.
//     Two FP32 accumulators in TMEM, C_re and C_im, 256 columns each = all 512 columns.
//       C_re += Xr Tr^T - Xi Ti^T        C_im += Xi Tr^T + Xr Ti^T       (the minus is the A-negate bit)
//   * precision: every x plane is split into BF16 hi + lo (16 mantissa bits) and the accumulation is FP32.
//     Templates that are exact in BF16 (any PSK alphabet) need 8 MMAs (128 x NT x 16) per 16 samples:
//     (x_hi + x_lo) t.  Other template sets are split too, t = t_hi + t_lo, and take 12: the t_lo part against
//     x_hi only (the x_lo t_lo term is below 2^-17 of the product).
//   * A operand = the Hankel matrix X[s, n] = x[s0 + s + n], never materialised: shared memory holds the
//     8x-expanded tile E[u][i][j] = x[s0 + 8u + i + j] (one 128-byte core matrix per u), and the core matrix
//     of (row group a, K chunk b) is E[a + b], so the UMMA descriptor simply uses LBO = SBO = 128 bytes --
//     overlapping core matrices.  (128 + L) / 8 core matrices per plane: 18 KB for L = 1024.
//   * B operand = templates, repacked once per handle into stage images in canonical K-major core-matrix
//     order, streamed by 1-D bulk copies (TMA engine) through a ring of stages, 16 samples per stage.
//   * warp roles: warps 0-3 build E of the NEXT tile (double buffered) while the MMAs of the current one run,
//     then run the epilogue (TMEM -> registers, |.|^2 / E_t, running max and lowest arg max, no cross-thread
//     reduction because a thread owns a shift); warp 4 lane 0 = bulk-copy producer, warp 5 lane 0 = MMA
//     issuer.  No CTA-wide barrier in the loop: stage full/empty, E full/empty and TMEM full/empty mbarriers.
#include <vector>

#include "internal.cuh"
#include "mt_tc.cuh"

namespace xc {

namespace {

constexpr int TC_M = 128;              // shifts per tile
constexpr int TC_KC = 32;              // template samples per pipeline stage (two K = 16 steps)
constexpr int TC_THREADS = 192;        // 4 epilogue/builder warps + producer warp + MMA warp
constexpr int TC_MAX_STAGES = 4;
constexpr int TC_TMEM_COLS = 512;
constexpr int TC_MAX_TEMPLATES = 2048; // 1/E_t table in shared memory

__host__ __device__ inline size_t tc_plane_bytes(int L32) { return (size_t)((TC_M + L32) / 8) * 128; }
__host__ __device__ inline size_t tc_stage_bytes(int NT) { return (size_t)NT * 128; }   // re + im planes of 32 samples

// ---- one-time repack of the templates into stage images; flags inexact BF16 conversions ----
__global__ void mt_tc_pack_kernel(const float2* __restrict__ templates, int T, int L, int L8, int L32, int NT, int NB,
                                  int NP, __nv_bfloat16* __restrict__ pack, int* __restrict__ inexact) {
    const int NKC = L32 / TC_KC;
    const size_t total = (size_t)NB * NKC * NP * NT * TC_KC;      // bf16 elements
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        // e = ((((nb * NKC + kc) * NP + p) * 4 + c8) * (NT / 8) + tg) * 64 + r * 8 + j
        // plane p: bit 0 = re / im, bit 1 = BF16 hi / lo part (NP = 4 only)
        size_t q = e;
        const int j = q % 8; q /= 8;
        const int r = q % 8; q /= 8;
        const int tg = q % (NT / 8); q /= (NT / 8);
        const int c8 = q % 4; q /= 4;
        const int p = q % NP; q /= NP;
        const int kc = q % NKC; q /= NKC;
        const int nb = (int)q;
        const int t = nb * NT + tg * 8 + r;
        const int n = kc * TC_KC + c8 * 8 + j;
        float v = 0.f;
        if (t < T && n < L) {
            const float2 s = templates[(size_t)t * L8 + n];
            v = (p & 1) ? s.y : s.x;
        }
        const __nv_bfloat16 hi = __float2bfloat16_rn(v);
        if (__bfloat162float(hi) != v) atomicOr(inexact, 1);
        pack[e] = (p & 2) ? __float2bfloat16_rn(v - __bfloat162float(hi)) : hi;
    }
}

__global__ void mt_tc_inv_energy_kernel(const float* __restrict__ energies, int T, int n, float* __restrict__ inv) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) inv[i] = (i < T && energies[i] > 0.f) ? 1.f / energies[i] : 0.f;
}


// shared-memory carve-up (bytes from the 128-byte aligned base)
constexpr int TC_SUB = 16;             // template samples per pipeline stage = one K step (half a stage image)
constexpr int TC_NBAR = 2 * TC_MAX_STAGES + 2 + 2 + 2;   // stage full/empty, E full/empty (x2), TMEM full/empty
struct TcSmem {
    uint32_t plane, stage, off_stage, off_inv, off_bar, total;
    int nstages, ne;
};
// Lp = longest contiguous run of template samples one x tile must cover (the whole template for a dense set,
// the longest group of a segmented one)
__host__ __device__ inline TcSmem tc_smem_layout(int Lp, int NT, int NB, int nstages, int ne, int np) {
    TcSmem m;
    m.plane = (uint32_t)tc_plane_bytes(Lp);
    m.stage = (uint32_t)NT * 32 * (uint32_t)np;        // np planes (re, im[, re_lo, im_lo]) of 16 samples
    m.nstages = nstages;
    m.ne = ne;
    m.off_stage = (uint32_t)ne * 4 * m.plane;
    m.off_inv = m.off_stage + (uint32_t)nstages * m.stage;
    m.off_bar = (m.off_inv + (uint32_t)(NB * NT) * 4 + 15) / 16 * 16;
    m.total = m.off_bar + TC_NBAR * 8 + 16;
    return m;
}

__global__ void __launch_bounds__(TC_THREADS, 1)
mt_tc_kernel(const uint8_t* __restrict__ pack, const float* __restrict__ inv_energy, const int NT, const int NB,
             const int L32, const int Lp, const int nseg, const int32_t* __restrict__ segs, const int nstages, const int ne,
             const int np, const float2* __restrict__ x, const int64_t xlen, const int64_t start, const int64_t K,
             const float* __restrict__ xenergy, float* __restrict__ best_out, int32_t* __restrict__ tidx_out,
             float* __restrict__ caf_out, float2* __restrict__ cplx_out, const int caf_cols, const int sqrt_out,
             const int seg_per_flush, float* __restrict__ scratch, const int64_t ntiles,
             const int64_t tiles_per_cta) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const TcSmem lay = tc_smem_layout(Lp, NT, NB, nstages, ne, np);
    const int NKS = L32 / TC_SUB;          // K steps (= pipeline stages consumed) per template block
    uint8_t* const sE = smem;
    float* const sInv = reinterpret_cast<float*>(smem + lay.off_inv);
    uint64_t* const bars = reinterpret_cast<uint64_t*>(smem + lay.off_bar);
    const uint32_t bar_full = tc::smem_u32(bars), bar_empty = tc::smem_u32(bars + TC_MAX_STAGES),
                   bar_efull = tc::smem_u32(bars + 2 * TC_MAX_STAGES), bar_eempty = tc::smem_u32(bars + 2 * TC_MAX_STAGES + 2),
                   bar_tfull = tc::smem_u32(bars + 2 * TC_MAX_STAGES + 4), bar_tempty = tc::smem_u32(bars + 2 * TC_MAX_STAGES + 5);
    uint32_t* const tmem_slot = reinterpret_cast<uint32_t*>(bars + TC_NBAR);
    const uint32_t sE_u32 = tc::smem_u32(sE), sStage_u32 = tc::smem_u32(smem + lay.off_stage);

    if (tid == 0) {
        for (int s = 0; s < TC_MAX_STAGES; ++s) {
            tc::mbar_init(bar_full + 8 * s, 1);
            tc::mbar_init(bar_empty + 8 * s, 1);
        }
        for (int b = 0; b < 2; ++b) {
            tc::mbar_init(bar_efull + 8 * b, 128);     // every builder thread arrives
            tc::mbar_init(bar_eempty + 8 * b, 1);      // tcgen05.commit of the last MMA that reads the buffer
        }
        tc::mbar_init(bar_tfull, 1);                   // tcgen05.commit of the last MMA of an accumulator generation
        tc::mbar_init(bar_tempty, 128);                // every epilogue thread arrives
        tc::mbar_fence_init();
    }
    if (warp == 0) tc::tmem_alloc(tc::smem_u32(tmem_slot), TC_TMEM_COLS);
    for (int i = tid; i < NB * NT; i += TC_THREADS) sInv[i] = inv_energy[i];
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    const int64_t tile0 = (int64_t)blockIdx.x * tiles_per_cta;
    int64_t tile1 = tile0 + tiles_per_cta;
    if (tile1 > ntiles) tile1 = ntiles;
    const int ntl = (tile1 > tile0) ? (int)(tile1 - tile0) : 0;     // tiles of this CTA
    const uint32_t idesc_pos = tc::instr_desc_bf16(TC_M, NT, 0), idesc_neg = tc::instr_desc_bf16(TC_M, NT, 1);
    const uint32_t b_lbo = (uint32_t)(NT / 8) * 128;

    if (warp < 4) {
        // ===== warps 0-3: build the expanded x tile of the NEXT tile while the MMAs of this one run, then the
        // epilogue of this tile.  thread = one shift (TMEM lane), columns = templates =====
        // x tile jobs: a dense template set has ONE per tile of shifts (shared by all template blocks); a segmented
        // set (templates that are non-zero only inside G groups, segs = {x offset, K steps} per group) has one per
        // (tile, template block, group), consumed in that order
        const int njobs = (nseg > 1) ? ntl * NB * nseg : ntl;
        auto build = [&](int job) {
            const int b = job % ne;
            const uint32_t use = (uint32_t)(job / ne);
            int til = job, xoff = 0, nks = NKS;
            if (nseg > 1) {
                const int sg = job % nseg;
                til = (job / nseg) / NB;
                xoff = __ldg(segs + 2 * sg);
                nks = __ldg(segs + 2 * sg + 1);
            }
            tc::mbar_wait(bar_eempty + 8 * b, (use & 1u) ^ 1u);      // the MMAs that read this buffer are done
            const int64_t s0 = start + (tile0 + til) * TC_M + xoff;  // x index of (row 0, n = 0)
            uint8_t* const E = sE + (size_t)b * 4 * lay.plane;
            // row (u, i) = BF16 hi / lo of x[s0 + 8u + i + j], j = 0..7, for re and im
            const int rows = TC_M + TC_SUB * nks;                     // = 8 U
            for (int row = tid; row < rows; row += 128) {
                const int64_t g = s0 + row;
                uint32_t w[4][4];                                     // [plane][pair of bf16]
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) {
                    float2 a = make_float2(0.f, 0.f), c = make_float2(0.f, 0.f);
                    if (g + 2 * jj < xlen) a = __ldg(x + g + 2 * jj);
                    if (g + 2 * jj + 1 < xlen) c = __ldg(x + g + 2 * jj + 1);
                    const __nv_bfloat16 arh = __float2bfloat16_rn(a.x), aih = __float2bfloat16_rn(a.y);
                    const __nv_bfloat16 crh = __float2bfloat16_rn(c.x), cih = __float2bfloat16_rn(c.y);
                    const __nv_bfloat16 arl = __float2bfloat16_rn(a.x - __bfloat162float(arh));
                    const __nv_bfloat16 ail = __float2bfloat16_rn(a.y - __bfloat162float(aih));
                    const __nv_bfloat16 crl = __float2bfloat16_rn(c.x - __bfloat162float(crh));
                    const __nv_bfloat16 cil = __float2bfloat16_rn(c.y - __bfloat162float(cih));
                    auto pk = [](__nv_bfloat16 lo, __nv_bfloat16 hi) {
                        return (uint32_t)__bfloat16_as_ushort(lo) | ((uint32_t)__bfloat16_as_ushort(hi) << 16);
                    };
                    w[0][jj] = pk(arh, crh);
                    w[1][jj] = pk(arl, crl);
                    w[2][jj] = pk(aih, cih);
                    w[3][jj] = pk(ail, cil);
                }
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    *reinterpret_cast<uint4*>(E + (size_t)q * lay.plane + (size_t)row * 16) =
                        make_uint4(w[q][0], w[q][1], w[q][2], w[q][3]);
            }
            tc::fence_proxy_async();                                  // generic-proxy stores -> tensor-core reads
            tc::mbar_arrive(bar_efull + 8 * b);
        };
        int built = 0;
        auto ensure = [&](int upto) {
            if (upto > njobs) upto = njobs;
            while (built < upto) build(built++);
        };
        ensure(1);
        uint32_t gen = 0;                                             // accumulator generations
        for (int i = 0; i < ntl; ++i) {
            float bestv = -1.f;
            int besti = 0;
            const int64_t o = (tile0 + i) * TC_M + warp * 32 + lane;
            const float inv_ex = (o < K) ? 1.f / xenergy[o] : 0.f;
            for (int nb = 0; nb < NB; ++nb) {
                // Long contractions are accumulated in chunks: the tensor core truncates the FP32 accumulator on every
                // read-modify-write (a bias of ~3e-8 of the sum per MMA), so after seg_per_flush segments the partial
                // sums leave TMEM and continue in an L2-resident FP32 scratch ([re|im][column][row] per CTA).
                const int nchunks = (nseg > 1 && seg_per_flush > 0) ? (nseg + seg_per_flush - 1) / seg_per_flush : 1;
                float* const sc_re = scratch + (size_t)blockIdx.x * 2 * NT * TC_M + (warp * 32 + lane);
                float* const sc_im = sc_re + (size_t)NT * TC_M;
                int sg = 0;
                for (int ch = 0; ch < nchunks; ++ch) {
                    // keep the tile builder one job ahead of the MMAs
                    if (nseg > 1) {
                        const int sg_end = (nchunks == 1) ? nseg : min(nseg, (ch + 1) * seg_per_flush);
                        for (; sg < sg_end; ++sg) ensure((i * NB + nb) * nseg + sg + 2);
                    } else if (nb == 0) {
                        ensure(i + 2);
                    }
                    const bool first = (ch == 0), last = (ch == nchunks - 1);
                    tc::mbar_wait(bar_tfull, gen & 1u);
                    tc::tc_fence_after();
                    const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
                    const float* inv = sInv + nb * NT;
                    for (int c = 0; c < NT / 16; ++c) {
                        float re[16], im[16];
                        tc::tmem_ld16(lane_base + 16 * c, re);
                        tc::tmem_ld16(lane_base + 256 + 16 * c, im);
                        tc::tmem_ld_wait();
                        if (!first) {
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                re[j] += sc_re[(size_t)(16 * c + j) * TC_M];
                                im[j] += sc_im[(size_t)(16 * c + j) * TC_M];
                            }
                        }
                        if (!last) {
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                sc_re[(size_t)(16 * c + j) * TC_M] = re[j];
                                sc_im[(size_t)(16 * c + j) * TC_M] = im[j];
                            }
                            continue;
                        }
                        if (cplx_out && o < K) {
                            const int col0 = nb * NT + 16 * c;
                            if (sqrt_out & 2) {
                                // normalised complex QF, template-major (T, K): TemplateCrossCorrelator.correlate
                                // (returnMax=False), xcorrRoutines.py:352-362; a warp writes 32 consecutive shifts per template
                                const float sx = sqrtf(inv_ex);
#pragma unroll
                                for (int j = 0; j < 16; ++j)
                                    if (col0 + j < caf_cols) {
                                        const float sc = sqrtf(inv[16 * c + j]) * sx;
                                        cplx_out[(size_t)(col0 + j) * (size_t)K + (size_t)o] = make_float2(re[j] * sc, im[j] * sc);
                                    }
                            } else {               // raw complex contraction d[s, t], row = shift
                                float2* dst = cplx_out + (size_t)o * (size_t)caf_cols + col0;
#pragma unroll
                                for (int j = 0; j < 16; ++j)
                                    if (col0 + j < caf_cols) dst[j] = make_float2(re[j], im[j]);
                            }
                        }
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            const float q = fmaf(re[j], re[j], im[j] * im[j]) * inv[16 * c + j];
                            if (q > bestv) {       // strict '>' keeps the lowest template on ties
                                bestv = q;
                                besti = nb * NT + 16 * c + j;
                            }
                            re[j] = q * inv_ex;
                        }
                        if (caf_out && o < K) {    // full surface: row = shift, column = template
                            const int col0 = nb * NT + 16 * c;
                            float* dst = caf_out + (size_t)o * (size_t)caf_cols + col0;
#pragma unroll
                            for (int j = 0; j < 16; ++j)
                                if (col0 + j < caf_cols) dst[j] = re[j];
                        }
                    }
                    tc::tc_fence_before();
                    tc::mbar_arrive(bar_tempty);                      // TMEM may be overwritten
                    ++gen;
                }
            }
            if (o < K) {
                const float q = bestv * inv_ex;
                if (best_out) best_out[o] = (sqrt_out & 1) ? sqrtf(q) : q;
                if (tidx_out) tidx_out[o] = besti;
            }
        }
    } else if (warp == 4) {
        if (lane == 0) {
            // ===== producer: half a stage image (16 samples of every plane) per stage =====
            uint32_t it = 0;
            const uint32_t part = (uint32_t)NT * 32;
            const size_t image = (size_t)NT * 64 * (size_t)np;       // 32 samples, all planes
            for (int i = 0; i < ntl; ++i)
                for (int nb = 0; nb < NB; ++nb) {
                    const uint8_t* src = pack + (size_t)nb * (NKS / 2) * image;
                    for (int ks = 0; ks < NKS; ++ks) {
                        const uint32_t st = it % (uint32_t)nstages, ph = (it / (uint32_t)nstages) & 1u;
                        tc::mbar_wait(bar_empty + 8 * st, ph ^ 1u);
                        tc::mbar_expect_tx(bar_full + 8 * st, (uint32_t)np * part);
                        const uint8_t* img = src + (size_t)(ks >> 1) * image + (size_t)(ks & 1) * part;
                        const uint32_t dst = sStage_u32 + st * lay.stage;
                        for (int p2 = 0; p2 < np; ++p2)
                            tc::bulk_g2s(dst + (uint32_t)p2 * part, img + (size_t)p2 * NT * 64, part, bar_full + 8 * st);
                        ++it;
                    }
                }
        }
    } else {
        if (lane == 0) {
            // ===== MMA issuer =====
            uint32_t it = 0, gen = 0;
            const uint32_t c_re = tmem, c_im = tmem + 256;
            int job = 0;
            for (int i = 0; i < ntl; ++i) {
                for (int nb = 0; nb < NB; ++nb) {
                    const int per = (nseg > 1 && seg_per_flush > 0) ? seg_per_flush : nseg;     // segments per accumulator generation
                    bool first = true;
                    for (int sg = 0; sg < nseg; ++sg) {
                        if (sg % per == 0) {
                            tc::mbar_wait(bar_tempty, (gen & 1u) ^ 1u);       // the epilogue has drained the accumulators
                            tc::tc_fence_after();
                            first = true;
                        }
                        const int b = job % ne;
                        const uint32_t use = (uint32_t)(job / ne);
                        if (nseg > 1 || nb == 0) tc::mbar_wait(bar_efull + 8 * b, use & 1u);
                        tc::tc_fence_after();
                        const uint32_t eb = sE_u32 + (uint32_t)b * 4 * lay.plane;
                        const int nks = (nseg > 1) ? __ldg(segs + 2 * sg + 1) : NKS;
                        for (int ks = 0; ks < nks; ++ks) {
                            const uint32_t st = it % (uint32_t)nstages, ph = (it / (uint32_t)nstages) & 1u;
                            tc::mbar_wait(bar_full + 8 * st, ph);
                            tc::tc_fence_after();
                            const uint32_t sb = sStage_u32 + st * lay.stage;
                            const uint32_t acc = first ? 0u : 1u;
                            first = false;
                            uint64_t a[4], bd[4];
#pragma unroll
                            for (int q = 0; q < 4; ++q)
                                a[q] = tc::smem_desc(eb + (uint32_t)q * lay.plane + 2 * (uint32_t)ks * 128, 128, 128);
#pragma unroll
                            for (int p2 = 0; p2 < 4; ++p2) bd[p2] = tc::smem_desc(sb + (uint32_t)p2 * (uint32_t)NT * 32, b_lbo, 128);
                            tc::umma_bf16(c_re, a[0], bd[0], idesc_pos, acc);
                            tc::umma_bf16(c_re, a[1], bd[0], idesc_pos, 1u);
                            tc::umma_bf16(c_re, a[2], bd[1], idesc_neg, 1u);
                            tc::umma_bf16(c_re, a[3], bd[1], idesc_neg, 1u);
                            tc::umma_bf16(c_im, a[2], bd[0], idesc_pos, acc);
                            tc::umma_bf16(c_im, a[3], bd[0], idesc_pos, 1u);
                            tc::umma_bf16(c_im, a[0], bd[1], idesc_pos, 1u);
                            tc::umma_bf16(c_im, a[1], bd[1], idesc_pos, 1u);
                            if (np == 4) {      // templates split into BF16 hi + lo: the lo part against x_hi
                                tc::umma_bf16(c_re, a[0], bd[2], idesc_pos, 1u);
                                tc::umma_bf16(c_re, a[2], bd[3], idesc_neg, 1u);
                                tc::umma_bf16(c_im, a[2], bd[2], idesc_pos, 1u);
                                tc::umma_bf16(c_im, a[0], bd[3], idesc_pos, 1u);
                            }
                            tc::umma_commit(bar_empty + 8 * st);      // stage free once these MMAs have read it
                            ++it;
                        }
                        if (nseg > 1 || nb == NB - 1) {
                            tc::umma_commit(bar_eempty + 8 * b);      // x tile buffer free
                            ++job;
                        }
                        if ((sg + 1) % per == 0 || sg == nseg - 1) {
                            tc::umma_commit(bar_tfull);               // accumulators of this generation complete
                            ++gen;
                        }
                    }
                }
            }
            if (job > 0)      // every commit has landed before the CTA tears down
                tc::mbar_wait(bar_eempty + 8 * ((job - 1) % ne), (uint32_t)((job - 1) / ne) & 1u);
        }
    }
    tc::tc_fence_before();
    __syncwarp();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, TC_TMEM_COLS);
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
// number of x-tile buffers (2 = the build of the next tile overlaps the MMAs) and of pipeline stages that fit
static int tc_pick_stages(int Lp, int NT, int NB, int np, int* ne_out) {
    for (int ne = 2; ne >= 1; --ne)
        for (int s = TC_MAX_STAGES; s >= (ne == 2 ? 3 : 2); --s)
            if ((int)tc_smem_layout(Lp, NT, NB, s, ne, np).total + 128 <= max_smem_optin()) {
                *ne_out = ne;
                return s;
            }
    *ne_out = 0;
    return 0;
}

void mt_tc_release(MtTcPlan* plan) {
    if (!plan) return;
    cudaFree(plan->d_pack);
    cudaFree(plan->d_inv_energy);
    cudaFree(plan->d_segs);
    plan->d_segs = nullptr;
    plan->d_pack = nullptr;
    plan->d_inv_energy = nullptr;
    plan->usable = false;
}

// Builds the BF16 stage images of a template set already on the device ([T, L8] float2, conjugated as the
// caller asked).  plan->usable stays false (and the caller keeps the FP32 kernel) when there are too many
// templates for the shared-memory energy table or when the expanded x tile plus two stages do not fit in
// shared memory.  plan->planes = 2 when every template value is exact in BF16, 4 (hi + lo split) otherwise.
// Segmented sets (segs != NULL, nseg pairs {x offset, K steps of 16 samples}): the template rows hold the
// concatenated group samples, every group padded to a multiple of 16, and L8 >= 16 * sum(K steps) rounded up to 32.
int32_t mt_tc_prepare(const float2* d_templates, int T, int L, int L8, const float* d_energies, MtTcPlan* plan,
                      const int32_t* segs, int nseg) {
    plan->usable = false;
    plan->T = T;
    plan->L = L;
    plan->L32 = (L + TC_KC - 1) / TC_KC * TC_KC;
    plan->Lp = plan->L32;
    plan->nseg = 1;
    plan->d_segs = nullptr;
    if (segs && nseg > 1) {
        int lp = 0, total = 0;
        for (int g = 0; g < nseg; ++g) {
            if (segs[2 * g + 1] * TC_SUB > lp) lp = segs[2 * g + 1] * TC_SUB;
            total += segs[2 * g + 1];
        }
        if (total * TC_SUB != plan->L32) return fail(XC_ERR_ARG, "segment K steps do not add up to the padded template length");
        plan->Lp = (lp + TC_KC - 1) / TC_KC * TC_KC;
        plan->nseg = nseg;
    }
    plan->NT = (T >= 256) ? 256 : (T + 15) / 16 * 16;
    plan->NB = (T + plan->NT - 1) / plan->NT;
    if (plan->NB * plan->NT > TC_MAX_TEMPLATES) return XC_OK;
    int ne_probe = 0;
    if (tc_pick_stages(plan->Lp, plan->NT, plan->NB, 4, &ne_probe) < 2) return XC_OK;
    int cc_major = 0;
    int dev = 0;
    XC_CUDA_TRY(cudaGetDevice(&dev));
    XC_CUDA_TRY(cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, dev));
    if (cc_major != 10) return XC_OK;
    int* d_flag = reinterpret_cast<int*>(workspace(SLOT_TMP3, 256));
    if (!d_flag) return XC_ERR_NOMEM;
    XC_CUDA_TRY(cudaMalloc(&plan->d_inv_energy, sizeof(float) * (size_t)plan->NB * plan->NT));
    const int n = plan->NB * plan->NT;
    mt_tc_inv_energy_kernel<<<(n + 255) / 256, 256>>>(d_energies, T, n, plan->d_inv_energy);
    XC_LAUNCH_CHECK();
    // first try the two-plane images (templates exact in BF16); when a value does not survive the conversion,
    // repack with the BF16 hi + lo split of the templates (four planes, a third precision pass)
    for (int np = 2; np <= 4; np += 2) {
        const size_t elems = (size_t)plan->NB * (plan->L32 / TC_KC) * np * plan->NT * TC_KC;
        XC_CUDA_TRY(cudaMemset(d_flag, 0, sizeof(int)));
        XC_CUDA_TRY(cudaMalloc(&plan->d_pack, elems * sizeof(__nv_bfloat16)));
        const int blocks = (int)((elems + 255) / 256 < 4096 ? (elems + 255) / 256 : 4096);
        mt_tc_pack_kernel<<<blocks, 256>>>(d_templates, T, L, L8, plan->L32, plan->NT, plan->NB, np,
                                           reinterpret_cast<__nv_bfloat16*>(plan->d_pack), d_flag);
        XC_LAUNCH_CHECK();
        int inexact = 0;
        XC_CUDA_TRY(cudaMemcpy(&inexact, d_flag, sizeof(int), cudaMemcpyDeviceToHost));
        plan->planes = np;
        if (!inexact || np == 4) break;
        cudaFree(plan->d_pack);
        plan->d_pack = nullptr;
    }
    if (plan->nseg > 1) {
        XC_CUDA_TRY(cudaMalloc(&plan->d_segs, sizeof(int32_t) * 2 * (size_t)nseg));
        XC_CUDA_TRY(cudaMemcpy(plan->d_segs, segs, sizeof(int32_t) * 2 * (size_t)nseg, cudaMemcpyHostToDevice));
    }
    plan->usable = true;
    return XC_OK;
}

int32_t mt_tc_run(const MtTcPlan& plan, const float2* x, int64_t xlen, int64_t start, int64_t K, const float* xenergy,
                  float* best, int32_t* tidx, int sqrt_out, cudaStream_t stream, float* caf, float2* cplx) {
    if (!plan.usable) return fail(XC_ERR_UNSUPPORTED, "tensor-core template plan is not usable");
    int ne = 0;
    const int nstages = tc_pick_stages(plan.Lp, plan.NT, plan.NB, plan.planes, &ne);
    const TcSmem lay = tc_smem_layout(plan.Lp, plan.NT, plan.NB, nstages, ne, plan.planes);
    const size_t smem = lay.total + 128;
    if ((int)smem > max_smem_optin()) return fail(XC_ERR_UNSUPPORTED, "tensor-core template kernel needs %zu B of shared memory", smem);
    static KernelOnce once;
    XC_TRY(once.configure(mt_tc_kernel, max_smem_optin()));
    const int64_t ntiles = (K + TC_M - 1) / TC_M;
    int64_t ctas = sm_count();
    if (ctas > ntiles) ctas = ntiles;
    const int64_t per = (ntiles + ctas - 1) / ctas;
    ctas = (ntiles + per - 1) / per;
    // accumulator generations of at most ~64 K steps for long segmented contractions (see the kernel's epilogue)
    int seg_per_flush = 0;
    float* scratch = nullptr;
    const int rmw = (plan.planes == 4) ? 6 : 4;                     // accumulator read-modify-writes per K step
    if (plan.nseg > 1 && (plan.L32 / TC_SUB) * rmw > 768) {
#ifdef XC_EXPERIMENTS
        static const int flush_ks = [] { const char* e = getenv("XC_TC_FLUSH_KS"); return e ? atoi(e) : 64; }();
#else
        constexpr int flush_ks = 64;          // K steps per accumulator generation (tools/prec_groupfreq.py)
#endif
        seg_per_flush = flush_ks / (plan.Lp / TC_SUB);
        if (seg_per_flush < 1) seg_per_flush = 1;
        scratch = reinterpret_cast<float*>(workspace(SLOT_TMP3, sizeof(float) * 2 * (size_t)plan.NT * TC_M * (size_t)ctas));
        if (!scratch) return XC_ERR_NOMEM;
    }
    profile_begin(stream);
    mt_tc_kernel<<<(unsigned)ctas, TC_THREADS, smem, stream>>>(reinterpret_cast<const uint8_t*>(plan.d_pack),
                                                               plan.d_inv_energy, plan.NT, plan.NB, plan.L32, plan.Lp, plan.nseg,
                                                               plan.d_segs, nstages, ne, plan.planes, x, xlen, start, K, xenergy,
                                                               best, tidx, caf, cplx, plan.T, sqrt_out, seg_per_flush, scratch, ntiles, per);
    profile_end(stream);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

}  // namespace xc
