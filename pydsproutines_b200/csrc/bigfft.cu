// bigfft.cu -- frequency search for cutouts that do not fit one block FFT (more than 16384 points, or 8192 when the
// length is not a power of two): the reference takes ANY length (xcorrRoutines.py:511-531, :98-151;
// IppXcorrFFT.cpp:94-178), and its only published numbers are at N = 1e6 and 1e7 (benchmarks/README.md:3-23).
//
// Four-step FFT over the block engine, L = R * C points per shift (R, C powers of two <= 16384, R = 2^floor(log2 L / 2)):
//   columns: n = n1 C + c; for every column c an R-point block FFT over n1, times W_L^(c k1), stored transposed
//            Y[k1][c] (the loader fuses the product rx[s + n] conj(cutout[n]) -- and the Bluestein pre-chirp);
//   rows:    for every k1 a C-point block FFT over c: X[k1 + R k2]; the epilogue is fused: |X|^2, per-row
//            (max, lowest bin) -> a per-shift reduction; or CAF rows; or the Bluestein middle step.
// Bluestein (any length N, and chirp-z grids): L >= N + bins - 1; second transform with the roles swapped
// (k = k2 R + k1: its "columns" are the contiguous rows the first transform left, no transpose pass in between):
//   columns(R; product) -> rows(C; times FFT_L(chirp), conj) -> columns(C) -> rows(R; |.|^2 for k < bins, arg max).
// Per shift the intermediate (8 L bytes) is written and read once per step; a batch of shifts takes about 1 GB of
// workspace.  This path is memory bound (about 7 x 8 L bytes per shift for Bluestein, 3 x for a
// power of two), unlike the fused single-SM kernels.
#include <math.h>

#include <map>
#include <vector>

#include "fused_fft_kernel.cuh"
#include "internal.cuh"
#include "plans.cuh"

namespace xc {

constexpr int BIG_MIN_LOG2 = 6;          // smallest block plan with an exchange buffer (64 points)
constexpr int BIG_MAX_LOG2L = 2 * MAX_LOG2N;
constexpr int TWB_BITS = 10;             // W_L^x = twA[x >> 10] * twB[x & 1023]
// CTAs per SM the kernels are compiled for: these passes wait on memory (ncu: long_scoreboard 7 per issue with one CTA of
// 172 registers per SM), so two CTAs of <= 128 registers beat one (benchmark_xcorrs 1e6: 9.1 -> 6.6 ms); three spill
#ifndef XC_BIG_MINB
#define XC_BIG_MINB 2
#endif
template <int THREADS>
constexpr int big_minb() { return THREADS <= 256 ? XC_BIG_MINB : 1; }

// ---------------------------------------------------------------------------------------------
// plans: twiddle tables per (device, log2 L); Bluestein tables per (device, N, bins, a, w)
// ---------------------------------------------------------------------------------------------
struct BigTw {
    float2* d_a = nullptr;   // [L >> 10] exp(-2 pi i (x << 10) / L)
    float2* d_b = nullptr;   // [1024]    exp(-2 pi i x / L)
};
static std::map<std::pair<int, int>, BigTw> g_bigtw;

static int32_t big_twiddles(int log2L, const BigTw** out) {
    int dev = 0;
    XC_CUDA_TRY(cudaGetDevice(&dev));
    auto key = std::make_pair(dev, log2L);
    auto it = g_bigtw.find(key);
    if (it != g_bigtw.end()) {
        *out = &it->second;
        return XC_OK;
    }
    const int64_t L = 1ll << log2L;
    const int64_t na = (L >> TWB_BITS) > 0 ? (L >> TWB_BITS) : 1, nb = 1 << TWB_BITS;
    std::vector<float2> ha((size_t)na), hb((size_t)nb);
    for (int64_t i = 0; i < na; ++i) {
        const double ang = -2.0 * M_PI * (double)((i << TWB_BITS) % L) / (double)L;
        ha[(size_t)i] = make_float2((float)cos(ang), (float)sin(ang));
    }
    for (int64_t i = 0; i < nb; ++i) {
        const double ang = -2.0 * M_PI * (double)(i % L) / (double)L;
        hb[(size_t)i] = make_float2((float)cos(ang), (float)sin(ang));
    }
    BigTw t;
    XC_CUDA_TRY(cudaMalloc(&t.d_a, sizeof(float2) * (size_t)na));
    XC_CUDA_TRY(cudaMalloc(&t.d_b, sizeof(float2) * (size_t)nb));
    XC_CUDA_TRY(cudaMemcpy(t.d_a, ha.data(), sizeof(float2) * (size_t)na, cudaMemcpyHostToDevice));
    XC_CUDA_TRY(cudaMemcpy(t.d_b, hb.data(), sizeof(float2) * (size_t)nb, cudaMemcpyHostToDevice));
    *out = &(g_bigtw[key] = t);
    return XC_OK;
}

// ---------------------------------------------------------------------------------------------
// column pass
// ---------------------------------------------------------------------------------------------
struct BigColsArgs {
    // input: MODE 0 = element (n, col) of batch entry b is in[b * in_batch + n * sn + col * sc];
    //        MODE 1 = product rx[s_b + idx] * cc[idx], idx = n * sn + col * sc, zero for idx >= nvalid
    const float2* in;
    int64_t in_batch, sn, sc;
    const float2* rx;
    int64_t rxlen;
    const float2* cc;          // [nvalid] conj(cutout) (times the Bluestein pre-chirp)
    int64_t nvalid;
    int64_t shift0, step;      // shift of list entry i: shifts ? shifts[i] : shift0 + i * step
    const int64_t* shifts;
    int64_t first;             // first shift-list entry of this batch
    int window;                // samples a valid window needs inside rx (the cutout length)
    float2* out;               // out[b * L + k * NC + col]
    int64_t L;
    int NC;                    // number of columns
    const float2* twa;
    const float2* twb;
    const float2* bp;          // MODE 2: FFT_L(chirp) in the register layout a PL-point row transform leaves ([row][slot][t])
};

// MODE 2 (Bluestein middle step fused with the second transform's column pass): the input "column" is a contiguous row of
// the first transform's column pass; it is transformed (PL points), multiplied by FFT_L(chirp), conjugated and brought back
// to input order through the exchange buffer, and then takes the ordinary path below (second PL-point transform, twiddle,
// transposed store).  Saves one write and one read of the 8 L bytes per shift against separate row / column passes.
template <class PL, int B, int MODE>
__global__ void __launch_bounds__(PL::TG* B, big_minb<PL::TG * B>()) big_cols_kernel(const BigColsArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int TG = PL::TG, V = PL::V;
    const int g = threadIdx.x / TG, t = threadIdx.x % TG;
    float2* const tab = reinterpret_cast<float2*>(smem_raw);
    build_fft_tables<PL>(tab, threadIdx.x, TG * B);
    float2* const buf = reinterpret_cast<float2*>(smem_raw + PL::TABLE_BYTES) + (size_t)g * PL::BUF;
    __syncthreads();
    FftThread<PL> th;
    th.init(t);
    const GroupBar<TG, B> bar{1 + g};
    const int b = blockIdx.y;
    const int ntiles = (a.NC + B - 1) / B;
    int64_t s = 0;
    bool in_range = true;
    if (MODE == 1) {
        s = a.shifts ? a.shifts[a.first + b] : a.shift0 + (a.first + b) * a.step;
        in_range = s >= 0 && s + a.window <= a.rxlen;
    }
    float2 v[V];
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int col = tile * B + g;
        const bool act = col < a.NC && in_range;
        constexpr int NQ = V / PL::R1;
        if (MODE == 1) {
            // slab load: the B adjacent columns of a row are 8 B contiguous bytes, so the CTA reads row by row with the
            // lanes along the columns (full sectors; one lane per row touched 32 sectors per load instruction and the
            // pass ran at the L1 tag rate) and hands the samples to the groups through their exchange buffers
            float2* const bufs = reinterpret_cast<float2*>(smem_raw + PL::TABLE_BYTES);
            const int ncol = (a.NC - tile * B < B) ? (a.NC - tile * B) : B;
            const float2* const xr = a.rx + s + (int64_t)tile * B;
            const float2* const cr = a.cc + (int64_t)tile * B;
#pragma unroll 4
            for (int i = threadIdx.x; i < PL::N * B; i += TG * B) {
                const int n = i / B, gg = i % B;
                const int64_t idx = (int64_t)n * a.sn + gg;
                float2 x = make_float2(0.f, 0.f);
                if (in_range && gg < ncol && (int64_t)tile * B + idx < a.nvalid) x = cmul(__ldg(xr + idx), __ldg(cr + idx));
                const int nr = (n + 2 * gg) & (PL::N - 1);       // rotated by 2 cells per column: the B stores of a row hit B banks
                bufs[(size_t)gg * PL::BUF + nr + ((nr >> 5) << 1)] = x;
            }
            __syncthreads();
#pragma unroll
            for (int n1 = 0; n1 < PL::R1; ++n1)
#pragma unroll
                for (int q = 0; q < NQ; ++q) {
                    const int n = (input_index<PL>(t, q, n1) + 2 * g) & (PL::N - 1);
                    v[q * PL::R1 + n1] = buf[n + ((n >> 5) << 1)];
                }
            bar();                                                 // reads done before the pass-1 stores of the transform
        } else {
#pragma unroll
        for (int n1 = 0; n1 < PL::R1; ++n1)
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                const int64_t n = input_index<PL>(t, q, n1);
                float2 x = make_float2(0.f, 0.f);
                if (act) {
                    const int64_t idx = n * a.sn + (int64_t)col * a.sc;
                    x = __ldg(a.in + (int64_t)b * a.in_batch + idx);
                }
                v[q * PL::R1 + n1] = x;
            }
        }
        if (MODE == 2) {
            block_fft<PL>(v, th, tab, buf, t, bar);
            const float2* const bpr = a.bp + (int64_t)(col < a.NC ? col : 0) * PL::N + t;
#pragma unroll
            for (int sl = 0; sl < V; ++sl) {
                const float2 p = cmul(v[sl], __ldg(bpr + sl * TG));
                v[sl] = make_float2(p.x, -p.y);
            }
            bar();                                                 // last-pass loads done: the buffer is free
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
                for (int q = 0; q < NQ; ++q) {
                    const int i = input_index<PL>(t, q, n1);
                    v[q * PL::R1 + n1] = buf[i + ((i >> 5) << 1)];
                }
            bar();
        }
        block_fft<PL>(v, th, tab, buf, t, bar);
        // ---- twiddle, then the transposed store through the exchange buffers: group g leaves its R outputs in natural
        // order in its own buffer (skewed by SK g cells against bank conflicts), and the whole CTA writes, for every k,
        // the B adjacent columns as one contiguous run of 8 B bytes (full 32-byte sectors instead of 8 bytes per lane at
        // a stride of NC samples) ----
        bar();                                                     // last-pass loads of the group are done
        constexpr int SK = (B > 1 && (PL::BUF - PL::N) / (B - 1) >= 4) ? 4 : ((B > 1 && (PL::BUF - PL::N) / (B - 1) >= 1) ? 1 : 0);
#pragma unroll
        for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
            for (int j = 0; j < PL::RL; ++j) {
                const int k = output_bin<PL>(t, q, j);
                const int64_t x = (int64_t)k * (col < a.NC ? col : 0);     // < L: no reduction needed
                const float2 w = cmul(__ldg(a.twa + (x >> TWB_BITS)), __ldg(a.twb + (x & ((1 << TWB_BITS) - 1))));
                buf[k + SK * g] = cmul(v[q * PL::RL + j], w);
            }
        __syncthreads();
        {
            const float2* const bufs = reinterpret_cast<const float2*>(smem_raw + PL::TABLE_BYTES);
            float2* const o = a.out + (int64_t)b * a.L + (int64_t)tile * B;
            const int ncol = (a.NC - tile * B < B) ? (a.NC - tile * B) : B;
            for (int i = threadIdx.x; i < PL::N * B; i += TG * B) {
                const int k = i / B, gg = i % B;
                if (gg < ncol) o[(int64_t)k * a.NC + gg] = bufs[(size_t)gg * PL::BUF + k + SK * gg];
            }
        }
        __syncthreads();                                           // buffers free for the next tile
    }
}

// ---------------------------------------------------------------------------------------------
// row pass with fused epilogues
// ---------------------------------------------------------------------------------------------
enum { EPI_RAWREG = 0, EPI_BLUE_MID = 1, EPI_ARGMAX = 2, EPI_CAF = 3 };

struct BigRowsArgs {
    const float2* in;          // in[b * L + row * LR + n]
    int64_t L;
    int NR;                    // number of rows; bin of (row, k) = row + NR * k
    float2* out;               // RAWREG: out[row * LR + slot * TG + t]; BLUE_MID: out[b * L + row * LR + k]
    const float2* bp;          // BLUE_MID: FFT_L(chirp) in the register layout RAWREG leaves
    unsigned long long* pmax;  // ARGMAX: [b * NR + row] packed (|X|^2 bits, ~bin)
    float2* pc;                // ARGMAX: [b * NR + row] X at that bin (NULL when not wanted)
    float* caf;                // CAF: [(first + b) * nbins + bin], float or float2
    int caf_complex;
    const float2* post;        // CAF complex: post-chirp of the Bluestein path (NULL for a plain FFT)
    const float* energy;       // [K] window energies, indexed by shift-list entry
    const float* refnormsq;
    int64_t first;
    int nbins;                 // bins >= nbins are ignored
    float inv_l2;              // 1 / L^2 for the Bluestein path (|.|^2 of the unnormalised inverse), else 1
};

template <class PL, int B, int EPI>
__global__ void __launch_bounds__(PL::TG* B, big_minb<PL::TG * B>()) big_rows_kernel(const BigRowsArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int TG = PL::TG, V = PL::V, LR = PL::N;
    const int g = threadIdx.x / TG, t = threadIdx.x % TG;
    float2* const tab = reinterpret_cast<float2*>(smem_raw);
    build_fft_tables<PL>(tab, threadIdx.x, TG * B);
    float2* const buf = reinterpret_cast<float2*>(smem_raw + PL::TABLE_BYTES) + (size_t)g * PL::BUF;
    unsigned long long* const gbest =
        reinterpret_cast<unsigned long long*>(smem_raw + PL::TABLE_BYTES + sizeof(float2) * (size_t)PL::BUF * B) + g;
    if (t == 0) *gbest = 0ull;
    __syncthreads();
    FftThread<PL> th;
    th.init(t);
    const GroupBar<TG, B> bar{1 + g};
    const int b = blockIdx.y;
    const int ntiles = (a.NR + B - 1) / B;
    float2 v[V];
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int row = tile * B + g;
        const bool act = row < a.NR;
        const float2* const src = a.in + (int64_t)b * a.L + (int64_t)row * LR;
        constexpr int NQ = V / PL::R1;
#pragma unroll
        for (int n1 = 0; n1 < PL::R1; ++n1)
#pragma unroll
            for (int q = 0; q < NQ; ++q)
                v[q * PL::R1 + n1] = act ? __ldg(src + input_index<PL>(t, q, n1)) : make_float2(0.f, 0.f);
        block_fft<PL>(v, th, tab, buf, t, bar);
        if (EPI == EPI_RAWREG) {
            if (act) {
#pragma unroll
                for (int sl = 0; sl < V; ++sl) a.out[(int64_t)row * LR + sl * TG + t] = v[sl];
            }
        } else if (EPI == EPI_BLUE_MID) {
            if (act) {
                float2* const o = a.out + (int64_t)b * a.L + (int64_t)row * LR;
#pragma unroll
                for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                    for (int j = 0; j < PL::RL; ++j) {
                        const int sl = q * PL::RL + j;
                        const float2 p = cmul(v[sl], __ldg(a.bp + (int64_t)row * LR + sl * TG + t));
                        o[output_bin<PL>(t, q, j)] = make_float2(p.x, -p.y);
                    }
            }
        } else if (EPI == EPI_CAF) {
            if (act) {
                const int64_t e = a.first + b;
                const float en = a.energy[e], refn = *a.refnormsq;
                const bool ok = en > 0.f;
                const float sc2 = ok ? a.inv_l2 / refn / en : 0.f;
                const float sc1 = ok ? sqrtf(a.inv_l2) / (sqrtf(refn) * sqrtf(en)) : 0.f;
#pragma unroll
                for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                    for (int j = 0; j < PL::RL; ++j) {
                        const int64_t bin = row + (int64_t)a.NR * output_bin<PL>(t, q, j);
                        if (bin < a.nbins) {
                            const float2 x = v[q * PL::RL + j];
                            if (a.caf_complex) {
                                float2 p = x;
                                if (a.post) {                      // X = conj(post * v) / L
                                    p = cmul(__ldg(a.post + bin), x);
                                    p.y = -p.y;
                                }
                                reinterpret_cast<float2*>(a.caf)[e * a.nbins + bin] = make_float2(p.x * sc1, p.y * sc1);
                            } else {
                                a.caf[e * a.nbins + bin] = cnorm2(x) * sc2;
                            }
                        }
                    }
            }
        } else {   // EPI_ARGMAX
            unsigned long long best = 0ull;
#pragma unroll
            for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                for (int j = 0; j < PL::RL; ++j) {
                    const int64_t bin = row + (int64_t)a.NR * output_bin<PL>(t, q, j);
                    if (act && bin < a.nbins) {
                        float m = cnorm2(v[q * PL::RL + j]);
                        if (!(m >= 0.f)) m = 0.f;
                        const unsigned long long p = pack_peak(m, 0xffffffffu - (uint32_t)bin);
                        best = p > best ? p : best;
                    }
                }
            atomicMax(gbest, best);
            bar();
            const unsigned long long gb = *gbest;
            if (act) {
                if (t == 0) a.pmax[(int64_t)b * a.NR + row] = gb;
                if (a.pc && best == gb && gb != 0ull) {
                    const int64_t wbin = (int64_t)(0xffffffffu - (uint32_t)(gb & 0xffffffffull));
#pragma unroll
                    for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                        for (int j = 0; j < PL::RL; ++j)
                            if (row + (int64_t)a.NR * output_bin<PL>(t, q, j) == wbin) a.pc[(int64_t)b * a.NR + row] = v[q * PL::RL + j];
                }
            }
            bar();
            if (t == 0) *gbest = 0ull;
        }
        bar();   // exchange buffer (and the group's slot) free for the next tile
    }
}

// per shift: winner over the rows -> qf2 / fidx / complex QF
__global__ void __launch_bounds__(256) big_finalize_kernel(const unsigned long long* __restrict__ pmax, const float2* __restrict__ pc,
                                                           int NR, int64_t first, int64_t shift0, int64_t step,
                                                           const int64_t* __restrict__ shifts, int window, int64_t rxlen,
                                                           const float* __restrict__ energy, const float* __restrict__ refnormsq,
                                                           float inv_l2, const float2* __restrict__ post, float* __restrict__ qf2,
                                                           uint32_t* __restrict__ fidx, float2* __restrict__ cplx) {
    __shared__ unsigned long long s_best[8];
    __shared__ int s_row[8];
    const int b = blockIdx.x;
    unsigned long long best = 0ull;
    int brow = 0;
    for (int r = threadIdx.x; r < NR; r += blockDim.x) {
        const unsigned long long p = pmax[(int64_t)b * NR + r];
        if (p > best) {
            best = p;
            brow = r;
        }
    }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
        const unsigned long long ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int orow = __shfl_xor_sync(0xffffffffu, brow, o);
        if (ob > best) {
            best = ob;
            brow = orow;
        }
    }
    if ((threadIdx.x & 31) == 0) {
        s_best[threadIdx.x >> 5] = best;
        s_row[threadIdx.x >> 5] = brow;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w)
            if (s_best[w] > best) {
                best = s_best[w];
                brow = s_row[w];
            }
        const int64_t e = first + b;
        const int64_t s = shifts ? shifts[e] : shift0 + e * step;
        const bool ok = s >= 0 && s + window <= rxlen;
        const float m = __uint_as_float((unsigned int)(best >> 32));
        const uint32_t bin = 0xffffffffu - (uint32_t)(best & 0xffffffffull);
        const float refn = *refnormsq;
        const float en = ok ? energy[e] : 1.f;
        if (qf2) qf2[e] = ok ? m * inv_l2 / refn / en : 0.f;
        if (fidx) fidx[e] = (ok && best != 0ull) ? bin : 0u;
        if (cplx) {
            float2 z = make_float2(0.f, 0.f);
            if (ok && best != 0ull) {
                float2 x = pc[(int64_t)b * NR + brow];
                if (post) {
                    x = cmul(post[bin], x);
                    x.y = -x.y;
                }
                const float sc = sqrtf(inv_l2) / (sqrtf(refn) * sqrtf(en));
                z = make_float2(x.x * sc, x.y * sc);
            }
            cplx[e] = z;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// launchers (dispatched on the block plan of the pass)
// ---------------------------------------------------------------------------------------------
template <int LOG2N>
struct BigGeom {
    using PL = typename PlanFor<LOG2N>::PL;
    // 4096- and 8192-point passes run 512 threads per CTA, so that the transposed store / slab load still sees 4 (2) adjacent
    // columns (benchmark_xcorrs 1e7, L = 4096 x 8192: 144 -> 122 ms)
    static constexpr int B = (PL::TG >= 512) ? 1 : (PL::TG >= 128) ? 512 / PL::TG : 256 / PL::TG;
    static constexpr size_t SMEM = PL::TABLE_BYTES + sizeof(float2) * (size_t)PL::BUF * B + sizeof(unsigned long long) * B + 16;
};

template <int MODE>
struct ColsRunner {
    const BigColsArgs& a;
    int batch;
    cudaStream_t stream;
    template <int LOG2N>
    int32_t run() {
        if constexpr (LOG2N >= BIG_MIN_LOG2) {
            using G = BigGeom<LOG2N>;
            auto kern = big_cols_kernel<typename G::PL, G::B, MODE>;
            static KernelOnce once;
            XC_TRY(once.configure(kern, max_smem_optin()));
            const int ntiles = (a.NC + G::B - 1) / G::B;
            int gx = ntiles;
            const int cap = sm_count() * 8;
            if ((int64_t)gx * batch > cap) gx = (cap + batch - 1) / batch;
            if (gx < 1) gx = 1;
            kern<<<dim3((unsigned)gx, (unsigned)batch), G::PL::TG * G::B, G::SMEM, stream>>>(a);
            XC_LAUNCH_CHECK();
            return XC_OK;
        } else {
            return fail(XC_ERR_UNSUPPORTED, "no column pass for 2^%d points", LOG2N);
        }
    }
};

template <int EPI>
struct RowsRunner {
    const BigRowsArgs& a;
    int batch;
    cudaStream_t stream;
    template <int LOG2N>
    int32_t run() {
        if constexpr (LOG2N >= BIG_MIN_LOG2) {
            using G = BigGeom<LOG2N>;
            auto kern = big_rows_kernel<typename G::PL, G::B, EPI>;
            static KernelOnce once;
            XC_TRY(once.configure(kern, max_smem_optin()));
            const int ntiles = (a.NR + G::B - 1) / G::B;
            int gx = ntiles;
            const int cap = sm_count() * 8;
            if ((int64_t)gx * batch > cap) gx = (cap + batch - 1) / batch;
            if (gx < 1) gx = 1;
            kern<<<dim3((unsigned)gx, (unsigned)batch), G::PL::TG * G::B, G::SMEM, stream>>>(a);
            XC_LAUNCH_CHECK();
            return XC_OK;
        } else {
            return fail(XC_ERR_UNSUPPORTED, "no row pass for 2^%d points", LOG2N);
        }
    }
};

// ---------------------------------------------------------------------------------------------
// Bluestein tables of the large path (device generated; phases in double, exact for the DFT case)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float2 big_chirp(double lin_cycles, double w, long long d, long long exact_n) {
    double frac;
    if (exact_n > 0) {
        const long long r = (long long)(((unsigned long long)(d < 0 ? -d : d) * (unsigned long long)(d < 0 ? -d : d)) % (unsigned long long)(2 * exact_n));
        frac = (double)r / (2.0 * (double)exact_n);
    } else {
        const double q = w * (double)d * (double)d * 0.5;
        frac = q - floor(q);
    }
    double sn, cs;
    sincospi(2.0 * (frac + (lin_cycles - floor(lin_cycles))), &sn, &cs);
    return make_float2((float)cs, (float)sn);
}
// pre[n] = exp(+2 pi i (a n + w n^2 / 2)), n < N; post[k] = exp(+2 pi i w k^2 / 2), k < nbins;
// ext[(d + L) % L] = exp(+2 pi i w d^2 / 2) for -(N - 1) <= d <= nbins - 1, 0 elsewhere
__global__ void big_chirp_kernel(float2* pre, float2* post, float2* ext, int64_t N, int64_t nbins, int64_t L, double a, double w,
                                 long long exact_n) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < L; i += stride) {
        if (i < N) pre[i] = big_chirp(a * (double)i, w, i, exact_n);
        if (i < nbins) post[i] = big_chirp(0.0, w, i, exact_n);
        float2 e = make_float2(0.f, 0.f);
        if (i <= nbins - 1) e = big_chirp(0.0, w, i, exact_n);
        else if (i >= L - (N - 1)) e = big_chirp(0.0, w, i - L, exact_n);
        ext[i] = e;
    }
}
// cc[n] = conj(cutout[n] * pre[n])  (pre == NULL: conj(cutout))
__global__ void big_cc_kernel(const float2* __restrict__ cutout, const float2* __restrict__ pre, int64_t N, float2* __restrict__ cc) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < N) {
        float2 c = cutout[i];
        if (pre) c = cmul(c, pre[i]);
        cc[i] = make_float2(c.x, -c.y);
    }
}

struct BigBlue {
    int64_t N = 0, nbins = 0;
    int log2L = 0;
    float2* d_pre = nullptr;    // [N]
    float2* d_post = nullptr;   // [nbins]
    float2* d_bp = nullptr;     // [L] FFT_L(ext) in the register layout of the C-point row pass
};
typedef std::tuple<int, int64_t, int64_t, double, double> BigBlueKey;
static std::map<BigBlueKey, BigBlue> g_bigblue;
static std::vector<BigBlueKey> g_bigblue_order;
constexpr size_t BIGBLUE_CACHE_MAX = 4;

static void bigblue_free(BigBlue& p) {
    if (p.d_pre) cudaFree(p.d_pre);
    if (p.d_post) cudaFree(p.d_post);
    if (p.d_bp) cudaFree(p.d_bp);
    p = BigBlue();
}

void release_big_plan_caches() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return;
    for (auto it = g_bigblue.begin(); it != g_bigblue.end();) {
        if (std::get<0>(it->first) == dev) {
            bigblue_free(it->second);
            it = g_bigblue.erase(it);
        } else {
            ++it;
        }
    }
    std::vector<BigBlueKey> keep;
    for (const BigBlueKey& k : g_bigblue_order)
        if (std::get<0>(k) != dev) keep.push_back(k);
    g_bigblue_order.swap(keep);
    for (auto it = g_bigtw.begin(); it != g_bigtw.end();) {
        if (it->first.first == dev) {
            cudaFree(it->second.d_a);
            cudaFree(it->second.d_b);
            it = g_bigtw.erase(it);
        } else {
            ++it;
        }
    }
}

static void big_split(int log2L, int* log2R, int* log2C) {
    *log2R = log2L / 2;
    *log2C = log2L - *log2R;
}

// one forward transform of `batch` arrays: columns(R) then rows(C)
template <int MODE, int EPI>
static int32_t big_forward(BigColsArgs ca, BigRowsArgs ra, int log2R, int log2C, int batch, cudaStream_t stream) {
    XC_TRY(dispatch_log2(log2R, ColsRunner<MODE>{ca, batch, stream}));
    XC_TRY(dispatch_log2(log2C, RowsRunner<EPI>{ra, batch, stream}));
    return XC_OK;
}

static int32_t bigblue_plan(int64_t N, const CztSpec* spec, cudaStream_t stream, const BigBlue** out) {
    int dev = 0;
    XC_CUDA_TRY(cudaGetDevice(&dev));
    const int64_t nbins = spec ? spec->nbins : N;
    const double a = spec ? spec->a_cycles : 0.0, w = spec ? spec->w_cycles : 1.0 / (double)N;
    const auto key = std::make_tuple(dev, N, nbins, a, w);
    auto it = g_bigblue.find(key);
    if (it != g_bigblue.end()) {
        *out = &it->second;
        return XC_OK;
    }
    BigBlue p;
    p.N = N;
    p.nbins = nbins;
    p.log2L = 2 * BIG_MIN_LOG2;
    while ((1ll << p.log2L) < N + nbins - 1) ++p.log2L;
    if (p.log2L > BIG_MAX_LOG2L) return fail(XC_ERR_UNSUPPORTED, "chirp-z of %lld samples onto %lld bins needs more than 2^%d points", (long long)N, (long long)nbins, BIG_MAX_LOG2L);
    const int64_t L = 1ll << p.log2L;
    int log2R, log2C;
    big_split(p.log2L, &log2R, &log2C);
    const BigTw* tw = nullptr;
    XC_TRY(big_twiddles(p.log2L, &tw));
    float2 *d_ext = nullptr, *d_y = nullptr;
    XC_CUDA_TRY(cudaMalloc(&p.d_pre, sizeof(float2) * (size_t)N));
    XC_CUDA_TRY(cudaMalloc(&p.d_post, sizeof(float2) * (size_t)nbins));
    XC_CUDA_TRY(cudaMalloc(&p.d_bp, sizeof(float2) * (size_t)L));
    XC_CUDA_TRY(cudaMalloc(&d_ext, sizeof(float2) * (size_t)L));
    XC_CUDA_TRY(cudaMalloc(&d_y, sizeof(float2) * (size_t)L));
    big_chirp_kernel<<<sm_count() * 8, 256, 0, stream>>>(p.d_pre, p.d_post, d_ext, N, nbins, L, a, w, spec ? 0 : (long long)N);
    XC_LAUNCH_CHECK();
    BigColsArgs ca{};
    ca.in = d_ext;
    ca.in_batch = L;
    ca.sn = 1ll << log2C;
    ca.sc = 1;
    ca.out = d_y;
    ca.L = L;
    ca.NC = 1 << log2C;
    ca.twa = tw->d_a;
    ca.twb = tw->d_b;
    BigRowsArgs ra{};
    ra.in = d_y;
    ra.L = L;
    ra.NR = 1 << log2R;
    ra.out = p.d_bp;
    XC_TRY((big_forward<0, EPI_RAWREG>(ca, ra, log2R, log2C, 1, stream)));
    XC_CUDA_TRY(cudaStreamSynchronize(stream));
    cudaFree(d_ext);
    cudaFree(d_y);
    if (g_bigblue_order.size() >= BIGBLUE_CACHE_MAX) {
        cudaDeviceSynchronize();
        auto old = g_bigblue.find(g_bigblue_order.front());
        if (old != g_bigblue.end()) {
            bigblue_free(old->second);
            g_bigblue.erase(old);
        }
        g_bigblue_order.erase(g_bigblue_order.begin());
    }
    g_bigblue_order.push_back(key);
    *out = &(g_bigblue[key] = p);
    return XC_OK;
}

// ---------------------------------------------------------------------------------------------
// entry: frequency search of one call through the four-step path
// ---------------------------------------------------------------------------------------------
bool bigfft_supported(int64_t N, const CztSpec* czt) {
    if (N < 2) return false;
    const int64_t nbins = czt ? czt->nbins : N;
    const bool pow2 = !czt && (N & (N - 1)) == 0;
    const int64_t need = pow2 ? N : N + nbins - 1;
    return need <= (1ll << BIG_MAX_LOG2L) && N < (1ll << 31) && nbins < (1ll << 31);
}

int32_t bigfft_freq_device(const float2* cutout, int64_t N, const float2* rx, int64_t rxlen, const ShiftList& sl,
                           const FusedOut& out, const CztSpec* czt, cudaStream_t stream) {
    const bool pow2 = !czt && (N & (N - 1)) == 0 && ilog2((int)N) >= 2 * BIG_MIN_LOG2;
    const BigBlue* blue = nullptr;
    int log2L;
    int64_t nbins;
    if (pow2) {
        log2L = ilog2((int)N);
        nbins = N;
    } else {
        XC_TRY(bigblue_plan(N, czt, stream, &blue));
        log2L = blue->log2L;
        nbins = blue->nbins;
    }
    if (log2L > BIG_MAX_LOG2L) return fail(XC_ERR_UNSUPPORTED, "frequency search needs 2^%d points, limit 2^%d", log2L, BIG_MAX_LOG2L);
    const int64_t L = 1ll << log2L;
    int log2R, log2C;
    big_split(log2L, &log2R, &log2C);
    const int R = 1 << log2R, C = 1 << log2C;
    const BigTw* tw = nullptr;
    XC_TRY(big_twiddles(log2L, &tw));
    // shifts per batch: about 1 GB of work arrays (one for a power of two, two for Bluestein).  Measured on the
    // reference's benchmark (N = 1e6, 128 shifts): 1 GB 9.9 ms, 200 MB 11.0, 96 MB (L2 resident) 12.5, 4 GB 11.2 --
    // the parallelism of a large batch is worth more than L2 residency.  A debug build takes XC_BIG_BATCH_MB.
    int64_t budget = 1024ll << 20;
#ifdef XC_EXPERIMENTS
    if (const char* e = getenv("XC_BIG_BATCH_MB")) budget = (int64_t)atoll(e) << 20;
#endif
    int64_t batch = budget / ((pow2 ? 1 : 2) * (int64_t)sizeof(float2) * L);
    if (batch < 1) batch = 1;
    if (batch > 16384) batch = 16384;
    if (batch > sl.K) batch = sl.K;
    float2* d_cc = reinterpret_cast<float2*>(workspace(SLOT_TMP2, sizeof(float2) * (size_t)N));
    const int NRmax = R > C ? R : C;
    // layout of the big slot: [Y: batch * L][Z: batch * L (Bluestein only)][pmax: batch * NRmax u64][pc: batch * NRmax]
    const size_t ybytes = sizeof(float2) * (size_t)batch * (size_t)L;
    const size_t pbytes = sizeof(unsigned long long) * (size_t)batch * (size_t)NRmax;
    unsigned char* big = reinterpret_cast<unsigned char*>(workspace(SLOT_BIG, ybytes * (pow2 ? 1 : 2) + 2 * pbytes + 256));
    if (!d_cc || !big) return XC_ERR_NOMEM;
    float2* d_y = reinterpret_cast<float2*>(big);
    float2* d_z = reinterpret_cast<float2*>(big + ybytes);
    unsigned long long* d_pmax = reinterpret_cast<unsigned long long*>(big + ybytes * (pow2 ? 1 : 2));
    float2* d_pc = reinterpret_cast<float2*>(big + ybytes * (pow2 ? 1 : 2) + pbytes);
    big_cc_kernel<<<(unsigned)((N + 255) / 256), 256, 0, stream>>>(cutout, blue ? blue->d_pre : nullptr, N, d_cc);
    XC_LAUNCH_CHECK();
    const float inv_l2 = pow2 ? 1.f : (float)(1.0 / ((double)L * (double)L));
    profile_begin(stream);
    for (int64_t first = 0; first < sl.K; first += batch) {
        const int nb = (int)((sl.K - first < batch) ? (sl.K - first) : batch);
        BigColsArgs ca{};
        ca.sn = C;
        ca.sc = 1;
        ca.rx = rx;
        ca.rxlen = rxlen;
        ca.cc = d_cc;
        ca.nvalid = N;
        ca.shift0 = sl.shift0;
        ca.step = sl.step;
        ca.shifts = sl.shifts;
        ca.first = first;
        ca.window = (int)N;
        ca.out = d_y;
        ca.L = L;
        ca.NC = C;
        ca.twa = tw->d_a;
        ca.twb = tw->d_b;
        BigRowsArgs ra{};
        ra.in = d_y;
        ra.L = L;
        ra.NR = R;
        ra.energy = out.energy;
        ra.refnormsq = out.refnormsq;
        ra.first = first;
        ra.nbins = (int)nbins;
        ra.inv_l2 = inv_l2;
        ra.caf = out.caf;
        ra.caf_complex = out.caf_complex;
        int NRfinal = R;
        if (pow2) {
            ra.pmax = d_pmax;
            ra.pc = out.cplx ? d_pc : nullptr;
            if (out.caf) XC_TRY((big_forward<1, EPI_CAF>(ca, ra, log2R, log2C, nb, stream)));
            else XC_TRY((big_forward<1, EPI_ARGMAX>(ca, ra, log2R, log2C, nb, stream)));
        } else {
            // first transform's column pass: product -> columns(R), transposed store Y[k1][c]
            XC_TRY(dispatch_log2(log2R, ColsRunner<1>{ca, nb, stream}));
            // its row pass (C points, times FFT_L(chirp), conjugated) fused with the column pass of the second transform
            // over k = k2 R + k1 (the same C-point rows): Y[k1][.] -> Z[j1][k1]
            BigColsArgs cb{};
            cb.in = d_y;
            cb.in_batch = L;
            cb.sn = 1;
            cb.sc = C;
            cb.out = d_z;
            cb.L = L;
            cb.NC = R;
            cb.twa = tw->d_a;
            cb.twb = tw->d_b;
            cb.bp = blue->d_bp;
            XC_TRY(dispatch_log2(log2C, ColsRunner<2>{cb, nb, stream}));
            BigRowsArgs rb = ra;
            rb.in = d_z;
            rb.NR = C;
            rb.out = nullptr;
            rb.bp = nullptr;
            rb.pmax = d_pmax;
            rb.pc = out.cplx ? d_pc : nullptr;
            rb.post = blue->d_post;
            NRfinal = C;
            if (out.caf) XC_TRY(dispatch_log2(log2R, RowsRunner<EPI_CAF>{rb, nb, stream}));
            else XC_TRY(dispatch_log2(log2R, RowsRunner<EPI_ARGMAX>{rb, nb, stream}));
        }
        if (!out.caf) {
            big_finalize_kernel<<<nb, 256, 0, stream>>>(d_pmax, out.cplx ? d_pc : nullptr, NRfinal, first, sl.shift0, sl.step,
                                                        sl.shifts, (int)N, rxlen, out.energy, out.refnormsq, inv_l2,
                                                        blue ? blue->d_post : nullptr, out.qf2, out.fidx, out.cplx);
            XC_LAUNCH_CHECK();
        }
    }
    profile_end(stream);
    return XC_OK;
}

}  // namespace xc
