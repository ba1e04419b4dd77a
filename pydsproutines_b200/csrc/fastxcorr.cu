// fastxcorr.cu -- fastXcorr on the B200.
//   freqsearch: fused product -> block FFT -> |X|^2 -> normalise -> arg max (fused_fft_kernel.cuh),
//               replacing the copy/multiply/cuFFT/argmax chain of cp_fastXcorr (xcorrRoutines.py:98-151)
//               and IppXcorrFFT_32fc::xcorr_thread (IppXcorrFFT.cpp:94-178);
//   no freqsearch: register-tiled sliding dot product (xcorrRoutines.py:483-509).
#include <math.h>

#include <map>
#include <tuple>
#include <vector>

#include "bluestein.cuh"
#include "ovs.cuh"
#include "fused_fft_kernel.cuh"
#include "fused_fft_pp.cuh"
#ifdef XC_EXPERIMENTS   // earlier layouts of the N = 8192 search, kept for A/B runs only (tools/staged_check.py)
#include "fused_fft_dual.cuh"
#include "fused_fft_staged.cuh"
#endif
#include "internal.cuh"
#include "plans.cuh"

namespace xc {

// search_pp.cu: the ping-pong layout of the N = 8192 search (fused_fft_pp.cuh); fold = 1 is the shipped variant
int32_t search_pp_device(const StagedArgs& a, const FusedOut& out, int64_t K, int N, int fold, cudaStream_t stream);
bool search_pp_takes(int N);
bool search_blue_pp_takes(int M);
int32_t search_blue_pp_device(const StagedArgs& a, const float2* bhat, const float2* post, int nvalid, int nbins, const FusedOut& out,
                              int64_t K, int M, cudaStream_t stream);
constexpr int PP_MAX_STEP = 16;       // (groups per SM - 1) * step stays well inside the 2048 samples of window slack

// ---------------------------------------------------------------------------------------------
// loader of the fused kernel: v[n] = rx[s + n] * conj(cutout[n])
// ---------------------------------------------------------------------------------------------
template <class PL, bool CUT_SMEM>
struct CutoutLoader {
    const float2* rx;
    int64_t rxlen;
    const float2* cutout;
    int64_t shift0, step;
    const int64_t* shifts;

    __host__ __device__ static constexpr size_t smem_bytes() { return CUT_SMEM ? sizeof(float2) * PL::N : 0; }
    size_t dyn_smem() const { return 0; }

    // conj(cutout) into shared memory
    __device__ __forceinline__ void prepare(int tid, int nthreads, unsigned char* smem) const {
        if (CUT_SMEM) {
            float2* cc = reinterpret_cast<float2*>(smem);
            for (int n = tid; n < PL::N; n += nthreads) {
                const float2 c = cutout[n];
                cc[n] = make_float2(c.x, -c.y);
            }
        }
    }
    // global loads of the window of shift-list entry i (false when it leaves rx)
    __device__ __forceinline__ bool issue(float2* v, int t, int64_t i, const unsigned char*) const {
        const int64_t s = shifts ? shifts[i] : shift0 + i * step;
        if (s < 0 || s + PL::N > rxlen) return false;
        const float2* x = rx + s + t;
        constexpr int NQ = PL::V / PL::R1;
#pragma unroll
        for (int n1 = 0; n1 < PL::R1; ++n1)
#pragma unroll
            for (int q = 0; q < NQ; ++q) v[q * PL::R1 + n1] = __ldg(x + input_index<PL>(0, q, n1));
        return true;
    }
    // product with conj(cutout)
    __device__ __forceinline__ void finish(float2* v, int t, const unsigned char* smem) const {
        constexpr int NQ = PL::V / PL::R1;
        if (CUT_SMEM) {
            const float2* cc = reinterpret_cast<const float2*>(smem) + t;
#pragma unroll
            for (int n1 = 0; n1 < PL::R1; ++n1) {
#pragma unroll
                for (int q = 0; q < NQ; ++q)
                    v[q * PL::R1 + n1] = pcmul(v[q * PL::R1 + n1], cc[input_index<PL>(0, q, n1)]);
                // at most 8 cutout values in flight: the 32 points already fill half the register file
                if (((n1 + 1) * NQ) % 8 == 0) asm volatile("" ::: "memory");
            }
        } else {
            const float2* cg = cutout + t;
#pragma unroll
            for (int n1 = 0; n1 < PL::R1; ++n1)
#pragma unroll
                for (int q = 0; q < NQ; ++q)
                    v[q * PL::R1 + n1] = pcmulc(v[q * PL::R1 + n1], __ldg(cg + input_index<PL>(0, q, n1)));
        }
    }
    // two windows at once (fused_fft_dual_kernel): every conj(cutout) value is read once for both products
    __device__ __forceinline__ void finish2(float2* v0, float2* v1, int t, const unsigned char* smem) const {
        constexpr int NQ = PL::V / PL::R1;
        if (CUT_SMEM) {
            const float2* cc = reinterpret_cast<const float2*>(smem) + t;
#pragma unroll
            for (int n1 = 0; n1 < PL::R1; ++n1) {
#pragma unroll
                for (int q = 0; q < NQ; ++q) {
                    const float2 c = cc[input_index<PL>(0, q, n1)];
                    v0[q * PL::R1 + n1] = pcmul(v0[q * PL::R1 + n1], c);
                    v1[q * PL::R1 + n1] = pcmul(v1[q * PL::R1 + n1], c);
                }
                if (((n1 + 1) * NQ) % 8 == 0) asm volatile("" ::: "memory");
            }
        } else {
            finish(v0, t, smem);
            finish(v1, t, smem);
        }
    }
    __device__ __forceinline__ bool load(float2* v, int t, int64_t i, const unsigned char* smem) const {
        if (!issue(v, t, i, smem)) return false;
        finish(v, t, smem);
        return true;
    }
};

// CutoutLoader with conj(cutout) in TENSOR MEMORY (one group of 512 threads: four warps per lane quarter, 64 columns each):
// for 16384 points the cutout is 128 KB, fits neither shared memory beside the exchange buffer nor the L1, and was re-read from
// the L2 for every shift.
template <class PL>
struct CutoutLoaderTmem {
    static_assert(PL::V == 32 && PL::TG % 128 == 0 && PL::TG <= 1024, "32 points per thread");
    static constexpr bool USES_TMEM = true;
    static constexpr int COPIES = PL::TG / 128;
    static constexpr int COLS = (64 * COPIES <= 64) ? 64 : (64 * COPIES <= 128) ? 128 : (64 * COPIES <= 256) ? 256 : 512;
    const float2* rx;
    int64_t rxlen;
    const float2* cutout;
    int64_t shift0, step;
    const int64_t* shifts;
    int nvalid = PL::N;            // Bluestein: only the first nvalid points of the padded window exist in rx

    __host__ __device__ static constexpr size_t smem_bytes() { return 16; }
    size_t dyn_smem() const { return 0; }
    __device__ __forceinline__ void prepare(int tid, int, unsigned char* smem) const {
        if (tid < 32) tc::tmem_alloc(tc::smem_u32(smem), COLS);
        tc::tc_fence_before();
    }
    __device__ __forceinline__ uint32_t base(int t, const unsigned char* smem) const {
        return *reinterpret_cast<const uint32_t*>(smem) + ((((uint32_t)(t >> 5) & 3u) * 32u) << 16) + (uint32_t)((t >> 7) * 64);
    }
    __device__ __forceinline__ void prepare2(int t, unsigned char* smem) const {      // one group per CTA: t = thread in the group
        tc::tc_fence_after();
        const uint32_t tb = base(t, smem);
        float r[16];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int j = 8 * c + i;
                const float2 cv = cutout[input_index<PL>(t, j / PL::R1, j % PL::R1)];
                r[2 * i] = cv.x;
                r[2 * i + 1] = -cv.y;
            }
            tc::tmem_st16(tb + 16 * c, r);
        }
        tc::tmem_st_wait();
        tc::tc_fence_before();
    }
    __device__ __forceinline__ void release(int tid, unsigned char* smem) const {
        tc::tc_fence_after();
        if (tid < 32) tc::tmem_dealloc(*reinterpret_cast<const uint32_t*>(smem), COLS);
    }
    __device__ __forceinline__ bool load(float2* v, int t, int64_t i, const unsigned char* smem) const {
        const int64_t s = shifts ? shifts[i] : shift0 + i * step;
        if (s < 0 || s + nvalid > rxlen) return false;
        const float2* x = rx + s + t;
        constexpr int NQ = PL::V / PL::R1;
        uint32_t cu[2][16];
        const uint32_t tb = base(t, smem);
        tc::tmem_ld16_issue(tb, cu[0]);
#pragma unroll
        for (int n1 = 0; n1 < PL::R1; ++n1)
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                const int n = input_index<PL>(0, q, n1);
                v[q * PL::R1 + n1] = (n + t < nvalid) ? __ldg(x + n) : make_float2(0.f, 0.f);
            }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            tc::tmem_ld_wait16(cu[c & 1]);
            if (c + 1 < 4) tc::tmem_ld16_issue(tb + 16 * (c + 1), cu[(c + 1) & 1]);
#pragma unroll
            for (int i = 0; i < 8; ++i)
                v[8 * c + i] = pcmul(v[8 * c + i], make_float2(__uint_as_float(cu[c & 1][2 * i]), __uint_as_float(cu[c & 1][2 * i + 1])));
        }
        return true;
    }
};

// loader of the Bluestein path: like CutoutLoader over the padded length M = PL::N, but only the first `nvalid`
// points exist in rx (the rest of the chirped product is zero) and `cutout` is cutout[n] * b[n], zero padded.
template <class PL, bool CUT_SMEM>
struct ChirpLoader {
    const float2* rx;
    int64_t rxlen;
    const float2* cutout;
    int64_t shift0, step;
    const int64_t* shifts;
    int nvalid;

    __host__ __device__ static constexpr size_t smem_bytes() { return CUT_SMEM ? sizeof(float2) * PL::N : 0; }
    size_t dyn_smem() const { return 0; }
    __device__ __forceinline__ void prepare(int tid, int nthreads, unsigned char* smem) const {
        if (CUT_SMEM) {
            float2* cc = reinterpret_cast<float2*>(smem);
            for (int n = tid; n < PL::N; n += nthreads) {
                const float2 c = cutout[n];
                cc[n] = make_float2(c.x, -c.y);
            }
        }
    }
    __device__ __forceinline__ bool load(float2* v, int t, int64_t i, const unsigned char* smem) const {
        const int64_t s = shifts ? shifts[i] : shift0 + i * step;
        if (s < 0 || s + nvalid > rxlen) return false;
        const float2* x = rx + s + t;
        constexpr int NQ = PL::V / PL::R1;
#pragma unroll
        for (int n1 = 0; n1 < PL::R1; ++n1)
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                const int n = input_index<PL>(0, q, n1);
                v[q * PL::R1 + n1] = (n + t < nvalid) ? __ldg(x + n) : make_float2(0.f, 0.f);
            }
        if (CUT_SMEM) {
            const float2* cc = reinterpret_cast<const float2*>(smem) + t;
#pragma unroll
            for (int n1 = 0; n1 < PL::R1; ++n1)
#pragma unroll
                for (int q = 0; q < NQ; ++q)
                    v[q * PL::R1 + n1] = pcmul(v[q * PL::R1 + n1], cc[input_index<PL>(0, q, n1)]);
        } else {
            const float2* cg = cutout + t;
#pragma unroll
            for (int n1 = 0; n1 < PL::R1; ++n1)
#pragma unroll
                for (int q = 0; q < NQ; ++q)
                    v[q * PL::R1 + n1] = pcmulc(v[q * PL::R1 + n1], __ldg(cg + input_index<PL>(0, q, n1)));
        }
        return true;
    }
};

constexpr int BLUE_MIN_LOG2M = 6;     // smallest plan with an exchange buffer

static int bluestein_log2m(int N) {
    int lg = BLUE_MIN_LOG2M;
    while ((1 << lg) < 2 * N - 1) ++lg;
    return lg;
}

// power-of-two lengths run the plain fused kernel; every other length 2 <= N <= 2^(MAX_LOG2N - 1) the Bluestein one
bool fft_length_pow2(int N) { return fft_length_supported(N); }
// one block FFT per shift (powers of two up to 2^14, Bluestein up to 2^13 - 1 points) or the four-step path of bigfft.cu
static bool fused_length(int N) {
    if (fft_length_supported(N)) return true;
    return N >= 3 && N <= (1 << (MAX_LOG2N - 1)) && bluestein_log2m(N) <= MAX_LOG2N;
}
bool fastxcorr_fft_supported(int N) { return fused_length(N) || bigfft_supported(N, nullptr); }

// Layout of the N = 8192 search.  The product always runs LAYOUT_PP; a build with -DXC_EXPERIMENTS keeps the
// earlier layouts selectable through XC_LAYOUT=b2|dual|staged|ppchain|pp for A/B runs (tools/staged_check.py).
enum { LAYOUT_B2 = 0, LAYOUT_DUAL, LAYOUT_STAGED, LAYOUT_PP_CHAIN, LAYOUT_PP };
static int search_layout() {
#ifdef XC_EXPERIMENTS
    static const int layout = [] {
        const char* e = getenv("XC_LAYOUT");
        if (!e) return (int)LAYOUT_PP;
        if (!strcmp(e, "b2")) return (int)LAYOUT_B2;
        if (!strcmp(e, "dual")) return (int)LAYOUT_DUAL;
        if (!strcmp(e, "staged")) return (int)LAYOUT_STAGED;
        if (!strcmp(e, "ppchain")) return (int)LAYOUT_PP_CHAIN;
        return (int)LAYOUT_PP;
    }();
    return layout;
#else
    return LAYOUT_PP;
#endif
}

struct FreqRunner {
    const float2* cutout;
    const float2* rx;
    int64_t rxlen;
    const ShiftList& sl;
    const FusedOut& out;
    cudaStream_t stream;
    template <int LOG2N>
    int32_t run() {
        using PF = PlanFor<LOG2N>;
        using LD = CutoutLoader<typename PF::PL, PF::CS>;
        LD ld{rx, rxlen, cutout, sl.shift0, sl.step, sl.shifts};
        if constexpr (LOG2N >= 7 && LOG2N <= 13) {
            // contiguous shift range, (max, arg max) outputs, 128 ... 8192 points (at 64 points the round-1 layout is faster): free-running thread groups per SM
            // over one staged rx window, constants in tensor memory (fused_fft_pp.cuh).  Short ranges, shift lists with
            // a step or a gather table and the CAF / complex outputs stay on the B-groups-per-CTA layout below.
            const int layout = search_layout();
            const bool plain = !out.caf;          // (max, arg max) and, optionally, the complex QF at the arg max
            // a shift range with a small step runs from the same staged window (it is refilled more often)
            const bool contiguous = !sl.shifts && sl.step >= 1 && sl.step <= PP_MAX_STEP && sl.K >= 16 * (int64_t)sm_count();
            if (layout == LAYOUT_PP && plain && contiguous)
                return search_pp_device(StagedArgs{rx, rxlen, cutout, sl.shift0, (int)sl.step}, out, sl.K, 1 << LOG2N, 1, stream);
#ifdef XC_EXPERIMENTS
            if constexpr (LOG2N == 13) {
                if (layout == LAYOUT_PP_CHAIN && plain && !out.cplx && contiguous)
                    return search_pp_device(StagedArgs{rx, rxlen, cutout, sl.shift0}, out, sl.K, 1 << LOG2N, 0, stream);
                if (layout == LAYOUT_STAGED && plain && !out.cplx && contiguous)
                    return launch_fused_staged<typename PF::PL>(StagedArgs{rx, rxlen, cutout, sl.shift0}, out, sl.K, stream);
                if (layout == LAYOUT_DUAL && plain && !out.cplx) return launch_fused_dual<typename PF::PL, LD>(ld, out, sl.K, stream);
            }
#endif
        }
#ifndef XC_NO_TMEM_CUTOUT
        if constexpr (LOG2N == 14) {
            // (max, arg max) searches at 16384 points: conj(cutout) from tensor memory instead of the L2
            if (!out.caf && !out.cplx) {
                using LT = CutoutLoaderTmem<typename PF::PL>;
                static_assert(PF::B == 1, "one group per CTA");
                return launch_fused<typename PF::PL, 1, LT>(LT{rx, rxlen, cutout, sl.shift0, sl.step, sl.shifts}, out, sl.K, stream);
            }
        }
#endif
        return launch_fused<typename PF::PL, PF::B, LD>(ld, out, sl.K, stream);
    }
};

// ---------------------------------------------------------------------------------------------
// Bluestein plan (per cutout length, cached for the life of the process): the chirp b[n] and FFT_M(b) laid
// out for the block FFT's output registers
// ---------------------------------------------------------------------------------------------
struct cd {
    double re, im;
};

struct BluePlan {
    int N = 0, nbins = 0, log2m = 0;
    float2* d_b = nullptr;          // [M]: pre-chirp conj(aa[n]) = exp(+i 2 pi (a n + w n^2 / 2)) for n < N, 0 beyond
    float2* d_bhat_perm = nullptr;  // [M]: FFT_M(W^(-d^2/2) wrapped)[output_bin(t, slot)] at slot * TG + t
    float2* d_bhat_nat = nullptr;   // [M]: the same in natural bin order (the ping-pong layout lays it out itself)
    float2* d_post = nullptr;       // [nbins]: exp(+i 2 pi w k^2 / 2); X[k] = conj(d_post[k] v) / M
};
// keyed by (device, length, bins, start frequency a, frequency step w) -- a, w in cycles per sample
typedef std::tuple<int, int, int, double, double> BlueKey;
static std::map<BlueKey, BluePlan> g_blue;
static std::vector<BlueKey> g_blue_order;         // insertion order: the cache keeps the BLUE_CACHE_MAX newest plans
constexpr size_t BLUE_CACHE_MAX = 16;

static void blue_free(BluePlan& p) {
    if (p.d_b) cudaFree(p.d_b);
    if (p.d_bhat_perm) cudaFree(p.d_bhat_perm);
    if (p.d_bhat_nat) cudaFree(p.d_bhat_nat);
    if (p.d_post) cudaFree(p.d_post);
    p = BluePlan();
}
// xc_release_workspace(): drops the plans of the current device (the device is idle: the caller synchronised)
void release_plan_caches() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return;
    for (auto it = g_blue.begin(); it != g_blue.end();) {
        if (std::get<0>(it->first) == dev) {
            blue_free(it->second);
            it = g_blue.erase(it);
        } else {
            ++it;
        }
    }
    std::vector<BlueKey> keep;
    for (const BlueKey& k : g_blue_order)
        if (std::get<0>(k) != dev) keep.push_back(k);
    g_blue_order.swap(keep);
    release_big_plan_caches();
}

// exp(+i 2 pi (lin + w d^2 / 2)); the DFT case (w = 1/N) reduces d^2 modulo 2N exactly
static cd chirp_phase(double lin_cycles, double w, long long d, int exact_n) {
    double frac;
    if (exact_n > 0) {
        const long long r = (d * d) % (2LL * exact_n);
        frac = (double)r / (2.0 * (double)exact_n);
    } else {
        const long double q = (long double)w * (long double)d * (long double)d / 2.0L;
        frac = (double)(q - floorl(q));
    }
    const double ang = 2.0 * M_PI * (frac + (lin_cycles - floor(lin_cycles)));
    return cd{cos(ang), sin(ang)};
}

static void fft_host_radix2(std::vector<cd>& a) {     // forward DFT, in place, length a power of two
    const size_t n = a.size();
    for (size_t i = 1, j = 0; i < n; ++i) {
        size_t bit = n >> 1;
        for (; j & bit; bit >>= 1) j ^= bit;
        j ^= bit;
        if (i < j) std::swap(a[i], a[j]);
    }
    for (size_t len = 2; len <= n; len <<= 1) {
        for (size_t i = 0; i < n; i += len)
            for (size_t k = 0; k < len / 2; ++k) {
                const double ang = -2.0 * M_PI * (double)k / (double)len;
                const cd w{cos(ang), sin(ang)};
                const cd u = a[i + k], x = a[i + k + len / 2];
                const cd vv{x.re * w.re - x.im * w.im, x.re * w.im + x.im * w.re};
                a[i + k] = cd{u.re + vv.re, u.im + vv.im};
                a[i + k + len / 2] = cd{u.re - vv.re, u.im - vv.im};
            }
    }
}

struct BluePermute {
    const std::vector<cd>& bhat;
    std::vector<float2>& perm;
    template <int LOG2N>
    int32_t run() {
        using PL = typename PlanFor<LOG2N>::PL;
        for (int t = 0; t < PL::TG; ++t)
            for (int q = 0; q < PL::V / PL::RL; ++q)
                for (int j = 0; j < PL::RL; ++j) {
                    const int bin = output_bin_base<PL>(t) + output_bin_off<PL>(q, j);
                    perm[(size_t)(q * PL::RL + j) * PL::TG + t] = make_float2((float)bhat[bin].re, (float)bhat[bin].im);
                }
        return XC_OK;
    }
};

// plan for X[k] = sum_n p[n] exp(-i 2 pi (a + w k) n), k < nbins (chirp-z).  The DFT of a length that is not a
// power of two is nbins = N, a = 0, w = 1/N.
static int czt_log2m(int N, int nbins) {
    int lg = BLUE_MIN_LOG2M;
    while ((1 << lg) < N + nbins - 1) ++lg;
    return lg;
}
static int32_t blue_plan_for(int N, const CztSpec* spec, const BluePlan** out) {
    int dev = 0;
    XC_CUDA_TRY(cudaGetDevice(&dev));
    const int nbins = spec ? spec->nbins : N;
    const double a = spec ? spec->a_cycles : 0.0, w = spec ? spec->w_cycles : 1.0 / (double)N;
    const int exact_n = spec ? 0 : N;
    const auto key = std::make_tuple(dev, N, nbins, a, w);
    auto it = g_blue.find(key);
    if (it != g_blue.end()) {
        *out = &it->second;
        return XC_OK;
    }
    BluePlan p;
    p.N = N;
    p.nbins = nbins;
    p.log2m = czt_log2m(N, nbins);
    if (p.log2m > MAX_LOG2N)
        return fail(XC_ERR_UNSUPPORTED, "chirp-z of %d samples onto %d bins needs more than %d points", N, nbins, 1 << MAX_LOG2N);
    const int M = 1 << p.log2m;
    std::vector<cd> ext(M, cd{0.0, 0.0});
    std::vector<float2> hb(M, make_float2(0.f, 0.f)), perm(M), post(nbins);
    for (int n = 0; n < N; ++n) {
        const cd pre = chirp_phase(a * (double)n, w, n, exact_n);
        hb[n] = make_float2((float)pre.re, (float)pre.im);
    }
    for (int d = -(N - 1); d <= nbins - 1; ++d) ext[(d + M) % M] = chirp_phase(0.0, w, d, exact_n);
    for (int k = 0; k < nbins; ++k) {
        const cd q = chirp_phase(0.0, w, k, exact_n);
        post[k] = make_float2((float)q.re, (float)q.im);
    }
    fft_host_radix2(ext);
    XC_TRY(dispatch_log2(p.log2m, BluePermute{ext, perm}));
    std::vector<float2> nat(M);
    for (int i = 0; i < M; ++i) nat[i] = make_float2((float)ext[i].re, (float)ext[i].im);
    XC_CUDA_TRY(cudaMalloc(&p.d_b, sizeof(float2) * M));
    XC_CUDA_TRY(cudaMalloc(&p.d_bhat_perm, sizeof(float2) * M));
    XC_CUDA_TRY(cudaMalloc(&p.d_post, sizeof(float2) * nbins));
    XC_CUDA_TRY(cudaMalloc(&p.d_bhat_nat, sizeof(float2) * M));
    XC_CUDA_TRY(cudaMemcpy(p.d_bhat_nat, nat.data(), sizeof(float2) * M, cudaMemcpyHostToDevice));
    XC_CUDA_TRY(cudaMemcpy(p.d_b, hb.data(), sizeof(float2) * M, cudaMemcpyHostToDevice));
    XC_CUDA_TRY(cudaMemcpy(p.d_bhat_perm, perm.data(), sizeof(float2) * M, cudaMemcpyHostToDevice));
    XC_CUDA_TRY(cudaMemcpy(p.d_post, post.data(), sizeof(float2) * nbins, cudaMemcpyHostToDevice));
    if (g_blue_order.size() >= BLUE_CACHE_MAX) {
        // callers that move the chirp-z window per detection would otherwise grow the cache without bound
        cudaDeviceSynchronize();              // the oldest plan may still be read by enqueued kernels
        auto old = g_blue.find(g_blue_order.front());
        if (old != g_blue.end()) {
            blue_free(old->second);
            g_blue.erase(old);
        }
        g_blue_order.erase(g_blue_order.begin());
    }
    g_blue_order.push_back(key);
    *out = &(g_blue[key] = p);
    return XC_OK;
}

int32_t bluestein_tables(int N, int* log2m, const float2** d_bhat_perm, const float2** d_chirp) {
    if (N < 3 || bluestein_log2m(N) > MAX_LOG2N)
        return fail(XC_ERR_UNSUPPORTED, "no Bluestein plan for length %d (limit %d)", N, 1 << (MAX_LOG2N - 1));
    const BluePlan* plan = nullptr;
    XC_TRY(blue_plan_for(N, nullptr, &plan));
    *log2m = plan->log2m;
    *d_bhat_perm = plan->d_bhat_perm;
    *d_chirp = plan->d_post;
    return XC_OK;
}

// cutB[n] = cutout[n] * b[n] for n < N, 0 for N <= n < M
__global__ void chirp_cutout_kernel(const float2* __restrict__ cutout, const float2* __restrict__ b, int N, int M,
                                    float2* __restrict__ out) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n < M) out[n] = (n < N) ? cmul(cutout[n], b[n]) : make_float2(0.f, 0.f);
}

struct BlueRunner {
    const float2* cutb;
    int N;
    const BluePlan* plan;
    const float2* rx;
    int64_t rxlen;
    const ShiftList& sl;
    const FusedOut& out;
    cudaStream_t stream;
    template <int LOG2N>
    int32_t run() {
        using PF = PlanFor<LOG2N>;
        using PL = typename PF::PL;
        if constexpr (PL::P >= 2) {
#ifndef XC_NO_TMEM_CUTOUT
            if constexpr (LOG2N == 14) {
                // the chirped cutout (128 KB) from tensor memory instead of the L2
                using LT = CutoutLoaderTmem<PL>;
                return launch_bluestein<PL, 1, LT>(LT{rx, rxlen, cutb, sl.shift0, sl.step, sl.shifts, N}, out, sl.K, plan->d_bhat_perm, plan->d_post,
                                                   plan->nbins, stream);
            }
#endif
            using LD = ChirpLoader<PL, PF::CS>;
            LD ld{rx, rxlen, cutb, sl.shift0, sl.step, sl.shifts, N};
            return launch_bluestein<PL, PF::B, LD>(ld, out, sl.K, plan->d_bhat_perm, plan->d_post, plan->nbins, stream);
        } else {
            return fail(XC_ERR_UNSUPPORTED, "no Bluestein plan for 2^%d points", LOG2N);
        }
    }
};

__global__ void pick_stride_kernel(const float* __restrict__ in, int64_t step, int64_t K, float* __restrict__ out) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < K; i += stride) out[i] = in[i * step];
}

int32_t fastxcorr_freq_device(const float2* cutout, int N, const float2* rx, int64_t rxlen, const ShiftList& sl,
                              float* qf2, uint32_t* fidx, float2* cplx, void* caf, int caf_complex,
                              cudaStream_t stream, const CztSpec* czt) {
    bool big = false;
    if (czt) {
        if (N < 1 || czt->nbins < 1) return fail(XC_ERR_ARG, "chirp-z search needs N >= 1 and at least one bin");
        big = (int64_t)N + czt->nbins - 1 > (1 << MAX_LOG2N);
        if (big && !bigfft_supported(N, czt))
            return fail(XC_ERR_UNSUPPORTED, "chirp-z search: cutout length + frequency bins - 1 must be <= 2^28, got %d + %d", N, czt->nbins);
    } else {
        big = !fused_length(N);
        if (big && !bigfft_supported(N, nullptr)) return fail(XC_ERR_UNSUPPORTED, "frequency search does not take a cutout of %d samples", N);
    }
    if (sl.K <= 0) return XC_OK;
    // normalisation terms: ||cutout||^2 (device scalar) and the window energies
    float* d_refn = reinterpret_cast<float*>(workspace(SLOT_TMP0, 256));
    float* d_energy = reinterpret_cast<float*>(workspace(SLOT_TMP1, sizeof(float) * (size_t)sl.K));
    if (!d_refn || !d_energy) return XC_ERR_NOMEM;
    XC_TRY(normsq_device(cutout, N, 1, N, d_refn, stream));
    if (!sl.shifts && sl.step == 1 && sl.shift0 >= 0 && sl.shift0 + sl.K - 1 + N <= rxlen) {
        XC_TRY(sliding_sum_device(rx + sl.shift0, true, rxlen - sl.shift0, 0, N, sl.K, d_energy, 0, stream));
    } else if (!sl.shifts && sl.step >= 2 && sl.step <= PP_MAX_STEP && sl.shift0 >= 0 && sl.shift0 + (sl.K - 1) * sl.step + N <= rxlen && N >= 256) {
        // a small step: the sliding sum over the covered range, then every step-th value (one warp per window would read
        // K N samples instead of K step)
        const int64_t nall = (sl.K - 1) * sl.step + 1;
        float* d_all = reinterpret_cast<float*>(workspace(SLOT_TMP2, sizeof(float) * (size_t)nall));
        if (!d_all) return XC_ERR_NOMEM;
        XC_TRY(sliding_sum_device(rx + sl.shift0, true, rxlen - sl.shift0, 0, N, nall, d_all, 0, stream));
        int64_t blocks = (sl.K + 255) / 256;
        if (blocks > (int64_t)sm_count() * 16) blocks = (int64_t)sm_count() * 16;
        pick_stride_kernel<<<(unsigned)blocks, 256, 0, stream>>>(d_all, sl.step, sl.K, d_energy);
        XC_LAUNCH_CHECK();
    } else {
        XC_TRY(window_energy_device(rx, rxlen, N, sl.shift0, sl.step, sl.shifts, sl.K, d_energy, stream));
    }
    FusedOut out{qf2, fidx, cplx, reinterpret_cast<float*>(caf), caf_complex, d_energy, d_refn};
    if (big) return bigfft_freq_device(cutout, N, rx, rxlen, sl, out, czt, stream);   // beyond one block FFT per shift
    if (czt || !fft_length_supported(N)) {
        // not a power of two, or a chirp-z frequency grid: Bluestein inside the fused kernel
        const BluePlan* plan = nullptr;
        XC_TRY(blue_plan_for(N, czt, &plan));
        const int M = 1 << plan->log2m;
        float2* d_cutb = reinterpret_cast<float2*>(workspace(SLOT_TMP2, sizeof(float2) * (size_t)M));
        if (!d_cutb) return XC_ERR_NOMEM;
        chirp_cutout_kernel<<<(M + 255) / 256, 256, 0, stream>>>(cutout, plan->d_b, N, M, d_cutb);
        XC_LAUNCH_CHECK();
        // shift ranges with (max, arg max) outputs: thread groups over one staged window, constants in tensor memory
        if (search_blue_pp_takes(M) && !out.caf && !sl.shifts && sl.step >= 1 && sl.step <= PP_MAX_STEP &&
            sl.K >= 16 * (int64_t)sm_count() && search_layout() == LAYOUT_PP)
            return search_blue_pp_device(StagedArgs{rx, rxlen, d_cutb, sl.shift0, (int)sl.step}, plan->d_bhat_nat, plan->d_post, N, plan->nbins,
                                         out, sl.K, M, stream);
        return dispatch_log2(plan->log2m, BlueRunner{d_cutb, N, plan, rx, rxlen, sl, out, stream});
    }
    return dispatch_log2(ilog2(N), FreqRunner{cutout, rx, rxlen, sl, out, stream});
}

}  // namespace xc

// =============================================================================================
// no frequency search: acc[s] = sum_n conj(rx[s+n]) * cutout[n]   (np.vdot(rx[s:s+N], cutout))
// =============================================================================================
namespace xc {

constexpr int OVS_MIN_N = 1024;    // overlap-save from this cutout length ...
constexpr int64_t OVS_MIN_K = 8192; // ... and this many shifts on
constexpr int ND_THREADS = 128;   // threads per CTA
constexpr int ND_R = 8;           // consecutive shifts per thread
constexpr int ND_TS = ND_THREADS * ND_R;   // shifts per CTA tile
constexpr int ND_NC = 256;        // cutout samples per shared-memory chunk

// step-1 fast path.  The rx chunk is kept as 16-byte units (two samples) with one pad unit per four,
// so the stride-4-unit LDS.128 of neighbouring threads land in distinct bank groups.
__device__ __forceinline__ int nd_phys_unit(int u) { return u + (u >> 2); }

__global__ void __launch_bounds__(ND_THREADS) nofreq_tiled_kernel(const float2* __restrict__ cutout, int N,
                                                                   const float2* __restrict__ rx, int64_t rxlen,
                                                                   int64_t shift0, int64_t K,
                                                                   const float* __restrict__ energy,
                                                                   const float* __restrict__ refnormsq,
                                                                   float* __restrict__ qf2, float2* __restrict__ cplx) {
    constexpr int RX_UNITS = (ND_TS + ND_NC) / 2;
    __shared__ float4 s_rx[RX_UNITS + RX_UNITS / 4 + 4];
    __shared__ float4 s_c[ND_NC / 2];
    const int tid = threadIdx.x;
    const int64_t i0 = (int64_t)blockIdx.x * ND_TS;          // first shift-list entry of the tile
    const int64_t base = shift0 + i0;                         // rx index of (shift i0, n = 0)
    // two-level sum: a chunk of ND_NC products in acc, the chunks in tot (a 1e5-sample cutout summed in one FP32 chain lost
    // 5e-5 of QF^2 at the peak)
    float2 acc[ND_R], tot[ND_R];
#pragma unroll
    for (int j = 0; j < ND_R; ++j) tot[j] = make_float2(0.f, 0.f);
    float2* s_rx2 = reinterpret_cast<float2*>(s_rx);
    float2* s_c2 = reinterpret_cast<float2*>(s_c);

    for (int n0 = 0; n0 < N; n0 += ND_NC) {
#pragma unroll
        for (int j = 0; j < ND_R; ++j) acc[j] = make_float2(0.f, 0.f);
        __syncthreads();
        for (int j = tid; j < ND_TS + ND_NC; j += ND_THREADS) {
            const int64_t g = base + n0 + j;
            float2 val = make_float2(0.f, 0.f);
            if (g >= 0 && g < rxlen) val = __ldg(rx + g);
            s_rx2[2 * nd_phys_unit(j >> 1) + (j & 1)] = val;
        }
        for (int j = tid; j < ND_NC; j += ND_THREADS) s_c2[j] = (n0 + j < N) ? cutout[n0 + j] : make_float2(0.f, 0.f);
        __syncthreads();
        // window registers: w[0..15] = rx chunk samples [b + 8m, b + 8m + 16), b = tid*8
        float2 w[16];
        {
            const int u0 = tid * 4;
#pragma unroll
            for (int vv = 0; vv < 4; ++vv) {
                const float4 f = s_rx[nd_phys_unit(u0 + vv)];
                w[2 * vv] = make_float2(f.x, f.y);
                w[2 * vv + 1] = make_float2(f.z, f.w);
            }
        }
#pragma unroll 1
        for (int m = 0; m < ND_NC / 8; ++m) {
            const int u1 = tid * 4 + 4 * m + 4;
#pragma unroll
            for (int vv = 0; vv < 4; ++vv) {
                const float4 f = s_rx[nd_phys_unit(u1 + vv)];
                w[8 + 2 * vv] = make_float2(f.x, f.y);
                w[8 + 2 * vv + 1] = make_float2(f.z, f.w);
            }
            float2 c[8];
#pragma unroll
            for (int vv = 0; vv < 4; ++vv) {
                const float4 f = s_c[4 * m + vv];
                c[2 * vv] = make_float2(f.x, f.y);
                c[2 * vv + 1] = make_float2(f.z, f.w);
            }
#pragma unroll
            for (int u = 0; u < 8; ++u)
#pragma unroll
                for (int j = 0; j < ND_R; ++j) {
                    const float2 x = w[u + j];
                    // acc += conj(x) * c in two packed instructions (the kernel is issue bound with scalar FFMA:
                    // profiles/r2_c1_c4_c6_ncu.md): (c.x, c.y) x.x + (c.y, -c.x) x.y -- the half swap with one negation
                    // and the scalar broadcast are operand modifiers
                    acc[j] = pfma(c[u], make_float2(x.x, x.x), acc[j]);
                    acc[j] = pfma(make_float2(c[u].y, -c[u].x), make_float2(x.y, x.y), acc[j]);
                }
#pragma unroll
            for (int j = 0; j < 8; ++j) w[j] = w[8 + j];
        }
#pragma unroll
        for (int j = 0; j < ND_R; ++j) tot[j] = padd(tot[j], acc[j]);
    }
#pragma unroll
    for (int j = 0; j < ND_R; ++j) acc[j] = tot[j];
    const float refn = *refnormsq;
#pragma unroll
    for (int j = 0; j < ND_R; ++j) {
        const int64_t i = i0 + tid * ND_R + j;
        if (i < K) {
            const float e = energy[i];
            if (qf2) qf2[i] = cnorm2(acc[j]) / refn / e;
            if (cplx) {
                const float sc = 1.f / (sqrtf(refn) * sqrtf(e));
                cplx[i] = make_float2(acc[j].x * sc, acc[j].y * sc);
            }
        }
    }
}

// general shift lists: one warp per shift
__global__ void nofreq_warp_kernel(const float2* __restrict__ cutout, int N, const float2* __restrict__ rx,
                                   int64_t rxlen, int64_t shift0, int64_t step, const int64_t* __restrict__ shifts,
                                   int64_t K, const float* __restrict__ energy, const float* __restrict__ refnormsq,
                                   float* __restrict__ qf2, float2* __restrict__ cplx) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const float refn = *refnormsq;
    for (int64_t i = warp; i < K; i += nwarps) {
        const int64_t s = shifts ? shifts[i] : shift0 + i * step;
        const bool ok = s >= 0 && s + N <= rxlen;
        float ar = 0.f, ai = 0.f;
        if (ok) {
            for (int n = lane; n < N; n += 32) {
                const float2 x = __ldg(rx + s + n);
                const float2 c = __ldg(cutout + n);
                ar = fmaf(x.x, c.x, fmaf(x.y, c.y, ar));
                ai = fmaf(x.x, c.y, fmaf(-x.y, c.x, ai));
            }
        }
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            ar += __shfl_xor_sync(0xffffffffu, ar, o);
            ai += __shfl_xor_sync(0xffffffffu, ai, o);
        }
        if (lane == 0) {
            const float e = ok ? energy[i] : 1.f;
            if (qf2) qf2[i] = ok ? fmaf(ar, ar, ai * ai) / refn / e : 0.f;
            if (cplx) {
                const float sc = ok ? 1.f / (sqrtf(refn) * sqrtf(e)) : 0.f;
                cplx[i] = make_float2(ar * sc, ai * sc);
            }
        }
    }
}

int32_t fastxcorr_nofreq_device(const float2* cutout, int N, const float2* rx, int64_t rxlen, const ShiftList& sl,
                                float* qf2, float2* cplx, cudaStream_t stream) {
    if (sl.K <= 0) return XC_OK;
    float* d_refn = reinterpret_cast<float*>(workspace(SLOT_TMP0, 256));
    float* d_energy = reinterpret_cast<float*>(workspace(SLOT_TMP1, sizeof(float) * (size_t)sl.K));
    if (!d_refn || !d_energy) return XC_ERR_NOMEM;
    XC_TRY(normsq_device(cutout, N, 1, N, d_refn, stream));
    const bool contiguous = !sl.shifts && sl.step == 1 && sl.shift0 >= 0 && sl.shift0 + sl.K - 1 + N <= rxlen;
    if (contiguous) {
        XC_TRY(sliding_sum_device(rx + sl.shift0, true, rxlen - sl.shift0, 0, N, sl.K, d_energy, 0, stream));
#ifndef XC_NO_OVS
        // long cutouts over many shifts: fast correlation through 8192- / 16384-point blocks (ovs.cuh) instead of 8 N flops per shift
        if (N >= OVS_MIN_N && N <= 8192 && sl.K >= OVS_MIN_K) {
            if (N <= 4096) return ovs_run<13>(cutout, N, rx, rxlen, sl.shift0, sl.K, d_energy, d_refn, qf2, cplx, stream);
            return ovs_run<14>(cutout, N, rx, rxlen, sl.shift0, sl.K, d_energy, d_refn, qf2, cplx, stream);
        }
#endif
        const int64_t blocks = (sl.K + ND_TS - 1) / ND_TS;
        nofreq_tiled_kernel<<<(unsigned)blocks, ND_THREADS, 0, stream>>>(cutout, N, rx, rxlen, sl.shift0, sl.K, d_energy,
                                                                          d_refn, qf2, cplx);
        XC_LAUNCH_CHECK();
    } else {
        XC_TRY(window_energy_device(rx, rxlen, N, sl.shift0, sl.step, sl.shifts, sl.K, d_energy, stream));
        int64_t blocks = (sl.K * 32 + 255) / 256;
        const int64_t cap = (int64_t)sm_count() * 8;
        if (blocks > cap) blocks = cap;
        nofreq_warp_kernel<<<(unsigned)blocks, 256, 0, stream>>>(cutout, N, rx, rxlen, sl.shift0, sl.step, sl.shifts, sl.K,
                                                                 d_energy, d_refn, qf2, cplx);
        XC_LAUNCH_CHECK();
    }
    return XC_OK;
}

}  // namespace xc

// =============================================================================================
// xc_fastxcorr_nofreq_noalloc_c64: the no-frequency-search correlator for callers that may not allocate
// (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  Same three kernels as
// fastxcorr_nofreq_device (||cutout||^2, sliding window energy, nofreq_tiled_kernel); the two normalisation
// temporaries live in a file-scope device array instead of the cached workspace.  No reference counterpart.
// =============================================================================================
namespace xc {

constexpr int64_t NOALLOC_MAX_SHIFTS = XC_NOALLOC_MAX_SHIFTS;
constexpr int NOALLOC_HEAD = 64;      // floats in front of the energies: [0] = ||cutout||^2, rest padding (256 bytes)
// 4 MiB + 256 bytes, held by every GPU that loads the library until the process exits
static __device__ __align__(256) float g_noalloc_scratch[NOALLOC_HEAD + NOALLOC_MAX_SHIFTS];
// In host code `g_noalloc_scratch` is only a handle for the runtime, not a device address: the address is looked up
// once per device (with that device current) and cached here.
static float* g_noalloc_ptr[XC_MAX_DEVICES] = {};

static int32_t noalloc_scratch(float** out) {
    int dev = 0;
    XC_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= XC_MAX_DEVICES) return fail(XC_ERR_UNSUPPORTED, "device ordinal %d out of range", dev);
    if (!g_noalloc_ptr[dev]) {
        float* p = nullptr;
        XC_CUDA_TRY(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
        if (!p) return fail(XC_ERR_CUDA, "cudaGetSymbolAddress returned a NULL scratch address on device %d", dev);
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return XC_OK;
}

}  // namespace xc

extern "C" int32_t xc_fastxcorr_nofreq_noalloc_c64(const void* cutout, int32_t N, const void* rx, int64_t rxlen,
                                                   int64_t shift0, int64_t K, float* qf2, void* stream) {
    using namespace xc;
    XC_API_GUARD;
    if (!cutout || !rx || !qf2) return fail(XC_ERR_ARG, "cutout/rx/qf2 is NULL");
    if (N < 1) return fail(XC_ERR_ARG, "cutout length must be >= 1");
    if (K < 0) return fail(XC_ERR_ARG, "number of shifts must be >= 0");
    if (K > NOALLOC_MAX_SHIFTS)
        return fail(XC_ERR_UNSUPPORTED, "xc_fastxcorr_nofreq_noalloc_c64 takes at most %lld shifts per call (the static scratch "
                    "array), got %lld: split the range or use xc_fastxcorr_c64", (long long)NOALLOC_MAX_SHIFTS, (long long)K);
    if (K == 0) return XC_OK;
    if (shift0 < 0 || shift0 + K - 1 + N > rxlen)
        return fail(XC_ERR_RANGE, "shifts [%lld, %lld] with cutout length %d leave rx of length %lld", (long long)shift0,
                    (long long)(shift0 + K - 1), N, (long long)rxlen);
    float* scratch = nullptr;
    XC_TRY(noalloc_scratch(&scratch));
    float* d_refn = scratch;                    // fully rewritten by every call: nothing has to start at zero
    float* d_energy = scratch + NOALLOC_HEAD;
    cudaStream_t st = (cudaStream_t)stream;
    const float2* dc = reinterpret_cast<const float2*>(cutout);
    const float2* drx = reinterpret_cast<const float2*>(rx);
    XC_TRY(normsq_device(dc, N, 1, N, d_refn, st, false));
    XC_TRY(sliding_sum_device(drx + shift0, true, rxlen - shift0, 0, N, K, d_energy, 0, st, false));
    const int64_t blocks = (K + ND_TS - 1) / ND_TS;
    nofreq_tiled_kernel<<<(unsigned)blocks, ND_THREADS, 0, st>>>(dc, N, drx, rxlen, shift0, K, d_energy, d_refn, qf2, nullptr);
    XC_LAUNCH_CHECK();
    return XC_OK;
}
