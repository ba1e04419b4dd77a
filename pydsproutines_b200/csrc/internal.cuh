// internal.cuh -- device-side entry points shared between the translation units of libxcorr_b200
// (all take DEVICE pointers and enqueue on `stream`; the C-ABI layer in api.cu does the staging).
#pragma once
#include "common.cuh"

namespace xc {

// energy.cu
// scratch = false keeps the call allocation free (very long windows then pay a serial first-window sum per tile)
int32_t sliding_sum_device(const void* x, bool complex_in, int64_t n_in, int64_t front_pad, int L, int64_t n_out,
                           float* out, int average, cudaStream_t stream, bool scratch = true);
int32_t window_energy_device(const float2* x, int64_t xlen, int L, int64_t shift0, int64_t step, const int64_t* shifts,
                             int64_t K, float* out, cudaStream_t stream);
int32_t normsq_device(const float2* x, int64_t n, int64_t rows, int64_t row_stride, float* out, cudaStream_t stream, bool scratch = true);
int32_t magnsq_device(const float2* x, int64_t n, void* out, int out_f64, cudaStream_t stream);

// fastxcorr.cu
struct ShiftList {
    int64_t shift0, step, K;
    const int64_t* shifts;   // device pointer or NULL
};
bool fastxcorr_fft_supported(int N);
bool fft_length_pow2(int N);   // a power of two with a block-FFT plan
// chirp-z frequency grid: bins k < nbins at (a_cycles + k * w_cycles) cycles per sample
struct CztSpec {
    int nbins;
    double a_cycles, w_cycles;
};
// czt == NULL: the N-point DFT grid (bin k at k / N cycles per sample)
int32_t fastxcorr_freq_device(const float2* cutout, int N, const float2* rx, int64_t rxlen, const ShiftList& sl,
                              float* qf2, uint32_t* fidx, float2* cplx, void* caf, int caf_complex,
                              cudaStream_t stream, const CztSpec* czt = nullptr);
int32_t fastxcorr_nofreq_device(const float2* cutout, int N, const float2* rx, int64_t rxlen, const ShiftList& sl,
                                float* qf2, float2* cplx, cudaStream_t stream);

// bigfft.cu: four-step frequency search for lengths beyond one block FFT (any N, chirp-z grids), FusedOut-style outputs
struct FusedOut;
bool bigfft_supported(int64_t N, const CztSpec* czt);
int32_t bigfft_freq_device(const float2* cutout, int64_t N, const float2* rx, int64_t rxlen, const ShiftList& sl,
                           const FusedOut& out, const CztSpec* czt, cudaStream_t stream);
void release_big_plan_caches();

// group_pp.cu: ping-pong GroupXcorrFFT kernel for fftlen 4096 over contiguous shifts (POD: lives inside xc_group)
struct GroupPPPlan {
    int usable;
    int G, L, W, NC, nq_live;
    unsigned char lens[32];
    float2* d_w;
    uint32_t* d_off;
    // rotated-row variant (starts that are not aligned to the 128-cell rows): sub-groups sorted by row
    int rot, B, F;               // B = thread groups per SM of the kernel variant, F = transform length
    int bin_shift;               // the kernel's transform is 2^bin_shift times the requested fftlen: only every 2^bin_shift-th bin counts
    int dense, G4;               // dense-fold variant for short transforms (group_dense.cuh): d_w = float4 [G4][F / 2], d_off = descriptors
    unsigned char rcnt[64];
    uint32_t rdesc[192];
};
int32_t group_pp_build(const float2* y, const int64_t* starts, int G, int L, int F, GroupPPPlan* p);
void group_pp_release(GroupPPPlan* p);
int32_t group_pp_run(const GroupPPPlan& p, const int64_t* d_starts, const float2* rx, int64_t rxlen, int64_t shift0, int64_t K,
                     int64_t lo, int64_t hi, const FusedOut& out, cudaStream_t stream);

// Bluestein tables of a transform length that is not a power of two (cached per length): padded plan length and
// FFT_M(chirp) laid out for the block FFT's output registers
int32_t bluestein_tables(int N, int* log2m, const float2** d_bhat_perm, const float2** d_chirp);

// misc.cu
int32_t peak_reduce_device(const float* qf2, const uint32_t* fidx, const int32_t* aux, int64_t K,
                           unsigned long long* d_packed, cudaStream_t stream, const int64_t* shifts = nullptr);
int32_t row_argmax_device(const float2* x, int64_t rows, int64_t cols, int use_normsq, uint32_t* idx, float* val,
                          cudaStream_t stream);
int32_t local_maxima_device(const float* x, int64_t n, float min_height, int32_t* idx, int32_t cap, int32_t* count,
                            cudaStream_t stream);

// mt_tc.cu -- tensor-core (tcgen05) multi-template correlator
struct MtTcPlan {
    int T = 0, L = 0;
    int L32 = 0;          // template length padded to a multiple of 32 (one pipeline stage = 32 samples)
    int Lp = 0;           // samples an x tile must cover: L32 (dense) or the longest group (segmented), multiple of 32
    int nseg = 1;         // 1 = dense; > 1 = the templates are non-zero only inside nseg groups
    int32_t* d_segs = nullptr;     // [nseg][2] {x offset of the group, K steps of 16 samples}
    int NT = 0;           // templates per block = UMMA N (multiple of 16, <= 256)
    int NB = 0;           // template blocks
    void* d_pack = nullptr;        // BF16 stage images [NB][L32/32][planes][canonical K-major core matrices]
    float* d_inv_energy = nullptr; // [NB * NT] 1 / E_t (0 for padding templates)
    int planes = 2;       // 2: templates exact in BF16 (re, im); 4: BF16 hi + lo split (re, im, re_lo, im_lo)
    bool usable = false;  // the tile fits in shared memory (otherwise the FP32 kernel is used)
};
int32_t mt_tc_prepare(const float2* d_templates, int T, int L, int L8, const float* d_energies, MtTcPlan* plan,
                      const int32_t* segs = nullptr, int nseg = 0);
void mt_tc_release(MtTcPlan* plan);
int32_t mt_tc_run(const MtTcPlan& plan, const float2* x, int64_t xlen, int64_t start, int64_t K, const float* xenergy,
                  float* best, int32_t* tidx, int sqrt_out, cudaStream_t stream, float* caf = nullptr,
                  float2* cplx = nullptr);

}  // namespace xc
