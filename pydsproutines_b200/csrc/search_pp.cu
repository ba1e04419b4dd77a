// search_pp.cu -- instantiations of the ping-pong layout of the frequency search (fused_fft_pp.cuh), kept in their own
// translation unit: 8192 points (two groups of 128 threads, 64 points per thread), 4096 (four groups of 128 threads, 32
// points per thread), 2048 (eight groups of 64 threads) and 1024 (sixteen warps, one FFT per warp).
#include "fused_fft_pp.cuh"
#include "fused_blue_pp.cuh"
#include "fused_fft_warp.cuh"
#include "fused_fft_subwarp.cuh"

#ifndef XC_PP_TG8192
#define XC_PP_TG8192 128
#endif
#ifndef XC_PP_B4096
#define XC_PP_B4096 4
#endif
#ifndef XC_PP_B2048
#define XC_PP_B2048 8
#endif
#ifndef XC_PP_B1024
#define XC_PP_B1024 16
#endif

namespace xc {

bool search_pp_takes(int N) { return N == 8192 || N == 4096 || N == 2048 || N == 1024 || N == 512 || N == 256 || N == 128; }

int32_t search_pp_device(const StagedArgs& a, const FusedOut& out, int64_t K, int N, int fold, cudaStream_t stream) {
#ifdef XC_EXPERIMENTS
    if (!fold && N == 8192) return launch_fused_pp<FftPlan<13, 128, 8, 32, 32>, 2, false>(a, out, K, stream);
#endif
    (void)fold;
    if (N == 8192) return launch_fused_pp<FftPlan<13, XC_PP_TG8192, 8, 32, 32>, 2, true>(a, out, K, stream);
    if (N == 4096) return launch_fused_pp<FftPlan<12, 128, 4, 32, 32>, XC_PP_B4096, true>(a, out, K, stream);
    if (N == 2048) return launch_fused_pp<FftPlan<11, 64, 2, 32, 32>, XC_PP_B2048, true>(a, out, K, stream);
    if (N == 1024) return launch_fused_warp<XC_PP_B1024>(a, out, K, stream);      // one warp per FFT (fused_fft_warp.cuh)
    if (N == 512) return launch_fused_subwarp<4, 16>(a, out, K, stream);          // 16 lanes per FFT, two FFTs per warp (fused_fft_subwarp.cuh)
    if (N == 256) return launch_fused_subwarp<3, 16>(a, out, K, stream);          // 8 lanes per FFT, four per warp
    if (N == 128) return launch_fused_subwarp<2, 16>(a, out, K, stream);          // 4 lanes per FFT, eight per warp
    return fail(XC_ERR_UNSUPPORTED, "no ping-pong layout for %d points", N);
}

// Bluestein (cutout lengths that are not a power of two, chirp-z grids) over a padded length M = 2048 / 4096 / 8192
#ifdef XC_NO_BLUE_PP
bool search_blue_pp_takes(int) { return false; }
#else
bool search_blue_pp_takes(int M) { return M == 8192 || M == 4096 || M == 2048; }
#endif
int32_t search_blue_pp_device(const StagedArgs& a, const float2* bhat, const float2* post, int nvalid, int nbins, const FusedOut& out,
                              int64_t K, int M, cudaStream_t stream) {
    const double m = (double)M;
    const BlueArgs ba{bhat, nvalid, nbins, (float)(1.0 / (m * m)), post};
    if (M == 8192) return launch_blue_pp<FftPlan<13, 256, 8, 32, 32>, 2>(a, ba, out, K, stream);     // 16 warps: 32 points per thread
    if (M == 4096) return launch_blue_pp<FftPlan<12, 128, 4, 32, 32>, XC_PP_B4096>(a, ba, out, K, stream);
    if (M == 2048) return launch_blue_pp<FftPlan<11, 64, 2, 32, 32>, XC_PP_B2048>(a, ba, out, K, stream);
    return fail(XC_ERR_UNSUPPORTED, "no ping-pong Bluestein layout for %d points", M);
}

}  // namespace xc
