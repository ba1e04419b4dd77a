// search_pp.cu -- instantiation of the ping-pong layout of the N = 8192 frequency search (fused_fft_pp.cuh), kept in
// its own translation unit.
#include "fused_fft_pp.cuh"

namespace xc {

int32_t search_pp_device(const StagedArgs& a, const FusedOut& out, int64_t K, int fold, cudaStream_t stream) {
    using PL = FftPlan<13, 128, 8, 32, 32>;
#ifdef XC_EXPERIMENTS
    if (!fold) return launch_fused_pp<PL, 2, false>(a, out, K, stream);
#endif
    (void)fold;
    return launch_fused_pp<PL, 2, true>(a, out, K, stream);
}

}  // namespace xc
