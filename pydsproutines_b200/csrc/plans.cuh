// plans.cuh -- the block-FFT plan chosen for each power-of-two length: threads per FFT, pass radices,
// FFTs in flight per CTA and whether the per-call constant vector (cutout) is kept in shared memory.
// From 1024 points up every thread holds 32 points and the last two passes are radix 32, so the
// first exchange is lane-contiguous on both sides and the second only needs the one-cell skew.
#pragma once
#include "fft_engine.cuh"

namespace xc {

template <int LOG2N> struct PlanFor;
template <> struct PlanFor<1> { using PL = FftPlan<1, 1, 2, 1, 1>; static constexpr int B = 128; static constexpr bool CS = true; };
template <> struct PlanFor<2> { using PL = FftPlan<2, 1, 4, 1, 1>; static constexpr int B = 128; static constexpr bool CS = true; };
template <> struct PlanFor<3> { using PL = FftPlan<3, 1, 8, 1, 1>; static constexpr int B = 128; static constexpr bool CS = true; };
template <> struct PlanFor<4> { using PL = FftPlan<4, 1, 16, 1, 1>; static constexpr int B = 128; static constexpr bool CS = true; };
template <> struct PlanFor<5> { using PL = FftPlan<5, 1, 32, 1, 1>; static constexpr int B = 128; static constexpr bool CS = true; };
template <> struct PlanFor<6> { using PL = FftPlan<6, 8, 8, 8, 1>; static constexpr int B = 32; static constexpr bool CS = true; };
template <> struct PlanFor<7> { using PL = FftPlan<7, 8, 16, 8, 1>; static constexpr int B = 32; static constexpr bool CS = true; };
template <> struct PlanFor<8> { using PL = FftPlan<8, 16, 16, 16, 1>; static constexpr int B = 16; static constexpr bool CS = true; };
template <> struct PlanFor<9> { using PL = FftPlan<9, 16, 16, 32, 1>; static constexpr int B = 16; static constexpr bool CS = true; };
template <> struct PlanFor<10> { using PL = FftPlan<10, 32, 32, 32, 1>; static constexpr int B = 8; static constexpr bool CS = true; };
template <> struct PlanFor<11> { using PL = FftPlan<11, 64, 2, 32, 32>; static constexpr int B = 8; static constexpr bool CS = true; };
template <> struct PlanFor<12> { using PL = FftPlan<12, 128, 4, 32, 32>; static constexpr int B = 4; static constexpr bool CS = true; };
template <> struct PlanFor<13> { using PL = FftPlan<13, 256, 8, 32, 32>; static constexpr int B = 2; static constexpr bool CS = true; };
template <> struct PlanFor<14> { using PL = FftPlan<14, 512, 16, 32, 32>; static constexpr int B = 1; static constexpr bool CS = false; };

constexpr int MAX_LOG2N = 14;

inline bool fft_length_supported(int N) { return N >= 2 && (N & (N - 1)) == 0 && ilog2(N) <= MAX_LOG2N; }

// calls f.template run<LOG2N>() for the runtime length
template <class F>
int32_t dispatch_log2(int log2n, F&& f) {
    switch (log2n) {
        case 1: return f.template run<1>();
        case 2: return f.template run<2>();
        case 3: return f.template run<3>();
        case 4: return f.template run<4>();
        case 5: return f.template run<5>();
        case 6: return f.template run<6>();
        case 7: return f.template run<7>();
        case 8: return f.template run<8>();
        case 9: return f.template run<9>();
        case 10: return f.template run<10>();
        case 11: return f.template run<11>();
        case 12: return f.template run<12>();
        case 13: return f.template run<13>();
        case 14: return f.template run<14>();
        default: break;
    }
    return fail(XC_ERR_UNSUPPORTED, "no block-FFT plan for 2^%d points", log2n);
}

}  // namespace xc
