// fft_engine.cuh -- in-register radix codelets and the shared-memory block FFT used by the fused
// cross-correlation kernels.  No cuFFT: a forward power-of-two FFT of N = R1*R2*R3 points is done by
// a group of TG threads holding V = N/TG points each, in at most three register passes joined by
// two in-place exchanges through one padded shared-memory buffer.
//
// Decomposition (decimation in frequency, natural-order input, digit-permuted output):
//   n = n1*M1 + c         (M1 = N/R1)      pass 1: DFT_R1 over n1, times W_N^(c*k1)
//   c = n2*M2 + d         (M2 = M1/R2)     pass 2: DFT_R2 over n2, times W_M1^(d*k2)
//   d in [0, R3)                           pass 3: DFT_R3 over d
//   bin k = k1 + R1*k2 + R1*R2*k3
// The buffer index of a point is always  i = k1*M1 + (k2 or n2)*M2 + d  stored at i + (i >> log2(R_last)),
// so a thread writes back exactly the cells it read (no barrier between read and write of a pass),
// pass-1 stores and pass-2 loads/stores are lane-contiguous, and the stride-R_last loads of the last
// pass hit distinct banks thanks to the one-cell skew per R_last cells.
#pragma once
#include "common.cuh"

namespace xc {

// ---------------------------------------------------------------------------------------------
// compile-time helpers
// ---------------------------------------------------------------------------------------------
__host__ __device__ constexpr int ilog2(int v) { return v <= 1 ? 0 : 1 + ilog2(v >> 1); }
__host__ __device__ constexpr int bitrev(int v, int bits) {
    int r = 0;
    for (int i = 0; i < bits; ++i) r |= ((v >> i) & 1) << (bits - 1 - i);
    return r;
}

// One decimation-in-time butterfly (u, v) <- (u + w v, u - w v), w = exp(-2*pi*i*idx/32), idx in [0, 16),
// in packed FP32: two instructions for w = 1 and w = -i (the half swap and the sign are operand
// modifiers), three otherwise: S = u + w v by two FFMA2, then u - w v = 2u - S by one more.
// idx is a compile-time constant after unrolling.
__device__ __forceinline__ void dit_butterfly(float2& u, float2& v, int idx) {
    // cos(2*pi*k/32), k = 0..8
    constexpr float C[9] = {1.f,
                            0.98078528040323044913f,
                            0.92387953251128675613f,
                            0.83146961230254523708f,
                            0.70710678118654752440f,
                            0.55557023301960222474f,
                            0.38268343236508977173f,
                            0.19509032201612826785f,
                            0.f};
    if (idx == 0) {
        const float2 s = padd(u, v);
        v = psub(u, v);
        u = s;
        return;
    }
    if (idx == 8) {   // w = -i: w v = (v.y, -v.x)
        const float2 r = make_float2(v.y, -v.x);
        const float2 s = padd(u, r);
        v = psub(u, r);
        u = s;
        return;
    }
    const float c = (idx <= 8) ? C[idx] : -C[16 - idx];   // w = c - i*s
    const float s = (idx <= 8) ? C[8 - idx] : C[idx - 8];
    // w v = (c v.x + s v.y, c v.y - s v.x)
    float2 S = pfma(v, make_float2(c, c), u);
    S = pfma(make_float2(v.y, v.x), make_float2(s, -s), S);
    v = pfma(u, make_float2(2.f, 2.f), make_float2(-S.x, -S.y));
    u = S;
}

// In-register forward DFT of R points (R = 2,4,8,16,32): natural-order input, a[j] holds X[bitrev(j, log2 R)]
// on return.  Decimation in time on the virtual array b[i] = a[bitrev(i)] (so its bit-reversed input
// order is the natural order of a); all index arithmetic folds at compile time.
// 194 packed instructions for R = 32, 26 for R = 8.
// H0 > 1 skips the stages below H0 (the caller has run them, see folded_butterfly()).
template <int R, int H0 = 1>
__device__ __forceinline__ void codelet_fft(float2* a) {
    constexpr int LG = ilog2(R);
#pragma unroll
    for (int h = H0; h < R; h <<= 1) {
#pragma unroll
        for (int blk = 0; blk < R; blk += 2 * h) {
#pragma unroll
            for (int j = 0; j < h; ++j)
                dit_butterfly(a[bitrev(blk + j, LG)], a[bitrev(blk + j + h, LG)], j * (16 / h));
        }
    }
}
template <>
__device__ __forceinline__ void codelet_fft<1, 1>(float2*) {}

// First stage (h = 1) of codelet_fft<R> with a pending twiddle on every input: the stage pairs a[i] with a[i + R/2]
// (w = 1), so (a[i], a[i + R/2]) <- (wu a[i] + wv a[i + R/2], wu a[i] - wv a[i + R/2]) costs five packed
// instructions (P = wu u; S = P + wv v; D = 2 P - S) instead of two complex multiplies and a butterfly (six).
__device__ __forceinline__ void folded_butterfly(float2& u, float2& v, float2 wu, float2 wv) {
    const float2 P = pcmul(u, wu);
    float2 S = pfma(v, make_float2(wv.x, wv.x), P);
    S = pfma(make_float2(v.y, v.x), make_float2(-wv.y, wv.y), S);
    v = pfma(P, make_float2(2.f, 2.f), make_float2(-S.x, -S.y));
    u = S;
}

// multiply the bit-reversed outputs of a codelet by successive powers of w: a[bitrev(k)] *= w^k
template <int R>
__device__ __forceinline__ void twiddle_powers(float2* a, float2 w) {
    float2 p = w;
#pragma unroll
    for (int k = 1; k < R; ++k) {
        const int j = bitrev(k, ilog2(R));
        a[j] = pcmul(a[j], p);
        if (k + 1 < R) p = pcmul(p, w);
    }
}

__device__ __forceinline__ float2 unit_root(int num, int den) {
    // exp(-2*pi*i*num/den), evaluated in double once per thread
    double s, c;
    sincospi(-2.0 * (double)num / (double)den, &s, &c);
    return make_float2((float)c, (float)s);
}

// ---------------------------------------------------------------------------------------------
// block FFT plan
// ---------------------------------------------------------------------------------------------
template <int LOG2N_, int TG_, int R1_, int R2_, int R3_>
struct FftPlan {
    static constexpr int LOG2N = LOG2N_;
    static constexpr int N = 1 << LOG2N_;
    static constexpr int TG = TG_;          // threads cooperating on one FFT
    static constexpr int V = N / TG_;       // points per thread
    static constexpr int R1 = R1_, R2 = R2_, R3 = R3_;
    static constexpr int P = (R3_ > 1) ? 3 : ((R2_ > 1) ? 2 : 1);
    static constexpr int RL = (P == 3) ? R3_ : ((P == 2) ? R2_ : R1_);   // radix of the last pass
    static constexpr int M1 = N / R1_;
    static constexpr int M2 = M1 / R2_;
    static constexpr int LOG2RL = ilog2(RL);
    // exchange buffer: cell a*RL + b (b < RL) lives at a*ROW + b.  ROW = RL + 2 keeps every row 16-byte
    // aligned (LDS.128 in the last pass) and its 16-byte pitch odd (conflict-free across lanes).
    static constexpr int ROW = RL + 2;
    static constexpr int BUF = (P == 1) ? 0 : (N / RL) * ROW;            // float2 cells per FFT buffer
    // twiddle tables shared by all groups of a CTA (three-pass plans): W_N^c for the M1 columns of
    // pass 1 (base of a per-FFT power chain) and the full R2 x M2 table W_M1^(d*k2) of pass 2.
    static constexpr int T1_CELLS = (P == 3) ? M1 : 0;
    static constexpr int T2_ROW = R2_ + 2;
    static constexpr int T2_CELLS = (P == 3) ? M2 * T2_ROW : 0;
    static constexpr size_t TABLE_BYTES = sizeof(float2) * (size_t)(T1_CELLS + T2_CELLS);
    // second exchange stays inside one warp when every thread owns one radix-32 column in passes 2 and 3
    static constexpr bool WARP_LOCAL_X2 = (P == 3) && R2_ == 32 && R3_ == 32 && V == 32;
    static_assert(R1_ * R2_ * R3_ == N, "radices must multiply to N");
    static_assert(V % R1_ == 0 && V % R2_ == 0 && V % R3_ == 0, "each pass must fill the registers");
    static_assert(R1_ <= 32 && R2_ <= 32 && R3_ <= 32, "codelets go up to radix 32");
    static_assert(R3_ == 1 || (TG_ % R3_ == 0 && TG_ % R2_ == 0), "three-pass plans need TG to be a multiple of R2 and R3");
    static_assert(R3_ > 1 || R2_ == 1 || TG_ * (N / TG_ / R1_) == R2_, "two-pass plans: the columns of pass 1 are the R2 points");
    static_assert(P == 1 || RL % 2 == 0, "LDS.128 of the last pass needs an even radix");
};

// Fills the CTA-wide twiddle tables (call once per CTA, before a __syncthreads()).
template <class PL>
__device__ __forceinline__ void build_fft_tables(float2* tab, int tid, int nthreads) {
    if (PL::P != 3) return;
    for (int c = tid; c < PL::T1_CELLS; c += nthreads) tab[c] = unit_root(c, PL::N);
    for (int i = tid; i < PL::M2 * PL::R2; i += nthreads) {
        const int d = i / PL::R2, k2 = i % PL::R2;
        tab[PL::T1_CELLS + d * PL::T2_ROW + k2] = unit_root((d * k2) % PL::M1, PL::M1);
    }
}

// Per-thread constants of a plan (two-pass plans keep the base twiddles of their one exchange here;
// three-pass plans read theirs from the shared tables and keep nothing).
template <class PL>
struct FftThread {
    float2 tw1[(PL::P == 2) ? PL::V / PL::R1 : 1];    // W_N^(t + TG*q)
    __device__ __forceinline__ void init(int t) {
        if (PL::P == 2) {
#pragma unroll
            for (int q = 0; q < PL::V / PL::R1; ++q) tw1[q] = unit_root(t + PL::TG * q, PL::N);
        }
    }
};

// Sample index held in register slot (q*R1 + n1) of thread t before pass 1.
template <class PL>
__device__ __forceinline__ int input_index(int t, int q, int n1) {
    return n1 * PL::M1 + t + PL::TG * q;
}

// Bin index held in register slot (q*RL + j) of thread t after the last pass, split into a per-thread
// base and a compile-time offset so that no per-slot value has to live in a register.
template <class PL>
__host__ __device__ __forceinline__ int output_bin_base(int t) {
    if (PL::P == 1) return 0;
    if (PL::P == 2) return t;
    return (t / PL::R2) + PL::R1 * (t % PL::R2);      // needs TG % R2 == 0 (asserted in the plan)
}
template <class PL>
__host__ __device__ constexpr int output_bin_off(int q, int j) {
    return (PL::P == 1)   ? bitrev(j, PL::LOG2RL)
           : (PL::P == 2) ? PL::TG * q + PL::R1 * bitrev(j, PL::LOG2RL)
                          : q * (PL::TG / PL::R2) + PL::R1 * PL::R2 * bitrev(j, PL::LOG2RL);
}
template <class PL>
__device__ __forceinline__ int output_bin(int t, int q, int j) {
    return output_bin_base<PL>(t) + output_bin_off<PL>(q, j);
}

// Runs the forward FFT on the V points the thread loaded according to input_index(); on return the
// registers hold the bins given by output_bin().  `buf` is the group's exchange buffer, `tab` the CTA's
// twiddle tables and BAR() a barrier over the TG threads of the group.  The caller must guarantee
// (with BAR) that the previous user of `buf` is done before calling.
// All shared-memory addresses are (per-thread base) + (compile-time constant).
template <class PL, class Bar>
__device__ __forceinline__ void block_fft(float2* v, const FftThread<PL>& th, const float2* tab, float2* buf, int t,
                                          Bar bar) {
    constexpr int RL = PL::RL, ROW = PL::ROW;
    // ---- pass 1 ----
#pragma unroll
    for (int q = 0; q < PL::V / PL::R1; ++q) codelet_fft<PL::R1>(v + q * PL::R1);
    if (PL::P == 1) return;
    {
        // cell k1*M1 + c, c = t + TG*q.  P == 2: M1 == RL and c < RL.  P == 3: M1 and TG are multiples of RL.
        constexpr int K1STEP = (PL::M1 / RL) * ROW;
        constexpr int QSTEP = (PL::P == 2) ? PL::TG : (PL::TG / RL) * ROW;
        float2* const b1 = buf + ((PL::P == 2) ? t : (t / RL) * ROW + (t % RL));
#pragma unroll
        for (int q = 0; q < PL::V / PL::R1; ++q) {
            const float2 w = (PL::P == 2) ? th.tw1[q] : tab[t + PL::TG * q];
            twiddle_powers<PL::R1>(v + q * PL::R1, w);
#pragma unroll
            for (int j = 0; j < PL::R1; ++j) b1[bitrev(j, ilog2(PL::R1)) * K1STEP + q * QSTEP] = v[q * PL::R1 + j];
        }
    }
    bar();
    if (PL::P == 3) {
        // ---- pass 2 (in place): cell k1*M1 + n2*M2 + d with e = t + TG*q = k1*M2 + d, M2 == RL ----
        constexpr int K1STEP = (PL::M1 / RL) * ROW;
        constexpr int QSTEP = (PL::TG / PL::M2) * K1STEP;
        float2* const b2 = buf + (t / PL::M2) * K1STEP + (t % PL::M2);
        const float4* const t2 = reinterpret_cast<const float4*>(tab + PL::T1_CELLS + (t % PL::M2) * PL::T2_ROW);
#pragma unroll
        for (int q = 0; q < PL::V / PL::R2; ++q)
#pragma unroll
            for (int n2 = 0; n2 < PL::R2; ++n2) v[q * PL::R2 + n2] = b2[q * QSTEP + n2 * ROW];
#pragma unroll
        for (int q = 0; q < PL::V / PL::R2; ++q) {
            codelet_fft<PL::R2>(v + q * PL::R2);
            // twiddles W_M1^(d*k2) from the table, two per 16-byte load (k2 = 2m, 2m+1)
#pragma unroll
            for (int m = 0; m < PL::R2 / 2; ++m) {
                const float4 w = t2[m];
                const int j0 = bitrev(2 * m, ilog2(PL::R2)), j1 = bitrev(2 * m + 1, ilog2(PL::R2));
                if (m > 0) v[q * PL::R2 + j0] = pcmul(v[q * PL::R2 + j0], make_float2(w.x, w.y));
                v[q * PL::R2 + j1] = pcmul(v[q * PL::R2 + j1], make_float2(w.z, w.w));
            }
#pragma unroll
            for (int j = 0; j < PL::R2; ++j) b2[q * QSTEP + bitrev(j, ilog2(PL::R2)) * ROW] = v[q * PL::R2 + j];
        }
        if (PL::WARP_LOCAL_X2) {
            __syncwarp();   // the 32 rows a warp wrote are exactly the 32 rows it reads next
        } else {
            bar();
        }
    }
    // ---- last pass: cell f*RL + d, f = t + TG*q; rows are 16-byte aligned ----
    {
        const float4* const b3 = reinterpret_cast<const float4*>(buf + t * ROW);
#pragma unroll
        for (int q = 0; q < PL::V / RL; ++q)
#pragma unroll
            for (int d2 = 0; d2 < RL / 2; ++d2) {
                const float4 x = b3[(q * PL::TG * ROW) / 2 + d2];
                v[q * RL + 2 * d2] = make_float2(x.x, x.y);
                v[q * RL + 2 * d2 + 1] = make_float2(x.z, x.w);
            }
    }
#pragma unroll
    for (int q = 0; q < PL::V / RL; ++q) codelet_fft<RL>(v + q * RL);
}

}  // namespace xc
