// fused_fft_staged.cuh -- the N = 8192 frequency search over a CONTIGUOUS shift range with the rx window resident
// in shared memory.  Built on the two-FFTs-per-thread layout of fused_fft_dual.cuh (one group of TG threads per SM,
// a thread carries the windows of shifts s and s + 1 in lockstep, sharing every conj(cutout) value, the pass-1
// twiddle chains and the pass-2 twiddle table), with two changes that make room for the window:
//   * ONE exchange buffer: the two FFTs take turns in the first exchange (the second exchange is warp local and
//     in place, so it needs no second buffer either);
//   * the samples rx[base, base + N + W) are staged once per W / 2 iterations by the whole group (plain coalesced
//     copies, zero filled outside the record) and every product reads its two windows from there: no global load,
//     hence no L2 latency, inside the per-shift loop, and the L2 -> L1 traffic drops from N to 1 sample per shift.
// (max, arg max) outputs only; shift lists with a step or a gather table stay on fused_fft_kernel.
#pragma once
#include "fused_fft_dual.cuh"

namespace xc {

template <class PL>
struct StagedLayout {
    static constexpr int W = 1792;                                        // extra samples behind the window (even)
    static constexpr size_t OFF_CUT = 0;
    static constexpr size_t OFF_TAB = sizeof(float2) * (size_t)PL::N;
    static constexpr size_t OFF_BUF = OFF_TAB + PL::TABLE_BYTES;
    static constexpr size_t OFF_STAGE = OFF_BUF + sizeof(float2) * (size_t)PL::BUF;
    static constexpr size_t OFF_WMAX = OFF_STAGE + sizeof(float2) * (size_t)(PL::N + W);
    static constexpr size_t OFF_BIN = OFF_WMAX + sizeof(float) * (PL::TG / 32) * 2;
    static constexpr size_t OFF_PEND = OFF_BIN + sizeof(unsigned int) * 2;
    static constexpr size_t BYTES = OFF_PEND + sizeof(float) * 2 + 16;
};

// Both FFTs through one exchange buffer.  On entry v0 / v1 hold the products in input_index() order, on return the
// bins of output_bin().  Three CTA-wide barriers inside; the caller's epilogue barrier frees the buffer.
template <class PL>
__device__ __forceinline__ void block_fft_dual_1buf(float2* v0, float2* v1, const float2* x, const float2* c, const float2* tab,
                                                    float2* buf, int t) {
    static_assert(PL::P == 3 && PL::WARP_LOCAL_X2, "staged layout: three passes, warp-local second exchange");
    constexpr int RL = PL::RL, ROW = PL::ROW;
    constexpr int K1STEP = (PL::M1 / RL) * ROW;
    // ---- products of both windows (x = stage + offset + t, c = conj(cutout) + t) and pass 1, column by column ----
#pragma unroll
    for (int q = 0; q < PL::V / PL::R1; ++q) {
#pragma unroll
        for (int n1 = 0; n1 < PL::R1; ++n1) {
            const int n = input_index<PL>(0, q, n1);
            const float2 cv = c[n];
            v0[q * PL::R1 + n1] = pcmul(x[n], cv);
            v1[q * PL::R1 + n1] = pcmul(x[n + 1], cv);
        }
        codelet_fft<PL::R1>(v0 + q * PL::R1);
        codelet_fft<PL::R1>(v1 + q * PL::R1);
    }
    constexpr int QSTEP1 = (PL::TG / RL) * ROW;
    const int o1 = (t / RL) * ROW + (t % RL);
#pragma unroll
    for (int q = 0; q < PL::V / PL::R1; ++q) {
        const float2 w = tab[t + PL::TG * q];
        twiddle_powers2<PL::R1>(v0 + q * PL::R1, v1 + q * PL::R1, w);
#pragma unroll
        for (int j = 0; j < PL::R1; ++j) buf[o1 + bitrev(j, ilog2(PL::R1)) * K1STEP + q * QSTEP1] = v0[q * PL::R1 + j];
    }
    __syncthreads();
    // ---- first exchange, one FFT after the other ----
    constexpr int QSTEP2 = (PL::TG / PL::M2) * K1STEP;
    const int o2 = (t / PL::M2) * K1STEP + (t % PL::M2);
#pragma unroll
    for (int q = 0; q < PL::V / PL::R2; ++q)
#pragma unroll
        for (int n2 = 0; n2 < PL::R2; ++n2) v0[q * PL::R2 + n2] = buf[o2 + q * QSTEP2 + n2 * ROW];
    __syncthreads();
#pragma unroll
    for (int q = 0; q < PL::V / PL::R1; ++q)
#pragma unroll
        for (int j = 0; j < PL::R1; ++j) buf[o1 + bitrev(j, ilog2(PL::R1)) * K1STEP + q * QSTEP1] = v1[q * PL::R1 + j];
    __syncthreads();
#pragma unroll
    for (int q = 0; q < PL::V / PL::R2; ++q)
#pragma unroll
        for (int n2 = 0; n2 < PL::R2; ++n2) v1[q * PL::R2 + n2] = buf[o2 + q * QSTEP2 + n2 * ROW];
    // ---- pass 2; from here on a warp only touches its own rows of the buffer ----
    const float4* const t2 = reinterpret_cast<const float4*>(tab + PL::T1_CELLS + (t % PL::M2) * PL::T2_ROW);
#pragma unroll
    for (int q = 0; q < PL::V / PL::R2; ++q) {
        codelet_fft<PL::R2>(v0 + q * PL::R2);
        codelet_fft<PL::R2>(v1 + q * PL::R2);
#pragma unroll
        for (int m = 0; m < PL::R2 / 2; ++m) {
            const float4 w = t2[m];
            const int j0 = bitrev(2 * m, ilog2(PL::R2)), j1 = bitrev(2 * m + 1, ilog2(PL::R2));
            if (m > 0) {
                v0[q * PL::R2 + j0] = pcmul(v0[q * PL::R2 + j0], make_float2(w.x, w.y));
                v1[q * PL::R2 + j0] = pcmul(v1[q * PL::R2 + j0], make_float2(w.x, w.y));
            }
            v0[q * PL::R2 + j1] = pcmul(v0[q * PL::R2 + j1], make_float2(w.z, w.w));
            v1[q * PL::R2 + j1] = pcmul(v1[q * PL::R2 + j1], make_float2(w.z, w.w));
        }
    }
    // ---- second exchange (in place, warp local), FFT 0 then FFT 1, and the last pass ----
    const float4* const b3 = reinterpret_cast<const float4*>(buf + t * ROW);
    __syncwarp();   // every lane holds its FFT-1 operands before the rows are overwritten
#pragma unroll
    for (int q = 0; q < PL::V / PL::R2; ++q)
#pragma unroll
        for (int j = 0; j < PL::R2; ++j) buf[o2 + q * QSTEP2 + bitrev(j, ilog2(PL::R2)) * ROW] = v0[q * PL::R2 + j];
    __syncwarp();
#pragma unroll
    for (int q = 0; q < PL::V / RL; ++q)
#pragma unroll
        for (int d2 = 0; d2 < RL / 2; ++d2) {
            const float4 x = b3[(q * PL::TG * ROW) / 2 + d2];
            v0[q * RL + 2 * d2] = make_float2(x.x, x.y);
            v0[q * RL + 2 * d2 + 1] = make_float2(x.z, x.w);
        }
    __syncwarp();
#pragma unroll
    for (int q = 0; q < PL::V / PL::R2; ++q)
#pragma unroll
        for (int j = 0; j < PL::R2; ++j) buf[o2 + q * QSTEP2 + bitrev(j, ilog2(PL::R2)) * ROW] = v1[q * PL::R2 + j];
#pragma unroll
    for (int q = 0; q < PL::V / RL; ++q) codelet_fft<RL>(v0 + q * RL);   // under the stores / loads of FFT 1
    __syncwarp();
#pragma unroll
    for (int q = 0; q < PL::V / RL; ++q)
#pragma unroll
        for (int d2 = 0; d2 < RL / 2; ++d2) {
            const float4 x = b3[(q * PL::TG * ROW) / 2 + d2];
            v1[q * RL + 2 * d2] = make_float2(x.x, x.y);
            v1[q * RL + 2 * d2 + 1] = make_float2(x.z, x.w);
        }
#pragma unroll
    for (int q = 0; q < PL::V / RL; ++q) codelet_fft<RL>(v1 + q * RL);
}

template <class PL>
__global__ void __launch_bounds__(PL::TG) fused_fft_staged_kernel(const StagedArgs a, const FusedOut out, const int64_t K,
                                                                  const int chunk) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    using L = StagedLayout<PL>;
    constexpr int TG = PL::TG, V = PL::V, NW = TG / 32, W = L::W;
    const int t = threadIdx.x;
    float2* const cc = reinterpret_cast<float2*>(smem_raw + L::OFF_CUT);
    for (int n = t; n < PL::N; n += TG) {
        const float2 c = a.cutout[n];
        cc[n] = make_float2(c.x, -c.y);
    }
    const float2* const tab = reinterpret_cast<const float2*>(smem_raw + L::OFF_TAB);
    build_fft_tables<PL>(reinterpret_cast<float2*>(smem_raw + L::OFF_TAB), t, TG);
    float2* const buf = reinterpret_cast<float2*>(smem_raw + L::OFF_BUF);
    float2* const stage = reinterpret_cast<float2*>(smem_raw + L::OFF_STAGE);
    float* const wmax = reinterpret_cast<float*>(smem_raw + L::OFF_WMAX);            // [2][NW]
    unsigned int* const binmin = reinterpret_cast<unsigned int*>(smem_raw + L::OFF_BIN);
    float* const pend = reinterpret_cast<float*>(smem_raw + L::OFF_PEND);
    if (t < 2) binmin[t] = 0xffffffffu;
    if (t < 4) reinterpret_cast<unsigned int*>(wmax)[t] = 0u;

    // this CTA owns shift-list entries [i0, i0 + 2 * chunk), taken in pairs
    const int64_t i0 = (int64_t)blockIdx.x * 2 * (int64_t)chunk;
    int64_t n_left = K - i0;
    if (n_left < 0) n_left = 0;
    if (n_left > 2 * (int64_t)chunk) n_left = 2 * (int64_t)chunk;
    const int n_act = (int)n_left;
    const int64_t s_first = a.shift0 + i0;
    float2 v0[V], v1[V];
    int o = W;                                        // offset of the current pair inside the stage

    for (int it = 0; it < chunk; ++it) {
        const int e0 = 2 * it, e1 = 2 * it + 1;
        if (o == W) {
            // refill: stage[j] = rx[s + j], j < N + W (zero outside the record).  Every thread is past the epilogue
            // barrier of the previous pair, so nobody still reads the old window.
            const int64_t base = s_first + e0;
            for (int j = t; j < PL::N + W; j += TG) {
                const int64_t g = base + j;
                stage[j] = (g >= 0 && g < a.rxlen) ? __ldg(a.rx + g) : make_float2(0.f, 0.f);
            }
            o = 0;
            __syncthreads();
        }
        const int64_t s0 = s_first + e0;
        const bool ok0 = e0 < n_act && s0 >= 0 && s0 + PL::N <= a.rxlen;
        const bool ok1 = e1 < n_act && s0 + 1 >= 0 && s0 + 1 + PL::N <= a.rxlen;
        // windows that leave the record still run (the stage is zero filled there); their results are dropped below
        const float2* const xw = stage + o + t;
        o += 2;
        // the barriers inside also order the previous pair's bin-minimum atomics before thread 0 reads them below
        block_fft_dual_1buf<PL>(v0, v1, xw, cc + t, tab, buf, t);
        if (t == 0 && it > 0) {                      // results of the previous pair
            const int p0 = 2 * (it - 1);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (p0 + h < n_act) {
                    const unsigned int b = binmin[h];
                    if (out.qf2) out.qf2[i0 + p0 + h] = pend[h];
                    if (out.fidx) out.fidx[i0 + p0 + h] = (b == 0xffffffffu) ? 0u : b;
                }
                binmin[h] = 0xffffffffu;
            }
        }
        // ---- |X|^2, group max of both ----
        float tmax0 = 0.f, tmax1 = 0.f;
#pragma unroll
        for (int j = 0; j < V; ++j) {
            tmax0 = fmaxf(tmax0, cnorm2(v0[j]));
            tmax1 = fmaxf(tmax1, cnorm2(v1[j]));
        }
        // |X|^2 >= +0, so the float order is the order of the bit patterns: one REDUX per warp, one shared atomic
        // per warp into the slot of this iteration's parity (the other slot is cleared for the next iteration)
        unsigned int* const gm = reinterpret_cast<unsigned int*>(wmax) + 2 * (it & 1);
        const unsigned int u0 = __reduce_max_sync(0xffffffffu, __float_as_uint(tmax0));
        const unsigned int u1 = __reduce_max_sync(0xffffffffu, __float_as_uint(tmax1));
        if ((t & 31) == 0) {
            atomicMax(gm, u0);
            atomicMax(gm + 1, u1);
        }
        __syncthreads();   // all last-pass loads done (buffer free) + group maxima complete
        const float m0 = __uint_as_float(gm[0]);
        const float m1 = __uint_as_float(gm[1]);
        if (t < 2) reinterpret_cast<unsigned int*>(wmax)[2 * ((it & 1) ^ 1) + t] = 0u;
        if (ok0 && tmax0 == m0) {
            unsigned int mybin = 0xffffffffu;
#pragma unroll
            for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                for (int j = 0; j < PL::RL; ++j)
                    if (cnorm2(v0[q * PL::RL + j]) == m0) mybin = min(mybin, (unsigned int)output_bin<PL>(t, q, j));
            atomicMin(binmin, mybin);
        }
        if (ok1 && tmax1 == m1) {
            unsigned int mybin = 0xffffffffu;
#pragma unroll
            for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                for (int j = 0; j < PL::RL; ++j)
                    if (cnorm2(v1[q * PL::RL + j]) == m1) mybin = min(mybin, (unsigned int)output_bin<PL>(t, q, j));
            atomicMin(binmin + 1, mybin);
        }
        if (t == 0) {
            const float refn = *out.refnormsq;
            if (e0 < n_act) pend[0] = ok0 ? m0 / refn / out.energy[i0 + e0] : 0.f;
            if (e1 < n_act) pend[1] = ok1 ? m1 / refn / out.energy[i0 + e1] : 0.f;
        }
    }
    __syncthreads();
    if (t == 0 && chunk > 0) {                       // the last pair
        const int p0 = 2 * (chunk - 1);
        for (int h = 0; h < 2; ++h)
            if (p0 + h < n_act) {
                const unsigned int b = binmin[h];
                if (out.qf2) out.qf2[i0 + p0 + h] = pend[h];
                if (out.fidx) out.fidx[i0 + p0 + h] = (b == 0xffffffffu) ? 0u : b;
            }
    }
}

template <class PL>
int32_t launch_fused_staged(const StagedArgs& a, const FusedOut& out, int64_t K, cudaStream_t stream) {
    auto kern = fused_fft_staged_kernel<PL>;
    constexpr size_t smem = StagedLayout<PL>::BYTES;
    if ((int)smem > max_smem_optin()) return fail(XC_ERR_UNSUPPORTED, "staged fused kernel needs %zu B of shared memory", smem);
    static KernelOnce once;
    XC_TRY(once.configure(kern, max_smem_optin()));
    const int64_t pairs = (K + 1) / 2;
    int64_t ctas = (int64_t)sm_count();
    if (ctas > pairs) ctas = pairs;
    if (ctas < 1) ctas = 1;
    const int64_t chunk = (pairs + ctas - 1) / ctas;
    if (chunk >= (1ll << 30)) return fail(XC_ERR_UNSUPPORTED, "too many shifts per thread group");
    profile_begin(stream);
    kern<<<(unsigned)ctas, PL::TG, smem, stream>>>(a, out, K, (int)chunk);
    profile_end(stream);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

}  // namespace xc
