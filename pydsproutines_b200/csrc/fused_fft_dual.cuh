// fused_fft_dual.cuh -- variant of the fused per-shift kernel in which every thread carries TWO FFTs (two
// consecutive entries of the shift list) through the three passes in lockstep.  What is shared between the two:
// the conj(cutout) value of every product, the pass-1 twiddle power chains and the pass-2 twiddle table loads
// -- a quarter of the shared-memory wavefronts of fused_fft_kernel -- and the two independent dependency chains
// give the scheduler the instruction-level parallelism that a second resident thread group gives the B = 2 layout.
// One CTA = one group of TG threads, two exchange buffers.  Three-pass plans with a warp-local second exchange
// only, (max, arg max) outputs only.
#pragma once
#include "fused_fft_kernel.cuh"

namespace xc {

template <int R>
__device__ __forceinline__ void twiddle_powers2(float2* a, float2* b, float2 w) {
    float2 p = w;
#pragma unroll
    for (int k = 1; k < R; ++k) {
        const int j = bitrev(k, ilog2(R));
        a[j] = pcmul(a[j], p);
        b[j] = pcmul(b[j], p);
        if (k + 1 < R) p = pcmul(p, w);
    }
}

template <class PL>
__device__ __forceinline__ void block_fft_dual(float2* v0, float2* v1, const float2* tab, float2* buf0, float2* buf1, int t) {
    static_assert(PL::P == 3 && PL::WARP_LOCAL_X2, "dual layout: three passes, warp-local second exchange");
    constexpr int RL = PL::RL, ROW = PL::ROW;
    // ---- pass 1 ----
#pragma unroll
    for (int q = 0; q < PL::V / PL::R1; ++q) {
        codelet_fft<PL::R1>(v0 + q * PL::R1);
        codelet_fft<PL::R1>(v1 + q * PL::R1);
    }
    {
        constexpr int K1STEP = (PL::M1 / RL) * ROW;
        constexpr int QSTEP = (PL::TG / RL) * ROW;
        const int o1 = (t / RL) * ROW + (t % RL);
#pragma unroll
        for (int q = 0; q < PL::V / PL::R1; ++q) {
            const float2 w = tab[t + PL::TG * q];
            twiddle_powers2<PL::R1>(v0 + q * PL::R1, v1 + q * PL::R1, w);
#pragma unroll
            for (int j = 0; j < PL::R1; ++j) {
                buf0[o1 + bitrev(j, ilog2(PL::R1)) * K1STEP + q * QSTEP] = v0[q * PL::R1 + j];
                buf1[o1 + bitrev(j, ilog2(PL::R1)) * K1STEP + q * QSTEP] = v1[q * PL::R1 + j];
            }
        }
    }
    __syncthreads();
    {
        // ---- pass 2 (in place) ----
        constexpr int K1STEP = (PL::M1 / RL) * ROW;
        constexpr int QSTEP = (PL::TG / PL::M2) * K1STEP;
        const int o2 = (t / PL::M2) * K1STEP + (t % PL::M2);
        const float4* const t2 = reinterpret_cast<const float4*>(tab + PL::T1_CELLS + (t % PL::M2) * PL::T2_ROW);
#pragma unroll
        for (int q = 0; q < PL::V / PL::R2; ++q)
#pragma unroll
            for (int n2 = 0; n2 < PL::R2; ++n2) {
                v0[q * PL::R2 + n2] = buf0[o2 + q * QSTEP + n2 * ROW];
                v1[q * PL::R2 + n2] = buf1[o2 + q * QSTEP + n2 * ROW];
            }
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
#pragma unroll
            for (int j = 0; j < PL::R2; ++j) {
                buf0[o2 + q * QSTEP + bitrev(j, ilog2(PL::R2)) * ROW] = v0[q * PL::R2 + j];
                buf1[o2 + q * QSTEP + bitrev(j, ilog2(PL::R2)) * ROW] = v1[q * PL::R2 + j];
            }
        }
        __syncwarp();
    }
    // ---- last pass ----
    {
        const float4* const b30 = reinterpret_cast<const float4*>(buf0 + t * ROW);
        const float4* const b31 = reinterpret_cast<const float4*>(buf1 + t * ROW);
#pragma unroll
        for (int q = 0; q < PL::V / RL; ++q)
#pragma unroll
            for (int d2 = 0; d2 < RL / 2; ++d2) {
                const float4 x = b30[(q * PL::TG * ROW) / 2 + d2];
                v0[q * RL + 2 * d2] = make_float2(x.x, x.y);
                v0[q * RL + 2 * d2 + 1] = make_float2(x.z, x.w);
                const float4 y = b31[(q * PL::TG * ROW) / 2 + d2];
                v1[q * RL + 2 * d2] = make_float2(y.x, y.y);
                v1[q * RL + 2 * d2 + 1] = make_float2(y.z, y.w);
            }
    }
#pragma unroll
    for (int q = 0; q < PL::V / RL; ++q) {
        codelet_fft<RL>(v0 + q * RL);
        codelet_fft<RL>(v1 + q * RL);
    }
}

// Loader concept additions: issue(v, t, i, smem) -> bool and finish2(v0, v1, t, smem) (both windows valid) / finish.
template <class PL, class Loader>
__global__ void __launch_bounds__(PL::TG) fused_fft_dual_kernel(const Loader loader, const FusedOut out, const int64_t K,
                                                                const int chunk) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int TG = PL::TG, V = PL::V, NW = TG / 32;
    constexpr size_t OFF_TAB = Loader::smem_bytes();
    constexpr size_t OFF_BUF = OFF_TAB + PL::TABLE_BYTES;
    constexpr size_t OFF_WMAX = OFF_BUF + sizeof(float2) * (size_t)PL::BUF * 2;
    constexpr size_t OFF_BIN = OFF_WMAX + sizeof(float) * NW * 2;
    constexpr size_t OFF_PEND = OFF_BIN + sizeof(unsigned int) * 2;
    const int t = threadIdx.x;
    loader.prepare(t, TG, smem_raw);
    const float2* const tab = reinterpret_cast<const float2*>(smem_raw + OFF_TAB);
    build_fft_tables<PL>(reinterpret_cast<float2*>(smem_raw + OFF_TAB), t, TG);
    float2* const buf0 = reinterpret_cast<float2*>(smem_raw + OFF_BUF);
    float2* const buf1 = buf0 + PL::BUF;
    float* const wmax = reinterpret_cast<float*>(smem_raw + OFF_WMAX);            // [2][NW]
    unsigned int* const binmin = reinterpret_cast<unsigned int*>(smem_raw + OFF_BIN);
    float* const pend = reinterpret_cast<float*>(smem_raw + OFF_PEND);
    if (t < 2) binmin[t] = 0xffffffffu;
    __syncthreads();

    // this CTA owns shift-list entries [i0, i0 + 2 * chunk), taken in pairs
    const int64_t i0 = (int64_t)blockIdx.x * 2 * (int64_t)chunk;
    int64_t n_left = K - i0;
    if (n_left < 0) n_left = 0;
    if (n_left > 2 * (int64_t)chunk) n_left = 2 * (int64_t)chunk;
    const int n_act = (int)n_left;               // active entries of this CTA
    float2 v0[V], v1[V];

    for (int it = 0; it < chunk; ++it) {
        const int e0 = 2 * it, e1 = 2 * it + 1;
        bool ok0 = false, ok1 = false;
        if (e0 < n_act) ok0 = loader.issue(v0, t, i0 + e0, smem_raw);
        if (e1 < n_act) ok1 = loader.issue(v1, t, i0 + e1, smem_raw);
        if (ok0 && ok1) {
            loader.finish2(v0, v1, t, smem_raw);
        } else {
            if (ok0) loader.finish(v0, t, smem_raw);
            if (ok1) loader.finish(v1, t, smem_raw);
        }
        if (!ok0) {
#pragma unroll
            for (int j = 0; j < V; ++j) v0[j] = make_float2(0.f, 0.f);
        }
        if (!ok1) {
#pragma unroll
            for (int j = 0; j < V; ++j) v1[j] = make_float2(0.f, 0.f);
        }
        // the barrier inside (after the pass-1 stores) also orders the previous iteration's bin-minimum atomics
        // before thread 0 reads them below
        block_fft_dual<PL>(v0, v1, tab, buf0, buf1, t);
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
        float m0 = tmax0, m1 = tmax1;
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, o));
            m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, o));
        }
        if ((t & 31) == 0) {
            wmax[t >> 5] = m0;
            wmax[NW + (t >> 5)] = m1;
        }
        __syncthreads();   // all last-pass loads done (buffers free) + warp maxima visible
        m0 = wmax[0];
        m1 = wmax[NW];
#pragma unroll
        for (int w = 1; w < NW; ++w) {
            m0 = fmaxf(m0, wmax[w]);
            m1 = fmaxf(m1, wmax[NW + w]);
        }
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

template <class PL, class Loader>
constexpr size_t fused_dual_smem_bytes() {
    return Loader::smem_bytes() + PL::TABLE_BYTES + sizeof(float2) * (size_t)PL::BUF * 2 + sizeof(float) * (PL::TG / 32) * 2 +
           sizeof(unsigned int) * 2 + sizeof(float) * 2 + 16;
}

template <class PL, class Loader>
int32_t launch_fused_dual(const Loader& loader, const FusedOut& out, int64_t K, cudaStream_t stream) {
    auto kern = fused_fft_dual_kernel<PL, Loader>;
    const size_t smem = fused_dual_smem_bytes<PL, Loader>();
    if ((int)smem > max_smem_optin()) return fail(XC_ERR_UNSUPPORTED, "dual fused kernel needs %zu B of shared memory", smem);
    static KernelOnce once;
    int first = 0;
    XC_TRY(once.configure(kern, max_smem_optin(), &first));
    int& occ = once.occupancy();
    if (first) {
        XC_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, PL::TG, smem));
        if (occ < 1) return fail(XC_ERR_CUDA, "dual fused kernel does not fit on an SM (smem %zu B)", smem);
    }
    const int64_t pairs = (K + 1) / 2;
    int64_t ctas = (int64_t)sm_count() * occ;
    if (ctas > pairs) ctas = pairs;
    if (ctas < 1) ctas = 1;
    const int64_t chunk = (pairs + ctas - 1) / ctas;
    if (chunk >= (1ll << 30)) return fail(XC_ERR_UNSUPPORTED, "too many shifts per thread group");
    profile_begin(stream);
    kern<<<(unsigned)ctas, PL::TG, smem, stream>>>(loader, out, K, (int)chunk);
    profile_end(stream);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

}  // namespace xc
