// bluestein.cuh -- frequency search for cutout lengths that are not a power of two (the reference takes any
// length: xcorrRoutines.py:511-531 calls scipy's mixed-radix FFT).  Chirp-z / Bluestein inside the same fused
// per-shift kernel: with b[n] = exp(+i pi n^2 / N),
//     X[k] = conj(b[k]) * sum_n (p[n] conj(b[n])) b[k - n],        p[n] = rx[s + n] conj(cutout[n])
// i.e. a length-N linear convolution, done as a circular one of power-of-two length M >= 2N - 1:
//     forward block FFT of a[n] = p[n] conj(b[n]) (zero padded), times FFT_M(b) (precomputed, laid out for the
//     thread mapping), back through a second FORWARD block FFT of the conjugate (IFFT(y) = conj(FFT(conj(y))) / M),
// and |X[k]|^2 = |.|^2 / M^2 for k < N -- the unit-modulus factor conj(b[k]) does not change the magnitude.
// The chirp conj(b[n]) is folded into the cutout copy held in shared memory, so the loader is the ordinary one.
#pragma once
#include "fused_fft_kernel.cuh"

namespace xc {

// cell i of the reorder exchange (between the two FFTs) lives at i + 2 * (i / 32): fits the block FFT's buffer
__device__ __forceinline__ int blue_cell(int i) { return i + ((i >> 5) << 1); }

// Loader concept as in fused_fft_kernel.cuh, plus: load() zero-fills the points n >= nvalid itself.
template <class PL, int B, class Loader>
__global__ void __launch_bounds__(PL::TG* B)
fused_bluestein_kernel(const Loader loader, const FusedOut out, const int64_t K, const int chunk,
                       const float2* __restrict__ bhat_perm, const float2* __restrict__ chirp, const int nbins,
                       const float inv_m2) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int TG = PL::TG, V = PL::V, NW = (TG + 31) / 32;
    static_assert(PL::P >= 2, "the Bluestein path needs an exchange buffer");
    constexpr size_t OFF_TAB = Loader::smem_bytes();
    constexpr size_t OFF_BUF = OFF_TAB + PL::TABLE_BYTES;
    constexpr size_t OFF_WMAX = OFF_BUF + sizeof(float2) * (size_t)PL::BUF * B;
    constexpr size_t OFF_BIN = OFF_WMAX + sizeof(float) * NW * B;
    constexpr size_t OFF_PEND = OFF_BIN + sizeof(unsigned int) * B;
    const int g = threadIdx.x / TG, t = threadIdx.x % TG;
    loader.prepare(threadIdx.x, TG * B, smem_raw);
    const float2* const tab = reinterpret_cast<const float2*>(smem_raw + OFF_TAB);
    build_fft_tables<PL>(reinterpret_cast<float2*>(smem_raw + OFF_TAB), threadIdx.x, TG * B);
    float2* const buf = reinterpret_cast<float2*>(smem_raw + OFF_BUF) + (size_t)g * PL::BUF;
    float* const wmax = reinterpret_cast<float*>(smem_raw + OFF_WMAX) + g * NW;
    unsigned int* const binmin = reinterpret_cast<unsigned int*>(smem_raw + OFF_BIN) + g;
    float* const pend = reinterpret_cast<float*>(smem_raw + OFF_PEND) + g;
    if (t == 0) *binmin = 0xffffffffu;
    __syncthreads();
    if constexpr (loader_uses_tmem<Loader>::value) {
        loader.prepare2(threadIdx.x, smem_raw);
        __syncthreads();
    }

    FftThread<PL> th;
    th.init(t);
    const GroupBar<TG, B> bar{1 + g};
    const int64_t i0 = ((int64_t)blockIdx.x * B + g) * (int64_t)chunk;
    int n_act = 0;
    if (i0 < K) n_act = (K - i0 < (int64_t)chunk) ? (int)(K - i0) : chunk;
    float2 v[V];

    for (int it = 0; it < chunk; ++it) {
        const bool active = it < n_act;
        bool ok = false;
        if (active) ok = loader.load(v, t, i0 + it, smem_raw);
        if (!ok) {
#pragma unroll
            for (int j = 0; j < V; ++j) v[j] = make_float2(0.f, 0.f);
        }
        // ---- FFT #1 of the chirped product ----
        block_fft<PL>(v, th, tab, buf, t, bar);
        if (t == 0 && it > 0 && it <= n_act) {     // results of the previous shift (ordered by the barrier above)
            const unsigned int b = *binmin;
            *binmin = 0xffffffffu;
            if (out.qf2) out.qf2[i0 + it - 1] = *pend;
            if (out.fidx) out.fidx[i0 + it - 1] = (b == 0xffffffffu) ? 0u : b;
        }
        // ---- times FFT_M(b), conjugated for the inverse transform ----
#pragma unroll
        for (int s = 0; s < V; ++s) {
            const float2 w = __ldg(bhat_perm + (size_t)s * TG + t);
            const float2 p = pcmul(v[s], w);
            v[s] = make_float2(p.x, -p.y);
        }
        bar();    // every thread has finished the last-pass loads of FFT #1: the buffer is free
        // ---- reorder: from output bins to the input order of the second FFT ----
#pragma unroll
        for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
            for (int j = 0; j < PL::RL; ++j) buf[blue_cell(output_bin<PL>(t, q, j))] = v[q * PL::RL + j];
        bar();
        constexpr int NQ = V / PL::R1;
#pragma unroll
        for (int n1 = 0; n1 < PL::R1; ++n1)
#pragma unroll
            for (int q = 0; q < NQ; ++q) v[q * PL::R1 + n1] = buf[blue_cell(input_index<PL>(t, q, n1))];
        bar();    // all reads done before the pass-1 stores of FFT #2
        // ---- FFT #2 ----
        block_fft<PL>(v, th, tab, buf, t, bar);
        const unsigned int bbase = (unsigned int)output_bin_base<PL>(t);
        if (out.caf) {
            // ---- CAF row (natural order, nbins columns): |X[k]|^2 scaled, or X[k] = conj(b[k]) conj(v) / M ----
            if (active) {
                const float refn = *out.refnormsq;
                const float denom_e = ok ? out.energy[i0 + it] : 1.f;
                const float sc2 = ok ? inv_m2 / refn / denom_e : 0.f;
                const float sc1 = ok ? sqrtf(inv_m2) / (sqrtf(refn) * sqrtf(denom_e)) : 0.f;
#pragma unroll
                for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                    for (int j = 0; j < PL::RL; ++j) {
                        const unsigned int bin = bbase + (unsigned int)output_bin_off<PL>(q, j);
                        if (bin < (unsigned int)nbins) {
                            const float2 x = v[q * PL::RL + j];
                            if (out.caf_complex) {
                                const float2 b = __ldg(chirp + bin);
                                // conj(b) * conj(x) = conj(b x)
                                const float2 p = cmul(b, x);
                                reinterpret_cast<float2*>(out.caf)[(size_t)(i0 + it) * nbins + bin] = make_float2(p.x * sc1, -p.y * sc1);
                            } else {
                                out.caf[(size_t)(i0 + it) * nbins + bin] = cnorm2(x) * sc2;
                            }
                        }
                    }
            }
            bar();   // buffer free for the next iteration
            continue;
        }
        // ---- |X[k]|^2 for k < N, group max, lowest arg max ----
        float tmax = -1.f;
#pragma unroll
        for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
            for (int j = 0; j < PL::RL; ++j)
                if (bbase + (unsigned int)output_bin_off<PL>(q, j) < (unsigned int)nbins)
                    tmax = fmaxf(tmax, cnorm2(v[q * PL::RL + j]));
        float m = tmax;
#pragma unroll
        for (int o = (TG < 32 ? TG : 32) / 2; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (TG > 32 && (t & 31) == 0) wmax[t >> 5] = m;
        bar();
        if (TG > 32) {
            m = wmax[0];
#pragma unroll
            for (int w = 1; w < NW; ++w) m = fmaxf(m, wmax[w]);
        }
        if (ok && tmax == m) {
            unsigned int mybin = 0xffffffffu;
#pragma unroll
            for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                for (int j = 0; j < PL::RL; ++j) {
                    const unsigned int bin = bbase + (unsigned int)output_bin_off<PL>(q, j);
                    if (bin < (unsigned int)nbins && cnorm2(v[q * PL::RL + j]) == m) mybin = min(mybin, bin);
                }
            atomicMin(binmin, mybin);
        }
        if (out.cplx) {   // complex QF at the winning bin: one more barrier to learn it
            bar();
            const unsigned int b = *binmin;
            if (ok && tmax == m) {
                const float sc = sqrtf(inv_m2) / (sqrtf(*out.refnormsq) * sqrtf(out.energy[i0 + it]));
#pragma unroll
                for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                    for (int j = 0; j < PL::RL; ++j)
                        if (bbase + (unsigned int)output_bin_off<PL>(q, j) == b) {
                            const float2 p = cmul(__ldg(chirp + b), v[q * PL::RL + j]);
                            out.cplx[i0 + it] = make_float2(p.x * sc, -p.y * sc);
                        }
            }
            if (active && !ok && t == 0) out.cplx[i0 + it] = make_float2(0.f, 0.f);
            bar();
        }
        if (t == 0 && active) *pend = ok ? m * inv_m2 / *out.refnormsq / out.energy[i0 + it] : 0.f;
    }
    bar();
    if (t == 0 && n_act == chunk && !out.caf) {
        const unsigned int b = *binmin;
        if (out.qf2) out.qf2[i0 + chunk - 1] = *pend;
        if (out.fidx) out.fidx[i0 + chunk - 1] = (b == 0xffffffffu) ? 0u : b;
    }
    if constexpr (loader_uses_tmem<Loader>::value) {
        __syncthreads();
        loader.release(threadIdx.x, smem_raw);
    }
}

// Persistent launch, as launch_fused_variant(): `nbins` = the true transform length N, PL::N = the padded length M.
template <class PL, int B, class Loader>
int32_t launch_bluestein(const Loader& ld, const FusedOut& out, int64_t K, const float2* bhat_perm, const float2* chirp,
                         int nbins, cudaStream_t stream) {
    auto kern = fused_bluestein_kernel<PL, B, Loader>;
    const size_t dyn = ld.dyn_smem();
    const size_t smem = fused_smem_bytes<PL, B, Loader>() + dyn;
    if ((int)smem > max_smem_optin()) return fail(XC_ERR_UNSUPPORTED, "Bluestein kernel needs %zu B of shared memory", smem);
    static KernelOnce once;           // per instantiation, flags per device
    int first = 0;
    XC_TRY(once.configure(kern, max_smem_optin(), &first));
    int& occ0 = once.occupancy();
    if (first) {
        XC_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ0, kern, PL::TG * B, fused_smem_bytes<PL, B, Loader>()));
        if (occ0 < 1) return fail(XC_ERR_CUDA, "Bluestein kernel does not fit on an SM (smem %zu B)", smem);
    }
    int occ = occ0;
    if (dyn > 0) {
        XC_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, PL::TG * B, smem));
        if (occ < 1) return fail(XC_ERR_CUDA, "Bluestein kernel does not fit on an SM (smem %zu B)", smem);
    }
    const int64_t max_ctas = (int64_t)sm_count() * occ;
    int64_t ctas = (K + B - 1) / B;
    if (ctas > max_ctas) ctas = max_ctas;
    if (ctas < 1) ctas = 1;
    const int64_t chunk = (K + ctas * B - 1) / (ctas * B);
    if (chunk >= (1ll << 31)) return fail(XC_ERR_UNSUPPORTED, "too many shifts per thread group");
    const double m = (double)PL::N;
    profile_begin(stream);
    kern<<<(unsigned)ctas, PL::TG * B, smem, stream>>>(ld, out, K, (int)chunk, bhat_perm, chirp, nbins, (float)(1.0 / (m * m)));
    profile_end(stream);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

}  // namespace xc
