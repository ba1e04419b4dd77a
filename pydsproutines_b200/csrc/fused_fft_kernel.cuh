// fused_fft_kernel.cuh -- the fused per-shift kernel skeleton shared by fastXcorr(freqsearch=True)
// and GroupXcorrFFT: a "loader" fills the registers with the product for one shift, the block FFT
// runs in shared memory, and the epilogue takes |X|^2, normalises by the energies and reduces to
// (max, lowest arg max) -- or writes the CAF row -- without the spectrum ever leaving the SM.
//
// One CTA hosts B independent groups of TG threads; each group walks a contiguous range of the
// shift list so that consecutive windows re-use the same L2 lines.  Three group barriers per FFT.
#pragma once
#include "fft_engine.cuh"

namespace xc {

template <int TG, int B>
struct GroupBar {
    int id;
    __device__ __forceinline__ void operator()() const {
        if (B == 1 || TG < 32) {
            __syncthreads();
        } else {
            asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(TG) : "memory");
        }
    }
};

struct FusedOut {
    float* qf2;          // [K] or NULL
    uint32_t* fidx;      // [K] or NULL
    float2* cplx;        // [K] or NULL: X[bin] / sqrt(scale) (complex QF)
    float* caf;          // [K, N] float (or float2 when caf_complex) or NULL
    int caf_complex;
    const float* energy; // [K] window energies (denominator), indexed like the shift list
    const float* refnormsq;  // device scalar: ||cutout||^2 (or ||ygroups||^2)
};

// Loader concept:
//   __device__ void prepare(int tid_in_cta, int nthreads, char* smem)   -- once per CTA (before a __syncthreads)
//   __device__ bool load(float2* v, int t, int64_t i)                   -- fills v for shift-list entry i,
//                                                                          returns false when the shift is out of range
//   static size_t smem_bytes()
template <class PL, int B, class Loader>
__global__ void __launch_bounds__(PL::TG* B) fused_fft_kernel(Loader loader, FusedOut out, int64_t K, int64_t chunk) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int TG = PL::TG, V = PL::V, NW = (TG + 31) / 32;
    const int tid = threadIdx.x;
    const int g = tid / TG;       // group within the CTA
    const int t = tid % TG;       // thread within the group
    // shared layout: [loader area][B exchange buffers][B * NW warp maxima][B bin minima][B flags]
    unsigned char* p = smem_raw;
    loader.prepare(tid, TG * B, p);
    p += Loader::smem_bytes();
    float2* buf = reinterpret_cast<float2*>(p) + (size_t)g * PL::BUF;
    p += sizeof(float2) * (size_t)PL::BUF * B;
    float* wmax = reinterpret_cast<float*>(p) + g * NW;
    p += sizeof(float) * NW * B;
    unsigned int* binmin = reinterpret_cast<unsigned int*>(p) + g;
    if (t == 0) *binmin = 0xffffffffu;
    __syncthreads();

    FftThread<PL> th;
    th.init(t);
    const GroupBar<TG, B> bar{1 + g};
    const float refn = *out.refnormsq;

    const int64_t gid = (int64_t)blockIdx.x * B + g;
    const int64_t i0 = gid * chunk;
    float2 v[V];
    float pend_max = 0.f;      // result of the previous FFT, written one barrier later (thread 0 only)
    int64_t pend_i = -1;

    for (int64_t it = 0; it < chunk; ++it) {
        const int64_t i = i0 + it;
        const bool active = i < K;
        bool ok = false;
        if (active) ok = loader.load(v, t, i);
        if (!ok) {
#pragma unroll
            for (int j = 0; j < V; ++j) v[j] = make_float2(0.f, 0.f);
        }
        // block FFT; the barrier inside (after the pass-1 stores) also orders the previous
        // iteration's bin-minimum atomics before thread 0 reads them below
        block_fft<PL>(v, th, buf, t, bar);
        if (PL::P == 1) bar();
        if (t == 0 && pend_i >= 0) {
            const unsigned int b = *binmin;
            *binmin = 0xffffffffu;
            if (out.qf2) out.qf2[pend_i] = pend_max;
            if (out.fidx) out.fidx[pend_i] = (b == 0xffffffffu) ? 0u : b;
            pend_i = -1;
        }
        const float denom_e = (active && ok) ? out.energy[i] : 1.f;
        if (out.caf) {
            // ---- CAF row: write all bins (natural order) ----
            if (active) {
                if (out.caf_complex) {
                    float2* row = reinterpret_cast<float2*>(out.caf) + (size_t)i * PL::N;
                    const float sc = ok ? 1.f / (sqrtf(refn) * sqrtf(denom_e)) : 0.f;
#pragma unroll
                    for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                        for (int j = 0; j < PL::RL; ++j) {
                            const float2 x = v[q * PL::RL + j];
                            row[output_bin<PL>(t, q, j)] = make_float2(x.x * sc, x.y * sc);
                        }
                } else {
                    float* row = out.caf + (size_t)i * PL::N;
#pragma unroll
                    for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                        for (int j = 0; j < PL::RL; ++j)
                            row[output_bin<PL>(t, q, j)] = ok ? cnorm2(v[q * PL::RL + j]) / refn / denom_e : 0.f;
                }
            }
            bar();   // buffer free for the next iteration
            continue;
        }
        // ---- |X|^2, group max ----
        float tmax = 0.f;
#pragma unroll
        for (int j = 0; j < V; ++j) tmax = fmaxf(tmax, cnorm2(v[j]));
        float m = tmax;
#pragma unroll
        for (int o = (TG < 32 ? TG : 32) / 2; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (TG > 32 && (t & 31) == 0) wmax[t >> 5] = m;
        bar();   // all last-pass loads done (buffer free) + warp maxima visible
        if (TG > 32) {
            m = wmax[0];
#pragma unroll
            for (int w = 1; w < NW; ++w) m = fmaxf(m, wmax[w]);
        }
        unsigned int mybin = 0xffffffffu;
        if (active && ok && tmax == m) {
#pragma unroll
            for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                for (int j = 0; j < PL::RL; ++j)
                    if (cnorm2(v[q * PL::RL + j]) == m) mybin = min(mybin, (unsigned int)output_bin<PL>(t, q, j));
            atomicMin(binmin, mybin);
        }
        if (out.cplx) {   // rare variant: one more barrier to learn the winning bin
            bar();
            const unsigned int b = *binmin;
            if (active && ok && mybin == b) {
                const float sc = 1.f / (sqrtf(refn) * sqrtf(denom_e));
#pragma unroll
                for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                    for (int j = 0; j < PL::RL; ++j)
                        if ((unsigned int)output_bin<PL>(t, q, j) == b) {
                            const float2 x = v[q * PL::RL + j];
                            out.cplx[i] = make_float2(x.x * sc, x.y * sc);
                        }
            }
            if (active && !ok && t == 0) out.cplx[i] = make_float2(0.f, 0.f);
            bar();
        }
        if (t == 0 && active) {
            pend_i = i;
            pend_max = ok ? m / refn / denom_e : 0.f;
        }
    }
    bar();
    if (t == 0 && pend_i >= 0) {
        const unsigned int b = *binmin;
        if (out.qf2) out.qf2[pend_i] = pend_max;
        if (out.fidx) out.fidx[pend_i] = (b == 0xffffffffu) ? 0u : b;
    }
}

template <class PL, int B, class Loader>
size_t fused_smem_bytes() {
    constexpr int NW = (PL::TG + 31) / 32;
    return Loader::smem_bytes() + sizeof(float2) * (size_t)PL::BUF * B + sizeof(float) * NW * B + sizeof(unsigned int) * B + 16;
}

// Launches the fused kernel as a persistent grid: as many CTAs as fit on the device (or fewer when
// K is small), each group taking `chunk` consecutive shift-list entries.
template <class PL, int B, class Loader>
int32_t launch_fused(const Loader& loader, const FusedOut& out, int64_t K, cudaStream_t stream) {
    auto kern = fused_fft_kernel<PL, B, Loader>;
    const size_t smem = fused_smem_bytes<PL, B, Loader>();
    if ((int)smem > max_smem_optin()) return fail(XC_ERR_UNSUPPORTED, "fused kernel needs %zu B of shared memory", smem);
    static bool configured = false;   // per instantiation
    static int occ = 1;
    if (!configured) {
        XC_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        XC_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, PL::TG * B, smem));
        if (occ < 1) return fail(XC_ERR_CUDA, "fused kernel does not fit on an SM (smem %zu B)", smem);
        configured = true;
    }
    const int64_t groups_needed = K;
    int64_t max_ctas = (int64_t)sm_count() * occ;
    int64_t ctas = (groups_needed + B - 1) / B;
    if (ctas > max_ctas) ctas = max_ctas;
    if (ctas < 1) ctas = 1;
    const int64_t chunk = (K + ctas * B - 1) / (ctas * B);
    profile_begin(stream);
    kern<<<(unsigned)ctas, PL::TG * B, smem, stream>>>(loader, out, K, chunk);
    profile_end(stream);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

}  // namespace xc
