// fused_fft_kernel.cuh -- the fused per-shift kernel skeleton shared by fastXcorr(freqsearch=True)
// and GroupXcorrFFT: a "loader" fills the registers with the product for one shift, the block FFT
// runs in shared memory, and the epilogue takes |X|^2, normalises by the energies and reduces to
// (max, lowest arg max) -- or writes the CAF row -- without the spectrum ever leaving the SM.
//
// One CTA hosts B independent groups of TG threads; each group walks a contiguous range of the
// shift list so that consecutive windows re-use the same L2 lines.  Three group barriers per FFT.
// Register discipline (the 8192-point plan runs 512 threads x 128 registers): nothing but the 32
// points, the base twiddles and a few 32-bit counters lives across an iteration; kernel arguments
// stay in the constant bank, pending results stay in shared memory, and the rarely used outputs
// (complex QF, CAF rows) are compiled into a separate GENERAL instantiation.
#pragma once
#include <type_traits>

#include "fft_engine.cuh"

namespace xc {

// loaders that keep their constants in tensor memory declare USES_TMEM and add prepare2() (after the first CTA barrier: the
// allocation is visible) and release() (after the last one)
template <class T, class = void>
struct loader_uses_tmem : std::false_type {};
template <class T>
struct loader_uses_tmem<T, std::void_t<decltype(T::USES_TMEM)>> : std::bool_constant<T::USES_TMEM> {};

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
    float2* cplx;        // [K] or NULL: X[bin] / sqrt(scale) (complex QF)            (GENERAL only)
    float* caf;          // [K, N] float (or float2 when caf_complex) or NULL          (GENERAL only)
    int caf_complex;
    const float* energy; // [K] window energies (denominator), indexed like the shift list
    const float* refnormsq;  // device scalar: ||cutout||^2 (or ||ygroups||^2)
};

// arguments of the layouts that search a CONTIGUOUS shift range from a staged window (fused_fft_pp.cuh)
struct StagedArgs {
    const float2* rx;
    int64_t rxlen;
    const float2* cutout;
    int64_t shift0;       // the shift of entry i is shift0 + i * step
    int step = 1;         // 1 <= step, (B - 1) step <= the window slack of the layout
};

// Loader concept (all const; the kernel argument is never copied into registers):
//   static constexpr size_t smem_bytes()
//   __device__ void prepare(int tid_in_cta, int nthreads, unsigned char* smem) const   -- once per CTA
//   __device__ bool load(float2* v, int t, int64_t i, const unsigned char* smem) const -- fills v for shift-list
//                                                    entry i, returns false when the shift is out of range
template <class PL, int B, class Loader, bool GENERAL>
__global__ void __launch_bounds__(PL::TG* B) fused_fft_kernel(const Loader loader, const FusedOut out, const int64_t K,
                                                              const int chunk) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int TG = PL::TG, V = PL::V, NW = (TG + 31) / 32;
    constexpr size_t OFF_TAB = Loader::smem_bytes();
    constexpr size_t OFF_BUF = OFF_TAB + PL::TABLE_BYTES;
    constexpr size_t OFF_WMAX = OFF_BUF + sizeof(float2) * (size_t)PL::BUF * B;
    constexpr size_t OFF_BIN = OFF_WMAX + sizeof(float) * NW * B;
    constexpr size_t OFF_PEND = OFF_BIN + sizeof(unsigned int) * B;     // B floats: pending QF^2 of the previous FFT
    const int g = threadIdx.x / TG;       // group within the CTA
    const int t = threadIdx.x % TG;       // thread within the group
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

    // first shift-list entry of this group; entries [i0, i0 + chunk) clipped to K
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
        // block FFT; the barrier inside (after the pass-1 stores) also orders the previous
        // iteration's bin-minimum atomics before thread 0 reads them below
        block_fft<PL>(v, th, tab, buf, t, bar);
        if (PL::P == 1) bar();
        if (t == 0 && it > 0 && it <= n_act) {     // results of the previous FFT
            const unsigned int b = *binmin;
            *binmin = 0xffffffffu;
            if (out.qf2) out.qf2[i0 + it - 1] = *pend;
            if (out.fidx) out.fidx[i0 + it - 1] = (b == 0xffffffffu) ? 0u : b;
        }
        if (GENERAL && out.caf) {
            // ---- CAF row: all bins in natural order.  The registers hold bins at a stride of R1 across the lanes, so the row
            // goes through the (now free) exchange buffer and leaves as coalesced runs (4-byte stores at a stride of R1
            // floats reached 0.13 of the HBM write rate) ----
            if (PL::P <= 2) {
                // one- and two-pass plans: lane t holds bins t + TG q + R1 k, i.e. the lanes already write consecutive bins
                if (active) {
                    const float refn = *out.refnormsq;
                    const float denom_e = ok ? out.energy[i0 + it] : 1.f;
                    const float sc = ok ? 1.f / (sqrtf(refn) * sqrtf(denom_e)) : 0.f;
                    const float sc2 = ok ? (1.f / refn) / denom_e : 0.f;
#pragma unroll
                    for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                        for (int j = 0; j < PL::RL; ++j) {
                            const float2 x = v[q * PL::RL + j];
                            const size_t o = (size_t)(i0 + it) * PL::N + output_bin<PL>(t, q, j);
                            if (out.caf_complex) reinterpret_cast<float2*>(out.caf)[o] = make_float2(x.x * sc, x.y * sc);
                            else out.caf[o] = cnorm2(x) * sc2;
                        }
                }
            } else {
                bar();   // every thread has finished the last-pass loads: the buffer is free
                const float refn = *out.refnormsq;
                const float denom_e = (active && ok) ? out.energy[i0 + it] : 1.f;
                if (out.caf_complex) {
                    const float sc = ok ? 1.f / (sqrtf(refn) * sqrtf(denom_e)) : 0.f;
#pragma unroll
                    for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                        for (int j = 0; j < PL::RL; ++j) {
                            const int bn = output_bin<PL>(t, q, j);
                            const float2 x = v[q * PL::RL + j];
                            buf[bn + ((bn / PL::RL) << 1)] = make_float2(x.x * sc, x.y * sc);      // rows of RL cells are ROW = RL + 2 apart
                        }
                    bar();
                    if (active) {
                        float2* row = reinterpret_cast<float2*>(out.caf) + (size_t)(i0 + it) * PL::N;
#pragma unroll
                        for (int i = 0; i < V; ++i) {
                            const int bn = i * TG + t;
                            row[bn] = buf[bn + ((bn / PL::RL) << 1)];
                        }
                    }
                } else {
                    float* const fb = reinterpret_cast<float*>(buf);
                    const float sc2 = ok ? (1.f / refn) / denom_e : 0.f;       // one reciprocal per row, not two divisions per bin
#pragma unroll
                    for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                        for (int j = 0; j < PL::RL; ++j) {
                            const int bn = output_bin<PL>(t, q, j);
                            fb[bn + (bn >> 5)] = cnorm2(v[q * PL::RL + j]) * sc2;
                        }
                    bar();
                    if (active) {
                        float* row = out.caf + (size_t)(i0 + it) * PL::N;
#pragma unroll
                        for (int i = 0; i < V; ++i) {
                            const int bn = i * TG + t;
                            row[bn] = fb[bn + (bn >> 5)];
                        }
                    }
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
        if (ok && tmax == m) {
#pragma unroll
            for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                for (int j = 0; j < PL::RL; ++j)
                    if (cnorm2(v[q * PL::RL + j]) == m) mybin = min(mybin, (unsigned int)output_bin<PL>(t, q, j));
            atomicMin(binmin, mybin);
        }
        if (GENERAL && out.cplx) {   // rare variant: one more barrier to learn the winning bin
            bar();
            const unsigned int b = *binmin;
            if (ok && mybin == b) {
                const float sc = 1.f / (sqrtf(*out.refnormsq) * sqrtf(out.energy[i0 + it]));
#pragma unroll
                for (int q = 0; q < V / PL::RL; ++q)
#pragma unroll
                    for (int j = 0; j < PL::RL; ++j)
                        if ((unsigned int)output_bin<PL>(t, q, j) == b) {
                            const float2 x = v[q * PL::RL + j];
                            out.cplx[i0 + it] = make_float2(x.x * sc, x.y * sc);
                        }
            }
            if (active && !ok && t == 0) out.cplx[i0 + it] = make_float2(0.f, 0.f);
            bar();
        }
        if (t == 0 && active) *pend = ok ? m / *out.refnormsq / out.energy[i0 + it] : 0.f;
    }
    bar();
    if (t == 0 && n_act == chunk && !(GENERAL && out.caf)) {   // the last FFT of a full chunk
        const unsigned int b = *binmin;
        if (out.qf2) out.qf2[i0 + chunk - 1] = *pend;
        if (out.fidx) out.fidx[i0 + chunk - 1] = (b == 0xffffffffu) ? 0u : b;
    }
    if constexpr (loader_uses_tmem<Loader>::value) {
        __syncthreads();
        loader.release(threadIdx.x, smem_raw);
    }
}

template <class PL, int B, class Loader>
constexpr size_t fused_smem_bytes() {
    constexpr int NW = (PL::TG + 31) / 32;
    return Loader::smem_bytes() + PL::TABLE_BYTES + sizeof(float2) * (size_t)PL::BUF * B + sizeof(float) * NW * B +
           sizeof(unsigned int) * B + sizeof(float) * B + 16;
}

template <class PL, int B, class Loader, bool GENERAL>
int32_t launch_fused_variant(const Loader& loader, const FusedOut& out, int64_t K, cudaStream_t stream) {
    auto kern = fused_fft_kernel<PL, B, Loader, GENERAL>;
    // static carve-up + whatever the loader keeps resident behind it (Loader::dyn_smem(), e.g. the group weights)
    const size_t dyn = loader.dyn_smem();
    const size_t smem = fused_smem_bytes<PL, B, Loader>() + dyn;
    if ((int)smem > max_smem_optin()) return fail(XC_ERR_UNSUPPORTED, "fused kernel needs %zu B of shared memory", smem);
    static KernelOnce once;           // per instantiation, flags per device
    int first = 0;
    XC_TRY(once.configure(kern, max_smem_optin(), &first));
    int& occ0 = once.occupancy();
    if (first) {
        XC_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ0, kern, PL::TG * B, fused_smem_bytes<PL, B, Loader>()));
        if (occ0 < 1) return fail(XC_ERR_CUDA, "fused kernel does not fit on an SM (smem %zu B)", smem);
    }
    int occ = occ0;
    if (dyn > 0) {
        XC_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, PL::TG * B, smem));
        if (occ < 1) return fail(XC_ERR_CUDA, "fused kernel does not fit on an SM (smem %zu B)", smem);
    }
    const int64_t max_ctas = (int64_t)sm_count() * occ;
    int64_t ctas = (K + B - 1) / B;
    if (ctas > max_ctas) ctas = max_ctas;
    if (ctas < 1) ctas = 1;
    const int64_t chunk = (K + ctas * B - 1) / (ctas * B);
    if (chunk >= (1ll << 31)) return fail(XC_ERR_UNSUPPORTED, "too many shifts per thread group");
    profile_begin(stream);
    kern<<<(unsigned)ctas, PL::TG * B, smem, stream>>>(loader, out, K, (int)chunk);
    profile_end(stream);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

// Launches the fused kernel as a persistent grid: as many CTAs as fit on the device (or fewer when
// K is small), each group taking `chunk` consecutive shift-list entries.
template <class PL, int B, class Loader>
int32_t launch_fused(const Loader& loader, const FusedOut& out, int64_t K, cudaStream_t stream) {
    if (out.caf || out.cplx) return launch_fused_variant<PL, B, Loader, true>(loader, out, K, stream);
    return launch_fused_variant<PL, B, Loader, false>(loader, out, K, stream);
}

}  // namespace xc
