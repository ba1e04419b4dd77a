// group.cu -- GroupXcorrFFT on the B200 (xcorrRoutines.py:1047-1262, GroupXcorrFFT.cpp:3-199).
//
// The reference does G zero-padded F-point FFTs per shift, rotates each by the group's start phase
// and sums them.  Because the starts are integers,
//     sum_g exp(-j 2 pi k s_g / F) FFT_F(p_g)[k]  ==  FFT_F(z)[k],   z[m] = sum_{(g,n): (s_g+n) mod F = m} p_g[n]
// so ONE F-point FFT of the products folded (time-aliased) into F cells gives the same surface.
// The fold is precomputed at create time as padded per-register-slot contributor lists laid out for
// the block FFT's thread mapping ("ELL by slot"): slot c of thread t sums
//     y[(base_c + j)*TG + t] * rx[shift + off[(base_c + j)*TG + t]],  j < len_c
// -- uniform trip counts inside a group of threads, coalesced loads, no atomics.
#include <vector>

#include <math.h>

#include "bluestein.cuh"
#include "fused_fft_kernel.cuh"
#include "internal.cuh"
#include "plans.cuh"

struct xc_group {
    int32_t G, L, F;
    int64_t span;            // max(starts) + L: samples needed after the shift
    int64_t min_start;
    int64_t* d_starts;       // [G]
    float2* d_ell_y;         // [rows * TG]
    int32_t* d_ell_off;      // [rows * TG]
    int32_t* d_slot_base;    // [V + 1]
    int32_t* d_row_base;     // [rows] when every contributor row is a run off = row_base + t, else NULL
    int32_t runs_inside;     // every row_base + t (t < TG) is a valid sample offset: no clamping needed
    float* d_refn;           // ||ygroups||^2
    int32_t rows;
    int32_t log2m;           // block-FFT plan: log2(F) when F is a power of two, else the padded Bluestein length
    int32_t blue;            // F is not a power of two: chirped weights, two FFTs per shift (bluestein.cuh)
    const float2* d_bhat_perm;   // Bluestein: FFT_M(chirp) in register order (owned by the per-length cache)
    const float2* d_chirp;       // Bluestein: b[k] = exp(+i pi k^2 / F), k < F
    int32_t device;              // the device that owns the arrays of this handle
    xc::GroupPPPlan pp;          // ping-pong layout for contiguous shift ranges (group_pp.cuh), usable or not
};

namespace xc {

constexpr int ELL_UNROLL = 16;

// v[c] = acc for a warp-uniform runtime c: a tree of uniform branches (log2 V of them) instead of V predicated
// selects; the empty asm keeps the leaves from being if-converted back into selects
template <int LO, int HI>
__device__ __forceinline__ void slot_store(float2* v, int c, float2 acc) {
    if constexpr (HI - LO == 1) {
        v[LO] = acc;
        asm volatile("" ::: "memory");
    } else {
        constexpr int MID = (LO + HI) / 2;
        if (c < MID) {
            slot_store<LO, MID>(v, c, acc);
        } else {
            slot_store<MID, HI>(v, c, acc);
        }
    }
}

template <class PL>
struct GroupLoader {
    const float2* rx;
    int64_t rxlen;
    int64_t shift0, step;
    const int64_t* shifts;
    const float2* ell_y;
    const int32_t* ell_off;
    const int32_t* slot_base_g;   // global [V + 1]
    int64_t lo, hi;               // valid shifts satisfy s + lo >= 0 and s + hi <= rxlen
    const int32_t* row_base;      // [rows] or NULL: offsets are row_base[j] + t (clamped into [lo, hi)), no per-thread table
    int32_t rows;                 // contributor rows (each TG weights)
    int32_t y_smem_off;           // byte offset of the shared-memory copy of ell_y behind the static carve-up, -1 = none
    int32_t runs_inside;          // row_base + t never leaves [lo, hi): skip the clamp

    size_t dyn_smem() const { return y_smem_off >= 0 ? sizeof(float2) * (size_t)rows * PL::TG + sizeof(int32_t) * (size_t)rows + 16 : 0; }   // weights, row bases, alignment slack

    // shared-memory table: one int4 {slot, first row, end row, -} per NON-EMPTY register slot, then their count
    __host__ __device__ static constexpr size_t smem_bytes() { return sizeof(int4) * (PL::V + 1); }

    __device__ __forceinline__ void prepare(int tid, int nthreads, unsigned char* smem) const {
        if (tid == 0) {
            int4* tab = reinterpret_cast<int4*>(smem);
            int n = 0;
            for (int c = 0; c < PL::V; ++c)
                if (slot_base_g[c + 1] > slot_base_g[c]) tab[n++] = make_int4(c, slot_base_g[c], slot_base_g[c + 1], 0);
            tab[PL::V] = make_int4(n, 0, 0, 0);
        }
        if (y_smem_off >= 0) {      // the weights never change: keep them next to the FFT buffers
            float2* ys = reinterpret_cast<float2*>(smem + y_smem_off);
            for (int i = tid; i < rows * PL::TG; i += nthreads) ys[i] = ell_y[i];
            if (row_base) {
                int32_t* rb = reinterpret_cast<int32_t*>(ys + (size_t)rows * PL::TG);
                // fast path (runs_inside): unsigned BYTE offsets from sample `lo`, so the address is one 64 + 32 bit add
                for (int i = tid; i < rows; i += nthreads) rb[i] = runs_inside ? (row_base[i] - (int32_t)lo) * 8 : row_base[i];
            }
        }
    }
    __device__ __forceinline__ bool issue(float2* v, int t, int64_t i, const unsigned char* smem) const {
        return load(v, t, i, smem);
    }
    __device__ __forceinline__ void finish(float2*, int, const unsigned char*) const {}
    // The gather is ONE rolled loop over the non-empty register slots (the code of a fully unrolled slot loop
    // does not fit the instruction cache); the finished sum is steered into its register by a uniform switch.
    __device__ __forceinline__ bool load(float2* v, int t, int64_t i, const unsigned char* smem) const {
        const int4* tab = reinterpret_cast<const int4*>(smem);
        const int64_t s = shifts ? shifts[i] : shift0 + i * step;
        if (s + lo < 0 || s + hi > rxlen) return false;
        const float2* x = rx + s;
        // weights: shared-memory copy when it fits (generic pointer either way)
        const float2* const yp = (y_smem_off >= 0) ? reinterpret_cast<const float2*>(smem + y_smem_off) : ell_y;
#pragma unroll
        for (int c = 0; c < PL::V; ++c) v[c] = make_float2(0.f, 0.f);
        const int n_ne = tab[PL::V].x;
        int4 nxt = tab[0];                  // the next slot's entry is fetched one iteration ahead
#pragma unroll 1
        for (int k = 0; k < n_ne; ++k) {
            const int4 ent = nxt;
            nxt = tab[k + 1];
            const int c = ent.x;
            float2 acc = make_float2(0.f, 0.f);
            const int j0 = ent.y, j1 = ent.z;
            // lists are padded to a multiple of ELL_UNROLL (zero weights): all index / weight / sample loads
            // of a batch are issued before the first dependent one is used
            if (row_base) {
                // run-structured lists (every group contributes consecutive samples to consecutive cells): the
                // sample address does not depend on a loaded offset, so weights and samples go out together
                const int32_t* const rb = (y_smem_off >= 0) ? reinterpret_cast<const int32_t*>(yp + (size_t)rows * PL::TG) : row_base;
                if (runs_inside && y_smem_off >= 0) {
                    // fast path: weights and row bases in shared memory (explicit ld.shared), unclamped offsets
                    const char* const xt = reinterpret_cast<const char*>(x + lo + t);
                    const uint32_t ys_s = (uint32_t)__cvta_generic_to_shared(smem + y_smem_off) + (uint32_t)t * 8u;
                    const uint32_t rb_s = (uint32_t)__cvta_generic_to_shared(smem + y_smem_off) + (uint32_t)rows * PL::TG * 8u;
#pragma unroll 1
                    for (int j = j0; j < j1; j += ELL_UNROLL) {
                        float2 r[ELL_UNROLL];
#pragma unroll
                        for (int u = 0; u < ELL_UNROLL; u += 4) {
                            uint32_t o[4];
                            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(o[0]), "=r"(o[1]), "=r"(o[2]), "=r"(o[3])
                                         : "r"(rb_s + 4u * (uint32_t)(j + u)));
#pragma unroll
                            for (int q = 0; q < 4; ++q) r[u + q] = __ldg(reinterpret_cast<const float2*>(xt + o[q]));
                        }
#pragma unroll
                        for (int u = 0; u < ELL_UNROLL; ++u) {
                            float2 y;
                            asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(y.x), "=f"(y.y) : "r"(ys_s + (uint32_t)(j + u) * (PL::TG * 8u)));
                            acc = pfma(make_float2(r[u].y, r[u].x), make_float2(-y.y, y.y), pfma(r[u], make_float2(y.x, y.x), acc));
                        }
                    }
                } else {
                    const int omin = (int)lo, omax = (int)hi - 1;
#pragma unroll 1
                    for (int j = j0; j < j1; j += ELL_UNROLL) {
                        float2 y[ELL_UNROLL], r[ELL_UNROLL];
#pragma unroll
                        for (int u = 0; u < ELL_UNROLL; ++u) {
                            const int o = min(max(rb[j + u] + t, omin), omax);
                            r[u] = __ldg(x + o);
                        }
#pragma unroll
                        for (int u = 0; u < ELL_UNROLL; ++u) y[u] = yp[(size_t)(j + u) * PL::TG + t];
#pragma unroll
                        for (int u = 0; u < ELL_UNROLL; ++u) acc = pfma(make_float2(r[u].y, r[u].x), make_float2(-y[u].y, y[u].y),
                                                                      pfma(r[u], make_float2(y[u].x, y[u].x), acc));
                    }
                }
            } else {
#pragma unroll 1
                for (int j = j0; j < j1; j += ELL_UNROLL) {
                    int off[ELL_UNROLL];
                    float2 y[ELL_UNROLL], r[ELL_UNROLL];
#pragma unroll
                    for (int u = 0; u < ELL_UNROLL; ++u) off[u] = __ldg(ell_off + (size_t)(j + u) * PL::TG + t);
#pragma unroll
                    for (int u = 0; u < ELL_UNROLL; ++u) y[u] = yp[(size_t)(j + u) * PL::TG + t];
#pragma unroll
                    for (int u = 0; u < ELL_UNROLL; ++u) r[u] = __ldg(x + off[u]);
#pragma unroll
                    for (int u = 0; u < ELL_UNROLL; ++u) acc = pfma(make_float2(r[u].y, r[u].x), make_float2(-y[u].y, y[u].y),
                                                                  pfma(r[u], make_float2(y[u].x, y[u].x), acc));
                }
            }
            slot_store<0, PL::V>(v, c, acc);
        }
        return true;
    }
};

// sum of the per-group window energies: E[i] = sum_g EL[shift_i + s_g]
__global__ void group_energy_kernel(const float* __restrict__ el, int64_t el_len, const int64_t* __restrict__ starts,
                                    int G, int64_t shift0, int64_t step, const int64_t* __restrict__ shifts, int64_t K,
                                    float* __restrict__ out) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < K; i += stride) {
        const int64_t s = shifts ? shifts[i] : shift0 + i * step;
        double acc = 0.0;
        for (int g = 0; g < G; ++g) {
            const int64_t p = s + starts[g];
            if (p >= 0 && p < el_len) acc += (double)__ldg(el + p);
        }
        out[i] = (float)acc;
    }
}

struct GroupBuilder {
    const float2* y;         // host, already conjugated if requested
    const int64_t* starts;   // host
    xc_group* h;
    template <int LOG2N>
    int32_t run() {
        using PL = typename PlanFor<LOG2N>::PL;
        // cells of the padded plan (PL::N >= F); the fold is modulo the TRANSFORM length F
        const int F = h->F, TG = PL::TG, V = PL::V, G = h->G, L = h->L;
        struct Contrib { int32_t off; float2 y; };
        std::vector<std::vector<Contrib>> rows(PL::N);
        for (int g = 0; g < G; ++g)
            for (int n = 0; n < L; ++n) {
                const int64_t a = starts[g] + n;
                const int64_t m = ((a % F) + F) % F;
                float2 w = y[(size_t)g * L + n];
                if (h->blue) {      // times conj(b[m]), b[m] = exp(+i pi m^2 / F)
                    const long long r = ((long long)m * m) % (2LL * F);
                    const double ang = -M_PI * (double)r / (double)F, c = cos(ang), sn = sin(ang);
                    w = make_float2((float)(w.x * c - w.y * sn), (float)(w.x * sn + w.y * c));
                }
                rows[(size_t)m].push_back(Contrib{(int32_t)a, w});
            }
        std::vector<int32_t> base(V + 1, 0);
        for (int c = 0; c < V; ++c) {
            const int q = c / PL::R1, n1 = c % PL::R1;
            size_t len = 0;
            for (int t = 0; t < TG; ++t) {
                const int m = n1 * PL::M1 + t + TG * q;
                if (rows[m].size() > len) len = rows[m].size();
            }
            len = (len + ELL_UNROLL - 1) / ELL_UNROLL * ELL_UNROLL;
            base[c + 1] = base[c] + (int32_t)len;
        }
        const int nrows = base[V];
        std::vector<float2> ey((size_t)(nrows > 0 ? nrows : 1) * TG, make_float2(0.f, 0.f));
        // padded entries point at a valid sample (the smallest start) and carry a zero weight
        std::vector<int32_t> eo((size_t)(nrows > 0 ? nrows : 1) * TG, (int32_t)h->min_start);
        for (int c = 0; c < V; ++c) {
            const int q = c / PL::R1, n1 = c % PL::R1;
            for (int t = 0; t < TG; ++t) {
                const int m = n1 * PL::M1 + t + TG * q;
                for (size_t j = 0; j < rows[m].size(); ++j) {
                    ey[((size_t)base[c] + j) * TG + t] = rows[m][j].y;
                    eo[((size_t)base[c] + j) * TG + t] = rows[m][j].off;
                }
            }
        }
        h->rows = nrows;
        // run structure: in every row the real contributors sit at off = base + t
        {
            std::vector<int32_t> rb((size_t)(nrows > 0 ? nrows : 1), (int32_t)h->min_start);
            std::vector<char> has((size_t)(nrows > 0 ? nrows : 1), 0);
            bool runs = true;
            for (int c = 0; c < V && runs; ++c) {
                const int q = c / PL::R1, n1 = c % PL::R1;
                for (int t = 0; t < TG && runs; ++t) {
                    const int m = n1 * PL::M1 + t + TG * q;
                    for (size_t j = 0; j < rows[m].size(); ++j) {
                        const size_t r = (size_t)base[c] + j;
                        const int32_t b0 = rows[m][j].off - t;
                        if (!has[r]) {
                            has[r] = 1;
                            rb[r] = b0;
                        } else if (rb[r] != b0) {
                            runs = false;
                            break;
                        }
                    }
                }
            }
            h->d_row_base = nullptr;
            h->runs_inside = 0;
            if (runs && nrows > 0) {
                bool inside = (h->span - h->min_start) < (int64_t(1) << 27);     // byte offsets fit 32 bits
                for (int r = 0; r < nrows; ++r)
                    if (rb[r] < h->min_start || (int64_t)rb[r] + TG > h->span) inside = false;
                h->runs_inside = inside ? 1 : 0;
                XC_CUDA_TRY(cudaMalloc(&h->d_row_base, sizeof(int32_t) * rb.size()));
                XC_CUDA_TRY(cudaMemcpy(h->d_row_base, rb.data(), sizeof(int32_t) * rb.size(), cudaMemcpyHostToDevice));
            }
        }
        XC_CUDA_TRY(cudaMalloc(&h->d_ell_y, sizeof(float2) * ey.size()));
        XC_CUDA_TRY(cudaMalloc(&h->d_ell_off, sizeof(int32_t) * eo.size()));
        XC_CUDA_TRY(cudaMalloc(&h->d_slot_base, sizeof(int32_t) * (V + 1)));
        XC_CUDA_TRY(cudaMemcpy(h->d_ell_y, ey.data(), sizeof(float2) * ey.size(), cudaMemcpyHostToDevice));
        XC_CUDA_TRY(cudaMemcpy(h->d_ell_off, eo.data(), sizeof(int32_t) * eo.size(), cudaMemcpyHostToDevice));
        XC_CUDA_TRY(cudaMemcpy(h->d_slot_base, base.data(), sizeof(int32_t) * (V + 1), cudaMemcpyHostToDevice));
        return XC_OK;
    }
};

struct GroupRunner {
    xc_group* h;
    const float2* rx;
    int64_t rxlen;
    const ShiftList& sl;
    const FusedOut& out;
    cudaStream_t stream;
    template <int LOG2N>
    int32_t run() {
        using PF = PlanFor<LOG2N>;
        using LD = GroupLoader<typename PF::PL>;
        // the weights move into shared memory when they fit behind the static carve-up
        const size_t stat = (fused_smem_bytes<typename PF::PL, PF::B, LD>() + 15) / 16 * 16;
        const size_t ybytes = sizeof(float2) * (size_t)h->rows * PF::PL::TG;
        const int32_t yoff = (h->rows > 0 && stat + ybytes + sizeof(int32_t) * (size_t)h->rows + 16 <= (size_t)max_smem_optin()) ? (int32_t)stat : -1;
        LD ld{rx, rxlen, sl.shift0, sl.step, sl.shifts, h->d_ell_y, h->d_ell_off, h->d_slot_base,
              h->min_start, h->span, h->d_row_base, h->rows, yoff, h->runs_inside};
        if (h->blue) {
            if constexpr (PF::PL::P >= 2) {
                return launch_bluestein<typename PF::PL, PF::B, LD>(ld, out, sl.K, h->d_bhat_perm, h->d_chirp, h->F, stream);
            } else {
                return fail(XC_ERR_UNSUPPORTED, "no Bluestein plan for 2^%d points", LOG2N);
            }
        }
        return launch_fused<typename PF::PL, PF::B, LD>(ld, out, sl.K, stream);
    }
};

// a build with -DXC_EXPERIMENTS keeps the round-1 kernel selectable (XC_GROUP_LAYOUT=ell) for A/B runs
static bool group_layout_pp() {
#ifdef XC_EXPERIMENTS
    static const bool pp = [] { const char* e = getenv("XC_GROUP_LAYOUT"); return !(e && !strcmp(e, "ell")); }();
    return pp;
#else
    return true;
#endif
}

}  // namespace xc

using namespace xc;

namespace xc {
int32_t collect_peak(const float* d_qf2, const uint32_t* d_fidx, const int32_t* d_aux, int64_t K, int64_t shift0,
                     int64_t step, const int64_t* d_shifts, xc_peak* peak, cudaStream_t st);
}

extern "C" {

int32_t xc_group_create(const void* ygroups, int32_t G, int32_t L, const int64_t* starts, int32_t F, int32_t auto_conj,
                        xc_group** handle) {
    XC_API_GUARD;
    if (!ygroups || !starts || !handle) return fail(XC_ERR_ARG, "ygroups/starts/handle is NULL");
    if (G < 1 || L < 1) return fail(XC_ERR_ARG, "need at least one group of at least one sample");
    if (F < L) return fail(XC_ERR_ARG, "INVALID_FFTLEN: fftlen %d is shorter than the group length %d", F, L);   // GroupXcorrFFT.cpp:11-14
    int log2m = 0;
    const float2* d_bhat_perm = nullptr;
    const float2* d_chirp = nullptr;
    const bool blue = !fft_length_supported(F);
    if (blue) {
        if (F < 3 || F > (1 << (MAX_LOG2N - 1)))
            return fail(XC_ERR_UNSUPPORTED, "GroupXcorrFFT supports fftlen 2..%d (powers of two up to %d), got %d",
                        1 << (MAX_LOG2N - 1), 1 << MAX_LOG2N, F);
    } else {
        log2m = ilog2(F);
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(XC_ERR_CUDA, "no CUDA device available; libxcorr_b200 has no CPU fallback");
    }
    int64_t mn = starts[0], mx = starts[0];
    for (int g = 1; g < G; ++g) {
        if (starts[g] < mn) mn = starts[g];
        if (starts[g] > mx) mx = starts[g];
    }
    if (mx + L >= (1ll << 31) || mn <= -(1ll << 31)) return fail(XC_ERR_UNSUPPORTED, "group starts must fit in 31 bits");
    xc_group* h = new xc_group();
    memset(h, 0, sizeof(*h));
    h->G = G; h->L = L; h->F = F;
    h->device = current_device();
    if (blue) {
        const int32_t bs = bluestein_tables(F, &log2m, &d_bhat_perm, &d_chirp);
        if (bs != XC_OK) {
            delete h;
            return bs;
        }
    }
    h->log2m = log2m;
    h->blue = blue ? 1 : 0;
    h->d_bhat_perm = d_bhat_perm;
    h->d_chirp = d_chirp;
    h->span = mx + L;
    h->min_start = mn;
    const float2* yin = reinterpret_cast<const float2*>(ygroups);
    std::vector<float2> y((size_t)G * L);
    double nsq = 0.0;
    for (size_t i = 0; i < y.size(); ++i) {
        y[i] = auto_conj ? make_float2(yin[i].x, -yin[i].y) : yin[i];
        nsq += (double)yin[i].x * yin[i].x + (double)yin[i].y * yin[i].y;
    }
    int32_t st = dispatch_log2(h->log2m, GroupBuilder{y.data(), starts, h});
    if (st == XC_OK && !blue) st = group_pp_build(y.data(), starts, G, L, F, &h->pp);
    if (st == XC_OK) {
        const float nf = (float)nsq;
        cudaError_t e = cudaMalloc(&h->d_refn, 256);
        if (e == cudaSuccess) e = cudaMemcpy(h->d_refn, &nf, sizeof(float), cudaMemcpyHostToDevice);
        if (e == cudaSuccess) e = cudaMalloc(&h->d_starts, sizeof(int64_t) * G);
        if (e == cudaSuccess) e = cudaMemcpy(h->d_starts, starts, sizeof(int64_t) * G, cudaMemcpyHostToDevice);
        if (e != cudaSuccess) st = fail(XC_ERR_CUDA, "group handle allocation failed: %s", cudaGetErrorString(e));
    }
    if (st != XC_OK) {
        xc_group_destroy(h);
        return st;
    }
    *handle = h;
    return XC_OK;
}

int32_t xc_group_destroy(xc_group* h) {
    XC_API_GUARD;
    if (!h) return XC_OK;
    cudaFree(h->d_ell_y);
    cudaFree(h->d_ell_off);
    cudaFree(h->d_slot_base);
    cudaFree(h->d_row_base);
    cudaFree(h->d_refn);
    cudaFree(h->d_starts);
    group_pp_release(&h->pp);
    delete h;
    return XC_OK;
}

int32_t xc_group_xcorr(xc_group* h, const void* rx, int64_t rxlen, int64_t shift0, int64_t step, int64_t K,
                       const int64_t* shifts, float* qf2, uint32_t* fidx, float* caf, xc_peak* peak, int32_t where,
                       void* stream) {
    XC_API_GUARD;
    if (where != XC_HOST && where != XC_DEVICE) return fail(XC_ERR_ARG, "where must be XC_HOST (0) or XC_DEVICE (1)");
    if (!h || !rx) return fail(XC_ERR_ARG, "handle/rx is NULL");
    XC_TRY(check_handle_device(h->device, "GroupXcorrFFT"));
    if (!qf2 && !caf) return fail(XC_ERR_ARG, "qf2 and caf are both NULL");
    if (K < 0 || rxlen < h->L) return fail(XC_ERR_ARG, "need K >= 0 and rx at least one group long");
    if (K == 0) {
        if (peak) { memset(peak, 0, sizeof(*peak)); peak->index = -1; }
        return XC_OK;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const size_t cafb = sizeof(float) * (size_t)K * (size_t)h->F;
    const float2* drx = reinterpret_cast<const float2*>(rx);
    const int64_t* dsh = shifts;
    float *dq = qf2, *dcaf = caf;
    uint32_t* df = fidx;
    if (where == XC_HOST) {
        float2* b = (float2*)workspace(SLOT_IN1, sizeof(float2) * (size_t)rxlen);
        if (!b) return XC_ERR_NOMEM;
        XC_CUDA_TRY(cudaMemcpyAsync(b, rx, sizeof(float2) * (size_t)rxlen, cudaMemcpyHostToDevice, st));
        drx = b;
        if (shifts) {
            int64_t* s = (int64_t*)workspace(SLOT_IN2, sizeof(int64_t) * (size_t)K);
            if (!s) return XC_ERR_NOMEM;
            XC_CUDA_TRY(cudaMemcpyAsync(s, shifts, sizeof(int64_t) * (size_t)K, cudaMemcpyHostToDevice, st));
            dsh = s;
        }
        if (qf2 || peak) { dq = (float*)workspace(SLOT_OUT0, sizeof(float) * (size_t)K); if (!dq) return XC_ERR_NOMEM; }
        if (fidx || peak) { df = (uint32_t*)workspace(SLOT_OUT1, sizeof(uint32_t) * (size_t)K); if (!df) return XC_ERR_NOMEM; }
        if (caf) { dcaf = (float*)workspace(SLOT_OUT2, cafb); if (!dcaf) return XC_ERR_NOMEM; }
    } else if (peak && (!dq || !df)) {
        if (!dq) dq = (float*)workspace(SLOT_OUT0, sizeof(float) * (size_t)K);
        if (!df) df = (uint32_t*)workspace(SLOT_OUT1, sizeof(uint32_t) * (size_t)K);
        if (!dq || !df) return XC_ERR_NOMEM;
    }
    // energies: length-L sliding energy of the whole record, then the sum over the groups per shift
    const int64_t el_len = rxlen - h->L + 1;
    float* d_el = (float*)workspace(SLOT_TMP2, sizeof(float) * (size_t)el_len);
    float* d_energy = (float*)workspace(SLOT_TMP1, sizeof(float) * (size_t)K);
    if (!d_el || !d_energy) return XC_ERR_NOMEM;
    XC_TRY(sliding_sum_device(drx, true, rxlen, 0, h->L, el_len, d_el, 0, st));
    {
        int64_t blocks = (K + 255) / 256;
        const int64_t cap = (int64_t)sm_count() * 16;
        if (blocks > cap) blocks = cap;
        group_energy_kernel<<<(unsigned)blocks, 256, 0, st>>>(d_el, el_len, h->d_starts, h->G, shift0, step, dsh, K, d_energy);
        XC_LAUNCH_CHECK();
    }
    ShiftList sl{shift0, step, K, dsh};
    // the fused kernel either reduces to (max, arg max) or writes CAF rows; both wanted = two passes
    if (dq || df) {
        FusedOut out{dq, df, nullptr, nullptr, 0, d_energy, h->d_refn};
        if (h->pp.usable && !dsh && step == 1 && K >= 16 * (int64_t)sm_count() && group_layout_pp())
            XC_TRY(group_pp_run(h->pp, h->d_starts, drx, rxlen, shift0, K, h->min_start, h->span, out, st));
        else
            XC_TRY(dispatch_log2(h->log2m, GroupRunner{h, drx, rxlen, sl, out, st}));
    }
    if (dcaf) {
        FusedOut out{nullptr, nullptr, nullptr, dcaf, 0, d_energy, h->d_refn};
        XC_TRY(dispatch_log2(h->log2m, GroupRunner{h, drx, rxlen, sl, out, st}));
    }
    if (where == XC_HOST) {
        if (qf2) XC_CUDA_TRY(cudaMemcpyAsync(qf2, dq, sizeof(float) * (size_t)K, cudaMemcpyDeviceToHost, st));
        if (fidx) XC_CUDA_TRY(cudaMemcpyAsync(fidx, df, sizeof(uint32_t) * (size_t)K, cudaMemcpyDeviceToHost, st));
        if (caf) XC_CUDA_TRY(cudaMemcpyAsync(caf, dcaf, cafb, cudaMemcpyDeviceToHost, st));
    }
    if (peak) XC_TRY(collect_peak(dq, df, nullptr, K, shift0, step, dsh, peak, st));
    if (where == XC_HOST) XC_CUDA_TRY(cudaStreamSynchronize(st));
    return XC_OK;
}

}  // extern "C"
