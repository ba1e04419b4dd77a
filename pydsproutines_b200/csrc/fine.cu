// fine.cu -- the refinement step that follows the detection path: sub-bin frequency and sub-sample time search around
// a found peak.  Replaces xcorrRoutines.fineFreqTimeSearch (xcorrRoutines.py:583-667),
// makeTimeScanSteervec (669-675) and GenXcorr.xcorr (679-719).
//
// The reference works in complex128 throughout and the answers are argmax picks over nearly flat peak tops, so
// this file is FP64 SIMT: small dense phase contractions
//     pp[j]   = sum_n y[n] conj(x[n]) exp(-j 2 pi f_j n / fs)                               (fine frequency scan)
//     X, Y    = DFT_N(x'), DFT_N(y)   with exact integer phase reduction (n m mod N)           (any length N)
//     cost[k] = sum_m X[m] conj(Y[m]) exp(-j 2 pi f_m td_k) / ||X|| / ||Y||                   (time scan)
// The steering matrix (len(td) x N complex128 in the reference) is generated on the fly unless the caller hands one in.
#include "internal.cuh"

namespace xc {

namespace {

constexpr int FT = 256;     // threads per block

__device__ __forceinline__ double2 zmul(double2 a, double2 b) {
    return make_double2(fma(a.x, b.x, -a.y * b.y), fma(a.x, b.y, a.y * b.x));
}
__device__ __forceinline__ double2 zfma(double2 a, double2 b, double2 c) {     // a * b + c
    return make_double2(fma(a.x, b.x, fma(-a.y, b.y, c.x)), fma(a.x, b.y, fma(a.y, b.x, c.y)));
}
// exp(-j 2 pi p), p reduced to [-0.5, 0.5] first
__device__ __forceinline__ double2 unit_phase(double p) {
    p -= rint(p);
    double s, c;
    sincospi(-2.0 * p, &s, &c);
    return make_double2(c, s);
}

__device__ __forceinline__ double2 block_sum(double2 v, double2* sh) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
        v.x += __shfl_xor_sync(0xffffffffu, v.x, o);
        v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
    }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) sh[w] = v;
    __syncthreads();
    double2 t = make_double2(0.0, 0.0);
    if (w == 0) {
        t = (l < FT / 32) ? sh[l] : make_double2(0.0, 0.0);
#pragma unroll
        for (int o = 4; o >= 1; o >>= 1) {
            t.x += __shfl_xor_sync(0xffffffffu, t.x, o);
            t.y += __shfl_xor_sync(0xffffffffu, t.y, o);
        }
    }
    __syncthreads();
    return t;      // valid in thread 0
}

// mode 0: out[n] = y[n] conj(x[n]);  mode 1: out[n] = x[n] exp(-j 2 pi fc n)
__global__ void fine_pointwise_kernel(const double2* __restrict__ x, const double2* __restrict__ y, int64_t N, int mode,
                                      double fc, double2* __restrict__ out) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < N; n += stride) {
        if (mode == 0) {
            const double2 a = y[n], b = x[n];
            out[n] = make_double2(fma(a.x, b.x, a.y * b.y), fma(a.y, b.x, -a.x * b.y));
        } else {
            out[n] = zmul(x[n], unit_phase(fc * (double)n));
        }
    }
}

// out[k] = scale * sum_m a[m] exp(-j 2 pi u[m] v[k]);  u == NULL: u[m] = m;  u_makefreq != 0: u[m] = makeFreq(N, fs)[m];
// scale = 1 / sqrt(norms[0] norms[1]) when norms != NULL.  One block per k.
__global__ void __launch_bounds__(FT) phase_sum_kernel(const double2* __restrict__ a, int64_t N, const double* __restrict__ u,
                                                       int u_makefreq, double fs, const double* __restrict__ v,
                                                       double2* __restrict__ out, const double* __restrict__ norms) {
    __shared__ double2 sh[FT / 32];
    const int64_t k = blockIdx.x;
    const double vk = v[k];
    double2 acc = make_double2(0.0, 0.0);
    for (int64_t m = threadIdx.x; m < N; m += FT) {
        double um;
        if (u) {
            um = u[m];
        } else if (u_makefreq) {                    // signalCreationRoutines.makeFreq (380-386)
            um = (double)m / (double)N * fs;
            if (um >= fs / 2) um -= fs;
        } else {
            um = (double)m;
        }
        acc = zfma(a[m], unit_phase(um * vk), acc);
    }
    const double2 t = block_sum(acc, sh);
    if (threadIdx.x == 0) {
        const double sc = norms ? 1.0 / (sqrt(norms[0]) * sqrt(norms[1])) : 1.0;
        out[k] = make_double2(t.x * sc, t.y * sc);
    }
}

// out[m] = sum_n a[n] exp(-j 2 pi (n m mod N) / N): exact phases, a thread walks n = tid, tid + FT, ... with a rotation
// recurrence that is re-seeded from the exact phase every 32 terms.  One block per output bin.
__global__ void __launch_bounds__(FT) dft_kernel(const double2* __restrict__ a, int64_t N, double2* __restrict__ out) {
    __shared__ double2 sh[FT / 32];
    const unsigned long long m = blockIdx.x, NN = (unsigned long long)N;
    const unsigned long long step = ((unsigned long long)FT % NN) * m % NN;       // phase increment between terms, in 1/N turns
    const double2 rot = unit_phase((double)step / (double)N);
    double2 acc = make_double2(0.0, 0.0);
    unsigned long long r = ((unsigned long long)threadIdx.x % NN) * m % NN;
    double2 w = unit_phase((double)r / (double)N);
    int since = 0;
    for (int64_t n = threadIdx.x; n < N; n += FT) {
        acc = zfma(a[n], w, acc);
        r += step;
        if (r >= NN) r -= NN;
        if (++since == 32) {
            w = unit_phase((double)r / (double)N);
            since = 0;
        } else {
            w = zmul(w, rot);
        }
    }
    const double2 t = block_sum(acc, sh);
    if (threadIdx.x == 0) out[m] = t;
}

// ---------------------------------------------------------------------------------------------
// FP64 transform of any length above FINE_DFT_MAX: radix-2 Stockham passes through global memory for a power of two,
// Bluestein (three such transforms of M >= 2 N - 1 points) otherwise.  The exact O(N^2) DFT above stays for short arrays.
// ---------------------------------------------------------------------------------------------
constexpr int64_t FINE_DFT_MAX = 1 << 14;

// one autosort pass: out[(j / Ns) 2 Ns + k (+ Ns)] = in[j] +- in[j + n / 2] exp(-i pi k / Ns), k = j mod Ns
__global__ void stockham_pass_kernel(const double2* __restrict__ in, double2* __restrict__ out, int64_t n, int64_t Ns) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x, half = n >> 1;
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < half; j += stride) {
        const int64_t k = j & (Ns - 1);
        const double2 w = unit_phase(0.5 * (double)k / (double)Ns);           // exp(-i pi k / Ns)
        const double2 u = in[j], v = zmul(in[j + half], w);
        const int64_t j0 = ((j - k) << 1) + k;
        out[j0] = make_double2(u.x + v.x, u.y + v.y);
        out[j0 + Ns] = make_double2(u.x - v.x, u.y - v.y);
    }
}
// chirp w[n] = exp(-i pi n^2 / N) (n^2 reduced modulo 2 N exactly): A = x w (zero padded), B = conj(w) wrapped around M
__global__ void blue64_setup_kernel(const double2* __restrict__ x, int64_t N, int64_t M, double2* __restrict__ A, double2* __restrict__ Bw) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < M; i += stride) {
        double2 a = make_double2(0.0, 0.0), b = make_double2(0.0, 0.0);
        const int64_t d = (i < N) ? i : ((M - i < N) ? M - i : -1);
        if (d >= 0) {
            const unsigned long long r = ((unsigned long long)d * (unsigned long long)d) % (unsigned long long)(2 * N);
            const double2 w = unit_phase(0.5 * (double)r / (double)N);            // exp(-i pi r / N)
            b = make_double2(w.x, -w.y);
            if (i < N) a = zmul(x[i], w);
        }
        A[i] = a;
        Bw[i] = b;
    }
}
// P = conj(FA FB) (input of the inverse transform as a forward one)
__global__ void blue64_mul_kernel(const double2* __restrict__ FA, const double2* __restrict__ FB, int64_t M, double2* __restrict__ P) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < M; i += stride) {
        const double2 p = zmul(FA[i], FB[i]);
        P[i] = make_double2(p.x, -p.y);
    }
}
// X[k] = w[k] conj(q[k]) / M, k < N
__global__ void blue64_finish_kernel(const double2* __restrict__ q, int64_t N, int64_t M, double2* __restrict__ X) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const double inv = 1.0 / (double)M;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < N; k += stride) {
        const unsigned long long r = ((unsigned long long)k * (unsigned long long)k) % (unsigned long long)(2 * N);
        const double2 w = unit_phase(0.5 * (double)r / (double)N);            // exp(-i pi r / N)
        const double2 v = zmul(w, make_double2(q[k].x, -q[k].y));
        X[k] = make_double2(v.x * inv, v.y * inv);
    }
}

unsigned grid_for(int64_t n);

// in-place-capable power-of-two transform: result ends in `dst` (a and b are two work arrays of n points, src may be a)
static int32_t stockham64(const double2* src, double2* a, double2* b, double2* dst, int64_t n, cudaStream_t st) {
    int passes = 0;
    for (int64_t Ns = 1; Ns < n; Ns <<= 1) ++passes;
    const double2* in = src;
    int p = 0;
    for (int64_t Ns = 1; Ns < n; Ns <<= 1, ++p) {
        double2* out = (p == passes - 1) ? dst : ((in == a) ? b : a);
        if (out == in) out = (in == a) ? b : a;      // never in place
        stockham_pass_kernel<<<grid_for(n / 2), FT, 0, st>>>(in, out, n, Ns);
        XC_LAUNCH_CHECK();
        in = out;
    }
    if (in != dst) XC_CUDA_TRY(cudaMemcpyAsync(dst, in, sizeof(double2) * (size_t)n, cudaMemcpyDeviceToDevice, st));
    return XC_OK;
}

// X = DFT_N(x) in FP64 for any N
static int32_t fine_transform(const double2* x, int64_t N, double2* X, cudaStream_t st) {
    if (N <= FINE_DFT_MAX) {
        dft_kernel<<<(unsigned)N, FT, 0, st>>>(x, N, X);
        XC_LAUNCH_CHECK();
        return XC_OK;
    }
    int64_t M = 1;
    const bool pow2 = (N & (N - 1)) == 0;
    while (M < (pow2 ? N : 2 * N - 1)) M <<= 1;
    double2* w = (double2*)workspace(SLOT_BIG, sizeof(double2) * (size_t)M * 5);
    if (!w) return XC_ERR_NOMEM;
    double2 *A = w, *T1 = w + M, *T2 = w + 2 * M, *FB = w + 3 * M, *FA = w + 4 * M;
    if (pow2) return stockham64(x, T1, T2, X, N, st);
    blue64_setup_kernel<<<grid_for(M), FT, 0, st>>>(x, N, M, A, T1);
    XC_LAUNCH_CHECK();
    XC_TRY(stockham64(T1, T1, T2, FB, M, st));          // FFT of the wrapped chirp
    XC_TRY(stockham64(A, T1, T2, FA, M, st));           // FFT of the chirped input
    blue64_mul_kernel<<<grid_for(M), FT, 0, st>>>(FA, FB, M, A);
    XC_LAUNCH_CHECK();
    XC_TRY(stockham64(A, T1, T2, FA, M, st));
    blue64_finish_kernel<<<grid_for(N), FT, 0, st>>>(FA, N, M, X);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

// r[m] = X[m] conj(Y[m]) (0 outside [bounds[0], bounds[1]) of makeFreq(N, fs));  norms[0] += |X|^2, norms[1] += |Y|^2
__global__ void __launch_bounds__(FT) spectrum_product_kernel(const double2* __restrict__ X, const double2* __restrict__ Y,
                                                              int64_t N, double fs, const double* __restrict__ bounds,
                                                              double2* __restrict__ r, double* __restrict__ norms) {
    __shared__ double2 sh[FT / 32];
    double2 e = make_double2(0.0, 0.0);
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; m < N; m += stride) {
        const double2 a = X[m], b = Y[m];
        e.x += a.x * a.x + a.y * a.y;
        e.y += b.x * b.x + b.y * b.y;
        double2 p = make_double2(fma(a.x, b.x, a.y * b.y), fma(a.y, b.x, -a.x * b.y));
        if (bounds) {
            double f = (double)m / (double)N * fs;
            if (f >= fs / 2) f -= fs;
            if (f < bounds[0] || f >= bounds[1]) p = make_double2(0.0, 0.0);      // xcorrRoutines.py:645-654
        }
        r[m] = p;
    }
    const double2 t = block_sum(e, sh);
    if (threadIdx.x == 0) {
        atomicAdd(norms, t.x);
        atomicAdd(norms + 1, t.y);
    }
}

// out[k] = scale * sum_m a[m] conj(S[k, m])   (np.dot(rx_vec, steeringvec.conj().T), xcorrRoutines.py:656-660)
__global__ void __launch_bounds__(FT) conj_matvec_kernel(const double2* __restrict__ a, int64_t N, const double2* __restrict__ S,
                                                         double2* __restrict__ out, const double* __restrict__ norms) {
    __shared__ double2 sh[FT / 32];
    const int64_t k = blockIdx.x;
    const double2* row = S + (size_t)k * (size_t)N;
    double2 acc = make_double2(0.0, 0.0);
    for (int64_t m = threadIdx.x; m < N; m += FT) {
        const double2 s = row[m];
        acc = zfma(a[m], make_double2(s.x, -s.y), acc);
    }
    const double2 t = block_sum(acc, sh);
    if (threadIdx.x == 0) {
        const double sc = 1.0 / (sqrt(norms[0]) * sqrt(norms[1]));
        out[k] = make_double2(t.x * sc, t.y * sc);
    }
}

// first index of the largest |v| (np.argmax(np.abs(v))); one block
__global__ void __launch_bounds__(FT) argmax_abs_kernel(const double2* __restrict__ v, int64_t n, int64_t* __restrict__ best) {
    __shared__ double sv[FT];
    __shared__ int64_t si[FT];
    double bv = -1.0;
    int64_t bi = 0;
    for (int64_t i = threadIdx.x; i < n; i += FT) {
        const double q = hypot(v[i].x, v[i].y);
        if (q > bv) { bv = q; bi = i; }
    }
    sv[threadIdx.x] = bv;
    si[threadIdx.x] = bi;
    __syncthreads();
    for (int o = FT / 2; o >= 1; o >>= 1) {
        if (threadIdx.x < o) {
            const double q = sv[threadIdx.x + o];
            const int64_t j = si[threadIdx.x + o];
            if (q > sv[threadIdx.x] || (q == sv[threadIdx.x] && j < si[threadIdx.x])) {
                sv[threadIdx.x] = q;
                si[threadIdx.x] = j;
            }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) *best = si[0];
}

int32_t fine_require_device() {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail(XC_ERR_CUDA, "no CUDA device available; libxcorr_b200 has no CPU fallback");
    }
    return XC_OK;
}

unsigned grid_for(int64_t n) {
    int64_t b = (n + FT - 1) / FT;
    const int64_t cap = (int64_t)sm_count() * 8;
    return (unsigned)(b > cap ? cap : (b < 1 ? 1 : b));
}

}  // namespace

}  // namespace xc

using namespace xc;

extern "C" {

int32_t xc_fine_freq_c128(const void* x, const void* y, int64_t N, const double* freq_cycles, int64_t NF, void* pp,
                          int64_t* best, int32_t where, void* stream) {
    XC_API_GUARD;
    if (where != XC_HOST && where != XC_DEVICE) return fail(XC_ERR_ARG, "where must be XC_HOST (0) or XC_DEVICE (1)");
    if (!x || !y || !freq_cycles || !pp) return fail(XC_ERR_ARG, "NULL argument");
    if (N < 1 || NF < 1 || NF > (int64_t(1) << 30)) return fail(XC_ERR_ARG, "need at least one sample and one frequency");
    XC_TRY(fine_require_device());
    cudaStream_t st = (cudaStream_t)stream;
    const size_t nb = sizeof(double2) * (size_t)N;
    const double2 *dx = (const double2*)x, *dy = (const double2*)y;
    double2* dpp = (double2*)pp;
    int64_t* dbest = best;
    double* df = (double*)workspace(SLOT_IN2, sizeof(double) * (size_t)NF);        // the frequency list is a host array
    if (!df) return XC_ERR_NOMEM;
    XC_CUDA_TRY(cudaMemcpyAsync(df, freq_cycles, sizeof(double) * (size_t)NF, cudaMemcpyHostToDevice, st));
    if (where == XC_HOST) {
        double2* bx = (double2*)workspace(SLOT_IN0, nb);
        double2* by = (double2*)workspace(SLOT_IN1, nb);
        dpp = (double2*)workspace(SLOT_OUT0, sizeof(double2) * (size_t)NF);
        dbest = (int64_t*)workspace(SLOT_OUT1, 256);
        if (!bx || !by || !dpp || !dbest) return XC_ERR_NOMEM;
        XC_CUDA_TRY(cudaMemcpyAsync(bx, x, nb, cudaMemcpyHostToDevice, st));
        XC_CUDA_TRY(cudaMemcpyAsync(by, y, nb, cudaMemcpyHostToDevice, st));
        dx = bx; dy = by;
    } else if (!dbest) {
        dbest = (int64_t*)workspace(SLOT_OUT1, 256);
        if (!dbest) return XC_ERR_NOMEM;
    }
    double2* d_p = (double2*)workspace(SLOT_TMP0, nb);
    if (!d_p) return XC_ERR_NOMEM;
    fine_pointwise_kernel<<<grid_for(N), FT, 0, st>>>(dx, dy, N, 0, 0.0, d_p);
    XC_LAUNCH_CHECK();
    phase_sum_kernel<<<(unsigned)NF, FT, 0, st>>>(d_p, N, nullptr, 0, 0.0, df, dpp, nullptr);
    XC_LAUNCH_CHECK();
    argmax_abs_kernel<<<1, FT, 0, st>>>(dpp, NF, dbest);
    XC_LAUNCH_CHECK();
    if (where == XC_HOST) {
        XC_CUDA_TRY(cudaMemcpyAsync(pp, dpp, sizeof(double2) * (size_t)NF, cudaMemcpyDeviceToHost, st));
        if (best) XC_CUDA_TRY(cudaMemcpyAsync(best, dbest, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
        XC_CUDA_TRY(cudaStreamSynchronize(st));
    }
    return XC_OK;
}

int32_t xc_time_scan_c128(const void* x, const void* y, int64_t N, double shift_cycles, const double* td, int64_t NT,
                          double fs, const double* freq_bounds, const void* steeringvec, void* cost, int64_t* best,
                          int32_t where, void* stream) {
    XC_API_GUARD;
    if (where != XC_HOST && where != XC_DEVICE) return fail(XC_ERR_ARG, "where must be XC_HOST (0) or XC_DEVICE (1)");
    if (!x || !y || !cost || (!td && !steeringvec)) return fail(XC_ERR_ARG, "NULL argument");
    if (N < 1 || NT < 1 || NT > (int64_t(1) << 30)) return fail(XC_ERR_ARG, "need at least one sample and one time offset");
    if (N > (int64_t(1) << 22)) return fail(XC_ERR_UNSUPPORTED, "time scan of more than 2^22 samples");
    XC_TRY(fine_require_device());
    cudaStream_t st = (cudaStream_t)stream;
    const size_t nb = sizeof(double2) * (size_t)N;
    const double2 *dx = (const double2*)x, *dy = (const double2*)y, *dS = (const double2*)steeringvec;
    const double *dtd = nullptr, *dbounds = nullptr;
    double2* dcost = (double2*)cost;
    int64_t* dbest = best;
    // temporaries: [X | Y | r | norms(2) bounds(2)] in one slot, the fine-shifted x in another
    double2* tmp = (double2*)workspace(SLOT_TMP0, 3 * nb + 64);
    if (!tmp) return XC_ERR_NOMEM;
    double2 *dX = tmp, *dY = tmp + N, *dr = tmp + 2 * (size_t)N;
    double* dnorms = reinterpret_cast<double*>(tmp + 3 * (size_t)N);
    if (where == XC_HOST) {
        double2* bx = (double2*)workspace(SLOT_IN0, nb);
        double2* by = (double2*)workspace(SLOT_IN1, nb);
        dcost = (double2*)workspace(SLOT_OUT0, sizeof(double2) * (size_t)NT);
        dbest = (int64_t*)workspace(SLOT_OUT1, 256);
        if (!bx || !by || !dcost || !dbest) return XC_ERR_NOMEM;
        XC_CUDA_TRY(cudaMemcpyAsync(bx, x, nb, cudaMemcpyHostToDevice, st));
        XC_CUDA_TRY(cudaMemcpyAsync(by, y, nb, cudaMemcpyHostToDevice, st));
        dx = bx; dy = by;
        if (steeringvec) {
            double2* b = (double2*)workspace(SLOT_OUT2, sizeof(double2) * (size_t)NT * (size_t)N);
            if (!b) return XC_ERR_NOMEM;
            XC_CUDA_TRY(cudaMemcpyAsync(b, steeringvec, sizeof(double2) * (size_t)NT * (size_t)N, cudaMemcpyHostToDevice, st));
            dS = b;
        }
    } else if (!dbest) {
        dbest = (int64_t*)workspace(SLOT_OUT1, 256);
        if (!dbest) return XC_ERR_NOMEM;
    }
    // td and freq_bounds are small HOST arrays in both modes
    if (td && !steeringvec) {
        double* b = (double*)workspace(SLOT_IN2, sizeof(double) * (size_t)NT);
        if (!b) return XC_ERR_NOMEM;
        XC_CUDA_TRY(cudaMemcpyAsync(b, td, sizeof(double) * (size_t)NT, cudaMemcpyHostToDevice, st));
        dtd = b;
    }
    if (freq_bounds) {
        XC_CUDA_TRY(cudaMemcpyAsync(dnorms + 2, freq_bounds, 2 * sizeof(double), cudaMemcpyHostToDevice, st));
        dbounds = dnorms + 2;
    }
    XC_CUDA_TRY(cudaMemsetAsync(dnorms, 0, 2 * sizeof(double), st));
    if (shift_cycles != 0.0) {          // x_aligned * fineshifts[fineFreq_ind] (xcorrRoutines.py:632)
        double2* dxs = (double2*)workspace(SLOT_TMP1, nb);
        if (!dxs) return XC_ERR_NOMEM;
        fine_pointwise_kernel<<<grid_for(N), FT, 0, st>>>(dx, nullptr, N, 1, shift_cycles, dxs);
        XC_LAUNCH_CHECK();
        dx = dxs;
    }
    XC_TRY(fine_transform(dx, N, dX, st));          // exact DFT up to 2^14 samples, Stockham / Bluestein in FP64 above
    XC_TRY(fine_transform(dy, N, dY, st));
    spectrum_product_kernel<<<grid_for(N), FT, 0, st>>>(dX, dY, N, fs, dbounds, dr, dnorms);
    XC_LAUNCH_CHECK();
    if (dS) {
        conj_matvec_kernel<<<(unsigned)NT, FT, 0, st>>>(dr, N, dS, dcost, dnorms);
    } else {
        phase_sum_kernel<<<(unsigned)NT, FT, 0, st>>>(dr, N, nullptr, 1, fs, dtd, dcost, dnorms);
    }
    XC_LAUNCH_CHECK();
    argmax_abs_kernel<<<1, FT, 0, st>>>(dcost, NT, dbest);
    XC_LAUNCH_CHECK();
    if (where == XC_HOST) {
        XC_CUDA_TRY(cudaMemcpyAsync(cost, dcost, sizeof(double2) * (size_t)NT, cudaMemcpyDeviceToHost, st));
        if (best) XC_CUDA_TRY(cudaMemcpyAsync(best, dbest, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
        XC_CUDA_TRY(cudaStreamSynchronize(st));
    }
    return XC_OK;
}

}  // extern "C"
