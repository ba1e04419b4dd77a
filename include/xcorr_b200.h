/* xcorr_b200.h -- C-ABI of libxcorr_b200.so: B200 (sm_100a) kernels for the cross-correlation
 * detection hot path of icyveins7/pydsproutines.
 *
 * Conventions (they follow the reference's own ctypes-DLL convention, cpuWola.py:38-66 and
 * cython_ext/compareIntPreambles/compareIntPreambles.py:7-30):
 *   - every buffer is caller-owned; outputs are caller-allocated;
 *   - complex64 data is passed as `const void*` to interleaved (re, im) float32 pairs;
 *   - `where` tells whether the data pointers are host (XC_HOST) or device (XC_DEVICE) memory;
 *     with XC_HOST the library stages through its own cached device workspace (H2D, kernels, D2H)
 *     and returns when the outputs are host-visible; with XC_DEVICE it only enqueues on `stream`
 *     (a cudaStream_t, NULL = default stream) and does NOT synchronise;
 *   - every function returns an int32 status (XC_OK == 0, negative on error) and leaves a
 *     thread-local message readable through xc_last_error(); no C++ exception crosses the ABI;
 *   - there is no CPU fallback: without a CUDA device every compute call returns XC_ERR_CUDA;
 *   - threading: every entry point takes one process-wide lock (calls from several host threads are serialised); the
 *     device workspace, the plan caches and the kernel attributes are kept PER DEVICE (xc_set_device() then calls on
 *     another device of the same process work); the workspace slots of a device are shared by all its streams, so
 *     keep ONE stream per device in flight (XC_DEVICE calls only enqueue; a second stream would reuse the
 *     normalisation scratch while the first still reads it).
 *
 * Each entry point cites the reference interface it replaces (file:line in pydsproutines @ e005737).
 */
#ifndef XCORR_B200_H
#define XCORR_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define XC_OK 0
#define XC_ERR_ARG (-1)         /* bad argument value (maps to ValueError)                       */
#define XC_ERR_RANGE (-2)       /* shift / index outside the input (maps to ValueError)          */
#define XC_ERR_UNSUPPORTED (-3) /* size not supported by the kernels (maps to NotImplementedError)*/
#define XC_ERR_NOMEM (-4)       /* device allocation failed (maps to MemoryError)                */
#define XC_ERR_CUDA (-5)        /* CUDA runtime / launch failure (maps to RuntimeError)          */

#define XC_HOST 0
#define XC_DEVICE 1

/* flags of xc_fastxcorr_c64 */
#define XC_FLAG_ZERO_OOB 1 /* shifts with s<0 or s+N>rxlen give 0.0 / 0 (IppXcorrFFT.cpp:125-130) */

/* The peak tuple (QF^2, frequency index, peak position) of one call; what ranks exchange in the
 * shift-sharded multi-GPU run (one allgather of world_size of these). 32 bytes. */
typedef struct xc_peak {
    float qf2;     /* largest QF^2 of the call                                     */
    uint32_t fidx; /* its frequency bin (0 when no frequency search)               */
    int64_t index; /* position in the call's shift list (lowest on ties)           */
    int64_t shift; /* the shift value itself (shift0 + index*step, or shifts[index])*/
    int32_t aux;   /* template index for the multi-template correlator, else 0     */
    int32_t rank;  /* filled by the host side when peaks of several ranks are merged*/
} xc_peak;

/* ---- library / device ------------------------------------------------------------------ */
int32_t xc_version(void);
const char* xc_last_error(void);
int32_t xc_device_count(void);
int32_t xc_set_device(int32_t device);
/* ordinal of the device that owns `p` (device or managed memory), -1 for host pointers: lets the Python side adopt any
 * object exposing __cuda_array_interface__ (the reference's gate is cupyHelpers.py:80-83 requireCupyArray) */
int32_t xc_pointer_device(const void* p);
/* frees the cached device workspace of the current device */
int32_t xc_release_workspace(void);
/* device properties the bench reports: out[0]=SM count, out[1]=max smem/block (opt-in), out[2]=clock kHz,
 * out[3]=compute capability major*10+minor */
int32_t xc_device_info(int32_t* out4);
/* number of kernels launched by this library in this process so far (bench.py's gpu_launches) */
int64_t xc_launch_count(void);
/* CUDA-event timing of the dominant (fused FFT) kernel: when enabled, every launch of it is bracketed
 * by two events on its own stream; xc_profile_last_ms() waits for the latest one and returns its
 * duration in milliseconds (-1 when nothing was recorded).  Used by bench.py for the roofline. */
int32_t xc_profile_enable(int32_t on);
double xc_profile_last_ms(void);
/* sustained FP32 FMA rate of the device in TFLOP/s measured with a register-resident FFMA loop
 * for `ms` milliseconds (roofline denominator of the FFT kernels; not a reference function) */
int32_t xc_measure_fp32_tflops(double ms, double* tflops);

/* ---- a11: sliding energy primitives ------------------------------------------------------ */
/* valid-mode windowed energy out[s] = sum_{n<L} |x[s+n]|^2, s in [0, n-L+1)
 * (what xcorrRoutines.py:342-355 slices out of cupyComplexMagnSq + cupyMovingAverage(sumInstead)) */
int32_t xc_sliding_energy_c64(const void* x, int64_t n, int32_t L, float* out, int32_t where,
                              void* stream);
/* cupyMovingAverage (filterRoutines.py:1167-1203 -> filter.cu:291-347): causal length-L window,
 * zero padded front, out has n elements; sum_instead != 0 gives the moving sum */
int32_t xc_moving_average_f32(const float* x, int64_t n, int32_t L, int32_t sum_instead, float* out,
                              int32_t where, void* stream);
/* cupyComplexMagnSq (cupyExtensions.py:337-387 -> complex_magn.cu:8-20): out[i] = re^2 + im^2
 * evaluated in the output type; out_f64 != 0 writes float64 */
int32_t xc_complex_magnsq_c64(const void* x, int64_t n, void* out, int32_t out_f64, int32_t where,
                              void* stream);

/* ---- a1, a2, a4, a5, a6: fastXcorr --------------------------------------------------------
 * xcorrRoutines.fastXcorr (xcorrRoutines.py:460-531), cp_fastXcorr (29-167), cp_fastXcorr_v2
 * (169-274), IppXcorrFFT_32fc::xcorr_array (IppXcorrFFT.cpp:94-178).
 * Shift list: shift0 + i*step, i in [0, K), or the explicit list `shifts` (int64[K], same memory
 * space as `rx`) when non-NULL.  The cutout is conjugated inside (pass it as the reference does).
 *   freqsearch == 0: qf2[i] = |sum_n conj(rx[s+n]) cutout[n]|^2 / ||cutout||^2 / ||rx[s:s+N]||^2
 *   freqsearch != 0: X = FFT_N(rx[s:s+N] * conj(cutout)); fidx[i] = argmax |X|^2 (lowest index on
 *                    ties); qf2[i] = |X[fidx]|^2 / ||cutout||^2 / ||rx[s:s+N]||^2
 * qf2: float32[K]; fidx: uint32[K] (may be NULL when freqsearch == 0); cplx: optional
 * complex64[K] receiving the un-squared complex QF of the absResult=False variants
 * (xcorrRoutines.py:497-509, 533-548); peak: optional, HOST pointer, filled with the call's
 * largest QF^2 (forces a stream synchronisation). */
int32_t xc_fastxcorr_c64(const void* cutout, int32_t N, const void* rx, int64_t rxlen, int64_t shift0,
                         int64_t step, int64_t K, const int64_t* shifts, int32_t freqsearch,
                         int32_t flags, float* qf2, uint32_t* fidx, void* cplx, xc_peak* peak,
                         int32_t where, void* stream);
/* a3: full CAF rows (xcorrRoutines.py:550-580): caf is float32 (K, N) QF^2 when complex_out == 0,
 * complex64 (K, N) un-squared QF when complex_out != 0. Bins are in natural FFT order. */
int32_t xc_fastxcorr_caf_c64(const void* cutout, int32_t N, const void* rx, int64_t rxlen,
                             int64_t shift0, int64_t step, int64_t K, const int64_t* shifts,
                             void* caf, int32_t complex_out, int32_t where, void* stream);
/* ---- 8f-1: chirp-z frequency zoom --------------------------------------------------------------
 * xcorrRoutines.cztXcorr (xcorrRoutines.py:413-457) with spectralRoutines.CZTCached (spectralRoutines.py:239-309):
 * per shift the product rx[s:s+N] * conj(cutout) is evaluated on the frequency grid
 *     f_k = (f0_cycles + k * df_cycles) cycles per sample,  k < nbins      (f / fs in the reference's units)
 * by the same fused kernel (Bluestein with a general chirp) while N + nbins - 1 <= 16384, by the four-step path up to
 * 2^28 points.  qf2[i] = max_k |X[k]|^2 /
 * ||cutout||^2 / ||rx slice||^2, fidx[i] = arg max, cplx[i] = X[arg max] / ||cutout|| / ||rx slice||,
 * caf (optional float32 (K, nbins)) = all bins.  Any of the outputs may be NULL. */
int32_t xc_cztxcorr_c64(const void* cutout, int32_t N, const void* rx, int64_t rxlen, int64_t shift0,
                        int64_t step, int64_t K, const int64_t* shifts, double f0_cycles, double df_cycles,
                        int32_t nbins, float* qf2, uint32_t* fidx, void* cplx, float* caf, int32_t where,
                        void* stream);
/* ---- allocation-free variant of the no-frequency-search correlator ------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arithmetic and the
 * same kernels as xc_fastxcorr_c64(freqsearch = 0) on the contiguous shifts shift0 .. shift0 + K - 1 (all inside rx),
 * DEVICE pointers only, enqueued on `stream` without synchronising:
 *     qf2[i] = |sum_n conj(rx[shift0 + i + n]) cutout[n]|^2 / ||cutout||^2 / ||rx[shift0 + i : + N]||^2,  float32[K].
 * The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine): ||cutout||^2 and the K window
 * energies live in a static device array of XC_NOALLOC_MAX_SHIFTS + 64 floats (4 MiB + 256 bytes, held by every GPU
 * that loads the library until the process exits).  K > XC_NOALLOC_MAX_SHIFTS returns XC_ERR_UNSUPPORTED (split the
 * range or use xc_fastxcorr_c64).  Every call on a device shares that array and fully rewrites the part it uses
 * (nothing has to start at zero); two calls that overlap on the GPU (different streams of one device) race on it --
 * keep them in stream order.  (Cutouts of 65536 samples and more: xc_fastxcorr_c64 sums ||cutout||^2 and the first window
 * of a tile through a small scratch allocation; this entry keeps the serial sums, the results agree to rounding.) */
#define XC_NOALLOC_MAX_SHIFTS 1048576
int32_t xc_fastxcorr_nofreq_noalloc_c64(const void* cutout, int32_t N, const void* rx, int64_t rxlen, int64_t shift0,
                                        int64_t K, float* qf2, void* stream);
/* 1 when the frequency search takes this cutout length: powers of two 2..16384 (one block FFT per shift on one SM), every
 * other length 3..8192 (Bluestein: two block FFTs per shift), and beyond that the four-step path: powers of two up to 2^28
 * points, any other length while 2 N - 1 <= 2^28 */
int32_t xc_fastxcorr_supported(int32_t N);

/* ---- a7, a8: GroupXcorrFFT ----------------------------------------------------------------
 * xcorrRoutines.GroupXcorrFFT (xcorrRoutines.py:1047-1262), GroupXcorrFFT.cpp:3-199.
 * ygroups: complex64 (G, L) HOST memory, conjugated inside when auto_conj != 0; starts: int64[G]
 * used as given (the Python side applies autoZeroStarts); F = fftlen >= L. */
typedef struct xc_group xc_group;
int32_t xc_group_create(const void* ygroups, int32_t G, int32_t L, const int64_t* starts, int32_t F,
                        int32_t auto_conj, xc_group** handle);
/* per shift s: Z[k] = sum_g exp(-j2pi k starts[g]/F) FFT_F(ygroups[g] * rx[s+starts[g] : +L])[k];
 * caf[k] = |Z[k]|^2 / sum_g ||rx_g||^2 / ||ygroups||^2.  qf2/fidx (float32[K], uint32[K]) receive
 * the max / argmax over k when non-NULL; caf (float32 (K, F)) the full rows when non-NULL. */
int32_t xc_group_xcorr(xc_group* handle, const void* rx, int64_t rxlen, int64_t shift0, int64_t step,
                       int64_t K, const int64_t* shifts, float* qf2, uint32_t* fidx, float* caf,
                       xc_peak* peak, int32_t where, void* stream);
int32_t xc_group_destroy(xc_group* handle);

/* ---- 8(f) rank 1 / 4: grouped correlator over an arbitrary frequency list ------------------
 * xcorrRoutines.GroupXcorr (xcorrRoutines.py:852-954: freqMat @ (rxconcat * yconcat), max over the list),
 * xcorrRoutines.GroupXcorrCZT (xcorrRoutines.py:957-1039: CAF over the grid f1 : binWidth : f2) and the
 * experimental GroupXcorrGPU RawKernels (xcorrRoutines.py:1693-2051).
 *   d[s, k]    = sum_g sum_t y_g[t] rx[s + starts[g] + t] exp(-j 2 pi freq_cycles[k] (starts[g] + t))
 *   QF2[s, k]  = |d[s, k]|^2 / sum_g ||rx[s + starts[g] : + lengths[g]]||^2 / ||y||^2
 * ycat = the group samples back to back (sum of lengths; conjugated on upload when conj != 0), freq_cycles =
 * f / fs.  start_freq_cycles (NULL = freq_cycles) is the frequency used for the starts[g] part of the phase:
 * GroupXcorrCZT rotates the groups with the nominal grid f1 + k binWidth while its CZT bins sit at
 * f1 + k (f2 - f1 + binWidth) / nbins (spectralRoutines.py:239-262).  Shifts are the contiguous range shift0 .. shift0 + K - 1.  qf2/fidx = max / arg max over the list per
 * shift, caf = the full [K, NF] float surface, cplx = the raw complex d[s, k] (device buffers only); any may be NULL.  Runs on the tensor cores (tcgen05);
 * XC_ERR_UNSUPPORTED on anything else. */
typedef struct xc_groupfreq xc_groupfreq;
int32_t xc_groupfreq_create(const void* ycat, const int64_t* starts, const int32_t* lengths, int32_t G,
                            const double* freq_cycles, const double* start_freq_cycles, int32_t NF, int32_t conj,
                            xc_groupfreq** handle);
int32_t xc_groupfreq_xcorr(xc_groupfreq* handle, const void* rx, int64_t rxlen, int64_t shift0, int64_t K,
                           float* qf2, uint32_t* fidx, float* caf, void* cplx, xc_peak* peak, int32_t where,
                           void* stream);
int32_t xc_groupfreq_destroy(xc_groupfreq* handle);
/* GroupXcorrCZT_Permutations (xcorrRoutines.py:1264-1690): every template of every group is correlated once
 * (xc_groupfreq_xcorr with cplx = the raw complex surface d[s, k], complex64 [K, NF], device only), then any
 * permutation is  caf[s, b] = |sum_g surfaces[template_idx[g], s, b]|^2 / sum_g rxnormsq[group_idx[g], s] / ynormsq
 * (getCAF_GPU, :1452-1483).  surfaces [NTemplates, K, NB] complex64, rxnormsq [NGroups, K] float32 and caf [K, NB]
 * float32 are DEVICE buffers; template_idx / group_idx are host arrays of G entries. */
int32_t xc_groupperm_combine(const void* surfaces, int64_t K, int32_t NB, const int32_t* template_idx,
                             const int32_t* group_idx, int32_t G, const float* rxnormsq, double ynormsq, float* caf,
                             void* stream);

/* stand-alone chirp-z transform, spectralRoutines.CZTCached(GPU).run / runMany (spectralRoutines.py:263-306, 347-391):
 * out[r, k] = sum_n x[r, n] exp(-j 2 pi (f0_cycles + k df_cycles) n), x = [rows, m] complex64, out = [rows, nbins]
 * complex64.  m + nbins - 1 <= 16384 (the transform runs inside the fused correlation kernel). */
int32_t xc_czt_c64(const void* x, int32_t m, int64_t rows, double f0_cycles, double df_cycles, int32_t nbins, void* out,
                   int32_t where, void* stream);

/* ---- 8(f) rank 3: capture ingestion ---------------------------------------------------------
 * usrpRoutines.simpleBinRead (usrpRoutines.py:51-66: np.fromfile(int16).astype(float32).view(complex64)) and the
 * copy-then-convert-on-device finding of benchmarks/benchmark_cupyCopyAndConvert.py:15-26.  iq = nsamp interleaved
 * (I, Q) int16 pairs, out = nsamp complex64; each side host or device (where_in / where_out).  With a host input
 * the raw int16 samples cross PCIe (half the bytes of complex64) and are widened on the device. */
int32_t xc_convert_i16_c64(const int16_t* iq, int64_t nsamp, void* out, int32_t where_in, int32_t where_out, void* stream);

/* ---- 8(f) rank 2: refinement around a found peak -------------------------------------------
 * xcorrRoutines.fineFreqTimeSearch (xcorrRoutines.py:583-667), makeTimeScanSteervec (669-675) and
 * GenXcorr.xcorr (679-719).  complex128 (interleaved double) like the reference; x / y / steeringvec / pp / cost /
 * best follow `where`, the short lists freq_cycles / td / freq_bounds are always HOST arrays.
 *   xc_fine_freq_c128: pp[j] = sum_n y[n] conj(x[n]) exp(-j 2 pi freq_cycles[j] n)  (= np.vdot(y.conj() * x, shift_j),
 *                      :620-629), *best = np.argmax(np.abs(pp)).
 *   xc_time_scan_c128: x' = x exp(-j 2 pi shift_cycles n) (:632), X = fft(x'), Y = fft(y) for ANY length N,
 *                      r = X conj(Y), zeroed where makeFreq(N, fs) is outside [freq_bounds[0], freq_bounds[1]) (:645-654),
 *                      cost[k] = sum_m r[m] conj(S[k, m]) / ||X|| / ||Y|| (:656-660) with S = steeringvec [NT, N] or, when
 *                      steeringvec is NULL, exp(+j 2 pi makeFreq(N, fs)[m] td[k]) generated on the fly;
 *                      *best = np.argmax(np.abs(cost)). */
int32_t xc_fine_freq_c128(const void* x, const void* y, int64_t N, const double* freq_cycles, int64_t NF, void* pp,
                          int64_t* best, int32_t where, void* stream);
int32_t xc_time_scan_c128(const void* x, const void* y, int64_t N, double shift_cycles, const double* td, int64_t NT,
                          double fs, const double* freq_bounds, const void* steeringvec, void* cost, int64_t* best,
                          int32_t where, void* stream);

/* ---- a9, a10: (multi-)template correlator -------------------------------------------------
 * xcorrRoutines.TemplateCrossCorrelator (xcorrRoutines.py:277-371) and
 * cupyExtensions.multiTemplateSlidingDotProduct (cupyExtensions.py:563-642 -> multiplySlices.cu:251-399).
 * templates: complex64 (T, L) HOST memory; conj is a flag word: XC_MT_CONJ conjugates them inside
 * (TemplateCrossCorrelator semantics; without it they are used as given, multiTemplateSlidingDotProduct semantics),
 * XC_MT_FP32 keeps the contraction on the FP32 SIMT kernel instead of the tensor cores (BF16 hi/lo split, FP32
 * accumulate; both agree to 1e-5 -- the flag exists for callers who want plain FP32 products and for A/B runs).
 * energies: optional float32[T] template energies (NULL = computed inside).
 * (best, template index) requests for templates of 512 ... 4096 samples over at least four blocks of 8193 - L shifts run as a fast
 * correlation (overlap-save through 8192-point FFT blocks, one transform of the x block for all templates, FP32); XC_MT_DIRECT
 * keeps the contraction.  Both agree with the reference to 1e-5. */
#define XC_MT_CONJ 1
#define XC_MT_FP32 2
#define XC_MT_DIRECT 4   /* keep the dense contraction (tensor cores) also where the fast-correlation path would run */
typedef struct xc_mt xc_mt;
int32_t xc_mt_create(const void* templates, int32_t T, int32_t L, int32_t conj, const float* energies,
                     xc_mt** handle);
/* for s = start + i, i in [0, K): d_t = sum_n tmpl_t[n] x[s+n];
 *   best[i] = max_t |d_t|^2 / E_t / E_x[s]  (sqrt of it when sqrt_out != 0), tidx[i] = argmax_t
 *   (lowest t on ties; index 0 when nothing is > 0, multiplySlices.cu:313-316,382-386);
 *   full (optional complex64 (T, K)) = d_t / sqrt(E_t) / sqrt(E_x[s])  (xcorrRoutines.py:352-357) */
int32_t xc_mt_correlate(xc_mt* handle, const void* x, int64_t xlen, int64_t start, int64_t K,
                        float* best, int32_t* tidx, void* full, int32_t sqrt_out, xc_peak* peak,
                        int32_t where, void* stream);
int32_t xc_mt_destroy(xc_mt* handle);
/* which kernel xc_mt_correlate runs for this handle when `full` is NULL: 1 = tensor cores (tcgen05: x split
 * into BF16 hi + lo, templates too unless they are exact in BF16, FP32 accumulation), 0 = FP32 SIMT (the
 * tile does not fit in shared memory, or more than 2048 templates).  No reference counterpart; the result
 * tuples are the same either way. */
int32_t xc_mt_uses_tensor_cores(const xc_mt* handle);

/* ---- a14: gather / copy helpers of the reference's GPU path -----------------------------------
 * Not on the hot path any more (the fused kernels never materialise windows or products in HBM); kept as thin,
 * bandwidth-bound compatibility functions.
 * cupyCopyGroups32fc (cupyExtensions.py:41-83): y[y_starts[g] + i] = x[x_starts[g] + i], i < lengths[g];
 * no bounds or overlap checks, as in the reference. xlen / ylen are only used to stage HOST buffers. */
int32_t xc_copy_groups_c64(const void* x, int64_t xlen, void* y, int64_t ylen, const int32_t* x_starts,
                           const int32_t* y_starts, const int32_t* lengths, int32_t G, int32_t where,
                           void* stream);
/* cupyCopyIncrementalEqualSlicesToMatrix_32fc (cupyExtensions.py:161-215 -> copying.cu:83-137):
 * out[r, c] = x[start + r * increment + c], complex64 (num_rows, row_length); 0 where the index leaves x */
int32_t xc_copy_slices_to_matrix_c64(const void* x, int64_t xlen, int64_t start, int64_t increment,
                                     int64_t row_length, int64_t num_rows, void* out, int32_t where,
                                     void* stream);
/* multiplySlidesNormalised (cupyExtensions.py:491-560 -> multiplySlices.cu:112-216):
 * out[r, t] = x[t] * y[start + r + t] / (float)(sqrt(sum_t |y[start + r + t]|^2) * coefficient), complex64
 * (idxlen, xlen); x is used as given (not conjugated); coefficient <= 0 means the norm of x (the default). */
int32_t xc_multiply_slides_normalised_c64(const void* x, int32_t xlen, const void* y, int64_t ylen,
                                          int64_t start, int64_t idxlen, double coefficient, void* out,
                                          int32_t where, void* stream);

/* ---- a12: row argmax (cupyExtensions.py:268-319 -> argmax.cu:93-152) ------------------------ */
/* per row of the complex64 (rows, cols) matrix: index of the largest abs(x) (use_normsq == 0) or
 * abs(x)^2 (use_normsq != 0), lowest index on ties; val (optional) receives that value */
int32_t xc_row_argmax_c64(const void* x, int64_t rows, int64_t cols, int32_t use_normsq, uint32_t* idx,
                          float* val, int32_t where, void* stream);

/* ---- a13: local maxima (cupyExtensions.py:651-686 -> peakfinding.cu:14-58) ------------------ */
/* indices i with x[i] > min_height, x[i] > x[i-1], x[i] > x[i+1] (outside = 0); at most `cap` are
 * stored (count still reports how many exist); sorted != 0 returns them in ascending order */
int32_t xc_find_local_maxima_f32(const float* x, int64_t n, float min_height, int32_t* idx, int32_t cap,
                                 int32_t* count, int32_t sorted, int32_t where, void* stream);

/* ---- 8e: the one exchange step of the shift-sharded multi-GPU search ---------------------------------------
 * (the reference has no multi-GPU code; north_star: work is partitioned by shift range, only the per-rank peak
 * tuples are combined with one small NCCL allgather).  One process per GPU: xc_set_device() first, rank 0 makes the
 * 128-byte NCCL unique id and hands it to the other ranks by any host channel, every rank calls xc_dist_init().
 * xc_dist_allgather_peak(): `mine` (host) -> `all` (host, world entries, rank-ordered; .rank filled in); *best_rank
 * (optional) = the winner of the deterministic reduction (largest QF^2, lowest shift on ties, negative shift = empty
 * shard).  The copies and the ncclAllGather are enqueued on `stream`; the call returns when `all` is host-visible.
 * NCCL is loaded at run time (libnccl.so.2); XC_ERR_UNSUPPORTED when it is absent.  xc_peak_argmax() is the same
 * reduction on caller-gathered tuples (what pydsproutines_b200/dist.py uses after torch.distributed's allgather). */
int32_t xc_dist_unique_id(void* out128);
int32_t xc_dist_init(int32_t rank, int32_t world, const void* unique_id128);
int32_t xc_dist_allgather_peak(const xc_peak* mine, xc_peak* all, int32_t* best_rank, void* stream);
int32_t xc_dist_finalize(void);
int32_t xc_peak_argmax(const xc_peak* all, int32_t world);

#ifdef __cplusplus
}
#endif
#endif /* XCORR_B200_H */
