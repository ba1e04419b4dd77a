// api.cu -- the C-ABI of libxcorr_b200 (include/xcorr_b200.h): argument checks, host<->device
// staging for XC_HOST callers, error strings.  All arithmetic is in the kernels; nothing here
// computes on the CPU.
#include <stdarg.h>

#include <atomic>

#include "internal.cuh"

namespace xc {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
int32_t fail(int32_t code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
void count_launch(int n) { g_launches += n; }

std::recursive_mutex& api_mutex() {
    static std::recursive_mutex m;
    return m;
}

// ---- per-device state ----------------------------------------------------------------------------
struct SlotBuf {
    void* ptr = nullptr;
    size_t bytes = 0;
};
constexpr int PIPE_MAX = 16;      // chunks of the pipelined host path
struct DeviceState {
    SlotBuf slots[SLOT_COUNT];
    int sm_count = 0, smem_optin = 0;
    cudaEvent_t prof_e0 = nullptr, prof_e1 = nullptr;
    bool prof_valid = false;
    // host pipeline (XC_HOST frequency search): copy-in / copy-out streams and per-chunk events
    cudaStream_t s_in = nullptr, s_out = nullptr;
    cudaEvent_t ev_in[PIPE_MAX] = {}, ev_k[PIPE_MAX] = {};
};
static DeviceState g_devs[XC_MAX_DEVICES];
static bool g_prof_on = false;

static DeviceState& dev_state() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= XC_MAX_DEVICES) dev = 0;
    DeviceState& d = g_devs[dev];
    if (!d.sm_count) {
        cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, dev);
        cudaDeviceGetAttribute(&d.smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    }
    return d;
}

void profile_begin(cudaStream_t st) {
    if (!g_prof_on) return;
    DeviceState& d = dev_state();
    if (!d.prof_e0) {
        cudaEventCreate(&d.prof_e0);
        cudaEventCreate(&d.prof_e1);
    }
    cudaEventRecord(d.prof_e0, st);
}
void profile_end(cudaStream_t st) {
    if (!g_prof_on) return;
    DeviceState& d = dev_state();
    cudaEventRecord(d.prof_e1, st);
    d.prof_valid = true;
}

int sm_count() {
    const int n = dev_state().sm_count;
    return n > 0 ? n : 1;
}
int max_smem_optin() { return dev_state().smem_optin; }
int current_device() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
    return dev;
}
int32_t check_handle_device(int handle_device, const char* what) {
    const int dev = current_device();
    if (dev != handle_device)
        return fail(XC_ERR_ARG, "%s handle was created on device %d, the current device is %d (xc_set_device)", what, handle_device, dev);
    return XC_OK;
}
void* workspace(Slot slot, size_t bytes) {
    SlotBuf& s = dev_state().slots[slot];
    if (bytes == 0) bytes = 256;
    if (s.bytes >= bytes) return s.ptr;
    if (s.ptr) {
        cudaDeviceSynchronize();   // the old block may still be in use by enqueued work
        cudaFree(s.ptr);
        s.ptr = nullptr;
        s.bytes = 0;
    }
    size_t want = bytes + bytes / 8;
    want = (want + 255) & ~(size_t)255;
    cudaError_t e = cudaMalloc(&s.ptr, want);
    if (e != cudaSuccess) {
        cudaGetLastError();
        e = cudaMalloc(&s.ptr, bytes);
        want = bytes;
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        s.ptr = nullptr;
        fail(XC_ERR_NOMEM, "device allocation of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
        return nullptr;
    }
    s.bytes = want;
    return s.ptr;
}
void release_plan_caches();   // fastxcorr.cu: Bluestein / chirp-z plans of the current device
int32_t release_workspace() {
    DeviceState& d = dev_state();
    cudaDeviceSynchronize();
    for (int i = 0; i < SLOT_COUNT; ++i) {
        if (d.slots[i].ptr) cudaFree(d.slots[i].ptr);
        d.slots[i] = SlotBuf();
    }
    release_plan_caches();
    return XC_OK;
}

// ---- staging helpers -------------------------------------------------------------------------------
// Input: returns a device pointer holding `bytes` bytes of `p` (p itself in device mode).
static const void* stage_in(Slot slot, const void* p, size_t bytes, int where, cudaStream_t st, int32_t* status) {
    if (where == XC_DEVICE || p == nullptr) return p;
    void* d = workspace(slot, bytes);
    if (!d) {
        *status = XC_ERR_NOMEM;
        return nullptr;
    }
    if (bytes) {
        cudaError_t e = cudaMemcpyAsync(d, p, bytes, cudaMemcpyHostToDevice, st);
        if (e != cudaSuccess) *status = fail(XC_ERR_CUDA, "H2D copy failed: %s", cudaGetErrorString(e));
    }
    return d;
}
static void* stage_out(Slot slot, void* p, size_t bytes, int where, int32_t* status) {
    if (where == XC_DEVICE || p == nullptr) return p;
    void* d = workspace(slot, bytes);
    if (!d) *status = XC_ERR_NOMEM;
    return d;
}
static int32_t fetch_out(void* host, const void* dev, size_t bytes, int where, cudaStream_t st) {
    if (where == XC_DEVICE || host == nullptr || bytes == 0) return XC_OK;
    XC_CUDA_TRY(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, st));
    return XC_OK;
}
static int32_t finish(int where, cudaStream_t st) {
    if (where == XC_HOST) XC_CUDA_TRY(cudaStreamSynchronize(st));
    return XC_OK;
}
static int32_t check_where(int where) {
    if (where != XC_HOST && where != XC_DEVICE) return fail(XC_ERR_ARG, "where must be XC_HOST (0) or XC_DEVICE (1)");
    return XC_OK;
}
static int32_t require_device() {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail(XC_ERR_CUDA, "no CUDA device available (%s); libxcorr_b200 has no CPU fallback",
                    e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    }
    return XC_OK;
}

// Fills *peak from the device arrays of one call (synchronises the stream).
int32_t collect_peak(const float* d_qf2, const uint32_t* d_fidx, const int32_t* d_aux, int64_t K, int64_t shift0,
                     int64_t step, const int64_t* d_shifts, xc_peak* peak, cudaStream_t st) {
    memset(peak, 0, sizeof(*peak));
    peak->index = -1;
    if (K <= 0) return XC_OK;
    unsigned long long* d_packed = reinterpret_cast<unsigned long long*>(workspace(SLOT_TMP3, 256));
    if (!d_packed) return XC_ERR_NOMEM;
    // one reduction, one 24-byte copy-out: {packed (value bits, ~index), (fidx, aux), shifts[index]}
    XC_TRY(peak_reduce_device(d_qf2, d_fidx, d_aux, K, d_packed, st, d_shifts));
    unsigned long long rec[3] = {0, 0, 0};
    XC_CUDA_TRY(cudaMemcpyAsync(rec, d_packed, sizeof(rec), cudaMemcpyDeviceToHost, st));
    XC_CUDA_TRY(cudaStreamSynchronize(st));
    const uint32_t vb = (uint32_t)(rec[0] >> 32);
    memcpy(&peak->qf2, &vb, 4);
    peak->index = (int64_t)(0xffffffffu - (uint32_t)(rec[0] & 0xffffffffull));
    peak->fidx = (uint32_t)(rec[1] & 0xffffffffull);
    peak->aux = (int32_t)(uint32_t)(rec[1] >> 32);
    peak->shift = d_shifts ? (int64_t)rec[2] : shift0 + peak->index * step;
    return XC_OK;
}

static int32_t fastxcorr_host_pipelined(const float2* cutout, int32_t N, const float2* rx, int64_t rxlen, int64_t shift0,
                                        int64_t K, float* qf2, uint32_t* fidx, xc_peak* peak, cudaStream_t st);

}  // namespace xc

using namespace xc;

extern "C" {

int32_t xc_version(void) { return 100; }
const char* xc_last_error(void) { return g_err; }
int32_t xc_device_count(void) {
    XC_API_GUARD;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}
int32_t xc_set_device(int32_t device) {
    XC_API_GUARD;
    XC_TRY(require_device());
    XC_CUDA_TRY(cudaSetDevice(device));
    return XC_OK;
}
int32_t xc_pointer_device(const void* p) {
    cudaPointerAttributes a;
    if (!p || cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return -1;
    }
    return (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) ? a.device : -1;
}
int32_t xc_release_workspace(void) {
    XC_API_GUARD;
    return release_workspace();
}
int32_t xc_device_info(int32_t* out4) {
    XC_API_GUARD;
    if (!out4) return fail(XC_ERR_ARG, "out4 is NULL");
    XC_TRY(require_device());
    int dev = 0, clock = 0, major = 0, minor = 0;
    XC_CUDA_TRY(cudaGetDevice(&dev));
    cudaDeviceGetAttribute(&clock, cudaDevAttrClockRate, dev);
    cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
    out4[0] = sm_count();
    out4[1] = max_smem_optin();
    out4[2] = clock;
    out4[3] = major * 10 + minor;
    return XC_OK;
}
int64_t xc_launch_count(void) { return g_launches.load(); }
int32_t xc_profile_enable(int32_t on) {
    XC_API_GUARD;
    g_prof_on = on != 0;
    dev_state().prof_valid = false;
    return XC_OK;
}
double xc_profile_last_ms(void) {
    XC_API_GUARD;
    DeviceState& d = dev_state();
    if (!d.prof_valid) return -1.0;
    if (cudaEventSynchronize(d.prof_e1) != cudaSuccess) return -1.0;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, d.prof_e0, d.prof_e1) != cudaSuccess) return -1.0;
    return (double)ms;
}

int32_t xc_sliding_energy_c64(const void* x, int64_t n, int32_t L, float* out, int32_t where, void* stream) {
    XC_API_GUARD;
    XC_TRY(check_where(where));
    if (!x || !out) return fail(XC_ERR_ARG, "x/out is NULL");
    if (L < 1 || n < L) return fail(XC_ERR_ARG, "need 1 <= L <= n (L=%d, n=%lld)", L, (long long)n);
    XC_TRY(require_device());
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t n_out = n - L + 1;
    int32_t status = XC_OK;
    const void* dx = stage_in(SLOT_IN0, x, sizeof(float2) * (size_t)n, where, st, &status);
    float* dout = (float*)stage_out(SLOT_OUT0, out, sizeof(float) * (size_t)n_out, where, &status);
    XC_TRY(status);
    XC_TRY(sliding_sum_device(dx, true, n, 0, L, n_out, dout, 0, st));
    XC_TRY(fetch_out(out, dout, sizeof(float) * (size_t)n_out, where, st));
    return finish(where, st);
}

int32_t xc_moving_average_f32(const float* x, int64_t n, int32_t L, int32_t sum_instead, float* out, int32_t where,
                              void* stream) {
    XC_API_GUARD;
    XC_TRY(check_where(where));
    if (!x || !out) return fail(XC_ERR_ARG, "x/out is NULL");
    if (L < 1) return fail(XC_ERR_ARG, "avgLength must be >= 1");
    if (n <= 0) return XC_OK;
    XC_TRY(require_device());
    cudaStream_t st = (cudaStream_t)stream;
    int32_t status = XC_OK;
    const void* dx = stage_in(SLOT_IN0, x, sizeof(float) * (size_t)n, where, st, &status);
    float* dout = (float*)stage_out(SLOT_OUT0, out, sizeof(float) * (size_t)n, where, &status);
    XC_TRY(status);
    XC_TRY(sliding_sum_device(dx, false, n, (int64_t)L - 1, L, n, dout, sum_instead ? 0 : 1, st));
    XC_TRY(fetch_out(out, dout, sizeof(float) * (size_t)n, where, st));
    return finish(where, st);
}

int32_t xc_complex_magnsq_c64(const void* x, int64_t n, void* out, int32_t out_f64, int32_t where, void* stream) {
    XC_API_GUARD;
    XC_TRY(check_where(where));
    if (!x || !out) return fail(XC_ERR_ARG, "x/out is NULL");
    if (n <= 0) return XC_OK;
    XC_TRY(require_device());
    cudaStream_t st = (cudaStream_t)stream;
    const size_t ob = (out_f64 ? 8 : 4) * (size_t)n;
    int32_t status = XC_OK;
    const void* dx = stage_in(SLOT_IN0, x, sizeof(float2) * (size_t)n, where, st, &status);
    void* dout = stage_out(SLOT_OUT0, out, ob, where, &status);
    XC_TRY(status);
    XC_TRY(magnsq_device((const float2*)dx, n, dout, out_f64, st));
    XC_TRY(fetch_out(out, dout, ob, where, st));
    return finish(where, st);
}

int32_t xc_fastxcorr_supported(int32_t N) { return fastxcorr_fft_supported(N) ? 1 : 0; }

static int32_t check_shift_args(int32_t N, int64_t rxlen, int64_t shift0, int64_t step, int64_t K,
                                const int64_t* shifts, int32_t flags) {
    if (N < 1) return fail(XC_ERR_ARG, "cutout length must be >= 1");
    if (K < 0) return fail(XC_ERR_ARG, "number of shifts must be >= 0");
    if (rxlen < 0) return fail(XC_ERR_ARG, "rxlen must be >= 0");
    if (!shifts && !(flags & XC_FLAG_ZERO_OOB) && K > 0) {
        const int64_t last = shift0 + (K - 1) * step;
        const int64_t lo = shift0 < last ? shift0 : last, hi = shift0 < last ? last : shift0;
        if (lo < 0 || hi + N > rxlen)
            return fail(XC_ERR_RANGE, "shifts [%lld, %lld] with cutout length %d leave rx of length %lld",
                        (long long)lo, (long long)hi, N, (long long)rxlen);
    }
    return XC_OK;
}

int32_t xc_fastxcorr_c64(const void* cutout, int32_t N, const void* rx, int64_t rxlen, int64_t shift0, int64_t step,
                         int64_t K, const int64_t* shifts, int32_t freqsearch, int32_t flags, float* qf2,
                         uint32_t* fidx, void* cplx, xc_peak* peak, int32_t where, void* stream) {
    XC_API_GUARD;
    XC_TRY(check_where(where));
    if (!cutout || !rx) return fail(XC_ERR_ARG, "cutout/rx is NULL");
    if (!qf2 && !cplx) return fail(XC_ERR_ARG, "qf2 and cplx are both NULL");
    XC_TRY(check_shift_args(N, rxlen, shift0, step, K, shifts, flags));
    if (freqsearch && !fastxcorr_fft_supported(N))
        return fail(XC_ERR_UNSUPPORTED, "fused frequency search does not take a cutout of %d samples", N);
    XC_TRY(require_device());
    if (K == 0) {
        if (peak) { memset(peak, 0, sizeof(*peak)); peak->index = -1; }
        return XC_OK;
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (where == XC_HOST && freqsearch && !shifts && step == 1 && !cplx && K >= (1 << 17) && shift0 >= 0 &&
        shift0 + K - 1 + N <= rxlen && fft_length_pow2(N))
        return fastxcorr_host_pipelined((const float2*)cutout, N, (const float2*)rx, rxlen, shift0, K, qf2, fidx, peak, st);
    int32_t status = XC_OK;
    const float2* dc = (const float2*)stage_in(SLOT_IN0, cutout, sizeof(float2) * (size_t)N, where, st, &status);
    const float2* drx = (const float2*)stage_in(SLOT_IN1, rx, sizeof(float2) * (size_t)rxlen, where, st, &status);
    const int64_t* dsh = (const int64_t*)stage_in(SLOT_IN2, shifts, sizeof(int64_t) * (size_t)K, where, st, &status);
    // qf2 is always produced on the device when a peak is wanted
    float* dq = (float*)stage_out(SLOT_OUT0, qf2, sizeof(float) * (size_t)K, where, &status);
    if (!dq && peak) dq = (float*)workspace(SLOT_OUT0, sizeof(float) * (size_t)K);
    uint32_t* df = (uint32_t*)stage_out(SLOT_OUT1, fidx, sizeof(uint32_t) * (size_t)K, where, &status);
    if (!df && peak && freqsearch) df = (uint32_t*)workspace(SLOT_OUT1, sizeof(uint32_t) * (size_t)K);
    float2* dz = (float2*)stage_out(SLOT_OUT2, cplx, sizeof(float2) * (size_t)K, where, &status);
    XC_TRY(status);
    ShiftList sl{shift0, step, K, dsh};
    if (freqsearch) {
        XC_TRY(fastxcorr_freq_device(dc, N, drx, rxlen, sl, dq, df, dz, nullptr, 0, st));
    } else {
        XC_TRY(fastxcorr_nofreq_device(dc, N, drx, rxlen, sl, dq, dz, st));
        if (fidx && df) XC_CUDA_TRY(cudaMemsetAsync(df, 0, sizeof(uint32_t) * (size_t)K, st));
    }
    XC_TRY(fetch_out(qf2, dq, sizeof(float) * (size_t)K, where, st));
    XC_TRY(fetch_out(fidx, df, sizeof(uint32_t) * (size_t)K, where, st));
    XC_TRY(fetch_out(cplx, dz, sizeof(float2) * (size_t)K, where, st));
    if (peak) XC_TRY(collect_peak(dq, freqsearch ? df : nullptr, nullptr, K, shift0, step, dsh, peak, st));
    return finish(where, st);
}

}  // extern "C" (reopened below)

namespace xc {
// XC_HOST frequency search over a contiguous in-range shift list, pipelined: the shift range is cut into chunks and
// H2D(rx part of chunk i + 1) | kernels(chunk i) | D2H(results of chunk i - 1) run on three streams, so that only the
// first copy-in and the last copy-out are exposed (pinned caller buffers overlap fully; pageable ones are staged by
// the driver and still copy while the previous chunk computes).  One cudaMemcpyAsync per chunk and direction.
static int32_t ensure_pipe(DeviceState& d) {
    if (d.s_in) return XC_OK;
    XC_CUDA_TRY(cudaStreamCreateWithFlags(&d.s_in, cudaStreamNonBlocking));
    XC_CUDA_TRY(cudaStreamCreateWithFlags(&d.s_out, cudaStreamNonBlocking));
    for (int i = 0; i < PIPE_MAX; ++i) {
        XC_CUDA_TRY(cudaEventCreateWithFlags(&d.ev_in[i], cudaEventDisableTiming));
        XC_CUDA_TRY(cudaEventCreateWithFlags(&d.ev_k[i], cudaEventDisableTiming));
    }
    return XC_OK;
}

static int32_t fastxcorr_host_pipelined(const float2* cutout, int32_t N, const float2* rx, int64_t rxlen, int64_t shift0,
                                        int64_t K, float* qf2, uint32_t* fidx, xc_peak* peak, cudaStream_t st) {
    DeviceState& d = dev_state();
    XC_TRY(ensure_pipe(d));
    int nch = (int)(K >> 18);                    // about 2^18 shifts (2 ms of search at N = 8192) per chunk
    if (nch < 2) nch = 2;
    if (nch > PIPE_MAX) nch = PIPE_MAX;
    float2* dc = (float2*)workspace(SLOT_IN0, sizeof(float2) * (size_t)N);
    float2* drx = (float2*)workspace(SLOT_IN1, sizeof(float2) * (size_t)rxlen);
    float* dq = (float*)workspace(SLOT_OUT0, sizeof(float) * (size_t)K);
    uint32_t* df = (uint32_t*)workspace(SLOT_OUT1, sizeof(uint32_t) * (size_t)K);
    if (!dc || !drx || !dq || !df) return XC_ERR_NOMEM;
    // earlier work of the caller's stream may still read the staging slots
    XC_CUDA_TRY(cudaEventRecord(d.ev_k[0], st));
    XC_CUDA_TRY(cudaStreamWaitEvent(d.s_in, d.ev_k[0], 0));
    XC_CUDA_TRY(cudaStreamWaitEvent(d.s_out, d.ev_k[0], 0));
    XC_CUDA_TRY(cudaMemcpyAsync(dc, cutout, sizeof(float2) * (size_t)N, cudaMemcpyHostToDevice, d.s_in));
    // short first and last chunks (a quarter of an even share): only the first copy-in and the last copy-out are exposed
    if (nch < 3) nch = 3;
    int64_t k0[PIPE_MAX + 1];
    const int64_t edge = K / (4 * (int64_t)nch);
    k0[0] = 0;
    k0[nch] = K;
    for (int c = 1; c < nch; ++c) k0[c] = edge + ((K - 2 * edge) * (c - 1)) / (nch - 2);
    int64_t copied = shift0;                     // rx[shift0, copied) is on its way
    for (int c = 0; c < nch; ++c) {
        const int64_t need = shift0 + k0[c + 1] + N - 1;     // chunk c reads rx[shift0 + k0[c], need)
        if (need > copied) {
            XC_CUDA_TRY(cudaMemcpyAsync(drx + copied, rx + copied, sizeof(float2) * (size_t)(need - copied),
                                        cudaMemcpyHostToDevice, d.s_in));
            copied = need;
        }
        XC_CUDA_TRY(cudaEventRecord(d.ev_in[c], d.s_in));
    }
    for (int c = 0; c < nch; ++c) {
        const int64_t kc = k0[c + 1] - k0[c];
        if (kc <= 0) continue;
        XC_CUDA_TRY(cudaStreamWaitEvent(st, d.ev_in[c], 0));
        ShiftList sl{shift0 + k0[c], 1, kc, nullptr};
        XC_TRY(fastxcorr_freq_device(dc, N, drx, rxlen, sl, dq + k0[c], df + k0[c], nullptr, nullptr, 0, st));
        XC_CUDA_TRY(cudaEventRecord(d.ev_k[c], st));
        XC_CUDA_TRY(cudaStreamWaitEvent(d.s_out, d.ev_k[c], 0));
        if (qf2) XC_CUDA_TRY(cudaMemcpyAsync(qf2 + k0[c], dq + k0[c], sizeof(float) * (size_t)kc, cudaMemcpyDeviceToHost, d.s_out));
        if (fidx) XC_CUDA_TRY(cudaMemcpyAsync(fidx + k0[c], df + k0[c], sizeof(uint32_t) * (size_t)kc, cudaMemcpyDeviceToHost, d.s_out));
    }
    if (peak) XC_TRY(collect_peak(dq, df, nullptr, K, shift0, 1, nullptr, peak, st));
    XC_CUDA_TRY(cudaStreamSynchronize(d.s_out));
    XC_CUDA_TRY(cudaStreamSynchronize(st));
    return XC_OK;
}
}  // namespace xc

extern "C" {

int32_t xc_fastxcorr_caf_c64(const void* cutout, int32_t N, const void* rx, int64_t rxlen, int64_t shift0,
                             int64_t step, int64_t K, const int64_t* shifts, void* caf, int32_t complex_out,
                             int32_t where, void* stream) {
    XC_API_GUARD;
    XC_TRY(check_where(where));
    if (!cutout || !rx || !caf) return fail(XC_ERR_ARG, "cutout/rx/caf is NULL");
    XC_TRY(check_shift_args(N, rxlen, shift0, step, K, shifts, 0));
    if (!fastxcorr_fft_supported(N))
        return fail(XC_ERR_UNSUPPORTED, "fused frequency search does not take a cutout of %d samples", N);
    XC_TRY(require_device());
    if (K == 0) return XC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t cb = (complex_out ? sizeof(float2) : sizeof(float)) * (size_t)K * (size_t)N;
    int32_t status = XC_OK;
    const float2* dc = (const float2*)stage_in(SLOT_IN0, cutout, sizeof(float2) * (size_t)N, where, st, &status);
    const float2* drx = (const float2*)stage_in(SLOT_IN1, rx, sizeof(float2) * (size_t)rxlen, where, st, &status);
    const int64_t* dsh = (const int64_t*)stage_in(SLOT_IN2, shifts, sizeof(int64_t) * (size_t)K, where, st, &status);
    void* dcaf = stage_out(SLOT_OUT0, caf, cb, where, &status);
    XC_TRY(status);
    ShiftList sl{shift0, step, K, dsh};
    XC_TRY(fastxcorr_freq_device(dc, N, drx, rxlen, sl, nullptr, nullptr, nullptr, dcaf, complex_out, st));
    XC_TRY(fetch_out(caf, dcaf, cb, where, st));
    return finish(where, st);
}

int32_t xc_cztxcorr_c64(const void* cutout, int32_t N, const void* rx, int64_t rxlen, int64_t shift0, int64_t step,
                        int64_t K, const int64_t* shifts, double f0_cycles, double df_cycles, int32_t nbins, float* qf2,
                        uint32_t* fidx, void* cplx, float* caf, int32_t where, void* stream) {
    XC_API_GUARD;
    XC_TRY(check_where(where));
    if (!cutout || !rx) return fail(XC_ERR_ARG, "cutout/rx is NULL");
    if (!qf2 && !cplx && !caf) return fail(XC_ERR_ARG, "qf2, cplx and caf are all NULL");
    if (nbins < 1) return fail(XC_ERR_ARG, "need at least one frequency bin");
    XC_TRY(check_shift_args(N, rxlen, shift0, step, K, shifts, 0));
    XC_TRY(require_device());
    if (K == 0) return XC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t cb = sizeof(float) * (size_t)K * (size_t)nbins;
    int32_t status = XC_OK;
    const float2* dc = (const float2*)stage_in(SLOT_IN0, cutout, sizeof(float2) * (size_t)N, where, st, &status);
    const float2* drx = (const float2*)stage_in(SLOT_IN1, rx, sizeof(float2) * (size_t)rxlen, where, st, &status);
    const int64_t* dsh = (const int64_t*)stage_in(SLOT_IN2, shifts, sizeof(int64_t) * (size_t)K, where, st, &status);
    float* dq = (float*)stage_out(SLOT_OUT0, qf2, sizeof(float) * (size_t)K, where, &status);
    uint32_t* df = (uint32_t*)stage_out(SLOT_OUT1, fidx, sizeof(uint32_t) * (size_t)K, where, &status);
    float2* dz = (float2*)stage_out(SLOT_OUT2, cplx, sizeof(float2) * (size_t)K, where, &status);
    float* dcaf = (float*)stage_out(SLOT_TMP3, caf, cb, where, &status);
    XC_TRY(status);
    ShiftList sl{shift0, step, K, dsh};
    const CztSpec spec{nbins, f0_cycles, df_cycles};
    if (dq || df || dz) XC_TRY(fastxcorr_freq_device(dc, N, drx, rxlen, sl, dq, df, dz, nullptr, 0, st, &spec));
    if (dcaf) XC_TRY(fastxcorr_freq_device(dc, N, drx, rxlen, sl, nullptr, nullptr, nullptr, dcaf, 0, st, &spec));
    XC_TRY(fetch_out(qf2, dq, sizeof(float) * (size_t)K, where, st));
    XC_TRY(fetch_out(fidx, df, sizeof(uint32_t) * (size_t)K, where, st));
    XC_TRY(fetch_out(cplx, dz, sizeof(float2) * (size_t)K, where, st));
    XC_TRY(fetch_out(caf, dcaf, cb, where, st));
    return finish(where, st);
}

// ---- stand-alone chirp-z transform of the rows of a matrix (CZTCached.run / runMany, spectralRoutines.py:263-306) ----
// The correlator kernel with an all-ones cutout computes X_r[k] / (sqrt(m) ||x_r||) for every row; the window energies
// it leaves in the workspace undo the normalisation.
namespace xc {
__global__ void czt_ones_kernel(float2* p, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = make_float2(1.f, 0.f);
}
__global__ void czt_unscale_kernel(float2* out, int64_t rows, int nbins, const float* __restrict__ energy, float sqrt_m) {
    const int64_t total = rows * nbins;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
        const float en = energy[e / nbins];
        const float sc = sqrt_m * sqrtf(en);
        const float2 v = out[e];
        out[e] = (en > 0.f) ? make_float2(v.x * sc, v.y * sc) : make_float2(0.f, 0.f);
    }
}
}  // namespace xc

int32_t xc_czt_c64(const void* x, int32_t m, int64_t rows, double f0_cycles, double df_cycles, int32_t nbins, void* out,
                   int32_t where, void* stream) {
    XC_API_GUARD;
    XC_TRY(check_where(where));
    if (!x || !out) return fail(XC_ERR_ARG, "x/out is NULL");
    if (m < 2 || nbins < 1 || rows < 0) return fail(XC_ERR_ARG, "need m >= 2, nbins >= 1, rows >= 0");
    XC_TRY(require_device());
    if (rows == 0) return XC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t ob = sizeof(float2) * (size_t)rows * (size_t)nbins;
    int32_t status = XC_OK;
    const float2* dx = (const float2*)stage_in(SLOT_IN1, x, sizeof(float2) * (size_t)rows * (size_t)m, where, st, &status);
    float2* dout = (float2*)stage_out(SLOT_OUT2, out, ob, where, &status);
    XC_TRY(status);
    float2* d_ones = (float2*)workspace(SLOT_TMP3, sizeof(float2) * (size_t)m);
    if (!d_ones) return XC_ERR_NOMEM;
    czt_ones_kernel<<<(m + 255) / 256, 256, 0, st>>>(d_ones, m);
    XC_LAUNCH_CHECK();
    ShiftList sl{0, (int64_t)m, rows, nullptr};                       // row r = the window at shift r * m
    const CztSpec spec{nbins, f0_cycles, df_cycles};
    XC_TRY(fastxcorr_freq_device(d_ones, m, dx, rows * (int64_t)m, sl, nullptr, nullptr, nullptr, dout, 1, st, &spec));
    const float* d_energy = (const float*)workspace(SLOT_TMP1, sizeof(float) * (size_t)rows);   // left there by the call above
    if (!d_energy) return XC_ERR_NOMEM;
    int64_t blocks = ((int64_t)rows * nbins + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    czt_unscale_kernel<<<(unsigned)blocks, 256, 0, st>>>(dout, rows, nbins, d_energy, sqrtf((float)m));
    XC_LAUNCH_CHECK();
    XC_TRY(fetch_out(out, dout, ob, where, st));
    return finish(where, st);
}

int32_t xc_row_argmax_c64(const void* x, int64_t rows, int64_t cols, int32_t use_normsq, uint32_t* idx, float* val,
                          int32_t where, void* stream) {
    XC_API_GUARD;
    XC_TRY(check_where(where));
    if (!x || !idx) return fail(XC_ERR_ARG, "x/idx is NULL");
    if (rows < 0 || cols < 1) return fail(XC_ERR_ARG, "need rows >= 0 and cols >= 1");
    XC_TRY(require_device());
    if (rows == 0) return XC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    int32_t status = XC_OK;
    const float2* dx = (const float2*)stage_in(SLOT_IN0, x, sizeof(float2) * (size_t)rows * (size_t)cols, where, st, &status);
    uint32_t* di = (uint32_t*)stage_out(SLOT_OUT0, idx, sizeof(uint32_t) * (size_t)rows, where, &status);
    float* dv = (float*)stage_out(SLOT_OUT1, val, sizeof(float) * (size_t)rows, where, &status);
    XC_TRY(status);
    XC_TRY(row_argmax_device(dx, rows, cols, use_normsq, di, dv, st));
    XC_TRY(fetch_out(idx, di, sizeof(uint32_t) * (size_t)rows, where, st));
    XC_TRY(fetch_out(val, dv, sizeof(float) * (size_t)rows, where, st));
    return finish(where, st);
}

int32_t xc_find_local_maxima_f32(const float* x, int64_t n, float min_height, int32_t* idx, int32_t cap, int32_t* count,
                                 int32_t sorted, int32_t where, void* stream) {
    XC_API_GUARD;
    (void)sorted;   // the compaction is ordered, so the result is always ascending
    XC_TRY(check_where(where));
    if (!x || !idx || !count) return fail(XC_ERR_ARG, "x/idx/count is NULL");
    if (cap < 0 || n < 0) return fail(XC_ERR_ARG, "cap and n must be >= 0");
    XC_TRY(require_device());
    cudaStream_t st = (cudaStream_t)stream;
    int32_t status = XC_OK;
    const float* dx = (const float*)stage_in(SLOT_IN0, x, sizeof(float) * (size_t)n, where, st, &status);
    int32_t* di = (int32_t*)stage_out(SLOT_OUT0, idx, sizeof(int32_t) * (size_t)(cap > 0 ? cap : 1), where, &status);
    int32_t* dcnt = (int32_t*)stage_out(SLOT_OUT1, count, sizeof(int32_t), where, &status);
    XC_TRY(status);
    XC_TRY(local_maxima_device(dx, n, min_height, di, cap, dcnt, st));
    XC_TRY(fetch_out(idx, di, sizeof(int32_t) * (size_t)cap, where, st));
    XC_TRY(fetch_out(count, dcnt, sizeof(int32_t), where, st));
    return finish(where, st);
}

}  // extern "C"
