// groupfreq.cu -- grouped correlator over an ARBITRARY frequency list, on the tensor cores.
//     d[s, k] = sum_g sum_t y_g[t] * rx[s + s_g + t] * exp(-j 2 pi f_k (s_g + t) / fs)
//     QF^2[s, k] = |d[s, k]|^2 / sum_g ||rx[s + s_g : + len_g]||^2 / ||y||^2
// Replaces the per-shift dense DFT of GroupXcorr.xcorr (xcorrRoutines.py:852-954: freqMat @ (rxconcat * yconcat),
// max over the list), the per-group CZT + group phase of GroupXcorrCZT.xcorr (xcorrRoutines.py:957-1039: the CAF
// over the grid f1 : binWidth : f2) and the experimental RawKernels of GroupXcorrGPU (xcorrRoutines.py:1693-2051).
//
// The frequency factor is folded into the template: T_k[g, t] = y_g[t] exp(-j 2 pi f_k (s_g + t) / fs), so the
// whole search is a multi-template correlation d[s, k] = sum_n rx[s + n] T_k[n] whose templates are non-zero only
// inside the G groups.  mt_tc_kernel runs it as a segmented contraction: M = 128 shifts, N = frequencies, K = the
// concatenated group samples; each group brings its own Hankel x tile (x offset s_g), the accumulators stay in
// TMEM across the groups.  The samples between the groups are never touched.
#include <algorithm>
#include <map>
#include <vector>

#include "internal.cuh"

struct xc_groupfreq {
    int32_t G, NF, Lcat;
    int64_t max_end;                 // max over groups of start + length
    int64_t start0;                  // single group: its start (the dense kernel has no per-group offset)
    float2* d_templates;             // [NF, Lcat] (released after the BF16 pack is built)
    float* d_energies;               // [NF] all = ||y||^2
    xc::MtTcPlan tc;
    int32_t device = 0;              // the device that owns the arrays of this handle
    // window energies: one sliding-sum pass per distinct group length
    std::vector<int32_t> lens;
    std::vector<int64_t*> d_starts;  // per distinct length: device start list
    std::vector<int32_t> counts;
};

namespace xc {

// out[i] (+)= sum_g el[shift0 + i + starts[g]]
__global__ void groupfreq_energy_kernel(const float* __restrict__ el, int64_t el_len, const int64_t* __restrict__ starts,
                                        int n, int64_t shift0, int64_t K, float* __restrict__ out, int accumulate) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < K; i += stride) {
        double acc = accumulate ? (double)out[i] : 0.0;
        for (int g = 0; g < n; ++g) {
            const int64_t p = shift0 + i + starts[g];
            if (p >= 0 && p < el_len) acc += (double)__ldg(el + p);
        }
        out[i] = (float)acc;
    }
}

// caf[s, b] = |sum_g surf[tmpl[g], s, b]|^2 / sum_g en[grp[g], s] / ynormsq   (GroupXcorrCZT_Permutations.getCAF)
__global__ void groupperm_combine_kernel(const float2* __restrict__ surf, int64_t K, int NB, const int32_t* __restrict__ sel,
                                         int G, const float* __restrict__ en, double inv_ynormsq, float* __restrict__ caf) {
    const int64_t total = K * NB;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
        const int64_t s = e / NB;
        float2 acc = make_float2(0.f, 0.f);
        double rn = 0.0;
        for (int g = 0; g < G; ++g) {
            const float2 v = __ldg(surf + (size_t)sel[2 * g] * (size_t)total + e);
            acc.x += v.x;
            acc.y += v.y;
            rn += (double)__ldg(en + (size_t)sel[2 * g + 1] * (size_t)K + s);
        }
        caf[e] = (float)(((double)acc.x * acc.x + (double)acc.y * acc.y) / rn * inv_ynormsq);
    }
}

int32_t collect_peak(const float* d_qf2, const uint32_t* d_fidx, const int32_t* d_aux, int64_t K, int64_t shift0,
                     int64_t step, const int64_t* d_shifts, xc_peak* peak, cudaStream_t st);

}  // namespace xc

using namespace xc;

extern "C" {

int32_t xc_groupperm_combine(const void* surfaces, int64_t K, int32_t NB, const int32_t* template_idx,
                             const int32_t* group_idx, int32_t G, const float* rxnormsq, double ynormsq, float* caf,
                             void* stream) {
    XC_API_GUARD;
    if (!surfaces || !template_idx || !group_idx || !rxnormsq || !caf) return fail(XC_ERR_ARG, "NULL argument");
    if (K < 0 || NB < 1 || G < 1 || G > 4096) return fail(XC_ERR_ARG, "bad sizes");
    if (!(ynormsq > 0.0)) return fail(XC_ERR_ARG, "template energy must be positive");
    if (K == 0) return XC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    int32_t* d_sel = (int32_t*)workspace(SLOT_IN2, sizeof(int32_t) * 2 * (size_t)G);
    if (!d_sel) return XC_ERR_NOMEM;
    std::vector<int32_t> sel(2 * (size_t)G);
    for (int g = 0; g < G; ++g) {
        sel[2 * g] = template_idx[g];
        sel[2 * g + 1] = group_idx[g];
    }
    XC_CUDA_TRY(cudaMemcpyAsync(d_sel, sel.data(), sizeof(int32_t) * sel.size(), cudaMemcpyHostToDevice, st));
    XC_CUDA_TRY(cudaStreamSynchronize(st));          // sel is a stack-lifetime host buffer
    int64_t blocks = (K * NB + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    groupperm_combine_kernel<<<(unsigned)blocks, 256, 0, st>>>(reinterpret_cast<const float2*>(surfaces), K, NB, d_sel, G,
                                                               rxnormsq, 1.0 / ynormsq, caf);
    XC_LAUNCH_CHECK();
    return XC_OK;
}

int32_t xc_groupfreq_destroy(xc_groupfreq* h) {
    XC_API_GUARD;
    if (!h) return XC_OK;
    mt_tc_release(&h->tc);
    cudaFree(h->d_templates);
    cudaFree(h->d_energies);
    for (int64_t* p : h->d_starts) cudaFree(p);
    delete h;
    return XC_OK;
}

int32_t xc_groupfreq_create(const void* ycat, const int64_t* starts, const int32_t* lengths, int32_t G,
                            const double* freq_cycles, const double* start_freq_cycles, int32_t NF, int32_t conj,
                            xc_groupfreq** handle) {
    XC_API_GUARD;
    if (!ycat || !starts || !lengths || !freq_cycles || !handle) return fail(XC_ERR_ARG, "NULL argument");
    if (G < 1 || NF < 1) return fail(XC_ERR_ARG, "need at least one group and one frequency");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(XC_ERR_CUDA, "no CUDA device available; libxcorr_b200 has no CPU fallback");
    }
    int64_t max_end = 0, total = 0;
    for (int g = 0; g < G; ++g) {
        if (lengths[g] < 1) return fail(XC_ERR_ARG, "group %d has no samples", g);
        if (starts[g] < 0 || starts[g] + lengths[g] >= (int64_t(1) << 31)) return fail(XC_ERR_ARG, "group %d: start out of range", g);
        max_end = std::max<int64_t>(max_end, starts[g] + lengths[g]);
        total += lengths[g];
    }
    // segments: groups cut into pieces of at most 512 samples (the x tile of a piece must fit in shared memory),
    // every piece padded to whole K steps of 16 samples; an even number of steps overall
    constexpr int PIECE = 512;
    std::vector<int32_t> segs;
    std::vector<int64_t> col0(G);
    int64_t steps = 0;
    for (int g = 0; g < G; ++g) {
        col0[g] = steps * 16;
        for (int off = 0; off < lengths[g]; off += PIECE) {
            const int len = std::min(PIECE, lengths[g] - off);
            const int nks = (len + 15) / 16;
            segs.push_back((int32_t)(starts[g] + off));
            segs.push_back(nks);
            steps += nks;
        }
    }
    if (steps & 1) {
        segs.back() += 1;
        steps += 1;
    }
    const int nseg = (int)(segs.size() / 2);
    if (steps * 16 > (int64_t(1) << 24)) return fail(XC_ERR_UNSUPPORTED, "more than 2^24 template samples");
    const int Lcat = (int)(steps * 16);
    if ((size_t)NF * (size_t)Lcat > (size_t(1) << 28))
        return fail(XC_ERR_UNSUPPORTED, "frequency list x group samples = %zu exceeds 2^28", (size_t)NF * (size_t)Lcat);
    // templates in double, stored FP32
    const float2* y = reinterpret_cast<const float2*>(ycat);
    std::vector<float2> tm((size_t)NF * Lcat, make_float2(0.f, 0.f));
    double ynorm = 0.0;
    {
        size_t p = 0;
        for (int g = 0; g < G; ++g)
            for (int t = 0; t < lengths[g]; ++t, ++p) ynorm += (double)y[p].x * y[p].x + (double)y[p].y * y[p].y;
    }
    for (int k = 0; k < NF; ++k) {
        const double fc = freq_cycles[k], fcs = start_freq_cycles ? start_freq_cycles[k] : fc;
        size_t p = 0;
        float2* row = tm.data() + (size_t)k * Lcat;
        for (int g = 0; g < G; ++g)
            for (int t = 0; t < lengths[g]; ++t, ++p) {
                double ph = fc * (double)t, phs = fcs * (double)starts[g];
                ph = (ph - floor(ph)) + (phs - floor(phs));
                const double c = cos(-2.0 * M_PI * ph), s = sin(-2.0 * M_PI * ph);
                const double yr = y[p].x, yi = conj ? -(double)y[p].y : (double)y[p].y;
                row[col0[g] + t] = make_float2((float)(yr * c - yi * s), (float)(yr * s + yi * c));
            }
    }
    xc_groupfreq* h = new xc_groupfreq();
    h->device = current_device();
    h->G = G; h->NF = NF; h->Lcat = Lcat; h->max_end = max_end;
    h->d_templates = nullptr; h->d_energies = nullptr;
    std::vector<float> en(NF, (float)ynorm);
    cudaError_t e = cudaMalloc(&h->d_templates, sizeof(float2) * tm.size());
    if (e == cudaSuccess) e = cudaMalloc(&h->d_energies, sizeof(float) * NF);
    if (e == cudaSuccess) e = cudaMemcpy(h->d_templates, tm.data(), sizeof(float2) * tm.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(h->d_energies, en.data(), sizeof(float) * NF, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cudaGetLastError();
        xc_groupfreq_destroy(h);
        return fail(e == cudaErrorMemoryAllocation ? XC_ERR_NOMEM : XC_ERR_CUDA, "template upload failed: %s", cudaGetErrorString(e));
    }
    int32_t st = mt_tc_prepare(h->d_templates, NF, Lcat, Lcat, h->d_energies, &h->tc, nseg > 1 ? segs.data() : nullptr, nseg);
    h->start0 = (nseg == 1) ? starts[0] : 0;
    cudaFree(h->d_templates);        // only the BF16 stage images are used from here on
    h->d_templates = nullptr;
    if (st == XC_OK && !h->tc.usable)
        st = fail(XC_ERR_UNSUPPORTED, "the grouped frequency-list correlator needs the sm_100a tensor-core path "
                                      "(<= 2048 frequencies)");
    if (st != XC_OK) {
        xc_groupfreq_destroy(h);
        return st;
    }
    // start lists per distinct group length (window energies)
    std::map<int32_t, std::vector<int64_t>> by_len;
    for (int g = 0; g < G; ++g) by_len[lengths[g]].push_back(starts[g]);
    for (auto& kv : by_len) {
        int64_t* d = nullptr;
        if (cudaMalloc(&d, sizeof(int64_t) * kv.second.size()) != cudaSuccess ||
            cudaMemcpy(d, kv.second.data(), sizeof(int64_t) * kv.second.size(), cudaMemcpyHostToDevice) != cudaSuccess) {
            cudaGetLastError();
            cudaFree(d);
            xc_groupfreq_destroy(h);
            return fail(XC_ERR_NOMEM, "group start upload failed");
        }
        h->lens.push_back(kv.first);
        h->d_starts.push_back(d);
        h->counts.push_back((int32_t)kv.second.size());
    }
    *handle = h;
    return XC_OK;
}

int32_t xc_groupfreq_xcorr(xc_groupfreq* h, const void* rx, int64_t rxlen, int64_t shift0, int64_t K, float* qf2,
                           uint32_t* fidx, float* caf, void* cplx, xc_peak* peak, int32_t where, void* stream) {
    XC_API_GUARD;
    if (where != XC_HOST && where != XC_DEVICE) return fail(XC_ERR_ARG, "where must be XC_HOST (0) or XC_DEVICE (1)");
    if (!h || !rx) return fail(XC_ERR_ARG, "handle/rx is NULL");
    XC_TRY(check_handle_device(h->device, "grouped frequency-list correlator"));
    if (K < 0 || shift0 < 0) return fail(XC_ERR_RANGE, "shifts must be >= 0");
    if (K > 0 && shift0 + K - 1 + h->max_end > rxlen)
        return fail(XC_ERR_RANGE, "the last shift reads past the end of rx");            // xcorrRoutines.py:922-925, 1000-1003
    if (K == 0) {
        if (peak) { memset(peak, 0, sizeof(*peak)); peak->index = -1; }
        return XC_OK;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const size_t cafb = sizeof(float) * (size_t)K * (size_t)h->NF;
    const float2* drx = reinterpret_cast<const float2*>(rx);
    float *dq = qf2, *dcaf = caf;
    float2* dcplx = reinterpret_cast<float2*>(cplx);
    uint32_t* df = fidx;
    if (cplx && where == XC_HOST) return fail(XC_ERR_UNSUPPORTED, "the complex surface is a device-side output");
    if (where == XC_HOST) {
        float2* b = (float2*)workspace(SLOT_IN1, sizeof(float2) * (size_t)rxlen);
        if (!b) return XC_ERR_NOMEM;
        XC_CUDA_TRY(cudaMemcpyAsync(b, rx, sizeof(float2) * (size_t)rxlen, cudaMemcpyHostToDevice, st));
        drx = b;
        if (qf2 || peak) { dq = (float*)workspace(SLOT_OUT0, sizeof(float) * (size_t)K); if (!dq) return XC_ERR_NOMEM; }
        if (fidx || peak) { df = (uint32_t*)workspace(SLOT_OUT1, sizeof(uint32_t) * (size_t)K); if (!df) return XC_ERR_NOMEM; }
        if (caf) { dcaf = (float*)workspace(SLOT_OUT2, cafb); if (!dcaf) return XC_ERR_NOMEM; }
    } else if (peak && (!dq || !df)) {
        if (!dq) dq = (float*)workspace(SLOT_OUT0, sizeof(float) * (size_t)K);
        if (!df) df = (uint32_t*)workspace(SLOT_OUT1, sizeof(uint32_t) * (size_t)K);
        if (!dq || !df) return XC_ERR_NOMEM;
    }
    // window energies: per distinct group length a sliding sum of |rx|^2, gathered at the group starts
    float* d_energy = (float*)workspace(SLOT_TMP1, sizeof(float) * (size_t)K);
    if (!d_energy) return XC_ERR_NOMEM;
    for (size_t q = 0; q < h->lens.size(); ++q) {
        const int64_t el_len = rxlen - h->lens[q] + 1;
        float* d_el = (float*)workspace(SLOT_TMP2, sizeof(float) * (size_t)el_len);
        if (!d_el) return XC_ERR_NOMEM;
        XC_TRY(sliding_sum_device(drx, true, rxlen, 0, h->lens[q], el_len, d_el, 0, st));
        int64_t blocks = (K + 255) / 256;
        const int64_t cap = (int64_t)sm_count() * 16;
        if (blocks > cap) blocks = cap;
        groupfreq_energy_kernel<<<(unsigned)blocks, 256, 0, st>>>(d_el, el_len, h->d_starts[q], h->counts[q], shift0, K,
                                                                  d_energy, q > 0 ? 1 : 0);
        XC_LAUNCH_CHECK();
    }
    XC_TRY(mt_tc_run(h->tc, drx, rxlen, shift0 + h->start0, K, d_energy, dq, reinterpret_cast<int32_t*>(df), 0, st, dcaf, dcplx));
    if (where == XC_HOST) {
        if (qf2) XC_CUDA_TRY(cudaMemcpyAsync(qf2, dq, sizeof(float) * (size_t)K, cudaMemcpyDeviceToHost, st));
        if (fidx) XC_CUDA_TRY(cudaMemcpyAsync(fidx, df, sizeof(uint32_t) * (size_t)K, cudaMemcpyDeviceToHost, st));
        if (caf) XC_CUDA_TRY(cudaMemcpyAsync(caf, dcaf, cafb, cudaMemcpyDeviceToHost, st));
    }
    if (peak) XC_TRY(collect_peak(dq, df, nullptr, K, shift0, 1, nullptr, peak, st));
    if (where == XC_HOST) XC_CUDA_TRY(cudaStreamSynchronize(st));
    return XC_OK;
}

}  // extern "C"
