// multitemplate.cu -- (multi-)template sliding correlator: for every shift s and template t
//     d_t[s] = sum_n tmpl_t[n] * x[s+n],   QF^2 = |d_t[s]|^2 / E_t / E_x[s],   best over t.
// Replaces multiTemplateSlidingDotProduct (multiplySlices.cu:251-399) and the FFT/IFFT formulation
// of TemplateCrossCorrelator.correlate (xcorrRoutines.py:312-371).
//
// FP32 SIMT formulation: a CTA owns 1024 consecutive shifts; the x tile (1024 + L samples) stays in
// shared memory for all templates, each template is streamed through a double-buffered shared
// copy, every thread keeps 8 consecutive shifts in registers and slides a 16-sample window, so the
// inner loop is 256 FFMA per 8 LDS.128.
#include <vector>

#include "internal.cuh"
#include "mt_ovs.cuh"

struct xc_mt {
    int32_t T, L, L8;
    float2* d_templates;   // [T, L8] zero padded rows, conjugated if requested at create time
    float* d_energies;     // [T]
    xc::MtTcPlan tc;       // tensor-core plan (usable when the templates are exact in BF16)
    int32_t device;        // the device that owns the arrays of this handle
    float2* d_H;           // [T][8192] spectra of the templates for the fast-correlation path (mt_ovs.cuh), made at first use
    int32_t fp32_only;     // XC_MT_FP32 at create time: plain FP32 products wanted, no fast correlation either
};

namespace xc {

constexpr int MT_OVS_MIN_L = 512;      // templates from this length on take the fast-correlation path over long records
constexpr int MT_THREADS = 128;
constexpr int MT_R = 8;
constexpr int MT_TS = MT_THREADS * MT_R;

__device__ __forceinline__ int mt_phys_unit(int u) { return u + (u >> 2); }

__host__ __device__ inline size_t mt_smem_bytes(int L8) {
    const size_t units = (MT_TS + L8) / 2;
    return sizeof(float4) * (units + units / 4 + 4) + 2 * sizeof(float2) * (size_t)L8;
}

__global__ void __launch_bounds__(MT_THREADS) mt_kernel(const float2* __restrict__ templates, int T, int L8,
                                                        const float* __restrict__ tenergy, const float2* __restrict__ x,
                                                        int64_t xlen, int64_t start, int64_t K,
                                                        const float* __restrict__ xenergy, float* __restrict__ best_out,
                                                        int32_t* __restrict__ tidx_out, float2* __restrict__ full,
                                                        int sqrt_out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int units = (MT_TS + L8) / 2;
    float4* s_x = reinterpret_cast<float4*>(smem_raw);
    float2* s_x2 = reinterpret_cast<float2*>(smem_raw);
    float4* s_t = s_x + (units + units / 4 + 4);      // two template buffers of L8 samples
    const int tid = threadIdx.x;
    const int64_t i0 = (int64_t)blockIdx.x * MT_TS;
    const int64_t base = start + i0;

    for (int j = tid; j < MT_TS + L8; j += MT_THREADS) {
        const int64_t g = base + j;
        float2 val = make_float2(0.f, 0.f);
        if (g >= 0 && g < xlen) val = __ldg(x + g);
        s_x2[2 * mt_phys_unit(j >> 1) + (j & 1)] = val;
    }
    {   // template 0 into buffer 0
        const float4* src = reinterpret_cast<const float4*>(templates);
        for (int j = tid; j < L8 / 2; j += MT_THREADS) s_t[j] = src[j];
    }
    float e_x[MT_R];
#pragma unroll
    for (int j = 0; j < MT_R; ++j) {
        const int64_t i = i0 + tid * MT_R + j;
        e_x[j] = i < K ? xenergy[i] : 1.f;
    }
    float best[MT_R];
    int bidx[MT_R];
#pragma unroll
    for (int j = 0; j < MT_R; ++j) { best[j] = 0.f; bidx[j] = 0; }
    __syncthreads();

    for (int t = 0; t < T; ++t) {
        const float4* cur = s_t + (size_t)(t & 1) * (L8 / 2);
        if (t + 1 < T) {   // prefetch the next template into the other buffer
            float4* nxt = s_t + (size_t)((t + 1) & 1) * (L8 / 2);
            const float4* src = reinterpret_cast<const float4*>(templates + (size_t)(t + 1) * L8);
            for (int j = tid; j < L8 / 2; j += MT_THREADS) nxt[j] = __ldg(src + j);
        }
        float2 acc[MT_R];
#pragma unroll
        for (int j = 0; j < MT_R; ++j) acc[j] = make_float2(0.f, 0.f);
        float2 w[16];
#pragma unroll
        for (int vv = 0; vv < 4; ++vv) {
            const float4 f = s_x[mt_phys_unit(tid * 4 + vv)];
            w[2 * vv] = make_float2(f.x, f.y);
            w[2 * vv + 1] = make_float2(f.z, f.w);
        }
#pragma unroll 1
        for (int m = 0; m < L8 / 8; ++m) {
            const int u1 = tid * 4 + 4 * m + 4;
#pragma unroll
            for (int vv = 0; vv < 4; ++vv) {
                const float4 f = s_x[mt_phys_unit(u1 + vv)];
                w[8 + 2 * vv] = make_float2(f.x, f.y);
                w[8 + 2 * vv + 1] = make_float2(f.z, f.w);
            }
            float2 c[8];
#pragma unroll
            for (int vv = 0; vv < 4; ++vv) {
                const float4 f = cur[4 * m + vv];
                c[2 * vv] = make_float2(f.x, f.y);
                c[2 * vv + 1] = make_float2(f.z, f.w);
            }
#pragma unroll
            for (int u = 0; u < 8; ++u)
#pragma unroll
                for (int j = 0; j < MT_R; ++j) {
                    const float2 xv = w[u + j];
                    acc[j].x = fmaf(c[u].x, xv.x, fmaf(-c[u].y, xv.y, acc[j].x));
                    acc[j].y = fmaf(c[u].x, xv.y, fmaf(c[u].y, xv.x, acc[j].y));
                }
#pragma unroll
            for (int j = 0; j < 8; ++j) w[j] = w[8 + j];
        }
        const float et = tenergy[t];
#pragma unroll
        for (int j = 0; j < MT_R; ++j) {
            const float q = cnorm2(acc[j]) / et / e_x[j];
            if (q > best[j]) { best[j] = q; bidx[j] = t; }      // strict '>' keeps the lowest template on ties
            if (full) {
                const int64_t i = i0 + tid * MT_R + j;
                if (i < K) {
                    const float sc = 1.f / (sqrtf(et) * sqrtf(e_x[j]));
                    full[(size_t)t * K + i] = make_float2(acc[j].x * sc, acc[j].y * sc);
                }
            }
        }
        __syncthreads();   // next template visible; current buffer free for the prefetch after next
    }
#pragma unroll
    for (int j = 0; j < MT_R; ++j) {
        const int64_t i = i0 + tid * MT_R + j;
        if (i < K) {
            if (best_out) best_out[i] = sqrt_out ? sqrtf(best[j]) : best[j];
            if (tidx_out) tidx_out[i] = bidx[j];
        }
    }
}

int32_t collect_peak(const float* d_qf2, const uint32_t* d_fidx, const int32_t* d_aux, int64_t K, int64_t shift0,
                     int64_t step, const int64_t* d_shifts, xc_peak* peak, cudaStream_t st);

}  // namespace xc

using namespace xc;

extern "C" {

int32_t xc_mt_create(const void* templates, int32_t T, int32_t L, int32_t conj, const float* energies, xc_mt** handle) {
    XC_API_GUARD;
    if (!templates || !handle) return fail(XC_ERR_ARG, "templates/handle is NULL");
    if (T < 1 || L < 1) return fail(XC_ERR_ARG, "need at least one template of at least one sample");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(XC_ERR_CUDA, "no CUDA device available; libxcorr_b200 has no CPU fallback");
    }
    const int L8 = (L + 7) / 8 * 8;
    if ((int)mt_smem_bytes(L8) > max_smem_optin())
        return fail(XC_ERR_UNSUPPORTED, "template length %d needs %zu B of shared memory (limit %d)", L,
                    mt_smem_bytes(L8), max_smem_optin());
    const float2* in = reinterpret_cast<const float2*>(templates);
    std::vector<float2> padded((size_t)T * L8, make_float2(0.f, 0.f));
    std::vector<float> en(T);
    for (int t = 0; t < T; ++t) {
        double e = 0.0;
        for (int n = 0; n < L; ++n) {
            const float2 v = in[(size_t)t * L + n];
            padded[(size_t)t * L8 + n] = (conj & XC_MT_CONJ) ? make_float2(v.x, -v.y) : v;
            e += (double)v.x * v.x + (double)v.y * v.y;
        }
        en[t] = energies ? energies[t] : (float)e;
    }
    xc_mt* h = new xc_mt();
    h->T = T; h->L = L; h->L8 = L8;
    h->device = current_device();
    h->d_templates = nullptr; h->d_energies = nullptr;
    h->d_H = nullptr;
    h->fp32_only = (conj & (XC_MT_FP32 | XC_MT_DIRECT)) ? 1 : 0;
    cudaError_t e = cudaMalloc(&h->d_templates, sizeof(float2) * padded.size());
    if (e == cudaSuccess) e = cudaMalloc(&h->d_energies, sizeof(float) * T);
    if (e == cudaSuccess) e = cudaMemcpy(h->d_templates, padded.data(), sizeof(float2) * padded.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(h->d_energies, en.data(), sizeof(float) * T, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cudaGetLastError();
        xc_mt_destroy(h);
        return fail(e == cudaErrorMemoryAllocation ? XC_ERR_NOMEM : XC_ERR_CUDA, "template upload failed: %s", cudaGetErrorString(e));
    }
    if (!(conj & XC_MT_FP32)) {          // XC_MT_FP32: keep the FP32 SIMT kernel (the tensor-core plan stays unusable)
        const int32_t st = mt_tc_prepare(h->d_templates, T, L, L8, h->d_energies, &h->tc);
        if (st != XC_OK) {
            xc_mt_destroy(h);
            return st;
        }
    }
    *handle = h;
    return XC_OK;
}

int32_t xc_mt_destroy(xc_mt* h) {
    XC_API_GUARD;
    if (!h) return XC_OK;
    mt_tc_release(&h->tc);
    cudaFree(h->d_templates);
    cudaFree(h->d_energies);
    cudaFree(h->d_H);
    delete h;
    return XC_OK;
}

int32_t xc_mt_uses_tensor_cores(const xc_mt* h) { return (h && h->tc.usable) ? 1 : 0; }

int32_t xc_mt_correlate(xc_mt* h, const void* x, int64_t xlen, int64_t start, int64_t K, float* best, int32_t* tidx,
                        void* full, int32_t sqrt_out, xc_peak* peak, int32_t where, void* stream) {
    XC_API_GUARD;
    if (where != XC_HOST && where != XC_DEVICE) return fail(XC_ERR_ARG, "where must be XC_HOST (0) or XC_DEVICE (1)");
    if (!h || !x) return fail(XC_ERR_ARG, "handle/x is NULL");
    XC_TRY(check_handle_device(h->device, "template correlator"));
    if (start < 0) return fail(XC_ERR_RANGE, "startIdx should be >= 0.");                                   // cupyExtensions.py:584-586
    if (K < 0 || start + K - 1 + h->L - 1 >= xlen)
        return fail(XC_ERR_RANGE, "final slide index should be within the bounds of d_x.");                   // cupyExtensions.py:587-591
    if (K == 0) {
        if (peak) { memset(peak, 0, sizeof(*peak)); peak->index = -1; }
        return XC_OK;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const float2* dx = reinterpret_cast<const float2*>(x);
    float* db = best;
    int32_t* dt = tidx;
    float2* dfull = reinterpret_cast<float2*>(full);
    const size_t fullb = sizeof(float2) * (size_t)h->T * (size_t)K;
    if (where == XC_HOST) {
        float2* b = (float2*)workspace(SLOT_IN1, sizeof(float2) * (size_t)xlen);
        if (!b) return XC_ERR_NOMEM;
        XC_CUDA_TRY(cudaMemcpyAsync(b, x, sizeof(float2) * (size_t)xlen, cudaMemcpyHostToDevice, st));
        dx = b;
        if (best || peak) { db = (float*)workspace(SLOT_OUT0, sizeof(float) * (size_t)K); if (!db) return XC_ERR_NOMEM; }
        if (tidx || peak) { dt = (int32_t*)workspace(SLOT_OUT1, sizeof(int32_t) * (size_t)K); if (!dt) return XC_ERR_NOMEM; }
        if (full) { dfull = (float2*)workspace(SLOT_OUT2, fullb); if (!dfull) return XC_ERR_NOMEM; }
    } else if (peak) {
        if (!db) db = (float*)workspace(SLOT_OUT0, sizeof(float) * (size_t)K);
        if (!dt) dt = (int32_t*)workspace(SLOT_OUT1, sizeof(int32_t) * (size_t)K);
        if (!db || !dt) return XC_ERR_NOMEM;
    }
    float* d_energy = (float*)workspace(SLOT_TMP1, sizeof(float) * (size_t)K);
    if (!d_energy) return XC_ERR_NOMEM;
    XC_TRY(sliding_sum_device(dx + start, true, xlen - start, 0, h->L, K, d_energy, 0, st));
    using OPL = PlanFor<13>::PL;                       // 8192-point blocks, one group of 256 threads per CTA
    constexpr int64_t OM = OPL::N;
    if (!dfull && !h->fp32_only && h->L >= MT_OVS_MIN_L && h->L <= OM / 2 && K >= 4 * (OM - h->L + 1)) {
        // (best, template) over a long record: fast correlation through 8192-point blocks (mt_ovs.cuh)
        constexpr size_t osmem = OPL::TABLE_BYTES + sizeof(float2) * (size_t)(OPL::BUF + OM + OM / 8);    // + the reorder buffer
        if (!h->d_H) {
            XC_CUDA_TRY(cudaMalloc(&h->d_H, sizeof(float2) * (size_t)h->T * OM));
            auto tk = mt_ovs_table_kernel<OPL>;
            static KernelOnce tonce;
            XC_TRY(tonce.configure(tk, max_smem_optin()));
            tk<<<(unsigned)h->T, OPL::TG, osmem, st>>>(h->d_templates, h->L, h->L8, h->d_H);
            XC_LAUNCH_CHECK();
            XC_CUDA_TRY(cudaStreamSynchronize(st));      // the handle may be used from another stream next time
        }
        const int S = (int)(OM - h->L + 1);
        const int64_t nblocks = (K + S - 1) / S;
        MtOvsArgs oa{dx, xlen, start, K, S, h->T, h->d_H, h->d_energies, d_energy, db, dt, sqrt_out ? 1 : 0,
                     (float)(1.0 / ((double)OM * (double)OM))};
#ifndef XC_MT_OVS1      // the shipped variant: two groups per SM, X and the running maxima in tensor memory (15.4 -> 12.9 ms at C4)
        auto kern = mt_ovs2_kernel<OPL, 2>;
        constexpr size_t ksmem = OPL::TABLE_BYTES + sizeof(float2) * (size_t)OPL::BUF * 2 + 16;
        constexpr int kthreads = OPL::TG * 2;
        int64_t ctas = (nblocks + 1) / 2 < (int64_t)sm_count() ? (nblocks + 1) / 2 : (int64_t)sm_count();
#else
        auto kern = mt_ovs_kernel<OPL>;
        constexpr size_t ksmem = osmem;
        constexpr int kthreads = OPL::TG;
        int64_t ctas = nblocks < (int64_t)sm_count() ? nblocks : (int64_t)sm_count();
#endif
        static KernelOnce once;
        XC_TRY(once.configure(kern, max_smem_optin()));
        profile_begin(st);
        kern<<<(unsigned)ctas, kthreads, ksmem, st>>>(oa, nblocks);
        profile_end(st);
        XC_LAUNCH_CHECK();
    } else if (h->tc.usable) {
        // dense complex contraction on the tensor cores (tcgen05); the full (T, K) complex QF comes out of the same
        // kernel's complex epilogue (flag 2), normalised and template-major
        XC_TRY(mt_tc_run(h->tc, dx, xlen, start, K, d_energy, db, dt, (sqrt_out ? 1 : 0) | (dfull ? 2 : 0), st, nullptr, dfull));
    } else {
        const size_t smem = mt_smem_bytes(h->L8);
        if ((int)smem > max_smem_optin()) return fail(XC_ERR_UNSUPPORTED, "template kernel needs %zu B of shared memory", smem);
        static KernelOnce once;
        XC_TRY(once.configure(mt_kernel, max_smem_optin()));
        const int64_t blocks = (K + MT_TS - 1) / MT_TS;
        mt_kernel<<<(unsigned)blocks, MT_THREADS, smem, st>>>(h->d_templates, h->T, h->L8, h->d_energies, dx, xlen, start,
                                                              K, d_energy, db, dt, dfull, sqrt_out);
        XC_LAUNCH_CHECK();
    }
    if (where == XC_HOST) {
        if (best) XC_CUDA_TRY(cudaMemcpyAsync(best, db, sizeof(float) * (size_t)K, cudaMemcpyDeviceToHost, st));
        if (tidx) XC_CUDA_TRY(cudaMemcpyAsync(tidx, dt, sizeof(int32_t) * (size_t)K, cudaMemcpyDeviceToHost, st));
        if (full) XC_CUDA_TRY(cudaMemcpyAsync(full, dfull, fullb, cudaMemcpyDeviceToHost, st));
    }
    if (peak) {
        XC_TRY(collect_peak(db, nullptr, dt, K, start, 1, nullptr, peak, st));
    }
    if (where == XC_HOST) XC_CUDA_TRY(cudaStreamSynchronize(st));
    return XC_OK;
}

}  // extern "C"
