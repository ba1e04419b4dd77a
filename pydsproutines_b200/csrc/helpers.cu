// helpers.cu -- the gather / copy helpers of the reference's GPU path.  The fused kernels
// make them unnecessary on the hot path (the windows and products never exist in HBM); they are kept as thin,
// bandwidth-bound compatibility functions so that code written against cupyExtensions keeps working.
//   cupyCopyGroups32fc                         cupyExtensions.py:41-83   -> copying.cu (copy_groups_kernel32fc)
//   cupyCopyIncrementalEqualSlicesToMatrix_32fc cupyExtensions.py:161-215 -> copying.cu:83-137
//   multiplySlidesNormalised                   cupyExtensions.py:491-560 -> multiplySlices.cu:112-216
#include "internal.cuh"

namespace xc {

// one CTA per group: y[yStarts[g] + i] = x[xStarts[g] + i], i < lengths[g]
__global__ void copy_groups_kernel(const float2* __restrict__ x, float2* __restrict__ y, const int32_t* __restrict__ xs,
                                   const int32_t* __restrict__ ys, const int32_t* __restrict__ len, int64_t xlen, int64_t ylen) {
    const int g = blockIdx.x;
    const int64_t a = xs[g], b = ys[g];
    const int n = len[g];
    if (n <= 0 || a < 0 || b < 0 || a + n > xlen || b + n > ylen) return;   // a group that leaves either buffer is skipped
    const float2* src = x + a;
    float2* dst = y + b;
    for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
}

// out[r, c] = x[start + r * inc + c] (0 outside x)
__global__ void slices_to_matrix_kernel(const float2* __restrict__ x, int64_t xlen, int64_t start, int64_t inc,
                                        int64_t rows, int64_t cols, float2* __restrict__ out) {
    const int64_t total = rows * cols;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
        const int64_t r = e / cols, c = e - r * cols;
        const int64_t i = start + r * inc + c;
        out[e] = (i >= 0 && i < xlen) ? __ldg(x + i) : make_float2(0.f, 0.f);
    }
}

// one warp per slide r: E = sum_t |y[start + r + t]|^2 in double (multiplySlices.cu:196-204), then
// z[r, t] = x[t] * y[start + r + t] / (float)(sqrt(E) * coefficient)
__global__ void multiply_slides_kernel(const float2* __restrict__ x, int xlen, const float2* __restrict__ y, int64_t ylen,
                                       int64_t start, int64_t idxlen, double coefficient, float2* __restrict__ z) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = warp; r < idxlen; r += nwarps) {
        const int64_t base = start + r;
        double e = 0.0;
        for (int t = lane; t < xlen; t += 32) {
            const int64_t i = base + t;
            if (i < ylen) {
                const float2 v = __ldg(y + i);
                e += (double)v.x * v.x + (double)v.y * v.y;
            }
        }
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) e += __shfl_xor_sync(0xffffffffu, e, o);
        const float norm = (float)(sqrt(e) * coefficient);
        for (int t = lane; t < xlen; t += 32) {
            const int64_t i = base + t;
            const float2 v = (i < ylen) ? __ldg(y + i) : make_float2(0.f, 0.f);
            const float2 p = cmul(__ldg(x + t), v);
            z[r * xlen + t] = make_float2(p.x / norm, p.y / norm);
        }
    }
}

// raw interleaved int16 I/Q -> complex64, 4 samples (16 B in, 32 B out) per thread and iteration
__global__ void convert_i16_c64_kernel(const int16_t* __restrict__ iq, int64_t nsamp, float2* __restrict__ out) {
    const int64_t nvec = nsamp / 4;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int4* in4 = reinterpret_cast<const int4*>(iq);
    float4* out4 = reinterpret_cast<float4*>(out);
    for (int64_t i = tid; i < nvec; i += stride) {
        const int4 v = __ldg(in4 + i);
        const int w[4] = {v.x, v.y, v.z, v.w};
        float f[8];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            f[2 * q] = (float)(int16_t)(w[q] & 0xffff);
            f[2 * q + 1] = (float)(int16_t)(w[q] >> 16);
        }
        out4[2 * i] = make_float4(f[0], f[1], f[2], f[3]);
        out4[2 * i + 1] = make_float4(f[4], f[5], f[6], f[7]);
    }
    for (int64_t n = nvec * 4 + tid; n < nsamp; n += stride) out[n] = make_float2((float)iq[2 * n], (float)iq[2 * n + 1]);
}

}  // namespace xc

using namespace xc;

static int32_t helpers_require_device() {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail(XC_ERR_CUDA, "no CUDA device available; libxcorr_b200 has no CPU fallback");
    }
    return XC_OK;
}

extern "C" {

int32_t xc_convert_i16_c64(const int16_t* iq, int64_t nsamp, void* out, int32_t where_in, int32_t where_out, void* stream) {
    XC_API_GUARD;
    if ((where_in != XC_HOST && where_in != XC_DEVICE) || (where_out != XC_HOST && where_out != XC_DEVICE))
        return fail(XC_ERR_ARG, "where must be XC_HOST (0) or XC_DEVICE (1)");
    if (!iq || !out) return fail(XC_ERR_ARG, "iq/out is NULL");
    if (nsamp < 0) return fail(XC_ERR_ARG, "negative size");
    XC_TRY(helpers_require_device());
    if (nsamp == 0) return XC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const int16_t* din = iq;
    float2* dout = (float2*)out;
    if (where_in == XC_HOST) {          // the raw samples cross PCIe: half the bytes of complex64
        int16_t* b = (int16_t*)workspace(SLOT_IN0, sizeof(int16_t) * 2 * (size_t)nsamp);
        if (!b) return XC_ERR_NOMEM;
        XC_CUDA_TRY(cudaMemcpyAsync(b, iq, sizeof(int16_t) * 2 * (size_t)nsamp, cudaMemcpyHostToDevice, st));
        din = b;
    }
    if (where_out == XC_HOST) {
        dout = (float2*)workspace(SLOT_OUT2, sizeof(float2) * (size_t)nsamp);
        if (!dout) return XC_ERR_NOMEM;
    }
    if ((reinterpret_cast<uintptr_t>(din) & 15) || (reinterpret_cast<uintptr_t>(dout) & 15))
        return fail(XC_ERR_ARG, "device buffers must be 16-byte aligned");
    int64_t blocks = (nsamp / 4 + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    convert_i16_c64_kernel<<<(unsigned)blocks, 256, 0, st>>>(din, nsamp, dout);
    XC_LAUNCH_CHECK();
    if (where_out == XC_HOST) {
        XC_CUDA_TRY(cudaMemcpyAsync(out, dout, sizeof(float2) * (size_t)nsamp, cudaMemcpyDeviceToHost, st));
        XC_CUDA_TRY(cudaStreamSynchronize(st));
    }
    return XC_OK;
}

int32_t xc_copy_groups_c64(const void* x, int64_t xlen, void* y, int64_t ylen, const int32_t* x_starts,
                           const int32_t* y_starts, const int32_t* lengths, int32_t G, int32_t where, void* stream) {
    XC_API_GUARD;
    if (where != XC_HOST && where != XC_DEVICE) return fail(XC_ERR_ARG, "where must be XC_HOST (0) or XC_DEVICE (1)");
    if (!x || !y || !x_starts || !y_starts || !lengths) return fail(XC_ERR_ARG, "NULL argument");
    if (G < 0 || xlen < 0 || ylen < 0) return fail(XC_ERR_ARG, "negative size");
    XC_TRY(helpers_require_device());
    if (G == 0) return XC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const float2* dx = (const float2*)x;
    float2* dy = (float2*)y;
    const int32_t *dxs = x_starts, *dys = y_starts, *dl = lengths;
    if (where == XC_HOST) {
        // the three index arrays are host memory here: a bad start or length would write into neighbouring workspace
        // blocks (XC_DEVICE callers get the same check per group inside the kernel)
        for (int g = 0; g < G; ++g) {
            if (lengths[g] < 0 || x_starts[g] < 0 || y_starts[g] < 0 || (int64_t)x_starts[g] + lengths[g] > xlen ||
                (int64_t)y_starts[g] + lengths[g] > ylen)
                return fail(XC_ERR_RANGE, "group %d: [%d, %d) of x (%lld samples) -> [%d, %d) of y (%lld samples) is out of bounds", g,
                            x_starts[g], x_starts[g] + lengths[g], (long long)xlen, y_starts[g], y_starts[g] + lengths[g], (long long)ylen);
        }
        float2* bx = (float2*)workspace(SLOT_IN0, sizeof(float2) * (size_t)xlen);
        float2* by = (float2*)workspace(SLOT_OUT0, sizeof(float2) * (size_t)ylen);
        int32_t* bi = (int32_t*)workspace(SLOT_IN2, sizeof(int32_t) * 3 * (size_t)G);
        if (!bx || !by || !bi) return XC_ERR_NOMEM;
        XC_CUDA_TRY(cudaMemcpyAsync(bx, x, sizeof(float2) * (size_t)xlen, cudaMemcpyHostToDevice, st));
        XC_CUDA_TRY(cudaMemcpyAsync(by, y, sizeof(float2) * (size_t)ylen, cudaMemcpyHostToDevice, st));
        XC_CUDA_TRY(cudaMemcpyAsync(bi, x_starts, sizeof(int32_t) * (size_t)G, cudaMemcpyHostToDevice, st));
        XC_CUDA_TRY(cudaMemcpyAsync(bi + G, y_starts, sizeof(int32_t) * (size_t)G, cudaMemcpyHostToDevice, st));
        XC_CUDA_TRY(cudaMemcpyAsync(bi + 2 * (size_t)G, lengths, sizeof(int32_t) * (size_t)G, cudaMemcpyHostToDevice, st));
        dx = bx; dy = by; dxs = bi; dys = bi + G; dl = bi + 2 * (size_t)G;
    }
    copy_groups_kernel<<<(unsigned)G, 256, 0, st>>>(dx, dy, dxs, dys, dl, xlen, ylen);
    XC_LAUNCH_CHECK();
    if (where == XC_HOST) {
        XC_CUDA_TRY(cudaMemcpyAsync(y, dy, sizeof(float2) * (size_t)ylen, cudaMemcpyDeviceToHost, st));
        XC_CUDA_TRY(cudaStreamSynchronize(st));
    }
    return XC_OK;
}

int32_t xc_copy_slices_to_matrix_c64(const void* x, int64_t xlen, int64_t start, int64_t increment, int64_t row_length,
                                     int64_t num_rows, void* out, int32_t where, void* stream) {
    XC_API_GUARD;
    if (where != XC_HOST && where != XC_DEVICE) return fail(XC_ERR_ARG, "where must be XC_HOST (0) or XC_DEVICE (1)");
    if (!x || !out) return fail(XC_ERR_ARG, "x/out is NULL");
    if (xlen < 0 || row_length < 0 || num_rows < 0) return fail(XC_ERR_ARG, "negative size");
    XC_TRY(helpers_require_device());
    const int64_t total = row_length * num_rows;
    if (total == 0) return XC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const float2* dx = (const float2*)x;
    float2* dout = (float2*)out;
    if (where == XC_HOST) {
        float2* bx = (float2*)workspace(SLOT_IN0, sizeof(float2) * (size_t)xlen);
        float2* bo = (float2*)workspace(SLOT_OUT2, sizeof(float2) * (size_t)total);
        if (!bx || !bo) return XC_ERR_NOMEM;
        XC_CUDA_TRY(cudaMemcpyAsync(bx, x, sizeof(float2) * (size_t)xlen, cudaMemcpyHostToDevice, st));
        dx = bx; dout = bo;
    }
    int64_t blocks = (total + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    slices_to_matrix_kernel<<<(unsigned)blocks, 256, 0, st>>>(dx, xlen, start, increment, num_rows, row_length, dout);
    XC_LAUNCH_CHECK();
    if (where == XC_HOST) {
        XC_CUDA_TRY(cudaMemcpyAsync(out, dout, sizeof(float2) * (size_t)total, cudaMemcpyDeviceToHost, st));
        XC_CUDA_TRY(cudaStreamSynchronize(st));
    }
    return XC_OK;
}

int32_t xc_multiply_slides_normalised_c64(const void* x, int32_t xlen, const void* y, int64_t ylen, int64_t start,
                                          int64_t idxlen, double coefficient, void* out, int32_t where, void* stream) {
    XC_API_GUARD;
    if (where != XC_HOST && where != XC_DEVICE) return fail(XC_ERR_ARG, "where must be XC_HOST (0) or XC_DEVICE (1)");
    if (!x || !y || !out) return fail(XC_ERR_ARG, "x/y/out is NULL");
    if (xlen < 1) return fail(XC_ERR_ARG, "the template needs at least one sample");
    if (start < 0 || idxlen < 0 || start + idxlen > ylen)
        return fail(XC_ERR_RANGE, "startIdx and idxlen should be within the bounds of d_y.");     // cupyExtensions.py:512-515
    XC_TRY(helpers_require_device());
    if (idxlen == 0) return XC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const float2* dx = (const float2*)x;
    const float2* dy = (const float2*)y;
    float2* dout = (float2*)out;
    const size_t outb = sizeof(float2) * (size_t)idxlen * (size_t)xlen;
    if (where == XC_HOST) {
        float2* bx = (float2*)workspace(SLOT_IN0, sizeof(float2) * (size_t)xlen);
        float2* by = (float2*)workspace(SLOT_IN1, sizeof(float2) * (size_t)ylen);
        float2* bo = (float2*)workspace(SLOT_OUT2, outb);
        if (!bx || !by || !bo) return XC_ERR_NOMEM;
        XC_CUDA_TRY(cudaMemcpyAsync(bx, x, sizeof(float2) * (size_t)xlen, cudaMemcpyHostToDevice, st));
        XC_CUDA_TRY(cudaMemcpyAsync(by, y, sizeof(float2) * (size_t)ylen, cudaMemcpyHostToDevice, st));
        dx = bx; dy = by; dout = bo;
    }
    if (!(coefficient > 0.0)) {          // default: the norm of the template (cupyExtensions.py:517-521)
        float* d_n = (float*)workspace(SLOT_TMP0, 256);
        if (!d_n) return XC_ERR_NOMEM;
        XC_TRY(normsq_device(dx, xlen, 1, xlen, d_n, st));
        float nsq = 0.f;
        XC_CUDA_TRY(cudaMemcpyAsync(&nsq, d_n, sizeof(float), cudaMemcpyDeviceToHost, st));
        XC_CUDA_TRY(cudaStreamSynchronize(st));
        coefficient = sqrt((double)nsq);
    }
    int64_t blocks = (idxlen * 32 + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    multiply_slides_kernel<<<(unsigned)blocks, 256, 0, st>>>(dx, xlen, dy, ylen, start, idxlen, coefficient, dout);
    XC_LAUNCH_CHECK();
    if (where == XC_HOST) {
        XC_CUDA_TRY(cudaMemcpyAsync(out, dout, outb, cudaMemcpyDeviceToHost, st));
        XC_CUDA_TRY(cudaStreamSynchronize(st));
    }
    return XC_OK;
}

}  // extern "C"
