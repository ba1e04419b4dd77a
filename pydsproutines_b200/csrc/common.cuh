// common.cuh -- shared helpers of libxcorr_b200: status/error plumbing, complex arithmetic,
// the cached device workspace used when the caller passes host pointers.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <mutex>

#include "../../include/xcorr_b200.h"

namespace xc {

// ---------------------------------------------------------------------------------------------
// error plumbing
// ---------------------------------------------------------------------------------------------
void set_error(const char* fmt, ...);
int32_t fail(int32_t code, const char* fmt, ...);
void count_launch(int n = 1);
// optional CUDA-event timing of the dominant (fused) kernel, read back by xc_profile_last_ms()
void profile_begin(cudaStream_t st);
void profile_end(cudaStream_t st);

#define XC_CUDA_TRY(expr)                                                                        \
    do {                                                                                         \
        cudaError_t _e = (expr);                                                                 \
        if (_e != cudaSuccess) {                                                                 \
            return xc::fail(_e == cudaErrorMemoryAllocation ? XC_ERR_NOMEM : XC_ERR_CUDA,        \
                            "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__,    \
                            __LINE__);                                                           \
        }                                                                                        \
    } while (0)

#define XC_TRY(expr)                  \
    do {                              \
        int32_t _s = (expr);          \
        if (_s != XC_OK) return _s;   \
    } while (0)

#define XC_LAUNCH_CHECK()                                                                        \
    do {                                                                                         \
        cudaError_t _e = cudaGetLastError();                                                     \
        if (_e != cudaSuccess)                                                                   \
            return xc::fail(XC_ERR_CUDA, "kernel launch failed: %s (%s:%d)",                     \
                            cudaGetErrorString(_e), __FILE__, __LINE__);                         \
        xc::count_launch();                                                                      \
    } while (0)

// Every extern "C" entry takes this lock: the workspace slots, plan caches and profiling events are shared
// process state, and ctypes releases the GIL around a call.  Recursive because entries call one another.
std::recursive_mutex& api_mutex();
#define XC_API_GUARD std::lock_guard<std::recursive_mutex> xc_api_guard_(xc::api_mutex())

// ---------------------------------------------------------------------------------------------
// device workspace: grow-only named slots PER DEVICE (indexed by the current device); calls on one
// device are serialised by the API lock and by stream order, see xcorr_b200.h "Threading"
// ---------------------------------------------------------------------------------------------
enum Slot {
    SLOT_IN0 = 0,  // staged inputs (host mode)
    SLOT_IN1,
    SLOT_IN2,
    SLOT_OUT0,     // staged outputs (host mode)
    SLOT_OUT1,
    SLOT_OUT2,
    SLOT_TMP0,     // kernel temporaries (energies, partial peaks, ...)
    SLOT_TMP1,
    SLOT_TMP2,
    SLOT_TMP3,
    SLOT_BIG,      // work arrays of the four-step frequency search (bigfft.cu)
    SLOT_SCAN,     // chunk sums of the energy kernels for very long windows (energy.cu)
    SLOT_COUNT
};
// returns a device pointer of at least `bytes` bytes (256-byte aligned), NULL + error on failure
void* workspace(Slot slot, size_t bytes);
int32_t release_workspace();
int sm_count();
int max_smem_optin();
int current_device();
// handles own device memory: using one on another device than the one it was created on is an error, not a fault
int32_t check_handle_device(int handle_device, const char* what);

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is per device: one flag per (kernel instantiation, device).
constexpr int XC_MAX_DEVICES = 64;
struct KernelOnce {
    bool done[XC_MAX_DEVICES] = {};
    int occ[XC_MAX_DEVICES] = {};
    // opts the kernel into `smem_bytes` of dynamic shared memory on the current device (first call per device)
    template <class Kern>
    int32_t configure(Kern kern, int smem_bytes, int* first = nullptr) {
        int dev = 0;
        XC_CUDA_TRY(cudaGetDevice(&dev));
        if (dev < 0 || dev >= XC_MAX_DEVICES) return fail(XC_ERR_UNSUPPORTED, "device ordinal %d out of range", dev);
        if (first) *first = done[dev] ? 0 : 1;
        if (!done[dev]) {
            XC_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
            done[dev] = true;
        }
        return XC_OK;
    }
    int& occupancy() {
        int dev = 0;
        cudaGetDevice(&dev);
        return occ[(dev >= 0 && dev < XC_MAX_DEVICES) ? dev : 0];
    }
};

// ---------------------------------------------------------------------------------------------
// complex helpers (float2 = interleaved complex64)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    return make_float2(fmaf(a.x, b.x, -a.y * b.y), fmaf(a.x, b.y, a.y * b.x));
}
// a * conj(b)
__device__ __forceinline__ float2 cmulc(float2 a, float2 b) {
    return make_float2(fmaf(a.x, b.x, a.y * b.y), fmaf(a.y, b.x, -a.x * b.y));
}
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }

// ---------------------------------------------------------------------------------------------
// packed FP32 (sm_100a FFMA2 / FADD2 / FMUL2): one instruction works on both halves of an aligned
// register pair, i.e. on one interleaved complex value.  Measured on B200 (tools/ubench/f32x2.cu):
// 2 warp-instructions/clk/SM = the full 128 FP32 lanes at half the issue slots; the half swap,
// per-half negation and scalar broadcast of an operand are free operand modifiers in SASS, so
// ptxas folds the make_float2() shuffles below into the arithmetic instruction.
// ---------------------------------------------------------------------------------------------
typedef unsigned long long u64;
__device__ __forceinline__ u64 f2u(float2 a) {
    u64 r;
    asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a.x), "f"(a.y));
    return r;
}
__device__ __forceinline__ float2 u2f(u64 r) {
    float2 v;
    asm("mov.b64 {%0,%1}, %2;" : "=f"(v.x), "=f"(v.y) : "l"(r));
    return v;
}
__device__ __forceinline__ float2 padd(float2 a, float2 b) {
    u64 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(f2u(a)), "l"(f2u(b)));
    return u2f(d);
}
__device__ __forceinline__ float2 psub(float2 a, float2 b) {
    u64 d;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(f2u(a)), "l"(f2u(b)));
    return u2f(d);
}
__device__ __forceinline__ float2 pmul(float2 a, float2 b) {
    u64 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(f2u(a)), "l"(f2u(b)));
    return u2f(d);
}
__device__ __forceinline__ float2 pfma(float2 a, float2 b, float2 c) {
    u64 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(f2u(a)), "l"(f2u(b)), "l"(f2u(c)));
    return u2f(d);
}
// a * b in two packed instructions
__device__ __forceinline__ float2 pcmul(float2 a, float2 b) {
    return pfma(make_float2(a.y, a.x), make_float2(-b.y, b.y), pmul(a, make_float2(b.x, b.x)));
}
// a * conj(b)
__device__ __forceinline__ float2 pcmulc(float2 a, float2 b) {
    return pfma(make_float2(a.y, a.x), make_float2(b.y, -b.y), pmul(a, make_float2(b.x, b.x)));
}
__device__ __forceinline__ float cnorm2(float2 a) { return fmaf(a.x, a.x, a.y * a.y); }

// per-call peak reduction record kept on the device: packs (value, index) so that a single
// 64-bit atomicMax picks the largest value and, on ties, the LOWEST index.
__device__ __forceinline__ unsigned long long pack_peak(float v, uint32_t inv_index) {
    // v >= 0 here (a QF^2), so its bit pattern is monotone in v
    return ((unsigned long long)__float_as_uint(v) << 32) | (unsigned long long)inv_index;
}

}  // namespace xc
