// Micro-benchmark: issue/throughput of packed FP32 (FFMA2/FADD2/FMUL2) vs scalar FFMA on sm_100a,
// alone and mixed with shared-memory loads.  Prints lane-FMA-ops per clock per SM for each variant.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/ubench_f32x2 tools/ubench/f32x2.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
#define DI __device__ __forceinline__
DI u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
DI float2 upk(u64 r) { float2 v; asm("mov.b64 {%0,%1}, %2;" : "=f"(v.x), "=f"(v.y) : "l"(r)); return v; }
DI u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
DI u64 mul2(u64 a, u64 b) { u64 d; asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
DI u64 add2(u64 a, u64 b) { u64 d; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
DI float fma1(float a, float b, float c) { float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }
DI float add1(float a, float b) { float d; asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b)); return d; }

constexpr int NACC = 12;
template <int MODE>
__global__ void __launch_bounds__(512) kern(float* out, int iters, float fa, float fb) {
    __shared__ float2 sm[1024];
    sm[threadIdx.x] = make_float2(fa, fb); sm[threadIdx.x + 512] = make_float2(fb, fa);
    __syncthreads();
    float2 x[NACC];
#pragma unroll
    for (int i = 0; i < NACC; ++i) x[i] = make_float2(threadIdx.x * 1e-3f + i, i * 0.5f);
    const float2 w = make_float2(fa, fb);
    const unsigned sa = (unsigned)__cvta_generic_to_shared(sm + threadIdx.x);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r) {
#pragma unroll
            for (int i = 0; i < NACC; ++i) {
                if (MODE == 0) {          // scalar FFMA, 3 register operands: 2 instr per complex
                    x[i].x = fma1(x[i].x, w.x, w.y); x[i].y = fma1(x[i].y, w.y, w.x);
                } else if (MODE == 1) {   // FFMA2 plain
                    x[i] = upk(fma2(pk(x[i].x, x[i].y), pk(w.x, w.y), pk(w.y, w.x)));
                } else if (MODE == 2) {   // FADD2
                    x[i] = upk(add2(pk(x[i].x, x[i].y), pk(w.x, w.y)));
                } else if (MODE == 3) {   // FMUL2
                    x[i] = upk(mul2(pk(x[i].x, x[i].y), pk(w.x, w.y)));
                } else if (MODE == 4) {   // complex multiply: FMUL2 + FFMA2 with swap/neg/broadcast modifiers
                    u64 D = pk(x[i].x, x[i].y), Ds = pk(x[i].y, x[i].x);
                    x[i] = upk(fma2(Ds, pk(-w.y, w.y), mul2(D, pk(w.x, w.x))));
                } else if (MODE == 5) {   // scalar FADD
                    x[i].x = add1(x[i].x, w.x); x[i].y = add1(x[i].y, w.y);
                } else if (MODE == 6) {   // FFMA2 + one LDS.64 per 4 FFMA2
                    x[i] = upk(fma2(pk(x[i].x, x[i].y), pk(w.x, w.y), pk(w.y, w.x)));
                    if ((i & 3) == 0) { float2 l; asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(l.x), "=f"(l.y) : "r"(sa + 8 * ((i * 37 + r * 64) & 511))); x[(i + 5) % NACC].x += l.x; }
                } else if (MODE == 7) {   // scalar FFMA + one LDS.64 per 8 FFMA (same ratio per lane-op as mode 6)
                    x[i].x = fma1(x[i].x, w.x, w.y); x[i].y = fma1(x[i].y, w.y, w.x);
                    if ((i & 3) == 0) { float2 l; asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(l.x), "=f"(l.y) : "r"(sa + 8 * ((i * 37 + r * 64) & 511))); x[(i + 5) % NACC].x += l.x; }
                } else if (MODE == 8) {   // radix-2 butterfly packed: add2, sub2(neg), mul2, fma2 (4 instr / 2 complex)
                    if (i + 1 < NACC && (i & 1) == 0) {
                        u64 U = pk(x[i].x, x[i].y), V = pk(x[i + 1].x, x[i + 1].y);
                        u64 S = add2(U, V);
                        float2 dd = upk(add2(U, pk(-x[i + 1].x, -x[i + 1].y)));
                        u64 T = fma2(pk(dd.y, dd.x), pk(-w.y, w.y), mul2(pk(dd.x, dd.y), pk(w.x, w.x)));
                        x[i] = upk(S); x[i + 1] = upk(T);
                    }
                }
            }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NACC; ++i) s += x[i].x + x[i].y;
    if (s == 123.456f) out[0] = s;
}

template <int MODE>
void run(const char* name, double laneops_per_inner, double instr_per_inner) {
    float* d; cudaMalloc(&d, 256);
    int dev = 0, sms = 0, khz = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
    const int iters = 2048, blocks = sms * 2, threads = 512;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    kern<MODE><<<blocks, threads>>>(d, iters, 0.999f, 0.001f);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        kern<MODE><<<blocks, threads>>>(d, iters, 0.999f, 0.001f);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float t; cudaEventElapsedTime(&t, e0, e1); if (t < best) best = t;
    }
    const double inner = (double)iters * 8 * NACC * threads * blocks;   // inner-body executions (per thread)
    const double clk = best * 1e-3 * khz * 1e3;                         // SM clocks at nominal max clock
    printf("%-44s %8.3f ms  lane-ops/clk/SM %7.1f   warp-instr/clk/SM %5.2f\n", name, best,
           inner * laneops_per_inner / clk / sms, inner * instr_per_inner / 32.0 / clk / sms);
    cudaFree(d);
}

int main() {
    run<0>("FFMA scalar (2 per complex)", 2, 2);
    run<5>("FADD scalar (2 per complex)", 2, 2);
    run<1>("FFMA2", 2, 1);
    run<2>("FADD2", 2, 1);
    run<3>("FMUL2", 2, 1);
    run<4>("cmul = FMUL2 + FFMA2 (swap/neg/bcast)", 4, 2);
    run<6>("FFMA2 + LDS.64 every 4th", 2, 1.25);
    run<7>("FFMA x2 + LDS.64 every 4th pair", 2, 2.25);
    run<8>("packed DIF butterfly (4 instr / 2 cplx)", 4, 2);
    return 0;
}
