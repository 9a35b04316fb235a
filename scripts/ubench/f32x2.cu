// Microbenchmark: issue rate of scalar FFMA/FADD vs the packed FFMA2/FADD2 (PTX *.f32x2, sm_100+),
// alone and mixed with integer/LDS filler, at the occupancy K1 runs at (16 warps / SM).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o f32x2 f32x2.cu ; run on a B200.
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 d; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ float fma1(float a, float b, float c) { float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }
__device__ __forceinline__ float add1(float a, float b) { float d; asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b)); return d; }
__device__ __forceinline__ int iadd(int a, int b) { int d; asm volatile("add.s32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }

template <int MODE>
__global__ void __launch_bounds__(256) bench(float* out, int iters, float seed) {
    __shared__ float sm[1024];
    for (int i = threadIdx.x; i < 1024; i += 256) sm[i] = seed * i;
    __syncthreads();
    float a[16];
    u64 p[8];
    int q[8];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = seed * (threadIdx.x + i);
#pragma unroll
    for (int i = 0; i < 8; ++i) { float2 t = make_float2(a[2 * i], a[2 * i + 1]); p[i] = *reinterpret_cast<u64*>(&t); q[i] = threadIdx.x + i; }
    const float b = seed + 1.0f, c = seed * 0.5f;
    float2 bt = make_float2(b, b), ct = make_float2(c, c);
    const u64 b2 = *reinterpret_cast<u64*>(&bt), c2 = *reinterpret_cast<u64*>(&ct);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            if (MODE == 0) {           // 16 scalar FFMA
#pragma unroll
                for (int i = 0; i < 16; ++i) a[i] = fma1(a[i], b, c);
            } else if (MODE == 1) {    // 8 FFMA2 (same flops)
#pragma unroll
                for (int i = 0; i < 8; ++i) p[i] = fma2(p[i], b2, c2);
            } else if (MODE == 2) {    // 16 scalar FADD
#pragma unroll
                for (int i = 0; i < 16; ++i) a[i] = add1(a[i], b);
            } else if (MODE == 3) {    // 8 FADD2
#pragma unroll
                for (int i = 0; i < 8; ++i) p[i] = add2(p[i], b2);
            } else if (MODE == 4) {    // 8 FFMA2 + 8 IADD
#pragma unroll
                for (int i = 0; i < 8; ++i) { p[i] = fma2(p[i], b2, c2); q[i] = iadd(q[i], it); }
            } else if (MODE == 5) {    // 16 FFMA + 8 IADD
#pragma unroll
                for (int i = 0; i < 8; ++i) { a[2 * i] = fma1(a[2 * i], b, c); a[2 * i + 1] = fma1(a[2 * i + 1], b, c); q[i] = iadd(q[i], it); }
            } else if (MODE == 6) {    // 8 FFMA2 + 4 LDS
#pragma unroll
                for (int i = 0; i < 8; ++i) { p[i] = fma2(p[i], b2, c2); if (i & 1) q[i] += __float_as_int(sm[(q[i] + it) & 1023]); }
            } else if (MODE == 7) {    // 8 FFMA (half the flops) -- latency/ILP check
#pragma unroll
                for (int i = 0; i < 8; ++i) a[i] = fma1(a[i], b, c);
            }
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += a[i];
#pragma unroll
    for (int i = 0; i < 8; ++i) { float2 t = *reinterpret_cast<float2*>(&p[i]); s += t.x + t.y + q[i]; }
    out[blockIdx.x * 256 + threadIdx.x] = s;
}

template <int MODE>
void run(const char* name, double flops_per_iter, double inst_per_iter, float* out, int sms, double ghz) {
    const int iters = 4000, blocks = sms * 2;
    bench<MODE><<<blocks, 256>>>(out, 10, 0.001f);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    bench<MODE><<<blocks, 256>>>(out, iters, 0.001f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double warps = blocks * 8.0;
    const double tf = flops_per_iter * 4 * iters * warps * 32 / (ms * 1e-3) / 1e12;
    const double ipc = inst_per_iter * 4 * iters * warps / (ms * 1e-3 * ghz * 1e9) / (sms * 4);
    printf("%-28s %8.3f ms  %7.2f TFLOP/s  %5.2f warp-inst/clk/SMSP (at %.3f GHz nominal)\n", name, ms, tf, ipc, ghz);
}

int main() {
    cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
    int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double ghz = khz * 1e-6;
    float* out; cudaMalloc(&out, pr.multiProcessorCount * 2 * 256 * 4);
    printf("%s, %d SMs, %.3f GHz\n", pr.name, pr.multiProcessorCount, ghz);
    run<0>("16 FFMA", 32, 16, out, pr.multiProcessorCount, ghz);
    run<1>("8 FFMA2", 32, 8, out, pr.multiProcessorCount, ghz);
    run<2>("16 FADD", 16, 16, out, pr.multiProcessorCount, ghz);
    run<3>("8 FADD2", 16, 8, out, pr.multiProcessorCount, ghz);
    run<4>("8 FFMA2 + 8 IADD", 32, 16, out, pr.multiProcessorCount, ghz);
    run<5>("16 FFMA + 8 IADD", 32, 24, out, pr.multiProcessorCount, ghz);
    run<6>("8 FFMA2 + 4 LDS(+IADD)", 32, 16, out, pr.multiProcessorCount, ghz);
    run<7>("8 FFMA", 16, 8, out, pr.multiProcessorCount, ghz);
    return 0;
}
