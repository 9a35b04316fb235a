// Shared declarations of libmixfe (plan object, error plumbing, small device helpers).
#pragma once

#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/mixfe.h"

namespace mixfe {

constexpr int kNumStems = MIXFE_NUM_STEMS;
constexpr unsigned kFullMask = 0xffffffffu;

// ---- error plumbing -----------------------------------------------------------------------
inline std::string& last_error_slot() {
    static thread_local std::string s;
    return s;
}

inline int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    last_error_slot() = buf;
    return code;
}

#define MIXFE_CUDA(expr)                                                                     \
    do {                                                                                     \
        cudaError_t _e = (expr);                                                             \
        if (_e != cudaSuccess)                                                               \
            return ::mixfe::fail(MIXFE_ECUDA, "%s failed: %s (%s:%d)", #expr,                \
                                 cudaGetErrorString(_e), __FILE__, __LINE__);                \
    } while (0)

inline std::atomic<long long>& launch_counter() {
    static std::atomic<long long> c{0};
    return c;
}

// ---- device helpers -------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFullMask, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(kFullMask, v, o));
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFullMask, v, o);
    return v;
}

}  // namespace mixfe
