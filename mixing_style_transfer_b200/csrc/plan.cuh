// The plan object behind the opaque mixfe_plan handle of include/mixfe.h.
#pragma once

#include <algorithm>
#include <cstring>
#include <map>
#include <mutex>
#include <utility>

#include "common.cuh"

struct mixfe_plan {
    int device = 0, sr = 0, n_fft = 0, hop = 0, n_mels = 0, n_mels_pad = 0, n_bins = 0, D = 0, n_freq = 0;
    std::vector<float> fb;              // host [n_freq][n_mels]
    // device constant tables
    float2* d_tw = nullptr;
    float2* d_twlin = nullptr;          // exp(-2 pi i j / 1024), j in [0, 1024)
    float2* d_twfine = nullptr;         // exp(-2 pi i e / 2^20), e in [0, 1024): fine twiddles of the long FFT
    unsigned short* d_flo = nullptr;    // [128] byte offset of each filter's first walked bin in the power column
    float* d_wsteps = nullptr;          // per group of 8 filters: [steps][8] weights walked by K1's mel projection
    int n_wsteps = 0;                   // total steps over all groups
    int wmeta[8][2][4] = {};            // per warp, per group slot: steps, weight offset, group index, 0
    // K1 shared-memory map and persistent grid
    int nxs = 0, smem_k1 = 0, n_sms = 0, k1_ctas = 0;
    int* d_k1_work = nullptr;           // K1's two work counters (k1_features.cuh), zero between launches
    int reserved_sms = 0;               // SMs K1's persistent grid leaves free (for a collective running beside it)
    int off_ex = 0, off_tw = 0, off_s = 0, off_wsteps = 0, off_flo = 0, off_red = 0;
    // K1G (k1g_features.cuh): generic power-of-two n_fft path; used when K1 does not cover the configuration
    bool use_generic = false;
    float2* d_twg = nullptr;            // [n_fft/2] exp(-2 pi i e / n_fft)
    float* d_win = nullptr;             // [n_fft] 0.5 * periodic hann
    int4* d_mel_meta = nullptr;         // CSR form of the filterbank: (first bin, bins, weight offset, -) per filter
    float* d_mel_w = nullptr;
    int4* d_bin_meta = nullptr;         // CSR of the transposed filterbank (K6): (first filter, filters, weight offset, -) per bin
    float* d_bin_w = nullptr;
    // K1E (k1e_logmel2048.cuh): log-mel of n_fft = 2048 plans on the one-warp transform; null items = not usable
    int4* d_eitems = nullptr;
    float* d_ewts = nullptr;
    int e_nwts = 0, e_nxs = 0, e_off_ex = 0, e_off_tw = 0, e_off_items = 0, e_off_wts = 0, e_off_bar = 0, smem_e = 0, e_ctas = 0;
    // K6E (k6e_logmel_bwd2048.cuh): backward of (2048, 512) plans on the one-warp transform; null table = not usable
    float4* d_ebinw = nullptr;          // [1025] per bin (first filter, its weight, the next filter's weight, -)
    float2* d_ew2048 = nullptr;         // [513] exp(-2 pi i k / 2048)
    int x_off_tw = 0, x_off_binw = 0, x_off_w = 0, x_off_dmel = 0, x_off_ola = 0, smem_x = 0;
    float4* d_binw = nullptr;           // K6F: per bin (first filter, its weight, the next filter's weight, -); null = K6F not usable
    int f_off_ex = 0, f_off_tw = 0, f_off_binw = 0, f_off_dmel = 0, f_off_ola = 0, smem_f = 0;      // K6F shared-memory map
    int b_nf = 0, b_span = 0, b_span_pad = 0, b_off_pw = 0, b_off_mel = 0, b_off_ola = 0, smem_b = 0;   // K6 smem map (0 = unsupported)
    int g_nf = 0, g_log2n = 0, g_pw_stride = 0, g_off_pw = 0, g_off_mel = 0, g_off_S = 0, g_off_red = 0, smem_g = 0;
    // plan-owned workspace (grows to the high-water mark, then never reallocates)
    void* ws = nullptr;
    size_t ws_bytes = 0;
    // pinned staging ring for the per-call parameter tables (clip decisions, filter sections, reverb slots): a
    // pageable cudaMemcpyAsync waits for the stream to drain before it copies, which serialises the host with
    // the GPU once per call; a pinned source does not.  One slot per call in flight, guarded by an event; the
    // bytes reach the workspace through k_param_fetch (below), not through a copy engine.
    static constexpr int kParamSlots = 8;
    char* param_host[kParamSlots] = {};
    char* param_host_dev[kParamSlots] = {};     // the same bytes as the device sees them (mapped)
    size_t param_cap[kParamSlots] = {};
    cudaEvent_t param_done[kParamSlots] = {};
    int param_next = 0;
    // optional per-kernel timing (bench.py): event pairs recorded around launches, drained on query
    bool timing = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev[MIXFE_NUM_KERNELS];
};

namespace mixfe {

inline int ensure_workspace(mixfe_plan* pl, size_t bytes, cudaStream_t stream) {
    if (bytes <= pl->ws_bytes) return MIXFE_OK;
    if (pl->ws) {
        MIXFE_CUDA(cudaStreamSynchronize(stream));      // growth only: the old buffer may still be in use
        MIXFE_CUDA(cudaFree(pl->ws));
        pl->ws = nullptr;
        pl->ws_bytes = 0;
    }
    MIXFE_CUDA(cudaMalloc(&pl->ws, bytes));
    pl->ws_bytes = bytes;
    return MIXFE_OK;
}

// Copies a call's parameter tables from mapped pinned host memory into the workspace with SM loads over PCIe.  A
// cudaMemcpyAsync would do, but it rides the host-to-device copy engine, and when a pipeline keeps that engine busy
// with waveform uploads on another stream the few KB of tables queue behind tens of MB (measured: the kernels of a
// chunk started 0.5 ms late, scripts/e2e_timeline.py) -- a kernel on the compute stream has no such neighbour.
__global__ void __launch_bounds__(256) k_param_fetch(uint4* __restrict__ dst, const uint4* __restrict__ src, int n16) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += gridDim.x * blockDim.x) dst[i] = src[i];
}

// One call's parameter upload: take() hands out pinned bytes of the current ring slot (after the copy that last
// used the slot has finished), upload() queues the copy to the device, and the destructor marks the slot busy
// until the stream has passed the copies.
struct ParamUpload {
    mixfe_plan* pl;
    cudaStream_t stream;
    int slot;
    size_t used = 0;
    bool queued = false;
    ParamUpload(mixfe_plan* p, cudaStream_t s) : pl(p), stream(s), slot(p->param_next) {
        pl->param_next = (slot + 1) % mixfe_plan::kParamSlots;
    }
    int reserve(size_t bytes) {
        if (pl->param_done[slot]) MIXFE_CUDA(cudaEventSynchronize(pl->param_done[slot]));
        // every slot is (re)sized together, on the first call that needs the size: cudaHostAlloc takes a millisecond and
        // waits for the device, which must not happen in call 6, 7, 8 of a steady-state loop that warmed up for 5
        if (bytes > pl->param_cap[slot]) {
            const size_t cap = (bytes + 65535) & ~(size_t)65535;
            for (int i = 0; i < mixfe_plan::kParamSlots; ++i) {
                if (pl->param_cap[i] >= cap) continue;
                if (pl->param_done[i]) MIXFE_CUDA(cudaEventSynchronize(pl->param_done[i]));
                else MIXFE_CUDA(cudaEventCreateWithFlags(&pl->param_done[i], cudaEventDisableTiming));
                if (pl->param_host[i]) MIXFE_CUDA(cudaFreeHost(pl->param_host[i]));
                pl->param_host[i] = nullptr;
                pl->param_cap[i] = 0;
                MIXFE_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&pl->param_host[i]), cap, cudaHostAllocMapped));
                MIXFE_CUDA(cudaHostGetDevicePointer(reinterpret_cast<void**>(&pl->param_host_dev[i]), pl->param_host[i], 0));
                pl->param_cap[i] = cap;
            }
        }
        return MIXFE_OK;
    }
    // the tables of one call sit in the slot at the offsets they have in the device workspace, so that ONE copy moves them
    int stage(size_t offset, const void* src, size_t bytes) {
        if (offset + bytes > pl->param_cap[slot]) return fail(MIXFE_EINVAL, "parameter ring slot overrun");
        std::memcpy(pl->param_host[slot] + offset, src, bytes);
        used = std::max(used, offset + bytes);
        return MIXFE_OK;
    }
    int upload(void* dst_base) {
        if (!used) return MIXFE_OK;
        const int n16 = (int)((used + 15) / 16);                 // slots and workspace offsets are 256-byte aligned
        k_param_fetch<<<std::min(32, (n16 + 255) / 256), 256, 0, stream>>>(
            reinterpret_cast<uint4*>(dst_base), reinterpret_cast<const uint4*>(pl->param_host_dev[slot]), n16);
        MIXFE_CUDA(cudaGetLastError());
        launch_counter()++;
        queued = true;
        return MIXFE_OK;
    }
    ~ParamUpload() { if (queued) cudaEventRecord(pl->param_done[slot], stream); }
};

// RAII bracket: records an event pair around the launches issued in its scope when timing is on.
struct KernelTimer {
    mixfe_plan* pl;
    int which;
    cudaStream_t stream;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    KernelTimer(mixfe_plan* p, int w, cudaStream_t s) : pl(p), which(w), stream(s) {
        if (pl->timing && cudaEventCreate(&e0) == cudaSuccess && cudaEventCreate(&e1) == cudaSuccess)
            cudaEventRecord(e0, stream);
    }
    ~KernelTimer() {
        if (e0 && e1) {
            cudaEventRecord(e1, stream);
            pl->ev[which].emplace_back(e0, e1);
        }
    }
};

// cudaFuncAttributeMaxDynamicSharedMemorySize is state of (kernel, device), shared by every plan: a plan with a
// smaller footprint must not lower the limit an earlier plan launches with.  Keep a monotone per-(kernel, device)
// maximum and only ever raise it.
template <class K>
inline int smem_optin(K* kernel, int device, int bytes) {
    static std::mutex mu;
    static std::map<std::pair<const void*, int>, int> high;
    std::lock_guard<std::mutex> lock(mu);
    int& cur = high[{reinterpret_cast<const void*>(kernel), device}];
    if (bytes <= cur) return MIXFE_OK;
    MIXFE_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    cur = bytes;
    return MIXFE_OK;
}

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev); else prev = -1;
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

}  // namespace mixfe
