// libmixfe: C ABI (include/mixfe.h) over the sm_100a kernels.  One translation unit; built by
// mixing_style_transfer_b200/_build.py with
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -Xcompiler -fPIC -shared
#include <algorithm>
#include <cmath>
#include <cstring>
#include <numeric>

#include "common.cuh"
#include "k1_features.cuh"
#include "plan.cuh"
#include "k3_augment.cuh"
#include "k4_reverb.cuh"

using namespace mixfe;


namespace {

template <class T>
int upload(T** dst, const std::vector<T>& v) {
    if (*dst) { cudaFree(*dst); *dst = nullptr; }
    if (v.empty()) return MIXFE_OK;
    MIXFE_CUDA(cudaMalloc((void**)dst, v.size() * sizeof(T)));
    MIXFE_CUDA(cudaMemcpy(*dst, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    return MIXFE_OK;
}

inline int align16(int x) { return (x + 15) & ~15; }

// torchaudio.functional.melscale_fbanks(n_freqs, 0, sr//2, n_mels, sr, norm=None, mel_scale='htk')
// evaluated in double precision (the Python layer can override it with the exact fp32 matrix).
std::vector<float> default_filterbank(int sr, int n_fft, int n_mels) {
    const int n_freq = n_fft / 2 + 1;
    const double f_max = (double)(sr / 2);
    const double m_max = 2595.0 * std::log10(1.0 + f_max / 700.0);
    std::vector<double> f_pts(n_mels + 2);
    for (int i = 0; i < n_mels + 2; ++i) {
        const double mel = m_max * (double)i / (double)(n_mels + 1);
        f_pts[i] = 700.0 * (std::pow(10.0, mel / 2595.0) - 1.0);
    }
    std::vector<float> fb((size_t)n_freq * n_mels, 0.0f);
    for (int k = 0; k < n_freq; ++k) {
        const double f = f_max * (double)k / (double)(n_freq - 1);
        for (int m = 0; m < n_mels; ++m) {
            const double down = (f - f_pts[m]) / (f_pts[m + 1] - f_pts[m]);
            const double up = (f_pts[m + 2] - f) / (f_pts[m + 2] - f_pts[m + 1]);
            fb[(size_t)k * n_mels + m] = (float)std::max(0.0, std::min(down, up));
        }
    }
    return fb;
}

// Turn the (n_freq, n_mels) matrix into the grouped 2-tap tables K1 walks.
int build_mel_tables(mixfe_plan* pl) {
    const int nf = pl->n_freq, nm = pl->n_mels, nmp = pl->n_mels_pad;
    auto w = [&](int k, int m) -> float { return (k < nf && m < nm) ? pl->fb[(size_t)k * nm + m] : 0.0f; };
    std::vector<int> frow0(nmp, 0), rows(nmp, 0);
    for (int m = 0; m < nm; ++m) {
        int lo = -1, hi = -1;
        for (int k = 0; k < nf; ++k)
            if (w(k, m) != 0.0f) { if (lo < 0) lo = k; hi = k; }
        if (lo >= 0) { frow0[m] = lo >> 1; rows[m] = (hi >> 1) - frow0[m] + 1; }
    }
    const int ng = nmp / 4;
    std::vector<int2> gdesc(ng);
    std::vector<float2> gw;
    std::vector<int> cost(ng);
    for (int g = 0; g < ng; ++g) {
        int nmax = 0;
        for (int q = 0; q < 4; ++q) nmax = std::max(nmax, rows[4 * g + q]);
        nmax = (nmax + 1) & ~1;                          // K1 walks two rows per iteration
        gdesc[g] = make_int2(nmax, (int)gw.size());
        for (int j = 0; j < nmax; ++j)
            for (int q = 0; q < 4; ++q) {
                const int m = 4 * g + q, r = frow0[m] + j;
                if (r >= kPRows) return fail(MIXFE_EUNSUPPORTED, "mel filter %d reaches past the P tile", m);
                gw.push_back(make_float2(w(2 * r, m), w(2 * r + 1, m)));
            }
        cost[g] = 6 * nmax + 14;
    }
    // longest-processing-time-first assignment of groups to the 8 warps
    std::vector<int> order(ng);
    std::iota(order.begin(), order.end(), 0);
    std::sort(order.begin(), order.end(), [&](int a, int b) { return cost[a] != cost[b] ? cost[a] > cost[b] : a < b; });
    std::vector<unsigned char> wtab(8 * kMaxGroupsPerWarp + 8, 0);
    int load[8] = {0, 0, 0, 0, 0, 0, 0, 0}, cnt[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int g : order) {
        int best = -1;
        for (int wi = 0; wi < 8; ++wi)
            if (cnt[wi] < kMaxGroupsPerWarp && (best < 0 || load[wi] < load[best])) best = wi;
        if (best < 0) return fail(MIXFE_EUNSUPPORTED, "too many mel filter groups (%d)", ng);
        wtab[best * kMaxGroupsPerWarp + cnt[best]++] = (unsigned char)g;
        load[best] += cost[g];
    }
    for (int wi = 0; wi < 8; ++wi) wtab[8 * kMaxGroupsPerWarp + wi] = (unsigned char)cnt[wi];
    pl->n_groups = ng;
    pl->nw2 = (int)gw.size();
    int rc;
    if ((rc = upload(&pl->d_frow0, frow0))) return rc;
    if ((rc = upload(&pl->d_gdesc, gdesc))) return rc;
    if ((rc = upload(&pl->d_gw, gw))) return rc;
    if ((rc = upload(&pl->d_wtab, wtab))) return rc;
    // shared-memory map of K1
    pl->nxs = (kTF - 1) * pl->hop + kNfft;
    const int region0 = std::max(pl->nxs * 8, 2 * nmp * 16 * 4);
    pl->off_ex = align16(region0);
    pl->off_tw = pl->off_ex + 8 * 1024 * 8;
    pl->off_s = pl->off_tw + 1024 * 8;
    pl->off_frow0 = pl->off_s + 3 * nmp * 8 * 4;
    pl->off_gdesc = align16(pl->off_frow0 + nmp * 4);
    pl->off_gw = align16(pl->off_gdesc + ng * 8);
    pl->off_wtab = align16(pl->off_gw + pl->nw2 * 8);
    pl->off_red = align16(pl->off_wtab + 8 * kMaxGroupsPerWarp + 8);
    pl->smem_k1 = pl->off_red + 8 * 4 * 4;
    if (pl->smem_k1 > 227 * 1024)
        return fail(MIXFE_EUNSUPPORTED, "K1 needs %d bytes of shared memory (hop too large)", pl->smem_k1);
    MIXFE_CUDA(cudaFuncSetAttribute(k1_features, cudaFuncAttributeMaxDynamicSharedMemorySize, pl->smem_k1));
    return MIXFE_OK;
}

int run_features(mixfe_plan* pl, const float* const stems[4], long long stride_b, long long stride_c, int B,
                 int T, float* features, float* logmel, long long lm_stride_b, float* melpow,
                 long long mp_stride_b, cudaStream_t stream) {
    if (!pl || !stems) return fail(MIXFE_EINVAL, "null plan or stems");
    for (int s = 0; s < 4; ++s)
        if (!stems[s]) return fail(MIXFE_EINVAL, "stems[%d] is NULL", s);
    if (B < 0 || T < 0) return fail(MIXFE_EINVAL, "negative B or T");
    if (B == 0) return MIXFE_OK;
    if (T <= pl->n_fft / 2)
        return fail(MIXFE_ESHORT, "Argument #4: Padding size should be less than the corresponding input dimension, "
                                  "but got: padding (%d, %d) at dimension 2 of input of length %d",
                    pl->n_fft / 2, pl->n_fft / 2, T);
    if (stride_c < T || (B > 1 && stride_b < T)) return fail(MIXFE_EINVAL, "strides smaller than T");
    DeviceGuard guard(pl->device);
    const int F = 1 + T / pl->hop;
    const int n_tiles = (F + kTF - 1) / kTF;
    if (B > 65535) return fail(MIXFE_EINVAL, "B > 65535 clips per call");
    const size_t spec_bytes = (size_t)B * n_tiles * kNumStems * pl->n_mels_pad * sizeof(float2);
    const size_t td_bytes = (size_t)B * n_tiles * kTdN * sizeof(float);
    if (features) {
        int rc = ensure_workspace(pl, spec_bytes + td_bytes, stream);
        if (rc) return rc;
    }
    K1Params p{};
    for (int s = 0; s < 4; ++s) p.stem[s] = stems[s];
    p.stride_b = stride_b; p.stride_c = stride_c;
    p.B = B; p.T = T; p.F = F; p.hop = pl->hop; p.n_mels = pl->n_mels; p.n_mels_pad = pl->n_mels_pad;
    p.n_tiles = n_tiles; p.n_groups = pl->n_groups; p.nxs = pl->nxs;
    p.logmel = logmel; p.lm_stride_b = lm_stride_b; p.melpow = melpow; p.mp_stride_b = mp_stride_b;
    p.part_spec = reinterpret_cast<float2*>(pl->ws);
    p.part_td = reinterpret_cast<float*>(reinterpret_cast<char*>(pl->ws) + spec_bytes);
    p.tw = pl->d_tw; p.frow0 = pl->d_frow0; p.gdesc = pl->d_gdesc; p.gw = pl->d_gw; p.nw2 = pl->nw2;
    p.wtab = pl->d_wtab; p.want_stats = features ? 1 : 0;
    p.off_ex = pl->off_ex; p.off_tw = pl->off_tw; p.off_s = pl->off_s; p.off_frow0 = pl->off_frow0;
    p.off_gdesc = pl->off_gdesc; p.off_gw = pl->off_gw; p.off_wtab = pl->off_wtab; p.off_red = pl->off_red;
    {
        KernelTimer timer(pl, MIXFE_KERNEL_FEATURES, stream);
        k1_features<<<dim3(n_tiles, B), kK1Threads, pl->smem_k1, stream>>>(p);
    }
    MIXFE_CUDA(cudaGetLastError());
    launch_counter()++;
    if (features) {
        K2Params q{};
        q.part_spec = p.part_spec; q.part_td = p.part_td; q.features = features;
        q.B = B; q.T = T; q.F = F; q.n_mels = pl->n_mels; q.n_mels_pad = pl->n_mels_pad; q.n_tiles = n_tiles;
        q.n_bins = pl->n_bins; q.D = pl->D;
        {
            KernelTimer timer(pl, MIXFE_KERNEL_FINALIZE, stream);
            k2_finalize<<<B, kK2Threads, 0, stream>>>(q);
        }
        MIXFE_CUDA(cudaGetLastError());
        launch_counter()++;
    }
    return MIXFE_OK;
}

}  // namespace

extern "C" {

int mixfe_abi_version(void) { return MIXFE_ABI_VERSION; }

const char* mixfe_last_error(void) { return last_error_slot().c_str(); }

long long mixfe_launch_count(void) { return launch_counter().load(); }

int mixfe_plan_create(mixfe_plan** out, int device, int sample_rate, int n_fft, int hop, int n_mels,
                      int n_spectral_bins) {
    if (!out) return fail(MIXFE_EINVAL, "out is NULL");
    *out = nullptr;
    if (sample_rate <= 0 || hop <= 0 || n_mels <= 0 || n_spectral_bins < 0)
        return fail(MIXFE_EINVAL, "bad plan arguments");
    if (n_fft != kNfft)
        return fail(MIXFE_EUNSUPPORTED, "n_fft=%d: this build's fused STFT kernel covers n_fft=1024 only", n_fft);
    if (n_mels > kMaxMels) return fail(MIXFE_EUNSUPPORTED, "n_mels=%d > %d", n_mels, kMaxMels);
    if (n_spectral_bins > n_mels)
        return fail(MIXFE_EINVAL, "n_spectral_bins=%d > n_mels=%d (the reference's torch.stack raises)",
                    n_spectral_bins, n_mels);
    int ndev = 0;
    MIXFE_CUDA(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(MIXFE_EINVAL, "device %d out of range (%d visible)", device, ndev);
    DeviceGuard guard(device);
    cudaDeviceProp prop;
    MIXFE_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail(MIXFE_EUNSUPPORTED, "device %d is sm_%d%d; libmixfe is built for sm_100a only", device,
                    prop.major, prop.minor);
    mixfe_plan* pl = new mixfe_plan();
    pl->device = device; pl->sr = sample_rate; pl->n_fft = n_fft; pl->hop = hop; pl->n_mels = n_mels;
    pl->n_mels_pad = (n_mels + 15) & ~15; pl->n_bins = n_spectral_bins; pl->n_freq = n_fft / 2 + 1;
    const int sd = n_spectral_bins > 0 ? n_spectral_bins + 2 : 5;
    pl->D = 4 * (6 + sd + 3) + 4 + 4;
    pl->fb = default_filterbank(sample_rate, n_fft, n_mels);
    std::vector<float2> tw(1024), twf(1024), twl(1024);
    for (int k1 = 0; k1 < 32; ++k1)
        for (int n2 = 0; n2 < 32; ++n2) {
            const double a = -2.0 * M_PI * (double)(k1 * n2) / 1024.0;
            tw[k1 * 32 + n2] = make_float2((float)std::cos(a), (float)std::sin(a));
        }
    for (int e = 0; e < 1024; ++e) {
        const double a = -2.0 * M_PI * (double)e / 1048576.0;
        twf[e] = make_float2((float)std::cos(a), (float)std::sin(a));
        const double c = -2.0 * M_PI * (double)e / 1024.0;
        twl[e] = make_float2((float)std::cos(c), (float)std::sin(c));
    }
    int rc = upload(&pl->d_tw, tw);
    if (!rc) rc = upload(&pl->d_twfine, twf);
    if (!rc) rc = upload(&pl->d_twlin, twl);
    if (!rc) rc = build_mel_tables(pl);
    if (rc) { mixfe_plan_destroy(pl); return rc; }
    *out = pl;
    return MIXFE_OK;
}

int mixfe_plan_destroy(mixfe_plan* pl) {
    if (!pl) return MIXFE_OK;
    DeviceGuard guard(pl->device);
    cudaFree(pl->d_tw); cudaFree(pl->d_twfine); cudaFree(pl->d_twlin); cudaFree(pl->d_frow0); cudaFree(pl->d_gdesc);
    cudaFree(pl->d_gw); cudaFree(pl->d_wtab); cudaFree(pl->ws); cudaFree(pl->aug_params_dev);
    if (pl->aug_params_host) cudaFreeHost(pl->aug_params_host);
    delete pl;
    return MIXFE_OK;
}

int mixfe_plan_set_filterbank(mixfe_plan* pl, const float* fb_host) {
    if (!pl || !fb_host) return fail(MIXFE_EINVAL, "null plan or filterbank");
    DeviceGuard guard(pl->device);
    MIXFE_CUDA(cudaDeviceSynchronize());            // tables may be in use by queued launches
    pl->fb.assign(fb_host, fb_host + (size_t)pl->n_freq * pl->n_mels);
    return build_mel_tables(pl);
}

int mixfe_plan_enable_timing(mixfe_plan* pl, int enable) {
    if (!pl) return fail(MIXFE_EINVAL, "null plan");
    pl->timing = enable != 0;
    return MIXFE_OK;
}

int mixfe_plan_kernel_time(mixfe_plan* pl, int which, double* total_ms, long long* launches) {
    if (!pl || which < 0 || which >= MIXFE_NUM_KERNELS) return fail(MIXFE_EINVAL, "bad plan or kernel id");
    DeviceGuard guard(pl->device);
    double tot = 0.0;
    long long n = 0;
    for (auto& pr : pl->ev[which]) {
        MIXFE_CUDA(cudaEventSynchronize(pr.second));
        float ms = 0.0f;
        MIXFE_CUDA(cudaEventElapsedTime(&ms, pr.first, pr.second));
        tot += ms;
        ++n;
        cudaEventDestroy(pr.first);
        cudaEventDestroy(pr.second);
    }
    pl->ev[which].clear();
    if (total_ms) *total_ms = tot;
    if (launches) *launches = n;
    return MIXFE_OK;
}

int mixfe_feature_dim(const mixfe_plan* pl) { return pl ? pl->D : fail(MIXFE_EINVAL, "null plan"); }

int mixfe_num_frames(const mixfe_plan* pl, int T) { return pl ? 1 + T / pl->hop : fail(MIXFE_EINVAL, "null plan"); }

size_t mixfe_workspace_bytes(const mixfe_plan* pl) { return pl ? pl->ws_bytes : 0; }

int mixfe_features_logmel(mixfe_plan* pl, const float* const stems[MIXFE_NUM_STEMS], long long stride_b,
                          long long stride_c, int B, int T, float* features, float* logmel,
                          long long logmel_stride_b, void* cuda_stream) {
    return run_features(pl, stems, stride_b, stride_c, B, T, features, logmel, logmel_stride_b, nullptr, 0,
                        static_cast<cudaStream_t>(cuda_stream));
}

int mixfe_debug_mel_power(mixfe_plan* pl, const float* const stems[MIXFE_NUM_STEMS], long long stride_b,
                          long long stride_c, int B, int T, float* melpow, long long melpow_stride_b,
                          void* cuda_stream) {
    return run_features(pl, stems, stride_b, stride_c, B, T, nullptr, nullptr, 0, melpow, melpow_stride_b,
                        static_cast<cudaStream_t>(cuda_stream));
}

int mixfe_augment(mixfe_plan* pl, const float* const in[MIXFE_NUM_STEMS], float* const out[MIXFE_NUM_STEMS],
                  long long stride_b, long long stride_c, int B, int T, const mixfe_aug_params* params,
                  const float* impulse, int L, void* cuda_stream) {
    return run_augment(pl, in, out, stride_b, stride_c, B, T, params, impulse, L,
                       static_cast<cudaStream_t>(cuda_stream));
}

}  // extern "C"
