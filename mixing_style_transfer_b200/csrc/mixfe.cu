// libmixfe: C ABI (include/mixfe.h) over the sm_100a kernels.  One translation unit; built by
// mixing_style_transfer_b200/_build.py with
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -Xcompiler -fPIC -shared
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <numeric>

#include "common.cuh"
#include "k1_features.cuh"
#include "k1g_features.cuh"
#include "k1e_logmel2048.cuh"
#include "k6_logmel_bwd.cuh"
#include "k6f_logmel_bwd.cuh"
#include "k6e_logmel_bwd2048.cuh"
#include "plan.cuh"
#include "k3_augment.cuh"
#include "k4_reverb.cuh"
#include "k7_loudness.cuh"
#include "k8_mrstft.cuh"
#include "k8f_mrstft1024.cuh"
#include "k8e_mrstft2048.cuh"
#include "k8h_mrstft512.cuh"
#include "k9_extensions.cuh"

using namespace mixfe;

// K11: the all-gather of BASELINE config 5 as ONE kernel on a few SMs: 128-bit loads of this rank's slice, 128-bit stores
// into the same offset of every peer's buffer over NVLink peer mappings (torch symmetric memory).  It runs beside K1's
// persistent grid (mixfe_plan_reserve_sms leaves it the SMs), so the gather of chunk c overlaps the kernels of chunk c + 1.
constexpr int kMaxPushPeers = 15;
constexpr int kPushThreads = 512;
struct PushParams {
    const void* src;
    void* dst[kMaxPushPeers];
    size_t bytes;
    int n_peers;
};
__global__ void __launch_bounds__(kPushThreads) k11_push_peers(const __grid_constant__ PushParams p) {
    const size_t n = p.bytes / 16, stride = (size_t)gridDim.x * kPushThreads;
    const int4* __restrict__ src = reinterpret_cast<const int4*>(p.src);
    for (size_t i = (size_t)blockIdx.x * kPushThreads + threadIdx.x; i < n; i += 4 * stride) {
        int4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (i + u * stride < n) v[u] = __ldg(src + i + u * stride);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (i + u * stride < n) {
                for (int q = 0; q < p.n_peers; ++q) reinterpret_cast<int4*>(p.dst[q])[i + u * stride] = v[u];
            }
        }
    }
}

// Scratch of mixfe_weighted_masking_noalloc: the per-(clip, 128-frame block) partial sums K9 exchanges between its two
// launches, held in the library image instead of the plan's cudaMalloc'ed workspace.  256 KB on every GPU that loads
// libmixfe, until exit.  This name is a HOST-SIDE STAND-IN in host code: the device address comes from
// cudaGetSymbolAddress (noalloc_setup below, once per device with that device current) and only that goes to a kernel.
constexpr int kNoAllocMaskEntries = MIXFE_NOALLOC_MASKING_MAX_BLOCKS;
static __device__ __align__(256) float g_noalloc_mask_part[kNoAllocMaskEntries * 8];

namespace {

constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mu;
float* g_noalloc_part[kNoAllocMaxDevices] = {};      // device address of g_noalloc_mask_part per device ordinal

// Set-up half (called from mixfe_plan_create with `device` current): resolve the array's address on this device once.
int noalloc_setup(int device) {
    if (device < 0 || device >= kNoAllocMaxDevices) return MIXFE_OK;     // such a device simply has no no-alloc entry
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (g_noalloc_part[device]) return MIXFE_OK;
    float* ptr = nullptr;
    const cudaError_t err = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_noalloc_mask_part);
    if (err != cudaSuccess || !ptr)
        return fail(MIXFE_ECUDA, "cudaGetSymbolAddress(g_noalloc_mask_part) on device %d: %s", device, cudaGetErrorString(err));
    g_noalloc_part[device] = ptr;
    return MIXFE_OK;
}

template <class T>
int upload(T** dst, const std::vector<T>& v) {
    if (*dst) { cudaFree(*dst); *dst = nullptr; }
    if (v.empty()) return MIXFE_OK;
    MIXFE_CUDA(cudaMalloc((void**)dst, v.size() * sizeof(T)));
    MIXFE_CUDA(cudaMemcpy(*dst, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    return MIXFE_OK;
}

inline int align16(int x) { return (x + 15) & ~15; }

inline float __int_as_float_host(int v) {
    float f;
    std::memcpy(&f, &v, sizeof f);
    return f;
}

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

// Turn the (n_freq, n_mels) matrix into the tables of K1's tile-level mel projection.  Filters are taken in
// groups of 8 consecutive ones; a warp owns two groups (a wide one and a narrow one, paired so that the warps'
// walks have similar lengths) and lane (q, i) walks filter i of each group over frames 2q, 2q+1 of the tile.
// All lanes of a warp walk a group for n_g steps = its longest support (rounded up to even), starting at
// start_m <= lo_m; weights outside the support are zero.
int build_mel_tables(mixfe_plan* pl) {
    const int nf = pl->n_freq, nm = pl->n_mels;
    const int n_groups = (nm + 7) / 8;
    if (n_groups > 2 * (kK1Threads / 32)) return fail(MIXFE_EUNSUPPORTED, "n_mels=%d > %d", nm, 16 * (kK1Threads / 32));
    auto w = [&](int k, int m) -> float { return (k >= 0 && k < nf && m < nm) ? pl->fb[(size_t)k * nm + m] : 0.0f; };
    std::vector<int> lo(n_groups * 8, 0), len(n_groups * 8, 0);
    for (int m = 0; m < nm; ++m) {
        int first = -1, last = -1;
        for (int k = 0; k < nf; ++k)
            if (w(k, m) != 0.0f) { if (first < 0) first = k; last = k; }
        if (first >= 0) { lo[m] = first; len[m] = last - first + 1; }
    }
    std::vector<unsigned short> fstart(kMaxMels, 0);
    std::vector<int> gn(n_groups, 0), gwoff(n_groups, 0);
    std::vector<float> wt;
    int n_steps = 0;
    for (int g = 0; g < n_groups; ++g) {
        int nmin = 0;
        for (int i = 0; i < 8; ++i) nmin = std::max(nmin, len[8 * g + i]);
        nmin = std::max(2, (nmin + 1) & ~1);              // K1 walks two steps per iteration
        // The 64-bit loads of one quarter-pair (16 lanes: 8 filters x 2 frame columns that sit 16 banks apart)
        // are free of bank conflicts when the 8 start bins have distinct residues mod 8 (equal starts broadcast).
        // Use the slack of the walk (and up to 4 extra steps) to get there: bipartite matching filter -> residue.
        int n = nmin;
        std::vector<int> start(8, 0);
        bool done = false;
        for (; n <= nmin + 4 && !done; n += 2) {
            auto range = [&](int i, int& lw, int& hi) {
                const int m = 8 * g + i;
                hi = len[m] ? std::min(lo[m], kPvEntries - n) : kPvEntries - n;
                lw = len[m] ? std::max(0, lo[m] + len[m] - n) : 0;
            };
            auto start_for = [&](int i, int r) -> int {
                int lw, hi;
                range(i, lw, hi);
                for (int st = hi; st >= lw; --st)
                    if ((st & 7) == r) return st;
                return -1;
            };
            std::vector<int> owner(8, -1);
            std::function<bool(int, std::vector<char>&)> place = [&](int i, std::vector<char>& seen) -> bool {
                for (int r = 0; r < 8; ++r) {
                    if (seen[r] || start_for(i, r) < 0) continue;
                    seen[r] = 1;
                    if (owner[r] < 0 || place(owner[r], seen)) { owner[r] = i; return true; }
                }
                return false;
            };
            int placed = 0;
            for (int i = 0; i < 8; ++i) {
                std::vector<char> seen(8, 0);
                if (place(i, seen)) ++placed;
            }
            if (placed == 8) {
                for (int r = 0; r < 8; ++r) start[owner[r]] = start_for(owner[r], r);
                done = true;
                break;
            }
        }
        if (!done) {                                       // exact starts; the loads may take extra wavefronts
            n = nmin;
            for (int i = 0; i < 8; ++i) {
                const int m = 8 * g + i;
                start[i] = len[m] ? std::min(lo[m], kPvEntries - n) : 0;
                if (len[m] && lo[m] + len[m] > start[i] + n)
                    return fail(MIXFE_EUNSUPPORTED, "mel filter %d (bins %d..%d) overruns the power column", m, lo[m], lo[m] + len[m] - 1);
            }
        }
        gn[g] = n;
        gwoff[g] = (int)wt.size();
        for (int i = 0; i < 8; ++i) fstart[8 * g + i] = (unsigned short)(8 * start[i]);
        for (int st = 0; st < n; ++st)
            for (int i = 0; i < 8; ++i) {
                const int m = 8 * g + i, k = start[i] + st;
                wt.push_back((len[m] && k >= lo[m] && k < lo[m] + len[m]) ? w(k, m) : 0.0f);
            }
        n_steps += n;
    }
    // pair the longest walk with the shortest, the second longest with the second shortest, ...
    std::vector<int> order(n_groups);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return gn[a] > gn[b]; });
    const int n_warps = kK1Threads / 32;
    for (int wv = 0; wv < n_warps; ++wv)
        for (int ab = 0; ab < 2; ++ab) {
            const int pos = ab == 0 ? wv : 2 * n_warps - 1 - wv;
            int* me = pl->wmeta[wv][ab];
            if (pos < n_groups) { me[0] = gn[order[pos]]; me[1] = gwoff[order[pos]]; me[2] = order[pos]; }
            else { me[0] = 0; me[1] = 0; me[2] = 2 * n_warps; }          // no group: filters >= n_mels_pad
            me[3] = 0;
        }
    pl->n_wsteps = n_steps;
    const int n_wt = (int)wt.size();
    int rc;
    if ((rc = upload(&pl->d_flo, fstart))) return rc;
    if ((rc = upload(&pl->d_wsteps, wt))) return rc;
    // shared-memory map of K1
    pl->nxs = (kTF - 1) * pl->hop + kNfft;
    pl->off_ex = (pl->nxs * 8 + 127) & ~127;          // exchange regions: bank 0, 16-byte aligned for the TMA statistics rows
    pl->off_tw = pl->off_ex + 8 * kWarpRegion;
    pl->off_s = pl->off_tw + 31 * 32 * 8;            // rows k1 = 1..31 of the twiddle table (row 0 is all ones)
    pl->off_wsteps = pl->off_s + 3 * kK1Threads * 16;   // S: one float4 per thread and stem 0..2
    pl->off_flo = align16(pl->off_wsteps + n_wt * 4);
    pl->off_red = align16(pl->off_flo + kMaxMels * 2);           // two mbarriers and the arrival counter
    pl->smem_k1 = pl->off_red + 32;
    if (pl->smem_k1 > 227 * 1024)
        return fail(MIXFE_EUNSUPPORTED, "K1 needs %d bytes of shared memory (hop too large or filters too wide)", pl->smem_k1);
    if (int rc = smem_optin(k1_features, pl->device, pl->smem_k1)) return rc;
    if (!pl->d_k1_work) {
        MIXFE_CUDA(cudaMalloc(&pl->d_k1_work, 2 * sizeof(int)));
        MIXFE_CUDA(cudaMemset(pl->d_k1_work, 0, 2 * sizeof(int)));
    }
    int per_sm = 0;
    MIXFE_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k1_features, kK1Threads, pl->smem_k1));
    pl->k1_ctas = std::max(1, per_sm) * pl->n_sms;
    // K6F (backward on the one-warp transform): hop 256 and at most two ADJACENT filters per bin (any triangular bank)
    if (pl->d_binw) { cudaFree(pl->d_binw); pl->d_binw = nullptr; }
    if (pl->hop == 256) {
        std::vector<float4> binw(nf);
        bool ok = true;
        for (int k = 0; k < nf && ok; ++k) {
            int first = -1, last = -1;
            for (int m = 0; m < nm; ++m)
                if (w(k, m) != 0.0f) { if (first < 0) first = m; last = m; }
            if (first < 0) { first = last = 0; }
            if (last - first > 1) { ok = false; break; }
            float4 e;
            e.x = __int_as_float_host(first);
            e.y = w(k, first);
            e.z = last > first ? w(k, last) : 0.0f;
            e.w = 0.0f;
            binw[k] = e;
        }
        if (ok) {
            if ((rc = upload(&pl->d_binw, binw))) return rc;
            pl->f_off_ex = 0;
            pl->f_off_tw = kBwWarps * kWarpRegion;
            pl->f_off_binw = pl->f_off_tw + 31 * 32 * 8;
            pl->f_off_dmel = align16(pl->f_off_binw + nf * 16);
            pl->f_off_ola = align16(pl->f_off_dmel + kBwWarps * (nm + 1) * 8);
            pl->smem_f = pl->f_off_ola + kBwSpan * 8;
            if (pl->smem_f > 227 * 1024) { cudaFree(pl->d_binw); pl->d_binw = nullptr; }
            else if ((rc = smem_optin(k6f_logmel_bwd, pl->device, pl->smem_f))) return rc;
        }
    }
    return MIXFE_OK;
}

// CSR form of the filterbank + window + twiddles + shared-memory map of the generic kernel K1G.
int build_generic_tables(mixfe_plan* pl) {
    const int nf = pl->n_freq, nm = pl->n_mels, n = pl->n_fft;
    std::vector<int> lo(nm, 0), len(nm, 0), off(nm, 0);
    std::vector<float> w;
    for (int m = 0; m < nm; ++m) {
        int first = -1, last = -1;
        for (int k = 0; k < nf; ++k)
            if (pl->fb[(size_t)k * nm + m] != 0.0f) { if (first < 0) first = k; last = k; }
        off[m] = (int)w.size();
        if (first >= 0) {
            lo[m] = first; len[m] = last - first + 1;
            for (int k = first; k <= last; ++k) w.push_back(pl->fb[(size_t)k * nm + m]);
        }
    }
    if (w.empty()) w.push_back(0.0f);
    std::vector<float2> tw(n / 2);
    for (int e = 0; e < n / 2; ++e) {
        const double a = -2.0 * M_PI * (double)e / (double)n;
        tw[e] = make_float2((float)std::cos(a), (float)std::sin(a));
    }
    // torch.hann_window(n, periodic=True): evaluated here in double precision and rounded once (the fp32 values torch
    // produces differ by at most one ulp, far below the parity tolerance).
    std::vector<float> win(n);
    for (int i = 0; i < n; ++i) win[i] = (float)(0.5 * (0.5 - 0.5 * std::cos(2.0 * M_PI * (double)i / (double)n)));
    int rc;
    std::vector<int4> mel_meta(nm);
    for (int m = 0; m < nm; ++m) mel_meta[m] = make_int4(lo[m], len[m], off[m], 0);
    if ((rc = upload(&pl->d_mel_meta, mel_meta)) ||
        (rc = upload(&pl->d_mel_w, w)) || (rc = upload(&pl->d_twg, tw)) || (rc = upload(&pl->d_win, win)))
        return rc;
    pl->g_log2n = 0;
    while ((1 << pl->g_log2n) < n) ++pl->g_log2n;
    pl->g_pw_stride = (n / 2 + 1 + (pl->g_log2n > 6 ? (n / 2) >> (pl->g_log2n - 5) : 0)) | 1;   // k + (k >> pad_sh), k <= n/2
    const int per_frame = fg_row(n) * 8 + pl->g_pw_stride * 8;
    int nfr = kTF;
    {
        const char* e = getenv("MIXFE_K1G_SMEM_KB");            // tuning knob: frame-buffer budget per CTA
        const int budget = (e ? atoi(e) : 72) * 1024;       // 72 KB: 2-3 CTAs per SM beat one CTA with all 8 frames in flight (profiles/README.md)
        while (nfr > 1 && nfr * per_frame > budget) nfr >>= 1;
    }
    pl->g_nf = nfr;
    pl->g_off_pw = nfr * fg_row(n) * 8;
    const int stage = pl->g_off_pw + nfr * pl->g_pw_stride * 8;
    pl->g_off_mel = align16(stage);
    pl->g_off_S = pl->g_off_mel + kTF * pl->n_mels_pad * 8;
    pl->g_off_red = pl->g_off_S + kNumStems * kTF * pl->n_mels_pad * 4;
    pl->smem_g = pl->g_off_red + 8 * 4 * 4;
    if (pl->smem_g > 227 * 1024) return fail(MIXFE_EUNSUPPORTED, "K1G needs %d bytes of shared memory", pl->smem_g);
    if (int rc = smem_optin(k1g_features, pl->device, pl->smem_g)) return rc;
    // K6 (log-mel backward): filterbank by bin + its shared-memory map
    {
        // Entry i describes bin k = bitrev_{log2n-1}(i) (i = n/2: the Nyquist bin): K6's lane i works on that bin
        // (k6_logmel_bwd.cuh), so the table and the weights are laid out in the order the lanes read them.
        std::vector<int4> bin_meta(nf);
        std::vector<float> bw;
        for (int i = 0; i < nf; ++i) {
            int k = n / 2;
            if (i < n / 2) {
                k = 0;
                for (int bit = 0; bit < pl->g_log2n - 1; ++bit) k |= ((i >> bit) & 1) << (pl->g_log2n - 2 - bit);
            }
            int first = -1, last = -1;
            for (int m = 0; m < nm; ++m)
                if (pl->fb[(size_t)k * nm + m] != 0.0f) { if (first < 0) first = m; last = m; }
            bin_meta[i] = make_int4(0, 0, (int)bw.size(), 0);
            if (first >= 0) {
                bin_meta[i].x = first; bin_meta[i].y = last - first + 1;
                for (int m = first; m <= last; ++m) bw.push_back(pl->fb[(size_t)k * nm + m]);
            }
        }
        if (bw.empty()) bw.push_back(0.0f);
        if ((rc = upload(&pl->d_bin_meta, bin_meta)) ||
            (rc = upload(&pl->d_bin_w, bw)))
            return rc;
        pl->b_span = (kTF - 1) * pl->hop + n;
        pl->b_span_pad = pl->b_span + (pl->g_log2n > 5 ? pl->b_span >> (pl->g_log2n - 5) : 0) + 1;
        const int per = fg_row(n) * 8 + pl->g_pw_stride * 8 + pl->n_mels_pad * 8;
        const int ola = align16(2 * pl->b_span_pad * 4);
        int bn = kTF;
        const char* e6 = getenv("MIXFE_K6_SMEM_KB");
        const int budget6 = (e6 ? atoi(e6) : 72) * 1024;
        while (bn > 1 && bn * per + ola + 64 > budget6) bn >>= 1;
        pl->b_nf = bn;
        pl->b_off_pw = bn * fg_row(n) * 8;
        pl->b_off_mel = pl->b_off_pw + bn * pl->g_pw_stride * 8;
        pl->b_off_ola = align16(pl->b_off_mel + bn * pl->n_mels_pad * 8);
        pl->smem_b = pl->b_off_ola + ola;
        if (pl->smem_b > 227 * 1024) pl->smem_b = 0;          // hop too large for the overlap-add buffer: backward unsupported
        else if (int rc6 = smem_optin(k6_logmel_bwd, pl->device, pl->smem_b)) return rc6;
    }
    return MIXFE_OK;
}

// Twiddles of the one-warp 1024-point transform (K1, K4) and of the 2^20-point four-step FFT of the reverb (K4).
int upload_fft_twiddles(mixfe_plan* pl) {
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
    return rc;
}

// K1E tables (n_fft = 2048): the filterbank by filter cut into <= 128 work items, longest supports first, the longest
// filters split into two halves in neighbouring items (k1e_logmel2048.cuh).
int build_k1e_tables(mixfe_plan* pl) {
    if (pl->d_eitems) { cudaFree(pl->d_eitems); pl->d_eitems = nullptr; }
    if (pl->d_ewts) { cudaFree(pl->d_ewts); pl->d_ewts = nullptr; }
    if (pl->n_fft != 2048 || pl->hop % 4 || pl->n_mels > kEItems) return MIXFE_OK;
    const int nf = pl->n_freq, nm = pl->n_mels;
    std::vector<int> lo(nm, 0), len(nm, 0), woff(nm, 0);
    std::vector<float> w;
    for (int m = 0; m < nm; ++m) {
        int first = -1, last = -1;
        for (int k = 0; k < nf; ++k)
            if (pl->fb[(size_t)k * nm + m] != 0.0f) { if (first < 0) first = k; last = k; }
        woff[m] = (int)w.size();
        if (first >= 0) {
            lo[m] = first; len[m] = last - first + 1;
            for (int k = first; k <= last; ++k) w.push_back(pl->fb[(size_t)k * nm + m]);
        }
    }
    if (w.empty()) w.push_back(0.0f);
    std::vector<int> order(nm);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return len[a] > len[b]; });
    // Cut the longest supports into 2 or 4 parts (neighbouring items, joined by shuffles) while items remain: greedy on
    // the longest piece.  Groups of 4 first, then pairs, then whole filters -- each kind sorted by piece length -- so
    // that groups stay aligned to their size and a warp's 32 lanes walk similar lengths.
    std::vector<int> parts(nm, 1);
    int used = nm;
    for (;;) {
        int best = -1, best_len = 7;                                       // pieces shorter than 8 bins are not worth a shuffle
        for (int m = 0; m < nm; ++m) {
            const int piece = (len[m] + parts[m] - 1) / parts[m];
            if (parts[m] < 4 && used + parts[m] <= kEItems && piece > best_len) { best = m; best_len = piece; }
        }
        if (best < 0) break;
        used += parts[best];
        parts[best] *= 2;
    }
    std::vector<int4> items;
    for (int kind = 4; kind >= 1; kind >>= 1) {
        std::vector<int> ms;
        for (int m : order)
            if (parts[m] == kind) ms.push_back(m);                         // `order` is by length, longest first
        for (int m : ms) {
            const int piece = (len[m] + kind - 1) / kind;
            for (int i = 0; i < kind; ++i) {
                const int o = std::min(len[m], i * piece), n = std::min(len[m], (i + 1) * piece) - o;
                const int lg = kind == 4 ? 2 : (kind == 2 ? 1 : 0);
                items.push_back(make_int4(m, lo[m] + o, n, (woff[m] + o) | (lg << 28) | (i ? (1 << 30) : 0)));
            }
        }
    }
    while ((int)items.size() < kEItems) items.push_back(make_int4(-1, 0, 0, 0));
    pl->e_nwts = (int)w.size();
    pl->e_nxs = (kETF - 1) * pl->hop + 2048;
    pl->e_off_ex = (pl->e_nxs * 8 + 127) & ~127;
    pl->e_off_tw = pl->e_off_ex + 8 * kWarpRegion;
    pl->e_off_items = pl->e_off_tw + 31 * 32 * 8;
    pl->e_off_wts = pl->e_off_items + kEItems * 16;
    pl->e_off_bar = align16(pl->e_off_wts + pl->e_nwts * 4);
    pl->smem_e = pl->e_off_bar + 32;
    if (pl->smem_e > 227 * 1024) return MIXFE_OK;                          // hop too large: K1G serves the plan
    int rc;
    if ((rc = upload(&pl->d_eitems, items)) || (rc = upload(&pl->d_ewts, w))) return rc;
    if (!pl->d_k1_work) {
        MIXFE_CUDA(cudaMalloc(&pl->d_k1_work, 2 * sizeof(int)));
        MIXFE_CUDA(cudaMemset(pl->d_k1_work, 0, 2 * sizeof(int)));
    }
    if ((rc = smem_optin(k1e_logmel2048<false>, pl->device, pl->smem_e)) || (rc = smem_optin(k1e_logmel2048<true>, pl->device, pl->smem_e)))
        return rc;
    int per_sm = 0;
    MIXFE_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k1e_logmel2048<true>, kEThreads, pl->smem_e));
    pl->e_ctas = std::max(1, per_sm) * pl->n_sms;
    return MIXFE_OK;
}

// Which kernel serves the plan: K1 for n_fft = 1024 when its shared-memory map fits, else K1G.
// K6E: per-bin filter pairs of the 1025 one-sided bins, the 2048-point twiddles, the shared-memory map.
int build_k6e_tables(mixfe_plan* pl) {
    if (pl->d_ebinw) { cudaFree(pl->d_ebinw); pl->d_ebinw = nullptr; }
    if (pl->d_ew2048) { cudaFree(pl->d_ew2048); pl->d_ew2048 = nullptr; }
    if (pl->n_fft != 2048 || pl->hop != 512) return MIXFE_OK;
    const int nf = pl->n_freq, nm = pl->n_mels;                     // nf = 1025
    std::vector<float4> binw(nf);
    for (int k = 0; k < nf; ++k) {
        int first = -1, last = -1;
        for (int m = 0; m < nm; ++m)
            if (pl->fb[(size_t)k * nm + m] != 0.0f) { if (first < 0) first = m; last = m; }
        if (first < 0) first = last = 0;
        if (last - first > 1) return MIXFE_OK;                       // not a bank of adjacent triangles: K6 serves the plan
        float4 e;
        e.x = __int_as_float_host(first);
        e.y = pl->fb[(size_t)k * nm + first];
        e.z = last > first ? pl->fb[(size_t)k * nm + last] : 0.0f;
        e.w = 0.0f;
        binw[k] = e;
    }
    std::vector<float2> w(513);
    for (int k = 0; k <= 512; ++k) {
        const double a = -2.0 * M_PI * (double)k / 2048.0;
        w[k] = make_float2((float)std::cos(a), (float)std::sin(a));
    }
    w[512] = make_float2(0.0f, -1.0f);
    pl->x_off_tw = kBwWarps * kWarpRegion;
    pl->x_off_binw = pl->x_off_tw + 31 * 32 * 8;
    pl->x_off_w = pl->x_off_binw + nf * 16;
    pl->x_off_dmel = align16(pl->x_off_w + 513 * 8);
    pl->x_off_ola = align16(pl->x_off_dmel + kBwWarps * (nm + 1) * 4);
    pl->smem_x = pl->x_off_ola + kBwSpan * 8;
    if (pl->smem_x > 227 * 1024) return MIXFE_OK;
    int rc;
    if ((rc = upload(&pl->d_ebinw, binw)) || (rc = upload(&pl->d_ew2048, w))) return rc;
    return smem_optin(k6e_logmel_bwd, pl->device, pl->smem_x);
}

int build_tables(mixfe_plan* pl) {
    int rc = build_generic_tables(pl);
    if (rc) return rc;
    if ((rc = build_k1e_tables(pl))) return rc;
    if ((rc = build_k6e_tables(pl))) return rc;
    const char* force = getenv("MIXFE_FORCE_GENERIC");
    pl->use_generic = true;
    if (pl->n_fft == kNfft && !(force && atoi(force))) {
        if (build_mel_tables(pl) == MIXFE_OK) pl->use_generic = false;
    }
    return MIXFE_OK;
}

int run_features(mixfe_plan* pl, const float* const stems[4], long long stride_b, long long stride_c, int B,
                 int T, float* features, float* logmel, long long lm_stride_b, float* melpow,
                 long long mp_stride_b, cudaStream_t stream) {
    if (!pl || !stems) return fail(MIXFE_EINVAL, "null plan or stems");
    if (!pl->n_fft) return fail(MIXFE_EINVAL, "augmentation-only plan (mixfe_plan_create_augment) has no feature tables");
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
    const size_t spec_bytes = (size_t)B * n_tiles * kNumStems * pl->n_mels_pad * sizeof(float2);
    const size_t td_bytes = (size_t)B * n_tiles * 2 * kTdN * sizeof(float);
    if (features) {
        int rc = ensure_workspace(pl, spec_bytes + td_bytes, stream);
        if (rc) return rc;
    }
    K1Params p{};
    for (int s = 0; s < 4; ++s) p.stem[s] = stems[s];
    p.stride_b = stride_b; p.stride_c = stride_c;
    p.B = B; p.T = T; p.F = F; p.hop = pl->hop; p.n_mels = pl->n_mels; p.n_mels_pad = pl->n_mels_pad;
    p.n_tiles = n_tiles; p.nxs = pl->nxs; p.total_items = (long long)B * n_tiles;
    p.logmel = logmel; p.lm_stride_b = lm_stride_b; p.melpow = melpow; p.mp_stride_b = mp_stride_b;
    p.part_spec = reinterpret_cast<float2*>(pl->ws);
    p.part_td = reinterpret_cast<float*>(reinterpret_cast<char*>(pl->ws) + spec_bytes);
    p.tw = pl->d_tw; p.wsteps = pl->d_wsteps; p.flo = pl->d_flo; p.n_wt = (pl->off_flo - pl->off_wsteps) / 4;
    memcpy(p.wmeta, pl->wmeta, sizeof(p.wmeta));
    p.want_stats = features ? 1 : 0;
    p.off_ex = pl->off_ex; p.off_tw = pl->off_tw; p.off_s = pl->off_s; p.off_wsteps = pl->off_wsteps;
    p.off_flo = pl->off_flo; p.off_red = pl->off_red;
    {
        // TMA bulk copies need 16-byte aligned global addresses and sizes.  Every row start is
        // stem + b * stride_b + c * stride_c + tile * kTF * hop (- 512 for the frame staging): with 16-byte aligned
        // stem pointers and strides that are multiples of 4 floats all of them are aligned at once (kTF * hop * 4 is a
        // multiple of 32 bytes), so the kernel only has to compare the tile index with the range that lies inside the
        // clip.
        bool aligned = (stride_b % 4 == 0 || B == 1) && stride_c % 4 == 0;
        for (int s = 0; s < 4; ++s) aligned = aligned && (reinterpret_cast<uintptr_t>(stems[s]) & 15u) == 0;
        const long long tile_span = (long long)kTF * pl->hop;
        p.tma_lo = 1; p.tma_hi = 0; p.rows_hi = -1;
        if (aligned && pl->nxs % 4 == 0) {
            p.tma_lo = (int)((kNfft / 2 + tile_span - 1) / tile_span);                 // tile * span - 512 >= 0
            const long long room = (long long)T + kNfft / 2 - pl->nxs;                 // tile * span - 512 + nxs <= T
            p.tma_hi = room < 0 ? -1 : (int)std::min<long long>(room / tile_span, n_tiles - 1);
        }
        if (aligned && pl->hop % 16 == 0 && tile_span * 4 <= kWarpRegion)
            p.rows_hi = (int)std::min<long long>((long long)T / tile_span - 1, n_tiles - 1);   // (tile + 1) * span <= T
    }
    const char* force_g = getenv("MIXFE_FORCE_GENERIC");
    if (pl->d_eitems && !melpow && (logmel || (features && pl->hop % 8 == 0)) && (!features || pl->hop % 8 == 0) &&
        !(force_g && atoi(force_g))) {
        // K1E: n_fft = 2048 on the one-warp 1024-point transform (one warp per (frame, channel)); log-mel and, for hops
        // that are multiples of 8, K1's statistics for the mixing features
        K1EParams e{};
        for (int s = 0; s < 4; ++s) e.stem[s] = stems[s];
        e.stride_b = stride_b; e.stride_c = stride_c; e.B = B; e.T = T; e.F = F; e.hop = pl->hop; e.n_mels = pl->n_mels;
        e.n_tiles = (F + kETF - 1) / kETF; e.nxs = pl->e_nxs; e.total_items = (long long)B * e.n_tiles;
        e.logmel = logmel; e.lm_stride_b = lm_stride_b;
        e.tw = pl->d_tw; e.items = pl->d_eitems; e.wts = pl->d_ewts; e.n_wts = pl->e_nwts;
        e.off_ex = pl->e_off_ex; e.off_tw = pl->e_off_tw; e.off_items = pl->e_off_items; e.off_wts = pl->e_off_wts;
        e.off_bar = pl->e_off_bar;
        bool aligned = (stride_b % 4 == 0 || B == 1) && stride_c % 4 == 0;
        for (int s = 0; s < 4; ++s) aligned = aligned && (reinterpret_cast<uintptr_t>(stems[s]) & 15u) == 0;
        const long long span = (long long)kETF * pl->hop;
        const size_t e_spec_bytes = (size_t)B * 2 * e.n_tiles * kNumStems * pl->n_mels_pad * sizeof(float2);
        const size_t e_td_bytes = (size_t)B * e.n_tiles * 2 * kTdN * sizeof(float);
        e.want_stats = features ? 1 : 0; e.n_mels_pad = pl->n_mels_pad; e.rows_hi = -1;
        const int e_grid = (int)std::min<long long>(e.total_items, (long long)pl->e_ctas);
        if (features) {
            const size_t s_bytes = (size_t)e_grid * 3 * kEThreads * sizeof(float2);
            int rc = ensure_workspace(pl, e_spec_bytes + e_td_bytes + s_bytes, stream);
            if (rc) return rc;
            e.part_spec = reinterpret_cast<float2*>(pl->ws);
            e.part_td = reinterpret_cast<float*>(reinterpret_cast<char*>(pl->ws) + e_spec_bytes);
            e.s_ws = reinterpret_cast<float2*>(reinterpret_cast<char*>(pl->ws) + e_spec_bytes + e_td_bytes);
            if (aligned && pl->hop % 32 == 0 && span * 4 <= kWarpRegion)
                e.rows_hi = (int)std::min<long long>((long long)T / span - 1, e.n_tiles - 1);      // (tile + 1) * span <= T
        }
        e.tma_lo = 1; e.tma_hi = 0;
        if (aligned) {
            e.tma_lo = (int)((1024 + span - 1) / span);                            // tile * span - 1024 >= 0
            const long long room = (long long)T + 1024 - pl->e_nxs;                // tile * span - 1024 + nxs <= T
            e.tma_hi = room < 0 ? -1 : (int)std::min<long long>(room / span, e.n_tiles - 1);
        }
        {
            KernelTimer timer(pl, MIXFE_KERNEL_FEATURES, stream);
            const int grid = e_grid;
            static const bool dynamic = !(std::getenv("MIXFE_K1_DYNAMIC") && std::atoi(std::getenv("MIXFE_K1_DYNAMIC")) == 0);
            e.work = dynamic ? pl->d_k1_work : nullptr;
            if (features) k1e_logmel2048<true><<<grid, kEThreads, pl->smem_e, stream>>>(e);
            else k1e_logmel2048<false><<<grid, kEThreads, pl->smem_e, stream>>>(e);
        }
        MIXFE_CUDA(cudaGetLastError());
        launch_counter()++;
        if (features) {
            K2Params q{};
            q.part_spec = e.part_spec; q.part_td = e.part_td; q.features = features;
            q.B = B; q.T = T; q.F = F; q.n_mels = pl->n_mels; q.n_mels_pad = pl->n_mels_pad; q.n_tiles = e.n_tiles;
            q.n_spec_tiles = 2 * e.n_tiles;
            q.n_bins = pl->n_bins; q.D = pl->D; q.mask_per_warp = 1;
            {
                KernelTimer timer(pl, MIXFE_KERNEL_FINALIZE, stream);
                k2_finalize<<<B, kK2Threads, 0, stream>>>(q);
            }
            MIXFE_CUDA(cudaGetLastError());
            launch_counter()++;
        }
        return MIXFE_OK;
    }
    if (pl->use_generic) {
        K1GParams g{};
        g.base = p;
        g.n_fft = pl->n_fft; g.log2n = pl->g_log2n; g.nf = pl->g_nf;
        g.twg = pl->d_twg; g.win = pl->d_win;
        g.mel_meta = pl->d_mel_meta; g.mel_w = pl->d_mel_w;
        g.pw_stride = pl->g_pw_stride; g.off_pw = pl->g_off_pw; g.off_mel = pl->g_off_mel; g.off_S = pl->g_off_S;
        g.off_red = pl->g_off_red;
        if (B > 65535) return fail(MIXFE_EUNSUPPORTED, "K1G: B=%d > 65535 clips per call", B);
        KernelTimer timer(pl, MIXFE_KERNEL_FEATURES, stream);
        k1g_features<<<dim3(n_tiles, B), kGThreads, pl->smem_g, stream>>>(g);
    } else {
        KernelTimer timer(pl, MIXFE_KERNEL_FEATURES, stream);
        const int per_sm = pl->n_sms ? pl->k1_ctas / pl->n_sms : 1;
        const int grid = (int)std::min<long long>(p.total_items, (long long)per_sm * (pl->n_sms - pl->reserved_sms));
        static const bool dynamic = !(std::getenv("MIXFE_K1_DYNAMIC") && std::atoi(std::getenv("MIXFE_K1_DYNAMIC")) == 0);
        p.work = dynamic ? pl->d_k1_work : nullptr;
        k1_features<<<grid, kK1Threads, pl->smem_k1, stream>>>(p);
    }
    MIXFE_CUDA(cudaGetLastError());
    launch_counter()++;
    if (features) {
        K2Params q{};
        q.part_spec = p.part_spec; q.part_td = p.part_td; q.features = features;
        q.B = B; q.T = T; q.F = F; q.n_mels = pl->n_mels; q.n_mels_pad = pl->n_mels_pad; q.n_tiles = n_tiles;
        q.n_spec_tiles = n_tiles;
        q.n_bins = pl->n_bins; q.D = pl->D; q.mask_per_warp = pl->use_generic ? 0 : 1;
        {
            KernelTimer timer(pl, MIXFE_KERNEL_FINALIZE, stream);
            k2_finalize<<<B, kK2Threads, 0, stream>>>(q);
        }
        MIXFE_CUDA(cudaGetLastError());
        launch_counter()++;
    }
    return MIXFE_OK;
}

int run_logmel_backward(mixfe_plan* pl, const float* const stems[4], long long stride_b, long long stride_c, int B, int T,
                        const float* logmel, long long lm_stride_b, const float* grad_logmel, long long g_stride_b,
                        float* const grad_stems[4], long long ostride_b, long long ostride_c, cudaStream_t stream) {
    if (!pl || !stems || !grad_stems || !grad_logmel) return fail(MIXFE_EINVAL, "null plan, stems or gradients");
    if (!pl->n_fft) return fail(MIXFE_EINVAL, "augmentation-only plan (mixfe_plan_create_augment) has no feature tables");
    for (int s = 0; s < 4; ++s)
        if (!stems[s] || !grad_stems[s]) return fail(MIXFE_EINVAL, "stems[%d] or grad_stems[%d] is NULL", s, s);
    if (B < 0 || T < 0) return fail(MIXFE_EINVAL, "negative B or T");
    if (B == 0) return MIXFE_OK;
    if (T <= pl->n_fft / 2)
        return fail(MIXFE_ESHORT, "Argument #4: Padding size should be less than the corresponding input dimension, "
                                  "but got: padding (%d, %d) at dimension 2 of input of length %d",
                    pl->n_fft / 2, pl->n_fft / 2, T);
    if (stride_c < T || (B > 1 && stride_b < T) || ostride_c < T || (B > 1 && ostride_b < T))
        return fail(MIXFE_EINVAL, "strides smaller than T");
    if (B > 32767) return fail(MIXFE_EUNSUPPORTED, "log-mel backward: B=%d > 32767 clips per call", B);
    DeviceGuard guard(pl->device);
    const int F = 1 + T / pl->hop;
    const char* force = getenv("MIXFE_FORCE_GENERIC_BWD");
    if (logmel && pl->d_binw && !pl->use_generic && !(force && atoi(force))) {
        // K6F: every sample written once; only the reflect-fold strips at the row ends are zeroed and take atomics
        K6FParams p{};
        for (int s = 0; s < 4; ++s) { p.stem[s] = stems[s]; p.out[s] = grad_stems[s]; }
        p.stride_b = stride_b; p.stride_c = stride_c; p.ostride_b = ostride_b; p.ostride_c = ostride_c;
        p.g = grad_logmel; p.g_stride_b = g_stride_b; p.lm = logmel; p.lm_stride_b = lm_stride_b;
        p.B = B; p.T = T; p.F = F; p.n_mels = pl->n_mels;
        const int hlast = (T + 511) / 256;                           // last hop region the padded signal reaches
        p.n_tiles = (hlast + 3 + kBwOwn - 1) / kBwOwn;
        p.tw = pl->d_tw; p.binw = pl->d_binw;
        p.off_ex = pl->f_off_ex; p.off_tw = pl->f_off_tw; p.off_binw = pl->f_off_binw; p.off_dmel = pl->f_off_dmel;
        p.off_ola = pl->f_off_ola;
        {
            KernelTimer timer(pl, MIXFE_KERNEL_LOGMEL_BWD, stream);
            k6f_zero_edges<<<dim3(2 * B, kNumStems), 128, 0, stream>>>(p);
            k6f_logmel_bwd<<<dim3(p.n_tiles, B, kNumStems), kBwThreads, pl->smem_f, stream>>>(p);
        }
        MIXFE_CUDA(cudaGetLastError());
        launch_counter() += 2;
        return MIXFE_OK;
    }
    if (logmel && pl->d_ebinw && !(force && atoi(force))) {
        // K6E: the same scheme for (2048, 512); a CTA serves one (clip, stem, channel)
        K6EParams p{};
        for (int s = 0; s < 4; ++s) { p.stem[s] = stems[s]; p.out[s] = grad_stems[s]; }
        p.stride_b = stride_b; p.stride_c = stride_c; p.ostride_b = ostride_b; p.ostride_c = ostride_c;
        p.g = grad_logmel; p.g_stride_b = g_stride_b; p.lm = logmel; p.lm_stride_b = lm_stride_b;
        p.B = B; p.T = T; p.F = F; p.n_mels = pl->n_mels;
        const int hlast = ((T + 1) / 2 + 511) / 256;                 // last hop region (units of 2 samples) the padded signal reaches
        p.n_tiles = (hlast + 3 + kBwOwn - 1) / kBwOwn;
        p.tw = pl->d_tw; p.binw = pl->d_ebinw; p.w2048 = pl->d_ew2048;
        p.off_ex = 0; p.off_tw = pl->x_off_tw; p.off_binw = pl->x_off_binw; p.off_w = pl->x_off_w;
        p.off_dmel = pl->x_off_dmel; p.off_ola = pl->x_off_ola;
        {
            KernelTimer timer(pl, MIXFE_KERNEL_LOGMEL_BWD, stream);
            k6e_zero_edges<<<dim3(2 * B, kNumStems), 128, 0, stream>>>(p);
            k6e_logmel_bwd<<<dim3(p.n_tiles, B, 2 * kNumStems), kBwThreads, pl->smem_x, stream>>>(p);
        }
        MIXFE_CUDA(cudaGetLastError());
        launch_counter() += 2;
        return MIXFE_OK;
    }
    if (!pl->smem_b) return fail(MIXFE_EUNSUPPORTED, "log-mel backward: hop=%d needs more shared memory than an SM has", pl->hop);
    // generic path: overlap-add across tiles with atomicAdd into zeroed gradients
    for (int s = 0; s < 4; ++s) {
        if (B == 1 || ostride_b == 2 * ostride_c) {            // rows evenly spaced: one 2-D memset per stem
            MIXFE_CUDA(cudaMemset2DAsync(grad_stems[s], sizeof(float) * ostride_c, 0, sizeof(float) * T, 2 * (size_t)B, stream));
        } else {
            for (int b = 0; b < B; ++b)
                MIXFE_CUDA(cudaMemset2DAsync(grad_stems[s] + (long long)b * ostride_b, sizeof(float) * ostride_c, 0,
                                             sizeof(float) * T, 2, stream));
        }
    }
    K6Params p{};
    for (int s = 0; s < 4; ++s) { p.stem[s] = stems[s]; p.out[s] = grad_stems[s]; }
    p.stride_b = stride_b; p.stride_c = stride_c; p.ostride_b = ostride_b; p.ostride_c = ostride_c;
    p.g = grad_logmel; p.g_stride_b = g_stride_b;
    p.B = B; p.T = T; p.F = F; p.hop = pl->hop; p.n_mels = pl->n_mels; p.n_mels_pad = pl->n_mels_pad;
    p.n_fft = pl->n_fft; p.log2n = pl->g_log2n; p.nf = pl->b_nf; p.span = pl->b_span; p.span_pad = pl->b_span_pad;
    p.twg = pl->d_twg; p.win = pl->d_win;
    p.mel_meta = pl->d_mel_meta; p.mel_w = pl->d_mel_w;
    p.bin_meta = pl->d_bin_meta; p.bin_w = pl->d_bin_w;
    p.pw_stride = pl->g_pw_stride;
    p.off_pw = pl->b_off_pw; p.off_mel = pl->b_off_mel; p.off_ola = pl->b_off_ola;
    const int n_tiles = (p.F + kTF - 1) / kTF;
    {
        KernelTimer timer(pl, MIXFE_KERNEL_LOGMEL_BWD, stream);
        k6_logmel_bwd<<<dim3(n_tiles, B, kNumStems), kGThreads, pl->smem_b, stream>>>(p);
    }
    MIXFE_CUDA(cudaGetLastError());
    launch_counter()++;
    return MIXFE_OK;
}

// BS.1770-4 K-weighting for an arbitrary sample rate: bilinear transform (with pre-warping) of the analog prototype
// behind the Recommendation's 48 kHz table; at 48 kHz this gives the table's coefficients (oracle/bs1770.py checks).
void k_weighting(double fs, double (&sos)[2][5]) {
    {
        const double f0 = 1681.974450955533, G = 3.999843853973347, Q = 0.7071752369554196;
        const double K = std::tan(M_PI * f0 / fs), Vh = std::pow(10.0, G / 20.0), Vb = std::pow(Vh, 0.4996667741545416);
        const double a0 = 1.0 + K / Q + K * K;
        sos[0][0] = (Vh + Vb * K / Q + K * K) / a0; sos[0][1] = 2.0 * (K * K - Vh) / a0; sos[0][2] = (Vh - Vb * K / Q + K * K) / a0;
        sos[0][3] = 2.0 * (K * K - 1.0) / a0; sos[0][4] = (1.0 - K / Q + K * K) / a0;
    }
    {
        const double f0 = 38.13547087602444, Q = 0.5003270373238773;
        const double K = std::tan(M_PI * f0 / fs), a0 = 1.0 + K / Q + K * K;
        sos[1][0] = 1.0; sos[1][1] = -2.0; sos[1][2] = 1.0;
        sos[1][3] = 2.0 * (K * K - 1.0) / a0; sos[1][4] = (1.0 - K / Q + K * K) / a0;
    }
}

// out: integrated loudness (k7_gate) or NULL; out_iqr: short-term-loudness IQR (k7_short_iqr) or NULL
int run_gated_loudness(mixfe_plan* pl, const float* const stems[4], long long stride_b, long long stride_c, int B, int T,
                       float* out, float* out_iqr, cudaStream_t stream) {
    if (!pl || !stems || (!out && !out_iqr)) return fail(MIXFE_EINVAL, "null plan, stems or output");
    if (!pl->n_fft) return fail(MIXFE_EINVAL, "augmentation-only plan (mixfe_plan_create_augment) has no feature tables");
    for (int s = 0; s < 4; ++s)
        if (!stems[s]) return fail(MIXFE_EINVAL, "stems[%d] is NULL", s);
    if (B < 0 || T < 0) return fail(MIXFE_EINVAL, "negative B or T");
    if (pl->sr % 10) return fail(MIXFE_EUNSUPPORTED, "gated loudness: sample rate %d is not a multiple of 10 Hz", pl->sr);
    const int sub = pl->sr / 10;
    if (T < 4 * sub) return fail(MIXFE_EINVAL, "Audio must have length greater than the block size (%d samples)", 4 * sub);
    if (out_iqr && T < kShortSubs * sub)
        return fail(MIXFE_EINVAL, "short-term loudness needs at least 3 s of audio (%d samples)", kShortSubs * sub);
    if (out_iqr && T / sub - (kShortSubs - 1) > kShortMaxWin)
        return fail(MIXFE_EUNSUPPORTED, "short-term loudness IQR: more than %d windows (clip longer than 6 min 52 s)", kShortMaxWin);
    if (B == 0) return MIXFE_OK;
    if (stride_c < T || (B > 1 && stride_b < T)) return fail(MIXFE_EINVAL, "strides smaller than T");
    if (B > 65535) return fail(MIXFE_EUNSUPPORTED, "gated loudness: B=%d > 65535 clips per call", B);
    DeviceGuard guard(pl->device);
    K7Params p{};
    for (int s = 0; s < 4; ++s) p.stem[s] = stems[s];
    p.stride_b = stride_b; p.stride_c = stride_c; p.B = B; p.T = T; p.sub = sub; p.n_sub = T / sub;
    p.spc = kLoudMaxSubs; p.n_chunks = (p.n_sub + p.spc - 1) / p.spc;
    k_weighting((double)pl->sr, p.sos);
    // warm-up: slowest pole of the two sections (the high-pass has a near-double real pole: n r^n, hence the margin)
    double r = 0.0;
    for (int i = 0; i < 2; ++i) {
        const double a1 = p.sos[i][3], a2 = p.sos[i][4], disc = a1 * a1 - 4.0 * a2;
        r = std::max(r, disc < 0.0 ? std::sqrt(a2) : (std::fabs(a1) + std::sqrt(disc)) / 2.0);
    }
    const double need = 1.25 * std::log(1e-13) / std::log(std::min(r, 0.999999));
    p.warm = ((int)std::ceil(need) + kAugBlock - 1) / kAugBlock * kAugBlock;
    const size_t bytes = sizeof(double) * (size_t)B * kLoudSignals * 2 * p.n_sub;
    int rc = ensure_workspace(pl, bytes, stream);
    if (rc) return rc;
    p.sub_energy = reinterpret_cast<double*>(pl->ws);
    p.out = out; p.out_iqr = out_iqr;
    {
        KernelTimer timer(pl, MIXFE_KERNEL_LOUDNESS, stream);
        k7_kweight_energy<<<dim3(p.n_chunks, 2, B), kLoudSignals * 32, 0, stream>>>(p);
        if (out) k7_gate<<<B, kLoudSignals * 32, 0, stream>>>(p);
        if (out_iqr) k7_short_iqr<<<dim3(kLoudSignals, B), kShortThreads, 0, stream>>>(p);
    }
    MIXFE_CUDA(cudaGetLastError());
    launch_counter() += 1 + (out ? 1 : 0) + (out_iqr ? 1 : 0);
    return MIXFE_OK;
}

int stft_loss_common(mixfe_plan* pl, const float* x, const float* y, int B, int C, int T, long long xs_b, long long xs_c,
                     long long ys_b, long long ys_c, K8Params& p) {
    if (!pl || !x || !y) return fail(MIXFE_EINVAL, "null plan, x or y");
    if (!pl->n_fft) return fail(MIXFE_EINVAL, "augmentation-only plan (mixfe_plan_create_augment) has no feature tables");
    if (B < 0 || C <= 0 || T < 0) return fail(MIXFE_EINVAL, "bad B, C or T");
    if (T <= pl->n_fft / 2)
        return fail(MIXFE_ESHORT, "Argument #4: Padding size should be less than the corresponding input dimension, "
                                  "but got: padding (%d, %d) at dimension 2 of input of length %d",
                    pl->n_fft / 2, pl->n_fft / 2, T);
    if (xs_c < T || ys_c < T) return fail(MIXFE_EINVAL, "channel strides smaller than T");
    if ((long long)B * C > 65535) return fail(MIXFE_EUNSUPPORTED, "STFT loss: B*C=%lld > 65535 rows per call", (long long)B * C);
    p = K8Params{};
    p.x = x; p.y = y; p.xs_b = xs_b; p.xs_c = xs_c; p.ys_b = ys_b; p.ys_c = ys_c;
    p.rows = B * C; p.C = C; p.T = T; p.F = 1 + T / pl->hop; p.hop = pl->hop; p.n_fft = pl->n_fft; p.log2n = pl->g_log2n;
    p.n_tiles = (p.F + kTF - 1) / kTF; p.span = (kTF - 1) * pl->hop + pl->n_fft;
    p.span_pad = p.span + (pl->g_log2n > 5 ? p.span >> (pl->g_log2n - 5) : 0) + 1;
    p.twg = pl->d_twg; p.win = pl->d_win;
    return MIXFE_OK;
}

// K8F serves the (1024, 256) resolution (MIXFE_FORCE_GENERIC_LOSS=1: the generic kernels, for A/B runs)
bool k8f_usable(const mixfe_plan* pl) {
    static const bool forced = std::getenv("MIXFE_FORCE_GENERIC_LOSS") && std::atoi(std::getenv("MIXFE_FORCE_GENERIC_LOSS"));
    return pl->n_fft == 1024 && pl->hop == 256 && pl->d_tw && !forced;
}
// K8H serves the (512, 128) resolution
bool k8h_usable(const mixfe_plan* pl) {
    static const bool forced = std::getenv("MIXFE_FORCE_GENERIC_LOSS") && std::atoi(std::getenv("MIXFE_FORCE_GENERIC_LOSS"));
    return pl->n_fft == 512 && pl->hop == 128 && pl->d_tw && !forced;
}
// K8E serves the (2048, 512) resolution
bool k8e_usable(const mixfe_plan* pl) {
    static const bool forced = std::getenv("MIXFE_FORCE_GENERIC_LOSS") && std::atoi(std::getenv("MIXFE_FORCE_GENERIC_LOSS"));
    return pl->n_fft == 2048 && pl->hop == 512 && pl->d_tw && !forced;
}

K8FParams k8f_params(const mixfe_plan* pl, const K8Params& p) {
    K8FParams f{};
    f.x = p.x; f.y = p.y; f.xs_b = p.xs_b; f.xs_c = p.xs_c; f.ys_b = p.ys_b; f.ys_c = p.ys_c;
    f.rows = p.rows; f.C = p.C; f.T = p.T; f.F = p.F; f.n_tiles = p.n_tiles; f.tw = pl->d_tw;
    return f;
}

int run_stft_loss_forward(mixfe_plan* pl, const float* x, const float* y, int B, int C, int T, long long xs_b, long long xs_c,
                          long long ys_b, long long ys_c, double* state, cudaStream_t stream) {
    K8Params p;
    int rc = stft_loss_common(pl, x, y, B, C, T, xs_b, xs_c, ys_b, ys_c, p);
    if (rc) return rc;
    if (!state) return fail(MIXFE_EINVAL, "state is NULL");
    if (B == 0) return fail(MIXFE_EINVAL, "empty batch");
    DeviceGuard guard(pl->device);
    if (k8h_usable(pl)) {
        K8FParams f = k8f_params(pl, p);
        f.off_tw = 4 * kHRegion; f.off_red = f.off_tw + 1024 * 8;
        const int smem_f = f.off_red + 128;
        if ((rc = smem_optin(k8h_partials, pl->device, smem_f))) return rc;
        rc = ensure_workspace(pl, sizeof(float) * 3 * (size_t)p.rows * p.n_tiles, stream);
        if (rc) return rc;
        f.part = p.part = reinterpret_cast<float*>(pl->ws);
        p.state = state;
        {
            KernelTimer timer(pl, MIXFE_KERNEL_STFT_LOSS, stream);
            k8h_partials<<<dim3(p.n_tiles, p.rows), 128, smem_f, stream>>>(f);
            k8_finish<<<1, 256, 0, stream>>>(p);
        }
        MIXFE_CUDA(cudaGetLastError());
        launch_counter() += 2;
        return MIXFE_OK;
    }
    if (k8f_usable(pl) || k8e_usable(pl)) {
        // K8F / K8E: the (1024, 256) / (2048, 512) resolution on the one-warp transform; same partial layout, same finish
        const bool e = k8e_usable(pl);
        K8FParams f = k8f_params(pl, p);
        f.off_tw = 8 * kWarpRegion; f.off_red = f.off_tw + 31 * 32 * 8;
        const int smem_f = f.off_red + 128 + (e ? 8 * 33 * 32 * 4 : 0);          // K8E: + the target magnitudes of the 8 frames
        if ((rc = e ? smem_optin(k8e_partials, pl->device, smem_f) : smem_optin(k8f_partials, pl->device, smem_f))) return rc;
        rc = ensure_workspace(pl, sizeof(float) * 3 * (size_t)p.rows * p.n_tiles, stream);
        if (rc) return rc;
        f.part = p.part = reinterpret_cast<float*>(pl->ws);
        p.state = state;
        {
            KernelTimer timer(pl, MIXFE_KERNEL_STFT_LOSS, stream);
            if (e) k8e_partials<<<dim3(p.n_tiles, p.rows), 256, smem_f, stream>>>(f);
            else k8f_partials<<<dim3(p.n_tiles, p.rows), 256, smem_f, stream>>>(f);
            k8_finish<<<1, 256, 0, stream>>>(p);
        }
        MIXFE_CUDA(cudaGetLastError());
        launch_counter() += 2;
        return MIXFE_OK;
    }
    int nf = kTF;
    while (nf > 1 && nf * fg_row(pl->n_fft) * 8 > 72 * 1024) nf >>= 1;
    p.nf = nf;
    p.off_red = nf * fg_row(pl->n_fft) * 8;
    const int smem = p.off_red + 8 * 3 * 4;
    if ((rc = smem_optin(k8_partials, pl->device, 100 * 1024)) || (rc = smem_optin(k8_backward, pl->device, 200 * 1024))) return rc;
    rc = ensure_workspace(pl, sizeof(float) * 3 * (size_t)p.rows * p.n_tiles, stream);
    if (rc) return rc;
    p.part = reinterpret_cast<float*>(pl->ws);
    p.state = state;
    {
        KernelTimer timer(pl, MIXFE_KERNEL_STFT_LOSS, stream);
        k8_partials<<<dim3(p.n_tiles, p.rows), kGThreads, smem, stream>>>(p);
        k8_finish<<<1, 256, 0, stream>>>(p);
    }
    MIXFE_CUDA(cudaGetLastError());
    launch_counter() += 2;
    return MIXFE_OK;
}

int run_stft_loss_backward(mixfe_plan* pl, const float* x, const float* y, int B, int C, int T, long long xs_b, long long xs_c,
                           long long ys_b, long long ys_c, const double* state, const float* gscale, float* gx,
                           long long gs_b, long long gs_c, cudaStream_t stream) {
    K8Params p;
    int rc = stft_loss_common(pl, x, y, B, C, T, xs_b, xs_c, ys_b, ys_c, p);
    if (rc) return rc;
    if (!state || !gscale || !gx) return fail(MIXFE_EINVAL, "state, gscale or gx is NULL");
    if (gs_c < T) return fail(MIXFE_EINVAL, "gradient channel stride smaller than T");
    if (B == 0) return MIXFE_OK;
    DeviceGuard guard(pl->device);
    if (k8f_usable(pl)) {
        K8FParams f = k8f_params(pl, p);
        const int hlast = (T + 511) / 256;                           // last hop region the padded signal reaches (K6F's tiling)
        f.n_tiles = (hlast + 3 + kBwOwn - 1) / kBwOwn;
        f.off_tw = kBwWarps * kWarpRegion; f.off_ola = f.off_tw + 31 * 32 * 8;
        const int smem_f = f.off_ola + kBwSpan * 4;
        if ((rc = smem_optin(k8f_backward, pl->device, smem_f))) return rc;
        f.state = state; f.gscale = gscale; f.gx = gx; f.gs_b = gs_b; f.gs_c = gs_c;
        {
            KernelTimer timer(pl, MIXFE_KERNEL_STFT_LOSS, stream);
            k8f_backward<<<dim3(f.n_tiles, p.rows), kBwThreads, smem_f, stream>>>(f);
        }
        MIXFE_CUDA(cudaGetLastError());
        launch_counter()++;
        return MIXFE_OK;
    }
    if (k8h_usable(pl)) {
        K8FParams f = k8f_params(pl, p);
        const int hlast = (T + 255) / 128;                           // last hop region the padded signal reaches
        f.n_tiles = (hlast + 3 + kHOwn - 1) / kHOwn;
        f.off_tw = 16 * kHRegion; f.off_ola = f.off_tw + 1024 * 8;
        const int smem_f = f.off_ola + kHSpan * 4;
        if ((rc = smem_optin(k8h_backward, pl->device, smem_f))) return rc;
        f.state = state; f.gscale = gscale; f.gx = gx; f.gs_b = gs_b; f.gs_c = gs_c;
        {
            KernelTimer timer(pl, MIXFE_KERNEL_STFT_LOSS, stream);
            k8h_backward<<<dim3(f.n_tiles, p.rows), 512, smem_f, stream>>>(f);
        }
        MIXFE_CUDA(cudaGetLastError());
        launch_counter()++;
        return MIXFE_OK;
    }
    if (k8e_usable(pl)) {
        K8FParams f = k8f_params(pl, p);
        const int hlast = ((T + 1) / 2 + 511) / 256;                 // K6E's tiling (units of two samples)
        f.n_tiles = (hlast + 3 + kBwOwn - 1) / kBwOwn;
        f.off_tw = kBwWarps * kWarpRegion; f.off_ola = f.off_tw + 31 * 32 * 8;
        const int smem_f = f.off_ola + kBwSpan * 8;
        if ((rc = smem_optin(k8e_backward, pl->device, smem_f))) return rc;
        f.state = state; f.gscale = gscale; f.gx = gx; f.gs_b = gs_b; f.gs_c = gs_c;
        {
            KernelTimer timer(pl, MIXFE_KERNEL_STFT_LOSS, stream);
            k8e_backward<<<dim3(f.n_tiles, p.rows), kBwThreads, smem_f, stream>>>(f);
        }
        MIXFE_CUDA(cudaGetLastError());
        launch_counter()++;
        return MIXFE_OK;
    }
    const int per = fg_row(pl->n_fft) * 8, ola = align16(p.span_pad * 4);
    if (per + ola > 200 * 1024) return fail(MIXFE_EUNSUPPORTED, "STFT loss backward: hop=%d needs too much shared memory", pl->hop);
    int nf = kTF;
    while (nf > 1 && nf * per + ola > 64 * 1024) nf >>= 1;
    p.nf = nf;
    p.off_ola = align16(nf * fg_row(pl->n_fft) * 8);
    const int smem = p.off_ola + ola;
    if ((rc = smem_optin(k8_partials, pl->device, 100 * 1024)) || (rc = smem_optin(k8_backward, pl->device, 200 * 1024))) return rc;
    p.state = const_cast<double*>(state); p.gscale = gscale; p.gx = gx; p.gs_b = gs_b; p.gs_c = gs_c;
    {
        KernelTimer timer(pl, MIXFE_KERNEL_STFT_LOSS, stream);
        k8_backward<<<dim3(p.n_tiles, p.rows), kGThreads, smem, stream>>>(p);
    }
    MIXFE_CUDA(cudaGetLastError());
    launch_counter()++;
    return MIXFE_OK;
}

// EXTENSION: activity- and band-weighted masking from the log-mel tensor (k9_extensions.cuh)
int run_weighted_masking(mixfe_plan* pl, const float* logmel, long long lm_stride_b, int B, int T, const float* band_weights,
                         double band_weight_sum, float beta, float tau, float* out, cudaStream_t stream) {
    if (!pl || !logmel || !band_weights || !out) return fail(MIXFE_EINVAL, "null plan, log-mel, band weights or output");
    if (!pl->n_fft) return fail(MIXFE_EINVAL, "augmentation-only plan (mixfe_plan_create_augment) has no feature tables");
    if (B < 0 || T < 0) return fail(MIXFE_EINVAL, "negative B or T");
    if (!(tau > 0.0f) || !(band_weight_sum > 0.0)) return fail(MIXFE_EINVAL, "tau and the band-weight sum must be positive");
    if (B == 0) return MIXFE_OK;
    if (B > 65535) return fail(MIXFE_EUNSUPPORTED, "weighted masking: B=%d > 65535 clips per call", B);
    DeviceGuard guard(pl->device);
    K9Params p{};
    p.logmel = logmel; p.lm_stride_b = lm_stride_b; p.B = B; p.F = 1 + T / pl->hop; p.n_mels = pl->n_mels;
    p.n_blocks = (p.F + kMaskThreads - 1) / kMaskThreads;
    p.beta = beta; p.inv_tau = 1.0f / tau;
    // E_0: mel energy of a -60 dBFS white-noise frame = sigma^2 * sum(hann^2) * n_freq, sum(hann^2) = 3 n_fft / 8
    p.e0 = 1e-6f * (3.0f * (float)pl->n_fft / 8.0f) * (float)pl->n_freq;
    p.aw = band_weights; p.aw_sum = band_weight_sum; p.out = out;
    int rc = ensure_workspace(pl, sizeof(float) * 8 * (size_t)B * p.n_blocks, stream);
    if (rc) return rc;
    p.part = reinterpret_cast<float*>(pl->ws);
    {
        KernelTimer timer(pl, MIXFE_KERNEL_EXTENSION, stream);
        k9_masking_partials<<<dim3(p.n_blocks, B), kMaskThreads, 0, stream>>>(p);
        k9_masking_finish<<<B, 32, 0, stream>>>(p);
    }
    MIXFE_CUDA(cudaGetLastError());
    launch_counter() += 2;
    return MIXFE_OK;
}

// EXTENSION: envelope-follower compressor on one tensor (k9_extensions.cuh, K10)
int run_envelope_compress(mixfe_plan* pl, const float* in, float* out, long long stride_b, long long stride_c, int B, int C,
                          int T, float threshold_db, float ratio, float attack_coeff, float release_coeff, cudaStream_t stream) {
    if (!pl || !in || !out) return fail(MIXFE_EINVAL, "null argument");
    if (in == out) return fail(MIXFE_EINVAL, "out must not alias in");
    if (B < 0 || C <= 0 || T <= 0) return fail(MIXFE_EINVAL, "bad B, C or T");
    if (!(ratio > 0.0f) || !std::isfinite(threshold_db)) return fail(MIXFE_EINVAL, "threshold %g dB, ratio %g", (double)threshold_db, (double)ratio);
    if (!(attack_coeff >= 0.0f && attack_coeff < 1.0f && release_coeff >= 0.0f && release_coeff < 1.0f))
        return fail(MIXFE_EINVAL, "attack / release coefficients must lie in [0, 1)");
    if (B == 0) return MIXFE_OK;
    if (B > 65535 || C > 65535) return fail(MIXFE_EINVAL, "B or C > 65535 per call");
    if ((C > 1 && stride_c < T) || (B > 1 && stride_b < T)) return fail(MIXFE_EINVAL, "strides smaller than T");
    DeviceGuard guard(pl->device);
    K10Params p{};
    p.in = in; p.out = out; p.stride_b = stride_b; p.stride_c = stride_c; p.B = B; p.C = C; p.T = T;
    p.a_rel = release_coeff; p.a_att = attack_coeff;
    float c1;
    compressor_constants(threshold_db, ratio, p.knee, p.c0, c1);
    p.c1m1 = c1 - 1.0f;                                  // gain = compressed magnitude / a
    {
        KernelTimer timer(pl, MIXFE_KERNEL_EXTENSION, stream);
        k10_envelope_compress<<<dim3(C, B), kEnvWarps * 32, 0, stream>>>(p);
    }
    MIXFE_CUDA(cudaGetLastError());
    launch_counter()++;
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
    if (n_fft < 64 || n_fft > 4096 || (n_fft & (n_fft - 1)))
        return fail(MIXFE_EUNSUPPORTED, "n_fft=%d: the STFT kernels cover powers of two in [64, 4096]", n_fft);
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
    pl->n_sms = prop.multiProcessorCount;
    pl->device = device; pl->sr = sample_rate; pl->n_fft = n_fft; pl->hop = hop; pl->n_mels = n_mels;
    pl->n_mels_pad = (n_mels + 15) & ~15; pl->n_bins = n_spectral_bins; pl->n_freq = n_fft / 2 + 1;
    const int sd = n_spectral_bins > 0 ? n_spectral_bins + 2 : 5;
    pl->D = 4 * (6 + sd + 3) + 4 + 4;
    pl->fb = default_filterbank(sample_rate, n_fft, n_mels);
    int rc = upload_fft_twiddles(pl);
    if (!rc) rc = build_tables(pl);
    if (!rc) rc = noalloc_setup(device);                 // address of the static scratch of mixfe_weighted_masking_noalloc
    if (rc) { mixfe_plan_destroy(pl); return rc; }
    *out = pl;
    return MIXFE_OK;
}

int mixfe_plan_create_augment(mixfe_plan** out, int device, int sample_rate) {
    if (!out) return fail(MIXFE_EINVAL, "out is NULL");
    *out = nullptr;
    if (sample_rate <= 0) return fail(MIXFE_EINVAL, "bad plan arguments");
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
    pl->n_sms = prop.multiProcessorCount;
    pl->device = device; pl->sr = sample_rate;              // n_fft = 0 marks the plan as augmentation-only
    const int rc = upload_fft_twiddles(pl);
    if (rc) { mixfe_plan_destroy(pl); return rc; }
    *out = pl;
    return MIXFE_OK;
}

int mixfe_plan_destroy(mixfe_plan* pl) {
    if (!pl) return MIXFE_OK;
    DeviceGuard guard(pl->device);
    cudaFree(pl->d_tw); cudaFree(pl->d_twfine); cudaFree(pl->d_twlin); cudaFree(pl->d_flo); cudaFree(pl->d_wsteps);
    cudaFree(pl->ws); cudaFree(pl->d_k1_work); cudaFree(pl->d_ebinw); cudaFree(pl->d_ew2048);
    cudaFree(pl->d_twg); cudaFree(pl->d_win); cudaFree(pl->d_mel_meta);
    cudaFree(pl->d_mel_w); cudaFree(pl->d_bin_meta); cudaFree(pl->d_bin_w); cudaFree(pl->d_binw); cudaFree(pl->d_eitems); cudaFree(pl->d_ewts);
    for (int i = 0; i < mixfe_plan::kParamSlots; ++i) {
        if (pl->param_done[i]) { cudaEventSynchronize(pl->param_done[i]); cudaEventDestroy(pl->param_done[i]); }
        if (pl->param_host[i]) cudaFreeHost(pl->param_host[i]);
    }
    delete pl;
    return MIXFE_OK;
}

int mixfe_plan_set_filterbank(mixfe_plan* pl, const float* fb_host) {
    if (!pl || !fb_host) return fail(MIXFE_EINVAL, "null plan or filterbank");
    if (!pl->n_fft) return fail(MIXFE_EINVAL, "augmentation-only plan (mixfe_plan_create_augment) has no feature tables");
    DeviceGuard guard(pl->device);
    MIXFE_CUDA(cudaDeviceSynchronize());            // tables may be in use by queued launches
    pl->fb.assign(fb_host, fb_host + (size_t)pl->n_freq * pl->n_mels);
    return build_tables(pl);
}

int mixfe_copy_async(void* dst, const void* src, size_t bytes, void* cuda_stream) {
    if (!dst || !src) return fail(MIXFE_EINVAL, "null dst or src");
    MIXFE_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(cuda_stream)));
    return MIXFE_OK;
}

int mixfe_copy_strided_async(void* dst, size_t dst_pitch, size_t dst_block_rows, const void* src, size_t src_pitch,
                             size_t src_block_rows, size_t width_bytes, size_t rows, size_t blocks, void* cuda_stream) {
    if (!dst || !src) return fail(MIXFE_EINVAL, "null dst or src");
    if (rows == 0 || width_bytes == 0 || blocks == 0) return MIXFE_OK;
    if (dst_pitch < width_bytes || src_pitch < width_bytes) return fail(MIXFE_EINVAL, "pitch smaller than the row width");
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    if (blocks == 1) {
        MIXFE_CUDA(cudaMemcpy2DAsync(dst, dst_pitch, src, src_pitch, width_bytes, rows, cudaMemcpyDefault, stream));
        return MIXFE_OK;
    }
    if (dst_block_rows < rows || src_block_rows < rows) return fail(MIXFE_EINVAL, "block stride smaller than the rows of a block");
    cudaMemcpy3DParms q{};
    q.srcPtr = make_cudaPitchedPtr(const_cast<void*>(src), src_pitch, width_bytes, src_block_rows);
    q.dstPtr = make_cudaPitchedPtr(dst, dst_pitch, width_bytes, dst_block_rows);
    q.extent = make_cudaExtent(width_bytes, rows, blocks);
    q.kind = cudaMemcpyDefault;
    MIXFE_CUDA(cudaMemcpy3DAsync(&q, stream));
    return MIXFE_OK;
}

int mixfe_push_peers(void* const* dst, int n_peers, const void* src, size_t bytes, int n_ctas, void* cuda_stream) {
    if (!dst || !src || n_peers < 0 || n_peers > kMaxPushPeers) return fail(MIXFE_EINVAL, "bad peers (at most %d)", kMaxPushPeers);
    if (n_peers == 0 || bytes == 0) return MIXFE_OK;
    if (n_ctas <= 0 || n_ctas > 1024) return fail(MIXFE_EINVAL, "bad CTA count");
    PushParams p{};
    p.src = src; p.bytes = bytes; p.n_peers = n_peers;
    bool vec = (reinterpret_cast<uintptr_t>(src) & 15u) == 0 && bytes % 16 == 0;
    for (int i = 0; i < n_peers; ++i) {
        if (!dst[i]) return fail(MIXFE_EINVAL, "peer pointer %d is NULL", i);
        p.dst[i] = dst[i];
        vec = vec && (reinterpret_cast<uintptr_t>(dst[i]) & 15u) == 0;
    }
    if (!vec) return fail(MIXFE_EINVAL, "push: pointers and size must be multiples of 16 bytes");
    k11_push_peers<<<n_ctas, kPushThreads, 0, static_cast<cudaStream_t>(cuda_stream)>>>(p);
    MIXFE_CUDA(cudaGetLastError());
    launch_counter()++;
    return MIXFE_OK;
}

int mixfe_plan_reserve_sms(mixfe_plan* pl, int n_sms) {
    if (!pl || n_sms < 0) return fail(MIXFE_EINVAL, "bad plan or SM count");
    pl->reserved_sms = std::min(n_sms, std::max(0, pl->n_sms - 1));
    return MIXFE_OK;
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

int mixfe_num_frames(const mixfe_plan* pl, int T) {
    return pl && pl->hop ? 1 + T / pl->hop : fail(MIXFE_EINVAL, "null or augmentation-only plan");
}

size_t mixfe_workspace_bytes(const mixfe_plan* pl) { return pl ? pl->ws_bytes : 0; }

int mixfe_debug_feature_kernel_config(const mixfe_plan* pl, int* ctas_per_sm, int* smem_bytes, int* mel_steps) {
    if (!pl) return fail(MIXFE_EINVAL, "null plan");
    if (pl->use_generic) {                      // K1G: one CTA per (clip, tile); ctas_per_sm = 0 marks the generic path
        if (ctas_per_sm) *ctas_per_sm = 0;
        if (smem_bytes) *smem_bytes = pl->smem_g;
        if (mel_steps) *mel_steps = 0;
        return MIXFE_OK;
    }
    if (ctas_per_sm) *ctas_per_sm = pl->n_sms ? pl->k1_ctas / pl->n_sms : 0;
    if (smem_bytes) *smem_bytes = pl->smem_k1;
    if (mel_steps) *mel_steps = pl->n_wsteps;
    return MIXFE_OK;
}

int mixfe_features_logmel(mixfe_plan* pl, const float* const stems[MIXFE_NUM_STEMS], long long stride_b,
                          long long stride_c, int B, int T, float* features, float* logmel,
                          long long logmel_stride_b, void* cuda_stream) {
    return run_features(pl, stems, stride_b, stride_c, B, T, features, logmel, logmel_stride_b, nullptr, 0,
                        static_cast<cudaStream_t>(cuda_stream));
}

int mixfe_logmel_backward(mixfe_plan* pl, const float* const stems[MIXFE_NUM_STEMS], long long stride_b,
                          long long stride_c, int B, int T, const float* logmel, long long logmel_stride_b,
                          const float* grad_logmel, long long grad_logmel_stride_b,
                          float* const grad_stems[MIXFE_NUM_STEMS], long long grad_stride_b, long long grad_stride_c,
                          void* cuda_stream) {
    return run_logmel_backward(pl, stems, stride_b, stride_c, B, T, logmel, logmel_stride_b, grad_logmel,
                               grad_logmel_stride_b, grad_stems, grad_stride_b, grad_stride_c,
                               static_cast<cudaStream_t>(cuda_stream));
}

int mixfe_gated_loudness(mixfe_plan* pl, const float* const stems[MIXFE_NUM_STEMS], long long stride_b, long long stride_c,
                         int B, int T, float* lufs, void* cuda_stream) {
    return run_gated_loudness(pl, stems, stride_b, stride_c, B, T, lufs, nullptr, static_cast<cudaStream_t>(cuda_stream));
}

int mixfe_short_term_loudness_iqr(mixfe_plan* pl, const float* const stems[MIXFE_NUM_STEMS], long long stride_b,
                                  long long stride_c, int B, int T, float* iqr, void* cuda_stream) {
    return run_gated_loudness(pl, stems, stride_b, stride_c, B, T, nullptr, iqr, static_cast<cudaStream_t>(cuda_stream));
}

int mixfe_stft_loss_forward(mixfe_plan* pl, const float* x, const float* y, int B, int C, int T, long long x_stride_b,
                            long long x_stride_c, long long y_stride_b, long long y_stride_c, double* state, void* cuda_stream) {
    return run_stft_loss_forward(pl, x, y, B, C, T, x_stride_b, x_stride_c, y_stride_b, y_stride_c, state,
                                 static_cast<cudaStream_t>(cuda_stream));
}

int mixfe_stft_loss_backward(mixfe_plan* pl, const float* x, const float* y, int B, int C, int T, long long x_stride_b,
                             long long x_stride_c, long long y_stride_b, long long y_stride_c, const double* state,
                             const float* grad_scale, float* grad_x, long long g_stride_b, long long g_stride_c,
                             void* cuda_stream) {
    return run_stft_loss_backward(pl, x, y, B, C, T, x_stride_b, x_stride_c, y_stride_b, y_stride_c, state, grad_scale, grad_x,
                                  g_stride_b, g_stride_c, static_cast<cudaStream_t>(cuda_stream));
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

int mixfe_weighted_masking(mixfe_plan* pl, const float* logmel, long long logmel_stride_b, int B, int T,
                           const float* band_weights, double band_weight_sum, float beta, float tau, float* out,
                           void* cuda_stream) {
    return run_weighted_masking(pl, logmel, logmel_stride_b, B, T, band_weights, band_weight_sum, beta, tau, out,
                                static_cast<cudaStream_t>(cuda_stream));
}

// mixfe_weighted_masking without any allocation: same two K9 launches, partial sums in g_noalloc_mask_part.  No
// cudaMalloc / cudaFree / event creation / workspace growth here or below; the plan is only read.
int mixfe_weighted_masking_noalloc(mixfe_plan* pl, const float* logmel, long long logmel_stride_b, int B, int T,
                                   const float* band_weights, double band_weight_sum, float beta, float tau, float* out,
                                   void* cuda_stream) {
    if (!pl || !logmel || !band_weights || !out) return fail(MIXFE_EINVAL, "null plan, log-mel, band weights or output");
    if (!pl->n_fft) return fail(MIXFE_EINVAL, "augmentation-only plan (mixfe_plan_create_augment) has no feature tables");
    if (B < 0 || T < 0) return fail(MIXFE_EINVAL, "negative B or T");
    if (!(tau > 0.0f) || !(band_weight_sum > 0.0)) return fail(MIXFE_EINVAL, "tau and the band-weight sum must be positive");
    if (B == 0) return MIXFE_OK;
    if (B > 65535) return fail(MIXFE_EUNSUPPORTED, "weighted masking: B=%d > 65535 clips per call", B);
    const int F = 1 + T / pl->hop;
    const int n_blocks = (F + kMaskThreads - 1) / kMaskThreads;
    if ((long long)B * n_blocks > kNoAllocMaskEntries)
        return fail(MIXFE_EUNSUPPORTED,
                    "mixfe_weighted_masking_noalloc: B=%d clips x %d blocks of %d frames = %lld partial sums exceed the static "
                    "scratch of %d (MIXFE_NOALLOC_MASKING_MAX_BLOCKS); split the batch or call mixfe_weighted_masking",
                    B, n_blocks, kMaskThreads, (long long)B * n_blocks, kNoAllocMaskEntries);
    float* part = nullptr;
    if (pl->device >= 0 && pl->device < kNoAllocMaxDevices) {
        std::lock_guard<std::mutex> lock(g_noalloc_mu);
        part = g_noalloc_part[pl->device];
    }
    if (!part) return fail(MIXFE_EUNSUPPORTED, "device %d has no static scratch (ordinal >= %d)", pl->device, kNoAllocMaxDevices);
    DeviceGuard guard(pl->device);
    K9Params p{};
    p.logmel = logmel; p.lm_stride_b = logmel_stride_b; p.B = B; p.F = F; p.n_mels = pl->n_mels; p.n_blocks = n_blocks;
    p.beta = beta; p.inv_tau = 1.0f / tau;
    p.e0 = 1e-6f * (3.0f * (float)pl->n_fft / 8.0f) * (float)pl->n_freq;          // as run_weighted_masking
    p.aw = band_weights; p.aw_sum = band_weight_sum; p.out = out;
    // every entry [0, B * n_blocks) is written by the first launch before the second reads it: nothing has to start at
    // zero, whatever an earlier call left behind
    p.part = part;
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    k9_masking_partials<<<dim3(n_blocks, B), kMaskThreads, 0, stream>>>(p);
    k9_masking_finish<<<B, 32, 0, stream>>>(p);
    MIXFE_CUDA(cudaGetLastError());
    launch_counter() += 2;
    return MIXFE_OK;
}

int mixfe_envelope_compress(mixfe_plan* pl, const float* in, float* out, long long stride_b, long long stride_c, int B, int C,
                            int T, float threshold_db, float ratio, float attack_coeff, float release_coeff, void* cuda_stream) {
    return run_envelope_compress(pl, in, out, stride_b, stride_c, B, C, T, threshold_db, ratio, attack_coeff, release_coeff,
                                 static_cast<cudaStream_t>(cuda_stream));
}

int mixfe_effects(mixfe_plan* pl, const float* in, float* out, long long stride_b, long long stride_c, int B, int C,
                  int T, const mixfe_aug_params* params, void* cuda_stream) {
    return run_effects(pl, in, out, stride_b, stride_c, B, C, T, params, static_cast<cudaStream_t>(cuda_stream));
}

int mixfe_reverb(mixfe_plan* pl, const float* in, float* out, long long stride_b, long long stride_c, int B, int C,
                 int T, const float* impulse, long long impulse_stride_b, int L, void* cuda_stream) {
    return run_reverb(pl, in, out, stride_b, stride_c, B, C, T, impulse, impulse_stride_b, L,
                      static_cast<cudaStream_t>(cuda_stream));
}

}  // extern "C"
