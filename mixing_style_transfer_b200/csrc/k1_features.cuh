// K1: fused STFT -> power -> sparse mel -> log, plus every statistic the 64-D mixing-feature
// vector needs, in ONE pass over the 8 channels of a clip.  K2: per-clip finalize.
//
// Reference behaviour being replaced (paths relative to the reference checkout):
//   src/model.py:41-67            MelSpectrogramPreprocessor.forward  (log-mel, 4 STFTs / clip)
//   src/mixing_utils.py:107-139   extract_dynamics   (rms, peak, crest, loudness)
//   src/mixing_utils.py:141-236   extract_spectral   (4 more STFTs / clip)
//   src/mixing_utils.py:238-268   extract_stereo     (ILD, L/R correlation, mid/side ratio)
//   src/mixing_utils.py:270-309   extract_masking    (4 more STFTs / clip)
//   src/mixing_utils.py:311-318   compute_loudness
//   src/mixing_utils.py:320-357   _flatten_features  (sorted-key layout, clamp, NaN scrub)
// The reference runs 12 STFTs per clip (each a (2,T) real FFT + dense 513x128 matmul); here each
// stem's L and R are packed into ONE complex 1024-point FFT per frame, shared by all consumers.
//
// Work decomposition.  grid = (frame tiles, clips); a CTA of 8 warps owns TF = 8 consecutive
// frames of one clip and walks the 4 stems.  Per stem:
//   stage   samples [t0*hop - 512, t0*hop + 7*hop + 512) of L and R, reflect-padded at the clip
//           edges, interleaved as float2 in shared memory (128-bit global loads in the interior)
//   FFT     one warp per frame: Hann window on the fly, 32 x 32 four-step FFT with the
//           transposition through a swizzled shared buffer, L/R untangled with warp shuffles,
//           power of bins 0..512 left in registers
//   P tile  powers written as rows of (2 bins x 16 columns), column = channel*8 + frame
//   mel     every quarter-warp owns one mel filter: LDS.128 of 2 bins x 2 columns, LDS.64 of the
//           two filter taps, 4 FMAs; result tile [mel][channel][frame] in shared memory
//   epilog  log(mel + 1e-10) -> coalesced global store; per-(stem, mel) sums of log and power for
//           the band / tilt / flatness features; S = (mel_L + mel_R)/2 kept for the masking term
// Time-domain statistics come from the tile's own hop region, read once for all 4 stems so the
// mixture never exists in memory.  Per-tile partials go to a workspace (no float atomics: results
// are run-to-run deterministic) and K2 reduces them in double precision.
#pragma once

#include "common.cuh"
#include "fft32.cuh"

namespace mixfe {

constexpr int kTF = 8;             // frames per CTA tile
constexpr int kK1Threads = 256;    // 8 warps, one frame each
constexpr int kNfft = 1024;
constexpr int kNfreq = 513;
constexpr int kPRS = 36;           // floats per P-tile row (32 used + pad to keep 16 B alignment)
constexpr int kPRows = 288;        // 257 rows of bin pairs + zeroed overshoot rows for ragged groups
constexpr int kTdN = 40;           // floats per tile of time-domain partials
constexpr int kMaxGroupsPerWarp = 8;

struct K1Params {
    const float* stem[kNumStems];
    long long stride_b, stride_c;
    int B, T, F, hop, n_mels, n_mels_pad, n_tiles, n_groups;
    int nxs;                         // staged samples per channel = (kTF-1)*hop + 1024
    float* logmel;
    long long lm_stride_b;
    float* melpow;                   // debug output (mel power before the log) or nullptr
    long long mp_stride_b;
    float2* part_spec;               // [B][n_tiles][4][n_mels_pad] (sum log, sum mel)
    float* part_td;                  // [B][n_tiles][kTdN]
    const float2* tw;                // [1024] tw[k1*32+n2] = exp(-2 pi i n2 k1 / 1024)
    const int* frow0;                // [n_mels_pad] first P-tile row of each filter
    const int2* gdesc;               // [n_groups] {rows to walk, offset into gw}
    const float2* gw;                // filter taps, [group][row][quarter] -> (w(even bin), w(odd bin))
    int nw2;
    const unsigned char* wtab;       // [8*kMaxGroupsPerWarp + 8] groups of each warp, then counts
    int want_stats;
    int off_ex, off_tw, off_s, off_frow0, off_gdesc, off_gw, off_wtab, off_red;
};

__device__ __forceinline__ float ln_fast(float x) {               // x >= 1e-10: never denormal
    float y;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y * 0.69314718055994531f;
}

__device__ __forceinline__ float fast_sigmoid_neg(float x) {     // 1 / (1 + exp(x))
    return __fdividef(1.0f, 1.0f + __expf(x));
}

__global__ void __launch_bounds__(kK1Threads, 2) k1_features(const __grid_constant__ K1Params p) {
    extern __shared__ __align__(16) unsigned char smem[];
    float2* const xs = reinterpret_cast<float2*>(smem);
    float* const meltile = reinterpret_cast<float*>(smem);                    // aliases xs (after the FFT phase)
    float* const logtile = meltile + p.n_mels_pad * 16;
    float2* const ex = reinterpret_cast<float2*>(smem + p.off_ex);
    float* const pt = reinterpret_cast<float*>(smem + p.off_ex);              // aliases ex (after the FFT phase)
    float2* const tw = reinterpret_cast<float2*>(smem + p.off_tw);
    float* const S = reinterpret_cast<float*>(smem + p.off_s);                // [3][n_mels_pad][8]
    int* const frow0 = reinterpret_cast<int*>(smem + p.off_frow0);
    int2* const gdesc = reinterpret_cast<int2*>(smem + p.off_gdesc);
    float2* const gw = reinterpret_cast<float2*>(smem + p.off_gw);
    unsigned char* const wtab = smem + p.off_wtab;
    float* const red = reinterpret_cast<float*>(smem + p.off_red);            // [8][4]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x, b = blockIdx.y;
    const int t0 = tile * kTF;

    // ---- constant tables -> shared memory ------------------------------------------------
    for (int i = tid; i < 1024; i += kK1Threads) tw[i] = p.tw[i];
    for (int i = tid; i < p.n_mels_pad; i += kK1Threads) frow0[i] = p.frow0[i];
    for (int i = tid; i < p.n_groups; i += kK1Threads) gdesc[i] = p.gdesc[i];
    for (int i = tid; i < p.nw2; i += kK1Threads) gw[i] = p.gw[i];
    if (tid < 8 * kMaxGroupsPerWarp + 8) wtab[tid] = p.wtab[tid];

    // ---- time-domain statistics over the hop region this tile owns --------------------------
    if (p.want_stats) {
        float acc[33];
#pragma unroll
        for (int i = 0; i < 33; ++i) acc[i] = 0.0f;
        const long long own0 = (long long)t0 * p.hop;
        long long rem = (long long)p.T - own0;
        const int own_n = rem < 0 ? 0 : (rem > (long long)kTF * p.hop ? kTF * p.hop : (int)rem);
        const long long cb = (long long)b * p.stride_b + own0;
        for (int i = tid; i < own_n; i += kK1Threads) {
            float mL = 0.0f, mR = 0.0f;
#pragma unroll
            for (int s = 0; s < kNumStems; ++s) {
                const float* g = p.stem[s] + cb + i;
                const float L = __ldg(g), R = __ldg(g + p.stride_c);
                acc[s * 8 + 0] += L;
                acc[s * 8 + 1] += R;
                acc[s * 8 + 2] = fmaf(L, L, acc[s * 8 + 2]);
                acc[s * 8 + 3] = fmaf(R, R, acc[s * 8 + 3]);
                const float d = L - R, m = L + R;
                acc[s * 8 + 4] = fmaf(d, d, acc[s * 8 + 4]);
                acc[s * 8 + 5] = fmaf(m, m, acc[s * 8 + 5]);
                acc[s * 8 + 6] = fmaxf(acc[s * 8 + 6], fabsf(L));
                acc[s * 8 + 7] = fmaxf(acc[s * 8 + 7], fabsf(R));
                mL += L;                                 // ((0 + v) + b) + d) + o, as sum(stems.values())
                mR += R;
            }
            acc[32] = fmaf(mL, mL, fmaf(mR, mR, acc[32]));
        }
        float* const scratch = reinterpret_cast<float*>(ex);       // [33][256], exchange region is idle
#pragma unroll
        for (int q = 0; q < 33; ++q) scratch[q * kK1Threads + tid] = acc[q];
        __syncthreads();
        for (int q = warp; q < 33; q += 8) {
            const bool is_max = q < 32 && (q & 7) >= 6;
            float v = is_max ? 0.0f : 0.0f;
            for (int j = 0; j < 8; ++j) {
                const float u = scratch[q * kK1Threads + lane + 32 * j];
                v = is_max ? fmaxf(v, u) : v + u;
            }
            v = is_max ? warp_max(v) : warp_sum(v);
            if (lane == 0) p.part_td[((long long)b * p.n_tiles + tile) * kTdN + q] = v;
        }
        // the next write into the exchange region happens after the post-staging __syncthreads
    }

    float msum[4] = {0.0f, 0.0f, 0.0f, 0.0f};      // masking partial sums of this thread

    for (int s = 0; s < kNumStems; ++s) {
        // ---- stage the samples of stem s -------------------------------------------------
        {
            const float* gL = p.stem[s] + (long long)b * p.stride_b;
            const float* gR = gL + p.stride_c;
            const long long base = (long long)t0 * p.hop - kNfft / 2;
            const bool interior = base >= 0 && base + p.nxs <= (long long)p.T && (p.nxs & 3) == 0 &&
                                  ((reinterpret_cast<unsigned long long>(gL + base) |
                                    reinterpret_cast<unsigned long long>(gR + base)) & 15ull) == 0;
            if (interior) {
                const float4* vL = reinterpret_cast<const float4*>(gL + base);
                const float4* vR = reinterpret_cast<const float4*>(gR + base);
                float4* dst = reinterpret_cast<float4*>(xs);
                for (int i = tid; i < (p.nxs >> 2); i += kK1Threads) {
                    const float4 a = __ldg(vL + i), c = __ldg(vR + i);
                    dst[2 * i] = make_float4(a.x, c.x, a.y, c.y);
                    dst[2 * i + 1] = make_float4(a.z, c.z, a.w, c.w);
                }
            } else {
                for (int i = tid; i < p.nxs; i += kK1Threads) {
                    long long j = base + i;
                    if (j < 0) j = -j;                               // torch 'reflect': -i -> i
                    if (j >= p.T) j = 2ll * (p.T - 1) - j;           // T-1+i -> T-1-i
                    j = j < 0 ? 0 : (j >= p.T ? p.T - 1 : j);        // frames past F only; keeps loads in bounds
                    xs[i] = make_float2(__ldg(gL + j), __ldg(gR + j));
                }
            }
        }
        __syncthreads();

        // ---- FFT of frame t0 + warp -------------------------------------------------------
        float PL[16], PR[16];
        float PnyL = 0.0f, PnyR = 0.0f;                 // bin 512 (lane 0 only)
        if (t0 + warp < p.F) {
            float re[32], im[32];
            {
                const float2 wl = tw[32 + lane];        // exp(-2 pi i lane/1024)
                const float cb = wl.x, sb = -wl.y;
                const float2* xp = xs + warp * p.hop + lane;
#pragma unroll
                for (int n1 = 0; n1 < 32; ++n1) {
                    // 0.5 * hann[32 n1 + lane]; cos(a + b) with a = 2 pi n1/32 (immediate), b per lane.
                    // The extra 0.5 makes |X +- conj(X')|^2 the channel powers without a final scale.
                    const float ca = n1 < 16 ? w32_re(n1) : -w32_re(n1 - 16);
                    const float sa = n1 < 16 ? w32_sin(n1) : -w32_sin(n1 - 16);
                    const float wv = fmaf(0.25f * sa, sb, fmaf(-0.25f * ca, cb, 0.25f));
                    const float2 v = xp[32 * n1];
                    re[n1] = v.x * wv;
                    im[n1] = v.y * wv;
                }
            }
            dft32_dif(re, im);
            float2* const exw = ex + warp * 1024;
#pragma unroll
            for (int k1 = 0; k1 < 32; ++k1) {
                const float yr = re[brev5(k1)], yi = im[brev5(k1)];
                float2 o;
                if (k1 == 0) {
                    o = make_float2(yr, yi);
                } else {
                    const float2 t = tw[k1 * 32 + lane];
                    o.x = fmaf(yr, t.x, -yi * t.y);
                    o.y = fmaf(yr, t.y, yi * t.x);
                }
                exw[k1 * 32 + (lane ^ k1)] = o;
            }
            __syncwarp();
#pragma unroll
            for (int n2 = 0; n2 < 32; ++n2) {
                const float2 v = exw[lane * 32 + (n2 ^ lane)];
                re[n2] = v.x;
                im[n2] = v.y;
            }
            dft32_dif(re, im);
            // lane holds X[lane + 32 k2] at index brev5(k2).  Untangle L (real part) and R (imag part):
            // |L_k|^2 = |X_k + conj(X_{N-k})|^2, |R_k|^2 = |X_k - conj(X_{N-k})|^2 (scale folded into the window).
            const int src = (32 - lane) & 31;
#pragma unroll
            for (int k2 = 0; k2 < 16; ++k2) {
                float qr = __shfl_sync(kFullMask, re[brev5(31 - k2)], src);
                float qi = __shfl_sync(kFullMask, im[brev5(31 - k2)], src);
                if (lane == 0) {                         // N - 32 k2 lives in lane 0 itself, register (32 - k2) & 31
                    qr = re[brev5((32 - k2) & 31)];
                    qi = im[brev5((32 - k2) & 31)];
                }
                const float xr = re[brev5(k2)], xi = im[brev5(k2)];
                const float a = xr + qr, bb = xi - qi, c = xr - qr, d = xi + qi;
                PL[k2] = fmaf(a, a, bb * bb);
                PR[k2] = fmaf(c, c, d * d);
            }
            {
                const float xr = re[brev5(16)], xi = im[brev5(16)];      // bin 512 is its own partner
                PnyL = 4.0f * xr * xr;
                PnyR = 4.0f * xi * xi;
            }
        } else {
#pragma unroll
            for (int k2 = 0; k2 < 16; ++k2) PL[k2] = PR[k2] = 0.0f;
        }
        __syncthreads();          // all warps are done with xs and with their exchange buffers

        // ---- P tile: row = bin >> 1, position = 2*column + (bin & 1), column = channel*8 + frame
        {
#pragma unroll
            for (int k2 = 0; k2 < 16; ++k2) {
                float* row = pt + ((lane >> 1) + 16 * k2) * kPRS + (lane & 1);
                row[2 * warp] = PL[k2];
                row[2 * (8 + warp)] = PR[k2];
            }
            if (lane == 0) {
                float* row = pt + 256 * kPRS;
                row[2 * warp] = PnyL;
                row[2 * (8 + warp)] = PnyR;
                row[2 * warp + 1] = 0.0f;               // "bin 513" slot
                row[2 * (8 + warp) + 1] = 0.0f;
            }
            for (int i = tid; i < (kPRows - 257) * kPRS; i += kK1Threads) pt[257 * kPRS + i] = 0.0f;
        }
        __syncthreads();

        // ---- sparse mel projection: one filter per quarter-warp ------------------------------
        {
            const int q = lane >> 3, j8 = lane & 7;
            const int ng = wtab[8 * kMaxGroupsPerWarp + warp];
            for (int gi = 0; gi < ng; ++gi) {
                const int g = wtab[warp * kMaxGroupsPerWarp + gi];
                const int2 gd = gdesc[g];
                const int m = 4 * g + q;
                const float* prow = pt + frow0[m] * kPRS + 4 * j8;
                const float2* wp = gw + gd.y + q;
                float a0 = 0.0f, a1 = 0.0f, c0 = 0.0f, c1 = 0.0f;
                for (int j = 0; j < gd.x; j += 2) {          // gd.x is even (host pads with zero taps)
                    const float4 v = *reinterpret_cast<const float4*>(prow + j * kPRS);
                    const float4 u = *reinterpret_cast<const float4*>(prow + (j + 1) * kPRS);
                    const float2 w = wp[4 * j];
                    const float2 x = wp[4 * j + 4];
                    a0 = fmaf(w.x, v.x, a0);
                    a1 = fmaf(w.x, v.z, a1);
                    c0 = fmaf(x.x, u.x, c0);
                    c1 = fmaf(x.x, u.z, c1);
                    a0 = fmaf(w.y, v.y, a0);
                    a1 = fmaf(w.y, v.w, a1);
                    c0 = fmaf(x.y, u.y, c0);
                    c1 = fmaf(x.y, u.w, c1);
                }
                a0 += c0;
                a1 += c1;
                *reinterpret_cast<float2*>(meltile + m * 16 + 2 * j8) = make_float2(a0, a1);
            }
        }
        __syncthreads();

        // ---- epilogue A: logs, spectral partial sums, S tiles / masking ----------------------
        const int nvalid = p.F - t0 < kTF ? p.F - t0 : kTF;          // frames of this tile that exist
        for (int r = tid; r < 2 * p.n_mels_pad; r += kK1Threads) {
            const int m = r >> 1, ch = r & 1;
            const float4 v0 = *reinterpret_cast<const float4*>(meltile + m * 16 + ch * 8);
            const float4 v1 = *reinterpret_cast<const float4*>(meltile + m * 16 + ch * 8 + 4);
            const float v[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
            float lg[8];
            float slog = 0.0f, smel = 0.0f;
#pragma unroll
            for (int t = 0; t < 8; ++t) lg[t] = ln_fast(v[t] + 1e-10f);
            if (nvalid == kTF) {
                slog = ((lg[0] + lg[1]) + (lg[2] + lg[3])) + ((lg[4] + lg[5]) + (lg[6] + lg[7]));
                smel = ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7]));
            } else {
#pragma unroll
                for (int t = 0; t < 8; ++t) {
                    if (t < nvalid) {
                        slog += lg[t];
                        smel += v[t];
                    }
                }
            }
            *reinterpret_cast<float4*>(logtile + m * 16 + ch * 8) = make_float4(lg[0], lg[1], lg[2], lg[3]);
            *reinterpret_cast<float4*>(logtile + m * 16 + ch * 8 + 4) = make_float4(lg[4], lg[5], lg[6], lg[7]);
            slog += __shfl_xor_sync(kFullMask, slog, 1);
            smel += __shfl_xor_sync(kFullMask, smel, 1);
            if (ch == 0 && p.want_stats)
                p.part_spec[(((long long)b * p.n_tiles + tile) * kNumStems + s) * p.n_mels_pad + m] =
                    make_float2(slog, smel);
        }
        if (p.want_stats) {
            for (int r = tid; r < 2 * p.n_mels_pad; r += kK1Threads) {
                const int m = r >> 1, h = r & 1;             // h selects frames 4h .. 4h+3
                const float4 vL = *reinterpret_cast<const float4*>(meltile + m * 16 + 4 * h);
                const float4 vR = *reinterpret_cast<const float4*>(meltile + m * 16 + 8 + 4 * h);
                const float4 Sc = make_float4(0.5f * (vL.x + vR.x), 0.5f * (vL.y + vR.y),
                                              0.5f * (vL.z + vR.z), 0.5f * (vL.w + vR.w));
                if (s < 3) {
                    *reinterpret_cast<float4*>(S + (s * p.n_mels_pad + m) * 8 + 4 * h) = Sc;
                } else if (m < p.n_mels) {
                    const float4 A0 = *reinterpret_cast<const float4*>(S + (0 * p.n_mels_pad + m) * 8 + 4 * h);
                    const float4 A1 = *reinterpret_cast<const float4*>(S + (1 * p.n_mels_pad + m) * 8 + 4 * h);
                    const float4 A2 = *reinterpret_cast<const float4*>(S + (2 * p.n_mels_pad + m) * 8 + 4 * h);
                    const float s0[4] = {A0.x, A0.y, A0.z, A0.w}, s1[4] = {A1.x, A1.y, A1.z, A1.w};
                    const float s2[4] = {A2.x, A2.y, A2.z, A2.w}, s3[4] = {Sc.x, Sc.y, Sc.z, Sc.w};
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        if (4 * h + i < nvalid) {
                            const float m01 = fmaxf(s0[i], s1[i]), m23 = fmaxf(s2[i], s3[i]);
                            // sigmoid(max_{j != i} S_j - S_i) = 1 / (1 + exp(S_i - max_others))
                            msum[0] += fast_sigmoid_neg(s0[i] - fmaxf(s1[i], m23));
                            msum[1] += fast_sigmoid_neg(s1[i] - fmaxf(s0[i], m23));
                            msum[2] += fast_sigmoid_neg(s2[i] - fmaxf(m01, s3[i]));
                            msum[3] += fast_sigmoid_neg(s3[i] - fmaxf(m01, s2[i]));
                        }
                    }
                }
            }
        }
        __syncthreads();

        // ---- epilogue B: coalesced store of the log-mel tile (32-byte runs along time) --------
        {
            const int col = tid & 15, ch = col >> 3, t = col & 7, m0 = tid >> 4;
            if (t < nvalid) {
                const long long o = ((long long)(2 * s + ch) * p.n_mels + m0) * p.F + t0 + t;
                const long long step = 16ll * p.F;
                if (p.logmel) {
                    float* dst = p.logmel + (long long)b * p.lm_stride_b + o;
                    const float* src = logtile + tid;
#pragma unroll 4
                    for (int m = m0; m < p.n_mels; m += 16, dst += step, src += kK1Threads) *dst = *src;
                }
                if (p.melpow) {
                    float* dst = p.melpow + (long long)b * p.mp_stride_b + o;
                    const float* src = meltile + tid;
                    for (int m = m0; m < p.n_mels; m += 16, dst += step, src += kK1Threads) *dst = *src;
                }
            }
        }
        __syncthreads();          // xs (= meltile/logtile) is restaged next
    }

    // ---- masking partial sums of the tile --------------------------------------------------
    if (p.want_stats) {
#pragma unroll
        for (int i = 0; i < 4; ++i) msum[i] = warp_sum(msum[i]);
        if (lane == 0) {
#pragma unroll
            for (int i = 0; i < 4; ++i) red[warp * 4 + i] = msum[i];
        }
        __syncthreads();
        if (tid < 4) {
            float v = 0.0f;
            for (int w = 0; w < 8; ++w) v += red[w * 4 + tid];
            p.part_td[((long long)b * p.n_tiles + tile) * kTdN + 33 + tid] = v;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// K2: reduce the per-tile partials of one clip and evaluate the feature formulas.
// ------------------------------------------------------------------------------------------------
struct K2Params {
    const float2* part_spec;
    const float* part_td;
    float* features;                 // [B][D]
    int B, T, F, n_mels, n_mels_pad, n_tiles;
    int n_bins;                      // 0: low/mid/high bands, else detailed curve of n_bins points
    int D;
};

__device__ __forceinline__ void put_feature(float* out, int idx, double v) {
    float f = (float)v;
    f = fminf(fmaxf(f, -100.0f), 100.0f);       // clamp (src/mixing_utils.py:339); fminf/fmaxf drop NaN ...
    if (v != v) f = 0.0f;                       // ... so scrub it explicitly (:343-348)
    out[idx] = f;
}

constexpr int kK2Threads = 512;
constexpr int kMaxMels = 128;

__global__ void __launch_bounds__(kK2Threads) k2_finalize(const K2Params p) {
    __shared__ double sE[kNumStems][kMaxMels];       // mean dB per mel band
    __shared__ double sCurve[kNumStems][kMaxMels];
    __shared__ double sLogTot[kNumStems][4], sMelTot[kNumStems][4];
    __shared__ double sTd[kTdN];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.x;
    const double kDb = 4.342944819032518;            // 10 / ln(10)

    // 1. spectral sums over tiles: thread = (stem, mel)
    {
        const int s = tid >> 7, m = tid & 127;
        double slog = 0.0, smel = 0.0;
        if (m < p.n_mels) {
            const float2* src = p.part_spec + ((long long)b * p.n_tiles * kNumStems + s) * p.n_mels_pad + m;
            const long long step = (long long)kNumStems * p.n_mels_pad;
#pragma unroll 8
            for (int t = 0; t < p.n_tiles; ++t) {
                const float2 v = __ldg(src + t * step);
                slog += (double)v.x;
                smel += (double)v.y;
            }
            sE[s][m] = kDb * slog / (2.0 * p.F);
        } else if (m < kMaxMels) {
            sE[s][m] = 0.0;
        }
        const double a = warp_sum(slog), c = warp_sum(smel);
        if (lane == 0) {
            sLogTot[s][warp & 3] = a;
            sMelTot[s][warp & 3] = c;
        }
    }
    // 2. time-domain partials: 8 threads per quantity
    if (tid < kTdN * 8) {
        const int q = tid >> 3, sub = tid & 7;
        const bool is_max = q < 32 && (q & 7) >= 6;
        double v = 0.0;
        for (int t = sub; t < p.n_tiles; t += 8) {
            const double u = (double)__ldg(p.part_td + ((long long)b * p.n_tiles + t) * kTdN + q);
            v = is_max ? fmax(v, u) : v + u;
        }
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) {
            const double u = __shfl_xor_sync(kFullMask, v, o);
            v = is_max ? fmax(v, u) : v + u;
        }
        if (sub == 0) sTd[q] = v;
    }
    __syncthreads();

    const int sd = p.n_bins > 0 ? p.n_bins + 2 : 5;
    const int per = 6 + 1 + sd + 3;
    float* out = p.features + (long long)b * p.D;
    const double N = (double)p.T;

    // 3. spectral block: warp s handles stem s
    if (warp < kNumStems) {
        const int s = warp;
        // sorted key order: bass, drums, (masking), other, vocals   (src/mixing_utils.py:324)
        const int base = s == 0 ? 3 * per + 4 : (s == 1 ? 0 : (s == 2 ? per : 2 * per + 4));
        const int ob = base + 7;
        const int n = p.n_mels;
        const double* E = sE[s];
        int nc;                                      // length of the curve the tilt is computed on
        const double* C;
        if (p.n_bins > 0 && p.n_bins < n) {          // F.interpolate(mode='linear', align_corners=True), :212-215
            nc = p.n_bins;
            const float scale = nc > 1 ? (float)(n - 1) / (float)(nc - 1) : 0.0f;
            for (int i = lane; i < nc; i += 32) {
                const float src = scale * (float)i;
                const int i0 = (int)src;
                const int i1 = i0 + (i0 < n - 1 ? 1 : 0);
                const double l1 = (double)(src - (float)i0);
                sCurve[s][i] = (1.0 - l1) * E[i0] + l1 * E[i1];
            }
            __syncwarp();
            C = sCurve[s];
        } else {
            nc = n;
            C = E;
        }
        if (p.n_bins == 0) {
            const int lo = n / 4, hi = n / 4 * 3;    // :172-173
            double a = 0.0, c = 0.0, d = 0.0;
            for (int i = lane; i < n; i += 32) {
                const double e = E[i];
                if (i < lo) a += e;
                else if (i < hi) c += e;
                else d += e;
            }
            a = warp_sum(a); c = warp_sum(c); d = warp_sum(d);
            if (lane == 0) {
                put_feature(out, ob + 0, a / lo);
                put_feature(out, ob + 1, c / (hi - lo));
                put_feature(out, ob + 2, d / (n - hi));
            }
        } else {
            for (int i = lane; i < nc; i += 32) put_feature(out, ob + i, C[i]);
        }
        // tilt: 0 when the unbiased std of the curve is < 1e-6, else Pearson r against arange (:185-190)
        double sum = 0.0;
        for (int i = lane; i < nc; i += 32) sum += C[i];
        const double mean = warp_sum(sum) / nc;
        const double imean = 0.5 * (nc - 1);
        double m2 = 0.0, cov = 0.0, iv = 0.0;
        for (int i = lane; i < nc; i += 32) {
            const double dc = C[i] - mean, di = (double)i - imean;
            m2 += dc * dc;
            cov += dc * di;
            iv += di * di;
        }
        m2 = warp_sum(m2); cov = warp_sum(cov); iv = warp_sum(iv);
        if (lane == 0) {
            const double sdev = nc > 1 ? sqrt(m2 / (nc - 1)) : 0.0;
            double tilt = 0.0;
            if (!(sdev < 1e-6)) {
                tilt = cov / sqrt(iv * m2);
                tilt = fmin(fmax(tilt, -1.0), 1.0);           // torch.corrcoef clips to [-1, 1]
            }
            const int ot = ob + (p.n_bins > 0 ? p.n_bins : 3);
            put_feature(out, ot, tilt);
            const double cnt = 2.0 * (double)p.F * (double)n;
            const double lt = sLogTot[s][0] + sLogTot[s][1] + sLogTot[s][2] + sLogTot[s][3];
            const double mt = sMelTot[s][0] + sMelTot[s][1] + sMelTot[s][2] + sMelTot[s][3];
            put_feature(out, ot + 1, exp(lt / cnt) / (mt / cnt + 1e-10));     // flatness, :195
        }
    }
    // 4. dynamics / relative loudness / stereo: one thread per stem; masking: 4 threads
    if (tid >= 128 && tid < 128 + kNumStems) {
        const int s = tid - 128;
        const int base = s == 0 ? 3 * per + 4 : (s == 1 ? 0 : (s == 2 ? per : 2 * per + 4));
        const double* q = sTd + s * 8;
        const double sL = q[0], sR = q[1], sLL = q[2], sRR = q[3], sDD = q[4], sSS = q[5], pkL = q[6], pkR = q[7];
        const double rmsL = sqrt(sLL / N), rmsR = sqrt(sRR / N);
        const double loud = -0.691 + 10.0 * log10((sLL + sRR) / (2.0 * N) + 1e-10);       // :311-318
        const double mixl = -0.691 + 10.0 * log10(sTd[32] / (2.0 * N) + 1e-10);
        put_feature(out, base + 0, rmsL);
        put_feature(out, base + 1, rmsR);
        put_feature(out, base + 2, 20.0 * log10(pkL / (rmsL + 1e-8)));                      // :127
        put_feature(out, base + 3, 20.0 * log10(pkR / (rmsR + 1e-8)));
        put_feature(out, base + 4, loud);
        put_feature(out, base + 5, loud);
        put_feature(out, base + 6, loud - mixl);                                            // :94-97
        const int os = base + 7 + sd;
        put_feature(out, os + 0, 20.0 * log10(rmsL / (rmsR + 1e-8)));                       // ILD, :246-248
        const double sLR = 0.25 * (sSS - sDD);
        const double cov = sLR - sL * sR / N, vL = sLL - sL * sL / N, vR = sRR - sR * sR / N;
        put_feature(out, os + 1, cov / (sqrt(fmax(vL, 0.0) * fmax(vR, 0.0)) + 1e-8));       // :251-256
        put_feature(out, os + 2, (sDD / (4.0 * N)) / (sSS / (4.0 * N) + 1e-8));             // :259-263
    }
    if (tid >= 160 && tid < 160 + kNumStems) {
        const int i = tid - 160;
        put_feature(out, 2 * per + i, sTd[33 + i] / ((double)p.n_mels * (double)p.F));      // :296-306
    }
}

}  // namespace mixfe
