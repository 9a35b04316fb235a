// K1G: the same fused STFT -> power -> mel -> log + feature statistics pass as K1 (k1_features.cuh), for the
// configurations K1's one-warp-per-frame 1024-point design does not cover: any power-of-two n_fft in
// [64, 4096], any hop, n_mels <= 128.  The reference's evaluation and style-transfer callers build
// MixingFeatureExtractor(44100, 2048, 512, 80) (inference/validate_retrieval.py:355,
// src/train_style_transfer.py:523-533); K1 stays the tuned path of the training configuration
// (1024, 256, 128).
//
// Reference behaviour: identical to K1's list (src/model.py:41-67, src/mixing_utils.py:107-357).
//
// Work decomposition.  One CTA (256 threads) per (clip, tile of 8 frames), walking the 4 stems.  Per stem the
// frames of the tile are transformed `nf` at a time (as many as fit in shared memory):
//   load    (L, R) sample pairs with torch 'reflect' indexing, times 0.5 * hann, as complex values in smem
//   FFT     in-place radix-4 decimation-in-frequency passes over shared memory (one radix-2 pass at the end
//           when log2(n_fft) is odd); X[k] ends at the bit-reversed position
//   power   L/R untangled from X[k] and X[n-k]; |L_k|^2, |R_k|^2 into a (frame, bin) column
//   mel     thread (frame, filter) walks the filter's contiguous support (CSR table in global memory)
// then, per stem, the same epilogue as K1: log(mel + 1e-10), per-(stem, mel) partial sums, the masking S tile,
// 32-byte-run stores of the log tile.  The per-tile partials have K1's layout, so K2 finalizes both.
#pragma once

#include "k1_features.cuh"

namespace mixfe {

constexpr int kGThreads = 256;

struct K1GParams {
    K1Params base;                   // stems, strides, B, T, F, hop, n_mels(_pad), n_tiles, outputs, partials
    int n_fft, log2n, nf;            // nf: frames transformed at once
    const float2* twg;               // [n_fft/2] exp(-2 pi i e / n_fft)
    const float* win;                // [n_fft] 0.5 * periodic hann
    const int* mel_lo;               // [n_mels] first bin of the filter's support
    const int* mel_len;              // [n_mels] bins in the support
    const int* mel_off;              // [n_mels] offset of the filter's weights in mel_w
    const float* mel_w;
    int pw_stride;                   // float2 entries per frame of the padded power column (odd)
    int off_pw, off_mel, off_S, off_red;
};

// idx / d for 0 <= idx and idx * d < 2^32, d >= 2, with magic = 0xFFFFFFFF / d + 1 (one multiply-high instead of the
// ~20-instruction runtime division).
__device__ __forceinline__ unsigned k1g_magic(int d) { return 0xFFFFFFFFu / (unsigned)d + 1u; }
__device__ __forceinline__ int k1g_div(int idx, unsigned magic) { return (int)__umulhi((unsigned)idx, magic); }

__device__ __forceinline__ float2 cmulf(float2 a, float2 w) {
    return make_float2(fmaf(a.x, w.x, -a.y * w.y), fmaf(a.x, w.y, a.y * w.x));
}

// One radix-4 DIF pass = the two radix-2 stages with half-sizes h and h/2, on `frames` rows of n points.
// n and h are powers of two: all index arithmetic is shifts and masks.
__device__ __forceinline__ void k1g_pass4(float2* z, const float2* __restrict__ tw, int n, int frames, int h, int tid) {
    const int hh = h >> 1, lq = 31 - __clz(n) - 2, lhh = 31 - __clz(hh), ls = lq + 1 - lhh - 1;   // s = n / (2 h) = 2^ls
    const int total = frames << lq;
    for (int idx = tid; idx < total; idx += kGThreads) {
        const int f = idx >> lq, q = idx & ((1 << lq) - 1);
        const int g = q >> lhh, j = q & (hh - 1);
        float2* r = z + ((size_t)f << (lq + 2)) + ((g << lhh) << 2) + j;         // f * n + g * 2h + j, 2h = 4 hh
        const float2 a = r[0], b = r[hh], c = r[h], d = r[h + hh];
        const float2 w1 = __ldg(tw + (j << ls)), w2 = __ldg(tw + (j << (ls + 1)));
        const float2 a1 = make_float2(a.x + c.x, a.y + c.y);
        const float2 b1 = make_float2(b.x + d.x, b.y + d.y);
        const float2 c1 = cmulf(make_float2(a.x - c.x, a.y - c.y), w1);
        const float2 t = cmulf(make_float2(b.x - d.x, b.y - d.y), w1);
        const float2 d1 = make_float2(t.y, -t.x);                    // times exp(-2 pi i (n/4) / n) = -i
        r[0] = make_float2(a1.x + b1.x, a1.y + b1.y);
        r[hh] = cmulf(make_float2(a1.x - b1.x, a1.y - b1.y), w2);
        r[h] = make_float2(c1.x + d1.x, c1.y + d1.y);
        r[h + hh] = cmulf(make_float2(c1.x - d1.x, c1.y - d1.y), w2);
    }
}

// Last stage (h = 1) when log2(n) is odd: neighbours, unit twiddle.
__device__ __forceinline__ void k1g_pass2_last(float2* z, int n, int frames, int tid) {
    float4* z4 = reinterpret_cast<float4*>(z);
    for (int idx = tid; idx < frames * (n >> 1); idx += kGThreads) {
        const float4 v = z4[idx];
        z4[idx] = make_float4(v.x + v.z, v.y + v.w, v.x - v.z, v.y - v.w);
    }
}

__global__ void __launch_bounds__(kGThreads) k1g_features(const __grid_constant__ K1GParams g) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const K1Params& p = g.base;
    float2* const z = reinterpret_cast<float2*>(smem);                         // [nf][n_fft]
    float2* const pw = reinterpret_cast<float2*>(smem + g.off_pw);             // [nf][pw_stride]
    float2* const melb = reinterpret_cast<float2*>(smem + g.off_mel);          // [8][n_mels_pad] (L, R)
    float* const S = reinterpret_cast<float*>(smem + g.off_S);                 // [4][8][n_mels_pad]
    float* const red = reinterpret_cast<float*>(smem + g.off_red);             // [8][4]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x, b = blockIdx.y;
    const int n = g.n_fft, half = n >> 1;
    const int t0 = tile * kTF;
    const int nvalid = p.F - t0 < kTF ? p.F - t0 : kTF;
    const unsigned mg_nb = k1g_magic(half + 1), mg_mp = k1g_magic(p.n_mels_pad);

    if (p.want_stats) {
        k1_time_stats(p, b, tile, tid);
    }

    for (int s = 0; s < kNumStems; ++s) {
        const float* gL = p.stem[s] + (long long)b * p.stride_b;
        const float* gR = gL + p.stride_c;
        for (int f0 = 0; f0 < kTF; f0 += g.nf) {
            const int frames = nvalid - f0 < g.nf ? (nvalid - f0 > 0 ? nvalid - f0 : 0) : g.nf;
            // ---- load + window ------------------------------------------------------------------
            for (int idx = tid; idx < frames * n; idx += kGThreads) {
                const int f = idx >> g.log2n, i = idx & (n - 1);
                long long j = (long long)(t0 + f0 + f) * p.hop - half + i;
                if (j < 0) j = -j;                               // torch 'reflect'
                if (j >= p.T) j = 2ll * (p.T - 1) - j;
                const float w = __ldg(g.win + i);
                z[idx] = make_float2(__ldg(gL + j) * w, __ldg(gR + j) * w);
            }
            __syncthreads();
            // ---- FFT ----------------------------------------------------------------------------
            int h = half;
            for (; h >= 2; h >>= 2) {
                k1g_pass4(z, g.twg, n, frames, h, tid);
                __syncthreads();
            }
            if (h == 1) {
                k1g_pass2_last(z, n, frames, tid);
                __syncthreads();
            }
            // ---- untangle L / R, power ------------------------------------------------------------
            // Lane i takes bin k = bitrev_{log2n-1}(i): X[k] then sits at position 2 i and X[n-k] at an odd position
            // walking the other way, so a warp reads neighbouring words instead of 32 words of one bank.  The power
            // column is padded (k + (k >> pad_sh)) so that the scattered k of a warp fall into different banks.
            const int sh = 32 - g.log2n, pad_sh = g.log2n > 6 ? g.log2n - 6 : 31;
            for (int idx = tid; idx < frames * (half + 1); idx += kGThreads) {
                const int f = k1g_div(idx, mg_nb), i = idx - f * (half + 1);
                const int k = i == half ? half : (int)(__brev((unsigned)i) >> (sh + 1));
                const float2 x = z[(size_t)f * n + (__brev((unsigned)k) >> sh)];
                const float2 y = z[(size_t)f * n + (__brev((unsigned)((n - k) & (n - 1))) >> sh)];
                const float a = x.x + y.x, bb = x.y - y.y, c = x.x - y.x, d = x.y + y.y;
                pw[(size_t)f * g.pw_stride + k + (k >> pad_sh)] = make_float2(fmaf(a, a, bb * bb), fmaf(d, d, c * c));
            }
            __syncthreads();
            // ---- mel projection -------------------------------------------------------------------
            for (int idx = tid; idx < g.nf * p.n_mels_pad; idx += kGThreads) {
                const int f = k1g_div(idx, mg_mp), m = idx - f * p.n_mels_pad;
                float aL = 0.0f, aR = 0.0f;
                if (f < frames && m < p.n_mels) {
                    const int lo = __ldg(g.mel_lo + m), len = __ldg(g.mel_len + m);
                    const float* w = g.mel_w + __ldg(g.mel_off + m);
                    const float2* src = pw + (size_t)f * g.pw_stride;
                    for (int i = 0; i < len; ++i) {
                        const float wi = __ldg(w + i);
                        const float2 v = src[lo + i + ((lo + i) >> pad_sh)];
                        aL = fmaf(wi, v.x, aL);
                        aR = fmaf(wi, v.y, aR);
                    }
                }
                melb[(f0 + f) * p.n_mels_pad + m] = make_float2(aL, aR);
            }
            __syncthreads();
        }

        // ---- epilogue: logs, partial sums, S tile, stores (idx = (m, ch, t), t fastest) ---------------
        const long long tile_id = (long long)b * p.n_tiles + tile;
        for (int idx = tid; idx < p.n_mels_pad * 16; idx += kGThreads) {
            const int t = idx & 7, ch = (idx >> 3) & 1, m = idx >> 4;
            const float2 lr = melb[t * p.n_mels_pad + m];
            const float v = ch ? lr.y : lr.x;
            const float lg = ln_fast(v + 1e-10f);
            const bool ok = t < nvalid && m < p.n_mels;
            if (ok) {
                const long long o = ((long long)(2 * s + ch) * p.n_mels + m) * p.F + t0 + t;
                if (p.logmel) p.logmel[(long long)b * p.lm_stride_b + o] = lg;
                if (p.melpow) p.melpow[(long long)b * p.mp_stride_b + o] = v;
            }
            if (p.want_stats) {
                float slog = t < nvalid ? lg : 0.0f, smel = t < nvalid ? v : 0.0f;
#pragma unroll
                for (int o = 1; o < 16; o <<= 1) {
                    slog += __shfl_xor_sync(kFullMask, slog, o);
                    smel += __shfl_xor_sync(kFullMask, smel, o);
                }
                if ((idx & 15) == 0) p.part_spec[(tile_id * kNumStems + s) * p.n_mels_pad + m] = make_float2(slog, smel);
                if (ch == 0) S[(s * kTF + t) * p.n_mels_pad + m] = 0.5f * (lr.x + lr.y);
            }
        }
        __syncthreads();
    }

    // ---- masking: mean over (mel, frame) of sigmoid(max_{j != i} S_j - S_i)  (src/mixing_utils.py:296-306) ----
    if (p.want_stats) {
        float msum[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        for (int idx = tid; idx < kTF * p.n_mels_pad; idx += kGThreads) {
            const int t = k1g_div(idx, mg_mp), m = idx - t * p.n_mels_pad;
            if (t < nvalid && m < p.n_mels) {
                const float s0 = S[idx], s1 = S[kTF * p.n_mels_pad + idx], s2 = S[2 * kTF * p.n_mels_pad + idx],
                            s3 = S[3 * kTF * p.n_mels_pad + idx];
                const float m01 = fmaxf(s0, s1), m23 = fmaxf(s2, s3);
                msum[0] += fast_sigmoid_neg(s0 - fmaxf(s1, m23));
                msum[1] += fast_sigmoid_neg(s1 - fmaxf(s0, m23));
                msum[2] += fast_sigmoid_neg(s2 - fmaxf(m01, s3));
                msum[3] += fast_sigmoid_neg(s3 - fmaxf(m01, s2));
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) msum[i] = warp_sum(msum[i]);
        if (lane == 0) {
#pragma unroll
            for (int i = 0; i < 4; ++i) red[warp * 4 + i] = msum[i];
        }
        __syncthreads();
        if (tid < 4) {
            float v = 0.0f;
            for (int w = 0; w < kGThreads / 32; ++w) v += red[w * 4 + tid];
            p.part_td[((long long)b * p.n_tiles + tile) * (2 * kTdN) + 33 + tid] = v;
        }
    }
}

}  // namespace mixfe
