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
//   FFT     in-place decimation in frequency over padded shared-memory rows, 3 or 4 stages per pass in registers
//           (fftg.cuh); X[k] ends at the bit-reversed position
//   power   L/R untangled from X[k] and X[n-k]; |L_k|^2, |R_k|^2 into a (frame, bin) column
//   mel     thread (frame, filter) walks the filter's contiguous support (CSR table in global memory)
// then, per stem, the same epilogue as K1: log(mel + 1e-10), per-(stem, mel) partial sums, the masking S tile,
// 32-byte-run stores of the log tile.  The per-tile partials have K1's layout, so K2 finalizes both.
#pragma once

#include "fftg.cuh"
#include "k1_features.cuh"

namespace mixfe {


struct K1GParams {
    K1Params base;                   // stems, strides, B, T, F, hop, n_mels(_pad), n_tiles, outputs, partials
    int n_fft, log2n, nf;            // nf: frames transformed at once
    const float2* twg;               // [n_fft/2] exp(-2 pi i e / n_fft)
    const float* win;                // [n_fft] 0.5 * periodic hann
    const int4* mel_meta;            // [n_mels] (first bin of the filter's support, bins in it, offset of its weights in mel_w, -)
    const float* mel_w;
    int pw_stride;                   // float2 entries per frame of the padded power column (odd)
    int off_pw, off_mel, off_S, off_red;
};

__global__ void __launch_bounds__(kGThreads, 3) k1g_features(const __grid_constant__ K1GParams g) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const K1Params& p = g.base;
    float2* const z = reinterpret_cast<float2*>(smem);                         // [nf][fg_row(n_fft)] padded rows
    float2* const pw = reinterpret_cast<float2*>(smem + g.off_pw);             // [nf][pw_stride]
    float2* const melb = reinterpret_cast<float2*>(smem + g.off_mel);          // [8][n_mels_pad] (L, R)
    float* const S = reinterpret_cast<float*>(smem + g.off_S);                 // [4][8][n_mels_pad]
    float* const red = reinterpret_cast<float*>(smem + g.off_red);             // [8][4]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x, b = blockIdx.y;
    const int n = g.n_fft, half = n >> 1, zs = fg_row(n);
    const int t0 = tile * kTF;
    const int nvalid = p.F - t0 < kTF ? p.F - t0 : kTF;
    const unsigned mg_nb = k1g_magic(half + 1), mg_mp = k1g_magic(p.n_mels_pad);

    if (p.want_stats) {
        k1_time_stats<1>(p, b, tile, tid);
    }

    for (int s = 0; s < kNumStems; ++s) {
        const float* gL = p.stem[s] + (long long)b * p.stride_b;
        const float* gR = gL + p.stride_c;
        for (int f0 = 0; f0 < kTF; f0 += g.nf) {
            const int frames = nvalid - f0 < g.nf ? (nvalid - f0 > 0 ? nvalid - f0 : 0) : g.nf;
            // ---- load + window ------------------------------------------------------------------
            fg_load_frames(z, gL, gR, g.win, g.log2n, p.hop, p.T, t0 + f0, frames, tid);
            __syncthreads();
            // ---- FFT (fftg.cuh; ends with a barrier) ----------------------------------------------------
            fg_fft(z, g.twg, g.log2n, frames, tid);
            // ---- untangle L / R, power ------------------------------------------------------------
            // Lane i takes bin k = bitrev_{log2n-1}(i): X[k] then sits at position 2 i and X[n-k] at an odd position
            // walking the other way, so a warp reads neighbouring words instead of 32 words of one bank.  The power
            // column is padded (k + (k >> pad_sh)) so that the scattered k of a warp fall into different banks.
            const int sh = 32 - g.log2n, pad_sh = g.log2n > 6 ? g.log2n - 5 : 31;
            for (int idx = tid; idx < frames * (half + 1); idx += kGThreads) {
                const int f = k1g_div(idx, mg_nb), i = idx - f * (half + 1);
                const int k = i == half ? half : (int)(__brev((unsigned)i) >> (sh + 1));
                const float2 x = z[f * zs + fg_pad((int)(__brev((unsigned)k) >> sh))];
                const float2 y = z[f * zs + fg_pad((int)(__brev((unsigned)((n - k) & (n - 1))) >> sh))];
                const float a = x.x + y.x, bb = x.y - y.y, c = x.x - y.x, d = x.y + y.y;
                pw[(size_t)f * g.pw_stride + k + (k >> pad_sh)] = make_float2(fmaf(a, a, bb * bb), fmaf(d, d, c * c));
            }
            __syncthreads();
            // ---- mel projection: 8 lanes per (frame, filter) walk the filter's support, then a 3-step reduction ----
            {
                const int sub = tid & 7, grp = tid >> 3, total = g.nf * p.n_mels_pad;
                for (int it0 = 0; it0 < total; it0 += kGThreads / 8) {
                    const int it = it0 + grp;
                    const int f = k1g_div(it, mg_mp), m = it - f * p.n_mels_pad;
                    float aL = 0.0f, aR = 0.0f;
                    if (it < total && f < frames && m < p.n_mels) {
                        const int4 mm = __ldg(g.mel_meta + m);
                        const int lo = mm.x, len = mm.y;
                        const float* w = g.mel_w + mm.z;
                        const float2* src = pw + (size_t)f * g.pw_stride;
                        for (int i = sub; i < len; i += 8) {
                            const float wi = __ldg(w + i);
                            const float2 v = src[lo + i + ((lo + i) >> pad_sh)];
                            aL = fmaf(wi, v.x, aL);
                            aR = fmaf(wi, v.y, aR);
                        }
                    }
#pragma unroll
                    for (int o = 1; o < 8; o <<= 1) {
                        aL += __shfl_xor_sync(kFullMask, aL, o);
                        aR += __shfl_xor_sync(kFullMask, aR, o);
                    }
                    if (sub == 0 && it < total) melb[(f0 + f) * p.n_mels_pad + m] = make_float2(aL, aR);
                }
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
