// K6: backward of the log-mel front end, d loss / d audio from d loss / d logmel.
//
// Reference behaviour: autograd through MelSpectrogramPreprocessor.forward (src/model.py:41-67), which the
// style-transfer loop relies on -- the TCN output flows through the frozen encoder's log-mel
// (src/train_style_transfer.py:206-215, inference/test_tcn_style_transfer.py:44-57): log -> mel matmul ->
// |X|^2 -> rfft -> Hann window -> framing with hop -> 'reflect' padding.
//
//   logmel = log(mel + 1e-10)                     dmel  = g / (mel + 1e-10)
//   mel[m] = sum_k fb[k][m] P[k]                  dP[k] = sum_m fb[k][m] dmel[m]
//   P[k]   = |X[k]|^2, k = 0..n/2                 G[k]  = 2 X[k] dP[k]
//   X      = rfft(w * frame)                      dframe[i] = w[i] * Re sum_k G[k] exp(+2 pi i k i / n)
//   frame  = xpad[t hop + i]                      overlap-add, then the adjoint of the reflect padding
//
// One CTA per (tile of 8 frames, clip, stem).  The forward spectrum is recomputed (no activations are saved: the
// forward writes only the log-mel), with the shared-memory FFT core of fftg.cuh.  L and R share both transforms: forward as
// one complex FFT of L + iR; backward as one complex FFT of conj(Y), Y = Hermitian extension of G_L plus i times that
// of G_R, whose inverse transform is dframe_L + i dframe_R.  conj(Y) is written in place over the forward spectrum,
// which fg_fft leaves in bit-reversed order, and the inverse runs as decimation in time (bit-reversed in, natural order
// out): one spectrum buffer, no permutation pass.  Frames are overlap-added in shared memory; the tile's
// span goes to global memory with atomicAdd (a sample receives at most two tiles' values when n_fft <= 9 hop, so
// the sum is order-independent there; within n_fft/2 samples of the clip edges the reflected terms add a third).
#pragma once

#include "k1g_features.cuh"

namespace mixfe {

struct K6Params {
    const float* stem[kNumStems];    // forward inputs
    long long stride_b, stride_c;
    float* out[kNumStems];           // gradients, zero-initialised by the caller
    long long ostride_b, ostride_c;
    const float* g;                  // d loss / d logmel: (b, ch, m, f) at g[b * g_stride_b + (ch * n_mels + m) * F + f]
    long long g_stride_b;
    int B, T, F, hop, n_mels, n_mels_pad, n_fft, log2n, nf, span;
    int span_pad;                    // floats per channel of the overlap-add span buffer (>= span)
    const float2* twg;               // [n_fft/2]
    const float* win;                // [n_fft] 0.5 * periodic hann
    const int4* mel_meta;            // filterbank by filter (CSR): (first bin, bins, weight offset, -)
    const float* mel_w;
    const int4* bin_meta;            // filterbank by bin (CSR of the transpose): (first filter, filters, weight offset, -),
                                     // entry i for bin bitrev_{log2n-1}(i), entry n/2 for the Nyquist bin
    const float* bin_w;
    int pw_stride;                   // entries per frame of the padded power column (k + (k >> pad_sh))
    int off_pw, off_mel, off_ola;
};

__global__ void __launch_bounds__(kGThreads, 3) k6_logmel_bwd(const __grid_constant__ K6Params p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    float2* const z = reinterpret_cast<float2*>(smem);                       // [nf][fg_row(n)] padded rows (fftg.cuh)
    float2* const pw = reinterpret_cast<float2*>(smem + p.off_pw);           // [nf][pw_stride] power
    float2* const melb = reinterpret_cast<float2*>(smem + p.off_mel);        // [nf][n_mels_pad] dmel
    float* const ola = reinterpret_cast<float*>(smem + p.off_ola);           // [2][span_pad]

    const int tid = threadIdx.x;
    const int tile = blockIdx.x, b = blockIdx.y, s = blockIdx.z;
    const int n = p.n_fft, half = n >> 1, nb = half + 1, zs = fg_row(n);
    const int t0 = tile * kTF;
    const int nvalid = p.F - t0 < kTF ? p.F - t0 : kTF;
    const int sh = 32 - p.log2n;
    const unsigned mg_nb = k1g_magic(nb), mg_mp = k1g_magic(p.n_mels_pad);
    const int pad_sh = p.log2n > 6 ? p.log2n - 5 : 31;
    const float* gL = p.stem[s] + (long long)b * p.stride_b;
    const float* gR = gL + p.stride_c;

    for (int i = tid; i < 2 * p.span_pad; i += kGThreads) ola[i] = 0.0f;

    for (int f0 = 0; f0 < nvalid; f0 += p.nf) {
        const int frames = nvalid - f0 < p.nf ? nvalid - f0 : p.nf;
        // ---- forward: load + window, FFT, untangle ------------------------------------------------
        fg_load_frames(z, gL, gR, p.win, p.log2n, p.hop, p.T, t0 + f0, frames, tid);
        __syncthreads();
        fg_fft(z, p.twg, p.log2n, frames, tid);
        for (int idx = tid; idx < frames * nb; idx += kGThreads) {
            // lane i takes bin k = bitrev_{log2n-1}(i) (k1g_features' untangle); the power column is padded.  The
            // spectrum itself stays in z: the gradient stage below reads it again and overwrites it in place.
            const int f = k1g_div(idx, mg_nb), i = idx - f * nb;
            const int k = i == half ? half : (int)(__brev((unsigned)i) >> (sh + 1));
            const float2 x = z[f * zs + fg_pad((int)(__brev((unsigned)k) >> sh))];
            const float2 y = z[f * zs + fg_pad((int)(__brev((unsigned)((n - k) & (n - 1))) >> sh))];
            const float a = x.x + y.x, bb = x.y - y.y, c = x.x - y.x, d = x.y + y.y;
            pw[(size_t)f * p.pw_stride + k + (k >> pad_sh)] = make_float2(fmaf(a, a, bb * bb), fmaf(d, d, c * c));
        }
        __syncthreads();
        // ---- mel, then dmel = g / (mel + 1e-10) ---------------------------------------------------
        for (int idx = tid; idx < frames * p.n_mels_pad; idx += kGThreads) {
            const int f = k1g_div(idx, mg_mp), m = idx - f * p.n_mels_pad;
            float dL = 0.0f, dR = 0.0f;
            if (m < p.n_mels) {
                const int4 mm = __ldg(p.mel_meta + m);
                const int lo = mm.x, len = mm.y;
                const float* w = p.mel_w + mm.z;
                const float2* src = pw + (size_t)f * p.pw_stride;
                float aL = 0.0f, aR = 0.0f;
                for (int i = 0; i < len; ++i) {
                    const float wi = __ldg(w + i);
                    const float2 v = src[lo + i + ((lo + i) >> pad_sh)];
                    aL = fmaf(wi, v.x, aL);
                    aR = fmaf(wi, v.y, aR);
                }
                const long long o = (long long)b * p.g_stride_b + ((long long)(2 * s) * p.n_mels + m) * p.F + t0 + f0 + f;
                dL = __ldg(p.g + o) / (aL + 1e-10f);
                dR = __ldg(p.g + o + (long long)p.n_mels * p.F) / (aR + 1e-10f);
            }
            melb[idx] = make_float2(dL, dR);
        }
        __syncthreads();
        // ---- dP[k] = sum_m fb[k][m] dmel[m]; G = 2 X dP; conj of the Hermitian extension of G_L + i G_R, written IN
        //      PLACE over X[k] and X[n-k] (element k of the inverse transform's input at position bitrev(k), which is
        //      what the decimation-in-time passes take).  One thread per bin, lane i on bin bitrev(i) as in the
        //      untangle; the bin's filter range is read once for all frames of the group. -------------------------
        for (int i = tid; i <= half; i += kGThreads) {
            const int k = i == half ? half : (int)(__brev((unsigned)i) >> (sh + 1));
            const int4 bm = __ldg(p.bin_meta + i);           // the table is in lane order (mixfe.cu)
            const float* w = p.bin_w + bm.z;
            const int px = fg_pad((int)(__brev((unsigned)k) >> sh));
            const int py = fg_pad((int)(__brev((unsigned)((n - k) & (n - 1))) >> sh));
            for (int f = 0; f < frames; ++f) {
                const float2* src = melb + (size_t)f * p.n_mels_pad + bm.x;
                float dL = 0.0f, dR = 0.0f;
                for (int q = 0; q < bm.y; ++q) {
                    const float wi = __ldg(w + q);
                    const float2 v = src[q];
                    dL = fmaf(wi, v.x, dL);
                    dR = fmaf(wi, v.y, dR);
                }
                float2* zf = z + f * zs;
                const float2 x = zf[px], y = zf[py];
                const float xa = x.x + y.x, xb = x.y - y.y, xc = x.x - y.x, xd = x.y + y.y;   // L = xa + i xb, R = xd - i xc
                const float a = 2.0f * xa * dL, bi = 2.0f * xb * dL;         // G_L = a + i bi
                const float c = 2.0f * xd * dR, di = -2.0f * xc * dR;        // G_R = c + i di
                if (k == 0 || k == half) {
                    zf[px] = make_float2(a, -c);
                } else {
                    zf[px] = make_float2(0.5f * (a - di), -0.5f * (bi + c));      // conj of (G_L + i G_R) / 2
                    zf[py] = make_float2(0.5f * (a + di), -0.5f * (c - bi));      // conj of (conj G_L + i conj G_R) / 2
                }
            }
        }
        __syncthreads();
        // ---- inverse transform = conj(FFT(conj Y)), bit-reversed in, natural order out -----------------------
        fg_fft_dit(z, p.twg, p.log2n, frames, tid);
        // ---- window and overlap-add, one frame after the other (frames overlap) ------------------------
        for (int f = 0; f < frames; ++f) {
            const int q0 = (f0 + f) * p.hop;
            for (int i = tid; i < n; i += kGThreads) {
                const float2 r = z[f * zs + fg_pad(i)];
                const float w = 2.0f * __ldg(p.win + i);
                const int o = q0 + i;
                ola[o] = fmaf(w, r.x, ola[o]);
                ola[p.span_pad + o] = fmaf(-w, r.y, ola[p.span_pad + o]);
            }
            __syncthreads();
        }
    }
    __syncthreads();
    // ---- the tile's span -> gradient, through the adjoint of the reflect padding ------------------------
    float* dLp = p.out[s] + (long long)b * p.ostride_b;
    float* dRp = dLp + p.ostride_c;
    const int used = nvalid > 0 ? (nvalid - 1) * p.hop + n : 0;
    for (int i = tid; i < used; i += kGThreads) {
        long long j = (long long)t0 * p.hop - half + i;
        if (j < 0) j = -j;
        if (j >= p.T) j = 2ll * (p.T - 1) - j;
        atomicAdd(dLp + j, ola[i]);
        atomicAdd(dRp + j, ola[p.span_pad + i]);
    }
}

}  // namespace mixfe
