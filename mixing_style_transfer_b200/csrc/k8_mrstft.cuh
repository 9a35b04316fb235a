// K8: one resolution of the multi-resolution STFT loss and its gradient.
//
// Reference behaviour: MultiResolutionSTFTLoss (src/loss.py:332-448), the cycle-consistency loss of the
// style-transfer loop (src/train_style_transfer.py:73,239).  Per resolution (n_fft = win, hop), on (B, C, T)
// prediction x and target y:
//   |X|, |Y| = abs(torch.stft(., n_fft, hop, win, hann_window(win), center=True, 'reflect'))        (:362-381)
//   sc  = || |Y| - |X| ||_F / (|| |Y| ||_F + 1e-8)                                                     (:395)
//   lm  = mean | log(|X| + 1e-5) - log(|Y| + 1e-5) |                                                  (:408)
// and the loss is the mean over resolutions of sc + lm (:433-446).
//
// k8_partials: one CTA per (tile of 8 frames, row = batch x channel).  x and y of a row share ONE complex FFT
// (x + i y; the shared-memory FFT core of fftg.cuh) and are untangled like L/R in K1.  Per tile: sum (|Y| - |X|)^2,
// sum |Y|^2, sum |log(|X| + 1e-5) - log(|Y| + 1e-5)| (fixed-order CTA reduction, no atomics).
// k8_finish: fp64 sum of the tile partials in a fixed order -> {A, Bn, L1, sc, lm, alpha, beta} with
//   alpha = 1 / ((sqrt(Bn) + 1e-8) sqrt(A)) (0 when A = 0), beta = 1 / count: the factors of d loss / d |X|.
// k8_backward: recompute the spectra, d loss / d |X| = alpha (|X| - |Y|) + beta sign(log..-log..) / (|X| + 1e-5),
//   G = that * X / |X| (0 at |X| = 0, as torch.abs), Hermitian extension, inverse FFT, window, overlap-add in shared
//   memory, adjoint of the reflect padding (K6's tail).  The upstream gradient and 1 / n_resolutions arrive as one
//   device scalar.
#pragma once

#include "k6_logmel_bwd.cuh"

namespace mixfe {

struct K8Params {
    const float* x;                  // prediction, element (row, t) at x[(row / C) * xs_b + (row % C) * xs_c + t]
    const float* y;                  // target, same addressing with ys_b, ys_c
    long long xs_b, xs_c, ys_b, ys_c;
    int rows, C, T, F, hop, n_fft, log2n, nf, n_tiles, span;
    const float2* twg;
    const float* win;                // 0.5 * periodic hann
    float* part;                     // [rows][n_tiles][3]
    double* state;                   // [8]: A, Bn, L1, sc, lm, alpha, beta, count
    const float* gscale;             // device scalar: upstream gradient / n_resolutions   (backward)
    float* gx;                       // gradient of x, zero-initialised by the caller, addressing gs_b, gs_c (backward)
    long long gs_b, gs_c;
    int span_pad;                    // floats of the overlap-add span buffer (>= span)
    int off_ola, off_red;
};

__device__ __forceinline__ void k8_load_fft(const K8Params& p, float2* z, const float* xr, const float* yr, int t0, int f0,
                                            int frames, int tid) {
    fg_load_frames(z, xr, yr, p.win, p.log2n, p.hop, p.T, t0 + f0, frames, tid);
    __syncthreads();
    fg_fft(z, p.twg, p.log2n, frames, tid);
}

__global__ void __launch_bounds__(kGThreads, 3) k8_partials(const __grid_constant__ K8Params p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    float2* const z = reinterpret_cast<float2*>(smem);
    float* const red = reinterpret_cast<float*>(smem + p.off_red);             // [8 warps][3]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x, row = blockIdx.y;
    const int n = p.n_fft, half = n >> 1, nb = half + 1, sh = 32 - p.log2n, zs = fg_row(n);
    const unsigned mg_nb = k1g_magic(nb);
    const int t0 = tile * kTF;
    const int nvalid = p.F - t0 < kTF ? p.F - t0 : kTF;
    const float* xr = p.x + (long long)(row / p.C) * p.xs_b + (long long)(row % p.C) * p.xs_c;
    const float* yr = p.y + (long long)(row / p.C) * p.ys_b + (long long)(row % p.C) * p.ys_c;
    float a2 = 0.0f, b2 = 0.0f, l1 = 0.0f;
    for (int f0 = 0; f0 < nvalid; f0 += p.nf) {
        const int frames = nvalid - f0 < p.nf ? nvalid - f0 : p.nf;
        k8_load_fft(p, z, xr, yr, t0, f0, frames, tid);
        for (int idx = tid; idx < frames * nb; idx += kGThreads) {
            // lane i takes bin k = bitrev_{log2n-1}(i): neighbouring words of z instead of one bank (k1g_features)
            const int f = k1g_div(idx, mg_nb), i = idx - f * nb;
            const int k = i == half ? half : (int)(__brev((unsigned)i) >> (sh + 1));
            const float2 u = z[f * zs + fg_pad((int)(__brev((unsigned)k) >> sh))];
            const float2 v = z[f * zs + fg_pad((int)(__brev((unsigned)((n - k) & (n - 1))) >> sh))];
            const float a = u.x + v.x, bb = u.y - v.y, c = u.x - v.x, d = u.y + v.y;   // X = a + i bb, Y = d - i c
            const float mx = sqrtf(fmaf(a, a, bb * bb)), my = sqrtf(fmaf(d, d, c * c));
            const float dm = my - mx;
            a2 = fmaf(dm, dm, a2);
            b2 = fmaf(my, my, b2);
            l1 += fabsf(logf(mx + 1e-5f) - logf(my + 1e-5f));
        }
        __syncthreads();
    }
    a2 = warp_sum(a2); b2 = warp_sum(b2); l1 = warp_sum(l1);
    if (lane == 0) { red[warp * 3] = a2; red[warp * 3 + 1] = b2; red[warp * 3 + 2] = l1; }
    __syncthreads();
    if (tid < 3) {
        float v = 0.0f;
        for (int w = 0; w < kGThreads / 32; ++w) v += red[w * 3 + tid];
        p.part[((long long)row * p.n_tiles + tile) * 3 + tid] = v;
    }
}

__global__ void __launch_bounds__(256) k8_finish(const K8Params p) {
    __shared__ double sh[3][8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long total = (long long)p.rows * p.n_tiles;
    double acc[3] = {0.0, 0.0, 0.0};
    for (long long i = tid; i < total; i += 256) {
#pragma unroll
        for (int q = 0; q < 3; ++q) acc[q] += (double)__ldg(p.part + i * 3 + q);
    }
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        acc[q] = warp_sum(acc[q]);
        if (lane == 0) sh[q][warp] = acc[q];
    }
    __syncthreads();
    if (tid == 0) {
        double v[3];
        for (int q = 0; q < 3; ++q) {
            v[q] = 0.0;
            for (int w = 0; w < 8; ++w) v[q] += sh[q][w];
        }
        const double count = (double)p.rows * (double)(p.n_fft / 2 + 1) * (double)p.F;
        const double nA = sqrt(v[0]), nB = sqrt(v[1]);
        p.state[0] = v[0]; p.state[1] = v[1]; p.state[2] = v[2];
        p.state[3] = nA / (nB + 1e-8);
        p.state[4] = v[2] / count;
        p.state[5] = nA > 0.0 ? 1.0 / ((nB + 1e-8) * nA) : 0.0;
        p.state[6] = 1.0 / count;
        p.state[7] = count;
    }
}

__global__ void __launch_bounds__(kGThreads, 3) k8_backward(const __grid_constant__ K8Params p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    float2* const z = reinterpret_cast<float2*>(smem);                       // [nf][fg_row(n)] padded rows (fftg.cuh)
    float* const ola = reinterpret_cast<float*>(smem + p.off_ola);           // [span_pad]
    const int tid = threadIdx.x;
    const int tile = blockIdx.x, row = blockIdx.y;
    const int n = p.n_fft, half = n >> 1, nb = half + 1, sh = 32 - p.log2n, zs = fg_row(n);
    const unsigned mg_nb = k1g_magic(nb);
    const int t0 = tile * kTF;
    const int nvalid = p.F - t0 < kTF ? p.F - t0 : kTF;
    const float* xr = p.x + (long long)(row / p.C) * p.xs_b + (long long)(row % p.C) * p.xs_c;
    const float* yr = p.y + (long long)(row / p.C) * p.ys_b + (long long)(row % p.C) * p.ys_c;
    const float gs = __ldg(p.gscale);
    const float alpha = (float)p.state[5] * gs, beta = (float)p.state[6] * gs;

    for (int i = tid; i < p.span_pad; i += kGThreads) ola[i] = 0.0f;
    for (int f0 = 0; f0 < nvalid; f0 += p.nf) {
        const int frames = nvalid - f0 < p.nf ? nvalid - f0 : p.nf;
        k8_load_fft(p, z, xr, yr, t0, f0, frames, tid);
        for (int idx = tid; idx < frames * nb; idx += kGThreads) {
            const int f = k1g_div(idx, mg_nb), i = idx - f * nb;
            const int k = i == half ? half : (int)(__brev((unsigned)i) >> (sh + 1));
            float2* zf = z + f * zs;
            const int px = fg_pad((int)(__brev((unsigned)k) >> sh));
            const int py = fg_pad((int)(__brev((unsigned)((n - k) & (n - 1))) >> sh));
            const float2 u = zf[px], v = zf[py];
            const float a = u.x + v.x, bb = u.y - v.y, c = u.x - v.x, d = u.y + v.y;   // X = a + i bb, Y = d - i c
            const float mx = sqrtf(fmaf(a, a, bb * bb)), my = sqrtf(fmaf(d, d, c * c));
            const float dl = logf(mx + 1e-5f) - logf(my + 1e-5f);
            const float sg = dl > 0.0f ? 1.0f : (dl < 0.0f ? -1.0f : 0.0f);
            const float gm = fmaf(alpha, mx - my, beta * sg / (mx + 1e-5f));            // d loss / d |X|
            const float inv = mx > 0.0f ? gm / mx : 0.0f;
            const float gx = inv * a, gy = inv * bb;                                    // G = d loss / d X
            // conj of the Hermitian extension of G, IN PLACE over the two spectrum entries it came from: element k of the
            // inverse transform's input at position bitrev(k) (the imaginary slot of the packed transform stays empty)
            if (k == 0 || k == half) {
                zf[px] = make_float2(gx, 0.0f);
            } else {
                zf[px] = make_float2(0.5f * gx, -0.5f * gy);
                zf[py] = make_float2(0.5f * gx, 0.5f * gy);
            }
        }
        __syncthreads();
        fg_fft_dit(z, p.twg, p.log2n, frames, tid);              // bit-reversed in, natural order out
        for (int f = 0; f < frames; ++f) {
            const int q0 = (f0 + f) * p.hop;
            for (int i = tid; i < n; i += kGThreads) {
                const float2 r = z[f * zs + fg_pad(i)];
                ola[q0 + i] = fmaf(2.0f * __ldg(p.win + i), r.x, ola[q0 + i]);
            }
            __syncthreads();
        }
    }
    __syncthreads();
    float* gr = p.gx + (long long)(row / p.C) * p.gs_b + (long long)(row % p.C) * p.gs_c;
    const int used = nvalid > 0 ? (nvalid - 1) * p.hop + n : 0;
    for (int i = tid; i < used; i += kGThreads) {
        long long j = (long long)t0 * p.hop - half + i;
        if (j < 0) j = -j;
        if (j >= p.T) j = 2ll * (p.T - 1) - j;
        atomicAdd(gr + j, ola[i]);
    }
}

}  // namespace mixfe
