// K4/K5: stereo reverb of the augmented mixture as an FFT-domain cross-correlation, and the
// energy-weighted redistribution onto the stems.
//
// Reference behaviour being replaced (src/mixing_utils.py of the reference):
//   :458-479  apply_reverb   F.conv1d(x[None,None], h[None,None], padding=L//2)[..., :T] per channel, i.e. the
//                            cross-correlation r[n] = sum_k x[n + k - L//2] h[k] with L = 22050 taps
//                            (22.8 s per call on 8 CPU cores), then 0.7 x + 0.3 r
//   :409-417  redistribution s_i += mix_rev * (mean(s_i^2) / (sum_j mean(s_j^2) + 1e-8)) * 0.3
//
// Method: the mixture's L and R are packed as one complex signal (h is real, so the correlation
// acts on both parts independently) and correlated with h through a 2^20-point FFT computed as
// 1024 x 1024 (four-step) with the same one-warp 1024-point transform as the STFT kernel:
//   columns  (stride-1024 transforms + inter-step twiddle)   k4_cols<kColsX | kColsH>
//   rows     X: forward, times H, forward again (= inverse of the conjugate), twiddle;  H: forward
//   columns  last transform, conjugate, 1/N, wet/dry mix and redistribution in the epilogue
// Frequency-domain data stays in the transposed order the four-step algorithm produces, so no
// transposition pass exists.  Clips longer than N - L + 1 samples are processed as overlap-save
// segments.
#pragma once

#include "fft32.cuh"
#include "k3_augment.cuh"
#include "plan.cuh"

namespace mixfe {

// The transform length is N = N1 * 1024 with N1 = 512 (N = 2^19: a 10 s clip plus the 0.5 s impulse fits one
// circular correlation) or N1 = 1024 (N = 2^20, longer clips, overlap-save segments).  Element n = r * 1024 + c sits in
// row r (of N1) and column c (of 1024); "columns" are the stride-1024 transforms of length N1, "rows" the contiguous
// transforms of length 1024.
constexpr int kRevN2 = 1024;
constexpr int kRevMaxN = 1024 * kRevN2;          // 2^20
constexpr int kColWarps = 8;

template <int LOG1> struct RevGeom {
    static constexpr int N1 = 1 << LOG1;                      // column length
    static constexpr int kColsPerWarp = 1024 / N1;            // a warp transforms 1024 points: one column of 1024 or two of 512
    static constexpr int kCols = kColWarps * kColsPerWarp;    // columns per CTA tile
    static constexpr int kStride = N1 + 4;                    // float2 elements per staged column
    static constexpr int kTwShift = 10 - LOG1;                // exp(-2 pi i e / N) = table entry e << kTwShift of the 2^20 tables
};

// One warp, 1024 points: in v[n1] = x[32 n1 + lane]; out v[k2] = X[lane + 32 k2].
// exw: 1024 float2 of warp-private shared memory; tw[k1*32 + n2] = exp(-2 pi i n2 k1/1024) in shared memory.
__device__ __forceinline__ void warp_fft1024(float (&re)[32], float (&im)[32], float2* exw, const float2* tw, int lane) {
    dft32_dif(re, im);
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
    __syncwarp();
    dft32_dif(re, im);
    float tr[32], ti[32];
#pragma unroll
    for (int k2 = 0; k2 < 32; ++k2) { tr[k2] = re[brev5(k2)]; ti[k2] = im[brev5(k2)]; }
#pragma unroll
    for (int k2 = 0; k2 < 32; ++k2) { re[k2] = tr[k2]; im[k2] = ti[k2]; }
}

// One warp, TWO 512-point transforms (columns A and B) at once, 512 = 16 x 32:
//   in  re[n1], im[n1]          = x_A[32 n1 + lane], n1 < 16;  re[16 + n1], im[16 + n1] = x_B[32 n1 + lane]
//   out re[k2], im[k2], k2 < 32 = X_A[lane + 16 k2] for lane < 16,  X_B[(lane - 16) + 16 k2] for lane >= 16.
// Every lane runs two 16-point DFTs (its 16 samples of each column), the inter-pass twiddle W_512^(n2 k1) comes from
// the even rows of the 1024-point table, the exchange buffer holds 32 rows (k1 of A, then k1 of B) of 32 entries with
// the XOR swizzle of warp_fft1024, and every lane finishes one 32-point DFT.
__device__ __forceinline__ void warp_fft512x2(float (&re)[32], float (&im)[32], float2* exw, const float2* tw, int lane) {
#pragma unroll
    for (int col = 0; col < 2; ++col) {
        float ar[16], ai[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) { ar[i] = re[16 * col + i]; ai[i] = im[16 * col + i]; }
        dft16_dif(ar, ai);
#pragma unroll
        for (int k1 = 0; k1 < 16; ++k1) {
            const float yr = ar[brev4(k1)], yi = ai[brev4(k1)];
            float2 o;
            if (k1 == 0) {
                o = make_float2(yr, yi);
            } else {
                const float2 t = tw[(2 * k1) * 32 + lane];            // exp(-2 pi i lane k1 / 512)
                o.x = fmaf(yr, t.x, -yi * t.y);
                o.y = fmaf(yr, t.y, yi * t.x);
            }
            const int row = 16 * col + k1;
            exw[row * 32 + (lane ^ row)] = o;
        }
    }
    __syncwarp();
#pragma unroll
    for (int n2 = 0; n2 < 32; ++n2) {
        const float2 v = exw[lane * 32 + (n2 ^ lane)];
        re[n2] = v.x;
        im[n2] = v.y;
    }
    __syncwarp();
    dft32_dif(re, im);
    float tr[32], ti[32];
#pragma unroll
    for (int k2 = 0; k2 < 32; ++k2) { tr[k2] = re[brev5(k2)]; ti[k2] = im[brev5(k2)]; }
#pragma unroll
    for (int k2 = 0; k2 < 32; ++k2) { re[k2] = tr[k2]; im[k2] = ti[k2]; }
}

// exp(-2 pi i e / 2^20) from two 1024-entry tables: e = 1024 hi + lo.
__device__ __forceinline__ float2 twiddle_long(const float2* __restrict__ tw_lin, const float2* __restrict__ tw_fine, unsigned e) {
    const float2 a = __ldg(tw_lin + (e >> 10)), c = __ldg(tw_fine + (e & 1023u));
    return make_float2(fmaf(a.x, c.x, -a.y * c.y), fmaf(a.x, c.y, a.y * c.x));
}

struct K4Params {
    float* out[kNumStems];                 // augmented stems (K3 output); updated in place by the last pass
    long long stride_b, stride_c;
    int T, L, half, seg_len, n_seg, n_chunks;
    // standalone mode (AudioAugmenter.apply_reverb on one tensor, :458-479): out[0] = 0.7 sa_in + 0.3 corr(sa_in, h),
    // no stem sum, no redistribution; n_ch = 1 packs a single channel (imaginary part zero)
    int standalone, n_ch;
    const float* sa_in;
    long long imp_stride;                  // floats between the impulses of consecutive clips (0: one impulse for all)
    const int* slot_clip;                  // [n_slots] clip index of each reverb slot
    const float* impulse;                  // [B][L] (row b at impulse + b * imp_stride)
    float2* X;                             // [n_slots][n_seg][N]
    float2* H;                             // [n_slots][N]
    const double* part_energy;             // [B][4][2][n_chunks]
    const float2* tw;                      // [1024] 32x32 twiddles
    const float2* tw_lin;                  // [1024] exp(-2 pi i j/1024)
    const float2* tw_fine;                 // [1024] exp(-2 pi i e/2^20)
};

enum { kColsX = 0, kColsH = 1, kColsInv = 2 };

// grid (1024 / kCols column groups, n_seg, n_slots); 8 warps, warp w transforms column(s) kColsPerWarp * w (+ 1) of the group.
// Register targets of the reverb passes, measured through MIXFE_LIB variants (12 reverbs of the bench step): rows / cols
// = 1 / 2 -> 0.306 ms, 1 / 1 -> 0.336, 2 / 1 -> 0.363, 1 / 3 -> 0.378, 2 / 3 -> 0.410, no bound (as before) -> 0.355.
#ifndef MIXFE_K4_ROWS_CTAS
#define MIXFE_K4_ROWS_CTAS 1            // resident CTAs per SM the register allocation of the row pass aims at
#endif
#ifndef MIXFE_K4_COLS_CTAS
#define MIXFE_K4_COLS_CTAS 2            // same for the column passes (74 KB of tile each)
#endif
template <int MODE, int LOG1>
__global__ void __launch_bounds__(256, MIXFE_K4_COLS_CTAS) k4_cols(const __grid_constant__ K4Params p) {
    using G = RevGeom<LOG1>;
    constexpr int N = G::N1 * kRevN2;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* const tile = reinterpret_cast<float2*>(smem_raw);                   // [kCols][kStride]
    float2* const tw = tile + G::kCols * G::kStride;                             // [1024]
    __shared__ float share[kNumStems];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int col0 = blockIdx.x * G::kCols, seg = blockIdx.y, slot = blockIdx.z;
    const int b = p.slot_clip[slot];
    const long long seg0 = (long long)seg * p.seg_len;
    float2* const Xs = p.X + ((long long)slot * p.n_seg + seg) * N;
    float2* const Hs = p.H + (long long)slot * N;

    for (int i = tid; i < 1024; i += 256) tw[i] = p.tw[i];
    if (MODE == kColsInv && !p.standalone) {
        // energy shares of the four stems (recomputed by every CTA: 8 * n_chunks doubles per stem)
        if (warp < kNumStems) {
            double e = 0.0;
            const double* src = p.part_energy + ((long long)b * kNumStems + warp) * 2 * p.n_chunks;
            for (int i = lane; i < 2 * p.n_chunks; i += 32) e += src[i];
            e = warp_sum(e);
            if (lane == 0) share[warp] = (float)(e / (2.0 * (double)p.T));       // mean over (2, T)
        }
        __syncthreads();
        if (tid == 0) {
            const float tot = ((share[0] + share[1]) + share[2]) + share[3] + 1e-8f;
            for (int s = 0; s < kNumStems; ++s) share[s] = share[s] / tot;
        }
    }
    // ---- stage N1 rows x kCols columns ------------------------------------------------------------
    // Rows are taken kBatch at a time with every global load of the batch issued before the first use: a thread then
    // has up to 32 loads in flight (r02w: one row at a time left the big passes waiting on the long scoreboard at
    // ~25 % of the DRAM rate).  Out-of-range samples load a clamped address and are zeroed afterwards.
    {
        constexpr int kRowStep = 256 / G::kCols, kBatch = 4;
        static_assert(G::N1 % (kRowStep * kBatch) == 0, "row batches");
        const int c = tid % G::kCols;
        for (int r0 = tid / G::kCols; r0 < G::N1; r0 += kRowStep * kBatch) {
            float2 v[kBatch];
            if (MODE == kColsX && !p.standalone) {
                float x[kBatch][2 * kNumStems];
                bool ok[kBatch];
#pragma unroll
                for (int u = 0; u < kBatch; ++u) {
                    const long long t = seg0 - p.half + (long long)(r0 + u * kRowStep) * kRevN2 + col0 + c;
                    ok[u] = t >= 0 && t < p.T;
                    const long long o = (long long)b * p.stride_b + (ok[u] ? t : 0);
#pragma unroll
                    for (int s = 0; s < kNumStems; ++s) {
                        x[u][2 * s] = p.out[s][o];
                        x[u][2 * s + 1] = p.out[s][o + p.stride_c];
                    }
                }
#pragma unroll
                for (int u = 0; u < kBatch; ++u) {
                    float mL = 0.0f, mR = 0.0f;
#pragma unroll
                    for (int s = 0; s < kNumStems; ++s) {              // ((0 + v) + b) + d) + o
                        mL += x[u][2 * s];
                        mR += x[u][2 * s + 1];
                    }
                    v[u] = ok[u] ? make_float2(mL, mR) : make_float2(0.0f, 0.0f);
                }
            } else {
#pragma unroll
                for (int u = 0; u < kBatch; ++u) {
                    const long long n = (long long)(r0 + u * kRowStep) * kRevN2 + col0 + c;
                    v[u] = make_float2(0.0f, 0.0f);
                    if (MODE == kColsX) {
                        const long long t = seg0 - p.half + n;
                        if (t >= 0 && t < p.T) {
                            const long long o = (long long)b * p.stride_b + t;
                            v[u].x = __ldg(p.sa_in + o);
                            if (p.n_ch == 2) v[u].y = __ldg(p.sa_in + o + p.stride_c);
                        }
                    } else if (MODE == kColsH) {
                        if (n < p.L) v[u].x = __ldg(p.impulse + (long long)b * p.imp_stride + n);
                    } else {
                        v[u] = Xs[n];
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < kBatch; ++u) tile[c * G::kStride + r0 + u * kRowStep] = v[u];
        }
    }
    __syncthreads();
    // ---- one (N1 = 1024) or two (N1 = 512) columns per warp -----------------------------------------
    {
        float re[32], im[32];
        if (G::kColsPerWarp == 1) {
            float2* const colbuf = tile + warp * G::kStride;
#pragma unroll
            for (int n1 = 0; n1 < 32; ++n1) {
                const float2 v = colbuf[32 * n1 + lane];
                re[n1] = v.x;
                im[n1] = v.y;
            }
            __syncwarp();
            warp_fft1024(re, im, colbuf, tw, lane);                   // the column's own storage is the exchange buffer
            const unsigned n2 = (unsigned)(col0 + warp);
#pragma unroll
            for (int k2 = 0; k2 < 32; ++k2) {
                const unsigned k1 = (unsigned)(lane + 32 * k2);       // output index of the column transform
                float2 o;
                if (MODE == kColsInv) {
                    o = make_float2(re[k2] * (1.0f / (float)N), -im[k2] * (1.0f / (float)N));
                } else {
                    const float2 w = twiddle_long(p.tw_lin, p.tw_fine, ((n2 * k1) << G::kTwShift) & (unsigned)(kRevMaxN - 1));
                    o = make_float2(fmaf(re[k2], w.x, -im[k2] * w.y), fmaf(re[k2], w.y, im[k2] * w.x));
                }
                colbuf[k1] = o;
            }
        } else {
            // columns 2 warp (A) and 2 warp + 1 (B); their two staged columns (2 x 516 float2 >= 1024) are the exchange buffer
            float2* const bufA = tile + (2 * warp) * G::kStride;
            float2* const bufB = bufA + G::kStride;
#pragma unroll
            for (int n1 = 0; n1 < 16; ++n1) {
                const float2 va = bufA[32 * n1 + lane], vb = bufB[32 * n1 + lane];
                re[n1] = va.x; im[n1] = va.y;
                re[16 + n1] = vb.x; im[16 + n1] = vb.y;
            }
            __syncwarp();
            warp_fft512x2(re, im, bufA, tw, lane);
            float2* const mine = lane < 16 ? bufA : bufB;
            const unsigned n2 = (unsigned)(col0 + 2 * warp + (lane >> 4));
#pragma unroll
            for (int k2 = 0; k2 < 32; ++k2) {
                const unsigned k1 = (unsigned)((lane & 15) + 16 * k2);
                float2 o;
                if (MODE == kColsInv) {
                    o = make_float2(re[k2] * (1.0f / (float)N), -im[k2] * (1.0f / (float)N));
                } else {
                    const float2 w = twiddle_long(p.tw_lin, p.tw_fine, ((n2 * k1) << G::kTwShift) & (unsigned)(kRevMaxN - 1));
                    o = make_float2(fmaf(re[k2], w.x, -im[k2] * w.y), fmaf(re[k2], w.y, im[k2] * w.x));
                }
                mine[k1] = o;
            }
        }
    }
    __syncthreads();
    // ---- write back (coalesced runs of kCols elements) / reverb epilogue ------------------------------
    {
        constexpr int kRowStep = 256 / G::kCols;
        const int c = tid % G::kCols;
        if (MODE != kColsInv) {
            float2* const dst = MODE == kColsX ? Xs : Hs;
#pragma unroll 4
            for (int r = tid / G::kCols; r < G::N1; r += kRowStep)
                dst[(long long)r * kRevN2 + col0 + c] = tile[c * G::kStride + r];
        } else if (p.standalone) {
#pragma unroll 4
            for (int r = tid / G::kCols; r < G::N1; r += kRowStep) {
                const long long n = (long long)r * kRevN2 + col0 + c;
                const long long t = seg0 + n;                            // c[m] is the output sample seg0 + m
                const float2 v = tile[c * G::kStride + r];
                if (n < p.seg_len && t < p.T) {
                    const long long o = (long long)b * p.stride_b + t;
                    p.out[0][o] = __ldg(p.sa_in + o) * 0.7f + v.x * 0.3f;                           // :478-479
                    if (p.n_ch == 2) p.out[0][o + p.stride_c] = __ldg(p.sa_in + o + p.stride_c) * 0.7f + v.y * 0.3f;
                }
            }
        } else {
            constexpr int kBatch = 2;                                    // 16 loads in flight, then 16 stores
            for (int r0 = tid / G::kCols; r0 < G::N1; r0 += kRowStep * kBatch) {
                float y[kBatch][kNumStems][2];
                long long o[kBatch];
                bool ok[kBatch];
#pragma unroll
                for (int u = 0; u < kBatch; ++u) {
                    const long long n = (long long)(r0 + u * kRowStep) * kRevN2 + col0 + c;
                    const long long t = seg0 + n;
                    ok[u] = n < p.seg_len && t < p.T;
                    o[u] = (long long)b * p.stride_b + (ok[u] ? t : 0);
#pragma unroll
                    for (int s = 0; s < kNumStems; ++s) {
                        y[u][s][0] = p.out[s][o[u]];
                        y[u][s][1] = p.out[s][o[u] + p.stride_c];
                    }
                }
#pragma unroll
                for (int u = 0; u < kBatch; ++u) {
                    if (!ok[u]) continue;
                    const float2 v = tile[c * G::kStride + r0 + u * kRowStep];
                    float mL = 0.0f, mR = 0.0f;
#pragma unroll
                    for (int s = 0; s < kNumStems; ++s) {
                        mL += y[u][s][0];
                        mR += y[u][s][1];
                    }
                    const float wetL = mL * 0.7f + v.x * 0.3f, wetR = mR * 0.7f + v.y * 0.3f;      // :478-479
#pragma unroll
                    for (int s = 0; s < kNumStems; ++s) {
                        p.out[s][o[u]] = y[u][s][0] + (wetL * share[s]) * 0.3f;                 // :417
                        p.out[s][o[u] + p.stride_c] = y[u][s][1] + (wetR * share[s]) * 0.3f;
                    }
                }
            }
        }
    }
}

enum { kRowsX = 0, kRowsH = 1 };

// grid (N1 / 8 row groups, n_seg, n_slots); 8 warps, warp w transforms row 8*group + w (1024 points).
template <int MODE, int LOG1>
__global__ void __launch_bounds__(256, MIXFE_K4_ROWS_CTAS) k4_rows(const __grid_constant__ K4Params p) {
    using G = RevGeom<LOG1>;
    constexpr int N = G::N1 * kRevN2;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* const exall = reinterpret_cast<float2*>(smem_raw);                  // [8][1024]
    float2* const tw = exall + 8 * 1024;                                        // [1024]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int row = blockIdx.x * 8 + warp, seg = blockIdx.y, slot = blockIdx.z;
    float2* const Xs = p.X + ((long long)slot * p.n_seg + seg) * N + (long long)row * kRevN2;
    float2* const Hs = p.H + (p.imp_stride == 0 ? 0ll : (long long)slot * N) + (long long)row * kRevN2;
    for (int i = tid; i < 1024; i += 256) tw[i] = p.tw[i];
    __syncthreads();
    float2* const exw = exall + warp * 1024;
    float2* const data = MODE == kRowsH ? Hs : Xs;
    float re[32], im[32];
#pragma unroll
    for (int n1 = 0; n1 < 32; ++n1) {
        const float2 v = data[32 * n1 + lane];
        re[n1] = v.x;
        im[n1] = v.y;
    }
    warp_fft1024(re, im, exw, tw, lane);                              // X[row][lane + 32 k2]
    if constexpr (MODE == kRowsH) {
#pragma unroll
        for (int k2 = 0; k2 < 32; ++k2) Hs[32 * k2 + lane] = make_float2(re[k2], im[k2]);
    } else {
        // correlation: R = X conj(H).  r = IFFT(R) = conj(FFT(conj(R)))/N and conj(R) = conj(X) H.
#pragma unroll
        for (int k2 = 0; k2 < 32; ++k2) {
            const float2 h = Hs[32 * k2 + lane];
            const float xr = re[k2], xi = -im[k2];
            re[k2] = fmaf(xr, h.x, -xi * h.y);
            im[k2] = fmaf(xr, h.y, xi * h.x);
        }
        warp_fft1024(re, im, exw, tw, lane);                          // now indexed by n2 = lane + 32 j
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const unsigned n2 = (unsigned)(lane + 32 * j);
            const float2 w = twiddle_long(p.tw_lin, p.tw_fine, ((n2 * (unsigned)row) << G::kTwShift) & (unsigned)(kRevMaxN - 1));
            Xs[n2] = make_float2(fmaf(re[j], w.x, -im[j] * w.y), fmaf(re[j], w.y, im[j] * w.x));
        }
    }
}

// ------------------------------------------------------------------------------------------------
// host side of mixfe_augment / mixfe_effects / mixfe_reverb
// ------------------------------------------------------------------------------------------------
inline double pole_radius(double a1, double a2) {
    const double disc = a1 * a1 - 4.0 * a2;
    if (disc < 0.0) return std::sqrt(a2);
    const double r1 = std::fabs((-a1 + std::sqrt(disc)) * 0.5), r2 = std::fabs((-a1 - std::sqrt(disc)) * 0.5);
    return r1 > r2 ? r1 : r2;
}

inline void mat2_mul(const double (&a)[4], const double (&b)[4], double (&c)[4]) {
    const double r[4] = {a[0] * b[0] + a[1] * b[2], a[0] * b[1] + a[1] * b[3], a[2] * b[0] + a[3] * b[2], a[2] * b[1] + a[3] * b[3]};
    for (int i = 0; i < 4; ++i) c[i] = r[i];
}

// One normalised section (b0 b1 b2 a1 a2) -> the tables of biquad_block (k3_augment.cuh).
inline void build_section(const double (&c)[5], AugSection& out) {
    std::memset(&out, 0, sizeof out);
    out.b0 = (float)c[0]; out.b1 = (float)c[1]; out.b2 = (float)c[2]; out.a1 = (float)c[3]; out.a2 = (float)c[4];
    out.form = (c[1] == 2.0 * c[0] && c[2] == c[0]) ? 1 : ((c[1] == -2.0 * c[0] && c[2] == c[0]) ? 2 : 0);
    // The lane recurrences run with the fp32 coefficients; the tables describe exactly that system.
    const double a1 = (double)out.a1, a2 = (double)out.a2;
    const double A[4] = {-a1, -a2, 1.0, 0.0};
    double P[4] = {1.0, 0.0, 0.0, 1.0};
    for (int n = 0; n < kAugV; ++n) mat2_mul(P, A, P);
    for (int i = 0; i < 4; ++i) out.mp[0][i] = P[i];       // M = A^16
    for (int d = 1; d < 5; ++d) mat2_mul(out.mp[d - 1], out.mp[d - 1], out.mp[d]);
    for (int d = 0; d < 5; ++d)
        for (int i = 0; i < 4; ++i) out.mpf[d][i] = (float)out.mp[d][i];
    out.scan64 = std::pow(pole_radius(a1, a2), (double)kAugV) > 0.2 ? 1 : 0;
}

// Static compressor constants (k3_augment.cuh): the knee in the amplitude domain and 2^(c0 + c1 lg2 a).
inline void compressor_constants(float thr_db, float ratio, float& knee, float& c0, float& c1) {
    // smallest fp32 a with fl32(20 * fl32(log10 a)) > thr, as the reference's fp32 mask `audio_db > threshold` (:440)
    auto db32 = [](float a) { return 20.0f * (float)std::log10((double)a); };
    float a = (float)std::pow(10.0, (double)thr_db / 20.0);
    for (int i = 0; i < 8 && db32(a) > thr_db; ++i) a = std::nextafterf(a, 0.0f);
    for (int i = 0; i < 16 && !(db32(a) > thr_db); ++i) a = std::nextafterf(a, INFINITY);
    knee = a;
    const double inv = 1.0 / (double)ratio;
    c1 = (float)inv;
    c0 = (float)((double)thr_db * (1.0 - inv) * 0.16609640474436813);        // log2(10) / 20
}

struct AugLaunch {                         // what the three entry points share
    std::vector<AugClipDev> clips;
    std::vector<AugSection> sections;      // [B][4][3]
    int chunk = 0, warm = 0, n_chunks = 0;
};

// decisions -> kernel form; pole radii size the warm-up of the chunked scan
inline int prepare_augment(const mixfe_aug_params* params, int B, int T, int n_stems, AugLaunch& L) {
    L.clips.assign(B, AugClipDev{});
    L.sections.assign((size_t)B * kNumStems * kAugSections, AugSection{});
    double rmax = 0.0;
    for (int b = 0; b < B; ++b) {
        AugClipDev& c = L.clips[b];
        const mixfe_aug_params& q = params[b];
        for (int s = 0; s < n_stems; ++s) {
            c.gain[s] = q.gain_lin[s];
            c.flags[s] = (q.tilt_on[s] ? 1 : 0) | (q.comp_on[s] ? 2 : 0) | (q.bw_on[s] ? 4 : 0);
            c.knee[s] = INFINITY; c.c0[s] = 0.0f; c.c1[s] = 1.0f;
            if (q.comp_on[s]) {
                if (!(q.comp_ratio[s] > 0.0f) || !std::isfinite(q.comp_threshold_db[s]))
                    return fail(MIXFE_EINVAL, "clip %d stem %d: compressor threshold %g dB, ratio %g", b, s,
                                (double)q.comp_threshold_db[s], (double)q.comp_ratio[s]);
                compressor_constants(q.comp_threshold_db[s], q.comp_ratio[s], c.knee[s], c.c0[s], c.c1[s]);
            }
            for (int sec = 0; sec < kAugSections; ++sec) {
                const double* sos = sec == 0 ? q.tilt_sos[s] : q.bw_sos[s] + 6 * (sec - 1);
                const bool on = sec == 0 ? q.tilt_on[s] != 0 : q.bw_on[s] != 0;
                if (!on) continue;
                const double a0 = sos[3];
                if (a0 == 0.0) return fail(MIXFE_EINVAL, "clip %d stem %d: a0 = 0", b, s);
                const double cn[5] = {sos[0] / a0, sos[1] / a0, sos[2] / a0, sos[4] / a0, sos[5] / a0};
                const double r = pole_radius(cn[3], cn[4]);
                if (!(r < 1.0)) return fail(MIXFE_EINVAL, "clip %d stem %d: unstable filter section (pole radius %g)", b, s, r);
                rmax = std::max(rmax, r);
                build_section(cn, L.sections[((size_t)b * kNumStems + s) * kAugSections + sec]);
            }
        }
    }
    // warm-up so that r^warm < 1e-10 (the fp32 output resolves 6e-8; 1e-13 cost the 500 Hz low-pass, r = 0.951, a second
    // 512-sample block per 8192-sample chunk: 12.5 % instead of 6.25 % of extra reads and arithmetic), in whole blocks;
    // chunk >= 8 * warm keeps the overhead <= 12.5 %
    int warm = 0;
    if (rmax > 0.0) {
        const double need = std::log(1e-10) / std::log(rmax);
        warm = need > 4.0e6 ? 1 << 30 : ((int)std::ceil(need) + kAugBlock - 1) / kAugBlock * kAugBlock;
    }
    int chunk = 8192;
    while (chunk < 8 * (long long)warm && chunk < T) chunk *= 2;
    if (chunk >= T || warm >= T) { chunk = (T + kAugBlock - 1) / kAugBlock * kAugBlock; warm = 0; }   // one exact chunk per row
    L.chunk = chunk; L.warm = warm; L.n_chunks = (T + chunk - 1) / chunk;
    return MIXFE_OK;
}

// Transform length for a clip of T samples and an impulse of L taps: 2^19 when one circular correlation of that size
// holds the whole clip (T + L - 1 <= 2^19: the 10 s clips of the training batch), else 2^20 with overlap-save segments.
struct ReverbLayout { int log1 = 10, n = kRevMaxN, n_seg = 0, seg_len = 0; };

inline ReverbLayout reverb_layout(int T, int L) {
    ReverbLayout r;
    if ((long long)T + L - 1 <= (1 << 19)) { r.log1 = 9; r.n = 1 << 19; }
    r.seg_len = r.n - L + 1;
    r.n_seg = (T + r.seg_len - 1) / r.seg_len;
    return r;
}

// the five launches of the FFT-domain correlation for the slots listed in p4.slot_clip
template <int LOG1>
inline int launch_reverb_n(mixfe_plan* pl, K4Params& p4, int n_slots, cudaStream_t stream) {
    using G = RevGeom<LOG1>;
    const int smem_cols = (G::kCols * G::kStride + 1024) * (int)sizeof(float2);
    const int smem_rows = (8 * 1024 + 1024) * (int)sizeof(float2);
    int rc;
    if ((rc = smem_optin(k4_cols<kColsX, LOG1>, pl->device, smem_cols)) || (rc = smem_optin(k4_cols<kColsH, LOG1>, pl->device, smem_cols)) ||
        (rc = smem_optin(k4_cols<kColsInv, LOG1>, pl->device, smem_cols)) || (rc = smem_optin(k4_rows<kRowsX, LOG1>, pl->device, smem_rows)) ||
        (rc = smem_optin(k4_rows<kRowsH, LOG1>, pl->device, smem_rows)))
        return rc;
    {
        KernelTimer timer(pl, MIXFE_KERNEL_REVERB, stream);
        // one impulse for all clips (imp_stride 0): its spectrum is computed once and every slot reads slot 0's
        const int h_slots = p4.imp_stride == 0 ? 1 : n_slots;
        const dim3 g1(kRevN2 / G::kCols, 1, h_slots), gs(kRevN2 / G::kCols, p4.n_seg, n_slots);
        k4_cols<kColsH, LOG1><<<g1, 256, smem_cols, stream>>>(p4);
        k4_rows<kRowsH, LOG1><<<dim3(G::N1 / 8, 1, h_slots), 256, smem_rows, stream>>>(p4);
        k4_cols<kColsX, LOG1><<<gs, 256, smem_cols, stream>>>(p4);
        k4_rows<kRowsX, LOG1><<<dim3(G::N1 / 8, p4.n_seg, n_slots), 256, smem_rows, stream>>>(p4);
        k4_cols<kColsInv, LOG1><<<gs, 256, smem_cols, stream>>>(p4);
    }
    MIXFE_CUDA(cudaGetLastError());
    launch_counter() += 5;
    return MIXFE_OK;
}

inline int launch_reverb(mixfe_plan* pl, K4Params& p4, const ReverbLayout& lay, int n_slots, cudaStream_t stream) {
    return lay.log1 == 9 ? launch_reverb_n<9>(pl, p4, n_slots, stream) : launch_reverb_n<10>(pl, p4, n_slots, stream);
}

inline int run_augment(mixfe_plan* pl, const float* const in[kNumStems], float* const out[kNumStems],
                       long long stride_b, long long stride_c, int B, int T, const mixfe_aug_params* params,
                       const float* impulse, int L, cudaStream_t stream) {
    if (!pl || !in || !out || !params) return fail(MIXFE_EINVAL, "null argument");
    for (int s = 0; s < kNumStems; ++s) {
        if (!in[s] || !out[s]) return fail(MIXFE_EINVAL, "stem pointer %d is NULL", s);
        if (in[s] == out[s]) return fail(MIXFE_EINVAL, "out must not alias in");
    }
    if (B < 0 || T <= 0) return fail(MIXFE_EINVAL, "bad B or T");
    if (B == 0) return MIXFE_OK;
    if (B > 65535) return fail(MIXFE_EINVAL, "B > 65535 clips per call");
    if (stride_c < T || (B > 1 && stride_b < T)) return fail(MIXFE_EINVAL, "strides smaller than T");
    DeviceGuard guard(pl->device);

    AugLaunch al;
    int rc = prepare_augment(params, B, T, kNumStems, al);
    if (rc) return rc;
    std::vector<int> slot_clip;
    for (int b = 0; b < B; ++b) {
        AugClipDev& c = al.clips[b];
        c.reverb_on = params[b].reverb_on ? 1 : 0;
        c.reverb_slot = -1;
        if (c.reverb_on) {
            if (!impulse || L <= 0) return fail(MIXFE_EINVAL, "clip %d wants reverb but no impulse was given", b);
            c.reverb_slot = (int)slot_clip.size();
            slot_clip.push_back(b);
        }
    }
    const int n_chunks = al.n_chunks;

    // ---- workspace -------------------------------------------------------------------------------------
    const int n_slots = (int)slot_clip.size();
    ReverbLayout lay;
    if (n_slots) {
        if (L > kRevMaxN / 2) return fail(MIXFE_EUNSUPPORTED, "impulse of %d taps (max %d)", L, kRevMaxN / 2);
        lay = reverb_layout(T, L);
    }
    const int n_seg = lay.n_seg, seg_len = lay.seg_len;
    size_t off = 0;
    auto carve = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    const size_t o_clips = carve(sizeof(AugClipDev) * B);
    const size_t o_sections = carve(sizeof(AugSection) * al.sections.size());
    const size_t o_slots = carve(sizeof(int) * (size_t)std::max(1, n_slots));
    const size_t o_order = carve(sizeof(int) * (size_t)B * kNumStems);
    const size_t o_tables = off;                                    // [0, o_tables) is uploaded from the host in one copy
    const size_t o_energy = carve(sizeof(double) * (size_t)B * kNumStems * 2 * n_chunks);
    const size_t o_X = carve(sizeof(float2) * (size_t)lay.n * n_slots * std::max(1, n_seg));
    const size_t o_H = carve(sizeof(float2) * (size_t)lay.n * n_slots);
    // The number of reverberated clips is drawn per call: reserve for ALL B clips (or, past 1 GiB, for the count rounded up
    // to 8) so that a call with more of them than any before does not stop the stream to reallocate -- measured as a
    // 2.2 -> 3.7 ms step in a 20-step loop whenever a new maximum fell into it.
    size_t reserve_bytes = off;
    if (impulse && L > 0 && L <= kRevMaxN / 2) {
        const ReverbLayout lw = reverb_layout(T, L);
        const size_t per_slot = sizeof(float2) * (size_t)lw.n * (size_t)(std::max(1, lw.n_seg) + 1) + 512;
        const size_t fixed = o_X + sizeof(int) * (size_t)B;
        const size_t n_res = per_slot * (size_t)B <= ((size_t)1 << 30) ? (size_t)B : (size_t)std::min(B, (n_slots + 7) / 8 * 8);
        reserve_bytes = std::max(reserve_bytes, fixed + per_slot * n_res);
    }
    rc = ensure_workspace(pl, reserve_bytes, stream);
    if (rc) return rc;
    char* ws = reinterpret_cast<char*>(pl->ws);
    ParamUpload up(pl, stream);
    rc = up.reserve(o_tables + sizeof(int) * (size_t)B + 256);
    if (rc) return rc;
    if ((rc = up.stage(o_clips, al.clips.data(), sizeof(AugClipDev) * B))) return rc;
    if ((rc = up.stage(o_sections, al.sections.data(), sizeof(AugSection) * al.sections.size()))) return rc;
    if ((rc = up.stage(o_slots, slot_clip.data(), sizeof(int) * (size_t)n_slots))) return rc;
    // (clip, stem) rows in order of decreasing work: the block scheduler hands out CTAs in grid order, so the rows with
    // the long effect chains start first and the launch ends on cheap copies instead of on a straggler
    static const int k3_mode = std::getenv("MIXFE_K3_SPLIT") ? std::atoi(std::getenv("MIXFE_K3_SPLIT")) : 2;
    const bool sorted = k3_mode == 2 && (long long)B * kNumStems <= 65535;
    if (sorted) {
        std::vector<int> order((size_t)B * kNumStems);
        std::iota(order.begin(), order.end(), 0);
        auto cost = [&](int pair) {
            const int f = al.clips[pair >> 2].flags[pair & 3];
            return ((f & 1) ? 3 : 0) + ((f & 2) ? 2 : 0) + ((f & 4) ? 6 : 0);
        };
        std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return cost(x) > cost(y); });
        if ((rc = up.stage(o_order, order.data(), sizeof(int) * order.size()))) return rc;
    }
    if ((rc = up.upload(ws))) return rc;

    K3Params p3{};
    p3.order = reinterpret_cast<const int*>(ws + o_order);
    for (int s = 0; s < kNumStems; ++s) { p3.in[s] = in[s]; p3.out[s] = out[s]; }
    p3.stride_b = stride_b; p3.stride_c = stride_c; p3.B = B; p3.T = T; p3.chunk = al.chunk; p3.warm = al.warm;
    p3.n_chunks = n_chunks; p3.n_stems = kNumStems; p3.clips = reinterpret_cast<const AugClipDev*>(ws + o_clips);
    p3.sections = reinterpret_cast<const AugSection*>(ws + o_sections);
    p3.part_energy = reinterpret_cast<double*>(ws + o_energy);
    {
        KernelTimer timer(pl, MIXFE_KERNEL_AUGMENT, stream);
        p3.stem_in_y = sorted ? 2 : (k3_mode != 0 ? 1 : 0);
        if (sorted) k3_augment<<<dim3(n_chunks, 2, B * kNumStems), 32, 0, stream>>>(p3);
        else if (p3.stem_in_y) k3_augment<<<dim3(n_chunks, 2 * kNumStems, B), 32, 0, stream>>>(p3);
        else k3_augment<<<dim3(n_chunks, 2, B), kAugWarps * 32, 0, stream>>>(p3);
    }
    MIXFE_CUDA(cudaGetLastError());
    launch_counter()++;
    if (!n_slots) return MIXFE_OK;

    K4Params p4{};
    for (int s = 0; s < kNumStems; ++s) p4.out[s] = out[s];
    p4.stride_b = stride_b; p4.stride_c = stride_c; p4.T = T; p4.L = L; p4.half = L / 2; p4.seg_len = seg_len;
    p4.n_seg = n_seg; p4.n_chunks = n_chunks; p4.slot_clip = reinterpret_cast<const int*>(ws + o_slots);
    p4.standalone = 0; p4.n_ch = 2; p4.sa_in = nullptr; p4.imp_stride = L;
    p4.impulse = impulse; p4.X = reinterpret_cast<float2*>(ws + o_X); p4.H = reinterpret_cast<float2*>(ws + o_H);
    p4.part_energy = p3.part_energy; p4.tw = pl->d_tw; p4.tw_lin = pl->d_twlin; p4.tw_fine = pl->d_twfine;
    return launch_reverb(pl, p4, lay, n_slots, stream);
}

// AudioAugmenter.apply_spectral_tilt / apply_compression / apply_bandwidth_limit (:421-456) and the gain of :389-392 on
// ONE tensor (B, C, T): the decisions of params[b] (stem index 0) are applied to every channel of clip b.
inline int run_effects(mixfe_plan* pl, const float* in, float* out, long long stride_b, long long stride_c, int B, int C,
                       int T, const mixfe_aug_params* params, cudaStream_t stream) {
    if (!pl || !in || !out || !params) return fail(MIXFE_EINVAL, "null argument");
    if (in == out) return fail(MIXFE_EINVAL, "out must not alias in");
    if (B < 0 || C <= 0 || T <= 0) return fail(MIXFE_EINVAL, "bad B, C or T");
    if (B == 0) return MIXFE_OK;
    if (B > 65535 || C > 65535) return fail(MIXFE_EINVAL, "B or C > 65535 per call");
    if ((C > 1 && stride_c < T) || (B > 1 && stride_b < T)) return fail(MIXFE_EINVAL, "strides smaller than T");
    DeviceGuard guard(pl->device);
    AugLaunch al;
    int rc = prepare_augment(params, B, T, 1, al);
    if (rc) return rc;
    size_t off = 0;
    auto carve = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    const size_t o_clips = carve(sizeof(AugClipDev) * B);
    const size_t o_sections = carve(sizeof(AugSection) * al.sections.size());
    rc = ensure_workspace(pl, off, stream);
    if (rc) return rc;
    char* ws = reinterpret_cast<char*>(pl->ws);
    ParamUpload up(pl, stream);
    rc = up.reserve(off);
    if (rc) return rc;
    if ((rc = up.stage(o_clips, al.clips.data(), sizeof(AugClipDev) * B))) return rc;
    if ((rc = up.stage(o_sections, al.sections.data(), sizeof(AugSection) * al.sections.size()))) return rc;
    if ((rc = up.upload(ws))) return rc;
    K3Params p3{};
    for (int s = 0; s < kNumStems; ++s) { p3.in[s] = in; p3.out[s] = out; }
    p3.stride_b = stride_b; p3.stride_c = stride_c; p3.B = B; p3.T = T; p3.chunk = al.chunk; p3.warm = al.warm;
    p3.n_chunks = al.n_chunks; p3.n_stems = 1; p3.clips = reinterpret_cast<const AugClipDev*>(ws + o_clips);
    p3.sections = reinterpret_cast<const AugSection*>(ws + o_sections);
    p3.part_energy = nullptr;
    {
        KernelTimer timer(pl, MIXFE_KERNEL_AUGMENT, stream);
        k3_augment<<<dim3(al.n_chunks, C, B), 32, 0, stream>>>(p3);
    }
    MIXFE_CUDA(cudaGetLastError());
    launch_counter()++;
    return MIXFE_OK;
}

// AudioAugmenter.apply_reverb (:458-479) on ONE tensor (B, C, T), C in {1, 2}: out = 0.7 in + 0.3 corr(in, h_b).
inline int run_reverb(mixfe_plan* pl, const float* in, float* out, long long stride_b, long long stride_c, int B, int C,
                      int T, const float* impulse, long long impulse_stride_b, int L, cudaStream_t stream) {
    if (!pl || !in || !out || !impulse) return fail(MIXFE_EINVAL, "null argument");
    if (in == out) return fail(MIXFE_EINVAL, "out must not alias in");
    if (B < 0 || T <= 0 || L <= 0) return fail(MIXFE_EINVAL, "bad B, T or L");
    if (C != 1 && C != 2) return fail(MIXFE_EINVAL, "C=%d: one or two channels per clip (split wider tensors into pairs)", C);
    if (B == 0) return MIXFE_OK;
    if (B > 65535) return fail(MIXFE_EINVAL, "B > 65535 clips per call");
    if ((C > 1 && stride_c < T) || (B > 1 && stride_b < T)) return fail(MIXFE_EINVAL, "strides smaller than T");
    if (impulse_stride_b != 0 && impulse_stride_b < L) return fail(MIXFE_EINVAL, "impulse stride smaller than L");
    if (L > kRevMaxN / 2) return fail(MIXFE_EUNSUPPORTED, "impulse of %d taps (max %d)", L, kRevMaxN / 2);
    DeviceGuard guard(pl->device);
    const ReverbLayout lay = reverb_layout(T, L);
    const int seg_len = lay.seg_len, n_seg = lay.n_seg;
    std::vector<int> slot_clip(B);
    for (int b = 0; b < B; ++b) slot_clip[b] = b;
    size_t off = 0;
    auto carve = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    const size_t o_slots = carve(sizeof(int) * (size_t)B);
    const size_t o_X = carve(sizeof(float2) * (size_t)lay.n * B * n_seg);
    const size_t o_H = carve(sizeof(float2) * (size_t)lay.n * (impulse_stride_b == 0 ? 1 : B));
    int rc = ensure_workspace(pl, off, stream);
    if (rc) return rc;
    char* ws = reinterpret_cast<char*>(pl->ws);
    ParamUpload up(pl, stream);
    rc = up.reserve(sizeof(int) * (size_t)B);
    if (rc) return rc;
    if ((rc = up.stage(o_slots, slot_clip.data(), sizeof(int) * (size_t)B))) return rc;
    if ((rc = up.upload(ws))) return rc;
    K4Params p4{};
    p4.out[0] = out;
    p4.stride_b = stride_b; p4.stride_c = stride_c; p4.T = T; p4.L = L; p4.half = L / 2; p4.seg_len = seg_len;
    p4.n_seg = n_seg; p4.n_chunks = 0; p4.slot_clip = reinterpret_cast<const int*>(ws + o_slots);
    p4.standalone = 1; p4.n_ch = C; p4.sa_in = in; p4.imp_stride = impulse_stride_b;
    p4.impulse = impulse; p4.X = reinterpret_cast<float2*>(ws + o_X); p4.H = reinterpret_cast<float2*>(ws + o_H);
    p4.part_energy = nullptr; p4.tw = pl->d_tw; p4.tw_lin = pl->d_twlin; p4.tw_fine = pl->d_twfine;
    return launch_reverb(pl, p4, lay, B, stream);
}

}  // namespace mixfe
