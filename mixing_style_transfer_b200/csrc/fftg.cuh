// Shared-memory FFT of `frames` rows of n complex points, n = 2^log2n in [64, 4096], used by K1G, K6 and K8.
//
// fg_fft: decimation in frequency, in place, result in bit-reversed order (X[k] at position bitrev(k)); fg_fft_dit:
// decimation in time, from bit-reversed to natural order (K6 runs its inverse transform with it, in place on the
// spectrum fg_fft left).  log2n is split into
// passes of 3 or 4 radix-2 stages; in a pass one thread owns the 2^R points that those stages combine (spacing hs),
// holds them in registers, and touches shared memory once to read and once to write them: 3 round trips for 512..4096
// points instead of one per two stages.  The twiddle of stage t for the point at offset j + mm * hs of its half is
//   exp(-2 pi i (j + mm hs) / (2 h_t)) = tw[j * s * 2^t] * exp(-2 pi i mm 2^t / 2^R):
// one table read per stage (R per pass) times a compile-time constant (a sixteenth of a turn).
//
// Rows are padded by one element per sixteen, position p at p + (p >> 4) (row stride n + n/16).  A 64-bit access is
// served half a warp at a time, 16 bank pairs wide: with this padding the 16 contiguous points of a half-warp (loads,
// untangle, overlap-add, the first passes) stay in 16 different bank pairs, and so do the points of the last pass,
// where a thread owns 8 or 16 contiguous points and the lanes are 8 or 16 apart (8 t + t/2, 17 t).  Counted over the
// passes, the untangle and the contiguous walks this is 1.3x to 1.8x fewer wavefronts than one pad per eight.
#pragma once

#include <cuda_runtime.h>

namespace mixfe {

constexpr int kGThreads = 256;

__device__ __forceinline__ int fg_pad(int p) { return p + (p >> 4); }
__host__ __device__ __forceinline__ int fg_row(int n) { return n + (n >> 4); }

// idx / d for 0 <= idx and idx * d < 2^32, d >= 2, with magic = 0xFFFFFFFF / d + 1 (one multiply-high instead of the
// ~20-instruction runtime division).
__device__ __forceinline__ unsigned k1g_magic(int d) { return 0xFFFFFFFFu / (unsigned)d + 1u; }
__device__ __forceinline__ int k1g_div(int idx, unsigned magic) { return (int)__umulhi((unsigned)idx, magic); }

__device__ __forceinline__ float2 cmulf(float2 a, float2 w) {
    return make_float2(fmaf(a.x, w.x, -a.y * w.y), fmaf(a.x, w.y, a.y * w.x));
}

// Rows f = 0 .. frames-1 of z <- (a[j], b[j]) * win[i], i = 0 .. n-1, j = (t_first + f) * hop - n/2 + i folded back
// into [0, T) as torch's 'reflect' padding does (T > n/2).  Tiles whose frames lie inside the signal take the loop
// without the folding and the 64-bit index arithmetic, four points per thread at a time.
__device__ __forceinline__ void fg_load_frames(float2* z, const float* __restrict__ a, const float* __restrict__ b,
                                               const float* __restrict__ win, int log2n, int hop, int T, int t_first,
                                               int frames, int tid) {
    const int n = 1 << log2n, half = n >> 1, zs = fg_row(n);
    const long long start = (long long)t_first * hop - half;
    if (start >= 0 && start + (long long)(frames - 1) * hop + n <= (long long)T) {
        const float* a0 = a + start;
        const float* b0 = b + start;
        for (int f = 0; f < frames; ++f) {
            const float* af = a0 + (size_t)f * hop;
            const float* bf = b0 + (size_t)f * hop;
            float2* d = z + f * zs;
            for (int i = tid; i < n; i += 4 * kGThreads) {        // 12 loads in flight, then the stores
                float w[4], va[4], vb[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int iu = i + u * kGThreads;
                    if (iu < n) { w[u] = __ldg(win + iu); va[u] = __ldg(af + iu); vb[u] = __ldg(bf + iu); }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int iu = i + u * kGThreads;
                    if (iu < n) d[fg_pad(iu)] = make_float2(va[u] * w[u], vb[u] * w[u]);
                }
            }
        }
    } else {
        for (int idx = tid; idx < (frames << log2n); idx += kGThreads) {
            const int f = idx >> log2n, i = idx & (n - 1);
            long long j = start + (long long)f * hop + i;
            if (j < 0) j = -j;
            if (j >= T) j = 2ll * (T - 1) - j;
            const float w = __ldg(win + i);
            z[f * zs + fg_pad(i)] = make_float2(__ldg(a + j) * w, __ldg(b + j) * w);
        }
    }
}

// d * exp(-2 pi i K / 16), K a compile-time constant after unrolling
__device__ __forceinline__ float2 fg_mul_w16(float2 d, int K) {
    constexpr float kC[8] = {1.0f, 0.92387953251128674f, 0.70710678118654752f, 0.38268343236508977f,
                             0.0f, -0.38268343236508977f, -0.70710678118654752f, -0.92387953251128674f};
    constexpr float kS[8] = {0.0f, 0.38268343236508977f, 0.70710678118654752f, 0.92387953251128674f,
                             1.0f, 0.92387953251128674f, 0.70710678118654752f, 0.38268343236508977f};
    if (K == 0) return d;
    if (K == 4) return make_float2(d.y, -d.x);
    if (K == 2) return make_float2(kC[2] * (d.x + d.y), kC[2] * (d.y - d.x));
    if (K == 6) return make_float2(kC[2] * (d.y - d.x), -kC[2] * (d.x + d.y));
    return make_float2(fmaf(d.x, kC[K], d.y * kS[K]), fmaf(d.y, kC[K], -d.x * kS[K]));
}

// One pass of R stages, the first with half-size h = 2^lh, the last with half-size hs = 2^(lh - R + 1).
template <int R>
__device__ __forceinline__ void fg_pass(float2* z, const float2* __restrict__ tw, int log2n, int frames, int lh, int tid) {
    constexpr int P = 1 << R;
    const int lhs = lh - (R - 1), hs = 1 << lhs;
    const int lq = log2n - R;                         // threads per row = n / P
    const int ls = log2n - 1 - lh;                    // table stride of the first stage, n / (2 h)
    const int zs = fg_row(1 << log2n);
    const int total = frames << lq;
    for (int idx = tid; idx < total; idx += kGThreads) {
        const int f = idx >> lq, q = idx & ((1 << lq) - 1);
        const int g = q >> lhs, j = q & (hs - 1);
        const int base = (g << (lh + 1)) + j;
        float2* row = z + f * zs;
        float2 v[P];
        if (lhs >= 4) {
            float2* r = row + fg_pad(base);
            const int sp = hs + (hs >> 4);
#pragma unroll
            for (int m = 0; m < P; ++m) v[m] = r[m * sp];
        } else {
#pragma unroll
            for (int m = 0; m < P; ++m) v[m] = row[fg_pad(base + (m << lhs))];
        }
#pragma unroll
        for (int t = 0; t < R; ++t) {
            const int dm = P >> (t + 1);
            const float2 wj = __ldg(tw + (j << (ls + t)));
#pragma unroll
            for (int a = 0; a < P; ++a) {
                if (a & dm) continue;
                const int b = a + dm, mm = a & (dm - 1);
                const float2 d = make_float2(v[a].x - v[b].x, v[a].y - v[b].y);
                v[a] = make_float2(v[a].x + v[b].x, v[a].y + v[b].y);
                v[b] = fg_mul_w16(cmulf(d, wj), (mm << t) * (16 / P));
            }
        }
        if (lhs >= 4) {
            float2* r = row + fg_pad(base);
            const int sp = hs + (hs >> 4);
#pragma unroll
            for (int m = 0; m < P; ++m) r[m * sp] = v[m];
        } else {
#pragma unroll
            for (int m = 0; m < P; ++m) row[fg_pad(base + (m << lhs))] = v[m];
        }
    }
}

// All passes, a block barrier after each.  Every thread of the block calls this.
__device__ __forceinline__ void fg_fft(float2* z, const float2* __restrict__ tw, int log2n, int frames, int tid) {
    const int passes = (log2n + 3) >> 2, base = log2n / passes, extra = log2n - base * passes;
    int lh = log2n - 1;
    for (int ps = 0; ps < passes; ++ps) {
        // the 4-stage passes go last (1024 points: 3 + 3 + 4): a 4-stage pass in the middle is the one arrangement whose
        // lanes collide two-way (scripts/bank_sim.py); fg_fft_dit mirrors this with its 4-stage passes first
        const int R = base + (ps >= passes - extra ? 1 : 0);
        if (R == 4) fg_pass<4>(z, tw, log2n, frames, lh, tid);
        else fg_pass<3>(z, tw, log2n, frames, lh, tid);
        lh -= R;
        __syncthreads();
    }
}

// Decimation in time: the mirror image.  Input in bit-reversed order (element k at position bitrev(k)), output in
// natural order.  One pass of R stages, the first with half-size h0 = 2^lh0, the last with h0 * 2^(R-1); a thread owns
// the 2^R points base + j + m h0.  Stage t multiplies the upper point of a pair by
//   exp(-2 pi i (j + mm h0) / (2 h0 2^t)) = tw[j * n / (2 h0 2^t)] * exp(-2 pi i mm / 2^(t+1))   before the butterfly.
template <int R>
__device__ __forceinline__ void fg_pass_dit(float2* z, const float2* __restrict__ tw, int log2n, int frames, int lh0, int tid) {
    constexpr int P = 1 << R;
    const int h0 = 1 << lh0;
    const int lq = log2n - R;
    const int zs = fg_row(1 << log2n);
    const int total = frames << lq;
    for (int idx = tid; idx < total; idx += kGThreads) {
        const int f = idx >> lq, q = idx & ((1 << lq) - 1);
        const int g = q >> lh0, j = q & (h0 - 1);
        const int base = (g << (lh0 + R)) + j;
        float2* row = z + f * zs;
        float2 v[P];
        if (lh0 >= 4) {
            float2* r = row + fg_pad(base);
            const int sp = h0 + (h0 >> 4);
#pragma unroll
            for (int m = 0; m < P; ++m) v[m] = r[m * sp];
        } else {
#pragma unroll
            for (int m = 0; m < P; ++m) v[m] = row[fg_pad(base + (m << lh0))];
        }
#pragma unroll
        for (int t = 0; t < R; ++t) {
            const int dm = 1 << t;
            const float2 wj = __ldg(tw + (j << (log2n - 1 - lh0 - t)));
#pragma unroll
            for (int a = 0; a < P; ++a) {
                if (a & dm) continue;
                const int b = a + dm, mm = a & (dm - 1);
                const float2 tb = fg_mul_w16(cmulf(v[b], wj), mm * (8 >> t));
                v[b] = make_float2(v[a].x - tb.x, v[a].y - tb.y);
                v[a] = make_float2(v[a].x + tb.x, v[a].y + tb.y);
            }
        }
        if (lh0 >= 4) {
            float2* r = row + fg_pad(base);
            const int sp = h0 + (h0 >> 4);
#pragma unroll
            for (int m = 0; m < P; ++m) r[m * sp] = v[m];
        } else {
#pragma unroll
            for (int m = 0; m < P; ++m) row[fg_pad(base + (m << lh0))] = v[m];
        }
    }
}

__device__ __forceinline__ void fg_fft_dit(float2* z, const float2* __restrict__ tw, int log2n, int frames, int tid) {
    const int passes = (log2n + 3) >> 2, base = log2n / passes, extra = log2n - base * passes;
    int lh0 = 0;
    for (int ps = 0; ps < passes; ++ps) {
        const int R = base + (ps < extra ? 1 : 0);
        if (R == 4) fg_pass_dit<4>(z, tw, log2n, frames, lh0, tid);
        else fg_pass_dit<3>(z, tw, log2n, frames, lh0, tid);
        lh0 += R;
        __syncthreads();
    }
}

}  // namespace mixfe
