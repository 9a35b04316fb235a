// K7: K-weighted, gated loudness (ITU-R BS.1770-4) of the four stems and of their mixture -- the "K-weighted, gated
// LUFS" dynamics feature the project proposal describes and the reference does not implement (it uses the
// un-weighted, un-gated -0.691 + 10 log10(mean x^2 + 1e-10), src/mixing_utils.py:311-318).  SURVEY.md section 8f
// row 4: an EXTENSION with parity unpinned; it is not part of the 64-vector.  Oracle: oracle/bs1770.py.
//
// k7_kweight_energy: one warp per (chunk, channel, clip, signal), signal = stem 0..3 or the mixture
// (((0 + v) + b) + d) + o in fp32, as src/mixing_utils.py:85 forms it).  The two K-weighting biquads (high shelf, RLB
// high-pass; fp64) run as K3's chunked parallel scan (k3_augment.cuh): lane-local zero-state response, warp-wide
// Kogge-Stone scan of the states, output correction; a chunk starts `warm` samples early from zero state with
// r^warm < 1e-13.  The squares of the filtered samples are summed per 100 ms sub-block (a gating block is four
// consecutive sub-blocks): two running bins per 512-sample step, fixed-order warp reductions, no atomics.
// k7_gate: one warp per (clip, signal): block loudness, absolute gate (-70 LUFS), relative gate (-10 LU), LUFS.
// k7_short_iqr: the proposal's dynamic-range descriptor DR_IQR = IQR(L_short(t)) (proposal p.2 section 2.1; not in the
//   reference's code either): short-term loudness over 3 s windows (30 sub-blocks) every 100 ms as EBU Tech 3341
//   defines it, windows under the absolute gate (-70 LUFS) dropped, inter-quartile range q75 - q25 of the rest with
//   linear interpolation between order statistics (numpy.quantile's default).  One CTA per (clip, signal): the windows
//   are formed from the sub-block energies, sorted in shared memory (bitonic), and the two quantiles read off.
#pragma once

#include "biquad_f64.cuh"

namespace mixfe {

constexpr int kLoudSignals = kNumStems + 1;
constexpr int kLoudMaxSubs = 8;            // sub-blocks per chunk
constexpr int kShortSubs = 30;             // sub-blocks per short-term window (3 s)
constexpr int kShortMaxWin = 4096;         // windows per clip the sort holds (6 min 52 s of audio)
constexpr int kShortThreads = 256;

struct K7Params {
    const float* stem[kNumStems];
    long long stride_b, stride_c;
    int B, T, sub, n_sub, spc, warm, n_chunks;       // sub: samples per 100 ms; spc: sub-blocks per chunk
    double sos[2][5];                                // b0 b1 b2 a1 a2: high shelf, high-pass
    double* sub_energy;                              // [B][5][2][n_sub]
    float* out;                                      // [B][5] LUFS
    float* out_iqr;                                  // [B][5] LU (k7_short_iqr)
};

__global__ void __launch_bounds__(kLoudSignals * 32) k7_kweight_energy(const __grid_constant__ K7Params p) {
    __shared__ SecTables tables[kLoudSignals][2];
    __shared__ double bins[kLoudSignals][kLoudMaxSubs];
    const int lane = threadIdx.x & 31, sig = threadIdx.x >> 5;
    const int chunk = blockIdx.x, ch = blockIdx.y, b = blockIdx.z;
    const long long row = (long long)b * p.stride_b + (long long)ch * p.stride_c;
    const long long c0 = (long long)chunk * p.spc * p.sub;
    const long long total = (long long)p.n_sub * p.sub;
    const long long c1 = c0 + (long long)p.spc * p.sub < total ? c0 + (long long)p.spc * p.sub : total;

    build_tables(tables[sig][0], p.sos[0][3], p.sos[0][4], lane);
    build_tables(tables[sig][1], p.sos[1][3], p.sos[1][4], lane);
    if (lane < 2) { tables[sig][lane].carry[0] = 0.0; tables[sig][lane].carry[1] = 0.0; }
    if (lane < kLoudMaxSubs) bins[sig][lane] = 0.0;
    __syncwarp();

    const float* r0 = p.stem[0] + row;
    const float* r1 = p.stem[1] + row;
    const float* r2 = p.stem[2] + row;
    const float* r3 = p.stem[3] + row;
    const float* mine = sig == 0 ? r0 : sig == 1 ? r1 : sig == 2 ? r2 : r3;
    long long start = c0 - p.warm;
    start = start < 0 ? 0 : (start & ~3ll);
    const bool vec_ok = ((reinterpret_cast<unsigned long long>(r0) | reinterpret_cast<unsigned long long>(r1) |
                          reinterpret_cast<unsigned long long>(r2) | reinterpret_cast<unsigned long long>(r3)) & 15ull) == 0;

    for (long long blk = start; blk < c1; blk += kAugBlock) {
        const long long i0 = blk + (long long)lane * kAugV;
        float x[kAugV];
        auto load_row = [&](const float* src, float (&dst)[kAugV]) {
            if (vec_ok && i0 + kAugV <= p.T) {
#pragma unroll
                for (int j = 0; j < kAugV / 4; ++j) {
                    const float4 q = __ldg(reinterpret_cast<const float4*>(src + i0) + j);
                    dst[4 * j] = q.x; dst[4 * j + 1] = q.y; dst[4 * j + 2] = q.z; dst[4 * j + 3] = q.w;
                }
            } else {
#pragma unroll
                for (int n = 0; n < kAugV; ++n) dst[n] = i0 + n < p.T ? __ldg(src + i0 + n) : 0.0f;
            }
        };
        if (sig < kNumStems) {
            load_row(mine, x);
        } else {
            float t[kAugV];
            load_row(r0, x);                               // (0 + v)
            load_row(r1, t);
#pragma unroll
            for (int n = 0; n < kAugV; ++n) x[n] += t[n];
            load_row(r2, t);
#pragma unroll
            for (int n = 0; n < kAugV; ++n) x[n] += t[n];
            load_row(r3, t);
#pragma unroll
            for (int n = 0; n < kAugV; ++n) x[n] += t[n];
        }
        double v[kAugV];
#pragma unroll
        for (int n = 0; n < kAugV; ++n) v[n] = (double)x[n];
        biquad_block64(v, p.sos[0], tables[sig][0], lane);
        biquad_block64(v, p.sos[1], tables[sig][1], lane);
        if (blk + kAugBlock > c0) {                        // the block reaches into the chunk: bin its squares
            const long long first = blk > c0 ? blk : c0;
            const int bA = (int)((first - c0) / p.sub);
            const long long edge = c0 + (long long)(bA + 1) * p.sub;       // first sample of bin bA + 1
            double eA = 0.0, eB = 0.0;
#pragma unroll
            for (int n = 0; n < kAugV; ++n) {
                const long long i = i0 + n;
                const double sq = v[n] * v[n];
                if (i >= c0 && i < c1) {
                    if (i < edge) eA += sq; else eB += sq;
                }
            }
            eA = warp_sum(eA);
            eB = warp_sum(eB);
            if (lane == 0) {
                bins[sig][bA] += eA;
                if (bA + 1 < p.spc) bins[sig][bA + 1] += eB;
            }
        }
    }
    __syncwarp();
    if (lane < p.spc) {
        const int j = chunk * p.spc + lane;
        if (j < p.n_sub) p.sub_energy[(((long long)b * kLoudSignals + sig) * 2 + ch) * p.n_sub + j] = bins[sig][lane];
    }
}

__global__ void __launch_bounds__(kLoudSignals * 32) k7_gate(const K7Params p) {
    const int lane = threadIdx.x & 31, sig = threadIdx.x >> 5, b = blockIdx.x;
    const double* eL = p.sub_energy + ((long long)b * kLoudSignals + sig) * 2 * p.n_sub;
    const double* eR = eL + p.n_sub;
    const int n_blocks = p.n_sub - 3;
    const double inv = 1.0 / (4.0 * (double)p.sub);
    double sL = 0.0, sR = 0.0, cnt = 0.0;
    for (int j = lane; j < n_blocks; j += 32) {
        const double zL = (eL[j] + eL[j + 1] + eL[j + 2] + eL[j + 3]) * inv;
        const double zR = (eR[j] + eR[j + 1] + eR[j + 2] + eR[j + 3]) * inv;
        const double l = -0.691 + 10.0 * log10(zL + zR);
        if (l >= -70.0) { sL += zL; sR += zR; cnt += 1.0; }
    }
    sL = warp_sum(sL); sR = warp_sum(sR); cnt = warp_sum(cnt);
    double lufs = -INFINITY;
    if (cnt > 0.0) {
        const double gamma_r = -0.691 + 10.0 * log10(sL / cnt + sR / cnt) - 10.0;
        sL = sR = cnt = 0.0;
        for (int j = lane; j < n_blocks; j += 32) {
            const double zL = (eL[j] + eL[j + 1] + eL[j + 2] + eL[j + 3]) * inv;
            const double zR = (eR[j] + eR[j + 1] + eR[j + 2] + eR[j + 3]) * inv;
            const double l = -0.691 + 10.0 * log10(zL + zR);
            if (l > gamma_r && l > -70.0) { sL += zL; sR += zR; cnt += 1.0; }
        }
        sL = warp_sum(sL); sR = warp_sum(sR); cnt = warp_sum(cnt);
        if (cnt > 0.0) lufs = -0.691 + 10.0 * log10(sL / cnt + sR / cnt);
    }
    if (lane == 0) p.out[(long long)b * kLoudSignals + sig] = (float)lufs;
}

__global__ void __launch_bounds__(kShortThreads) k7_short_iqr(const K7Params p) {
    __shared__ double val[kShortMaxWin];
    __shared__ int n_valid;
    const int tid = threadIdx.x, sig = blockIdx.x, b = blockIdx.y;
    const double* eL = p.sub_energy + ((long long)b * kLoudSignals + sig) * 2 * p.n_sub;
    const double* eR = eL + p.n_sub;
    const int n_win = p.n_sub - (kShortSubs - 1);
    int N = 2;
    while (N < n_win) N <<= 1;
    if (tid == 0) n_valid = 0;
    __syncthreads();
    const double inv = 1.0 / ((double)kShortSubs * (double)p.sub);
    int mine = 0;
    for (int j = tid; j < N; j += kShortThreads) {
        double l = INFINITY;                                   // padding and gated windows sort to the end
        if (j < n_win) {
            double zL = 0.0, zR = 0.0;
            for (int q = 0; q < kShortSubs; ++q) { zL += eL[j + q]; zR += eR[j + q]; }
            const double v = -0.691 + 10.0 * log10((zL + zR) * inv);
            if (v >= -70.0) { l = v; ++mine; }
        }
        val[j] = l;
    }
    if (mine) atomicAdd(&n_valid, mine);
    __syncthreads();
    for (int k = 2; k <= N; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < N; i += kShortThreads) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const double a = val[i], c = val[ixj];
                    const bool up = (i & k) == 0;
                    if ((a > c) == up) { val[i] = c; val[ixj] = a; }
                }
            }
            __syncthreads();
        }
    }
    if (tid == 0) {
        const int n = n_valid;
        float r = 0.0f;
        if (n >= 2) {
            auto quant = [&](double q) {
                const double pos = q * (double)(n - 1);
                const int i0 = (int)pos;
                const int i1 = i0 + 1 < n ? i0 + 1 : i0;
                const double fr = pos - (double)i0;
                return val[i0] + (val[i1] - val[i0]) * fr;
            };
            r = (float)(quant(0.75) - quant(0.25));
        }
        p.out_iqr[(long long)b * kLoudSignals + sig] = r;
    }
}

}  // namespace mixfe
