// K3: per-stem augmentation chain  gain -> tilt EQ biquad -> static compressor -> 4th-order low-pass
// as ONE streaming pass, with the IIR recurrences evaluated as chunked parallel scans.
//
// Reference behaviour being replaced (src/mixing_utils.py of the reference):
//   :389-392  gain          audio * 10^(gain_db/20)                       (fp32)
//   :421-433  tilt          scipy.signal.sosfilt(butter(2, ...)), fp64 inside, .float() after
//   :435-447  compression   memoryless dB curve, threshold -20 dB, ratio 4 (fp32)
//   :449-456  bandwidth     scipy.signal.sosfilt(butter(4, fc, 'low')), two sections, fp64, .float()
// The reference pulls CUDA tensors back to the host for both sosfilt calls (.cpu().numpy(), :431,:454).
//
// Parallel form of a biquad (transposed direct form II, state z = (z1, z2)):
//   z' = A z + B x,  y = z1 + b0 x,  A = [[-a1, 1], [-a2, 0]].
// One warp owns a chunk of one (clip, channel) row and walks it in blocks of 32 lanes x 16 samples
// (four 128-bit loads per lane).  Per block and section each lane
//   1. runs the recurrence over its 16 samples from zero state (y0, end state e),
//   2. joins a warp-wide Kogge-Stone scan of the affine maps z -> M z + e with M = A^16
//      (5 shuffle steps with the precomputed powers M^(2^d)), adds M^(lane+1) * carry,
//   3. corrects its outputs with the true start state: y[n] = y0[n] + (C A^n) z_start.
// Chunks are independent: a warp starts `warm` samples early from zero state; with pole radius r
// the start-state error decays as r^warm (the host sizes `warm` so that r^warm < 1e-13; a chunk
// that starts at sample 0 is exact).  Everything inside the filters is fp64 like scipy's.
#pragma once

#include "plan.cuh"

namespace mixfe {

constexpr int kAugLogV = 4;
constexpr int kAugV = 1 << kAugLogV;       // samples per lane per block
constexpr int kAugBlock = 32 * kAugV;     // 512 samples per warp step
constexpr int kAugWarps = kNumStems;      // warps per CTA, one stem each

struct AugClipDev {                        // per-clip decisions in kernel form
    float gain[kNumStems];
    int flags[kNumStems];                  // bit0 tilt, bit1 compress, bit2 band-limit
    double sos[kNumStems][3][5];           // normalised b0 b1 b2 a1 a2: section 0 = tilt, 1..2 = band-limit
    int reverb_on;
    int reverb_slot;                       // index into the reverb workspace, -1 when off
};

struct K3Params {
    const float* in[kNumStems];
    float* out[kNumStems];
    long long stride_b, stride_c;
    int B, T, chunk, warm, n_chunks;
    const AugClipDev* clips;
    double* part_energy;                   // [B][4][2][n_chunks]
};

struct SecTables {                         // per warp, per section, in shared memory
    double mp[6][4];                       // M^(2^d), d = 0..5, row-major 2x2
    double h[kAugV][2];                    // C A^n
    double ml[4][32];                      // M^(lane+1), component-major
    double carry[2];                       // state at the start of the block being processed (warp-uniform)
};

__device__ __forceinline__ void mat2_mul(const double (&a)[4], const double (&b)[4], double (&c)[4]) {
    c[0] = a[0] * b[0] + a[1] * b[2];
    c[1] = a[0] * b[1] + a[1] * b[3];
    c[2] = a[2] * b[0] + a[3] * b[2];
    c[3] = a[2] * b[1] + a[3] * b[3];
}

// Every lane computes the (tiny) tables redundantly in registers; lanes then store their share.
__device__ __forceinline__ void build_tables(SecTables& tb, double a1, double a2, int lane) {
    double ap[kAugLogV + 1][4];             // A^(2^d), d = 0..log2 V  (A^V = M)
    ap[0][0] = -a1; ap[0][1] = 1.0; ap[0][2] = -a2; ap[0][3] = 0.0;
#pragma unroll
    for (int d = 1; d <= kAugLogV; ++d) mat2_mul(ap[d - 1], ap[d - 1], ap[d]);
    double mp[6][4];
#pragma unroll
    for (int i = 0; i < 4; ++i) mp[0][i] = ap[kAugLogV][i];
#pragma unroll
    for (int d = 1; d < 6; ++d) mat2_mul(mp[d - 1], mp[d - 1], mp[d]);
    if (lane < 6) {
#pragma unroll
        for (int d = 0; d < 6; ++d)
            if (lane == d) {
#pragma unroll
                for (int i = 0; i < 4; ++i) tb.mp[d][i] = mp[d][i];
            }
    }
    // M^(lane+1)
    {
        double r[4] = {1.0, 0.0, 0.0, 1.0}, t[4];
        const int e = lane + 1;
#pragma unroll
        for (int d = 0; d < 6; ++d) {
            if ((e >> d) & 1) {
                mat2_mul(r, mp[d], t);
#pragma unroll
                for (int i = 0; i < 4; ++i) r[i] = t[i];
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) tb.ml[i][lane] = r[i];
    }
    // h[n] = first row of A^n, n = lane < V
    if (lane < kAugV) {
        double r0 = 1.0, r1 = 0.0;
#pragma unroll
        for (int d = 0; d < kAugLogV; ++d) {
            if ((lane >> d) & 1) {
                const double n0 = r0 * ap[d][0] + r1 * ap[d][2];
                const double n1 = r0 * ap[d][1] + r1 * ap[d][3];
                r0 = n0;
                r1 = n1;
            }
        }
        tb.h[lane][0] = r0;
        tb.h[lane][1] = r1;
    }
}

// One biquad over the lane's 16 samples (in place) with the warp-wide state scan.
// Zero-state response in direct form, w[n] = b0 x[n] + b1 x[n-1] + b2 x[n-2] (no dependences) and
// y[n] = (w[n] - a2 y[n-2]) - a1 y[n-1]: ONE dependent DFMA per sample instead of the two of the DF2T update; the
// DF2T state the scan works on is rebuilt from the last two inputs and outputs.
__device__ __forceinline__ void biquad_block(double (&v)[kAugV], const double* c /*b0 b1 b2 a1 a2*/,
                                             SecTables& tb, int lane) {
    const double b0 = c[0], b1 = c[1], b2 = c[2], a1 = c[3], a2 = c[4];
    const double carry1 = tb.carry[0], carry2 = tb.carry[1];
    double xm1 = 0.0, xm2 = 0.0, ym1 = 0.0, ym2 = 0.0;
#pragma unroll
    for (int n = 0; n < kAugV; ++n) {
        const double x = v[n];
        const double w = fma(b0, x, fma(b1, xm1, b2 * xm2));
        const double y = fma(-a1, ym1, fma(-a2, ym2, w));
        xm2 = xm1; xm1 = x;
        ym2 = ym1; ym1 = y;
        v[n] = y;
    }
    // DF2T state after the last sample: z1 = b1 x[N-1] - a1 y[N-1] + (b2 x[N-2] - a2 y[N-2]), z2 = b2 x[N-1] - a2 y[N-1]
    double z1 = fma(b1, xm1, fma(-a1, ym1, fma(b2, xm2, -a2 * ym2)));
    double z2 = fma(b2, xm1, -a2 * ym1);
    // inclusive scan of end states: e_l <- sum_{j<=l} M^(l-j) e_j
#pragma unroll
    for (int d = 0; d < 5; ++d) {
        const double t1 = __shfl_up_sync(kFullMask, z1, 1 << d);
        const double t2 = __shfl_up_sync(kFullMask, z2, 1 << d);
        if (lane >= (1 << d)) {
            z1 += tb.mp[d][0] * t1 + tb.mp[d][1] * t2;
            z2 += tb.mp[d][2] * t1 + tb.mp[d][3] * t2;
        }
    }
    z1 += tb.ml[0][lane] * carry1 + tb.ml[1][lane] * carry2;
    z2 += tb.ml[2][lane] * carry1 + tb.ml[3][lane] * carry2;
    double s1 = __shfl_up_sync(kFullMask, z1, 1);
    double s2 = __shfl_up_sync(kFullMask, z2, 1);
    if (lane == 0) { s1 = carry1; s2 = carry2; }
    __syncwarp();                                        // every lane has read the old carry
    if (lane == 31) { tb.carry[0] = z1; tb.carry[1] = z2; }
    __syncwarp();
#pragma unroll
    for (int n = 0; n < kAugV; ++n) v[n] = fma(tb.h[n][0], s1, fma(tb.h[n][1], s2, v[n]));
}

__device__ __forceinline__ float compress_static(float x) {
    // db = 20 log10(|x| + 1e-8); above -20 dB: -20 + (db + 20)/4; y = sign(x) 10^(db/20)
    float db = 20.0f * log10f(fabsf(x) + 1e-8f);
    if (db > -20.0f) db = -20.0f + (db - (-20.0f)) / 4.0f;
    const float mag = exp10f(db / 20.0f);
    return x > 0.0f ? mag : (x < 0.0f ? -mag : (x == 0.0f ? 0.0f * mag : x));     // torch.sign: 0 -> 0, NaN -> NaN
}

__global__ void __launch_bounds__(kAugWarps * 32, 4) k3_augment(const __grid_constant__ K3Params p) {
    __shared__ SecTables tables[kAugWarps][3];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // one warp per (chunk, channel, clip, stem): the four warps of a CTA take the four stems of one chunk
    // (4x the warps of a stem-sequential walk; the scan is latency-bound and needs them)
    const int chunk = blockIdx.x;
    const int ch = blockIdx.y, b = blockIdx.z;
    const AugClipDev& cp = p.clips[b];
    const long long row = (long long)b * p.stride_b + (long long)ch * p.stride_c;
    const long long c0 = (long long)chunk * p.chunk;
    const long long c1 = c0 + p.chunk < p.T ? c0 + p.chunk : p.T;

    {
        const int s = warp;
        // (selects, not p.in[s]: a runtime index into kernel parameters would copy them to local memory)
        const float* __restrict__ src = (s == 0 ? p.in[0] : s == 1 ? p.in[1] : s == 2 ? p.in[2] : p.in[3]) + row;
        float* __restrict__ dst = (s == 0 ? p.out[0] : s == 1 ? p.out[1] : s == 2 ? p.out[2] : p.out[3]) + row;
        const int flags = cp.flags[s];
        const float gain = cp.gain[s];
        const bool tilt = flags & 1, comp = flags & 2, bw = flags & 4;
        const bool filtered = tilt || bw;
        __syncwarp();
        if (tilt) build_tables(tables[warp][0], cp.sos[s][0][3], cp.sos[s][0][4], lane);
        if (bw) {
            build_tables(tables[warp][1], cp.sos[s][1][3], cp.sos[s][1][4], lane);
            build_tables(tables[warp][2], cp.sos[s][2][3], cp.sos[s][2][4], lane);
        }
        __syncwarp();
        long long start = filtered ? c0 - p.warm : c0;      // c0 and warm are multiples of the block
        if (start < 0) start = 0;
        const bool vec_ok = ((reinterpret_cast<unsigned long long>(src) | reinterpret_cast<unsigned long long>(dst)) & 15ull) == 0;
        if (lane < 3) { tables[warp][lane].carry[0] = 0.0; tables[warp][lane].carry[1] = 0.0; }
        __syncwarp();
        double energy = 0.0;
        // the next block's samples are requested before this block's (long, serial) arithmetic starts
        auto fetch = [&](long long blk, float4 (&q)[kAugV / 4]) {
            const long long i0 = blk + (long long)lane * kAugV;
            if (blk < c1 && vec_ok && i0 + kAugV <= p.T) {
#pragma unroll
                for (int j = 0; j < kAugV / 4; ++j)          // volatile: ptxas must not sink the loads to their first use
                    asm volatile("ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%4];"
                                 : "=f"(q[j].x), "=f"(q[j].y), "=f"(q[j].z), "=f"(q[j].w)
                                 : "l"(reinterpret_cast<const float4*>(src + i0) + j));
            }
        };
        float4 qn[kAugV / 4];
        fetch(start, qn);
        for (long long blk = start; blk < c1; blk += kAugBlock) {
            const long long i0 = blk + (long long)lane * kAugV;
            float x[kAugV];
            if (vec_ok && i0 + kAugV <= p.T) {
#pragma unroll
                for (int j = 0; j < kAugV / 4; ++j) {
                    x[4 * j] = qn[j].x; x[4 * j + 1] = qn[j].y; x[4 * j + 2] = qn[j].z; x[4 * j + 3] = qn[j].w;
                }
            } else {
#pragma unroll
                for (int n = 0; n < kAugV; ++n) x[n] = i0 + n < p.T ? __ldg(src + i0 + n) : 0.0f;
            }
            // consume this block's loads BEFORE the next block's are issued: a wait on the loads' scoreboard would
            // otherwise also wait for the new ones and expose the full memory latency every block
#pragma unroll
            for (int n = 0; n < kAugV; ++n) x[n] *= gain;
#pragma unroll
            for (int n = 0; n < kAugV; n += 4)
                asm volatile("" : "+f"(x[n]), "+f"(x[n + 1]), "+f"(x[n + 2]), "+f"(x[n + 3]));
            fetch(blk + kAugBlock, qn);
            if (filtered) {
                // one copy of the biquad in the instruction stream: sections 0 (tilt), 1 and 2 (band-limit), the
                // compressor between 0 and 1; fp32 rounding where the reference's .float() sits (:433, :456)
                double v[kAugV];
#pragma unroll
                for (int n = 0; n < kAugV; ++n) v[n] = (double)x[n];
#pragma unroll 1
                for (int sec = 0; sec < 3; ++sec) {
                    if (sec == 1 && comp) {
#pragma unroll
                        for (int n = 0; n < kAugV; ++n) v[n] = (double)compress_static((float)v[n]);
                    }
                    if (sec == 0 ? tilt : bw) {
                        biquad_block(v, cp.sos[s][sec], tables[warp][sec], lane);
                        if (sec == 0) {
#pragma unroll
                            for (int n = 0; n < kAugV; ++n) v[n] = (double)(float)v[n];
                        }
                    }
                }
#pragma unroll
                for (int n = 0; n < kAugV; ++n) x[n] = (float)v[n];
            } else if (comp) {
#pragma unroll
                for (int n = 0; n < kAugV; ++n) x[n] = compress_static(x[n]);
            }
            if (blk >= c0) {                                  // warm-up blocks are computed but not stored
                if (vec_ok && i0 + kAugV <= p.T) {
#pragma unroll
                    for (int j = 0; j < kAugV / 4; ++j)
                        reinterpret_cast<float4*>(dst + i0)[j] = make_float4(x[4 * j], x[4 * j + 1], x[4 * j + 2], x[4 * j + 3]);
#pragma unroll
                    for (int n = 0; n < kAugV; ++n) energy = fma((double)x[n], (double)x[n], energy);
                } else {
#pragma unroll
                    for (int n = 0; n < kAugV; ++n) {
                        if (i0 + n < p.T) {
                            dst[i0 + n] = x[n];
                            energy = fma((double)x[n], (double)x[n], energy);
                        }
                    }
                }
            }
        }
        energy = warp_sum(energy);
        if (lane == 0) p.part_energy[(((long long)b * kNumStems + s) * 2 + ch) * p.n_chunks + chunk] = energy;
    }
}

}  // namespace mixfe
