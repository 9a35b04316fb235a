// fp64 chunked-scan biquad (transposed direct form II state), used by K7's K-weighting (k7_loudness.cuh): the 38 Hz
// high-pass of BS.1770 has a near-double pole at r = 0.9973 and needs double precision throughout.  (K3 runs its
// Butterworth sections in fp32 inside a lane, see k3_augment.cuh.)
//
// Parallel form (state z = (z1, z2)):  z' = A z + B x,  y = z1 + b0 x,  A = [[-a1, 1], [-a2, 0]].
// A warp walks blocks of 32 lanes x 16 samples; per block each lane
//   1. runs the recurrence over its 16 samples from zero state (y0, end state e),
//   2. joins a warp-wide Kogge-Stone scan of the affine maps z -> M z + e with M = A^16
//      (5 shuffle steps with the precomputed powers M^(2^d)), adds M^(lane+1) * carry,
//   3. corrects its outputs with the true start state: y[n] = y0[n] + (C A^n) z_start.
#pragma once

#include "k3_augment.cuh"

namespace mixfe {

struct SecTables {                         // per warp, per section, in shared memory
    double mp[6][4];                       // M^(2^d), d = 0..5, row-major 2x2
    double h[kAugV][2];                    // C A^n
    double ml[4][32];                      // M^(lane+1), component-major
    double carry[2];                       // state at the start of the block being processed (warp-uniform)
};

__device__ __forceinline__ void dmat2_mul(const double (&a)[4], const double (&b)[4], double (&c)[4]) {
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
    for (int d = 1; d <= kAugLogV; ++d) dmat2_mul(ap[d - 1], ap[d - 1], ap[d]);
    double mp[6][4];
#pragma unroll
    for (int i = 0; i < 4; ++i) mp[0][i] = ap[kAugLogV][i];
#pragma unroll
    for (int d = 1; d < 6; ++d) dmat2_mul(mp[d - 1], mp[d - 1], mp[d]);
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
                dmat2_mul(r, mp[d], t);
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
__device__ __forceinline__ void biquad_block64(double (&v)[kAugV], const double* c /*b0 b1 b2 a1 a2*/,
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

}  // namespace mixfe
