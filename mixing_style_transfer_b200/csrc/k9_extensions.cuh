// K9 / K10: two extensions the reference's proposal and the brief name but the reference's code does not implement
// (VERDICT round 1 items 5 and 6).  PARITY UNPINNED: there is no reference implementation;
// the oracle (oracle/extensions.py) restates the definitions below and the kernels are tested against it.
//
// K9  activity- and band-weighted inter-stem masking (Mixing_Style_Representation_Learning_Proposal.pdf p.2 section 2.4):
//       Delta_i(k,t) = S_i(k,t) - max_{j != i} S_j(k,t),   mu_i(k,t) = sigmoid((beta - Delta_i(k,t)) / tau),
//       M_i = sum_{k,t} a_i(t) A(k) mu_i(k,t) / sum_{k,t} a_i(t) A(k).
//     S_i = channel-mean mel POWER as in the reference's un-weighted form (src/mixing_utils.py:278-306, beta = 0, tau = 1
//     there).  The proposal leaves a_i and A open; this implementation's choices: a_i(t) = E_i(t) / (E_i(t) + E_0), the
//     frame's mel energy E_i(t) = sum_k S_i(k,t) against the energy E_0 of a -60 dBFS white-noise frame (a soft activity
//     gate; exactly 0 under 1e-6 E_0, i.e. -120 dBFS, where the log-mel's 1e-10 floor is all that is left); A(k) = the IEC 61672 A-weighting (as a power ratio, normalised to its maximum) at the centre of mel band k,
//     passed in by the host.  M_i = 0 for a stem that is never active.
//     Input: the log-mel tensor K1 wrote (S is recovered as exp(logmel) - 1e-10); thread = frame, two passes over the bands.
//
// K10 envelope-follower compressor (the brief's "compressor envelope follower as chunked parallel scan"; the reference's
//     apply_compression is memoryless, src/mixing_utils.py:435-447).  Decoupled smooth peak detector (Giannoulis, Massberg,
//     Reiss, "Digital dynamic range compressor design", JAES 2012, eq. 17):
//       p[n] = max(|x[n]|, a_R p[n-1])                 instant attack, exponential release
//       s[n] = a_A s[n-1] + (1 - a_A) p[n]             attack smoothing
//     then the reference's static curve on the level of s: above the threshold the level is thr + (L - thr)/ratio,
//     y[n] = x[n] g(s[n]).  With a_A = a_R = 0 this is the reference's memoryless compressor.
//     Both recurrences are scans: the first over the semiring (max, x) -- the map p -> max(c, a^k p) is closed under
//     composition -- the second affine.  One CTA per (clip, channel) row walks super-blocks of 8 warps x 32 lanes x 16
//     samples: lane-local recurrence from zero state, Kogge-Stone scan over the lanes, a serial pass over the 8 warp
//     totals in shared memory, correction; the carry runs exactly along the whole row (no warm-up).
#pragma once

#include "plan.cuh"

namespace mixfe {

// ------------------------------------------------------------------------------------------------ K9
constexpr int kMaskThreads = 128;

struct K9Params {
    const float* logmel;               // (B, 8, n_mels, F)
    long long lm_stride_b;
    int B, F, n_mels, n_blocks;
    float beta, inv_tau, e0;
    const float* aw;                   // [n_mels] band weights
    double aw_sum;
    float* part;                       // [B][n_blocks][8]: 4 numerators, 4 activity sums
    float* out;                        // [B][4]
};

__global__ void __launch_bounds__(kMaskThreads) k9_masking_partials(const __grid_constant__ K9Params p) {
    __shared__ float red[kMaskThreads / 32][8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.y, t = blockIdx.x * kMaskThreads + tid;
    const float* lm = p.logmel + (long long)b * p.lm_stride_b + t;
    const long long ch = (long long)p.n_mels * p.F;
    float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (t < p.F) {
        float E[4] = {0.f, 0.f, 0.f, 0.f};
        for (int m = 0; m < p.n_mels; ++m) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float s = 0.5f * (__expf(__ldg(lm + (2 * i) * ch + (long long)m * p.F)) + __expf(__ldg(lm + (2 * i + 1) * ch + (long long)m * p.F))) - 1e-10f;
                E[i] += fmaxf(s, 0.0f);
            }
        }
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        for (int m = 0; m < p.n_mels; ++m) {
            float S[4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
                S[i] = fmaxf(0.5f * (__expf(__ldg(lm + (2 * i) * ch + (long long)m * p.F)) + __expf(__ldg(lm + (2 * i + 1) * ch + (long long)m * p.F))) - 1e-10f, 0.0f);
            const float w = __ldg(p.aw + m);
            const float m01 = fmaxf(S[0], S[1]), m23 = fmaxf(S[2], S[3]);
            const float other[4] = {fmaxf(S[1], m23), fmaxf(S[0], m23), fmaxf(m01, S[3]), fmaxf(m01, S[2])};
#pragma unroll
            for (int i = 0; i < 4; ++i)
                acc[i] = fmaf(w, __fdividef(1.0f, 1.0f + __expf(-(p.beta - (S[i] - other[i])) * p.inv_tau)), acc[i]);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float a = E[i] > 1e-6f * p.e0 ? E[i] / (E[i] + p.e0) : 0.0f;       // under -120 dBFS: not active at all
            v[i] = a * acc[i];
            v[4 + i] = a;
        }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = warp_sum(v[i]);
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < 8; ++i) red[warp][i] = v[i];
    }
    __syncthreads();
    if (tid < 8) {
        float s = 0.0f;
        for (int w = 0; w < kMaskThreads / 32; ++w) s += red[w][tid];
        p.part[((long long)b * p.n_blocks + blockIdx.x) * 8 + tid] = s;
    }
}

__global__ void __launch_bounds__(32) k9_masking_finish(const __grid_constant__ K9Params p) {
    const int b = blockIdx.x, lane = threadIdx.x;
    if (lane < 8) {
        double s = 0.0;
        for (int k = 0; k < p.n_blocks; ++k) s += (double)p.part[((long long)b * p.n_blocks + k) * 8 + lane];
        const double den = __shfl_down_sync(0xffu, s, 4, 8);
        if (lane < 4) p.out[b * 4 + lane] = den > 0.0 ? (float)(s / (den * p.aw_sum)) : 0.0f;
    }
}

// ------------------------------------------------------------------------------------------------ K10
constexpr int kEnvV = 16;
constexpr int kEnvWarps = 8;
constexpr int kEnvSuper = kEnvWarps * 32 * kEnvV;          // 4096 samples per CTA step

struct K10Params {
    const float* in;
    float* out;
    long long stride_b, stride_c;
    int B, C, T;
    float a_rel, a_att;                // per-sample coefficients exp(-1 / (tau sr)), 0 = none
    float knee, c0, c1m1;              // gain(a) = a >= knee ? 2^(c0 + c1m1 lg2 a) : 1, a = s + 1e-8
};

struct EnvTables {                     // powers of one coefficient a
    float pn[kEnvV];                   // a^(n+1)
    float pl[32];                      // a^(16 l)
    float ps[5];                       // a^(16 2^d): Kogge-Stone steps
    float pw[kEnvWarps + 1];           // a^(512 w)
};

__device__ __forceinline__ void env_build(EnvTables& t, float a, int tid) {
    // exact repeated squaring in double, rounded once
    auto pw = [a](int e) { return (float)pow((double)a, (double)e); };
    if (tid < kEnvV) t.pn[tid] = pw(tid + 1);
    if (tid < 32) t.pl[tid] = pw(16 * tid);
    if (tid < 5) t.ps[tid] = pw(16 << tid);
    if (tid <= kEnvWarps) t.pw[tid] = pw(512 * tid);
}

__global__ void __launch_bounds__(kEnvWarps * 32) k10_envelope_compress(const __grid_constant__ K10Params p) {
    __shared__ EnvTables trel, tatt;
    __shared__ float wend[2][kEnvWarps];          // per-warp zero-carry end values of the two stages
    __shared__ float carry[2];                    // state at the start of the super-block
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int c = blockIdx.x, b = blockIdx.y;
    const float* __restrict__ src = p.in + (long long)b * p.stride_b + (long long)c * p.stride_c;
    float* __restrict__ dst = p.out + (long long)b * p.stride_b + (long long)c * p.stride_c;
    env_build(trel, p.a_rel, tid);
    env_build(tatt, p.a_att, tid);
    if (tid < 2) carry[tid] = 0.0f;
    __syncthreads();
    const float a_rel = p.a_rel, a_att = p.a_att, one_m = 1.0f - p.a_att;
    for (long long s0 = 0; s0 < p.T; s0 += kEnvSuper) {
        const long long i0 = s0 + (long long)warp * (32 * kEnvV) + lane * kEnvV;
        float x[kEnvV], v[kEnvV];
#pragma unroll
        for (int n = 0; n < kEnvV; ++n) x[n] = i0 + n < p.T ? __ldg(src + i0 + n) : 0.0f;
        // ---- stage 1: p[n] = max(|x[n]|, a_R p[n-1]) --------------------------------------------------
        float st = 0.0f;
#pragma unroll
        for (int n = 0; n < kEnvV; ++n) { st = fmaxf(fabsf(x[n]), a_rel * st); v[n] = st; }
        float e = st;
#pragma unroll
        for (int d = 0; d < 5; ++d) {
            const float o = __shfl_up_sync(kFullMask, e, 1 << d);
            if (lane >= (1 << d)) e = fmaxf(e, trel.ps[d] * o);
        }
        float sl = __shfl_up_sync(kFullMask, e, 1);
        if (lane == 0) sl = 0.0f;
        if (lane == 31) wend[0][warp] = e;
        __syncthreads();
        {
            float cw = trel.pw[warp] * carry[0];                     // carry into this warp
            for (int u = 0; u < warp; ++u) cw = fmaxf(cw, trel.pw[warp - 1 - u] * wend[0][u]);
            sl = fmaxf(sl, trel.pl[lane] * cw);
#pragma unroll
            for (int n = 0; n < kEnvV; ++n) v[n] = fmaxf(v[n], trel.pn[n] * sl);
        }
        // ---- stage 2: s[n] = a_A s[n-1] + (1 - a_A) p[n] ----------------------------------------------
        st = 0.0f;
#pragma unroll
        for (int n = 0; n < kEnvV; ++n) { st = fmaf(a_att, st, one_m * v[n]); v[n] = st; }
        e = st;
#pragma unroll
        for (int d = 0; d < 5; ++d) {
            const float o = __shfl_up_sync(kFullMask, e, 1 << d);
            if (lane >= (1 << d)) e = fmaf(tatt.ps[d], o, e);
        }
        sl = __shfl_up_sync(kFullMask, e, 1);
        if (lane == 0) sl = 0.0f;
        if (lane == 31) wend[1][warp] = e;
        __syncthreads();
        {
            float cw = tatt.pw[warp] * carry[1];
            for (int u = 0; u < warp; ++u) cw = fmaf(tatt.pw[warp - 1 - u], wend[1][u], cw);
            sl = fmaf(tatt.pl[lane], cw, sl);
#pragma unroll
            for (int n = 0; n < kEnvV; ++n) v[n] = fmaf(tatt.pn[n], sl, v[n]);
        }
        // ---- gain computer + store --------------------------------------------------------------------
#pragma unroll
        for (int n = 0; n < kEnvV; ++n) {
            const float a = v[n] + 1e-8f;
            float g = 1.0f;
            if (a >= p.knee) {
                float l;
                asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(a));
                asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(g) : "f"(fmaf(p.c1m1, l, p.c0)));
            }
            if (i0 + n < p.T) dst[i0 + n] = x[n] * g;
        }
        // ---- carry to the next super-block: the states after its last sample ----------------------------
        __syncthreads();                                          // every warp has consumed carry[] and wend[]
        if (tid < 2) {
            const EnvTables& tb = tid == 0 ? trel : tatt;
            float cw = tb.pw[kEnvWarps] * carry[tid];
            for (int u = 0; u < kEnvWarps; ++u)
                cw = tid == 0 ? fmaxf(cw, tb.pw[kEnvWarps - 1 - u] * wend[0][u]) : fmaf(tb.pw[kEnvWarps - 1 - u], wend[1][u], cw);
            carry[tid] = cw;
        }
        __syncthreads();
    }
}

}  // namespace mixfe
