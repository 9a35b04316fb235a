// K3: per-stem augmentation chain  gain -> tilt EQ biquad -> static compressor -> 4th-order low-pass
// as ONE streaming pass, with the IIR recurrences evaluated as chunked parallel scans.
//
// Reference behaviour being replaced (src/mixing_utils.py of the reference):
//   :389-392  gain          audio * 10^(gain_db/20)                       (fp32)
//   :421-433  tilt          scipy.signal.sosfilt(butter(2, ...)), fp64 inside, .float() after
//   :435-447  compression   memoryless dB curve, threshold / ratio arguments (fp32)
//   :449-456  bandwidth     scipy.signal.sosfilt(butter(4, fc, 'low')), two sections, fp64, .float()
// The reference pulls CUDA tensors back to the host for both sosfilt calls (.cpu().numpy(), :431,:454).
//
// Parallel form of a biquad.  The section is split into its FIR half  w[n] = b0 x[n] + b1 x[n-1] + b2 x[n-2]
// (no recursion: the two history samples of a lane's first outputs are the previous lane's last inputs, fetched
// with a shuffle) and its all-pole half  y[n] = w[n] - a1 y[n-1] - a2 y[n-2]  with state s = (y[n-1], y[n-2]):
//   s' = A s + (w, 0),  A = [[-a1, -a2], [1, 0]].
// One warp owns a chunk of one (clip, channel) row and walks it in blocks of 32 lanes x 16 samples (four 128-bit
// loads per lane).  Per block and section each lane
//   1. runs the all-pole recurrence over its 16 samples from ZERO state in fp32 (one dependent FFMA per sample) and
//      keeps only the end state, i.e. its last two outputs,
//   2. joins a warp-wide Kogge-Stone scan of the affine maps s -> M s + e, M = A^16, in fp64 (5 shuffle steps with
//      the precomputed powers M^(2^d); the carry of the previous block enters at lane 0),
//   3. runs the recurrence again from its true start state, in place (no correction tables).
// Why this split: with the FIR half taken on true history, every quantity the fp32 recurrence handles is of the
// size of the OUTPUT.  (Running the whole biquad from zero state and correcting afterwards makes a high-pass see a
// step at every lane boundary: transient and correction are then of the size of the INPUT and cancel, which costs
// the low-frequency rejection of the 2 kHz high-pass its accuracy.)  Butterworth numerators are g (1, +-2, 1): the high-pass FIR is evaluated as
// the second difference g ((x - x1) - (x1 - x2)), exact for slowly varying input.
// fp32 inside a lane, fp64 for everything that lives longer than 16 samples (the scanned states).  Measured against
// scipy's fp64 sosfilt: <= 5e-6 of the output peak for the slowest filter (500 Hz low-pass, poles at r = 0.951);
// tolerance 1e-4.
// Chunks are independent: a warp starts `warm` samples early from zero state; with pole radius r the start-state
// error decays as r^warm (the host sizes `warm` so that r^warm < 1e-10; a chunk that starts at sample 0 is exact).
//
// Compressor (memoryless): a = |x| + 1e-8; above the knee  sign(x) 10^((thr + (20 log10 a - thr)/ratio)/20)
// = sign(x) 2^(c0 + lg2(a)/ratio); below it the reference evaluates sign(x) 10^(log10 a) = sign(x) a up to its own
// fp32 rounding (4e-7).  The knee is passed in the amplitude domain (the smallest fp32 a whose fp32 dB value exceeds
// the threshold), so the mask is a single compare.
#pragma once

#include "plan.cuh"

namespace mixfe {

constexpr int kAugLogV = 4;
constexpr int kAugV = 1 << kAugLogV;       // samples per lane per block
constexpr int kAugBlock = 32 * kAugV;     // 512 samples per warp step
constexpr int kAugWarps = kNumStems;      // warps per CTA, one stem each
constexpr int kAugSections = 3;           // 0 = tilt, 1..2 = band-limit
#ifndef MIXFE_K3_PREFETCH
#define MIXFE_K3_PREFETCH 1               // request block i + 1 before the arithmetic of block i (16 more registers)
#endif
#ifndef MIXFE_K3_CTAS
#define MIXFE_K3_CTAS 6                   // resident CTAs per SM the register allocation aims at
#endif

struct AugSection {                        // one biquad in kernel form, built on the host (run_augment)
    double mp[5][4];                       // M^(2^d), d = 0..4, row-major 2x2, M = A^16
    float mpf[5][4];                       // the same in fp32 (sections whose poles decay fast scan in fp32)
    float b0, b1, b2, a1, a2;              // normalised coefficients
    int form;                              // 0: generic FIR; 1: b = b0 (1, 2, 1); 2: b = b0 (1, -2, 1)
    int scan64;                            // 1: slow poles (|lambda|^16 > 0.2, the 500 Hz low-pass): scan the states in fp64
    int pad;
};
static_assert(sizeof(AugSection) == 272, "AugSection layout");

struct AugClipDev {                        // per-clip decisions in kernel form
    float gain[kNumStems];
    int flags[kNumStems];                  // bit0 tilt, bit1 compress, bit2 band-limit
    float knee[kNumStems];                 // compressor: smallest a = |x| + 1e-8 that is compressed
    float c0[kNumStems], c1[kNumStems];    // compressed magnitude = 2^(c0 + c1 lg2 a)
    int reverb_on;
    int reverb_slot;                       // index into the reverb workspace, -1 when off
    int pad[2];
};

struct K3Params {
    const float* in[kNumStems];
    float* out[kNumStems];
    long long stride_b, stride_c;
    int B, T, chunk, warm, n_chunks, n_stems;
    int stem_in_y;                         // 1: one-warp CTAs, grid.y = 2 * stem + channel; 0: warp = stem, grid.y = channel;
                                           // 2: one-warp CTAs, grid.z indexes `order`, grid.y = channel
    const int* order;                      // [B * 4] (clip * 4 + stem), heaviest rows first (mode 2)
    const AugClipDev* clips;
    const AugSection* sections;            // [B][kNumStems][kAugSections]
    double* part_energy;                   // [B][4][2][n_chunks], or NULL
};

struct AugWarpState {                      // per warp and section, in shared memory
    AugSection sec;
    double carry[2];                       // all-pole state at the start of the block being processed (warp-uniform)
    float xh[2];                           // the section's input one and two samples before the block
    float pad[2];
};

// One biquad over the lane's 16 samples, in place.
__device__ __forceinline__ void biquad_block(float (&v)[kAugV], AugWarpState& ws, int lane) {
    const AugSection& sc = ws.sec;
    const float b0 = sc.b0, na1 = -sc.a1, na2 = -sc.a2;
    const int form = sc.form;
    const double c1 = ws.carry[0], c2 = ws.carry[1];
    float x1 = __shfl_up_sync(kFullMask, v[kAugV - 1], 1), x2 = __shfl_up_sync(kFullMask, v[kAugV - 2], 1);
    if (lane == 0) { x1 = ws.xh[0]; x2 = ws.xh[1]; }
    const float in_last = v[kAugV - 1], in_prev = v[kAugV - 2];
    // FIR half on true history, in place: v <- w
    if (form == 2) {                       // high-pass numerator b0 (1, -2, 1): second difference
        float d1 = x1 - x2;
#pragma unroll
        for (int n = 0; n < kAugV; ++n) {
            const float x = v[n], d0 = x - x1;
            v[n] = b0 * (d0 - d1);
            d1 = d0; x1 = x;
        }
    } else if (form == 1) {                // low-pass numerator b0 (1, 2, 1)
#pragma unroll
        for (int n = 0; n < kAugV; ++n) {
            const float x = v[n];
            v[n] = b0 * fmaf(2.0f, x1, x + x2);
            x2 = x1; x1 = x;
        }
    } else {
        const float b1 = sc.b1, b2 = sc.b2;
#pragma unroll
        for (int n = 0; n < kAugV; ++n) {
            const float x = v[n];
            v[n] = fmaf(b0, x, fmaf(b1, x1, b2 * x2));
            x2 = x1; x1 = x;
        }
    }
    // all-pole half from ZERO state: only the end state (the last two outputs) is kept
    float y1 = 0.0f, y2 = 0.0f;
#pragma unroll
    for (int n = 0; n < kAugV; ++n) {
        const float y = fmaf(na1, y1, fmaf(na2, y2, v[n]));
        y2 = y1; y1 = y;
    }
    // inclusive scan of the lanes' zero-state end states: e_l <- sum_{j<=l} M^(l-j) e_j, the block carry entering at lane 0.
    // fp64 for sections with slow poles; fp32 otherwise (half the shuffles, no conversions: a state error of 6e-8 then
    // dies within a lane or two, measured <= 1e-6 of the output peak for the 2 kHz high-pass and the 4 - 12 kHz low-passes).
    double s1, s2;
    if (sc.scan64) {
        double e1 = (double)y1, e2 = (double)y2;
        if (lane == 0) {
            e1 += sc.mp[0][0] * c1 + sc.mp[0][1] * c2;
            e2 += sc.mp[0][2] * c1 + sc.mp[0][3] * c2;
        }
#pragma unroll
        for (int d = 0; d < 5; ++d) {
            const double t1 = __shfl_up_sync(kFullMask, e1, 1 << d);
            const double t2 = __shfl_up_sync(kFullMask, e2, 1 << d);
            if (lane >= (1 << d)) {
                e1 += sc.mp[d][0] * t1 + sc.mp[d][1] * t2;
                e2 += sc.mp[d][2] * t1 + sc.mp[d][3] * t2;
            }
        }
        s1 = __shfl_up_sync(kFullMask, e1, 1); s2 = __shfl_up_sync(kFullMask, e2, 1);
        if (lane == 0) { s1 = c1; s2 = c2; }
        __syncwarp();                          // every lane has read the old carry and history
        if (lane == 31) {
            ws.carry[0] = e1; ws.carry[1] = e2;
            ws.xh[0] = in_last; ws.xh[1] = in_prev;
        }
    } else {
        const float fc1 = (float)c1, fc2 = (float)c2;
        float e1 = y1, e2 = y2;
        if (lane == 0) {
            e1 = fmaf(sc.mpf[0][0], fc1, fmaf(sc.mpf[0][1], fc2, e1));
            e2 = fmaf(sc.mpf[0][2], fc1, fmaf(sc.mpf[0][3], fc2, e2));
        }
#pragma unroll
        for (int d = 0; d < 5; ++d) {
            const float t1 = __shfl_up_sync(kFullMask, e1, 1 << d);
            const float t2 = __shfl_up_sync(kFullMask, e2, 1 << d);
            if (lane >= (1 << d)) {
                e1 = fmaf(sc.mpf[d][0], t1, fmaf(sc.mpf[d][1], t2, e1));
                e2 = fmaf(sc.mpf[d][2], t1, fmaf(sc.mpf[d][3], t2, e2));
            }
        }
        float f1 = __shfl_up_sync(kFullMask, e1, 1), f2 = __shfl_up_sync(kFullMask, e2, 1);
        if (lane == 0) { f1 = fc1; f2 = fc2; }
        s1 = (double)f1; s2 = (double)f2;
        __syncwarp();
        if (lane == 31) {
            ws.carry[0] = (double)e1; ws.carry[1] = (double)e2;
            ws.xh[0] = in_last; ws.xh[1] = in_prev;
        }
    }
    __syncwarp();
    // all-pole half again, from the lane's true start state (y[-1], y[-2]), in place
    y1 = (float)s1; y2 = (float)s2;
#pragma unroll
    for (int n = 0; n < kAugV; ++n) {
        const float y = fmaf(na1, y1, fmaf(na2, y2, v[n]));
        y2 = y1; y1 = y;
        v[n] = y;
    }
}

__device__ __forceinline__ float compress_static(float x, float knee, float c0, float c1) {
    const float a = fabsf(x) + 1e-8f;
    float mag = a;
    if (a >= knee) {
        float l, m;
        asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(a));
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(m) : "f"(fmaf(c1, l, c0)));
        mag = m;
    }
    return x > 0.0f ? mag : (x < 0.0f ? -mag : (x == 0.0f ? 0.0f : x));      // torch.sign: 0 -> 0, NaN -> NaN
}

__global__ void __launch_bounds__(kAugWarps * 32, MIXFE_K3_CTAS) k3_augment(const __grid_constant__ K3Params p) {
    __shared__ __align__(16) AugWarpState state[kAugWarps][kAugSections];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // one warp per (chunk, channel, clip, stem): the warps of a CTA take the stems of one chunk
    // (stem_in_y: one-warp CTAs, the stem rides in blockIdx.y -- a finished warp then frees its registers at once instead
    // of waiting for the slowest of four stems whose effect chains differ by up to 4x)
    // (stem_in_y == 2: blockIdx.z walks the (clip, stem) rows in the host's order, heaviest effect chains first)
    const int chunk = blockIdx.x;
    const int pair = p.stem_in_y == 2 ? __ldg(p.order + blockIdx.z) : 0;
    const int ch = p.stem_in_y == 1 ? (int)(blockIdx.y & 1) : (int)blockIdx.y;
    const int b = p.stem_in_y == 2 ? pair >> 2 : (int)blockIdx.z;
    const AugClipDev& cp = p.clips[b];
    const long long row = (long long)b * p.stride_b + (long long)ch * p.stride_c;
    const long long c0 = (long long)chunk * p.chunk;
    const long long c1 = c0 + p.chunk < p.T ? c0 + p.chunk : p.T;
    const int s = p.stem_in_y == 2 ? (pair & 3) : (p.stem_in_y ? (int)(blockIdx.y >> 1) : warp);
    if (s >= p.n_stems) return;

    // (selects, not p.in[s]: a runtime index into kernel parameters would copy them to local memory)
    const float* __restrict__ src = (s == 0 ? p.in[0] : s == 1 ? p.in[1] : s == 2 ? p.in[2] : p.in[3]) + row;
    float* __restrict__ dst = (s == 0 ? p.out[0] : s == 1 ? p.out[1] : s == 2 ? p.out[2] : p.out[3]) + row;
    const int flags = cp.flags[s];
    const float gain = cp.gain[s];
    const bool tilt = flags & 1, comp = flags & 2, bw = flags & 4;
    const bool filtered = tilt || bw;
    const float knee = cp.knee[s], cc0 = cp.c0[s], cc1 = cp.c1[s];
    if (filtered) {
        const AugSection* gsec = p.sections + ((long long)b * kNumStems + s) * kAugSections;
        for (int sec = 0; sec < kAugSections; ++sec) {
            if (sec == 0 ? !tilt : !bw) continue;
            const int4* g4 = reinterpret_cast<const int4*>(gsec + sec);
            int4* s4 = reinterpret_cast<int4*>(&state[warp][sec].sec);
            if (lane < (int)(sizeof(AugSection) / 16)) s4[lane] = __ldg(g4 + lane);
            if (lane == 0) {
                state[warp][sec].carry[0] = 0.0; state[warp][sec].carry[1] = 0.0;
                state[warp][sec].xh[0] = 0.0f; state[warp][sec].xh[1] = 0.0f;
            }
        }
    }
    __syncwarp();
    long long start = filtered ? c0 - p.warm : c0;      // c0 and warm are multiples of the block
    if (start < 0) start = 0;
    const bool vec_ok = ((reinterpret_cast<unsigned long long>(src) | reinterpret_cast<unsigned long long>(dst)) & 15ull) == 0;
    double energy = 0.0;
    // the next block's samples are requested before this block's (serial) arithmetic starts
    auto fetch = [&](long long blk, float4 (&q)[kAugV / 4]) {
        const long long i0 = blk + (long long)lane * kAugV;
        if (blk < c1 && vec_ok && i0 + kAugV <= p.T) {
#pragma unroll
            for (int j = 0; j < kAugV / 4; ++j)          // volatile: ptxas must not sink the loads to their first use
                asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
                             : "=f"(q[j].x), "=f"(q[j].y), "=f"(q[j].z), "=f"(q[j].w)
                             : "l"(reinterpret_cast<const float4*>(src + i0) + j));
        }
    };
    float4 qn[kAugV / 4];
    if (MIXFE_K3_PREFETCH) fetch(start, qn);
    for (long long blk = start; blk < c1; blk += kAugBlock) {
        const long long i0 = blk + (long long)lane * kAugV;
        const bool full = vec_ok && i0 + kAugV <= p.T;
        float x[kAugV];
        if (!MIXFE_K3_PREFETCH) fetch(blk, qn);
        if (full) {
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
        if (MIXFE_K3_PREFETCH) fetch(blk + kAugBlock, qn);
        // fp32 values between the stages, where the reference's .float() sits (:433, :456)
        if (tilt) biquad_block(x, state[warp][0], lane);
        if (comp) {
#pragma unroll
            for (int n = 0; n < kAugV; ++n) x[n] = compress_static(x[n], knee, cc0, cc1);
        }
        if (bw) {
#pragma unroll 1
            for (int sec = 1; sec < kAugSections; ++sec) biquad_block(x, state[warp][sec], lane);
        }
        if (blk >= c0) {                                  // warm-up blocks are computed but not stored
            float e0 = 0.0f, e1 = 0.0f;
            if (full) {
#pragma unroll
                for (int j = 0; j < kAugV / 4; ++j)
                    reinterpret_cast<float4*>(dst + i0)[j] = make_float4(x[4 * j], x[4 * j + 1], x[4 * j + 2], x[4 * j + 3]);
#pragma unroll
                for (int n = 0; n < kAugV; n += 2) { e0 = fmaf(x[n], x[n], e0); e1 = fmaf(x[n + 1], x[n + 1], e1); }
            } else {
#pragma unroll
                for (int n = 0; n < kAugV; ++n) {
                    if (i0 + n < p.T) {
                        dst[i0 + n] = x[n];
                        e0 = fmaf(x[n], x[n], e0);
                    }
                }
            }
            energy += (double)(e0 + e1);
        }
    }
    if (p.part_energy) {
        energy = warp_sum(energy);
        if (lane == 0) p.part_energy[(((long long)b * kNumStems + s) * 2 + ch) * p.n_chunks + chunk] = energy;
    }
}

}  // namespace mixfe
