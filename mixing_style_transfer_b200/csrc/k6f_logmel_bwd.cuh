// K6F: backward of the log-mel front end for the (n_fft, hop) = (1024, 256) configuration -- the one the style-transfer
// loop differentiates (src/train_style_transfer.py:206-215) -- on K1's one-warp 1024-point transform.
// (k6_logmel_bwd.cuh stays the path of every other configuration; same mathematics, see its header.)
//
//   dmel[m]  = g[m] / (mel[m] + 1e-10) = g[m] exp(-logmel[m])         the forward's OUTPUT is an input here: no mel pass
//   dP[k]    = sum_m fb[k][m] dmel[m]                                  <= 2 adjacent filters per bin (HTK triangles)
//   G[k]     = 2 X[k] dP[k] per channel, X recomputed (frame load, window, FFT as in K1)
//   dframe   = hann * Re / Im of the inverse transform of the Hermitian extension of G_L + i G_R
//   dx       = overlap-add of the frames, adjoint of the 'reflect' padding
//
// Every gradient sample is written ONCE: a CTA owns 13 consecutive hop regions (3328 samples) of one (clip, stem) and
// computes the 16 frames that touch them (one frame per warp, 16 warps), so the overlap-add is complete in shared memory
// before anything goes to global memory -- no zero-fill of the gradients, no atomics, 16/13 of the frame arithmetic.
// Only the 513 samples at either end of a row, where the reflect padding folds a second padded position onto the same
// sample (possibly owned by another CTA), take two commutative atomicAdd contributions into a strip the launcher zeroes.
//
// The inverse transform reuses the forward four-step verbatim: the warp holds the spectrum bin lane + 32 k2 in register
// slot k2, which is exactly the input layout (element 32 n1 + lane in slot n1) of the same routine applied to the
// conjugated spectrum; its output, element lane + 32 j in slot j, is the frame sample the lane loaded in the first place.
#pragma once

#include "k1_features.cuh"

namespace mixfe {

constexpr int kBwWarps = 16;                      // frames per CTA
constexpr int kBwOwn = 13;                        // hop regions a CTA owns
constexpr int kBwThreads = kBwWarps * 32;
constexpr int kBwSpan = (kBwWarps - 1) * 256 + 1024;          // 4864 positions covered by the 16 frames
constexpr int kBwEdge = 513;                      // samples at either end of a row that take atomics

struct K6FParams {
    const float* stem[kNumStems];    // forward inputs
    long long stride_b, stride_c;
    float* out[kNumStems];           // gradients (written once; only the edge strips must be zero on entry)
    long long ostride_b, ostride_c;
    const float* g;                  // d loss / d logmel, (b, ch, m, f) at g[b * g_stride_b + (ch * n_mels + m) * F + f]
    long long g_stride_b;
    const float* lm;                 // the forward's log-mel output, same shape
    long long lm_stride_b;
    int B, T, F, n_mels, n_tiles;
    const float2* tw;                // [1024] 32 x 32 twiddles
    const float4* binw;              // [513] per bin: (first filter as int bits, weight of it, weight of the next, -)
    int off_ex, off_tw, off_binw, off_dmel, off_ola;
};

// zero the strips the reflect padding folds onto
__global__ void k6f_zero_edges(const K6FParams p) {
    const int row = blockIdx.x, s = blockIdx.y;          // row = 2 b + channel
    float* dst = p.out[s] + (long long)(row >> 1) * p.ostride_b + (long long)(row & 1) * p.ostride_c;
    const int n = p.T < kBwEdge ? p.T : kBwEdge;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        dst[i] = 0.0f;
        dst[p.T - 1 - i] = 0.0f;
    }
}

__global__ void __launch_bounds__(kBwThreads, 1) k6f_logmel_bwd(const __grid_constant__ K6FParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char* const ex = smem + p.off_ex;                                  // 16 warp regions (exchange)
    float2* const tw = reinterpret_cast<float2*>(smem + p.off_tw);              // [31][32], rows k1 = 1..31
    float4* const binw = reinterpret_cast<float4*>(smem + p.off_binw);          // [513]
    float2* const dmel = reinterpret_cast<float2*>(smem + p.off_dmel);          // [16][n_mels + 1] (L, R)
    float2* const ola = reinterpret_cast<float2*>(smem + p.off_ola);            // [kBwSpan] (L, R)

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x, b = blockIdx.y, s = blockIdx.z;
    const int h_first = -2 + kBwOwn * tile;              // first owned hop region (padded coordinates)
    const int t_first = h_first - 1;                     // frame of warp 0
    const long long p0 = 256ll * t_first - 512;          // padded position of ola[0]
    const int dm_stride = p.n_mels + 1;

    for (int i = tid; i < 31 * 32; i += kBwThreads) tw[i] = p.tw[32 + i];
    for (int i = tid; i < kNfreq; i += kBwThreads) binw[i] = p.binw[i];
    // ---- this warp's frame -> registers with torch's 'reflect' indexing.  Issued BEFORE the dmel stage so that the 64
    //      loads fly under it (one CTA per SM: nothing else hides their latency) -----------------------------------------
    const int t = t_first + warp;                        // this warp's frame
    const bool active = t >= 0 && t < p.F;
    float fl[32], fr[32];
    if (active) {
        const float* gL = p.stem[s] + (long long)b * p.stride_b;
        const float* gR = gL + p.stride_c;
        const long long base = 256ll * t - 512 + lane;
        const bool inside = base - lane >= 0 && base - lane + 1024 <= (long long)p.T;
        if (inside) {                                    // interior frame: one pointer per row, immediate offsets
            const float* const pl = gL + base;
            const float* const pr = gR + base;
#pragma unroll
            for (int n1 = 0; n1 < 32; ++n1) {
                fl[n1] = __ldg(pl + 32 * n1);
                fr[n1] = __ldg(pr + 32 * n1);
            }
        } else {
#pragma unroll
            for (int n1 = 0; n1 < 32; ++n1) {
                long long j = base + 32 * n1;
                if (j < 0) j = -j;
                if (j >= p.T) j = 2ll * (p.T - 1) - j;
                j = j < 0 ? 0 : (j >= p.T ? p.T - 1 : j);
                fl[n1] = __ldg(gL + j);
                fr[n1] = __ldg(gR + j);
            }
        }
    }
    // ---- dmel of the 16 frames: thread = (frame, filter) with the frame fastest (64-byte runs along time) ----------
    {
        const float* const gbL = p.g + (long long)b * p.g_stride_b + (long long)(2 * s) * p.n_mels * p.F;
        const float* const lbL = p.lm + (long long)b * p.lm_stride_b + (long long)(2 * s) * p.n_mels * p.F;
        const float* const gbR = gbL + (long long)p.n_mels * p.F;
        const float* const lbR = lbL + (long long)p.n_mels * p.F;
        for (int idx = tid; idx < kBwWarps * dm_stride; idx += kBwThreads) {
            const int f = idx & (kBwWarps - 1), m = idx >> 4;
            const int tf = t_first + f;
            float2 v = make_float2(0.0f, 0.0f);
            if (m < p.n_mels && tf >= 0 && tf < p.F) {
                const int o = m * p.F + tf;                              // one (clip, channel) plane: fits an int
                v.x = __ldg(gbL + o) * __expf(-__ldg(lbL + o));
                v.y = __ldg(gbR + o) * __expf(-__ldg(lbR + o));
            }
            dmel[f * dm_stride + m] = v;
        }
    }
    __syncthreads();

    float2* const exw = reinterpret_cast<float2*>(ex + warp * kWarpRegion);
    cplx64 z[32];
    if (active) {
        // ---- Hann window on the fly (as K1) -----------------------------------------------------------------------------
        const float2 wl = tw[lane];
        const float cb = wl.x, sb = -wl.y;
#pragma unroll
        for (int n1 = 0; n1 < 32; ++n1) {
            const float ca = n1 < 16 ? w32_re(n1) : -w32_re(n1 - 16);
            const float sa = n1 < 16 ? w32_sin(n1) : -w32_sin(n1 - 16);
            const float wv = fmaf(0.25f * sa, sb, fmaf(-0.25f * ca, cb, 0.25f));          // 0.5 * hann[32 n1 + lane]
            z[n1] = cpack(fl[n1] * wv, fr[n1] * wv);
        }
    }
    // two transforms: forward (pass pair 0), then, after the gradient stage, the same routine on the conjugated spectrum
#pragma unroll 1
    for (int dir = 0; dir < 2; ++dir) {
        if (active) {
#pragma unroll 1
            for (int pass = 0; pass < 2; ++pass) {
                dft32_dif_packed(z);
                if (pass == 0) {
                    const unsigned twa = smem_u32(tw + lane) - 256;
                    const unsigned wa = smem_u32(exw) + lane * 8;
#pragma unroll
                    for (int k1 = 0; k1 < 32; ++k1) {
                        if (k1 == 0) {
                            sts_c<0>(wa, z[0]);
                        } else {
                            float yr, yi;
                            cunpack(z[brev5(k1)], yr, yi);
                            const float2 tq = lds64<0>(twa + k1 * 256);
                            sts64<0>(wa + k1 * kExRow, fmaf(yr, tq.x, -yi * tq.y), fmaf(yr, tq.y, yi * tq.x));
                        }
                    }
                    __syncwarp();
                    const unsigned ra = smem_u32(exw) + lane * kExRow;
#pragma unroll
                    for (int n2 = 0; n2 < 32; ++n2) z[n2] = lds_c<0>(ra + n2 * 8);
                    __syncwarp();
                }
            }
            if (dir == 0) {
                // ---- gradient stage.  Slot k2 (register brev5(k2)) holds X[kap], kap = lane + 32 k2; its partner
                //      X[1024 - kap] sits in lane (32 - lane) & 31, slot 31 - k2 (lane 0: slot (32 - k2) & 31).  With
                //      x = X[k], y = X[1024 - k], k <= 512:  L = xa + i xb, R = xd - i xc (k6_logmel_bwd.cuh);
                //      G_L = 2 L dP_L, G_R = 2 R dP_R; the inverse transform's conjugated input is
                //      (a - di, -(bi + c)) / 2 at k and (a + di, -(c - bi)) / 2 at 1024 - k; (a, -c) at k = 0 and 512.
                const int src = (32 - lane) & 31;
                const bool l0 = lane == 0;
                const float2* dm = dmel + warp * dm_stride;
                // One PAIR of bins per evaluation: o = X[kk] (kk <= 512), q = X[1024 - kk]  ->  the conjugated inverse-
                // transform inputs at kk (lo) and at 1024 - kk (hi).  With xa.. as above and the factors 2 and 1/2 folded:
                //   lo = ( xa dL + xc dR, -(xb dL + xd dR) ),   hi = ( xa dL - xc dR, xb dL - xd dR ).
                auto grad_pair = [&](int kk, cplx64 o, cplx64 q, cplx64& lo, cplx64& hi) {
                    float xx, xy, yx, yy;
                    cunpack(o, xx, xy);
                    cunpack(q, yx, yy);
                    const float4 bw = binw[kk];
                    const int m0 = __float_as_int(bw.x);
                    const float2 d0 = dm[m0], d1 = dm[m0 + 1];
                    const float dL = fmaf(bw.y, d0.x, bw.z * d1.x), dR = fmaf(bw.y, d0.y, bw.z * d1.y);
                    const float xa = xx + yx, xb = xy - yy, xc = xx - yx, xd = xy + yy;
                    const float pa = xa * dL, pb = xb * dL;
                    lo = cpack(fmaf(xc, dR, pa), -fmaf(xd, dR, pb));
                    hi = cpack(fmaf(-xc, dR, pa), fmaf(-xd, dR, pb));
                };
                auto shfl_c = [&](cplx64 v) -> cplx64 {
                    float x, y;
                    cunpack(v, x, y);
                    return cpack(__shfl_sync(kFullMask, x, src), __shfl_sync(kFullMask, y, src));
                };
                auto twice = [](cplx64 v) -> cplx64 {
                    float x, y;
                    cunpack(v, x, y);
                    return cpack(2.0f * x, 2.0f * y);
                };
                // Slot k2 <= 15 of every lane is the lower member of a pair; the upper member sits in the partner lane's
                // slot 31 - k2 and receives `hi` by shuffle.  Lane 0 pairs with itself, slot k2 with slot 32 - k2: it feeds
                // its own slot 32 - k2 as the partner value and files what comes back one slot later (carry); its slots 0
                // and 16 (bins 0 and 512) are their own partners and take twice the `lo` value.
                cplx64 zn[32];
                cplx64 carry, unused;
                grad_pair(512, z[brev5(16)], z[brev5(16)], carry, unused);
                carry = twice(carry);
#pragma unroll
                for (int k2 = 15; k2 >= 0; --k2) {
                    const cplx64 o = z[brev5(k2)];
                    const cplx64 q = shfl_c(l0 ? z[brev5((32 - k2) & 31)] : z[brev5(31 - k2)]);
                    cplx64 lo, hi;
                    grad_pair(lane + 32 * k2, o, q, lo, hi);
                    const cplx64 recv = shfl_c(hi);
                    zn[k2] = (k2 == 0 && l0) ? twice(lo) : lo;
                    zn[31 - k2] = l0 ? carry : recv;
                    carry = recv;
                }
                // natural slot order for the second transform
#pragma unroll
                for (int k2 = 0; k2 < 32; ++k2) z[k2] = zn[k2];
            }
        }
    }
    // ---- window, overlap-add.  Quarter q of frame w lands in hop region w + q, so in turn q every warp adds its q-th
    //      quarter and no two warps meet (all 16 warps busy in every turn; the first cut let 4 of 16 work per turn).
    //      Turn 0 is the first writer of every region and stores (no zero-fill); regions outside the 13 owned ones
    //      (3 .. 15) are never read and are skipped. -----------------------------------------------------------------------
    {
        const float2 wl = tw[lane];
        const float cb = wl.x, sb = -wl.y;
        float2* const dst = ola + 256 * warp + lane;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int h = warp + q;
            if (h >= 3 && h < 3 + kBwOwn) {
#pragma unroll
                for (int j = 8 * q; j < 8 * q + 8; ++j) {
                    const float ca = j < 16 ? w32_re(j) : -w32_re(j - 16);
                    const float sa = j < 16 ? w32_sin(j) : -w32_sin(j - 16);
                    const float hw = fmaf(0.5f * sa, sb, fmaf(-0.5f * ca, cb, 0.5f));             // hann[32 j + lane]
                    float rx, ry;
                    cunpack(z[brev5(j)], rx, ry);
                    if (q == 0) {
                        dst[32 * j] = active ? make_float2(hw * rx, -hw * ry) : make_float2(0.0f, 0.0f);
                    } else if (active) {
                        float2 acc = dst[32 * j];
                        acc.x = fmaf(hw, rx, acc.x);
                        acc.y = fmaf(-hw, ry, acc.y);
                        dst[32 * j] = acc;
                    }
                }
            }
            __syncthreads();
        }
    }
    // ---- the 13 owned hop regions -> gradient, each sample once; edges through the adjoint of the reflect padding -----
    {
        float* dL = p.out[s] + (long long)b * p.ostride_b;
        float* dR = dL + p.ostride_c;
        const long long first = p0 + 768;                            // padded position of the first owned sample
        const bool interior = first >= kBwEdge && first + kBwOwn * 256 <= (long long)p.T - kBwEdge
                              && ((reinterpret_cast<unsigned long long>(dL + first) | reinterpret_cast<unsigned long long>(dR + first)) & 15) == 0;
        if (interior) {                                              // no reflection, no strip: four samples per thread
            float4* const vL = reinterpret_cast<float4*>(dL + first);
            float4* const vR = reinterpret_cast<float4*>(dR + first);
            const float4* const src4 = reinterpret_cast<const float4*>(ola + 768);
            for (int i = tid; i < kBwOwn * 64; i += kBwThreads) {
                const float4 a = src4[2 * i], c = src4[2 * i + 1];   // (L0 R0 L1 R1), (L2 R2 L3 R3)
                vL[i] = make_float4(a.x, a.z, c.x, c.z);
                vR[i] = make_float4(a.y, a.w, c.y, c.w);
            }
            return;
        }
        for (int i = tid; i < kBwOwn * 256; i += kBwThreads) {
            const int o = 768 + i;
            const long long pp = p0 + o;                              // padded position
            if (pp < -512 || pp >= (long long)p.T + 512) continue;
            long long j = pp;
            if (j < 0) j = -j;
            if (j >= p.T) j = 2ll * (p.T - 1) - j;
            const float2 v = ola[o];
            if (j < kBwEdge || j >= (long long)p.T - kBwEdge) {
                atomicAdd(dL + j, v.x);
                atomicAdd(dR + j, v.y);
            } else {
                dL[j] = v.x;
                dR[j] = v.y;
            }
        }
    }
}

}  // namespace mixfe
