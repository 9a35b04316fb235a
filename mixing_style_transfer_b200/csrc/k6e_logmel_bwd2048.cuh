// K6E: backward of the log-mel front end for (n_fft, hop) = (2048, 512) -- the configuration the reference trains its
// encoder on (src/train_style_transfer.py:523-533) -- on the one-warp 1024-point transform, built like K6F
// (k6f_logmel_bwd.cuh: owned hop regions, overlap-add complete in shared memory, every gradient sample written once).
//
// A 2048-sample REAL frame of one channel is one 1024-point COMPLEX transform of z[m] = x[2m] + i x[2m+1] (the real-FFT
// packing), so a warp handles one (frame, channel) with no exchange between warps.  With Z = FFT_1024(z), x^ = Z[k],
// y^ = Z[1024 - k], w = exp(-2 pi i k / 2048) and the window 0.5 hann folded into z:
//     Se = x^ + conj y^   (spectrum of the even samples),   So = (x^ - conj y^) / i   (of the odd samples)
//     X[k] = Se + w So,   X[1024 - k] = conj(Se - w So)                                   0 < k < 512
// The gradient of sum_k P[k] dP[k], P = |X|^2, with respect to the frame is the real part of a one-sided inverse DFT of
// H[k] = 2 dP[k] X[k]; split into even / odd output samples and packed again as ge + i go it is ONE inverse 1024-point
// transform of  Y[k] = P' + i conj(w) Q',  Y[1024 - k] = conj(P') + i w conj(Q'),  where with ds = dP[k] + dP[1024 - k],
// dd = dP[k] - dP[1024 - k]:  P' = ds Se + dd w So,  Q' = dd Se + ds w So.  The kernel feeds conj(Y) to the forward routine:
//     hi = conj(Y[1024 - k]) = P' - i V,   lo = conj(Y[k]) = conj(P' + i V),   V = conj(w) Q' = dd conj(w) Se + ds So
// (bin 0 pairs with bin 1024 and takes twice lo; bin 512 is its own partner and takes lo).  The output is conj(ge + i go).
//
// Geometry in units of float2 = two samples: frame = 1024 units, hop = 256 units, reflect padding = 512 units -- K6F's
// numbers, so the tiling (16 frames per CTA, 13 owned hop regions) and the overlap-add by quarters are K6F's.  A CTA serves
// one (clip, stem, channel).  The strips at the row ends that the reflect padding folds onto are 1025 samples wide.
#pragma once

#include "k6f_logmel_bwd.cuh"

namespace mixfe {

constexpr int kBeEdge = 1025;                     // samples at either end of a row that take atomics
constexpr int kBeBins = 1025;                     // one-sided bins of the 2048-point transform

struct K6EParams {
    const float* stem[kNumStems];    // forward inputs
    long long stride_b, stride_c;
    float* out[kNumStems];           // gradients (written once; only the edge strips must be zero on entry)
    long long ostride_b, ostride_c;
    const float* g;                  // d loss / d logmel, (b, ch, m, f) at g[b * g_stride_b + (ch * n_mels + m) * F + f]
    long long g_stride_b;
    const float* lm;                 // the forward's log-mel output, same shape
    long long lm_stride_b;
    int B, T, F, n_mels, n_tiles;
    const float2* tw;                // [1024] 32 x 32 twiddles of the 1024-point transform
    const float4* binw;              // [1025] per bin: (first filter as int bits, weight of it, weight of the next, -)
    const float2* w2048;             // [512] exp(-2 pi i k / 2048)
    int off_ex, off_tw, off_binw, off_w, off_dmel, off_ola;
};

// zero the strips the reflect padding folds onto
__global__ void k6e_zero_edges(const K6EParams p) {
    const int row = blockIdx.x, s = blockIdx.y;          // row = 2 b + channel
    float* dst = p.out[s] + (long long)(row >> 1) * p.ostride_b + (long long)(row & 1) * p.ostride_c;
    const int n = p.T < kBeEdge ? p.T : kBeEdge;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        dst[i] = 0.0f;
        dst[p.T - 1 - i] = 0.0f;
    }
}

__global__ void __launch_bounds__(kBwThreads, 1) k6e_logmel_bwd(const __grid_constant__ K6EParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char* const ex = smem + p.off_ex;                                  // 16 warp regions (exchange)
    float2* const tw = reinterpret_cast<float2*>(smem + p.off_tw);              // [31][32], rows k1 = 1..31
    float4* const binw = reinterpret_cast<float4*>(smem + p.off_binw);          // [1025]
    float2* const w2048 = reinterpret_cast<float2*>(smem + p.off_w);            // [512]
    float* const dmel = reinterpret_cast<float*>(smem + p.off_dmel);            // [16][n_mels + 1]
    float2* const ola = reinterpret_cast<float2*>(smem + p.off_ola);            // [kBwSpan] (even, odd sample)

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x, b = blockIdx.y, s = blockIdx.z >> 1, ch = blockIdx.z & 1;
    const int h_first = -2 + kBwOwn * tile;              // first owned hop region (padded coordinates, units of 2 samples)
    const int t_first = h_first - 1;                     // frame of warp 0
    const long long p0 = 256ll * t_first - 512;          // padded position of ola[0], in units of 2 samples
    const int dm_stride = p.n_mels + 1;

    for (int i = tid; i < 31 * 32; i += kBwThreads) tw[i] = p.tw[32 + i];
    for (int i = tid; i < kBeBins; i += kBwThreads) binw[i] = p.binw[i];
    for (int i = tid; i < 512; i += kBwThreads) w2048[i] = p.w2048[i];
    // ---- this warp's frame -> registers with torch's 'reflect' indexing, issued before the dmel stage (as K6F) ----------
    const int t = t_first + warp;                        // this warp's frame
    const bool active = t >= 0 && t < p.F;
    float fe[32], fo[32];                                // even / odd samples: element 32 n1 + lane of z
    if (active) {
        const float* gx = p.stem[s] + (long long)b * p.stride_b + (long long)ch * p.stride_c;
        const long long base = 2 * (256ll * t - 512 + lane);                 // sample index of this lane's first even sample
        const long long first = 2 * (256ll * t - 512);
        const bool inside = first >= 0 && first + 2048 <= (long long)p.T && (reinterpret_cast<unsigned long long>(gx) & 7) == 0;
        if (inside) {
            const float2* const px = reinterpret_cast<const float2*>(gx + base);
#pragma unroll
            for (int n1 = 0; n1 < 32; ++n1) {
                const float2 v = __ldg(px + 32 * n1);
                fe[n1] = v.x;
                fo[n1] = v.y;
            }
        } else {
#pragma unroll
            for (int n1 = 0; n1 < 32; ++n1) {
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    long long j = base + 64 * n1 + c;
                    if (j < 0) j = -j;
                    if (j >= p.T) j = 2ll * (p.T - 1) - j;
                    j = j < 0 ? 0 : (j >= p.T ? p.T - 1 : j);
                    const float v = __ldg(gx + j);
                    if (c == 0) fe[n1] = v; else fo[n1] = v;
                }
            }
        }
    }
    // ---- dmel of the 16 frames: thread = (frame, filter) with the frame fastest (64-byte runs along time) ----------
    {
        const float* const gb = p.g + (long long)b * p.g_stride_b + (long long)(2 * s + ch) * p.n_mels * p.F;
        const float* const lb = p.lm + (long long)b * p.lm_stride_b + (long long)(2 * s + ch) * p.n_mels * p.F;
        for (int idx = tid; idx < kBwWarps * dm_stride; idx += kBwThreads) {
            const int f = idx & (kBwWarps - 1), m = idx >> 4;
            const int tf = t_first + f;
            float v = 0.0f;
            if (m < p.n_mels && tf >= 0 && tf < p.F) {
                const int o = m * p.F + tf;                              // one (clip, channel) plane: fits an int
                v = __ldg(gb + o) * __expf(-__ldg(lb + o));
            }
            dmel[f * dm_stride + m] = v;
        }
    }
    __syncthreads();

    float2* const exw = reinterpret_cast<float2*>(ex + warp * kWarpRegion);
    // window phases of this lane: hann2048[64 n1 + 2 lane + c] = 0.5 - 0.5 cos(2 pi n1 / 32 + B_c), B_c = pi (2 lane + c) / 1024
    float cb0, sb0, cb1, sb1;
    {
        const float2 wl = tw[lane];                      // exp(-2 pi i lane / 1024): B_0
        cb0 = wl.x; sb0 = -wl.y;
        sincospif((float)(2 * lane + 1) * (1.0f / 1024.0f), &sb1, &cb1);
    }
    cplx64 z[32];
    if (active) {
#pragma unroll
        for (int n1 = 0; n1 < 32; ++n1) {
            const float ca = n1 < 16 ? w32_re(n1) : -w32_re(n1 - 16);
            const float sa = n1 < 16 ? w32_sin(n1) : -w32_sin(n1 - 16);
            const float w0 = fmaf(0.25f * sa, sb0, fmaf(-0.25f * ca, cb0, 0.25f));        // 0.5 * hann[even sample]
            const float w1 = fmaf(0.25f * sa, sb1, fmaf(-0.25f * ca, cb1, 0.25f));        // 0.5 * hann[odd sample]
            z[n1] = cpack(fe[n1] * w0, fo[n1] * w1);
        }
    }
    // two transforms: forward, then, after the gradient stage, the same routine on the conjugated spectrum (as K6F)
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
                // ---- gradient stage (header).  Slot k2 (register brev5(k2)) holds Z[kap], kap = lane + 32 k2; its partner
                //      Z[1024 - kap] sits in lane (32 - lane) & 31, slot 31 - k2 (lane 0: slot (32 - k2) & 31). -------------
                const int src = (32 - lane) & 31;
                const bool l0 = lane == 0;
                const float* dm = dmel + warp * dm_stride;
                auto dP = [&](int bin) -> float {
                    const float4 bw = binw[bin];
                    const int m0 = __float_as_int(bw.x);
                    return fmaf(bw.y, dm[m0], bw.z * dm[m0 + 1]);
                };
                // One PAIR of bins: o = Z[kk] (kk <= 512), q = Z[1024 - kk] -> conjugated inverse-transform inputs at kk (lo)
                // and at 1024 - kk (hi)
                auto grad_pair = [&](int kk, cplx64 o, cplx64 q, cplx64& lo, cplx64& hi) {
                    float xx, xy, yx, yy;
                    cunpack(o, xx, xy);
                    cunpack(q, yx, yy);
                    const float sex = xx + yx, sey = xy - yy;            // Se
                    const float sox = xy + yy, soy = yx - xx;            // So = (x^ - conj y^) / i
                    const float2 w = w2048[kk];
                    const float tx = fmaf(w.x, sox, -w.y * soy), ty = fmaf(w.x, soy, w.y * sox);      // w So
                    const float ux = fmaf(w.x, sex, w.y * sey), uy = fmaf(w.x, sey, -w.y * sex);      // conj(w) Se
                    const float dk = dP(kk), dn = dP(1024 - kk);
                    const float ds = dk + dn, dd = dk - dn;
                    const float px = fmaf(ds, sex, dd * tx), py = fmaf(ds, sey, dd * ty);             // P'
                    const float vx = fmaf(dd, ux, ds * sox), vy = fmaf(dd, uy, ds * soy);             // V
                    hi = cpack(px + vy, py - vx);
                    lo = cpack(px - vy, -(py + vx));
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
                // Pairing as in K6F: slot k2 <= 15 of every lane is the lower member; the upper member sits in the partner
                // lane's slot 31 - k2 and receives `hi` by shuffle.  Lane 0 pairs with itself, slot k2 with slot 32 - k2
                // (one-slot carry); its slot 16 (bin 512) is its own partner, its slot 0 is the pair (0, 1024): twice lo.
                cplx64 zn[32];
                cplx64 carry, unused;
                grad_pair(512, z[brev5(16)], z[brev5(16)], carry, unused);
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
#pragma unroll
                for (int k2 = 0; k2 < 32; ++k2) z[k2] = zn[k2];
            }
        }
    }
    // ---- window, overlap-add by quarters (K6F): in turn q every warp adds the q-th quarter of its frame into hop region
    //      warp + q; turn 0 stores; regions outside the 13 owned ones are skipped.  Element = (even, odd) sample. ---------
    {
        float2* const dst = ola + 256 * warp + lane;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int h = warp + q;
            if (h >= 3 && h < 3 + kBwOwn) {
#pragma unroll
                for (int j = 8 * q; j < 8 * q + 8; ++j) {
                    const float ca = j < 16 ? w32_re(j) : -w32_re(j - 16);
                    const float sa = j < 16 ? w32_sin(j) : -w32_sin(j - 16);
                    const float h0 = fmaf(0.5f * sa, sb0, fmaf(-0.5f * ca, cb0, 0.5f));           // hann[even sample]
                    const float h1 = fmaf(0.5f * sa, sb1, fmaf(-0.5f * ca, cb1, 0.5f));           // hann[odd sample]
                    float rx, ry;
                    cunpack(z[brev5(j)], rx, ry);
                    if (q == 0) {
                        dst[32 * j] = active ? make_float2(h0 * rx, -h1 * ry) : make_float2(0.0f, 0.0f);
                    } else if (active) {
                        float2 acc = dst[32 * j];
                        acc.x = fmaf(h0, rx, acc.x);
                        acc.y = fmaf(-h1, ry, acc.y);
                        dst[32 * j] = acc;
                    }
                }
            }
            __syncthreads();
        }
    }
    // ---- the 13 owned hop regions -> gradient, each sample once; edges through the adjoint of the reflect padding -----
    {
        float* const dx = p.out[s] + (long long)b * p.ostride_b + (long long)ch * p.ostride_c;
        const long long first = 2 * (p0 + 768);                      // sample position of the first owned sample
        const bool interior = first >= kBeEdge && first + kBwOwn * 512 <= (long long)p.T - kBeEdge
                              && (reinterpret_cast<unsigned long long>(dx + first) & 15) == 0;
        if (interior) {                                              // no reflection, no strip: four samples per thread
            float4* const v4 = reinterpret_cast<float4*>(dx + first);
            const float4* const src4 = reinterpret_cast<const float4*>(ola + 768);
            for (int i = tid; i < kBwOwn * 128; i += kBwThreads) v4[i] = src4[i];
            return;
        }
        const float* const olaf = reinterpret_cast<const float*>(ola + 768);
        for (int i = tid; i < kBwOwn * 512; i += kBwThreads) {
            const long long pp = first + i;                           // padded sample position
            if (pp < -1024 || pp >= (long long)p.T + 1024) continue;
            long long j = pp;
            if (j < 0) j = -j;
            if (j >= p.T) j = 2ll * (p.T - 1) - j;
            const float v = olaf[i];
            if (j < kBeEdge || j >= (long long)p.T - kBeEdge) atomicAdd(dx + j, v);
            else dx[j] = v;
        }
    }
}

}  // namespace mixfe
