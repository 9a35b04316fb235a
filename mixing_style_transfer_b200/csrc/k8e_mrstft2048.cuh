// K8E: the (n_fft, hop) = (2048, 512) resolution of the multi-resolution STFT loss on the one-warp 1024-point transform
// (k8_mrstft.cuh has the definitions; k8f_mrstft1024.cuh is the (1024, 256) sibling).  A 2048-sample real frame is one
// 1024-point complex transform of z[m] = s[2m] + i s[2m+1] (the real-FFT packing of k6e_logmel_bwd2048.cuh), so a warp
// transforms the target's frame, keeps its 1025 magnitudes (33 per lane: the lane that holds the lower bin k of a pair
// also forms bin 1024 - k), then transforms the prediction's frame and compares bin by bin in the same lanes.
//   Se = u + conj v,  So = (u - conj v) / i,  S[k] = Se + w So,  S[1024 - k] = conj(Se - w So),  w = exp(-2 pi i k / 2048)
// k8e_partials: one CTA per (tile of 8 frames, row): k8_partials' three sums in its layout (k8_finish is shared).
// k8e_backward: K6E's scheme in units of two samples (16 frames, 13 owned hop regions per CTA); the gradient stage is K6E's
//   with dP[bin] = (d loss / d|X|) / (2 |X|) formed in the lane; every gradient sample is ADDED ONCE to grad_x, atomics only
//   on the 1025 samples at either end of a row.
#pragma once

#include "k1e_logmel2048.cuh"
#include "k6e_logmel_bwd2048.cuh"
#include "k8f_mrstft1024.cuh"

namespace mixfe {

// frame t of ONE signal as z[m] = s[2m] + i s[2m+1], m = 32 n1 + lane in slot n1, window 0.5 hann2048 (reflect indexing)
__device__ __forceinline__ void k8e_load_frame(const float* g, int t, int T, int lane, float cb0, float sb0, float cb1, float sb1,
                                               cplx64 (&z)[32]) {
    const long long first = 2 * (256ll * t - 512);
    const long long base = first + 2 * lane;
    const bool inside = first >= 0 && first + 2048 <= (long long)T && (reinterpret_cast<unsigned long long>(g) & 7) == 0;
    float fe[32], fo[32];
    if (inside) {
        const float2* const px = reinterpret_cast<const float2*>(g + base);
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
                if (j >= T) j = 2ll * (T - 1) - j;
                j = j < 0 ? 0 : (j >= T ? T - 1 : j);
                const float v = __ldg(g + j);
                if (c == 0) fe[n1] = v; else fo[n1] = v;
            }
        }
    }
#pragma unroll
    for (int n1 = 0; n1 < 32; ++n1) {
        const float ca = n1 < 16 ? w32_re(n1) : -w32_re(n1 - 16);
        const float sa = n1 < 16 ? w32_sin(n1) : -w32_sin(n1 - 16);
        const float w0 = fmaf(0.25f * sa, sb0, fmaf(-0.25f * ca, cb0, 0.25f));
        const float w1 = fmaf(0.25f * sa, sb1, fmaf(-0.25f * ca, cb1, 0.25f));
        z[n1] = cpack(fe[n1] * w0, fo[n1] * w1);
    }
}

// One pair of bins of a real-packed spectrum: o = Z[k] (k <= 512), q = Z[1024 - k], w = exp(-2 pi i k / 2048)
//   -> Se, So, T = w So;  S[k] = Se + T,  S[1024 - k] = conj(Se - T)
struct K8EPair {
    float sex, sey, sox, soy, tx, ty;
};
__device__ __forceinline__ K8EPair k8e_pair(cplx64 o, cplx64 q, float wr, float wi) {
    float xx, xy, yx, yy;
    cunpack(o, xx, xy);
    cunpack(q, yx, yy);
    K8EPair r;
    r.sex = xx + yx; r.sey = xy - yy; r.sox = xy + yy; r.soy = yx - xx;
    r.tx = fmaf(wr, r.sox, -wi * r.soy);
    r.ty = fmaf(wr, r.soy, wi * r.sox);
    return r;
}

// per-lane constants: window phases of the even / odd samples and W_2048^lane
struct K8ELane {
    float cb0, sb0, cb1, sb1, wlx, wly;
};
__device__ __forceinline__ K8ELane k8e_lane(int lane) {
    K8ELane c;
    sincospif((float)(2 * lane) * (1.0f / 1024.0f), &c.sb0, &c.cb0);
    sincospif((float)(2 * lane + 1) * (1.0f / 1024.0f), &c.sb1, &c.cb1);
    sincospif((float)lane * (1.0f / 1024.0f), &c.wly, &c.wlx);
    c.wly = -c.wly;                                      // exp(-i pi lane / 1024)
    return c;
}

// magnitudes of the 33 one-sided bins this lane forms, handed to fn(i, magnitude) with compile-time i:
// i = 2 k2: |S[lane + 32 k2]|, i = 2 k2 + 1: |S[1024 - lane - 32 k2]|, i = 32: |S[512]| (meaningful in lane 0 only)
template <class Fn>
__device__ __forceinline__ void k8e_magnitudes(const cplx64 (&z)[32], const K8ELane& c, int lane, Fn&& fn) {
    const int src = (32 - lane) & 31;
    const bool l0 = lane == 0;
#pragma unroll
    for (int k2 = 0; k2 < 16; ++k2) {
        float qx, qy;
        cunpack(l0 ? z[brev5((32 - k2) & 31)] : z[brev5(31 - k2)], qx, qy);
        qx = __shfl_sync(kFullMask, qx, src);
        qy = __shfl_sync(kFullMask, qy, src);
        const float cs = cos64(k2), sn = sin64(k2);
        const float wr = k2 == 0 ? c.wlx : fmaf(c.wlx, cs, c.wly * sn), wi = k2 == 0 ? c.wly : fmaf(c.wly, cs, -c.wlx * sn);
        const K8EPair r = k8e_pair(z[brev5(k2)], cpack(qx, qy), wr, wi);
        const float ar = r.sex + r.tx, ai = r.sey + r.ty, br = r.sex - r.tx, bi = r.sey - r.ty;
        fn(2 * k2, sqrtf(fmaf(ar, ar, ai * ai)));
        fn(2 * k2 + 1, sqrtf(fmaf(br, br, bi * bi)));
    }
    float xx, xy;
    cunpack(z[brev5(16)], xx, xy);
    fn(32, 2.0f * sqrtf(fmaf(xx, xx, xy * xy)));         // S[512] = 2 conj Z[512]
}

// grid (n_tiles of 8 frames, rows); 8 warps
__global__ void __launch_bounds__(256, 2) k8e_partials(const __grid_constant__ K8FParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    float2* const tw = reinterpret_cast<float2*>(smem + p.off_tw);              // [31][32], rows k1 = 1..31
    float* const red = reinterpret_cast<float*>(smem + p.off_red);              // [8 warps][3]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x, row = blockIdx.y;
    const float* xr = p.x + (long long)(row / p.C) * p.xs_b + (long long)(row % p.C) * p.xs_c;
    const float* yr = p.y + (long long)(row / p.C) * p.ys_b + (long long)(row % p.C) * p.ys_c;
    for (int i = tid; i < 31 * 32; i += 256) tw[i] = p.tw[32 + i];
    __syncthreads();
    const int t = tile * kTF + warp;
    float a2 = 0.0f, b2 = 0.0f, l1 = 0.0f;
    if (t < p.F) {
        float2* const exw = reinterpret_cast<float2*>(smem + warp * kWarpRegion);
        const K8ELane c = k8e_lane(lane);
        cplx64 z[32];
        // the target's 33 magnitudes per lane wait in shared memory ([i][lane]: conflict-free) while the prediction's frame is
        // loaded and transformed: as registers they spilled (448 bytes of stack)
        float* const my = reinterpret_cast<float*>(smem + p.off_red + 128) + warp * (33 * 32) + lane;
        k8e_load_frame(yr, t, p.T, lane, c.cb0, c.sb0, c.cb1, c.sb1, z);
        k8f_fft1024(z, exw, tw, lane);
        k8e_magnitudes(z, c, lane, [&](int i, float m) { my[i * 32] = m; });
        k8e_load_frame(xr, t, p.T, lane, c.cb0, c.sb0, c.cb1, c.sb1, z);
        k8f_fft1024(z, exw, tw, lane);
        k8e_magnitudes(z, c, lane, [&](int i, float mx) {
            if (i < 32 || lane == 0) {
                const float myv = my[i * 32];
                const float dm = myv - mx;
                a2 = fmaf(dm, dm, a2);
                b2 = fmaf(myv, myv, b2);
                l1 += fabsf(logf(mx + 1e-5f) - logf(myv + 1e-5f));
            }
        });
    }
    a2 = warp_sum(a2); b2 = warp_sum(b2); l1 = warp_sum(l1);
    if (lane == 0) { red[warp * 3] = a2; red[warp * 3 + 1] = b2; red[warp * 3 + 2] = l1; }
    __syncthreads();
    if (tid < 3) {
        float v = 0.0f;
        for (int w = 0; w < 8; ++w) v += red[w * 3 + tid];
        p.part[((long long)row * p.n_tiles + tile) * 3 + tid] = v;
    }
}

// grid (n_tiles of 13 hop regions, rows); 16 warps (K6E's geometry in units of two samples)
__global__ void __launch_bounds__(kBwThreads, 1) k8e_backward(const __grid_constant__ K8FParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    float2* const tw = reinterpret_cast<float2*>(smem + p.off_tw);              // [31][32], rows k1 = 1..31
    float2* const ola = reinterpret_cast<float2*>(smem + p.off_ola);            // [kBwSpan] (even, odd sample)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x, row = blockIdx.y;
    const int h_first = -2 + kBwOwn * tile;              // first owned hop region (padded coordinates, units of 2 samples)
    const int t_first = h_first - 1;                     // frame of warp 0
    const long long p0 = 256ll * t_first - 512;          // padded position of ola[0], in units of 2 samples
    const float* xr = p.x + (long long)(row / p.C) * p.xs_b + (long long)(row % p.C) * p.xs_c;
    const float* yr = p.y + (long long)(row / p.C) * p.ys_b + (long long)(row % p.C) * p.ys_c;
    const float gs = __ldg(p.gscale);
    const float alpha = (float)p.state[5] * gs, beta = (float)p.state[6] * gs;

    for (int i = tid; i < 31 * 32; i += kBwThreads) tw[i] = p.tw[32 + i];
    __syncthreads();
    const int t = t_first + warp;
    const bool active = t >= 0 && t < p.F;
    float2* const exw = reinterpret_cast<float2*>(smem + warp * kWarpRegion);
    const K8ELane c = k8e_lane(lane);
    cplx64 z[32];
    if (active) {
        float my[33];
        k8e_load_frame(yr, t, p.T, lane, c.cb0, c.sb0, c.cb1, c.sb1, z);
        k8f_fft1024(z, exw, tw, lane);
        k8e_magnitudes(z, c, lane, [&](int i, float m) { my[i] = m; });
        k8e_load_frame(xr, t, p.T, lane, c.cb0, c.sb0, c.cb1, c.sb1, z);
        k8f_fft1024(z, exw, tw, lane);
        // ---- gradient stage: K6E's (k6e_logmel_bwd2048.cuh) with dP[bin] = g / (2 |X|), g = d loss / d|X| -------------
        const int src = (32 - lane) & 31;
        const bool l0 = lane == 0;
        auto half_w = [&](float mx, float myv) -> float {
            const float dl = logf(mx + 1e-5f) - logf(myv + 1e-5f);
            const float sg = dl > 0.0f ? 1.0f : (dl < 0.0f ? -1.0f : 0.0f);
            const float gm = fmaf(alpha, mx - myv, beta * sg / (mx + 1e-5f));
            return mx > 0.0f ? 0.5f * gm / mx : 0.0f;
        };
        auto grad_pair = [&](int k2, cplx64 o, cplx64 q, float my_k, float my_n, cplx64& lo, cplx64& hi) {
            const float cs = cos64(k2 & 31), sn = sin64(k2 & 31);
            // k2 = 16 stands for bin 512 of lane 0: w = -i
            const float wr = k2 == 0 ? c.wlx : (k2 == 16 ? 0.0f : fmaf(c.wlx, cs, c.wly * sn));
            const float wi = k2 == 0 ? c.wly : (k2 == 16 ? -1.0f : fmaf(c.wly, cs, -c.wlx * sn));
            const K8EPair r = k8e_pair(o, q, wr, wi);
            const float ux = fmaf(wr, r.sex, wi * r.sey), uy = fmaf(wr, r.sey, -wi * r.sex);      // conj(w) Se
            const float ar = r.sex + r.tx, ai = r.sey + r.ty, br = r.sex - r.tx, bi = r.sey - r.ty;
            const float dk = half_w(sqrtf(fmaf(ar, ar, ai * ai)), my_k), dn = half_w(sqrtf(fmaf(br, br, bi * bi)), my_n);
            const float ds = dk + dn, dd = dk - dn;
            const float px = fmaf(ds, r.sex, dd * r.tx), py = fmaf(ds, r.sey, dd * r.ty);         // P'
            const float vx = fmaf(dd, ux, ds * r.sox), vy = fmaf(dd, uy, ds * r.soy);             // V
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
        cplx64 zn[32];
        cplx64 carry, unused;
        grad_pair(16, z[brev5(16)], z[brev5(16)], my[32], my[32], carry, unused);       // bin 512 (meaningful in lane 0 only)
#pragma unroll
        for (int k2 = 15; k2 >= 0; --k2) {
            const cplx64 o = z[brev5(k2)];
            const cplx64 q = shfl_c(l0 ? z[brev5((32 - k2) & 31)] : z[brev5(31 - k2)]);
            cplx64 lo, hi;
            grad_pair(k2, o, q, my[2 * k2], my[2 * k2 + 1], lo, hi);
            const cplx64 recv = shfl_c(hi);
            zn[k2] = (k2 == 0 && l0) ? twice(lo) : lo;
            zn[31 - k2] = l0 ? carry : recv;
            carry = recv;
        }
#pragma unroll
        for (int k2 = 0; k2 < 32; ++k2) z[k2] = zn[k2];
        k8f_fft1024(z, exw, tw, lane);                   // the same routine on the conjugated spectrum = the inverse
    }
    // ---- window, overlap-add by quarters (K6E); element = (even, odd) sample ---------------------------------------------
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
                    const float h0 = fmaf(0.5f * sa, c.sb0, fmaf(-0.5f * ca, c.cb0, 0.5f));       // hann[even sample]
                    const float h1 = fmaf(0.5f * sa, c.sb1, fmaf(-0.5f * ca, c.cb1, 0.5f));       // hann[odd sample]
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
    // ---- the 13 owned hop regions are ADDED to the gradient, each sample once; strips at the row ends by atomics ------------
    {
        float* const gr = p.gx + (long long)(row / p.C) * p.gs_b + (long long)(row % p.C) * p.gs_c;
        const long long first = 2 * (p0 + 768);                      // sample position of the first owned sample
        const float* const olaf = reinterpret_cast<const float*>(ola + 768);
        for (int i = tid; i < kBwOwn * 512; i += kBwThreads) {
            const long long pp = first + i;                           // padded sample position
            if (pp < -1024 || pp >= (long long)p.T + 1024) continue;
            long long j = pp;
            if (j < 0) j = -j;
            if (j >= p.T) j = 2ll * (p.T - 1) - j;
            const float v = olaf[i];
            if (j < kBeEdge || j >= (long long)p.T - kBeEdge) atomicAdd(gr + j, v);
            else gr[j] += v;
        }
    }
}

}  // namespace mixfe
