// K8F: the (n_fft, hop) = (1024, 256) resolution of the multi-resolution STFT loss (k8_mrstft.cuh has the definitions and
// stays the path of every other resolution) on K1's one-warp 1024-point transform -- "reusing K1's FFT block", as SURVEY.md
// section 8f row 3 puts it.  Prediction x and target y of a row share ONE complex transform (x + i y), untangled like L / R:
// with u = Z[k], v = Z[1024 - k]:  X = (u.x + v.x) + i (u.y - v.y),  Y = (u.y + v.y) - i (u.x - v.x)  (window 0.5 hann folded in).
//
// k8f_partials: one CTA per (tile of 8 frames, row), one warp per frame; the partner bin comes by shuffle from lane
//   (32 - lane) & 31; per tile the three sums of k8_partials, in its layout, so k8_finish is shared.
// k8f_backward: K6F's scheme (k6f_logmel_bwd.cuh): a CTA of 16 warps owns 13 hop regions of one row and computes the 16 frames
//   that touch them; the gradient stage is K6F's with dP_L = (d loss / d|X|) / (2 |X|) computed in the lane from the untangled
//   bin and dP_R = 0 (the target gets no gradient); the overlap-add is complete in shared memory and every gradient sample
//   is ADDED ONCE to grad_x (the resolutions accumulate into one tensor the caller zeroed), with atomics only on the 513
//   samples at either end of a row that the reflect padding folds onto.
#pragma once

#include "k6f_logmel_bwd.cuh"
#include "k8_mrstft.cuh"

namespace mixfe {

struct K8FParams {
    const float* x;                  // prediction, element (row, t) at x[(row / C) * xs_b + (row % C) * xs_c + t]
    const float* y;                  // target, same addressing with ys_b, ys_c
    long long xs_b, xs_c, ys_b, ys_c;
    int rows, C, T, F, n_tiles;      // n_tiles: of 8 frames (forward) or of 13 hop regions (backward)
    const float2* tw;                // [1024] 32 x 32 twiddles
    float* part;                     // [rows][n_tiles][3]   (forward)
    const double* state;             // k8_finish's output   (backward)
    const float* gscale;             // device scalar: upstream gradient / n_resolutions
    float* gx;                       // gradient of x, accumulated; addressing gs_b, gs_c
    long long gs_b, gs_c;
    int off_tw, off_red, off_ola;
};

// frame t of a row pair -> this lane's 32 windowed (x, y) values: element 32 n1 + lane in slot n1 (torch 'reflect' indexing)
__device__ __forceinline__ void k8f_load_frame(const float* gx_, const float* gy_, int t, int T, int lane, const float2* tw,
                                               cplx64 (&z)[32]) {
    const long long base = 256ll * t - 512 + lane;
    const bool inside = base - lane >= 0 && base - lane + 1024 <= (long long)T;
    float fx[32], fy[32];
    if (inside) {
        const float* const px = gx_ + base;
        const float* const py = gy_ + base;
#pragma unroll
        for (int n1 = 0; n1 < 32; ++n1) {
            fx[n1] = __ldg(px + 32 * n1);
            fy[n1] = __ldg(py + 32 * n1);
        }
    } else {
#pragma unroll
        for (int n1 = 0; n1 < 32; ++n1) {
            long long j = base + 32 * n1;
            if (j < 0) j = -j;
            if (j >= T) j = 2ll * (T - 1) - j;
            j = j < 0 ? 0 : (j >= T ? T - 1 : j);
            fx[n1] = __ldg(gx_ + j);
            fy[n1] = __ldg(gy_ + j);
        }
    }
    const float2 wl = tw[lane];
    const float cb = wl.x, sb = -wl.y;
#pragma unroll
    for (int n1 = 0; n1 < 32; ++n1) {
        const float ca = n1 < 16 ? w32_re(n1) : -w32_re(n1 - 16);
        const float sa = n1 < 16 ? w32_sin(n1) : -w32_sin(n1 - 16);
        const float wv = fmaf(0.25f * sa, sb, fmaf(-0.25f * ca, cb, 0.25f));          // 0.5 * hann[32 n1 + lane]
        z[n1] = cpack(fx[n1] * wv, fy[n1] * wv);
    }
}

// K1's four-step transform on the warp's exchange region; on return slot k2 (register brev5(k2)) holds bin lane + 32 k2
__device__ __forceinline__ void k8f_fft1024(cplx64 (&z)[32], float2* exw, const float2* tw, int lane) {
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
}

// grid (n_tiles of 8 frames, rows); 8 warps
__global__ void __launch_bounds__(256, 2) k8f_partials(const __grid_constant__ K8FParams p) {
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
        cplx64 z[32];
        k8f_load_frame(xr, yr, t, p.T, lane, tw, z);
        k8f_fft1024(z, exw, tw, lane);
        const int src = (32 - lane) & 31;
        const bool l0 = lane == 0;
        auto bin = [&](cplx64 o, cplx64 q) {
            float ux, uy, vx, vy;
            cunpack(o, ux, uy);
            cunpack(q, vx, vy);
            const float a = ux + vx, bb = uy - vy, c = ux - vx, d = uy + vy;             // X = a + i bb, Y = d - i c
            const float mx = sqrtf(fmaf(a, a, bb * bb)), my = sqrtf(fmaf(d, d, c * c));
            const float dm = my - mx;
            a2 = fmaf(dm, dm, a2);
            b2 = fmaf(my, my, b2);
            l1 += fabsf(logf(mx + 1e-5f) - logf(my + 1e-5f));
        };
        // slot k2 <= 15 holds the one-sided bin lane + 32 k2; its partner 1024 - k sits in lane (32 - lane) & 31, slot 31 - k2
        // (lane 0: its own slot 32 - k2; bins 0 and 512 are their own partners)
#pragma unroll
        for (int k2 = 0; k2 < 16; ++k2) {
            float qx, qy;
            cunpack(l0 ? z[brev5((32 - k2) & 31)] : z[brev5(31 - k2)], qx, qy);
            qx = __shfl_sync(kFullMask, qx, src);
            qy = __shfl_sync(kFullMask, qy, src);
            bin(z[brev5(k2)], cpack(qx, qy));
        }
        if (l0) bin(z[brev5(16)], z[brev5(16)]);
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

// grid (n_tiles of 13 hop regions, rows); 16 warps (K6F's geometry: kBwWarps frames, kBwOwn owned regions)
__global__ void __launch_bounds__(kBwThreads, 1) k8f_backward(const __grid_constant__ K8FParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    float2* const tw = reinterpret_cast<float2*>(smem + p.off_tw);              // [31][32], rows k1 = 1..31
    float* const ola = reinterpret_cast<float*>(smem + p.off_ola);              // [kBwSpan]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x, row = blockIdx.y;
    const int h_first = -2 + kBwOwn * tile;              // first owned hop region (padded coordinates)
    const int t_first = h_first - 1;                     // frame of warp 0
    const long long p0 = 256ll * t_first - 512;          // padded position of ola[0]
    const float* xr = p.x + (long long)(row / p.C) * p.xs_b + (long long)(row % p.C) * p.xs_c;
    const float* yr = p.y + (long long)(row / p.C) * p.ys_b + (long long)(row % p.C) * p.ys_c;
    const float gs = __ldg(p.gscale);
    const float alpha = (float)p.state[5] * gs, beta = (float)p.state[6] * gs;

    for (int i = tid; i < 31 * 32; i += kBwThreads) tw[i] = p.tw[32 + i];
    __syncthreads();
    const int t = t_first + warp;                        // this warp's frame
    const bool active = t >= 0 && t < p.F;
    float2* const exw = reinterpret_cast<float2*>(smem + warp * kWarpRegion);
    cplx64 z[32];
    if (active) {
        k8f_load_frame(xr, yr, t, p.T, lane, tw, z);
        k8f_fft1024(z, exw, tw, lane);
        // ---- gradient stage (K6F's pairing).  o = Z[k] (k <= 512), q = Z[1024 - k]: X = xa + i xb, Y = xd - i xc;
        //      g = d loss / d|X| = alpha (|X| - |Y|) + beta sign(log(|X| + 1e-5) - log(|Y| + 1e-5)) / (|X| + 1e-5);
        //      G = g X / |X| (0 at |X| = 0, as torch.abs); the conjugated inverse-transform inputs are
        //      lo = (xa, -xb) w at k and hi = (xa, xb) w at 1024 - k with w = g / (2 |X|); twice lo at bins 0 and 512.
        const int src = (32 - lane) & 31;
        const bool l0 = lane == 0;
        auto grad_pair = [&](cplx64 o, cplx64 q, cplx64& lo, cplx64& hi) {
            float xx, xy, yx, yy;
            cunpack(o, xx, xy);
            cunpack(q, yx, yy);
            const float xa = xx + yx, xb = xy - yy, xc = xx - yx, xd = xy + yy;
            const float mx = sqrtf(fmaf(xa, xa, xb * xb)), my = sqrtf(fmaf(xd, xd, xc * xc));
            const float dl = logf(mx + 1e-5f) - logf(my + 1e-5f);
            const float sg = dl > 0.0f ? 1.0f : (dl < 0.0f ? -1.0f : 0.0f);
            const float gm = fmaf(alpha, mx - my, beta * sg / (mx + 1e-5f));
            const float w = mx > 0.0f ? 0.5f * gm / mx : 0.0f;
            lo = cpack(xa * w, -xb * w);
            hi = cpack(xa * w, xb * w);
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
        grad_pair(z[brev5(16)], z[brev5(16)], carry, unused);
        carry = twice(carry);
#pragma unroll
        for (int k2 = 15; k2 >= 0; --k2) {
            const cplx64 o = z[brev5(k2)];
            const cplx64 q = shfl_c(l0 ? z[brev5((32 - k2) & 31)] : z[brev5(31 - k2)]);
            cplx64 lo, hi;
            grad_pair(o, q, lo, hi);
            const cplx64 recv = shfl_c(hi);
            zn[k2] = (k2 == 0 && l0) ? twice(lo) : lo;
            zn[31 - k2] = l0 ? carry : recv;
            carry = recv;
        }
#pragma unroll
        for (int k2 = 0; k2 < 32; ++k2) z[k2] = zn[k2];
        k8f_fft1024(z, exw, tw, lane);                   // the same routine on the conjugated spectrum = the inverse
    }
    // ---- window, overlap-add by quarters (K6F): turn q = every warp's q-th quarter into hop region warp + q; turn 0 stores
    {
        const float2 wl = tw[lane];
        const float cb = wl.x, sb = -wl.y;
        float* const dst = ola + 256 * warp + lane;
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
                    if (q == 0) dst[32 * j] = active ? hw * rx : 0.0f;
                    else if (active) dst[32 * j] = fmaf(hw, rx, dst[32 * j]);
                }
            }
            __syncthreads();
        }
    }
    // ---- the 13 owned hop regions are ADDED to the gradient, each sample once; strips at the row ends by atomics ------------
    {
        float* const gr = p.gx + (long long)(row / p.C) * p.gs_b + (long long)(row % p.C) * p.gs_c;
        for (int i = tid; i < kBwOwn * 256; i += kBwThreads) {
            const int o = 768 + i;
            const long long pp = p0 + o;                              // padded position
            if (pp < -512 || pp >= (long long)p.T + 512) continue;
            long long j = pp;
            if (j < 0) j = -j;
            if (j >= p.T) j = 2ll * (p.T - 1) - j;
            const float v = ola[o];
            if (j < kBwEdge || j >= (long long)p.T - kBwEdge) atomicAdd(gr + j, v);
            else gr[j] += v;
        }
    }
}

}  // namespace mixfe
