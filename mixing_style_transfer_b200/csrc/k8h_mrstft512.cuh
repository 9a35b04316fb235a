// K8H: the (n_fft, hop) = (512, 128) resolution of the multi-resolution STFT loss (k8_mrstft.cuh has the definitions) with
// TWO frames per warp on the 512-point warp transform of the reverb's column pass (warp_fft512x2, k4_reverb.cuh: every lane
// runs two 16-point DFTs and, after one exchange, one 32-point DFT).  Prediction x and target y of a row share a complex
// transform (x + i y) as in K8F.  Layouts: in, slot 16 c + i of lane L = element 32 i + L of frame c (c = 0, 1); out, slot k2
// of lane 16 c + l = bin l + 16 k2 of frame c.  The partner bin 512 - k of bin k = l + 16 k2 sits in lane 16 c + ((16 - l) & 15),
// slot 31 - k2 (l = 0: the lane's own slot 32 - k2): K8F's pairing inside each half-warp.
//
// k8h_partials: one CTA of 4 warps per (tile of 8 frames, row): k8_partials' three sums in its layout (k8_finish is shared).
// k8h_backward: K6F's scheme with 32 frames and 29 owned hop regions per CTA of 16 warps.  The out layout is not the in layout,
//   so the conjugated gradient spectrum goes once through the warp's exchange buffer before the same routine runs as the
//   inverse; its output, sample l + 16 k2 in slot k2, is windowed and overlap-added by quarters (slots 8 q .. 8 q + 7).
//   Every gradient sample is ADDED ONCE to grad_x; atomics only on the 257 samples at either end of a row.
#pragma once

#include "k4_reverb.cuh"
#include "k8f_mrstft1024.cuh"

namespace mixfe {

constexpr int kHFrames = 32;                      // frames per CTA of the backward kernel (16 warps x 2)
constexpr int kHOwn = 29;                         // hop regions it owns
constexpr int kHSpan = (kHFrames - 1) * 128 + 512;            // 4480 positions covered by the 32 frames
constexpr int kHEdge = 257;                       // samples at either end of a row that take atomics
constexpr int kHRegion = 8448;                    // bytes of a warp's exchange buffer (1024 float2 + padding for the re-layout)

// cos / sin of 2 pi i / 16, i = 0..15 (compile-time after unrolling)
__device__ __forceinline__ constexpr float cos16(int i) { return i <= 8 ? w32_re(2 * i) : w32_re(32 - 2 * i); }
__device__ __forceinline__ constexpr float sin16(int i) { return i <= 8 ? w32_sin(2 * i) : -w32_sin(32 - 2 * i); }

// frames t and t + 1 of a row pair -> slots 16 c + i = 0.5 hann[32 i + lane] (x + i y)[128 (t + c) - 256 + 32 i + lane]
__device__ __forceinline__ void k8h_load_frames(const float* gx_, const float* gy_, int t, int T, int lane, float cbf, float sbf,
                                                float (&re)[32], float (&im)[32]) {
#pragma unroll
    for (int c = 0; c < 2; ++c) {
        const long long base = 128ll * (t + c) - 256 + lane;
        const bool inside = base - lane >= 0 && base - lane + 512 <= (long long)T;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            long long j = base + 32 * i;
            if (!inside) {
                if (j < 0) j = -j;
                if (j >= T) j = 2ll * (T - 1) - j;
                j = j < 0 ? 0 : (j >= T ? T - 1 : j);
            }
            re[16 * c + i] = __ldg(gx_ + j);
            im[16 * c + i] = __ldg(gy_ + j);
        }
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const float wv = fmaf(0.25f * sin16(i), sbf, fmaf(-0.25f * cos16(i), cbf, 0.25f));        // 0.5 * hann512[32 i + lane]
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            re[16 * c + i] *= wv;
            im[16 * c + i] *= wv;
        }
    }
}

// grid (n_tiles of 8 frames, rows); 4 warps, two frames each
__global__ void __launch_bounds__(128) k8h_partials(const __grid_constant__ K8FParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    float2* const tw = reinterpret_cast<float2*>(smem + p.off_tw);              // [32][32] twiddles of the 1024-point table
    float* const red = reinterpret_cast<float*>(smem + p.off_red);              // [4 warps][3]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x, row = blockIdx.y;
    const float* xr = p.x + (long long)(row / p.C) * p.xs_b + (long long)(row % p.C) * p.xs_c;
    const float* yr = p.y + (long long)(row / p.C) * p.ys_b + (long long)(row % p.C) * p.ys_c;
    for (int i = tid; i < 1024; i += 128) tw[i] = p.tw[i];
    __syncthreads();
    const int t = tile * kTF + 2 * warp;                 // frames t (lanes 0..15 after the transform) and t + 1 (lanes 16..31)
    float a2 = 0.0f, b2 = 0.0f, l1 = 0.0f;
    if (t < p.F) {
        float2* const exw = reinterpret_cast<float2*>(smem + warp * kHRegion);
        float cbf, sbf;
        sincospif((float)lane * (1.0f / 256.0f), &sbf, &cbf);                   // window phase 2 pi lane / 512
        float re[32], im[32];
        k8h_load_frames(xr, yr, t, p.T, lane, cbf, sbf, re, im);
        warp_fft512x2(re, im, exw, tw, lane);
        const int l = lane & 15;
        const int src = (lane & 16) | ((16 - l) & 15);
        const bool l0 = l == 0;
        const bool mine = t + (lane >> 4) < p.F;         // the second frame of the last pair may not exist
        auto bin = [&](float ux, float uy, float vx, float vy) {
            const float a = ux + vx, bb = uy - vy, c = ux - vx, d = uy + vy;             // X = a + i bb, Y = d - i c
            const float mx = sqrtf(fmaf(a, a, bb * bb)), my = sqrtf(fmaf(d, d, c * c));
            const float dm = my - mx;
            a2 = fmaf(dm, dm, a2);
            b2 = fmaf(my, my, b2);
            l1 += fabsf(logf(mx + 1e-5f) - logf(my + 1e-5f));
        };
#pragma unroll
        for (int k2 = 0; k2 < 16; ++k2) {
            const float qx = __shfl_sync(kFullMask, l0 ? re[(32 - k2) & 31] : re[31 - k2], src);
            const float qy = __shfl_sync(kFullMask, l0 ? im[(32 - k2) & 31] : im[31 - k2], src);
            if (mine) bin(re[k2], im[k2], qx, qy);
        }
        if (l0 && mine) bin(re[16], im[16], re[16], im[16]);                             // bin 256 is its own partner
    }
    a2 = warp_sum(a2); b2 = warp_sum(b2); l1 = warp_sum(l1);
    if (lane == 0) { red[warp * 3] = a2; red[warp * 3 + 1] = b2; red[warp * 3 + 2] = l1; }
    __syncthreads();
    if (tid < 3) {
        float v = 0.0f;
        for (int w = 0; w < 4; ++w) v += red[w * 3 + tid];
        p.part[((long long)row * p.n_tiles + tile) * 3 + tid] = v;
    }
}

// grid (n_tiles of 29 hop regions, rows); 16 warps, two frames each
__global__ void __launch_bounds__(512, 1) k8h_backward(const __grid_constant__ K8FParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    float2* const tw = reinterpret_cast<float2*>(smem + p.off_tw);              // [32][32]
    float* const ola = reinterpret_cast<float*>(smem + p.off_ola);              // [kHSpan]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x, row = blockIdx.y;
    const int h_first = -2 + kHOwn * tile;               // first owned hop region (padded coordinates)
    const int t_first = h_first - 1;                     // first frame of the CTA
    const long long p0 = 128ll * t_first - 256;          // padded position of ola[0]
    const float* xr = p.x + (long long)(row / p.C) * p.xs_b + (long long)(row % p.C) * p.xs_c;
    const float* yr = p.y + (long long)(row / p.C) * p.ys_b + (long long)(row % p.C) * p.ys_c;
    const float gs = __ldg(p.gscale);
    const float alpha = (float)p.state[5] * gs, beta = (float)p.state[6] * gs;

    for (int i = tid; i < 1024; i += 512) tw[i] = p.tw[i];
    __syncthreads();
    const int t = t_first + 2 * warp;                    // this warp's frames t and t + 1
    const int l = lane & 15, c = lane >> 4;
    const int tm = t + c;                                // the frame this lane holds after a transform
    const bool mine = tm >= 0 && tm < p.F;
    float2* const exw = reinterpret_cast<float2*>(smem + warp * kHRegion);
    float re[32], im[32];
    if (t + 1 >= 0 && t < p.F) {                         // at least one of the two frames exists (warp-uniform)
        float cbf, sbf;
        sincospif((float)lane * (1.0f / 256.0f), &sbf, &cbf);
        k8h_load_frames(xr, yr, t, p.T, lane, cbf, sbf, re, im);
        warp_fft512x2(re, im, exw, tw, lane);
        // ---- gradient stage: K8F's, inside each half-warp.  lo = (xa, -xb) w at bin k, hi = (xa, xb) w at 512 - k,
        //      w = g / (2 |X|); twice lo at bins 0 and 256 -----------------------------------------------------------
        const int src = (lane & 16) | ((16 - l) & 15);
        const bool l0 = l == 0;
        auto grad_pair = [&](float xx, float xy, float yx, float yy, float& lox, float& loy, float& hix, float& hiy) {
            const float xa = xx + yx, xb = xy - yy, xc = xx - yx, xd = xy + yy;
            const float mx = sqrtf(fmaf(xa, xa, xb * xb)), my = sqrtf(fmaf(xd, xd, xc * xc));
            const float dl = logf(mx + 1e-5f) - logf(my + 1e-5f);
            const float sg = dl > 0.0f ? 1.0f : (dl < 0.0f ? -1.0f : 0.0f);
            const float gm = fmaf(alpha, mx - my, beta * sg / (mx + 1e-5f));
            const float w = (mine && mx > 0.0f) ? 0.5f * gm / mx : 0.0f;
            lox = xa * w; loy = -xb * w; hix = xa * w; hiy = xb * w;
        };
        float nr[32], ni[32];
        float cx, cy, ux, uy;
        grad_pair(re[16], im[16], re[16], im[16], cx, cy, ux, uy);                       // bin 256 (meaningful for l = 0)
        cx *= 2.0f; cy *= 2.0f;
#pragma unroll
        for (int k2 = 15; k2 >= 0; --k2) {
            const float qx = __shfl_sync(kFullMask, l0 ? re[(32 - k2) & 31] : re[31 - k2], src);
            const float qy = __shfl_sync(kFullMask, l0 ? im[(32 - k2) & 31] : im[31 - k2], src);
            float lox, loy, hix, hiy;
            grad_pair(re[k2], im[k2], qx, qy, lox, loy, hix, hiy);
            const float rx = __shfl_sync(kFullMask, hix, src), ry = __shfl_sync(kFullMask, hiy, src);
            nr[k2] = (k2 == 0 && l0) ? 2.0f * lox : lox;
            ni[k2] = (k2 == 0 && l0) ? 2.0f * loy : loy;
            nr[31 - k2] = l0 ? cx : rx;
            ni[31 - k2] = l0 ? cy : ry;
            cx = rx; cy = ry;
        }
        // ---- the spectrum sits in the OUT layout (lane 16 c + l, slot k2 = bin l + 16 k2 of frame c); the routine wants the
        //      IN layout (lane L, slot 16 c + i = element 32 i + L): once through the exchange buffer (frame stride 528) ------
        {
            float2* const w0 = exw + c * 528 + l;
#pragma unroll
            for (int k2 = 0; k2 < 32; ++k2) w0[16 * k2] = make_float2(nr[k2], ni[k2]);
            __syncwarp();
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const float2 v = exw[cc * 528 + 32 * i + lane];
                    re[16 * cc + i] = v.x;
                    im[16 * cc + i] = v.y;
                }
            }
            __syncwarp();
        }
        warp_fft512x2(re, im, exw, tw, lane);            // the same routine on the conjugated spectrum = the inverse
    }
    // ---- window, overlap-add by quarters (K6F): sample l + 16 k2 of the lane's frame is in slot k2; quarter q = slots
    //      8 q .. 8 q + 7 of frame f lands in hop region f + q; turn 0 stores; only the owned regions 3 .. 31 are kept -----
    {
        float cbo, sbo;
        sincospif((float)l * (1.0f / 256.0f), &sbo, &cbo);                      // window phase 2 pi l / 512
        const int f = 2 * warp + c;
        float* const dst = ola + 128 * f + l;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int h = f + q;
            if (h >= 3 && h < 3 + kHOwn) {
#pragma unroll
                for (int j = 8 * q; j < 8 * q + 8; ++j) {
                    const float ca = j < 16 ? w32_re(j) : -w32_re(j - 16);
                    const float sa = j < 16 ? w32_sin(j) : -w32_sin(j - 16);
                    const float hw = fmaf(0.5f * sa, sbo, fmaf(-0.5f * ca, cbo, 0.5f));           // hann512[l + 16 j]
                    if (q == 0) dst[16 * j] = mine ? hw * re[j] : 0.0f;
                    else if (mine) dst[16 * j] = fmaf(hw, re[j], dst[16 * j]);
                }
            }
            __syncthreads();
        }
    }
    // ---- the 29 owned hop regions are ADDED to the gradient, each sample once; strips at the row ends by atomics ------------
    {
        float* const gr = p.gx + (long long)(row / p.C) * p.gs_b + (long long)(row % p.C) * p.gs_c;
        for (int i = tid; i < kHOwn * 128; i += 512) {
            const int o = 384 + i;
            const long long pp = p0 + o;                              // padded position
            if (pp < -256 || pp >= (long long)p.T + 256) continue;
            long long j = pp;
            if (j < 0) j = -j;
            if (j >= p.T) j = 2ll * (p.T - 1) - j;
            const float v = ola[o];
            if (j < kHEdge || j >= (long long)p.T - kHEdge) atomicAdd(gr + j, v);
            else gr[j] += v;
        }
    }
}

}  // namespace mixfe
