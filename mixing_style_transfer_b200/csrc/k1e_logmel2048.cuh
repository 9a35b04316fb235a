// K1E: log-mel of the n_fft = 2048 configurations -- the reference's trained encoder and evaluation extractor use
// (2048, 512, 80) (scripts/train_baseline.sh:42-44, src/train_style_transfer.py:523-533, inference/validate_retrieval.py:355)
// -- on the one-warp 1024-point transform of K1.  (Features at n_fft = 2048, and every other n_fft, stay on K1G.)
//
// Reference behaviour being replaced: MelSpectrogramPreprocessor.forward (src/model.py:41-67) for that configuration.
//
// A 2048-sample REAL frame of one channel is one 1024-point COMPLEX transform of z[m] = x[2m] + i x[2m+1] (the real-FFT
// packing), so ONE WARP handles one (frame, channel) on K1's register-resident transform.  With Z = FFT_1024(z),
// x^ = Z[k], y^ = Z[1024 - k], w = exp(-2 pi i k / 2048) and the window 0.5 hann folded into z:
//   Se = x^ + conj y^,  So = (x^ - conj y^) / i,  X[k] = Se + w So,  X[1024 - k] = conj(Se - w So),  P = |X|^2;
// the partner bin comes from lane (32 - lane) & 31 by shuffle and the lane that holds the lower bin writes both powers
// into its channel's component of the frame's (L, R) power column.  The two warps of a frame (L, R) meet at ONE pair
// barrier before their exchange regions become that column.  (The first cut split a frame by decimation in time over an
// even and an odd warp with L + iR packing: parked spectra, two pair barriers, 64 shared-memory loads per lane: 2.09 ms
// against 1.73 ms per 72 clips.)
// A CTA (8 warps, 2 per SM, persistent) therefore handles tiles of 4 frames; staging (TMA bulk copies one (tile, stem) ahead,
// cooperative loads with torch's 'reflect' at the clip edges), the exchange layout and the log epilogue are K1's.
// Mel projection: thread = (work item, frame pair).  The filterbank by filter (CSR) is cut into at most 128 items, the
// longest filters in 2 or 4 parts that sit in neighbouring lanes and are joined by shuffles, items sorted by
// length so that a warp's lanes walk similar supports; sums run in a fixed order (deterministic).
#pragma once

#include "k1_features.cuh"

namespace mixfe {

constexpr int kETF = 4;                  // frames per tile (2 warps each: L, R)
constexpr int kEThreads = 256;
constexpr int kEItems = 128;             // mel work items per frame pair
constexpr int kEColOff = kWarpRegion - 4096;      // the power column starts here in the even warp's region and runs on
                                                  // into the odd warp's: 512 entries + 513 entries, contiguous

struct K1EParams {
    const float* stem[kNumStems];
    long long stride_b, stride_c;
    int B, T, F, hop, n_mels, n_tiles;
    int nxs;                         // staged samples per channel = (kETF - 1) * hop + 2048
    int tma_lo, tma_hi;              // tiles staged by TMA (empty range = none)
    long long total_items;
    int* work;                       // K1's work counters (k1_features.cuh: dynamic item distribution), or NULL
    float* logmel;
    long long lm_stride_b;
    const float2* tw;                // [1024] 32 x 32 twiddles of the 1024-point transform
    const int4* items;               // [kEItems] (filter, first bin, bins, weight offset | log2(parts) << 28 | not-first << 30)
    const float* wts;                // CSR weights
    int n_wts;
    int off_ex, off_tw, off_items, off_wts, off_bar;
    // statistics for the mixing features (want_stats): K1's partials, reduced by K2
    int want_stats, n_mels_pad;
    int rows_hi;                     // tiles 0..rows_hi get their time-statistics rows by TMA (-1 = none)
    float2* part_spec;               // [B][2 n_tiles][4][n_mels_pad] (sum log, sum mel): one entry per frame PAIR
    float* part_td;                  // [B][n_tiles][2][kTdN] (k1_features.cuh)
    float2* s_ws;                    // [grid][3][256]: S values of stems 0..2 for the masking term, one slab per CTA (global,
                                     // L2-resident: 6 KB more of shared memory would cost the second CTA per SM)
};

// cos / sin of 2 pi j / 64, j = 0..31 (compile-time after unrolling)
__device__ __forceinline__ constexpr float cos64(int j) {
    constexpr float c[17] = {1.0f, 0.99518472667219688624f, 0.98078528040323044913f, 0.95694033573220886494f,
                             0.92387953251128675613f, 0.88192126434835502971f, 0.83146961230254523708f, 0.77301045336273696081f,
                             0.70710678118654752440f, 0.63439328416364549822f, 0.55557023301960222474f, 0.47139673682599764856f,
                             0.38268343236508977173f, 0.29028467725446236764f, 0.19509032201612826785f, 0.09801714032956060199f, 0.0f};
    return j <= 16 ? c[j] : -c[32 - j];
}
__device__ __forceinline__ constexpr float sin64(int j) { return j <= 16 ? cos64(16 - j) : cos64(j - 16); }

__device__ __forceinline__ void named_bar(int id, int threads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

__device__ __forceinline__ const float* k1e_row(const K1EParams& p, int b, int tile, int s) {      // first staged sample (left)
    const float* g = s == 0 ? p.stem[0] : s == 1 ? p.stem[1] : s == 2 ? p.stem[2] : p.stem[3];
    return g + (long long)b * p.stride_b + ((long long)tile * kETF * p.hop - 1024);
}

// one thread: arm the barrier and start both row copies (interior tiles, 16-byte aligned rows)
__device__ __forceinline__ void k1e_stage_tma(const K1EParams& p, int b, int tile, int s, float* xsL, unsigned bar) {
    const float* gL = k1e_row(p, b, tile, s);
    const unsigned bytes = (unsigned)p.nxs * 4u;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    mbar_expect_tx(bar, 2u * bytes);
    tma_bulk_g2s(smem_u32(xsL), gL, bytes, bar);
    tma_bulk_g2s(smem_u32(xsL + p.nxs), gL + p.stride_c, bytes, bar);
}

// all threads: torch's 'reflect' indexing (clip edges) or plain loads (unaligned rows)
__device__ __forceinline__ void k1e_stage_coop(const K1EParams& p, int b, int tile, int s, float* xsL, int tid) {
    const float* g = s == 0 ? p.stem[0] : s == 1 ? p.stem[1] : s == 2 ? p.stem[2] : p.stem[3];
    const float* gL = g + (long long)b * p.stride_b;
    const float* gR = gL + p.stride_c;
    const long long base = (long long)tile * kETF * p.hop - 1024;
    for (int i = tid; i < p.nxs; i += kEThreads) {
        long long j = base + i;
        if (j < 0) j = -j;
        if (j >= p.T) j = 2ll * (p.T - 1) - j;
        j = j < 0 ? 0 : (j >= p.T ? p.T - 1 : j);
        xsL[i] = __ldg(gL + j);
        xsL[p.nxs + i] = __ldg(gR + j);
    }
}

// STATS: also K1's partials for the mixing features (compile-time: the log-mel-only kernel keeps its register budget)
template <bool STATS>
__global__ void __launch_bounds__(kEThreads, 2) k1e_logmel2048(const __grid_constant__ K1EParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    float* const xsL = reinterpret_cast<float*>(smem);                          // [nxs] left, then [nxs] right
    unsigned char* const ex = smem + p.off_ex;                                  // 8 warp regions
    float2* const tw = reinterpret_cast<float2*>(smem + p.off_tw);              // [31][32], rows k1 = 1..31
    int4* const items = reinterpret_cast<int4*>(smem + p.off_items);            // [kEItems]
    float* const wts = reinterpret_cast<float*>(smem + p.off_wts);
    const unsigned bar = smem_u32(smem + p.off_bar);
    int* const arrivals = reinterpret_cast<int*>(smem + p.off_bar + 8);
    const unsigned bar2 = smem_u32(smem + p.off_bar + 24);                      // mbarrier of the statistics rows
    float2* const S = STATS ? p.s_ws + (size_t)blockIdx.x * (3 * kEThreads) : nullptr;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int q = warp >> 1, par = warp & 1;            // frame of the tile, even / odd half
    long long item = blockIdx.x;
    if (item >= p.total_items) return;
    int* const next_slot = reinterpret_cast<int*>(smem + p.off_bar + 16);

    for (int i = tid; i < 31 * 32; i += kEThreads) tw[i] = p.tw[32 + i];
    for (int i = tid; i < kEItems; i += kEThreads) items[i] = p.items[i];
    for (int i = tid; i < p.n_wts; i += kEThreads) wts[i] = p.wts[i];
    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_init(bar2, 1);
        *arrivals = 0;
    }
    // per-lane constants: the window phases B_c = pi (2 lane + c) / 1024 of the even (c = 0) and odd (c = 1) samples, and
    // W_2048^lane = exp(-i pi lane / 1024)
    float cb0, sb0, cb1, sb1, wlx, wly;
    sincospif((float)(2 * lane) * (1.0f / 1024.0f), &sb0, &cb0);
    sincospif((float)(2 * lane + 1) * (1.0f / 1024.0f), &sb1, &cb1);
    sincospif((float)lane * (1.0f / 1024.0f), &wly, &wlx);
    wly = -wly;
    __syncthreads();

    unsigned char* const region = ex + warp * kWarpRegion;
    float2* const exw = reinterpret_cast<float2*>(region);
    unsigned char* const regE = ex + (2 * q) * kWarpRegion;        // first of the two regions of the frame's warps (L, R)

    int b = (int)(item / p.n_tiles), tile = (int)(item - (long long)b * p.n_tiles);
    bool cur_tma = tile >= p.tma_lo && tile <= p.tma_hi;
    unsigned tma_parity = 0, stats_parity = 0;
    if (cur_tma) { if (tid == 0) k1e_stage_tma(p, b, tile, 0, xsL, bar); } else k1e_stage_coop(p, b, tile, 0, xsL, tid);

    for (;;) {
        const int u0 = tile * kETF;
        const int nvalid = p.F - u0 < kETF ? p.F - u0 : kETF;
        long long next_item = item + gridDim.x;
        int drawn = 0;
        if (p.work && tid == 0) drawn = atomicAdd(p.work, 1);         // in flight until the last stem needs it
        float msum[4] = {0.0f, 0.0f, 0.0f, 0.0f};                     // masking partial sums of this thread
        for (int s = 0; s < kNumStems; ++s) {
            if (cur_tma) {
                mbar_wait(bar, tma_parity);
                tma_parity ^= 1u;
            }
            if (p.work && s == kNumStems - 1 && tid == 0) *next_slot = drawn;
            __syncthreads();                 // B0: xs holds (tile, s); the previous mel walk has left the power columns
            if (p.work && s == kNumStems - 1) next_item = (long long)gridDim.x + *next_slot;
            bool stats_rows = false;
            if (STATS && s == 0) {
                // time-domain statistics of the tile's own region (kETF * hop samples = eight "hops" of hop / 2 for K1's
                // role code): the 8 rows by TMA into the idle exchange regions, or role loads from global memory
                if (tile <= p.rows_hi) {
                    if (tid == 0) {
                        const long long cb = (long long)b * p.stride_b + (long long)tile * kETF * p.hop;
                        const unsigned bytes = (unsigned)(kETF * p.hop) * 4u;
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        mbar_expect_tx(bar2, 8u * bytes);
#pragma unroll
                        for (int r = 0; r < kNumStems; ++r) {
                            tma_bulk_g2s(smem_u32(ex + (2 * r) * kWarpRegion), p.stem[r] + cb, bytes, bar2);
                            tma_bulk_g2s(smem_u32(ex + (2 * r + 1) * kWarpRegion), p.stem[r] + cb + p.stride_c, bytes, bar2);
                        }
                    }
                    stats_rows = true;       // consumed below, after this warp's frame is in registers
                } else {
                    k1_time_stats<2>(p, b, tile, tid);
                }
            }
            cplx64 z[32];
            const bool active = q < nvalid;
            if (active) {
                // channel `par` of frame q as z[m] = x[2m] + i x[2m+1], m = 32 n1 + lane, window 0.5 * hann2048 folded in
                const unsigned xa = smem_u32(xsL + par * p.nxs + q * p.hop + 2 * lane);
                float2 f[32];
#pragma unroll
                for (int n1 = 0; n1 < 32; ++n1) f[n1] = lds64<0>(xa + n1 * 256);
#pragma unroll
                for (int n1 = 0; n1 < 32; ++n1) {
                    const float ca = n1 < 16 ? w32_re(n1) : -w32_re(n1 - 16);
                    const float sa = n1 < 16 ? w32_sin(n1) : -w32_sin(n1 - 16);
                    const float w0 = fmaf(0.25f * sa, sb0, fmaf(-0.25f * ca, cb0, 0.25f));
                    const float w1 = fmaf(0.25f * sa, sb1, fmaf(-0.25f * ca, cb1, 0.25f));
                    z[n1] = cpack(f[n1].x * w0, f[n1].y * w1);
                }
            }
            // ---- prefetch the next (tile, stem): the last warp to have its samples in registers starts the copy -------
            {
                int nb = b, ntile = tile, ns = s + 1;
                if (ns == kNumStems) {
                    nb = (int)(next_item / p.n_tiles);
                    ntile = (int)(next_item - (long long)nb * p.n_tiles);
                    ns = 0;
                }
                const bool have_next = s + 1 < kNumStems || next_item < p.total_items;
                bool next_tma = false;
                if (have_next) {
                    next_tma = ntile >= p.tma_lo && ntile <= p.tma_hi;
                    if (next_tma) {
                        __syncwarp();
                        if (lane == 0) {
                            __threadfence_block();
                            if (atomicAdd(arrivals, 1) == kEThreads / 32 - 1) {
                                *arrivals = 0;
                                k1e_stage_tma(p, nb, ntile, ns, xsL, bar);
                            }
                        }
                    } else {
                        __syncthreads();
                        k1e_stage_coop(p, nb, ntile, ns, xsL, tid);
                    }
                }
                cur_tma = next_tma;
            }
            if (stats_rows) {                // the copy had the frame load to land
                mbar_wait(bar2, stats_parity);
                stats_parity ^= 1u;
                k1_time_stats_smem<2>(p, ex, b, tile, tid);
                __syncthreads();             // the rows are dead: the exchange regions go back to their warps
            }
            if (active) {
#pragma unroll 1
                for (int pass = 0; pass < 2; ++pass) {
                    dft32_dif_packed(z);
                    if (pass == 0) {
                        const unsigned twa = smem_u32(tw + lane) - 256;
                        const unsigned wa = smem_u32(exw) + lane * 8;
                        float2 tq[2][8];
#pragma unroll
                        for (int g = 1; g < 8; ++g) tq[0][g] = lds64<0>(twa + g * 256);
#pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            if (g < 3) {
#pragma unroll
                                for (int r = 0; r < 8; ++r) tq[(g + 1) & 1][r] = lds64<0>(twa + (8 * g + 8 + r) * 256);
                            }
#pragma unroll
                            for (int r = 0; r < 8; ++r) {
                                const int k1 = 8 * g + r;
                                if (k1 == 0) {
                                    sts_c<0>(wa, z[0]);
                                } else {
                                    float yr, yi;
                                    cunpack(z[brev5(k1)], yr, yi);
                                    const float2 t = tq[g & 1][r];
                                    sts64<0>(wa + k1 * kExRow, fmaf(yr, t.x, -yi * t.y), fmaf(yr, t.y, yi * t.x));
                                }
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
            // Slot k2 (register brev5(k2)) holds Z[k], k = lane + 32 k2; its partner Z[1024 - k] sits in lane (32 - lane) & 31,
            // slot 31 - k2 (lane 0: its own slot 32 - k2).  With x^ = Z[k], y^ = Z[1024 - k]:  Se = x^ + conj y^,
            // So = (x^ - conj y^) / i,  X[k] = Se + w So,  X[1024 - k] = conj(Se - w So),  w = W_2048^k = W_2048^lane W_64^k2.
            // The lane that holds the lower bin writes both powers into this channel's component of the frame's (L, R) column.
            named_bar(1 + q, 64);            // both warps of the frame are past their transposition: their regions become the column
            if (active) {
                float* const colf = reinterpret_cast<float*>(regE + kEColOff) + par;
                const int src = (32 - lane) & 31;
                const bool l0 = lane == 0;
#pragma unroll
                for (int k2 = 0; k2 < 16; ++k2) {
                    float xx, xy, yx, yy;
                    cunpack(z[brev5(k2)], xx, xy);
                    cunpack(l0 ? z[brev5((32 - k2) & 31)] : z[brev5(31 - k2)], yx, yy);
                    yx = __shfl_sync(kFullMask, yx, src);
                    yy = __shfl_sync(kFullMask, yy, src);
                    const float sex = xx + yx, sey = xy - yy, sox = xy + yy, soy = yx - xx;
                    const float c = cos64(k2), sn = sin64(k2);                 // W_64^k2 = c - i sn
                    const float wr = k2 == 0 ? wlx : fmaf(wlx, c, wly * sn), wi = k2 == 0 ? wly : fmaf(wly, c, -wlx * sn);
                    const float tx = fmaf(wr, sox, -wi * soy), ty = fmaf(wr, soy, wi * sox);
                    const float ar = sex + tx, ai = sey + ty, br = sex - tx, bi = sey - ty;
                    const int k = lane + 32 * k2;
                    colf[2 * k] = fmaf(ar, ar, ai * ai);
                    colf[2 * (1024 - k)] = fmaf(br, br, bi * bi);              // (lane 0, k2 = 0: bins 0 and 1024)
                }
                if (l0) {                    // bin 512 is its own partner: X[512] = conj Z[512] * 2
                    float xx, xy;
                    cunpack(z[brev5(16)], xx, xy);
                    colf[2 * 512] = 4.0f * fmaf(xx, xx, xy * xy);
                }
            }
            __syncthreads();                 // B1: the 4 power columns of the tile are complete

            // ---- mel projection + log + store: thread = (item, frame pair) ----------------------------------------------
            {
                const int fp = tid >> 7, it = tid & (kEItems - 1);
                const int4 im = items[it];
                const int m = im.x, b0 = im.y, len = im.z;
                const int woff = im.w & 0x0fffffff;
                const int lg_parts = (im.w >> 28) & 3;                 // the filter is cut into 1, 2 or 4 neighbouring items
                const bool second = (im.w >> 30) & 1;                  // not the first part: does not store
                const float2* c0 = reinterpret_cast<const float2*>(ex + (4 * fp) * kWarpRegion + kEColOff) + b0;
                const float2* c1 = reinterpret_cast<const float2*>(ex + (4 * fp + 2) * kWarpRegion + kEColOff) + b0;
                const float* w = wts + woff;
                float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f, a3 = 0.0f;
#pragma unroll 4
                for (int i = 0; i < len; ++i) {
                    const float wi = w[i];
                    const float2 u = c0[i], v = c1[i];
                    a0 = fmaf(wi, u.x, a0);
                    a1 = fmaf(wi, u.y, a1);
                    a2 = fmaf(wi, v.x, a2);
                    a3 = fmaf(wi, v.y, a3);
                }
                // the parts of a cut filter sit in 2 or 4 neighbouring, aligned lanes: pairwise sums, the same order in every lane
#pragma unroll
                for (int d = 1; d <= 2; d <<= 1) {
                    const float o0 = __shfl_xor_sync(kFullMask, a0, d), o1 = __shfl_xor_sync(kFullMask, a1, d);
                    const float o2 = __shfl_xor_sync(kFullMask, a2, d), o3 = __shfl_xor_sync(kFullMask, a3, d);
                    if (lg_parts >= d) { a0 += o0; a1 += o1; a2 += o2; a3 += o3; }
                }
                if constexpr (!STATS) {
                    if (!second && m >= 0 && m < p.n_mels && p.logmel) {
                        const int u = u0 + 2 * fp;
                        float* d0 = p.logmel + (long long)b * p.lm_stride_b + ((long long)(2 * s) * p.n_mels + m) * p.F + u;
                        float* d1 = d0 + (long long)p.n_mels * p.F;
                        if (u < p.F) { d0[0] = ln_fast(a0 + 1e-10f); d1[0] = ln_fast(a1 + 1e-10f); }
                        if (u + 1 < p.F) { d0[1] = ln_fast(a2 + 1e-10f); d1[1] = ln_fast(a3 + 1e-10f); }
                    }
                } else {
                    const int u = u0 + 2 * fp;
                    const bool owner = !second && m >= 0 && m < p.n_mels;      // one thread per (filter, frame pair)
                    const bool ok0 = u < p.F, ok1 = u + 1 < p.F;
                    const float lg0 = ln_fast(a0 + 1e-10f), lg1 = ln_fast(a1 + 1e-10f);
                    const float lg2 = ln_fast(a2 + 1e-10f), lg3 = ln_fast(a3 + 1e-10f);
                    if (owner && p.logmel) {
                        float* d0 = p.logmel + (long long)b * p.lm_stride_b + ((long long)(2 * s) * p.n_mels + m) * p.F + u;
                        float* d1 = d0 + (long long)p.n_mels * p.F;
                        if (ok0) { d0[0] = lg0; d1[0] = lg1; }
                        if (ok1) { d0[1] = lg2; d1[1] = lg3; }
                    }
                    // K1's partials: per (stem, filter) sums of log and power over the frame pair; S = (mel_L + mel_R) / 2
                    // kept per thread for stems 0..2, the masking sigmoids at the last stem
                    const long long tile_id = (long long)b * p.n_tiles + tile;
                    if (owner) {
                        const float slog = (ok0 ? lg0 + lg1 : 0.0f) + (ok1 ? lg2 + lg3 : 0.0f);
                        const float smel = (ok0 ? a0 + a1 : 0.0f) + (ok1 ? a2 + a3 : 0.0f);
                        p.part_spec[((2 * tile_id + fp) * kNumStems + s) * p.n_mels_pad + m] = make_float2(slog, smel);
                    }
                    const float sc0 = 0.5f * (a0 + a1), sc1 = 0.5f * (a2 + a3);
                    if (s < 3) {
                        S[s * kEThreads + tid] = make_float2(sc0, sc1);
                    } else {
                        const float2 A0 = S[tid], A1 = S[kEThreads + tid], A2 = S[2 * kEThreads + tid];
                        const float s0[2] = {A0.x, A0.y}, s1[2] = {A1.x, A1.y}, s2[2] = {A2.x, A2.y}, s3[2] = {sc0, sc1};
                        const bool ok[2] = {owner && ok0, owner && ok1};
#pragma unroll
                        for (int i = 0; i < 2; ++i) {
                            if (ok[i]) {
                                const float m01 = fmaxf(s0[i], s1[i]), m23 = fmaxf(s2[i], s3[i]);
                                msum[0] += fast_sigmoid_neg(s0[i] - fmaxf(s1[i], m23));
                                msum[1] += fast_sigmoid_neg(s1[i] - fmaxf(s0[i], m23));
                                msum[2] += fast_sigmoid_neg(s2[i] - fmaxf(m01, s3[i]));
                                msum[3] += fast_sigmoid_neg(s3[i] - fmaxf(m01, s2[i]));
                            }
                        }
#pragma unroll
                        for (int i = 0; i < 4; ++i) msum[i] = warp_sum(msum[i]);
                        if (lane == 0)           // the warp's own slots of the tile's records (K1's layout)
                            *reinterpret_cast<float4*>(p.part_td + tile_id * (2 * kTdN) + (warp >> 2) * kTdN + kTdMask + (warp & 3) * 4) =
                                make_float4(msum[0], msum[1], msum[2], msum[3]);
                    }
                }
            }
        }
        if (next_item >= p.total_items) break;
        item = next_item;
        b = (int)(item / p.n_tiles);
        tile = (int)(item - (long long)b * p.n_tiles);
    }
    if (p.work && tid == 0) {            // the last CTA out leaves the counters at zero for the next launch
        if (atomicAdd(p.work + 1, 1) == (int)gridDim.x - 1) {
            p.work[0] = 0;
            p.work[1] = 0;
        }
    }
}

}  // namespace mixfe
