// K1: fused STFT -> power -> sparse mel -> log, plus every statistic the 64-D mixing-feature
// vector needs, in ONE pass over the 8 channels of a clip.  K2: per-clip finalize.
//
// Reference behaviour being replaced (paths relative to the reference checkout):
//   src/model.py:41-67            MelSpectrogramPreprocessor.forward  (log-mel, 4 STFTs / clip)
//   src/mixing_utils.py:107-139   extract_dynamics   (rms, peak, crest, loudness)
//   src/mixing_utils.py:141-236   extract_spectral   (4 more STFTs / clip)
//   src/mixing_utils.py:238-268   extract_stereo     (ILD, L/R correlation, mid/side ratio)
//   src/mixing_utils.py:270-309   extract_masking    (4 more STFTs / clip)
//   src/mixing_utils.py:311-318   compute_loudness
//   src/mixing_utils.py:320-357   _flatten_features  (sorted-key layout, clamp, NaN scrub)
// The reference runs 12 STFTs per clip (each a (2,T) real FFT + dense 513x128 matmul); here each
// stem's L and R are packed into ONE complex 1024-point FFT per frame, shared by all consumers.
//
// Work decomposition.  Persistent CTAs (2 per SM) of 8 warps stride over (clip, frame-tile) items;
// a tile is TF = 8 consecutive frames and the CTA walks its 4 stems.  Per (tile, stem):
//   stage   samples [t0*hop - 512, t0*hop + 7*hop + 512) of L and R as two planar rows in shared memory.
//           Interior tiles: two TMA bulk copies (cp.async.bulk + mbarrier) issued by ONE thread, one
//           (tile, stem) AHEAD -- by the last warp that has pulled its frame into registers, without a
//           CTA barrier -- so the copy overlaps the FFT phase of the current stem.  Clip-edge tiles
//           (torch 'reflect' padding) and unaligned inputs: cooperative loads.
//   FFT     one warp per frame: Hann window on the fly, 32 x 32 four-step FFT on packed (re, im) register
//           pairs (FADD2 butterflies), the transposition through a swizzled warp-private shared buffer
//           with the twiddle loads software-pipelined, L/R untangled with warp shuffles; the 513 (L, R)
//           powers go to a warp-private column in the same buffer
//   mel     WARP-LOCAL: lane l owns filters l, 32+l, 64+l, 96+l of its warp's frame and walks their
//           contiguous support (weights from a [step][lane] table, powers at lo(filter)+step);
//           no CTA barrier between FFT and mel.  Result: 128 x (L, R) mel powers per warp.
//   epilog  thread (mel, channel) gathers the 8 frames of the 8 warps: log(mel + 1e-10), per-(stem,
//           mel) sums of log and power for the band / tilt / flatness features, S = (mel_L+mel_R)/2
//           kept for the masking term; the log tile is then stored with coalesced 32-byte runs.
// Time-domain statistics come from the tile's own hop region, read once for all 4 stems (128-bit
// loads) so the mixture never exists in memory.  Per-tile partials go to a workspace (no float
// atomics: results are run-to-run deterministic) and K2 reduces them in double precision.
#pragma once

#include "common.cuh"
#include "fft32.cuh"

namespace mixfe {

constexpr int kTF = 8;             // frames per CTA tile
constexpr int kK1Threads = 256;    // 8 warps, one frame each
constexpr int kNfft = 1024;
constexpr int kNfreq = 513;
constexpr int kTdN = 56;           // floats per record of time-domain partials (two records per tile)
constexpr int kTdMask = 40;        // first of the 16 masking slots of a record (4 warps x 4 stems)
constexpr int kMaxMels = 128;
// warp-private shared region: the 32 x 32 exchange of the FFT, element (row, col) at byte
//   264 row + 8 col                     (row stride 264 bytes = 66 words: the transposed read, lane = row, touches 16
//                                        different bank pairs per half-warp; every offset is an immediate, no XOR swizzle)
// i.e. 8440 bytes; afterwards
//   [skew, skew + 8 * 544)        (L, R) power column of the warp's frame, bins 0..512 + zero-weight overshoot;
//                                 skew = 64 bytes for frames 2, 3, 6, 7 (the frame pairs that are read in the
//                                 same half-warp phase of the mel walk then sit 16 banks apart)
constexpr int kWarpRegion = 8448;                   // a multiple of 128 bytes: all regions start on bank 0
constexpr int kExRow = 264;                         // bytes per exchange row
constexpr int kPvEntries = 544;

struct K1Params {
    const float* stem[kNumStems];
    long long stride_b, stride_c;
    int B, T, F, hop, n_mels, n_mels_pad, n_tiles;
    int nxs;                         // staged samples per channel = (kTF-1)*hop + 1024
    // Staging decisions the host makes once per launch (pointers, strides and hop decide the alignment of every
    // (clip, tile, stem) at once; the tile index alone decides whether it lies inside the clip):
    int tma_lo, tma_hi;              // tiles tma_lo..tma_hi are staged by TMA bulk copies (empty range = none)
    int rows_hi;                     // tiles 0..rows_hi get their time-statistics rows by TMA (-1 = none)
    long long total_items;           // B * n_tiles
    int* work;                       // [2] zeroed counters (items handed out beyond the first per CTA, CTAs finished), or NULL:
                                     // NULL = CTA i takes items i, i + grid, ...; else every item after the first comes from
                                     // an atomic counter, so a CTA that drew slow tiles (clip edges) does not end the launch late
    float* logmel;
    long long lm_stride_b;
    float* melpow;                   // debug output (mel power before the log) or nullptr
    long long mp_stride_b;
    float2* part_spec;               // [B][n_tiles][4][n_mels_pad] (sum log, sum mel)
    float* part_td;                  // [B][n_tiles][2][kTdN]
    const float2* tw;                // [1024] tw[k1*32+n2] = exp(-2 pi i n2 k1 / 1024)
    const float* wsteps;             // per group of 8 filters: [steps][8] weights
    const unsigned short* flo;       // [kMaxMels] byte offset (8 * first walked bin) of each filter in the power column
    int n_wt;                        // floats in wsteps
    int wmeta[8][2][4];              // per warp, per group slot: steps (even), weight offset, group index, 0
    int want_stats;
    int off_ex, off_tw, off_s, off_wsteps, off_flo, off_red;
};

__device__ __forceinline__ float ln_fast(float x) {               // x >= 1e-10: never denormal
    float y;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y * 0.69314718055994531f;
}

__device__ __forceinline__ float fast_sigmoid_neg(float x) {     // 1 / (1 + exp(x))
    return __fdividef(1.0f, 1.0f + __expf(x));
}

// Shared-memory accesses whose ORDER matters for latency hiding: volatile asm keeps ptxas from sinking a
// load next to its first use (it does, to save registers, and every load then costs a full LDS latency).
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
template <int IMM>
__device__ __forceinline__ float2 lds64(unsigned addr) {
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2+%3];" : "=f"(v.x), "=f"(v.y) : "r"(addr), "n"(IMM));
    return v;
}
template <int IMM>
__device__ __forceinline__ void sts64(unsigned addr, float x, float y) {
    asm volatile("st.shared.v2.f32 [%0+%1], {%2, %3};" ::"r"(addr), "n"(IMM), "f"(x), "f"(y) : "memory");
}
__device__ __forceinline__ float lds32(unsigned addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
    return v;
}
template <int IMM>
__device__ __forceinline__ cplx64 lds_c(unsigned addr) {
    cplx64 v;
    asm volatile("ld.shared.b64 %0, [%1+%2];" : "=l"(v) : "r"(addr), "n"(IMM));
    return v;
}
template <int IMM>
__device__ __forceinline__ void sts_c(unsigned addr, cplx64 v) {
    asm volatile("st.shared.b64 [%0+%1], %2;" ::"r"(addr), "n"(IMM), "l"(v) : "memory");
}

// ---- staging of one (clip, tile, stem): planar L / R sample rows in shared memory --------------------
// Interior tiles whose rows are 16-byte aligned in global memory are staged by the TMA engine: ONE thread
// issues two bulk copies (cp.async.bulk, 11 KB each for hop 256) that complete on an mbarrier; no thread
// spends instructions on the copy and the data does not pass through registers or L1.  Clip-edge tiles
// (reflect padding) and unaligned inputs fall back to cooperative loads.
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "MIXFE_WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra MIXFE_DONE_%=;\n"
        "bra MIXFE_WAIT_%=;\n"
        "MIXFE_DONE_%=:\n"
        "}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

struct K1Stage {                     // where the samples of one (clip, tile, stem) come from
    const float* gL;
    const float* gR;
    long long base;                  // first staged sample (may be negative / run past T at the clip edges)
    bool tma;
};

__device__ __forceinline__ bool k1_stage_is_tma(const K1Params& p, int tile) {
    return tile >= p.tma_lo && tile <= p.tma_hi;
}

__device__ __forceinline__ K1Stage k1_stage_of(const K1Params& p, int b, int tile, int s) {
    K1Stage st;
    st.gL = p.stem[s] + (long long)b * p.stride_b;
    st.gR = st.gL + p.stride_c;
    st.base = (long long)tile * kTF * p.hop - kNfft / 2;
    st.tma = k1_stage_is_tma(p, tile);
    return st;
}

// one thread: arm the barrier and start both row copies
__device__ __forceinline__ void k1_stage_tma(const K1Params& p, const K1Stage& st, float* xsL, unsigned bar) {
    const unsigned bytes = (unsigned)p.nxs * 4u;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // earlier generic reads of xs vs the async writes
    mbar_expect_tx(bar, 2u * bytes);
    tma_bulk_g2s(smem_u32(xsL), st.gL + st.base, bytes, bar);
    tma_bulk_g2s(smem_u32(xsL + p.nxs), st.gR + st.base, bytes, bar);
}

// all threads: loads with torch's 'reflect' indexing (clip edges) or plain loads (unaligned rows)
__device__ __forceinline__ void k1_stage_coop(const K1Params& p, const K1Stage& st, float* xsL, int tid) {
    float* xsR = xsL + p.nxs;
    for (int i = tid; i < p.nxs; i += kK1Threads) {
        long long j = st.base + i;
        if (j < 0) j = -j;                               // torch 'reflect': -i -> i
        if (j >= p.T) j = 2ll * (p.T - 1) - j;           // T-1+i -> T-1-i
        j = j < 0 ? 0 : (j >= p.T ? p.T - 1 : j);        // frames past F only; keeps loads in bounds
        xsL[i] = __ldg(st.gL + j);
        xsR[i] = __ldg(st.gR + j);
    }
}

// Time-domain partials of the hop region a tile owns.  A tile writes TWO records of kTdN floats, one per half
// of its region: per stem (slots 8s..8s+7) sum L, sum R, sum L^2, sum R^2, sum (L-R)^2, sum (L+R)^2, max|L|,
// max|R|; the mixture energy of the half in four pieces (slots 32, 37, 38, 39); the masking sums of the whole
// tile in slots 33..36 of record 0.
//
// Roles instead of a CTA-wide reduction: warp w sums stem (w & 3) over half (w >> 2) of the region -- two rows,
// 14 running values, one warp-shuffle reduction, no shared memory and no barrier -- and then the mixture energy
// of eighth w (all 8 rows).  Sums run on pairs of consecutive samples (packed f32x2 adds / FMAs).
struct K1StemAcc {
    cplx64 sL, sR, sLL, sRR, sDD, sSS;
    float mxL, mxR;
};

__device__ __forceinline__ void k1_stem_acc(K1StemAcc& a, float l0, float l1, float r0, float r1) {
    const cplx64 L = cpack(l0, l1), R = cpack(r0, r1);
    a.sL = cadd(a.sL, L);
    a.sR = cadd(a.sR, R);
    a.sLL = cfma(L, L, a.sLL);
    a.sRR = cfma(R, R, a.sRR);
    const cplx64 d = csub(L, R), m = cadd(L, R);
    a.sDD = cfma(d, d, a.sDD);
    a.sSS = cfma(m, m, a.sSS);
    a.mxL = fmaxf(fmaxf(a.mxL, fabsf(l0)), fabsf(l1));
    a.mxR = fmaxf(fmaxf(a.mxR, fabsf(r0)), fabsf(r1));
}

__device__ __forceinline__ float cfold(cplx64 v) {
    float x, y;
    cunpack(v, x, y);
    return x + y;
}

// (P: K1Params or K1EParams -- any parameter block with stem, stride_b, stride_c, T, hop, n_tiles, part_td; an eighth of the
// region a tile owns is p.hop / DIV samples: DIV = 1 for K1's tiles of 8 frames, 2 for K1E's tiles of 4.  The value is
// re-read from the parameter bank at every use: as a register it pushed K1 over its 128.)
template <int DIV, class P>
__device__ __forceinline__ void k1_time_stats(const P& p, int b, int tile, int tid) {
    const int lane = tid & 31, warp = tid >> 5;
    const long long own0 = (long long)tile * kTF * (p.hop / DIV);
    const long long rem = (long long)p.T - own0;
    const int own_n = rem < 0 ? 0 : (rem > (long long)kTF * (p.hop / DIV) ? kTF * (p.hop / DIV) : (int)rem);
    const long long cb = (long long)b * p.stride_b + own0;
    unsigned long long amask = (unsigned long long)((p.hop / DIV) & 3);
#pragma unroll
    for (int s = 0; s < kNumStems; ++s)
        amask |= reinterpret_cast<unsigned long long>(p.stem[s] + cb) |
                 reinterpret_cast<unsigned long long>(p.stem[s] + cb + p.stride_c);
    const bool vec = (amask & 15ull) == 0;       // every row start 16-byte aligned, role ranges start on float4 boundaries
    float* rec = p.part_td + ((long long)b * p.n_tiles + tile) * (2 * kTdN) + (warp >> 2) * kTdN;

    // ---- role A: stem (warp & 3), half (warp >> 2) -------------------------------------------------
    {
        const int s = warp & 3;
        const int r0 = (warp >> 2) * (kTF / 2) * (p.hop / DIV);
        const int r1 = min(own_n, r0 + (kTF / 2) * (p.hop / DIV));
        const float* gL = p.stem[s] + cb;
        const float* gR = gL + p.stride_c;
        K1StemAcc a;
        a.sL = a.sR = a.sLL = a.sRR = a.sDD = a.sSS = 0ull;
        a.mxL = a.mxR = 0.0f;
        int i = r0;
        if (vec) {
            const int r1v = r0 + ((r1 - r0) > 0 ? ((r1 - r0) & ~3) : 0);
#pragma unroll 1
            for (int i4 = r0 + 4 * lane; i4 < r1v; i4 += 8 * 128) {
                float4 vL[8], vR[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {                // up to 16 x 128-bit loads in flight per lane
                    const bool in = i4 + 128 * u < r1v;
                    vL[u] = in ? __ldg(reinterpret_cast<const float4*>(gL + i4 + 128 * u)) : make_float4(0.f, 0.f, 0.f, 0.f);
                    vR[u] = in ? __ldg(reinterpret_cast<const float4*>(gR + i4 + 128 * u)) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    k1_stem_acc(a, vL[u].x, vL[u].y, vR[u].x, vR[u].y);
                    k1_stem_acc(a, vL[u].z, vL[u].w, vR[u].z, vR[u].w);
                }
            }
            i = r1v;
        }
        for (int j = i + lane; j < r1; j += 32) k1_stem_acc(a, __ldg(gL + j), 0.0f, __ldg(gR + j), 0.0f);
        float v[6] = {cfold(a.sL), cfold(a.sR), cfold(a.sLL), cfold(a.sRR), cfold(a.sDD), cfold(a.sSS)};
#pragma unroll
        for (int q = 0; q < 6; ++q) v[q] = warp_sum(v[q]);
        const float mL = warp_max(a.mxL), mR = warp_max(a.mxR);
        if (lane == 0) {
            float4* dst = reinterpret_cast<float4*>(rec + s * 8);
            dst[0] = make_float4(v[0], v[1], v[2], v[3]);
            dst[1] = make_float4(v[4], v[5], mL, mR);
        }
    }
    // ---- role M: mixture energy of eighth `warp` ----------------------------------------------------
    {
        const int r0 = warp * (p.hop / DIV);
        const int r1 = min(own_n, r0 + (p.hop / DIV));
        cplx64 e = 0ull;
        int i = r0;
        if (vec) {
            const int r1v = r0 + ((r1 - r0) > 0 ? ((r1 - r0) & ~3) : 0);
#pragma unroll 1
            for (int i4 = r0 + 4 * lane; i4 < r1v; i4 += 2 * 128) {
                float4 vL[2][kNumStems], vR[2][kNumStems];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    const bool in = i4 + 128 * u < r1v;
#pragma unroll
                    for (int s = 0; s < kNumStems; ++s) {
                        vL[u][s] = in ? __ldg(reinterpret_cast<const float4*>(p.stem[s] + cb + i4 + 128 * u)) : make_float4(0.f, 0.f, 0.f, 0.f);
                        vR[u][s] = in ? __ldg(reinterpret_cast<const float4*>(p.stem[s] + cb + p.stride_c + i4 + 128 * u)) : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                }
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    // ((0 + v) + b) + d) + o, as sum(stems.values())
                    const cplx64 l01 = cadd(cadd(cadd(cpack(vL[u][0].x, vL[u][0].y), cpack(vL[u][1].x, vL[u][1].y)), cpack(vL[u][2].x, vL[u][2].y)), cpack(vL[u][3].x, vL[u][3].y));
                    const cplx64 l23 = cadd(cadd(cadd(cpack(vL[u][0].z, vL[u][0].w), cpack(vL[u][1].z, vL[u][1].w)), cpack(vL[u][2].z, vL[u][2].w)), cpack(vL[u][3].z, vL[u][3].w));
                    const cplx64 q01 = cadd(cadd(cadd(cpack(vR[u][0].x, vR[u][0].y), cpack(vR[u][1].x, vR[u][1].y)), cpack(vR[u][2].x, vR[u][2].y)), cpack(vR[u][3].x, vR[u][3].y));
                    const cplx64 q23 = cadd(cadd(cadd(cpack(vR[u][0].z, vR[u][0].w), cpack(vR[u][1].z, vR[u][1].w)), cpack(vR[u][2].z, vR[u][2].w)), cpack(vR[u][3].z, vR[u][3].w));
                    e = cfma(l01, l01, cfma(q01, q01, e));
                    e = cfma(l23, l23, cfma(q23, q23, e));
                }
            }
            i = r1v;
        }
        float es = cfold(e);
        for (int j = i + lane; j < r1; j += 32) {
            float mL = 0.0f, mR = 0.0f;
#pragma unroll
            for (int s = 0; s < kNumStems; ++s) {
                mL += __ldg(p.stem[s] + cb + j);
                mR += __ldg(p.stem[s] + cb + p.stride_c + j);
            }
            es = fmaf(mL, mL, fmaf(mR, mR, es));
        }
        es = warp_sum(es);
        if (lane == 0) rec[(warp & 3) == 0 ? 32 : 36 + (warp & 3)] = es;
    }
}

// The same roles on rows the TMA engine has put into shared memory: row (2 stem + channel) holds the tile's own
// region (kTF * hop <= 2048 samples, all of them inside the clip) at rows + 8 KB * row.
template <int DIV, class P>
__device__ __forceinline__ void k1_time_stats_smem(const P& p, const unsigned char* rows, int b, int tile, int tid) {
    const int lane = tid & 31, warp = tid >> 5;
    float* rec = p.part_td + ((long long)b * p.n_tiles + tile) * (2 * kTdN) + (warp >> 2) * kTdN;
    {
        const int s = warp & 3;
        const int n4 = (kTF / 2) * (p.hop / DIV) / 4;                    // float4 per row of this half
        const float4* rL = reinterpret_cast<const float4*>(rows + (2 * s) * kWarpRegion) + (warp >> 2) * n4;
        const float4* rR = rL + kWarpRegion / 16;
        K1StemAcc a;
        a.sL = a.sR = a.sLL = a.sRR = a.sDD = a.sSS = 0ull;
        a.mxL = a.mxR = 0.0f;
#pragma unroll 4
        for (int i = lane; i < n4; i += 32) {
            const float4 vL = rL[i], vR = rR[i];
            k1_stem_acc(a, vL.x, vL.y, vR.x, vR.y);
            k1_stem_acc(a, vL.z, vL.w, vR.z, vR.w);
        }
        float v[6] = {cfold(a.sL), cfold(a.sR), cfold(a.sLL), cfold(a.sRR), cfold(a.sDD), cfold(a.sSS)};
#pragma unroll
        for (int q = 0; q < 6; ++q) v[q] = warp_sum(v[q]);
        const float mL = warp_max(a.mxL), mR = warp_max(a.mxR);
        if (lane == 0) {
            float4* dst = reinterpret_cast<float4*>(rec + s * 8);
            dst[0] = make_float4(v[0], v[1], v[2], v[3]);
            dst[1] = make_float4(v[4], v[5], mL, mR);
        }
    }
    {
        const int n4 = (p.hop / DIV) / 4;                                // float4 per row of this eighth
        const float4* r0 = reinterpret_cast<const float4*>(rows) + warp * n4;
        cplx64 e = 0ull;
#pragma unroll 2
        for (int i = lane; i < n4; i += 32) {
            float4 vL[kNumStems], vR[kNumStems];
#pragma unroll
            for (int s = 0; s < kNumStems; ++s) {
                vL[s] = r0[(2 * s) * (kWarpRegion / 16) + i];
                vR[s] = r0[(2 * s + 1) * (kWarpRegion / 16) + i];
            }
            // ((0 + v) + b) + d) + o, as sum(stems.values())
            const cplx64 l01 = cadd(cadd(cadd(cpack(vL[0].x, vL[0].y), cpack(vL[1].x, vL[1].y)), cpack(vL[2].x, vL[2].y)), cpack(vL[3].x, vL[3].y));
            const cplx64 l23 = cadd(cadd(cadd(cpack(vL[0].z, vL[0].w), cpack(vL[1].z, vL[1].w)), cpack(vL[2].z, vL[2].w)), cpack(vL[3].z, vL[3].w));
            const cplx64 q01 = cadd(cadd(cadd(cpack(vR[0].x, vR[0].y), cpack(vR[1].x, vR[1].y)), cpack(vR[2].x, vR[2].y)), cpack(vR[3].x, vR[3].y));
            const cplx64 q23 = cadd(cadd(cadd(cpack(vR[0].z, vR[0].w), cpack(vR[1].z, vR[1].w)), cpack(vR[2].z, vR[2].w)), cpack(vR[3].z, vR[3].w));
            e = cfma(l01, l01, cfma(q01, q01, e));
            e = cfma(l23, l23, cfma(q23, q23, e));
        }
        const float es = warp_sum(cfold(e));
        if (lane == 0) rec[(warp & 3) == 0 ? 32 : 36 + (warp & 3)] = es;
    }
}

__global__ void __launch_bounds__(kK1Threads, 2) k1_features(const __grid_constant__ K1Params p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    float* const xsL = reinterpret_cast<float*>(smem);                        // [nxs] left, then [nxs] right
    unsigned char* const ex = smem + p.off_ex;                                // 8 warp regions of 8 KB
    float2* const tw = reinterpret_cast<float2*>(smem + p.off_tw);              // [31][32], rows k1 = 1..31
    float4* const S = reinterpret_cast<float4*>(smem + p.off_s);              // [3][256]: thread-private S values of stems 0..2
    float* const wsteps = reinterpret_cast<float*>(smem + p.off_wsteps);      // [n_wt]
    unsigned short* const flo = reinterpret_cast<unsigned short*>(smem + p.off_flo);   // [kMaxMels]
    const unsigned bar = smem_u32(smem + p.off_red);                          // mbarrier of the TMA staging
    int* const arrivals = reinterpret_cast<int*>(smem + p.off_red + 8);       // warps done reading xs
    const unsigned bar2 = smem_u32(smem + p.off_red + 16);                    // mbarrier of the statistics rows

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    long long item = blockIdx.x;
    if (item >= p.total_items) return;                  // (the launcher never starts more CTAs than items)
    int* const next_slot = reinterpret_cast<int*>(smem + p.off_red + 24);

    // ---- constant tables -> shared memory (once per persistent CTA) ----------------------------
    for (int i = tid; i < 31 * 32; i += kK1Threads) tw[i] = p.tw[32 + i];      // rows k1 = 1..31
    for (int i = tid; i < p.n_wt; i += kK1Threads) wsteps[i] = p.wsteps[i];
    for (int i = tid; i < kMaxMels; i += kK1Threads) flo[i] = p.flo[i];

    unsigned char* const region = ex + warp * kWarpRegion;
    float2* const exw = reinterpret_cast<float2*>(region);
    float2* const pv = reinterpret_cast<float2*>(region + ((warp >> 1) & 1) * 64);

    int b = (int)(item / p.n_tiles), tile = (int)(item - (long long)b * p.n_tiles);
    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_init(bar2, 1);
        *arrivals = 0;
    }
    __syncthreads();
    bool cur_tma;                    // how the stage about to be consumed was filled (uniform over the CTA)
    unsigned tma_parity = 0, stats_parity = 0;
    {
        const K1Stage st = k1_stage_of(p, b, tile, 0);
        cur_tma = st.tma;
        if (st.tma) {
            if (tid == 0) k1_stage_tma(p, st, xsL, bar);
        } else {
            k1_stage_coop(p, st, xsL, tid);
        }
    }

    for (;;) {
        const int t0 = tile * kTF;
        const int nvalid = p.F - t0 < kTF ? p.F - t0 : kTF;          // frames of this tile that exist
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
            __syncthreads();                 // B0: xs holds (tile, s); the mel walk of the previous stem has left the power columns
            if (p.work && s == kNumStems - 1) next_item = (long long)gridDim.x + *next_slot;
            bool stats_rows = false;
            if (s == 0 && p.want_stats) {
                // The tile's own region of all 8 rows (64 KB = the idle exchange regions) by TMA when it lies inside
                // the clip and every row is 16-byte aligned; else role loads straight from global memory.
                if (tile <= p.rows_hi) {
                    if (tid == 0) {
                        const long long cb = (long long)b * p.stride_b + (long long)tile * kTF * p.hop;
                        const unsigned bytes = (unsigned)(kTF * p.hop) * 4u;
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        mbar_expect_tx(bar2, 8u * bytes);
#pragma unroll
                        for (int q = 0; q < kNumStems; ++q) {
                            tma_bulk_g2s(smem_u32(ex + (2 * q) * kWarpRegion), p.stem[q] + cb, bytes, bar2);
                            tma_bulk_g2s(smem_u32(ex + (2 * q + 1) * kWarpRegion), p.stem[q] + cb + p.stride_c, bytes, bar2);
                        }
                    }
                    stats_rows = true;       // consumed below, after this warp's frame is in registers
                } else {
                    k1_time_stats<1>(p, b, tile, tid);
                }
            }

            // ---- frame t0 + warp -> registers, Hann window on the fly ------------------------
            cplx64 z[32];                    // (L, R) sample pairs = one complex value each
            const bool active = warp < nvalid;
            if (active) {
                const float2 wl = tw[lane];             // row k1 = 1: exp(-2 pi i lane/1024)
                const float cb = wl.x, sb = -wl.y;
                const unsigned xa = smem_u32(xsL + warp * p.hop + lane);
                const unsigned xr_off = (unsigned)p.nxs * 4u;
                float fl[32], fr[32];
#pragma unroll
                for (int n1 = 0; n1 < 32; ++n1) {       // all 64 loads in flight before the first multiply
                    fl[n1] = lds32(xa + n1 * 128);
                    fr[n1] = lds32(xa + xr_off + n1 * 128);
                }
#pragma unroll
                for (int n1 = 0; n1 < 32; ++n1) z[n1] = cpack(fl[n1], fr[n1]);
#pragma unroll
                for (int n1 = 0; n1 < 32; ++n1) {
                    // 0.5 * hann[32 n1 + lane]; cos(a + b) with a = 2 pi n1/32 (immediate), b per lane.
                    // The extra 0.5 makes |X +- conj(X')|^2 the channel powers without a final scale.
                    const float ca = n1 < 16 ? w32_re(n1) : -w32_re(n1 - 16);
                    const float sa = n1 < 16 ? w32_sin(n1) : -w32_sin(n1 - 16);
                    const float wv = fmaf(0.25f * sa, sb, fmaf(-0.25f * ca, cb, 0.25f));
                    float xl, xr;
                    cunpack(z[n1], xl, xr);
                    z[n1] = cpack(xl * wv, xr * wv);
                }
            }
            // ---- prefetch the next (tile, stem) while this one is transformed -------------------
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
                    next_tma = k1_stage_is_tma(p, ntile);
                    if (next_tma) {
                        // no CTA barrier: the last warp to finish reading xs starts the copy
                        __syncwarp();
                        if (lane == 0) {
                            __threadfence_block();
                            if (atomicAdd(arrivals, 1) == kK1Threads / 32 - 1) {
                                *arrivals = 0;
                                k1_stage_tma(p, k1_stage_of(p, nb, ntile, ns), xsL, bar);
                            }
                        }
                    } else {
                        __syncthreads();     // B0': every warp has its frame in registers, xs is dead
                        k1_stage_coop(p, k1_stage_of(p, nb, ntile, ns), xsL, tid);
                    }
                }
                cur_tma = next_tma;
                if (ns == 0 && next_item < p.total_items && p.want_stats) {
                    // pull the hop region the next tile owns towards L2 for stems 1..3 (stem 0 comes with its
                    // staging): the statistics loads at the top of the next tile then hit L2 instead of HBM
                    const long long own0 = (long long)ntile * kTF * p.hop;
                    const int lines_per_row = (kTF * p.hop * 4) >> 7;
                    for (int l = tid; l < lines_per_row; l += kK1Threads) {
                        if (own0 + l * 32 < p.T) {
                            const float* g0 = p.stem[1] + (long long)nb * p.stride_b + own0 + l * 32;
                            const long long ds = p.stem[2] - p.stem[1], dd = p.stem[3] - p.stem[1];
                            asm volatile("prefetch.global.L2 [%0];" ::"l"(g0));
                            asm volatile("prefetch.global.L2 [%0];" ::"l"(g0 + p.stride_c));
                            asm volatile("prefetch.global.L2 [%0];" ::"l"(g0 + ds));
                            asm volatile("prefetch.global.L2 [%0];" ::"l"(g0 + ds + p.stride_c));
                            asm volatile("prefetch.global.L2 [%0];" ::"l"(g0 + dd));
                            asm volatile("prefetch.global.L2 [%0];" ::"l"(g0 + dd + p.stride_c));
                        }
                    }
                }
            }

            if (stats_rows) {                // the copy had the frame load to land
                mbar_wait(bar2, stats_parity);
                stats_parity ^= 1u;
                k1_time_stats_smem<1>(p, ex, b, tile, tid);
                __syncthreads();             // the rows are dead: the exchange regions go back to their warps
            }

            if (active) {
#pragma unroll 1
                for (int pass = 0; pass < 2; ++pass) {          // one copy of the 32-point DFT in the instruction stream
                    dft32_dif_packed(z);
                    if (pass == 0) {
                        // inter-pass twiddles in batches of 8, the next batch in flight while this one is used;
                        // element (k1, lane) of the exchange goes to row k1, column lane of the padded layout
                        const unsigned twa = smem_u32(tw + lane) - 256;            // the table starts at row k1 = 1
                        const unsigned wa = smem_u32(exw) + lane * 8;
                        float2 tq[2][8];
#pragma unroll
                        for (int q = 1; q < 8; ++q) tq[0][q] = lds64<0>(twa + q * 256);
#pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            if (g < 3) {
#pragma unroll
                                for (int q = 0; q < 8; ++q) tq[(g + 1) & 1][q] = lds64<0>(twa + (8 * g + 8 + q) * 256);
                            }
#pragma unroll
                            for (int q = 0; q < 8; ++q) {
                                const int k1 = 8 * g + q;
                                if (k1 == 0) {
                                    sts_c<0>(wa, z[0]);
                                } else {
                                    float yr, yi;
                                    cunpack(z[brev5(k1)], yr, yi);
                                    const float2 t = tq[g & 1][q];
                                    sts64<0>(wa + k1 * kExRow, fmaf(yr, t.x, -yi * t.y), fmaf(yr, t.y, yi * t.x));
                                }
                            }
                        }
                        __syncwarp();
                        const unsigned ra = smem_u32(exw) + lane * kExRow;
#pragma unroll
                        for (int n2 = 0; n2 < 32; ++n2) z[n2] = lds_c<0>(ra + n2 * 8);
                        __syncwarp();            // the exchange buffer becomes the power column
                    }
                }
                // lane holds X[lane + 32 k2] at index brev5(k2).  Untangle L (real part) and R (imag part):
                // |L_k|^2 = |X_k + conj(X_{N-k})|^2, |R_k|^2 = |X_k - conj(X_{N-k})|^2 (scale folded into the window).
                const int src = (32 - lane) & 31;
#pragma unroll
                for (int k2 = 0; k2 < 16; ++k2) {
                    float pr, pi;
                    cunpack(z[brev5(31 - k2)], pr, pi);
                    float qr = __shfl_sync(kFullMask, pr, src);
                    float qi = __shfl_sync(kFullMask, pi, src);
                    if (lane == 0) {                         // N - 32 k2 lives in lane 0 itself, register (32 - k2) & 31
                        cunpack(z[brev5((32 - k2) & 31)], qr, qi);
                    }
                    const cplx64 x = z[brev5(k2)], q = cpack(qr, -qi);       // conj(X[N-k])
                    float a, bb, c, d;
                    cunpack(cadd(x, q), a, bb);                              // 2 L_k
                    cunpack(csub(x, q), c, d);                               // 2 i R_k
                    pv[lane + 32 * k2] = make_float2(fmaf(a, a, bb * bb), fmaf(c, c, d * d));
                }
                if (lane == 0) {
                    float xr, xi;
                    cunpack(z[brev5(16)], xr, xi);                           // bin 512 is its own partner
                    pv[512] = make_float2(4.0f * xr * xr, 4.0f * xi * xi);
                }
            }
            __syncthreads();                 // B1: the 8 power columns of the tile are complete

            // ---- tile-level mel projection fused with the epilogue ------------------------------------
            // lane (q = lane >> 3, i = lane & 7) of warp w walks filter i of the warp's two groups over frames
            // 2q and 2q + 1: per step one weight (shared by the 4 quarters) and two 64-bit (L, R) powers.
            {
                const int q = lane >> 3, fi = lane & 7;
                const bool ok0 = 2 * q < nvalid, ok1 = 2 * q + 1 < nvalid;
                const unsigned char* col = ex + 2 * q * kWarpRegion + (q & 1) * 64;
                const long long tile_id = (long long)b * p.n_tiles + tile;
                float* const lm_tile = p.logmel + (long long)b * p.lm_stride_b + t0 + 2 * q;      // + (ch n_mels + m) F
                float sc[4];                                  // S = (mel_L + mel_R) / 2 of (group a, b) x (frame 2q, 2q+1)
#pragma unroll
                for (int ab = 0; ab < 2; ++ab) {
                    const int n = p.wmeta[warp][ab][0];
                    const int m = 8 * p.wmeta[warp][ab][2] + fi;
                    float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f, a3 = 0.0f, c0 = 0.0f, c1 = 0.0f, c2 = 0.0f, c3 = 0.0f;
                    if (n > 0) {
                        const unsigned char* p0 = col + flo[m];
                        const float* wp = wsteps + p.wmeta[warp][ab][1] + fi;
#pragma unroll 4
                        for (int i = 0; i < n; i += 2) {
                            const float w0 = wp[8 * i], w1 = wp[8 * i + 8];
                            const float2 u0 = *reinterpret_cast<const float2*>(p0 + 8 * i);
                            const float2 u1 = *reinterpret_cast<const float2*>(p0 + 8 * i + 8);
                            const float2 v0 = *reinterpret_cast<const float2*>(p0 + kWarpRegion + 8 * i);
                            const float2 v1 = *reinterpret_cast<const float2*>(p0 + kWarpRegion + 8 * i + 8);
                            a0 = fmaf(w0, u0.x, a0);
                            a1 = fmaf(w0, u0.y, a1);
                            a2 = fmaf(w0, v0.x, a2);
                            a3 = fmaf(w0, v0.y, a3);
                            c0 = fmaf(w1, u1.x, c0);
                            c1 = fmaf(w1, u1.y, c1);
                            c2 = fmaf(w1, v1.x, c2);
                            c3 = fmaf(w1, v1.y, c3);
                        }
                    }
                    const float v[4] = {a0 + c0, a1 + c1, a2 + c2, a3 + c3};     // (frame 2q: L, R), (frame 2q+1: L, R)
                    float lg[4];
#pragma unroll
                    for (int k = 0; k < 4; ++k) lg[k] = ln_fast(v[k] + 1e-10f);
                    if (p.logmel && m < p.n_mels) {
                        // The thread's two frames are neighbours in time and the 4 quarters of the warp cover the tile's 8
                        // frames: one store instruction writes 8 filters x 32 contiguous bytes (whole sectors), straight
                        // from the registers -- no staging tile, no store phase, no barrier.
                        float* d0 = lm_tile + ((long long)(2 * s) * p.n_mels + m) * p.F;
                        float* d1 = d0 + (long long)p.n_mels * p.F;
                        if (ok0) { d0[0] = lg[0]; d1[0] = lg[1]; }
                        if (ok1) { d0[1] = lg[2]; d1[1] = lg[3]; }
                    }
                    if (p.melpow && m < p.n_mels) {
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const int t = 2 * q + (k >> 1), ch = k & 1;
                            if (t < nvalid)
                                p.melpow[(long long)b * p.mp_stride_b + ((long long)(2 * s + ch) * p.n_mels + m) * p.F + t0 + t] = v[k];
                        }
                    }
                    if (p.want_stats) {
                        float slog = (ok0 ? lg[0] + lg[1] : 0.0f) + (ok1 ? lg[2] + lg[3] : 0.0f);
                        float smel = (ok0 ? v[0] + v[1] : 0.0f) + (ok1 ? v[2] + v[3] : 0.0f);
                        slog += __shfl_xor_sync(kFullMask, slog, 8);
                        smel += __shfl_xor_sync(kFullMask, smel, 8);
                        slog += __shfl_xor_sync(kFullMask, slog, 16);
                        smel += __shfl_xor_sync(kFullMask, smel, 16);
                        if (q == 0 && m < p.n_mels_pad)
                            p.part_spec[(tile_id * kNumStems + s) * p.n_mels_pad + m] = make_float2(slog, smel);
                    }
                    sc[2 * ab] = 0.5f * (v[0] + v[1]);
                    sc[2 * ab + 1] = 0.5f * (v[2] + v[3]);
                }
                if (p.want_stats) {
                    if (s < 3) {
                        S[s * kK1Threads + tid] = make_float4(sc[0], sc[1], sc[2], sc[3]);
                    } else {
                        const float4 A0 = S[tid], A1 = S[kK1Threads + tid], A2 = S[2 * kK1Threads + tid];
                        const float s0[4] = {A0.x, A0.y, A0.z, A0.w}, s1[4] = {A1.x, A1.y, A1.z, A1.w};
                        const float s2[4] = {A2.x, A2.y, A2.z, A2.w};
                        const bool fa = 8 * p.wmeta[warp][0][2] + fi < p.n_mels, fb = 8 * p.wmeta[warp][1][2] + fi < p.n_mels;
                        const bool ok[4] = {fa && ok0, fa && ok1, fb && ok0, fb && ok1};
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            if (ok[i]) {
                                const float m01 = fmaxf(s0[i], s1[i]), m23 = fmaxf(s2[i], sc[i]);
                                // sigmoid(max_{j != i} S_j - S_i) = 1 / (1 + exp(S_i - max_others))
                                msum[0] += fast_sigmoid_neg(s0[i] - fmaxf(s1[i], m23));
                                msum[1] += fast_sigmoid_neg(s1[i] - fmaxf(s0[i], m23));
                                msum[2] += fast_sigmoid_neg(s2[i] - fmaxf(m01, sc[i]));
                                msum[3] += fast_sigmoid_neg(sc[i] - fmaxf(m01, s2[i]));
                            }
                        }
#pragma unroll
                        for (int i = 0; i < 4; ++i) msum[i] = warp_sum(msum[i]);
                        if (lane == 0)           // the warp's own slots of the tile's records: no CTA-wide reduction
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
    // the last CTA to finish leaves the counters at zero for the next launch (launches of one plan are stream-ordered)
    if (p.work && tid == 0) {
        if (atomicAdd(p.work + 1, 1) == (int)gridDim.x - 1) {
            p.work[0] = 0;
            p.work[1] = 0;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// K2: reduce the per-tile partials of one clip and evaluate the feature formulas.
// ------------------------------------------------------------------------------------------------
struct K2Params {
    const float2* part_spec;
    const float* part_td;
    float* features;                 // [B][D]
    int B, T, F, n_mels, n_mels_pad, n_tiles;
    int n_bins;                      // 0: low/mid/high bands, else detailed curve of n_bins points
    int D;
    int n_spec_tiles;                // entries of part_spec per clip (n_tiles for K1 / K1G, 2 n_tiles for K1E: one per frame pair)
    int mask_per_warp;               // masking sums: 1 = slots kTdMask.. of both records (K1), 0 = slots 33..36 of record 0 (K1G)
};

__device__ __forceinline__ void put_feature(float* out, int idx, double v) {
    float f = (float)v;
    f = fminf(fmaxf(f, -100.0f), 100.0f);       // clamp (src/mixing_utils.py:339); fminf/fmaxf drop NaN ...
    if (v != v) f = 0.0f;                       // ... so scrub it explicitly (:343-348)
    out[idx] = f;
}

constexpr int kK2Threads = 512;

__global__ void __launch_bounds__(kK2Threads) k2_finalize(const K2Params p) {
    __shared__ double sE[kNumStems][kMaxMels];       // mean dB per mel band
    __shared__ double sCurve[kNumStems][kMaxMels];
    __shared__ double sLogTot[kNumStems][4], sMelTot[kNumStems][4];
    __shared__ double sTd[kTdN];
    __shared__ double sAcc[2][kNumStems][kMaxMels][2];   // (sum log, sum mel) per half of the tiles

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.x;
    const double kDb = 4.342944819032518;            // 10 / ln(10)

    // 1. spectral sums over tiles.  The loop is bound by the latency of its loads (72 CTAs have few bytes in flight), so
    //    a thread takes two filters (one 16-byte load) of HALF the tiles, 12 loads in flight; the halves meet in shared memory.
    {
        const int h = tid >> 8, s = (tid >> 6) & 3, m2 = 2 * (tid & 63);
        double acc[4] = {0.0, 0.0, 0.0, 0.0};                 // (sum log, sum mel) of filters m2 and m2 + 1
        if (m2 < p.n_mels_pad) {                              // n_mels_pad is a multiple of 16: m2 + 1 lies inside the row too
            const float4* src = reinterpret_cast<const float4*>(
                p.part_spec + ((long long)b * p.n_spec_tiles * kNumStems + s) * p.n_mels_pad + m2);
            const long long step = (long long)kNumStems * p.n_mels_pad / 2;       // float4 per tile
            const int t_mid = (p.n_spec_tiles + 1) / 2;
            const int t0 = h ? t_mid : 0, t1 = h ? p.n_spec_tiles : t_mid;
            constexpr int kU = 12;
            for (int t = t0; t < t1; t += kU) {
                float4 v[kU];
#pragma unroll
                for (int u = 0; u < kU; ++u) v[u] = t + u < t1 ? __ldg(src + (t + u) * step) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int u = 0; u < kU; ++u) {
                    acc[0] += (double)v[u].x; acc[1] += (double)v[u].y;
                    acc[2] += (double)v[u].z; acc[3] += (double)v[u].w;
                }
            }
        }
        if (m2 < kMaxMels) {
            sAcc[h][s][m2][0] = acc[0]; sAcc[h][s][m2][1] = acc[1];
            sAcc[h][s][m2 + 1][0] = acc[2]; sAcc[h][s][m2 + 1][1] = acc[3];
        }
    }
    __syncthreads();
    {   // thread = (stem, mel); filters beyond n_mels (padding the kernels may leave unwritten) do not count
        const int s = tid >> 7, m = tid & 127;
        double slog = 0.0, smel = 0.0;
        if (m < p.n_mels) {
            slog = sAcc[0][s][m][0] + sAcc[1][s][m][0];
            smel = sAcc[0][s][m][1] + sAcc[1][s][m][1];
            sE[s][m] = kDb * slog / (2.0 * p.F);
        } else if (m < kMaxMels) {
            sE[s][m] = 0.0;
        }
        const double a = warp_sum(slog), c = warp_sum(smel);
        if (lane == 0) {
            sLogTot[s][warp & 3] = a;
            sMelTot[s][warp & 3] = c;
        }
    }
    // 2. time-domain partials: 8 threads per quantity, two records per tile
    if (tid < kTdN * 8) {
        const int q = tid >> 3, sub = tid & 7;
        const bool is_max = q < 32 && (q & 7) >= 6;
        double v = 0.0;
        const bool used = p.mask_per_warp ? (q < 33 || q >= kTdMask) : q < 37;
        if (used) {
            const float* src = p.part_td + (long long)b * p.n_tiles * (2 * kTdN);
            const int n_rec = 2 * p.n_tiles, step = (q >= 33 && q < 37) ? 2 : 1;          // K1G's masking sums live in record 0 only
#pragma unroll 6
            for (int t = sub * step; t < n_rec; t += 8 * step) {
                const float* r = src + (long long)t * kTdN;
                double u = (double)__ldg(r + q);
                if (q == 32) u += (double)__ldg(r + 37) + (double)__ldg(r + 38) + (double)__ldg(r + 39);
                v = is_max ? fmax(v, u) : v + u;
            }
        }
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) {
            const double u = __shfl_xor_sync(kFullMask, v, o);
            v = is_max ? fmax(v, u) : v + u;
        }
        if (sub == 0) sTd[q] = v;
    }
    __syncthreads();

    const int sd = p.n_bins > 0 ? p.n_bins + 2 : 5;
    const int per = 6 + 1 + sd + 3;
    float* out = p.features + (long long)b * p.D;
    const double N = (double)p.T;

    // 3. spectral block: warp s handles stem s
    if (warp < kNumStems) {
        const int s = warp;
        // sorted key order: bass, drums, (masking), other, vocals   (src/mixing_utils.py:324)
        const int base = s == 0 ? 3 * per + 4 : (s == 1 ? 0 : (s == 2 ? per : 2 * per + 4));
        const int ob = base + 7;
        const int n = p.n_mels;
        const double* E = sE[s];
        int nc;                                      // length of the curve the tilt is computed on
        const double* C;
        if (p.n_bins > 0 && p.n_bins < n) {          // F.interpolate(mode='linear', align_corners=True), :212-215
            nc = p.n_bins;
            const float scale = nc > 1 ? (float)(n - 1) / (float)(nc - 1) : 0.0f;
            for (int i = lane; i < nc; i += 32) {
                const float src = scale * (float)i;
                const int i0 = (int)src;
                const int i1 = i0 + (i0 < n - 1 ? 1 : 0);
                const double l1 = (double)(src - (float)i0);
                sCurve[s][i] = (1.0 - l1) * E[i0] + l1 * E[i1];
            }
            __syncwarp();
            C = sCurve[s];
        } else {
            nc = n;
            C = E;
        }
        if (p.n_bins == 0) {
            const int lo = n / 4, hi = n / 4 * 3;    // :172-173
            double a = 0.0, c = 0.0, d = 0.0;
            for (int i = lane; i < n; i += 32) {
                const double e = E[i];
                if (i < lo) a += e;
                else if (i < hi) c += e;
                else d += e;
            }
            a = warp_sum(a); c = warp_sum(c); d = warp_sum(d);
            if (lane == 0) {
                put_feature(out, ob + 0, a / lo);
                put_feature(out, ob + 1, c / (hi - lo));
                put_feature(out, ob + 2, d / (n - hi));
            }
        } else {
            for (int i = lane; i < nc; i += 32) put_feature(out, ob + i, C[i]);
        }
        // tilt: 0 when the unbiased std of the curve is < 1e-6, else Pearson r against arange (:185-190)
        double sum = 0.0;
        for (int i = lane; i < nc; i += 32) sum += C[i];
        const double mean = warp_sum(sum) / nc;
        const double imean = 0.5 * (nc - 1);
        double m2 = 0.0, cov = 0.0, iv = 0.0;
        for (int i = lane; i < nc; i += 32) {
            const double dc = C[i] - mean, di = (double)i - imean;
            m2 += dc * dc;
            cov += dc * di;
            iv += di * di;
        }
        m2 = warp_sum(m2); cov = warp_sum(cov); iv = warp_sum(iv);
        if (lane == 0) {
            const double sdev = nc > 1 ? sqrt(m2 / (nc - 1)) : 0.0;
            double tilt = 0.0;
            if (!(sdev < 1e-6)) {
                tilt = cov / sqrt(iv * m2);
                tilt = fmin(fmax(tilt, -1.0), 1.0);           // torch.corrcoef clips to [-1, 1]
            }
            const int ot = ob + (p.n_bins > 0 ? p.n_bins : 3);
            put_feature(out, ot, tilt);
            const double cnt = 2.0 * (double)p.F * (double)n;
            const double lt = sLogTot[s][0] + sLogTot[s][1] + sLogTot[s][2] + sLogTot[s][3];
            const double mt = sMelTot[s][0] + sMelTot[s][1] + sMelTot[s][2] + sMelTot[s][3];
            put_feature(out, ot + 1, exp(lt / cnt) / (mt / cnt + 1e-10));     // flatness, :195
        }
    }
    // 4. dynamics / relative loudness / stereo: one thread per stem; masking: 4 threads
    if (tid >= 128 && tid < 128 + kNumStems) {
        const int s = tid - 128;
        const int base = s == 0 ? 3 * per + 4 : (s == 1 ? 0 : (s == 2 ? per : 2 * per + 4));
        const double* q = sTd + s * 8;
        const double sL = q[0], sR = q[1], sLL = q[2], sRR = q[3], sDD = q[4], sSS = q[5], pkL = q[6], pkR = q[7];
        const double rmsL = sqrt(sLL / N), rmsR = sqrt(sRR / N);
        const double loud = -0.691 + 10.0 * log10((sLL + sRR) / (2.0 * N) + 1e-10);       // :311-318
        const double mixl = -0.691 + 10.0 * log10(sTd[32] / (2.0 * N) + 1e-10);
        put_feature(out, base + 0, rmsL);
        put_feature(out, base + 1, rmsR);
        put_feature(out, base + 2, 20.0 * log10(pkL / (rmsL + 1e-8)));                      // :127
        put_feature(out, base + 3, 20.0 * log10(pkR / (rmsR + 1e-8)));
        put_feature(out, base + 4, loud);
        put_feature(out, base + 5, loud);
        put_feature(out, base + 6, loud - mixl);                                            // :94-97
        const int os = base + 7 + sd;
        put_feature(out, os + 0, 20.0 * log10(rmsL / (rmsR + 1e-8)));                       // ILD, :246-248
        const double sLR = 0.25 * (sSS - sDD);
        const double cov = sLR - sL * sR / N, vL = sLL - sL * sL / N, vR = sRR - sR * sR / N;
        put_feature(out, os + 1, cov / (sqrt(fmax(vL, 0.0) * fmax(vR, 0.0)) + 1e-8));       // :251-256
        put_feature(out, os + 2, (sDD / (4.0 * N)) / (sSS / (4.0 * N) + 1e-8));             // :259-263
    }
    if (tid >= 160 && tid < 160 + kNumStems) {
        const int i = tid - 160;
        const double msum = p.mask_per_warp ? sTd[kTdMask + i] + sTd[kTdMask + 4 + i] + sTd[kTdMask + 8 + i] + sTd[kTdMask + 12 + i]
                                            : sTd[33 + i];
        put_feature(out, 2 * per + i, msum / ((double)p.n_mels * (double)p.F));      // :296-306
    }
}

}  // namespace mixfe
