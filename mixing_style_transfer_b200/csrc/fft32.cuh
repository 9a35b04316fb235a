// Register-resident 32-point complex DFT, the building block of every FFT in libmixfe.
//
// A 1024-point transform is computed by ONE WARP as 32 x 32 (four-step): each lane runs a
// 32-point DFT on values it holds in registers, multiplies by the inter-pass twiddle, the warp
// transposes through shared memory, and each lane runs a second 32-point DFT.  This header holds
// only the per-lane arithmetic; it is __host__ __device__ so tests/host/test_fft32.cpp can check
// it against a naive DFT without a GPU.
//
// Replaces (reference side): torch.stft's rfft inside torchaudio.transforms.Spectrogram, reached
// from src/mixing_utils.py:159,280 and src/model.py:58 of the reference.
#pragma once

#ifndef __CUDACC__
#define MIXFE_HD inline
#else
#define MIXFE_HD __host__ __device__ __forceinline__
#endif

namespace mixfe {

// cos(2*pi*j/32), j = 0..8 ; sin(2*pi*j/32) = cos(2*pi*(8-j)/32)
MIXFE_HD constexpr float cos32(int j) {
    switch (j) {
        case 0: return 1.0f;
        case 1: return 0.98078528040323044913f;
        case 2: return 0.92387953251128675613f;
        case 3: return 0.83146961230254523708f;
        case 4: return 0.70710678118654752440f;
        case 5: return 0.55557023301960222474f;
        case 6: return 0.38268343236508977173f;
        case 7: return 0.19509032201612826785f;
        default: return 0.0f;
    }
}

// W_32^t = exp(-2*pi*i*t/32) for t in [0,16): real and (negated) imaginary parts.
MIXFE_HD constexpr float w32_re(int t) { return t <= 8 ? cos32(t) : -cos32(16 - t); }
MIXFE_HD constexpr float w32_sin(int t) { return t <= 8 ? cos32(8 - t) : cos32(t - 8); }   // sin(2 pi t/32) >= 0

MIXFE_HD constexpr int brev5(int i) {
    return ((i & 1) << 4) | ((i & 2) << 2) | (i & 4) | ((i & 8) >> 2) | ((i & 16) >> 4);
}

// In-place forward DFT of 32 complex values, radix-2 decimation in frequency, fully unrolled.
// On return X[k] is found at index brev5(k).
MIXFE_HD void dft32_dif(float (&re)[32], float (&im)[32]) {
#pragma unroll
    for (int h = 16; h >= 1; h >>= 1) {
#pragma unroll
        for (int blk = 0; blk < 32; blk += 2 * h) {
#pragma unroll
            for (int j = 0; j < h; ++j) {
                const int i0 = blk + j, i1 = i0 + h;
                const int t = j * (16 / h);                 // twiddle W_32^t, t in [0,16)
                const float ar = re[i0], ai = im[i0], br = re[i1], bi = im[i1];
                re[i0] = ar + br;
                im[i0] = ai + bi;
                const float dr = ar - br, di = ai - bi;
                if (t == 0) {
                    re[i1] = dr;
                    im[i1] = di;
                } else if (t == 8) {                        // times -i
                    re[i1] = di;
                    im[i1] = -dr;
                } else {
                    const float c = w32_re(t), s = w32_sin(t);
                    re[i1] = dr * c + di * s;               // (dr + i di)(c - i s)
                    im[i1] = di * c - dr * s;
                }
            }
        }
    }
}

#ifdef __CUDACC__
// Packed variant for sm_100: a complex value lives in one 64-bit register pair (re, im) and the butterfly
// adds/subtracts are single FADD2 instructions (PTX add/sub.rn.f32x2, new with sm_100); twiddle products
// stay scalar on the two halves.  ptxas coalesces the pack/unpack moves away: 160 FADD2 + 122 FMUL/FFMA
// per 32-point DFT instead of 320 FADD + 122.
typedef unsigned long long cplx64;
__device__ __forceinline__ cplx64 cpack(float x, float y) {
    cplx64 d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(x), "f"(y));
    return d;
}
__device__ __forceinline__ void cunpack(cplx64 d, float& x, float& y) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(d));
}
__device__ __forceinline__ cplx64 cadd(cplx64 a, cplx64 b) {
    cplx64 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ cplx64 csub(cplx64 a, cplx64 b) {
    cplx64 d;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

__device__ __forceinline__ cplx64 cfma(cplx64 a, cplx64 b, cplx64 c) {      // lane-wise a * b + c
    cplx64 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

// In-place forward DFT of 32 packed complex values; X[k] is found at index brev5(k) (as dft32_dif).
__device__ __forceinline__ void dft32_dif_packed(cplx64 (&z)[32]) {
#pragma unroll
    for (int h = 16; h >= 1; h >>= 1) {
#pragma unroll
        for (int blk = 0; blk < 32; blk += 2 * h) {
#pragma unroll
            for (int j = 0; j < h; ++j) {
                const int i0 = blk + j, i1 = i0 + h;
                const int t = j * (16 / h);
                const cplx64 a = z[i0], b = z[i1];
                z[i0] = cadd(a, b);
                const cplx64 d = csub(a, b);
                if (t == 0) {
                    z[i1] = d;
                } else {
                    float dr, di;
                    cunpack(d, dr, di);
                    if (t == 8) {                           // times -i
                        z[i1] = cpack(di, -dr);
                    } else {
                        const float c = w32_re(t), s = w32_sin(t);
                        z[i1] = cpack(dr * c + di * s, di * c - dr * s);
                    }
                }
            }
        }
    }
}
#endif  // __CUDACC__

// Same for 16 values (used by the 512-point = 32 x 16 column transforms of the long FFT).
// On return X[k] is found at index brev4(k).
MIXFE_HD constexpr int brev4(int i) {
    return ((i & 1) << 3) | ((i & 2) << 1) | ((i & 4) >> 1) | ((i & 8) >> 3);
}

MIXFE_HD void dft16_dif(float (&re)[16], float (&im)[16]) {
#pragma unroll
    for (int h = 8; h >= 1; h >>= 1) {
#pragma unroll
        for (int blk = 0; blk < 16; blk += 2 * h) {
#pragma unroll
            for (int j = 0; j < h; ++j) {
                const int i0 = blk + j, i1 = i0 + h;
                const int t = j * (8 / h) * 2;              // in units of 2*pi/32
                const float ar = re[i0], ai = im[i0], br = re[i1], bi = im[i1];
                re[i0] = ar + br;
                im[i0] = ai + bi;
                const float dr = ar - br, di = ai - bi;
                if (t == 0) {
                    re[i1] = dr;
                    im[i1] = di;
                } else if (t == 8) {
                    re[i1] = di;
                    im[i1] = -dr;
                } else {
                    const float c = w32_re(t), s = w32_sin(t);
                    re[i1] = dr * c + di * s;
                    im[i1] = di * c - dr * s;
                }
            }
        }
    }
}

}  // namespace mixfe
