// Host check of the per-lane FFT arithmetic and of the warp-level index maps used by the CUDA
// kernels (no GPU needed).  Emulates the 32 lanes of a warp with plain loops:
//   pass 1 (lane = n2): v[n1] = x[32*n1 + n2] -> dft32 -> Y[k1]; times W_1024^(n2*k1)
//   exchange through a swizzled buffer  idx = k1*32 + (n2 ^ k1)
//   pass 2 (lane = k1): v[n2] -> dft32 -> X[k1 + 32*k2]
//   untangle of two real signals packed as L + iR, via the partner map the kernel uses.
// Prints "OK" and exits 0 when every check is within tolerance of a double-precision DFT.
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../mixing_style_transfer_b200/csrc/fft32.cuh"

using namespace mixfe;
typedef std::complex<double> cd;

static std::vector<cd> naive_dft(const std::vector<cd>& x) {
    const int n = (int)x.size();
    std::vector<cd> X(n);
    for (int k = 0; k < n; ++k) {
        cd acc = 0;
        for (int j = 0; j < n; ++j) acc += x[j] * std::polar(1.0, -2.0 * M_PI * (double)((long)j * k % n) / n);
        X[k] = acc;
    }
    return X;
}

int main() {
    int bad = 0;
    srand(7);
    // --- 32- and 16-point kernels -----------------------------------------------------------
    {
        float re[32], im[32];
        std::vector<cd> x(32);
        for (int i = 0; i < 32; ++i) { re[i] = rand() / (float)RAND_MAX - 0.5f; im[i] = rand() / (float)RAND_MAX - 0.5f; x[i] = cd(re[i], im[i]); }
        auto X = naive_dft(x);
        dft32_dif(re, im);
        double e = 0;
        for (int k = 0; k < 32; ++k) e = std::max(e, std::abs(cd(re[brev5(k)], im[brev5(k)]) - X[k]));
        printf("dft32 max err %.3g\n", e);
        if (e > 5e-6) ++bad;
    }
    {
        float re[16], im[16];
        std::vector<cd> x(16);
        for (int i = 0; i < 16; ++i) { re[i] = rand() / (float)RAND_MAX - 0.5f; im[i] = rand() / (float)RAND_MAX - 0.5f; x[i] = cd(re[i], im[i]); }
        auto X = naive_dft(x);
        dft16_dif(re, im);
        double e = 0;
        for (int k = 0; k < 16; ++k) e = std::max(e, std::abs(cd(re[brev4(k)], im[brev4(k)]) - X[k]));
        printf("dft16 max err %.3g\n", e);
        if (e > 5e-6) ++bad;
    }
    // --- warp-level 1024-point transform with L/R untangle ------------------------------------
    {
        const int N = 1024;
        std::vector<float> L(N), R(N);
        std::vector<cd> xl(N), xr(N);
        for (int i = 0; i < N; ++i) {
            L[i] = rand() / (float)RAND_MAX - 0.5f; R[i] = rand() / (float)RAND_MAX - 0.5f;
            xl[i] = L[i]; xr[i] = R[i];
        }
        auto XL = naive_dft(xl), XR = naive_dft(xr);
        std::vector<float> exr(N), exi(N);
        // pass 1
        for (int lane = 0; lane < 32; ++lane) {
            float re[32], im[32];
            for (int n1 = 0; n1 < 32; ++n1) { re[n1] = 0.5f * L[32 * n1 + lane]; im[n1] = 0.5f * R[32 * n1 + lane]; }
            dft32_dif(re, im);
            for (int k1 = 0; k1 < 32; ++k1) {
                const double a = -2.0 * M_PI * (double)(lane * k1) / N;
                const float tr = (float)cos(a), ti = (float)sin(a);
                const float yr = re[brev5(k1)], yi = im[brev5(k1)];
                const int idx = k1 * 32 + (lane ^ k1);
                exr[idx] = yr * tr - yi * ti;
                exi[idx] = yr * ti + yi * tr;
            }
        }
        // pass 2: lane = k1; keep registers of every lane for the shuffle emulation
        std::vector<float> Xre(32 * 32), Xim(32 * 32);      // [lane][k2]
        for (int lane = 0; lane < 32; ++lane) {
            float re[32], im[32];
            for (int n2 = 0; n2 < 32; ++n2) { re[n2] = exr[lane * 32 + (n2 ^ lane)]; im[n2] = exi[lane * 32 + (n2 ^ lane)]; }
            dft32_dif(re, im);
            for (int k2 = 0; k2 < 32; ++k2) { Xre[lane * 32 + k2] = re[brev5(k2)]; Xim[lane * 32 + k2] = im[brev5(k2)]; }
        }
        // untangle, partner map: lane l != 0 takes register (31-k2) of lane (32-l)&31; lane 0 its own (32-k2)&31
        double e = 0, pmax = 0;
        for (int lane = 0; lane < 32; ++lane) {
            for (int k2 = 0; k2 <= 16; ++k2) {
                if (k2 == 16 && lane != 0) continue;
                const int k = lane + 32 * k2;
                float pr, pi;
                if (lane == 0) { pr = Xre[((32 - k2) & 31)]; pi = Xim[((32 - k2) & 31)]; }
                else { const int src = (32 - lane) & 31; pr = Xre[src * 32 + (31 - k2)]; pi = Xim[src * 32 + (31 - k2)]; }
                const float xr_ = Xre[lane * 32 + k2], xi_ = Xim[lane * 32 + k2];
                const float a = xr_ + pr, b = xi_ - pi, c = xr_ - pr, d = xi_ + pi;
                const float PL = a * a + b * b, PR = c * c + d * d;      // |L_k|^2 , |R_k|^2 (0.5 folded into the input)
                e = std::max(e, std::fabs((double)PL - std::norm(XL[k])));
                e = std::max(e, std::fabs((double)PR - std::norm(XR[k])));
                pmax = std::max(pmax, std::norm(XL[k]));
            }
        }
        printf("warp fft1024 + untangle: max |dP| %.3g (max P %.3g)\n", e, pmax);
        if (e > 1e-4 * pmax) ++bad;
    }
    if (bad) { printf("FAILED %d\n", bad); return 1; }
    printf("OK\n");
    return 0;
}
