/* CPU oracle helpers (TEST INFRASTRUCTURE ONLY -- never linked into the product).
 *
 * oracle_sosfilt_df2t : what scipy.signal.sosfilt computes for the reference at
 *     src/mixing_utils.py:432,455 -- a cascade of biquads in transposed direct form II,
 *     double precision, zero initial state, float32 input promoted to double.
 * oracle_xcorr_same   : what F.conv1d(x, h, padding=L/2)[..., :T] computes at
 *     src/mixing_utils.py:467-473,479 -- a cross-correlation (conv1d does not flip the kernel),
 *     accumulated here in double so it can serve as ground truth for both implementations.
 */
#include <stddef.h>

void oracle_sosfilt_df2t(const double *sos, int n_sections, const double *x, double *y, long n)
{
    for (long i = 0; i < n; ++i) y[i] = x[i];
    for (int s = 0; s < n_sections; ++s) {
        const double b0 = sos[6 * s + 0], b1 = sos[6 * s + 1], b2 = sos[6 * s + 2];
        const double a0 = sos[6 * s + 3], a1 = sos[6 * s + 4], a2 = sos[6 * s + 5];
        const double nb0 = b0 / a0, nb1 = b1 / a0, nb2 = b2 / a0, na1 = a1 / a0, na2 = a2 / a0;
        double z1 = 0.0, z2 = 0.0;
        for (long i = 0; i < n; ++i) {
            const double in = y[i];
            const double out = nb0 * in + z1;
            z1 = nb1 * in - na1 * out + z2;
            z2 = nb2 * in - na2 * out;
            y[i] = out;
        }
    }
}

void oracle_xcorr_same(const float *x, long T, const float *h, long L, float *out)
{
    const long half = L / 2;
    for (long n = 0; n < T; ++n) {
        long k0 = half - n;           /* first k with n + k - half >= 0 */
        if (k0 < 0) k0 = 0;
        long k1 = T - 1 - n + half;   /* last k with n + k - half <= T-1 */
        if (k1 > L - 1) k1 = L - 1;
        double acc = 0.0;
        const float *xs = x + (n - half);
        for (long k = k0; k <= k1; ++k) acc += (double)xs[k] * (double)h[k];
        out[n] = (float)acc;
    }
}
