/* loss_oracle.c -- TEST INFRASTRUCTURE (checker): plain-C CPU restatement of the reference's photometric loss.
 * Never linked into, imported by or called from scorp_b200/.
 *
 * Follows /root/reference/gs3dgs/utils/loss_utils.py:
 *   :17-18  l1_loss   = mean |a - b|
 *   :24-26  gaussian  = exp(-(x - 5)^2 / (2 * 1.5^2)) stored as float32, normalised by its float32 sum
 *   :39-43  create_window = outer product of the 1-D window (float32), one copy per channel (groups = channel)
 *   :56-73  _ssim     = five F.conv2d(padding = 5, groups = C) -> mu1, mu2, E11, E22, E12; the SSIM map; its mean
 * and train_3dgs.py:106-107: loss = (1 - lambda) * l1 + lambda * (1 - ssim).
 * The sums are accumulated in double from the float32 inputs and the float32 2-D window (a 121-tap direct
 * convolution, as the reference's conv2d is -- not the separable form the product kernels use); the gradients are
 * the analytic adjoint (what autograd computes for the reference), also in double.
 * Parity status: PINNED by tests/golden/l*.npz = outputs and autograd gradients of the reference module itself,
 * imported from /root/reference and run on CPU fp32 by tools/make_golden_loss.py.
 */
#include "oracle_common.h"
#include <stdlib.h>

#define WS 11
#define WR 5

static void window2d(float w2[WS][WS]) {
    float g[WS], s = 0.0f;
    for (int i = 0; i < WS; ++i) { int x = i - WS / 2; g[i] = (float)exp(-(double)(x * x) / (2.0 * 1.5 * 1.5)); s += g[i]; }
    for (int i = 0; i < WS; ++i) g[i] = g[i] / s;
    for (int i = 0; i < WS; ++i) for (int j = 0; j < WS; ++j) w2[i][j] = g[i] * g[j];     /* _1D.mm(_1D.t()) in float32 */
}

/* out[0] = loss, out[1] = l1, out[2] = ssim; dl1 / dssim (may be NULL): gradients of l1 and of ssim w.r.t. img */
void loss_oracle_l1_ssim(int C, int H, int W, const float* img, const float* gt, float lambda, double* out,
                         float* dl1, float* dssim) {
    float w2[WS][WS];
    window2d(w2);
    const size_t n = (size_t)C * H * W;
    const double C1 = 0.01 * 0.01, C2 = 0.03 * 0.03;
    double* dmu1 = (double*)malloc(n * sizeof(double));
    double* ds1 = (double*)malloc(n * sizeof(double));
    double* ds12 = (double*)malloc(n * sizeof(double));
    double sum_l1 = 0.0, sum_ss = 0.0;
    for (int c = 0; c < C; ++c) {
        const float* a = img + (size_t)c * H * W;
        const float* b = gt + (size_t)c * H * W;
#pragma omp parallel for reduction(+ : sum_l1, sum_ss) schedule(static)
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x) {
                double mu1 = 0, mu2 = 0, e11 = 0, e22 = 0, e12 = 0;
                for (int i = 0; i < WS; ++i) {
                    const int yy = y + i - WR;
                    if (yy < 0 || yy >= H) continue;               /* zero padding */
                    for (int j = 0; j < WS; ++j) {
                        const int xx = x + j - WR;
                        if (xx < 0 || xx >= W) continue;
                        const double w = w2[i][j], p = a[(size_t)yy * W + xx], q = b[(size_t)yy * W + xx];
                        mu1 += w * p; mu2 += w * q; e11 += w * p * p; e22 += w * q * q; e12 += w * p * q;
                    }
                }
                const double s1 = e11 - mu1 * mu1, s2 = e22 - mu2 * mu2, s12 = e12 - mu1 * mu2;
                const double A = 2 * mu1 * mu2 + C1, B = 2 * s12 + C2, Cc = mu1 * mu1 + mu2 * mu2 + C1, D = s1 + s2 + C2;
                const double m = (A * B) / (Cc * D);
                const size_t o = (size_t)c * H * W + (size_t)y * W + x;
                /* d map / d mu1 (through A, B, Cc, D), d map / d E11, d map / d E12 */
                dmu1[o] = (2 * mu2 * B - 2 * mu2 * A) / (Cc * D) - (A * B) * (2 * mu1) / (Cc * Cc * D)
                        + (A * B) * (2 * mu1) / (Cc * D * D);
                ds1[o] = -(A * B) / (Cc * D * D);
                ds12[o] = 2 * A / (Cc * D);
                sum_ss += m;
                sum_l1 += fabs((double)a[(size_t)y * W + x] - (double)b[(size_t)y * W + x]);
            }
    }
    out[1] = sum_l1 / (double)n; out[2] = sum_ss / (double)n;
    out[0] = (1.0 - (double)lambda) * out[1] + (double)lambda * (1.0 - out[2]);
    if (dl1)
        for (size_t i = 0; i < n; ++i) {
            const double d = (double)img[i] - (double)gt[i];
            dl1[i] = (float)((d > 0 ? 1.0 : (d < 0 ? -1.0 : 0.0)) / (double)n);
        }
    if (dssim)
        for (int c = 0; c < C; ++c) {
#pragma omp parallel for schedule(static)
            for (int y = 0; y < H; ++y)
                for (int x = 0; x < W; ++x) {
                    const size_t o = (size_t)c * H * W + (size_t)y * W + x;
                    double acc = 0;
                    for (int i = 0; i < WS; ++i) {
                        const int yy = y - (i - WR);                 /* adjoint: output pixels whose window covers (y, x) */
                        if (yy < 0 || yy >= H) continue;
                        for (int j = 0; j < WS; ++j) {
                            const int xx = x - (j - WR);
                            if (xx < 0 || xx >= W) continue;
                            const size_t q = (size_t)c * H * W + (size_t)yy * W + xx;
                            acc += (double)w2[i][j] * (dmu1[q] + 2.0 * (double)img[o] * ds1[q] + (double)gt[o] * ds12[q]);
                        }
                    }
                    dssim[o] = (float)(acc / (double)n);
                }
        }
    free(dmu1); free(ds1); free(ds12);
}
