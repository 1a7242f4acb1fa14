// loss.cu -- fused photometric loss of the training loops: L1 + 11x11 Gaussian-window SSIM, forward and backward
// (the step immediately after the rasterizer boundary).
//
// Replaces, for loss = (1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image, gt))
// (reference train_3dgs.py:106-107, train_2dgs.py; gs3dgs/utils/loss_utils.py:17-18 l1_loss, :24-26 gaussian,
// :39-43 create_window, :45-54 ssim, :56-73 _ssim), the five grouped F.conv2d calls plus ~20 elementwise torch
// kernels of the forward and their autograd mirrors with two kernels:
//   l1_ssim_forward_kernel   reads image and gt once, writes the three per-pixel partial derivatives of the SSIM map
//                            (w.r.t. mu1, E[x^2], E[xy]) and per-CTA partial sums of |x - y| and of the SSIM map
//   l1_ssim_backward_kernel  convolves the three derivative maps with the same window and adds the L1 sign term
// Both are separable (11 + 11 taps instead of 121) over a 32x32 shared-memory tile with a 5-pixel zero-padded halo;
// a thread produces four neighbouring outputs per pass from a 14-element sliding window held in registers. The halo is exactly
// the zero padding of F.conv2d(padding = 5). The reduction is deterministic: per-CTA partials, then one CTA sums them
// in double in a fixed order.
#include "common.cuh"
#include <cmath>

namespace gsr {

constexpr int LW = 11, LR = 5;                 // window size, radius (loss_utils.py:45 window_size = 11)
constexpr int LTX = 32, LTY = 32;              // output tile per CTA
constexpr int LV = 4;                          // outputs per thread: 4 along x in the horizontal pass, 4 along y in the vertical
constexpr int LTHREADS = LTX * LTY / LV;       // 256
constexpr int LHX = LTX + 2 * LR, LHY = LTY + 2 * LR;
constexpr int LSPAN = LV + LW - 1;             // inputs a thread reads for its LV outputs (sliding window)
constexpr float SSIM_C1 = 0.01f * 0.01f, SSIM_C2 = 0.03f * 0.03f;      // loss_utils.py:67-68

struct LossWindow { float g[LW]; };

// gaussian(11, 1.5) (loss_utils.py:24-26): exp() in double per tap, stored as float32, normalised in float32
static LossWindow make_window() {
    LossWindow w;
    float s = 0.0f;
    for (int i = 0; i < LW; ++i) {
        const int x = i - LW / 2;
        w.g[i] = (float)std::exp(-(double)(x * x) / (2.0 * 1.5 * 1.5));
        s += w.g[i];
    }
    for (int i = 0; i < LW; ++i) w.g[i] /= s;
    return w;
}

__device__ __forceinline__ float block_sum(float v, float* s_red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();
    if (lane == 0) s_red[warp] = v;
    __syncthreads();
    float t = 0.0f;
    if (warp == 0) {
        t = lane < (LTHREADS / 32) ? s_red[lane] : 0.0f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    }
    return t;       // valid in thread 0
}

// LV sliding-window sums of one quantity: out[j] = sum_k g[k] * v[j + k], taps in the order k = 0..10
__device__ __forceinline__ void window4(const LossWindow& win, const float (&v)[LSPAN], float (&out)[LV]) {
#pragma unroll
    for (int j = 0; j < LV; ++j) out[j] = 0.0f;
#pragma unroll
    for (int k = 0; k < LW; ++k)
#pragma unroll
        for (int j = 0; j < LV; ++j) out[j] = fmaf(win.g[k], v[j + k], out[j]);
}

// Shared-memory layout: halo rows have an odd stride (LHX + 1 = 43) and the horizontally filtered rows a stride of
// LTX + 1 = 33, which makes every access pattern below bank-conflict free: the horizontal pass maps a warp to
// 4 rows x 8 groups of 4 columns, the vertical pass to 32 consecutive columns.
__global__ void __launch_bounds__(LTHREADS)
l1_ssim_forward_kernel(int H, int W, const float* __restrict__ img, const float* __restrict__ gt, LossWindow win,
                       float* __restrict__ dm_dmu1, float* __restrict__ dm_dsig1, float* __restrict__ dm_dsig12,
                       float* __restrict__ partial) {
    __shared__ float s_x[LHY][LHX + 1], s_y[LHY][LHX + 1];
    __shared__ float s_h[5][LHY][LTX + 1];
    __shared__ float s_red[LTHREADS / 32];
    const int c = blockIdx.z;
    const int x0 = blockIdx.x * LTX, y0 = blockIdx.y * LTY;
    const size_t plane = (size_t)c * H * W;
    const int tid = threadIdx.x;
    // halo load: a warp takes whole rows (lanes = columns 0..31, then 32..41), so there is no index division and a
    // row is one or two contiguous requests
#pragma unroll
    for (int r = tid >> 5; r < LHY; r += LTHREADS / 32) {      // at most 6 rows per warp: all loads in flight at once
        const int gy = y0 + r - LR, lane = tid & 31;
        const bool rowin = gy >= 0 && gy < H;
        const size_t rowoff = plane + (size_t)(rowin ? gy : 0) * W;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int q = lane + 32 * h, gx = x0 + q - LR;
            if (q < LHX) {
                const bool in = rowin && gx >= 0 && gx < W;
                s_x[r][q] = in ? __ldg(img + rowoff + gx) : 0.0f;
                s_y[r][q] = in ? __ldg(gt + rowoff + gx) : 0.0f;
            }
        }
    }
    __syncthreads();
    for (int it = tid; it < LHY * (LTX / LV); it += LTHREADS) {     // horizontal pass over all halo rows
        const int r = it / (LTX / LV), c0 = (it - r * (LTX / LV)) * LV;
        float xv[LSPAN], yv[LSPAN], t[LSPAN], o[LV];
#pragma unroll
        for (int i = 0; i < LSPAN; ++i) { xv[i] = s_x[r][c0 + i]; yv[i] = s_y[r][c0 + i]; }
        window4(win, xv, o);
#pragma unroll
        for (int j = 0; j < LV; ++j) s_h[0][r][c0 + j] = o[j];
        window4(win, yv, o);
#pragma unroll
        for (int j = 0; j < LV; ++j) s_h[1][r][c0 + j] = o[j];
#pragma unroll
        for (int i = 0; i < LSPAN; ++i) t[i] = xv[i] * xv[i];
        window4(win, t, o);
#pragma unroll
        for (int j = 0; j < LV; ++j) s_h[2][r][c0 + j] = o[j];
#pragma unroll
        for (int i = 0; i < LSPAN; ++i) t[i] = yv[i] * yv[i];
        window4(win, t, o);
#pragma unroll
        for (int j = 0; j < LV; ++j) s_h[3][r][c0 + j] = o[j];
#pragma unroll
        for (int i = 0; i < LSPAN; ++i) t[i] = xv[i] * yv[i];
        window4(win, t, o);
#pragma unroll
        for (int j = 0; j < LV; ++j) s_h[4][r][c0 + j] = o[j];
    }
    __syncthreads();
    const int tx = tid % LTX, ty0 = (tid / LTX) * LV;
    const int px = x0 + tx;
    float f[5][LV];
#pragma unroll
    for (int q = 0; q < 5; ++q) {
        float col[LSPAN];
#pragma unroll
        for (int i = 0; i < LSPAN; ++i) col[i] = s_h[q][ty0 + i][tx];
        window4(win, col, f[q]);
    }
    float l1 = 0.0f, sm_sum = 0.0f;
#pragma unroll
    for (int j = 0; j < LV; ++j) {
        const int py = y0 + ty0 + j;
        if (px < W && py < H) {
            const float mu1 = f[0][j], mu2 = f[1][j], e11 = f[2][j], e22 = f[3][j], e12 = f[4][j];
            // loss_utils.py:59-70
            const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu1_mu2 = mu1 * mu2;
            const float sigma1_sq = e11 - mu1_sq, sigma2_sq = e22 - mu2_sq, sigma12 = e12 - mu1_mu2;
            const float A = 2.0f * mu1_mu2 + SSIM_C1, B = 2.0f * sigma12 + SSIM_C2;
            const float Cc = mu1_sq + mu2_sq + SSIM_C1, D = sigma1_sq + sigma2_sq + SSIM_C2;
            const float inv_cd = 1.0f / (Cc * D);
            const float sm = A * B * inv_cd;
            // partial derivatives of the map w.r.t. the three windowed sums that depend on img1:
            //   mu1 (also inside sigma1_sq = E11 - mu1^2 and sigma12 = E12 - mu1 mu2), E11, E12
            const size_t o = plane + (size_t)py * W + px;
            dm_dmu1[o] = 2.0f * inv_cd * (mu2 * (B - A) - mu1 * sm * (D - Cc));
            dm_dsig1[o] = -sm / D;
            dm_dsig12[o] = 2.0f * A * inv_cd;
            sm_sum += sm;
            l1 += fabsf(s_x[ty0 + j + LR][tx + LR] - s_y[ty0 + j + LR][tx + LR]);
        }
    }
    const float t1 = block_sum(l1, s_red);
    const float t2 = block_sum(sm_sum, s_red);
    if (tid == 0) {
        const size_t b = ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        partial[2 * b] = t1; partial[2 * b + 1] = t2;
    }
}

// one CTA: sums the per-CTA partials in double (fixed order) and writes out[0] = loss, out[1] = l1, out[2] = ssim
__global__ void __launch_bounds__(256)
l1_ssim_finalize_kernel(int nblocks, double inv_n, float lambda, const float* __restrict__ partial, float* __restrict__ out) {
    __shared__ double s1[256], s2[256];
    double a = 0.0, b = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += 256) { a += (double)partial[2 * i]; b += (double)partial[2 * i + 1]; }
    s1[threadIdx.x] = a; s2[threadIdx.x] = b;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) { s1[threadIdx.x] += s1[threadIdx.x + o]; s2[threadIdx.x] += s2[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const float l1 = (float)(s1[0] * inv_n), ss = (float)(s2[0] * inv_n);
        out[0] = (1.0f - lambda) * l1 + lambda * (1.0f - ss);      // train_3dgs.py:107
        out[1] = l1; out[2] = ss;
    }
}

__global__ void __launch_bounds__(LTHREADS)
l1_ssim_backward_kernel(int H, int W, const float* __restrict__ img, const float* __restrict__ gt, LossWindow win,
                        const float* __restrict__ dm_dmu1, const float* __restrict__ dm_dsig1,
                        const float* __restrict__ dm_dsig12, const float* __restrict__ upstream,
                        float w_l1, float w_ssim, float inv_n, float* __restrict__ dL_dimg) {
    __shared__ float s_m[3][LHY][LHX + 1];
    __shared__ float s_h[3][LHY][LTX + 1];
    const int c = blockIdx.z;
    const int x0 = blockIdx.x * LTX, y0 = blockIdx.y * LTY;
    const size_t plane = (size_t)c * H * W;
    const int tid = threadIdx.x;
    // (flat element loop here: measured faster than the row-wise load of the forward for three maps, 0.041 vs 0.047 ms)
    for (int e = tid; e < LHY * LHX; e += LTHREADS) {
        const int r = e / LHX, q = e - r * LHX;
        const int gy = y0 + r - LR, gx = x0 + q - LR;
        const bool in = gy >= 0 && gy < H && gx >= 0 && gx < W;
        const size_t o = plane + (size_t)gy * W + gx;
        s_m[0][r][q] = in ? __ldg(dm_dmu1 + o) : 0.0f;
        s_m[1][r][q] = in ? __ldg(dm_dsig1 + o) : 0.0f;
        s_m[2][r][q] = in ? __ldg(dm_dsig12 + o) : 0.0f;
    }
    __syncthreads();
    for (int it = tid; it < LHY * (LTX / LV); it += LTHREADS) {
        const int r = it / (LTX / LV), c0 = (it - r * (LTX / LV)) * LV;
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            float v[LSPAN], o[LV];
#pragma unroll
            for (int i = 0; i < LSPAN; ++i) v[i] = s_m[q][r][c0 + i];
            window4(win, v, o);
#pragma unroll
            for (int j = 0; j < LV; ++j) s_h[q][r][c0 + j] = o[j];
        }
    }
    __syncthreads();
    const int tx = tid % LTX, ty0 = (tid / LTX) * LV;
    const int px = x0 + tx;
    float f[3][LV];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        float col[LSPAN];
#pragma unroll
        for (int i = 0; i < LSPAN; ++i) col[i] = s_h[q][ty0 + i][tx];
        window4(win, col, f[q]);
    }
    const float up = upstream ? __ldg(upstream) : 1.0f;
#pragma unroll
    for (int j = 0; j < LV; ++j) {
        const int py = y0 + ty0 + j;
        if (px < W && py < H) {
            const size_t o = plane + (size_t)py * W + px;
            const float x = __ldg(img + o), y = __ldg(gt + o);
            const float dssim = f[0][j] + 2.0f * x * f[1][j] + y * f[2][j];       // d (sum of the SSIM map) / d img[o]
            const float diff = x - y;
            const float dl1 = (diff > 0.0f) ? 1.0f : ((diff < 0.0f) ? -1.0f : 0.0f);    // torch.abs backward: sign, 0 at 0
            dL_dimg[o] = up * inv_n * (w_l1 * dl1 + w_ssim * dssim);
        }
    }
}

static size_t loss_blocks(int C, int H, int W) {
    return (size_t)((W + LTX - 1) / LTX) * ((H + LTY - 1) / LTY) * C;
}

}  // namespace gsr

using namespace gsr;

extern "C" {

size_t gsr_l1_ssim_workspace_bytes(int C, int H, int W) {
    if (C <= 0 || H <= 0 || W <= 0) return 256;
    const size_t n = (size_t)C * H * W;
    return 3 * align_up(n * sizeof(float)) + align_up(loss_blocks(C, H, W) * 2 * sizeof(float));
}

int gsr_l1_ssim_forward(int C, int H, int W, const float* img, const float* gt, float lambda_dssim,
                        float* out3, void* workspace, size_t workspace_bytes, void* stream)
{
    cudaStream_t st = (cudaStream_t)stream;
    if (C <= 0 || H <= 0 || W <= 0 || C > 65535) { set_error("bad image shape C=%d H=%d W=%d", C, H, W); return GSR_EINVAL; }
    if (!img || !gt || !out3 || !workspace) { set_error("null pointer in gsr_l1_ssim_forward"); return GSR_EINVAL; }
    if (workspace_bytes < gsr_l1_ssim_workspace_bytes(C, H, W)) { set_error("loss workspace too small"); return GSR_EINVAL; }
    static const LossWindow win = make_window();
    const size_t n = (size_t)C * H * W, stride = align_up(n * sizeof(float));
    char* base = (char*)workspace;
    float* m0 = (float*)base; float* m1 = (float*)(base + stride); float* m2 = (float*)(base + 2 * stride);
    float* partial = (float*)(base + 3 * stride);
    const dim3 grid((W + LTX - 1) / LTX, (H + LTY - 1) / LTY, C);
    if (grid.y > 65535) { set_error("image too tall"); return GSR_EUNSUPPORTED; }
    l1_ssim_forward_kernel<<<grid, LTHREADS, 0, st>>>(H, W, img, gt, win, m0, m1, m2, partial);
    int rc;
    if ((rc = check_cuda(cudaGetLastError(), "l1_ssim forward launch"))) return rc;
    l1_ssim_finalize_kernel<<<1, 256, 0, st>>>((int)loss_blocks(C, H, W), 1.0 / (double)n, lambda_dssim, partial, out3);
    return check_cuda(cudaGetLastError(), "l1_ssim finalize launch");
}

int gsr_l1_ssim_backward(int C, int H, int W, const float* img, const float* gt, const float* upstream,
                         float w_l1, float w_ssim, const void* workspace, float* dL_dimg, void* stream)
{
    cudaStream_t st = (cudaStream_t)stream;
    if (C <= 0 || H <= 0 || W <= 0 || C > 65535) { set_error("bad image shape C=%d H=%d W=%d", C, H, W); return GSR_EINVAL; }
    if (!img || !gt || !workspace || !dL_dimg) { set_error("null pointer in gsr_l1_ssim_backward"); return GSR_EINVAL; }
    static const LossWindow win = make_window();
    const size_t n = (size_t)C * H * W, stride = align_up(n * sizeof(float));
    const char* base = (const char*)workspace;
    const dim3 grid((W + LTX - 1) / LTX, (H + LTY - 1) / LTY, C);
    if (grid.y > 65535) { set_error("image too tall"); return GSR_EUNSUPPORTED; }
    l1_ssim_backward_kernel<<<grid, LTHREADS, 0, st>>>(H, W, img, gt, win, (const float*)base, (const float*)(base + stride),
                                                        (const float*)(base + 2 * stride), upstream, w_l1, w_ssim,
                                                        (float)(1.0 / (double)n), dL_dimg);
    return check_cuda(cudaGetLastError(), "l1_ssim backward launch");
}

}  // extern "C"
