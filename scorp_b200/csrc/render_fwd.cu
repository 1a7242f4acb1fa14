// render_fwd.cu -- per-tile front-to-back alpha blend (3DGS forward).
//
// Replaces renderCUDA (reference diff-gaussian-rasterization/cuda_rasterizer/forward.cu:263-381).
// One CTA per 16x16 tile; each of the 8 warps owns an 8x4 pixel block. Per round of 256 list
// entries: (1) 48-byte records are gathered into shared memory with cp.async, double-buffered
// against the blend of the previous round; (2) every thread tests ITS record against the 8 warp
// blocks (exact minimum of the conic form over the block rectangle vs. the record's power
// threshold) and the warps exchange 8 ballots; (3) each warp blends only the entries whose bit
// is set. Skipped (pixel, Gaussian) pairs are exactly pairs the reference also skips
// (power > 0 or alpha < 1/255), so the image, alpha, depth and n_contrib are bit-identical.
#include "common.cuh"

namespace gsr {

// Lower bound test: can record (x,y,a,b,c,thr) reach power >= thr anywhere in the pixel
// rectangle [px0,px1]x[py0,py1]?  q(d) = 0.5(a dx^2 + c dy^2) + b dx dy, power = -q.
__device__ __forceinline__ bool block_may_contribute(float x, float y, float a, float b, float c, float thr,
                                                     float inv_a_b, float inv_c_b, bool convex,
                                                     float px0, float px1, float py0, float py1) {
    if (!convex) return true;
    // d = mean - pixel; pixel in [p0,p1]  =>  d in [m - p1, m - p0]
    const float lx = x - px1, hx = x - px0, ly = y - py1, hy = y - py0;
    const float cx = fminf(fmaxf(0.0f, lx), hx), cy = fminf(fmaxf(0.0f, ly), hy);
    if (cx == 0.0f && cy == 0.0f) return true;
    float qmin = 3.0e38f;
    if (cx != 0.0f) {   // facing vertical edge dx = cx
        const float dy = fminf(fmaxf(-inv_c_b * cx, ly), hy);
        qmin = 0.5f * (a * cx * cx + c * dy * dy) + b * cx * dy;
    }
    if (cy != 0.0f) {   // facing horizontal edge dy = cy
        const float dx = fminf(fmaxf(-inv_a_b * cy, lx), hx);
        qmin = fminf(qmin, 0.5f * (a * dx * dx + c * cy * cy) + b * dx * cy);
    }
    // keep unless provably below threshold (slack covers rounding of this test; NaN keeps)
    return !(-qmin < thr - 1e-3f - 1e-5f * fabsf(qmin));
}

__global__ void __launch_bounds__(256)
render_forward_kernel(int W, int H, int gx,
                      const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order, const uint32_t* __restrict__ point_list,
                      const GaussRec* __restrict__ recs, const float* __restrict__ bg,
                      uint32_t* __restrict__ n_contrib, float* __restrict__ out_color,
                      float* __restrict__ out_depth, float* __restrict__ out_alpha) {
    __shared__ __align__(16) GaussRec s_rec[2][BATCH];
    __shared__ uint32_t s_bal[8][8];          // [consumer warp block][producer warp]

    const int tile = (int)tile_order[blockIdx.x];      // longest lists first (binning.cu)
    const int tile_x = (tile % gx) * TILE, tile_y = (tile / gx) * TILE;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int bx = (wid & 1) * 8, by = (wid >> 1) * 4;
    const int pxi = tile_x + bx + (lane & 7), pyi = tile_y + by + (lane >> 3);
    const bool inside = pxi < W && pyi < H;
    const float pxf = (float)pxi, pyf = (float)pyi;
    const uint32_t pix_id = (uint32_t)pyi * W + pxi;

    const uint2 range = ranges[tile];
    const int total = (int)(range.y - range.x);
    const int rounds = (total + BATCH - 1) / BATCH;

    float T = 1.0f, C0 = 0.0f, C1 = 0.0f, C2 = 0.0f, Dacc = 0.0f, Wacc = 0.0f;
    uint32_t last_contributor = 0;
    bool done = !inside;

    auto issue = [&](int r) {
        const int p = r * BATCH + tid;
        if (p < total) {
            const uint32_t id = __ldg(point_list + range.x + p);
            const char* src = reinterpret_cast<const char*>(recs + id);
            char* dst = reinterpret_cast<char*>(&s_rec[r & 1][tid]);
            cp_async16(dst, src); cp_async16(dst + 16, src + 16); cp_async16(dst + 32, src + 32);
        }
        cp_async_commit();
    };

    if (rounds > 0) issue(0);
    for (int r = 0; r < rounds; ++r) {
        // all warps finished? (also orders the previous round's reads before buffer reuse)
        if (__syncthreads_and(done)) break;
        if (r + 1 < rounds) { issue(r + 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
        __syncthreads();

        const int n_here = min(BATCH, total - r * BATCH);
        const GaussRec* buf = s_rec[r & 1];
        // per-record cull masks -> ballots
        {
            uint32_t mask = 0;
            if (tid < n_here) {
                const float4 g0 = *reinterpret_cast<const float4*>(&buf[tid]);
                const float4 g1 = *(reinterpret_cast<const float4*>(&buf[tid]) + 1);
                const float x = g0.x, y = g0.y, a = g0.z, b = g0.w, c = g1.x, thr = g1.w;
                const bool convex = (a > 0.0f) && (c > 0.0f) && (a * c - b * b > 0.0f);
                const float inv_a_b = convex ? b / a : 0.0f, inv_c_b = convex ? b / c : 0.0f;
#pragma unroll
                for (int w = 0; w < 8; ++w) {
                    const float x0 = (float)(tile_x + (w & 1) * 8), y0 = (float)(tile_y + (w >> 1) * 4);
                    if (block_may_contribute(x, y, a, b, c, thr, inv_a_b, inv_c_b, convex, x0, x0 + 7.0f, y0, y0 + 3.0f))
                        mask |= 1u << w;
                }
            }
#pragma unroll
            for (int w = 0; w < 8; ++w) {
                const uint32_t bal = __ballot_sync(0xffffffffu, (mask >> w) & 1u);
                if (lane == 0) s_bal[w][wid] = bal;
            }
        }
        __syncthreads();

        if (!__all_sync(0xffffffffu, done)) {
            for (int v = 0; v < 8; ++v) {
                uint32_t bits = s_bal[wid][v];
                while (bits) {
                    const int j = v * 32 + __ffs(bits) - 1;
                    bits &= bits - 1;
                    if (done) continue;
                    const float4 g0 = *reinterpret_cast<const float4*>(&buf[j]);
                    const float4 g1 = *(reinterpret_cast<const float4*>(&buf[j]) + 1);
                    const float dx = subr(g0.x, pxf), dy = subr(g0.y, pyf);
                    // forward.cu:334 with the reference build's contraction
                    const float power = fmar(fmar(dx, mulr(dx, g0.z), mulr(dy, mulr(dy, g1.x))), -0.5f,
                                             -mulr(dy, mulr(dx, g0.w)));
                    if (power > 0.0f) continue;
                    if (power < g1.w) continue;                  // implies alpha < 1/255
                    const float alpha = fminf(0.99f, mulr(g1.y, expf(power)));
                    if (alpha < 1.0f / 255.0f) continue;
                    const float test_T = mulr(T, subr(1.0f, alpha));
                    if (test_T < 0.0001f) { done = true; continue; }
                    const float4 g2 = *(reinterpret_cast<const float4*>(&buf[j]) + 2);
                    C0 = fmar(T, mulr(alpha, g2.x), C0);
                    C1 = fmar(T, mulr(alpha, g2.y), C1);
                    C2 = fmar(T, mulr(alpha, g2.z), C2);
                    Wacc = fmar(T, alpha, Wacc);
                    Dacc = fmar(T, mulr(alpha, g1.z), Dacc);
                    T = test_T;
                    last_contributor = (uint32_t)(r * BATCH + j + 1);
                }
                if (__all_sync(0xffffffffu, done)) break;
            }
        }
    }
    cp_async_wait<0>();

    if (inside) {
        const size_t hw = (size_t)H * W;
        n_contrib[pix_id] = last_contributor;
        out_color[pix_id] = fmar(T, __ldg(bg), C0);
        out_color[hw + pix_id] = fmar(T, __ldg(bg + 1), C1);
        out_color[2 * hw + pix_id] = fmar(T, __ldg(bg + 2), C2);
        out_alpha[pix_id] = Wacc;
        out_depth[pix_id] = Dacc;
    }
}

cudaError_t launch_render_forward(int W, int H, int gx, int gy, const GeomState& g, const ImageState& img,
                                  const BinningState& bin, const float* bg,
                                  float* out_color, float* out_depth, float* out_alpha, cudaStream_t st) {
    render_forward_kernel<<<gx * gy, 256, 0, st>>>(W, H, gx, img.ranges, img.tile_order, bin.point_list, g.rec, bg,
                                                   img.n_contrib, out_color, out_depth, out_alpha);
    return cudaGetLastError();
}

}  // namespace gsr
