// render_bwd.cu -- back-to-front re-traversal of each tile's list (3DGS backward blend).
//
// Replaces renderCUDA (reference diff-gaussian-rasterization/cuda_rasterizer/backward.cu:417-601).
// Same CTA/warp layout and cull ballots as the forward. Differences from the reference that do
// not change the mathematics:
//  * the traversal starts at the tile's largest n_contrib instead of the end of the list;
//  * the ten per-pair gradient terms are reduced over the warp with a 12-shuffle multi-value
//    butterfly and committed with ONE red.global.add.f32 per term per (warp, Gaussian), instead
//    of ten atomics per (pixel, Gaussian) pair;
//  * accumulators live in one 48-byte slot per Gaussian (ACC_* in common.cuh) that the fused
//    preprocess-backward kernel consumes.
#include "common.cuh"
#include <cstdlib>

namespace gsr {

__device__ __forceinline__ bool block_may_contribute_b(float x, float y, float a, float b, float c, float thr,
                                                       float inv_a_b, float inv_c_b, bool convex,
                                                       float px0, float px1, float py0, float py1) {
    if (!convex) return true;
    const float lx = x - px1, hx = x - px0, ly = y - py1, hy = y - py0;
    const float cx = fminf(fmaxf(0.0f, lx), hx), cy = fminf(fmaxf(0.0f, ly), hy);
    if (cx == 0.0f && cy == 0.0f) return true;
    float qmin = 3.0e38f;
    if (cx != 0.0f) {
        const float dy = fminf(fmaxf(-inv_c_b * cx, ly), hy);
        qmin = 0.5f * (a * cx * cx + c * dy * dy) + b * cx * dy;
    }
    if (cy != 0.0f) {
        const float dx = fminf(fmaxf(-inv_a_b * cy, lx), hx);
        qmin = fminf(qmin, 0.5f * (a * dx * dx + c * cy * cy) + b * dx * cy);
    }
    return !(-qmin < thr - 1e-3f - 1e-5f * fabsf(qmin));
}

// pair step of the multi-value butterfly: lanes with (lane & m) == 0 end up with p summed over
// {lane, lane^m}, the others with q.
__device__ __forceinline__ float bfly_pair(float p, float q, int m, bool hi) {
    const float send = hi ? p : q, keep = hi ? q : p;
    return keep + __shfl_xor_sync(0xffffffffu, send, m);
}
__device__ __forceinline__ float bfly_one(float p, int m) { return p + __shfl_xor_sync(0xffffffffu, p, m); }

__global__ void __launch_bounds__(256, 4)      // 64 registers: 4 CTAs/SM; measured 1.19 -> 1.06 ms at C2 (5 CTAs/SM: 1.09)
render_backward_kernel(int W, int H, int gx,
                       const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order, const uint32_t* __restrict__ point_list,
                       const GaussRec* __restrict__ recs, const float* __restrict__ bg,
                       const float* __restrict__ alphas, const uint32_t* __restrict__ n_contrib,
                       const float* __restrict__ dL_dpix, const float* __restrict__ dL_ddepth_pix,
                       const float* __restrict__ dL_dalpha_pix, float* __restrict__ acc, int direct_max) {
    __shared__ __align__(16) GaussRec s_rec[2][BATCH];
    __shared__ uint32_t s_bal[8][8];
    __shared__ uint32_t s_max[8];

    const int tile = (int)tile_order[blockIdx.x];      // longest lists first (binning.cu)
    const int tile_x = (tile % gx) * TILE, tile_y = (tile / gx) * TILE;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int bx = (wid & 1) * 8, by = (wid >> 1) * 4;
    const int pxi = tile_x + bx + (lane & 7), pyi = tile_y + by + (lane >> 3);
    const bool inside = pxi < W && pyi < H;
    const float pxf = (float)pxi, pyf = (float)pyi;
    const uint32_t pix_id = (uint32_t)pyi * W + pxi;
    const size_t hw = (size_t)H * W;

    const uint2 range = ranges[tile];
    const int total = (int)(range.y - range.x);

    const uint32_t last_contributor = inside ? n_contrib[pix_id] : 0u;
    const float T_final = inside ? (1.0f - alphas[pix_id]) : 0.0f;
    float dpix0 = 0.f, dpix1 = 0.f, dpix2 = 0.f, ddepth = 0.f, dalpha = 0.f;
    if (inside) {
        dpix0 = dL_dpix[pix_id]; dpix1 = dL_dpix[hw + pix_id]; dpix2 = dL_dpix[2 * hw + pix_id];
        ddepth = dL_ddepth_pix[pix_id]; dalpha = dL_dalpha_pix[pix_id];
    }
    const float bg_dot_dpixel = __ldg(bg) * dpix0 + __ldg(bg + 1) * dpix1 + __ldg(bg + 2) * dpix2;

    // tile-wide maximum of n_contrib
    {
        const uint32_t wm = __reduce_max_sync(0xffffffffu, last_contributor);
        if (lane == 0) s_max[wid] = wm;
    }
    __syncthreads();
    uint32_t max_contrib = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) max_contrib = max(max_contrib, s_max[w]);
    if (max_contrib == 0 || total == 0) return;
    const uint32_t warp_max = s_max[wid];

    float T = T_final;
    float ar0 = 0.f, ar1 = 0.f, ar2 = 0.f, lc0 = 0.f, lc1 = 0.f, lc2 = 0.f;
    float adr = 0.f, ldp = 0.f, aar = 0.f, last_alpha = 0.f;
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;

    // which reduced value this lane commits (see bfly order below)
    const bool b1 = lane & 2, b2 = lane & 4, b3 = lane & 8, b4 = lane & 16;
    const bool writer = !(lane & 1) && (!b1 || (!b2 && !b3));
    const int slot = b1 ? (8 + (b4 ? 1 : 0)) : ((b2 ? 4 : 0) + (b3 ? 2 : 0) + (b4 ? 1 : 0));

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

    const int r_start = (int)((max_contrib - 1) / BATCH);
    issue(r_start);
    for (int r = r_start; r >= 0; --r) {
        __syncthreads();                       // previous round's readers are done with both buffers' reuse
        if (r > 0) { issue(r - 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
        __syncthreads();

        const int n_here = min(BATCH, total - r * BATCH);
        const GaussRec* buf = s_rec[r & 1];
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
                    if (block_may_contribute_b(x, y, a, b, c, thr, inv_a_b, inv_c_b, convex, x0, x0 + 7.0f, y0, y0 + 3.0f))
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

        if ((uint32_t)(r * BATCH) >= warp_max) continue;     // nothing at or before this round for this warp
        for (int v = 7; v >= 0; --v) {
            uint32_t bits = s_bal[wid][v];
            while (bits) {
                const int jb = 31 - __clz(bits);
                bits &= ~(1u << jb);
                const int j = v * 32 + jb;
                const uint32_t pos = (uint32_t)(r * BATCH + j);      // 0-based list position
                if (pos >= warp_max) continue;

                const float4 g0 = *reinterpret_cast<const float4*>(&buf[j]);
                const float4 g1 = *(reinterpret_cast<const float4*>(&buf[j]) + 1);
                const float dx = subr(g0.x, pxf), dy = subr(g0.y, pyf);
                const float power = fmar(fmar(dx, mulr(dx, g0.z), mulr(dy, mulr(dy, g1.x))), -0.5f,
                                         -mulr(dy, mulr(dx, g0.w)));
                bool act = (pos < last_contributor) && !(power > 0.0f) && !(power < g1.w);
                float G = 0.f, alpha = 0.f;
                if (act) {
                    G = expf(power);
                    alpha = fminf(0.99f, mulr(g1.y, G));
                    act = !(alpha < 1.0f / 255.0f);
                }
                const uint32_t act_mask = __ballot_sync(0xffffffffu, act);
                if (act_mask == 0) continue;

                float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f, v4 = 0.f, v5 = 0.f, v6 = 0.f, v7 = 0.f, v8 = 0.f, v9 = 0.f;
                if (act) {
                    const float4 g2 = *(reinterpret_cast<const float4*>(&buf[j]) + 2);
                    const float inv_1ma = __frcp_rn(1.f - alpha);     // shared by the two divisions below (gradients: rtol 1e-4)
                    T = T * inv_1ma;
                    const float w = alpha * T;                       // dchannel_dcolor
                    float dL_dopa = 0.0f;
                    ar0 = last_alpha * lc0 + (1.f - last_alpha) * ar0; lc0 = g2.x; dL_dopa += (g2.x - ar0) * dpix0;
                    ar1 = last_alpha * lc1 + (1.f - last_alpha) * ar1; lc1 = g2.y; dL_dopa += (g2.y - ar1) * dpix1;
                    ar2 = last_alpha * lc2 + (1.f - last_alpha) * ar2; lc2 = g2.z; dL_dopa += (g2.z - ar2) * dpix2;
                    v6 = w * dpix0; v7 = w * dpix1; v8 = w * dpix2;
                    adr = last_alpha * ldp + (1.f - last_alpha) * adr; ldp = g1.z;
                    dL_dopa += (g1.z - adr) * ddepth;
                    v9 = w * ddepth;
                    aar = last_alpha + (1.f - last_alpha) * aar;
                    dL_dopa += (1.f - aar) * dalpha;
                    dL_dopa *= T;
                    last_alpha = alpha;
                    dL_dopa += (-T_final * inv_1ma) * bg_dot_dpixel;
                    const float dL_dG = g1.y * dL_dopa;
                    const float gdx = G * dx, gdy = G * dy;
                    const float dG_ddelx = -gdx * g0.z - gdy * g0.w;
                    const float dG_ddely = -gdy * g1.x - gdx * g0.w;
                    v0 = dL_dG * dG_ddelx * ddelx_dx;
                    v1 = dL_dG * dG_ddely * ddely_dy;
                    v2 = -0.5f * gdx * dx * dL_dG;
                    v3 = -0.5f * gdx * dy * dL_dG;
                    v4 = -0.5f * gdy * dy * dL_dG;
                    v5 = G * dL_dopa;
                }
                // Few contributing pixels (splat edges): the active lanes commit their terms directly, 10 predicated
                // RED instructions instead of the ~45-instruction butterfly below.
                if (__popc(act_mask) <= direct_max) {
                    if (act) {
                        float* a = acc + (size_t)__ldg(point_list + range.x + pos) * ACC_STRIDE;
                        atomicAdd(a + 0, v0); atomicAdd(a + 1, v1); atomicAdd(a + 2, v2); atomicAdd(a + 3, v3); atomicAdd(a + 4, v4);
                        atomicAdd(a + 5, v5); atomicAdd(a + 6, v6); atomicAdd(a + 7, v7); atomicAdd(a + 8, v8); atomicAdd(a + 9, v9);
                    }
                    continue;
                }
                // 10 values -> 12 shuffles; final owner lanes given by (writer, slot)
                const float a0 = bfly_pair(v0, v1, 16, b4), a1 = bfly_pair(v2, v3, 16, b4), a2 = bfly_pair(v4, v5, 16, b4);
                const float a3 = bfly_pair(v6, v7, 16, b4), a4 = bfly_pair(v8, v9, 16, b4);
                const float c0 = bfly_pair(a0, a1, 8, b3), c1 = bfly_pair(a2, a3, 8, b3), c2 = bfly_one(a4, 8);
                const float d0 = bfly_pair(c0, c1, 4, b2), d1 = bfly_one(c2, 4);
                const float e0 = bfly_pair(d0, d1, 2, b1);
                const float f0 = bfly_one(e0, 1);
                if (writer && f0 != 0.0f) {
                    const uint32_t id = __ldg(point_list + range.x + pos);
                    atomicAdd(acc + (size_t)id * ACC_STRIDE + slot, f0);
                }
            }
        }
    }
}

// Visits with at most this many contributing pixels skip the warp reduction (tunable: GSR_DIRECT_MAX).
// Measured on B200 at C2: blend backward 1.244 ms (0) -> 1.192 ms (8); the 18-term surfel variant only loses (default 0).
int blend_direct_max(int dflt) {
    static int env = -2;
    if (env == -2) { const char* e = getenv("GSR_DIRECT_MAX"); env = e ? atoi(e) : -1; if (env > 32) env = 32; }
    return env >= 0 ? env : dflt;
}

cudaError_t launch_render_backward(int W, int H, int gx, int gy, const GeomState& g, const ImageState& img,
                                   const BinningState& bin, const float* bg, const float* out_alpha,
                                   const float* dL_dpix, const float* dL_ddepth, const float* dL_dalpha,
                                   float* acc, cudaStream_t st) {
    render_backward_kernel<<<gx * gy, 256, 0, st>>>(W, H, gx, img.ranges, img.tile_order, bin.point_list, g.rec, bg, out_alpha,
                                                    img.n_contrib, dL_dpix, dL_ddepth, dL_dalpha, acc, blend_direct_max(8));
    return cudaGetLastError();
}

}  // namespace gsr
