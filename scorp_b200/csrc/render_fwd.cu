// render_fwd.cu -- per-tile front-to-back alpha blend (3DGS forward).
//
// Replaces renderCUDA (reference diff-gaussian-rasterization/cuda_rasterizer/forward.cu:263-381).
// One CTA of 4 warps per 16x16 tile; each warp owns an 8x8 pixel block and every thread two vertically
// adjacent pixels, so the list walk, the record loads and the x-dependent terms are shared by the pair and
// the rest runs as packed FFMA2/FMUL2/FADD2 with the record fields as scalar operands. Per round of 256 list
// entries: (1) 48-byte records are gathered into shared memory with cp.async, double-buffered
// against the blend of the previous round; (2) every thread tests ITS record against the 8 warp
// blocks (exact minimum of the conic form over the block rectangle vs. the record's power
// threshold) and the warps exchange 8 ballots; (3) each warp blends only the entries whose bit
// is set. Skipped (pixel, Gaussian) pairs are exactly pairs the reference also skips
// (power > 0 or alpha < 1/255), so the image, alpha, depth and n_contrib are bit-identical.
#include "blend.cuh"
#include <cstdlib>

namespace gsr {

constexpr int FWD_THREADS = BLEND_THREADS;

__global__ void __launch_bounds__(FWD_THREADS, 8)
render_forward_kernel(int W, int H, int gx, int tiles_per_view,
                      const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order, const uint32_t* __restrict__ point_list,
                      const GaussRec* __restrict__ recs, const float* __restrict__ bg,
                      uint32_t* __restrict__ n_contrib, float* __restrict__ out_color,
                      float* __restrict__ out_depth, float* __restrict__ out_alpha, uint32_t* __restrict__ contrib) {
    __shared__ __align__(16) GaussRec s_rec[2][BATCH];
    __shared__ uint32_t s_bal[4][8];          // [consumer warp block][32-record word of the round]

    const int tile = (int)tile_order[blockIdx.x];      // longest lists first (binning.cu)
    // batched views: tile = view * tiles_per_view + local tile; every per-pixel output of view v starts v images in
    const int view_id = tile / tiles_per_view, ltile = tile - view_id * tiles_per_view;
    const int tile_x = (ltile % gx) * TILE, tile_y = (ltile / gx) * TILE;
    if (view_id) {
        const size_t vo = (size_t)view_id * H * W;
        n_contrib += vo; out_color += 3 * vo; out_depth += vo; out_alpha += vo;
    }
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int bx = (wid & 1) * 8, by = (wid >> 1) * 8;
    const int pxi = tile_x + bx + (lane & 7), pyi = tile_y + by + 2 * (lane >> 3);     // pixels (pxi, pyi) and (pxi, pyi+1)
    const bool in0 = pxi < W && pyi < H, in1 = pxi < W && pyi + 1 < H;
    const float pxf = (float)pxi;
    const float2 npy = make_float2(-(float)pyi, -(float)(pyi + 1));
    const uint32_t pix_id = (uint32_t)pyi * W + pxi;

    const uint2 range = ranges[tile];
    const int total = (int)(range.y - range.x);
    const int rounds = (total + BATCH - 1) / BATCH;
    // contribution words of this warp block for the backward (zero-initialised by the caller): bit = list entry
    uint32_t* my_contrib = contrib + contrib_base(range.x, tile) * 4 + wid;

    // per-pixel state, .x = upper pixel, .y = lower pixel
    float2 T = bc(1.0f), C0 = bc(0.0f), C1 = bc(0.0f), C2 = bc(0.0f), Dacc = bc(0.0f), Wacc = bc(0.0f);
    uint32_t last0 = 0, last1 = 0;
    uint32_t live = (in0 ? 1u : 0u) | (in1 ? 2u : 0u);      // bit h: pixel h still blending

    auto issue = [&](int r) {
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int slot = tid + k * FWD_THREADS;
            const int p = r * BATCH + slot;
            if (p < total) {
                const uint32_t id = __ldg(point_list + range.x + p);
                const char* src = reinterpret_cast<const char*>(recs + id);
                char* dst = reinterpret_cast<char*>(&s_rec[r & 1][slot]);
                cp_async16(dst, src); cp_async16(dst + 16, src + 16); cp_async16(dst + 32, src + 32);
            }
        }
        cp_async_commit();
    };

    if (rounds > 0) issue(0);
    for (int r = 0; r < rounds; ++r) {
        // all pixels finished? (also orders the previous round's reads before buffer reuse)
        if (__syncthreads_and(live == 0u)) break;
        if (r + 1 < rounds) { issue(r + 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
        __syncthreads();

        const int n_here = min(BATCH, total - r * BATCH);
        const GaussRec* buf = s_rec[r & 1];
        // per-record cull masks against the four 8x8 warp blocks -> ballots
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int slot = tid + k * FWD_THREADS;
            uint32_t mask = 0;
            if (slot < n_here) {
                const float4 g0 = *reinterpret_cast<const float4*>(&buf[slot]);
                const float4 g1 = *(reinterpret_cast<const float4*>(&buf[slot]) + 1);
                const CullRec c = cull_prepare(g0.x, g0.y, g0.z, g0.w, g1.x, g1.w);
#pragma unroll
                for (int w = 0; w < 4; ++w) {
                    const float x0 = (float)(tile_x + (w & 1) * 8), y0 = (float)(tile_y + (w >> 1) * 8);
                    if (block_may_contribute(c, x0, x0 + 7.0f, y0, y0 + 7.0f)) mask |= 1u << w;
                }
            }
#pragma unroll
            for (int w = 0; w < 4; ++w) {
                const uint32_t bal = __ballot_sync(0xffffffffu, (mask >> w) & 1u);
                if (lane == 0) s_bal[w][wid + 4 * k] = bal;
            }
        }
        __syncthreads();

        if (__any_sync(0xffffffffu, live != 0u)) {
            const uint32_t pos_base = (uint32_t)(r * BATCH) + 1u;
            for (int v = 0; v < 8; ++v) {
                uint32_t bits = s_bal[wid][v];
                uint32_t took_bits = 0;           // per lane; OR-reduced over the warp after the word
                while (bits) {
                    const int j = v * 32 + __ffs(bits) - 1;
                    const uint32_t rest = bits & (bits - 1);
                    const uint32_t bit = bits ^ rest;
                    bits = rest;
                    if (live == 0u) continue;
                    const float4 g0 = *reinterpret_cast<const float4*>(&buf[j]);
                    const float4 g1 = *(reinterpret_cast<const float4*>(&buf[j]) + 1);
                    float dx; float2 dy;
                    const float2 power = pair_power(g0.x, g0.y, g0.z, g0.w, g1.x, pxf, npy, dx, dy);
                    // candidate: still blending, power <= 0 and power >= thr (below thr implies alpha < 1/255)
                    bool c0 = (live & 1u) && !(power.x > 0.0f) && !(power.x < g1.w);
                    bool c1 = (live & 2u) && !(power.y > 0.0f) && !(power.y < g1.w);
                    if (!(c0 || c1)) continue;
                    const float2 ga = __fmul2_rn(bc(g1.y), expf2_nonpos(power));
                    float2 alpha = make_float2(fminf(0.99f, ga.x), fminf(0.99f, ga.y));
                    c0 = c0 && !(alpha.x < 1.0f / 255.0f);
                    c1 = c1 && !(alpha.y < 1.0f / 255.0f);
                    alpha.x = c0 ? alpha.x : 0.0f;
                    alpha.y = c1 ? alpha.y : 0.0f;
                    // test_T = T * (1 - alpha); a half with alpha == 0 keeps T exactly
                    const float2 test_T = __fmul2_rn(T, __ffma2_rn(alpha, bc(-1.0f), bc(1.0f)));
                    if (test_T.x < 0.0001f) { live &= ~1u; c0 = false; alpha.x = 0.0f; }     // T >= 1e-4 always, so only a
                    if (test_T.y < 0.0001f) { live &= ~2u; c1 = false; alpha.y = 0.0f; }     // contributing half can trip this
                    if (!(c0 || c1)) continue;
                    const float4 g2 = *(reinterpret_cast<const float4*>(&buf[j]) + 2);
                    // a half with alpha == 0 adds T*0 (finite colours and depth assumed)
                    C0 = __ffma2_rn(T, __fmul2_rn(alpha, bc(g2.x)), C0);
                    C1 = __ffma2_rn(T, __fmul2_rn(alpha, bc(g2.y)), C1);
                    C2 = __ffma2_rn(T, __fmul2_rn(alpha, bc(g2.z)), C2);
                    Wacc = __ffma2_rn(T, alpha, Wacc);
                    Dacc = __ffma2_rn(T, __fmul2_rn(alpha, bc(g1.z)), Dacc);
                    if (c0) { T.x = test_T.x; last0 = pos_base + (uint32_t)j; }
                    if (c1) { T.y = test_T.y; last1 = pos_base + (uint32_t)j; }
                    took_bits |= bit;
                }
                // the backward walks exactly the (warp block, entry) visits that blended something here
                took_bits = __reduce_or_sync(0xffffffffu, took_bits);
                if (lane == 0 && took_bits) my_contrib[(size_t)(r * 8 + v) * 4] = took_bits;
                if (!__any_sync(0xffffffffu, live != 0u)) break;
            }
        }
    }
    cp_async_wait<0>();

    const size_t hw = (size_t)H * W;
    const float b0 = __ldg(bg), b1 = __ldg(bg + 1), b2 = __ldg(bg + 2);
    if (in0) {
        n_contrib[pix_id] = last0;
        out_color[pix_id] = fmar(T.x, b0, C0.x);
        out_color[hw + pix_id] = fmar(T.x, b1, C1.x);
        out_color[2 * hw + pix_id] = fmar(T.x, b2, C2.x);
        out_alpha[pix_id] = Wacc.x;
        out_depth[pix_id] = Dacc.x;
    }
    if (in1) {
        const uint32_t p1 = pix_id + (uint32_t)W;
        n_contrib[p1] = last1;
        out_color[p1] = fmar(T.y, b0, C0.y);
        out_color[hw + p1] = fmar(T.y, b1, C1.y);
        out_color[2 * hw + p1] = fmar(T.y, b2, C2.y);
        out_alpha[p1] = Wacc.y;
        out_depth[p1] = Dacc.y;
    }
}

// ---- one warp per CTA (round 2b) --------------------------------------------------------------------------------------
// The four warp blocks of a tile share only the staged records. As a CTA they pay for that with three barriers per 256
// entries, a tile-wide "everyone done?" vote, and a CTA slot held until the slowest block finishes. Here a CTA is ONE
// warp and one 8x8 block: per 32 list entries, lane l stages entry l (id -> three 16-byte cp.async, double-buffered, ids
// fetched two words ahead), tests it against the warp's own block, and the warp blends the entries of the ballot. Same
// visit arithmetic, same order per pixel, so the outputs are bit-identical; a block whose pixels are saturated stops
// staging and testing at once. Cost: the records are read from L2 once per block instead of once per tile.
__global__ void __launch_bounds__(32, 32)
render_forward_warp_kernel(int W, int H, int gx, int tiles_per_view,
                           const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order, const uint32_t* __restrict__ point_list,
                           const GaussRec* __restrict__ recs, const float* __restrict__ bg,
                           uint32_t* __restrict__ n_contrib, float* __restrict__ out_color,
                           float* __restrict__ out_depth, float* __restrict__ out_alpha, uint32_t* __restrict__ contrib) {
    __shared__ __align__(16) GaussRec s_rec[2][32];
    constexpr uint32_t FULL = 0xffffffffu;
    const int tile = (int)tile_order[blockIdx.x >> 2];      // longest lists first (binning.cu)
    const int wid = blockIdx.x & 3, lane = threadIdx.x;
    const int view_id = tile / tiles_per_view, ltile = tile - view_id * tiles_per_view;
    const int tile_x = (ltile % gx) * TILE, tile_y = (ltile / gx) * TILE;
    if (view_id) {
        const size_t vo = (size_t)view_id * H * W;
        n_contrib += vo; out_color += 3 * vo; out_depth += vo; out_alpha += vo;
    }
    const int bx = (wid & 1) * 8, by = (wid >> 1) * 8;
    const int pxi = tile_x + bx + (lane & 7), pyi = tile_y + by + 2 * (lane >> 3);     // pixels (pxi, pyi) and (pxi, pyi+1)
    const bool in0 = pxi < W && pyi < H, in1 = pxi < W && pyi + 1 < H;
    const float pxf = (float)pxi;
    const float2 npy = make_float2(-(float)pyi, -(float)(pyi + 1));
    const uint32_t pix_id = (uint32_t)pyi * W + pxi;
    const float blk_x0 = (float)(tile_x + bx), blk_y0 = (float)(tile_y + by);

    const uint2 range = ranges[tile];
    const int total = (int)(range.y - range.x);
    const int words = (total + 31) >> 5;
    uint32_t* my_contrib = contrib + contrib_base(range.x, tile) * 4 + wid;
    const uint32_t* ids = point_list + range.x;

    float2 T = bc(1.0f), C0 = bc(0.0f), C1 = bc(0.0f), C2 = bc(0.0f), Dacc = bc(0.0f), Wacc = bc(0.0f);
    uint32_t last0 = 0, last1 = 0;
    uint32_t live = (in0 ? 1u : 0u) | (in1 ? 2u : 0u);      // bit h: pixel h still blending

    auto load_id = [&](int kk) { const int p = kk * 32 + lane; return p < total ? __ldg(ids + p) : 0xffffffffu; };
    auto stage = [&](int kk, uint32_t id) {
        if (id != 0xffffffffu) {
            const char* src = reinterpret_cast<const char*>(recs + id);
            char* dst = reinterpret_cast<char*>(&s_rec[kk & 1][lane]);
            cp_async16(dst, src); cp_async16(dst + 16, src + 16); cp_async16(dst + 32, src + 32);
        }
        cp_async_commit();
    };
    if (words > 0 && __any_sync(FULL, live != 0u)) {
        stage(0, load_id(0));
        uint32_t id_next = words > 1 ? load_id(1) : 0xffffffffu;
        for (int k = 0; k < words; ++k) {
            if (k + 1 < words) {
                stage(k + 1, id_next);
                id_next = (k + 2 < words) ? load_id(k + 2) : 0xffffffffu;
                cp_async_wait<1>();
            } else cp_async_wait<0>();
            __syncwarp();
            const GaussRec* buf = s_rec[k & 1];
            bool pass = false;
            if (k * 32 + lane < total) {
                const float4 g0 = *reinterpret_cast<const float4*>(&buf[lane]);
                const float4 g1 = *(reinterpret_cast<const float4*>(&buf[lane]) + 1);
                const CullRec c = cull_prepare(g0.x, g0.y, g0.z, g0.w, g1.x, g1.w);
                pass = block_may_contribute(c, blk_x0, blk_x0 + 7.0f, blk_y0, blk_y0 + 7.0f);
            }
            uint32_t bits = __ballot_sync(FULL, pass);
            const uint32_t pos_base = (uint32_t)(k * 32) + 1u;
            uint32_t took_bits = 0;           // per lane; OR-reduced over the warp after the word
            while (bits) {
                const int j = __ffs(bits) - 1;
                const uint32_t rest = bits & (bits - 1);
                const uint32_t bit = bits ^ rest;
                bits = rest;
                if (live == 0u) continue;
                const float4 g0 = *reinterpret_cast<const float4*>(&buf[j]);
                const float4 g1 = *(reinterpret_cast<const float4*>(&buf[j]) + 1);
                float dx; float2 dy;
                const float2 power = pair_power(g0.x, g0.y, g0.z, g0.w, g1.x, pxf, npy, dx, dy);
                bool c0 = (live & 1u) && !(power.x > 0.0f) && !(power.x < g1.w);
                bool c1 = (live & 2u) && !(power.y > 0.0f) && !(power.y < g1.w);
                if (!(c0 || c1)) continue;
                const float2 ga = __fmul2_rn(bc(g1.y), expf2_nonpos(power));
                float2 alpha = make_float2(fminf(0.99f, ga.x), fminf(0.99f, ga.y));
                c0 = c0 && !(alpha.x < 1.0f / 255.0f);
                c1 = c1 && !(alpha.y < 1.0f / 255.0f);
                alpha.x = c0 ? alpha.x : 0.0f;
                alpha.y = c1 ? alpha.y : 0.0f;
                const float2 test_T = __fmul2_rn(T, __ffma2_rn(alpha, bc(-1.0f), bc(1.0f)));
                if (test_T.x < 0.0001f) { live &= ~1u; c0 = false; alpha.x = 0.0f; }
                if (test_T.y < 0.0001f) { live &= ~2u; c1 = false; alpha.y = 0.0f; }
                if (!(c0 || c1)) continue;
                const float4 g2 = *(reinterpret_cast<const float4*>(&buf[j]) + 2);
                C0 = __ffma2_rn(T, __fmul2_rn(alpha, bc(g2.x)), C0);
                C1 = __ffma2_rn(T, __fmul2_rn(alpha, bc(g2.y)), C1);
                C2 = __ffma2_rn(T, __fmul2_rn(alpha, bc(g2.z)), C2);
                Wacc = __ffma2_rn(T, alpha, Wacc);
                Dacc = __ffma2_rn(T, __fmul2_rn(alpha, bc(g1.z)), Dacc);
                if (c0) { T.x = test_T.x; last0 = pos_base + (uint32_t)j; }
                if (c1) { T.y = test_T.y; last1 = pos_base + (uint32_t)j; }
                took_bits |= bit;
            }
            took_bits = __reduce_or_sync(FULL, took_bits);
            if (lane == 0 && took_bits) my_contrib[(size_t)k * 4] = took_bits;
            if (!__any_sync(FULL, live != 0u)) break;
            __syncwarp();                   // every lane is done with this buffer before the next stage overwrites it
        }
        cp_async_wait<0>();
    }

    const size_t hw = (size_t)H * W;
    const float b0 = __ldg(bg), b1 = __ldg(bg + 1), b2 = __ldg(bg + 2);
    if (in0) {
        n_contrib[pix_id] = last0;
        out_color[pix_id] = fmar(T.x, b0, C0.x);
        out_color[hw + pix_id] = fmar(T.x, b1, C1.x);
        out_color[2 * hw + pix_id] = fmar(T.x, b2, C2.x);
        out_alpha[pix_id] = Wacc.x;
        out_depth[pix_id] = Dacc.x;
    }
    if (in1) {
        const uint32_t p1 = pix_id + (uint32_t)W;
        n_contrib[p1] = last1;
        out_color[p1] = fmar(T.y, b0, C0.y);
        out_color[hw + p1] = fmar(T.y, b1, C1.y);
        out_color[2 * hw + p1] = fmar(T.y, b2, C2.y);
        out_alpha[p1] = Wacc.y;
        out_depth[p1] = Dacc.y;
    }
}

cudaError_t launch_render_forward(int W, int H, int gx, int gy, const GeomState& g, const ImageState& img,
                                  const BinningState& bin, const float* bg,
                                  float* out_color, float* out_depth, float* out_alpha, cudaStream_t st, int views) {
    static int per_warp = -1;           // GSR_FWD_WARP=0: the four-warp CTA with barriers (round 1/2a)
    if (per_warp < 0) { const char* e = getenv("GSR_FWD_WARP"); per_warp = e ? (atoi(e) ? 1 : 0) : 1; }
    if (per_warp)
        render_forward_warp_kernel<<<gx * gy * views * 4, 32, 0, st>>>(W, H, gx, gx * gy, img.ranges, img.tile_order, bin.point_list, g.rec, bg,
                                                                      img.n_contrib, out_color, out_depth, out_alpha,
                                                                      reinterpret_cast<uint32_t*>(bin.contrib));
    else
    render_forward_kernel<<<gx * gy * views, FWD_THREADS, 0, st>>>(W, H, gx, gx * gy, img.ranges, img.tile_order, bin.point_list, g.rec, bg,
                                                   img.n_contrib, out_color, out_depth, out_alpha,
                                                           reinterpret_cast<uint32_t*>(bin.contrib));
    return cudaGetLastError();
}

}  // namespace gsr
