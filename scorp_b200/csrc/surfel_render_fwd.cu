// surfel_render_fwd.cu -- per-tile front-to-back blend of 2DGS surfels with ray-splat intersection.
//
// Replaces renderCUDA (reference diff-surfel-rasterization/cuda_rasterizer/forward.cu:258-448).
// Same CTA layout as the 3DGS blend (render_fwd.cu): one CTA per 16x16 tile, each warp owns an 8x4
// pixel block, 96-byte records gathered with cp.async and double-buffered, per-record cull ballots
// (here: the record's conservative contribution box against the warp's pixel rectangle).
// The per-pair arithmetic is the reference's expression for expression (default contraction).
#include "surfel.cuh"
#include <cstdlib>

namespace gsr {

constexpr float kNear = 0.2f, kFar = 100.0f;     // near_n, far_n (auxiliary.h:37-38)
constexpr int SURFEL_SMEM = 2 * BATCH * (int)sizeof(SurfelRec) + 8 * 8 * 4;

__global__ void __launch_bounds__(256)
surfel_render_forward_kernel(int W, int H, int gx,
                             const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order, const uint32_t* __restrict__ point_list,
                             const SurfelRec* __restrict__ recs, const float* __restrict__ bg,
                             uint32_t* __restrict__ n_contrib, uint32_t* __restrict__ median_contrib,
                             float* __restrict__ final_T, float* __restrict__ out_M1, float* __restrict__ out_M2,
                             float* __restrict__ out_color, float* __restrict__ out_others, uint32_t* __restrict__ contrib) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SurfelRec (*s_rec)[BATCH] = reinterpret_cast<SurfelRec (*)[BATCH]>(smem_raw);
    uint32_t (*s_bal)[8] = reinterpret_cast<uint32_t (*)[8]>(smem_raw + 2 * BATCH * sizeof(SurfelRec));

    const int tile = (int)tile_order[blockIdx.x];      // longest lists first (binning.cu)
    const int tile_x = (tile % gx) * TILE, tile_y = (tile / gx) * TILE;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int bx = (wid & 1) * 8, by = (wid >> 1) * 4;
    const uint32_t pxi = tile_x + bx + (lane & 7), pyi = tile_y + by + (lane >> 3);
    const bool inside = pxi < (uint32_t)W && pyi < (uint32_t)H;
    const float pxf = (float)pxi, pyf = (float)pyi;
    const uint32_t pix_id = pyi * W + pxi;

    const uint2 range = ranges[tile];
    const int total = (int)(range.y - range.x);
    const int rounds = (total + BATCH - 1) / BATCH;
    // contribution words of this warp block for the backward (zero-initialised by the caller): bit = list entry that
    // blended into some pixel of the block, i.e. exactly the (warp block, entry) visits the backward has work for
    uint32_t* my_contrib = contrib + contrib_base(range.x, tile) * 8 + wid;

    float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f;
    float N0 = 0.f, N1 = 0.f, N2 = 0.f, D = 0.f, M1 = 0.f, M2 = 0.f, distortion = 0.f, median_depth = 0.f;
    uint32_t last_contributor = 0, median_contributor = 0;
    bool done = !inside;

    auto issue = [&](int r) {
        const int p = r * BATCH + tid;
        if (p < total) {
            const uint32_t id = __ldg(point_list + range.x + p);
            const char* src = reinterpret_cast<const char*>(recs + id);
            char* dst = reinterpret_cast<char*>(&s_rec[r & 1][tid]);
#pragma unroll
            for (int q = 0; q < 6; ++q) cp_async16(dst + 16 * q, src + 16 * q);
        }
        cp_async_commit();
    };

    if (rounds > 0) issue(0);
    for (int r = 0; r < rounds; ++r) {
        if (__syncthreads_and(done)) break;
        if (r + 1 < rounds) { issue(r + 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
        __syncthreads();

        const int n_here = min(BATCH, total - r * BATCH);
        const SurfelRec* buf = s_rec[r & 1];
        {
            uint32_t mask = 0;
            if (tid < n_here) {
                const float4 box = *(reinterpret_cast<const float4*>(&buf[tid]) + 5);
#pragma unroll
                for (int w = 0; w < 8; ++w) {
                    const float x0 = (float)(tile_x + (w & 1) * 8), y0 = (float)(tile_y + (w >> 1) * 4);
                    if (box.x <= x0 + 7.0f && box.y >= x0 && box.z <= y0 + 3.0f && box.w >= y0) mask |= 1u << w;
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
                uint32_t took_bits = 0;           // per lane; OR-reduced over the warp after the word
                while (bits) {
                    const int j = v * 32 + __ffs(bits) - 1;
                    const uint32_t rest = bits & (bits - 1);
                    const uint32_t bit = bits ^ rest;
                    bits = rest;
                    if (done) continue;
                    const float4* q = reinterpret_cast<const float4*>(&buf[j]);
                    const float4 q0 = q[0], q1 = q[1], q2 = q[2];
                    const float3 Tu = make_float3(q0.z, q0.w, q1.x), Tv = make_float3(q1.y, q1.z, q1.w);
                    const float3 Tw = make_float3(q2.x, q2.y, q2.z);
                    // forward.cu:355-377
                    const float3 k = make_float3(pxf * Tw.x - Tu.x, pxf * Tw.y - Tu.y, pxf * Tw.z - Tu.z);
                    const float3 l = make_float3(pyf * Tw.x - Tv.x, pyf * Tw.y - Tv.y, pyf * Tw.z - Tv.z);
                    const float3 p = make_float3(k.y * l.z - k.z * l.y, k.z * l.x - k.x * l.z, k.x * l.y - k.y * l.x);
                    if (p.z == 0.0f) continue;
                    const float sx = p.x / p.z, sy = p.y / p.z;
                    const float rho3d = (sx * sx + sy * sy);
                    const float dx = q0.x - pxf, dy = q0.y - pyf;
                    const float rho2d = 2.0f * (dx * dx + dy * dy);
                    const float rho = fminf(rho3d, rho2d);
                    const float depth = (sx * Tw.x + sy * Tw.y) + Tw.z;
                    if (depth < kNear) continue;
                    const float4 q3 = q[3];
                    const float power = -0.5f * rho;
                    if (power > 0.0f) continue;
                    const float alpha = fminf(0.99f, q3.z * expf(power));
                    if (alpha < 1.0f / 255.0f) continue;
                    const float test_T = T * (1 - alpha);
                    if (test_T < 0.0001f) { done = true; continue; }
                    const float w = alpha * T;
                    const uint32_t contributor = (uint32_t)(r * BATCH + j + 1);
                    // forward.cu:391-411
                    const float A = 1 - T;
                    const float m = kFar / (kFar - kNear) * (1 - kNear / depth);
                    distortion += (m * m * A + M2 - 2 * m * M1) * w;
                    D += depth * w;
                    M1 += m * w;
                    M2 += m * m * w;
                    if (T > 0.5f) { median_depth = depth; median_contributor = contributor; }
                    N0 += q2.w * w; N1 += q3.x * w; N2 += q3.y * w;
                    const float4 q4 = q[4];
                    C0 += q3.w * w; C1 += q4.x * w; C2 += q4.y * w;
                    T = test_T;
                    last_contributor = contributor;
                    took_bits |= bit;
                }
                took_bits = __reduce_or_sync(0xffffffffu, took_bits);
                if (lane == 0 && took_bits) my_contrib[(size_t)(r * 8 + v) * 8] = took_bits;
                if (__all_sync(0xffffffffu, done)) break;
            }
        }
    }
    cp_async_wait<0>();

    if (inside) {
        const size_t hw = (size_t)H * W;
        final_T[pix_id] = T;
        n_contrib[pix_id] = last_contributor;
        median_contrib[pix_id] = median_contributor;
        out_M1[pix_id] = M1;
        out_M2[pix_id] = M2;
        out_color[pix_id] = C0 + T * __ldg(bg);
        out_color[hw + pix_id] = C1 + T * __ldg(bg + 1);
        out_color[2 * hw + pix_id] = C2 + T * __ldg(bg + 2);
        out_others[pix_id] = D;                         // DEPTH_OFFSET 0
        out_others[hw + pix_id] = 1 - T;                // ALPHA_OFFSET 1
        out_others[2 * hw + pix_id] = N0;               // NORMAL_OFFSET 2..4
        out_others[3 * hw + pix_id] = N1;
        out_others[4 * hw + pix_id] = N2;
        out_others[5 * hw + pix_id] = median_depth;     // MIDDEPTH_OFFSET 5
        out_others[6 * hw + pix_id] = distortion;       // DISTORTION_OFFSET 6
    }
}

// ---- one warp per CTA (round 2b; see render_fwd.cu, render_forward_warp_kernel) ------------------------------------------
// A CTA is one warp and one 8x4 pixel block: per 32 list entries lane l stages entry l (id -> six 16-byte cp.async,
// double-buffered, ids fetched two words ahead), tests its contribution box against the warp's own rectangle, and the warp
// blends the entries of the ballot. No CTA barriers; a saturated block stops at once; same arithmetic and order per pixel.
__global__ void __launch_bounds__(32, 32)
surfel_render_forward_warp_kernel(int W, int H, int gx,
                                  const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order, const uint32_t* __restrict__ point_list,
                                  const SurfelRec* __restrict__ recs, const float* __restrict__ bg,
                                  uint32_t* __restrict__ n_contrib, uint32_t* __restrict__ median_contrib,
                                  float* __restrict__ final_T, float* __restrict__ out_M1, float* __restrict__ out_M2,
                                  float* __restrict__ out_color, float* __restrict__ out_others, uint32_t* __restrict__ contrib) {
    __shared__ __align__(16) SurfelRec s_rec[2][32];
    constexpr uint32_t FULL = 0xffffffffu;
    const int tile = (int)tile_order[blockIdx.x >> 3];      // longest lists first (binning.cu)
    const int wid = blockIdx.x & 7, lane = threadIdx.x;
    const int tile_x = (tile % gx) * TILE, tile_y = (tile / gx) * TILE;
    const int bx = (wid & 1) * 8, by = (wid >> 1) * 4;
    const uint32_t pxi = tile_x + bx + (lane & 7), pyi = tile_y + by + (lane >> 3);
    const bool inside = pxi < (uint32_t)W && pyi < (uint32_t)H;
    const float pxf = (float)pxi, pyf = (float)pyi;
    const uint32_t pix_id = pyi * W + pxi;
    const float blk_x0 = (float)(tile_x + bx), blk_y0 = (float)(tile_y + by);

    const uint2 range = ranges[tile];
    const int total = (int)(range.y - range.x);
    const int words = (total + 31) >> 5;
    uint32_t* my_contrib = contrib + contrib_base(range.x, tile) * 8 + wid;
    const uint32_t* ids = point_list + range.x;

    float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f;
    float N0 = 0.f, N1 = 0.f, N2 = 0.f, D = 0.f, M1 = 0.f, M2 = 0.f, distortion = 0.f, median_depth = 0.f;
    uint32_t last_contributor = 0, median_contributor = 0;
    bool done = !inside;

    auto load_id = [&](int kk) { const int p = kk * 32 + lane; return p < total ? __ldg(ids + p) : 0xffffffffu; };
    auto stage = [&](int kk, uint32_t id) {
        if (id != 0xffffffffu) {
            const char* src = reinterpret_cast<const char*>(recs + id);
            char* dst = reinterpret_cast<char*>(&s_rec[kk & 1][lane]);
#pragma unroll
            for (int q = 0; q < 6; ++q) cp_async16(dst + 16 * q, src + 16 * q);
        }
        cp_async_commit();
    };
    if (words > 0 && !__all_sync(FULL, done)) {
        stage(0, load_id(0));
        uint32_t id_next = words > 1 ? load_id(1) : 0xffffffffu;
        for (int kw = 0; kw < words; ++kw) {
            if (kw + 1 < words) {
                stage(kw + 1, id_next);
                id_next = (kw + 2 < words) ? load_id(kw + 2) : 0xffffffffu;
                cp_async_wait<1>();
            } else cp_async_wait<0>();
            __syncwarp();
            const SurfelRec* buf = s_rec[kw & 1];
            bool pass = false;
            if (kw * 32 + lane < total) {
                const float4 box = *(reinterpret_cast<const float4*>(&buf[lane]) + 5);
                pass = box.x <= blk_x0 + 7.0f && box.y >= blk_x0 && box.z <= blk_y0 + 3.0f && box.w >= blk_y0;
            }
            uint32_t bits = __ballot_sync(FULL, pass);
            const uint32_t pos_base = (uint32_t)(kw * 32) + 1u;
            uint32_t took_bits = 0;           // per lane; OR-reduced over the warp after the word
            while (bits) {
                const int j = __ffs(bits) - 1;
                const uint32_t rest = bits & (bits - 1);
                const uint32_t bit = bits ^ rest;
                bits = rest;
                if (done) continue;
                const float4* q = reinterpret_cast<const float4*>(&buf[j]);
                const float4 q0 = q[0], q1 = q[1], q2 = q[2];
                const float3 Tu = make_float3(q0.z, q0.w, q1.x), Tv = make_float3(q1.y, q1.z, q1.w);
                const float3 Tw = make_float3(q2.x, q2.y, q2.z);
                // forward.cu:355-377
                const float3 k = make_float3(pxf * Tw.x - Tu.x, pxf * Tw.y - Tu.y, pxf * Tw.z - Tu.z);
                const float3 l = make_float3(pyf * Tw.x - Tv.x, pyf * Tw.y - Tv.y, pyf * Tw.z - Tv.z);
                const float3 p = make_float3(k.y * l.z - k.z * l.y, k.z * l.x - k.x * l.z, k.x * l.y - k.y * l.x);
                if (p.z == 0.0f) continue;
                const float sx = p.x / p.z, sy = p.y / p.z;
                const float rho3d = (sx * sx + sy * sy);
                const float dx = q0.x - pxf, dy = q0.y - pyf;
                const float rho2d = 2.0f * (dx * dx + dy * dy);
                const float rho = fminf(rho3d, rho2d);
                const float depth = (sx * Tw.x + sy * Tw.y) + Tw.z;
                if (depth < kNear) continue;
                const float4 q3 = q[3];
                const float power = -0.5f * rho;
                if (power > 0.0f) continue;
                const float alpha = fminf(0.99f, q3.z * expf(power));
                if (alpha < 1.0f / 255.0f) continue;
                const float test_T = T * (1 - alpha);
                if (test_T < 0.0001f) { done = true; continue; }
                const float w = alpha * T;
                const uint32_t contributor = pos_base + (uint32_t)j;
                // forward.cu:391-411
                const float A = 1 - T;
                const float m = kFar / (kFar - kNear) * (1 - kNear / depth);
                distortion += (m * m * A + M2 - 2 * m * M1) * w;
                D += depth * w;
                M1 += m * w;
                M2 += m * m * w;
                if (T > 0.5f) { median_depth = depth; median_contributor = contributor; }
                N0 += q2.w * w; N1 += q3.x * w; N2 += q3.y * w;
                const float4 q4 = q[4];
                C0 += q3.w * w; C1 += q4.x * w; C2 += q4.y * w;
                T = test_T;
                last_contributor = contributor;
                took_bits |= bit;
            }
            took_bits = __reduce_or_sync(FULL, took_bits);
            if (lane == 0 && took_bits) my_contrib[(size_t)kw * 8] = took_bits;
            if (__all_sync(FULL, done)) break;
            __syncwarp();                   // every lane is done with this buffer before the next stage overwrites it
        }
        cp_async_wait<0>();
    }

    if (inside) {
        const size_t hw = (size_t)H * W;
        final_T[pix_id] = T;
        n_contrib[pix_id] = last_contributor;
        median_contrib[pix_id] = median_contributor;
        out_M1[pix_id] = M1;
        out_M2[pix_id] = M2;
        out_color[pix_id] = C0 + T * __ldg(bg);
        out_color[hw + pix_id] = C1 + T * __ldg(bg + 1);
        out_color[2 * hw + pix_id] = C2 + T * __ldg(bg + 2);
        out_others[pix_id] = D;                         // DEPTH_OFFSET 0
        out_others[hw + pix_id] = 1 - T;                // ALPHA_OFFSET 1
        out_others[2 * hw + pix_id] = N0;               // NORMAL_OFFSET 2..4
        out_others[3 * hw + pix_id] = N1;
        out_others[4 * hw + pix_id] = N2;
        out_others[5 * hw + pix_id] = median_depth;     // MIDDEPTH_OFFSET 5
        out_others[6 * hw + pix_id] = distortion;       // DISTORTION_OFFSET 6
    }
}

cudaError_t launch_surfel_render_forward(int W, int H, int gx, int gy, const SurfelGeom& g, const SurfelImage& img,
                                         const BinningState& bin, const float* bg, float* out_color, float* out_others,
                                         cudaStream_t st) {
    // GSR_SURFEL_FWD_WARP=1: the one-warp kernel. Measured at C4: 0.738 ms against 0.721 for the eight-warp CTA below --
    // with eight blocks per tile and 96-byte records the per-block staging reads 8x the records from L2, which costs
    // more than the barriers did -- so the CTA kernel stays the default here (the 3DGS forward gains 5 %: four blocks, 48 bytes).
    static int per_warp = -1;
    if (per_warp < 0) { const char* e = getenv("GSR_SURFEL_FWD_WARP"); per_warp = (e && atoi(e)) ? 1 : 0; }
    if (per_warp) {
        surfel_render_forward_warp_kernel<<<gx * gy * 8, 32, 0, st>>>(W, H, gx, img.ranges, img.tile_order, bin.point_list, g.rec, bg,
                                                                    img.n_contrib, img.median_contrib, img.final_T, img.M1,
                                                                    img.M2, out_color, out_others,
                                                                    reinterpret_cast<uint32_t*>(bin.contrib));
        return cudaGetLastError();
    }
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(surfel_render_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SURFEL_SMEM);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    surfel_render_forward_kernel<<<gx * gy, 256, SURFEL_SMEM, st>>>(W, H, gx, img.ranges, img.tile_order, bin.point_list, g.rec, bg,
                                                                    img.n_contrib, img.median_contrib, img.final_T, img.M1,
                                                                    img.M2, out_color, out_others,
                                                                    reinterpret_cast<uint32_t*>(bin.contrib));
    return cudaGetLastError();
}

}  // namespace gsr
