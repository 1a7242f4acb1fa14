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
#include "blend.cuh"
#include <cstdlib>

namespace gsr {

// pair step of the multi-value butterfly: lanes with (lane & m) == 0 end up with p summed over
// {lane, lane^m}, the others with q.
__device__ __forceinline__ float bfly_pair(float p, float q, int m, bool hi) {
    const float send = hi ? p : q, keep = hi ? q : p;
    return keep + __shfl_xor_sync(0xffffffffu, send, m);
}
__device__ __forceinline__ float bfly_one(float p, int m) { return p + __shfl_xor_sync(0xffffffffu, p, m); }

__device__ __forceinline__ float hsum(float2 v) { return v.x + v.y; }

__global__ void __launch_bounds__(BLEND_THREADS, 8)
render_backward_kernel(int W, int H, int gx,
                       const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order, const uint32_t* __restrict__ point_list,
                       const GaussRec* __restrict__ recs, const float* __restrict__ bg,
                       const float* __restrict__ alphas, const uint32_t* __restrict__ n_contrib,
                       const float* __restrict__ dL_dpix, const float* __restrict__ dL_ddepth_pix,
                       const float* __restrict__ dL_dalpha_pix, float* __restrict__ acc, int direct_max,
                       const uint32_t* __restrict__ contrib) {
    __shared__ __align__(16) GaussRec s_rec[2][BATCH];
    __shared__ uint32_t s_bal[4][8];          // [consumer warp block][32-record word of the round]
    __shared__ uint32_t s_max[4];

    const int tile = (int)tile_order[blockIdx.x];      // longest lists first (binning.cu)
    const int tile_x = (tile % gx) * TILE, tile_y = (tile / gx) * TILE;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int bx = (wid & 1) * 8, by = (wid >> 1) * 8;
    const int pxi = tile_x + bx + (lane & 7), pyi = tile_y + by + 2 * (lane >> 3);     // pixels (pxi, pyi) and (pxi, pyi+1)
    const bool in0 = pxi < W && pyi < H, in1 = pxi < W && pyi + 1 < H;
    const float pxf = (float)pxi;
    const float2 npy = make_float2(-(float)pyi, -(float)(pyi + 1));
    const uint32_t p0 = (uint32_t)pyi * W + pxi, p1 = p0 + (uint32_t)W;
    const size_t hw = (size_t)H * W;

    const uint2 range = ranges[tile];
    const int total = (int)(range.y - range.x);
    const uint32_t* my_contrib = contrib + contrib_base(range.x, tile) * 4;

    // per-pixel constants, .x = upper pixel, .y = lower pixel (zeros outside the image)
    const uint32_t last0 = in0 ? n_contrib[p0] : 0u, last1 = in1 ? n_contrib[p1] : 0u;
    const float2 T_final = make_float2(in0 ? 1.0f - alphas[p0] : 0.0f, in1 ? 1.0f - alphas[p1] : 0.0f);
    auto ld2 = [&](const float* base) { return make_float2(in0 ? base[p0] : 0.0f, in1 ? base[p1] : 0.0f); };
    const float2 dpix0 = ld2(dL_dpix), dpix1 = ld2(dL_dpix + hw), dpix2 = ld2(dL_dpix + 2 * hw);
    const float2 ddepth = ld2(dL_ddepth_pix), dalpha = ld2(dL_dalpha_pix);
    const float bg0 = __ldg(bg), bg1 = __ldg(bg + 1), bg2 = __ldg(bg + 2);
    // -T_final * (bg . dL_dpixel): the background term of dL/dalpha is this times 1/(1-alpha)  (backward.cu:545-549)
    const float2 nTf_bgdot = make_float2(-T_final.x * (bg0 * dpix0.x + bg1 * dpix1.x + bg2 * dpix2.x),
                                         -T_final.y * (bg0 * dpix0.y + bg1 * dpix1.y + bg2 * dpix2.y));

    // tile-wide maximum of n_contrib
    {
        const uint32_t wm = __reduce_max_sync(0xffffffffu, max(last0, last1));
        if (lane == 0) s_max[wid] = wm;
    }
    __syncthreads();
    const uint32_t max_contrib = max(max(s_max[0], s_max[1]), max(s_max[2], s_max[3]));
    if (max_contrib == 0 || total == 0) return;
    const uint32_t warp_max = s_max[wid];

    // Running state per pixel. The reference carries (last_alpha, last_color, accum_rec) and forms
    // accum_rec = last_alpha*last_color + (1-last_alpha)*accum_rec at the next contributor (backward.cu:512-528);
    // B below is that value formed one step earlier, B <- B + alpha*(c - B), which is the identity for a
    // non-contributing pixel (alpha = 0), so the pair needs no per-pixel selects.
    float2 T = T_final, B0 = bc(0.f), B1 = bc(0.f), B2 = bc(0.f), Bd = bc(0.f), Ba = bc(0.f);
    const float nddelx = -0.5f * W, nddely = -0.5f * H;

    // which reduced value this lane commits (see bfly order below)
    // the lane id in an opaque register: the role bits are then one LOP3-with-predicate each where they are used, instead
    // of being re-derived from SR_TID.X inside every visit (ptxas cannot keep five predicates alive across the loop)
    uint32_t lane_o;
    asm volatile("mov.u32 %0, %1;" : "=r"(lane_o) : "r"((uint32_t)lane));
    const bool b1 = lane_o & 2, b2 = lane_o & 4, b3 = lane_o & 8, b4 = lane_o & 16;
    const bool writer = !(lane_o & 1) && (!b1 || (!b2 && !b3));
    const int slot = b1 ? (8 + (b4 ? 1 : 0)) : ((b2 ? 4 : 0) + (b3 ? 2 : 0) + (b4 ? 1 : 0));

    auto issue = [&](int r) {
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int sl = tid + k * BLEND_THREADS;
            const int p = r * BATCH + sl;
            if (p < total) {
                const uint32_t id = __ldg(point_list + range.x + p);
                const char* src = reinterpret_cast<const char*>(recs + id);
                char* dst = reinterpret_cast<char*>(&s_rec[r & 1][sl]);
                cp_async16(dst, src); cp_async16(dst + 16, src + 16); cp_async16(dst + 32, src + 32);
            }
        }
        cp_async_commit();
    };

    const int r_start = (int)((max_contrib - 1) / BATCH);
    issue(r_start);
    for (int r = r_start; r >= 0; --r) {
        __syncthreads();                       // previous round's readers are done with both buffers' reuse
        if (r > 0) { issue(r - 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
        __syncthreads();

        // shared-window address of this round's buffer, made opaque so that it stays in ONE register for the round: left
        // to itself ptxas re-derives it (S2R SR_CgaCtaId, LEA, 2 x IMAD ...) in front of every visit's loads
        uint32_t buf_s;
        asm volatile("mov.u32 %0, %1;" : "=r"(buf_s) : "r"((uint32_t)__cvta_generic_to_shared(s_rec[r & 1])));
        // which entries to visit: the forward recorded, per warp block, the entries that blended into some pixel
        // of the block (render_fwd.cu); everything else contributes nothing here either
        if (tid < 32) {
            const int w = tid >> 3, v = tid & 7;
            s_bal[w][v] = (r * BATCH + v * 32 < total) ? __ldg(my_contrib + (size_t)(r * 8 + v) * 4 + w) : 0u;
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

                const uint32_t rec_s = buf_s + (uint32_t)j * (uint32_t)sizeof(GaussRec);
                const float4 g0 = lds128(rec_s);
                const float4 g1 = lds128(rec_s + 16);
                float dx; float2 dy;
                const float2 power = pair_power(g0.x, g0.y, g0.z, g0.w, g1.x, pxf, npy, dx, dy);
                bool a0 = (pos < last0) && !(power.x > 0.0f) && !(power.x < g1.w);
                bool a1 = (pos < last1) && !(power.y > 0.0f) && !(power.y < g1.w);
                float2 G = bc(0.f), alpha = bc(0.f);
                if (a0 || a1) {
                    G = expf2_nonpos(power);                          // the forward's expf, bit for bit
                    const float2 ga = __fmul2_rn(bc(g1.y), G);
                    alpha = make_float2(fminf(0.99f, ga.x), fminf(0.99f, ga.y));
                    a0 = a0 && !(alpha.x < 1.0f / 255.0f);
                    a1 = a1 && !(alpha.y < 1.0f / 255.0f);
                }
                const bool act = a0 || a1;
                const uint32_t act_mask = __ballot_sync(0xffffffffu, act);
                if (act_mask == 0) continue;

                float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f, v4 = 0.f, v5 = 0.f, v6 = 0.f, v7 = 0.f, v8 = 0.f, v9 = 0.f;
                if (act) {
                    // a pixel that does not take this Gaussian runs with G = alpha = 0, 1/(1-alpha) = 1: every term
                    // below is then zero and its state unchanged (finite colours and depth assumed)
                    if (!a0) { G.x = 0.f; alpha.x = 0.f; }
                    if (!a1) { G.y = 0.f; alpha.y = 0.f; }
                    const float4 g2 = lds128(rec_s + 32);
                    const float2 oma = __ffma2_rn(alpha, bc(-1.0f), bc(1.0f));
                    const float2 inv = make_float2(a0 ? rcp_approx(oma.x) : 1.0f, a1 ? rcp_approx(oma.y) : 1.0f);
                    T = __fmul2_rn(T, inv);
                    const float2 w = __fmul2_rn(alpha, T);                  // dchannel_dcolor
                    const float2 d0 = __ffma2_rn(B0, bc(-1.0f), bc(g2.x)), d1 = __ffma2_rn(B1, bc(-1.0f), bc(g2.y));
                    const float2 d2 = __ffma2_rn(B2, bc(-1.0f), bc(g2.z)), dd = __ffma2_rn(Bd, bc(-1.0f), bc(g1.z));
                    const float2 da = __ffma2_rn(Ba, bc(-1.0f), bc(1.0f));
                    float2 dL = __fmul2_rn(d0, dpix0);
                    dL = __ffma2_rn(d1, dpix1, dL); dL = __ffma2_rn(d2, dpix2, dL);
                    dL = __ffma2_rn(dd, ddepth, dL); dL = __ffma2_rn(da, dalpha, dL);
                    B0 = __ffma2_rn(alpha, d0, B0); B1 = __ffma2_rn(alpha, d1, B1); B2 = __ffma2_rn(alpha, d2, B2);
                    Bd = __ffma2_rn(alpha, dd, Bd); Ba = __ffma2_rn(alpha, da, Ba);
                    v6 = hsum(__fmul2_rn(w, dpix0)); v7 = hsum(__fmul2_rn(w, dpix1)); v8 = hsum(__fmul2_rn(w, dpix2));
                    v9 = hsum(__fmul2_rn(w, ddepth));
                    dL = __ffma2_rn(dL, T, __fmul2_rn(nTf_bgdot, inv));     // dL/dalpha
                    const float2 dLdG = __fmul2_rn(bc(g1.y), dL);
                    const float2 gdx = __fmul2_rn(G, bc(dx)), gdy = __fmul2_rn(G, dy);
                    const float2 dGx = __ffma2_rn(gdx, bc(g0.z), __fmul2_rn(gdy, bc(g0.w)));      // -dG/ddelx
                    const float2 dGy = __ffma2_rn(gdy, bc(g1.x), __fmul2_rn(gdx, bc(g0.w)));      // -dG/ddely
                    v0 = hsum(__fmul2_rn(__fmul2_rn(dLdG, dGx), bc(nddelx)));
                    v1 = hsum(__fmul2_rn(__fmul2_rn(dLdG, dGy), bc(nddely)));
                    const float2 hg = __fmul2_rn(dLdG, bc(-0.5f));
                    const float2 X = __fmul2_rn(hg, gdx);
                    v2 = hsum(__fmul2_rn(X, bc(dx)));
                    v3 = hsum(__fmul2_rn(X, dy));
                    v4 = hsum(__fmul2_rn(__fmul2_rn(hg, gdy), dy));
                    v5 = hsum(__fmul2_rn(G, dL));
                }
                // Few contributing threads (splat edges): they commit their terms directly, 10 predicated RED
                // instructions instead of the ~45-instruction butterfly below.
                if (__popc(act_mask) <= direct_max) {
                    if (act) {
                        float* a = acc + (size_t)__ldg(point_list + range.x + pos) * ACC_STRIDE;
                        atomicAdd(a + 0, v0); atomicAdd(a + 1, v1); atomicAdd(a + 2, v2); atomicAdd(a + 3, v3); atomicAdd(a + 4, v4);
                        atomicAdd(a + 5, v5); atomicAdd(a + 6, v6); atomicAdd(a + 7, v7); atomicAdd(a + 8, v8); atomicAdd(a + 9, v9);
                    }
                    continue;
                }
                // 10 values -> 12 shuffles; final owner lanes given by (writer, slot)
                const float a0_ = bfly_pair(v0, v1, 16, b4), a1_ = bfly_pair(v2, v3, 16, b4), a2_ = bfly_pair(v4, v5, 16, b4);
                const float a3_ = bfly_pair(v6, v7, 16, b4), a4_ = bfly_pair(v8, v9, 16, b4);
                const float c0 = bfly_pair(a0_, a1_, 8, b3), c1 = bfly_pair(a2_, a3_, 8, b3), c2 = bfly_one(a4_, 8);
                const float e0 = bfly_pair(bfly_pair(c0, c1, 4, b2), bfly_one(c2, 4), 2, b1);
                const float f0 = bfly_one(e0, 1);
                if (writer && f0 != 0.0f) {
                    const uint32_t id = __ldg(point_list + range.x + pos);
                    atomicAdd(acc + (size_t)id * ACC_STRIDE + slot, f0);
                }
            }
        }
    }
}

// ---- one warp per CTA (round 2b) --------------------------------------------------------------------------------------
// The kernel above couples the four warp blocks of a tile with three CTA barriers per 256 list entries although, in the
// backward, they share nothing but the staged records: each block visits only the entries the forward flagged for it
// (contribution words), commits its own atomics, and finishes when ITS pixels are done. ncu: barrier waits were the
// second-largest stall of the kernel and the slowest block of a tile held the CTA's slot. Here a CTA is ONE warp and one
// 8x8 block: it walks its own contribution words back to front, packs the flagged entries 32 at a time (one per lane: list
// position -> Gaussian id -> three 16-byte cp.async of the record into the warp's own buffer), and blends a chunk while
// the next one is in flight. No CTA barrier, no s_bal, records that this block never touches are never staged, and the
// block scheduler backfills a finished block's slot at once. The visit arithmetic is the one above.
constexpr int WCHUNK = 32;
__global__ void __launch_bounds__(32, 32)
render_backward_warp_kernel(int W, int H, int gx,
                            const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order, const uint32_t* __restrict__ point_list,
                            const GaussRec* __restrict__ recs, const float* __restrict__ bg,
                            const float* __restrict__ alphas, const uint32_t* __restrict__ n_contrib,
                            const float* __restrict__ dL_dpix, const float* __restrict__ dL_ddepth_pix,
                            const float* __restrict__ dL_dalpha_pix, float* __restrict__ acc, int direct_max,
                            const uint32_t* __restrict__ contrib) {
    __shared__ __align__(16) GaussRec s_rec[2][WCHUNK];
    constexpr uint32_t FULL = 0xffffffffu;
    const int tile = (int)tile_order[blockIdx.x >> 2];      // longest lists first (binning.cu)
    const int wid = blockIdx.x & 3, lane = threadIdx.x;
    const int tile_x = (tile % gx) * TILE, tile_y = (tile / gx) * TILE;
    const int bx = (wid & 1) * 8, by = (wid >> 1) * 8;
    const int pxi = tile_x + bx + (lane & 7), pyi = tile_y + by + 2 * (lane >> 3);     // pixels (pxi, pyi) and (pxi, pyi+1)
    const bool in0 = pxi < W && pyi < H, in1 = pxi < W && pyi + 1 < H;
    const float pxf = (float)pxi;
    const float2 npy = make_float2(-(float)pyi, -(float)(pyi + 1));
    const uint32_t p0 = (uint32_t)pyi * W + pxi, p1 = p0 + (uint32_t)W;
    const size_t hw = (size_t)H * W;

    const uint2 range = ranges[tile];
    const int total = (int)(range.y - range.x);
    const uint32_t last0 = in0 ? n_contrib[p0] : 0u, last1 = in1 ? n_contrib[p1] : 0u;
    const uint32_t warp_max = min(__reduce_max_sync(FULL, max(last0, last1)), (uint32_t)total);
    if (warp_max == 0) return;
    const uint32_t* my_contrib = contrib + contrib_base(range.x, tile) * 4 + wid;      // word k covers entries [32k, 32k+32)
    const uint32_t* ids = point_list + range.x;

    const float2 T_final = make_float2(in0 ? 1.0f - alphas[p0] : 0.0f, in1 ? 1.0f - alphas[p1] : 0.0f);
    auto ld2 = [&](const float* base) { return make_float2(in0 ? base[p0] : 0.0f, in1 ? base[p1] : 0.0f); };
    const float2 dpix0 = ld2(dL_dpix), dpix1 = ld2(dL_dpix + hw), dpix2 = ld2(dL_dpix + 2 * hw);
    const float2 ddepth = ld2(dL_ddepth_pix), dalpha = ld2(dL_dalpha_pix);
    const float bg0 = __ldg(bg), bg1 = __ldg(bg + 1), bg2 = __ldg(bg + 2);
    const float2 nTf_bgdot = make_float2(-T_final.x * (bg0 * dpix0.x + bg1 * dpix1.x + bg2 * dpix2.x),
                                         -T_final.y * (bg0 * dpix0.y + bg1 * dpix1.y + bg2 * dpix2.y));
    float2 T = T_final, B0 = bc(0.f), B1 = bc(0.f), B2 = bc(0.f), Bd = bc(0.f), Ba = bc(0.f);
    const float nddelx = -0.5f * W, nddely = -0.5f * H;

    uint32_t lane_o;
    asm volatile("mov.u32 %0, %1;" : "=r"(lane_o) : "r"((uint32_t)lane));
    const bool b1 = lane_o & 2, b2 = lane_o & 4, b3 = lane_o & 8, b4 = lane_o & 16;
    const bool writer = !(lane_o & 1) && (!b1 || (!b2 && !b3));
    const int slot = b1 ? (8 + (b4 ? 1 : 0)) : ((b2 ? 4 : 0) + (b3 ? 2 : 0) + (b4 ? 1 : 0));

    // position of the (n+1)-th highest set bit of `bits` (n < popc(bits)): largest b with popc(bits >> b) > n
    auto nth_from_top = [](uint32_t bits, int n) {
        int b = 0;
#pragma unroll
        for (int s = 16; s; s >>= 1) if (__popc(bits >> (b + s)) > n) b += s;
        return b;
    };
    // walk of the contribution words, back to front: lane l of `words` holds word kbase - l
    int k = (int)((warp_max - 1) >> 5), kbase = k;
    auto load_words = [&]() { const int kk = kbase - lane; return kk >= 0 ? __ldg(my_contrib + (size_t)kk * 4) : 0u; };
    uint32_t words = load_words();
    uint32_t bits = __shfl_sync(FULL, words, 0);
    if (((warp_max - 1) & 31u) != 31u) bits &= (2u << ((warp_max - 1) & 31u)) - 1u;       // entries at or beyond warp_max
    // next chunk: up to 32 flagged positions, highest first, one per lane; returns how many
    auto fill = [&](uint32_t& mypos) {
        int have = 0;
        mypos = 0u;
        while (have < WCHUNK) {
            if (bits == 0u) {
                if (k == 0) break;
                --k;
                if (kbase - k == 32) { kbase = k; words = load_words(); }
                bits = __shfl_sync(FULL, words, kbase - k);
                continue;
            }
            const int n = __popc(bits), take = min(n, WCHUNK - have), mine = lane - have;
            if (mine >= 0 && mine < take) mypos = (uint32_t)(k * 32 + nth_from_top(bits, mine));
            if (take == n) bits = 0u;
            else bits &= (1u << nth_from_top(bits, take - 1)) - 1u;       // drop the `take` highest
            have += take;
        }
        return have;
    };
    auto stage = [&](int buf, int have, uint32_t mypos, uint32_t& myid) {
        if (lane < have) {
            myid = __ldg(ids + mypos);
            const char* src = reinterpret_cast<const char*>(recs + myid);
            char* dst = reinterpret_cast<char*>(&s_rec[buf][lane]);
            cp_async16(dst, src); cp_async16(dst + 16, src + 16); cp_async16(dst + 32, src + 32);
        }
        cp_async_commit();
    };

    uint32_t pos_c, id_c = 0u, pos_n, id_n = 0u;
    int have_c = fill(pos_c), buf = 0;
    stage(0, have_c, pos_c, id_c);
    while (have_c > 0) {
        const int have_n = fill(pos_n);
        stage(buf ^ 1, have_n, pos_n, id_n);        // (an empty group when nothing is left)
        cp_async_wait<1>();
        __syncwarp();
        uint32_t buf_s;
        asm volatile("mov.u32 %0, %1;" : "=r"(buf_s) : "r"((uint32_t)__cvta_generic_to_shared(s_rec[buf])));
        for (int j = 0; j < have_c; ++j) {
            const uint32_t pos = __shfl_sync(FULL, pos_c, j);
            const uint32_t rec_s = buf_s + (uint32_t)j * (uint32_t)sizeof(GaussRec);
            const float4 g0 = lds128(rec_s);
            const float4 g1 = lds128(rec_s + 16);
            float dx; float2 dy;
            const float2 power = pair_power(g0.x, g0.y, g0.z, g0.w, g1.x, pxf, npy, dx, dy);
            bool a0 = (pos < last0) && !(power.x > 0.0f) && !(power.x < g1.w);
            bool a1 = (pos < last1) && !(power.y > 0.0f) && !(power.y < g1.w);
            float2 G = bc(0.f), alpha = bc(0.f);
            if (a0 || a1) {
                G = expf2_nonpos(power);                          // the forward's expf, bit for bit
                const float2 ga = __fmul2_rn(bc(g1.y), G);
                alpha = make_float2(fminf(0.99f, ga.x), fminf(0.99f, ga.y));
                a0 = a0 && !(alpha.x < 1.0f / 255.0f);
                a1 = a1 && !(alpha.y < 1.0f / 255.0f);
            }
            const bool act = a0 || a1;
            const uint32_t act_mask = __ballot_sync(FULL, act);
            if (act_mask == 0) continue;

            float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f, v4 = 0.f, v5 = 0.f, v6 = 0.f, v7 = 0.f, v8 = 0.f, v9 = 0.f;
            if (act) {
                if (!a0) { G.x = 0.f; alpha.x = 0.f; }
                if (!a1) { G.y = 0.f; alpha.y = 0.f; }
                const float4 g2 = lds128(rec_s + 32);
                const float2 oma = __ffma2_rn(alpha, bc(-1.0f), bc(1.0f));
                const float2 inv = make_float2(a0 ? rcp_approx(oma.x) : 1.0f, a1 ? rcp_approx(oma.y) : 1.0f);
                T = __fmul2_rn(T, inv);
                const float2 w = __fmul2_rn(alpha, T);                  // dchannel_dcolor
                const float2 d0 = __ffma2_rn(B0, bc(-1.0f), bc(g2.x)), d1 = __ffma2_rn(B1, bc(-1.0f), bc(g2.y));
                const float2 d2 = __ffma2_rn(B2, bc(-1.0f), bc(g2.z)), dd = __ffma2_rn(Bd, bc(-1.0f), bc(g1.z));
                const float2 da = __ffma2_rn(Ba, bc(-1.0f), bc(1.0f));
                float2 dL = __fmul2_rn(d0, dpix0);
                dL = __ffma2_rn(d1, dpix1, dL); dL = __ffma2_rn(d2, dpix2, dL);
                dL = __ffma2_rn(dd, ddepth, dL); dL = __ffma2_rn(da, dalpha, dL);
                B0 = __ffma2_rn(alpha, d0, B0); B1 = __ffma2_rn(alpha, d1, B1); B2 = __ffma2_rn(alpha, d2, B2);
                Bd = __ffma2_rn(alpha, dd, Bd); Ba = __ffma2_rn(alpha, da, Ba);
                v6 = hsum(__fmul2_rn(w, dpix0)); v7 = hsum(__fmul2_rn(w, dpix1)); v8 = hsum(__fmul2_rn(w, dpix2));
                v9 = hsum(__fmul2_rn(w, ddepth));
                dL = __ffma2_rn(dL, T, __fmul2_rn(nTf_bgdot, inv));     // dL/dalpha
                const float2 dLdG = __fmul2_rn(bc(g1.y), dL);
                const float2 gdx = __fmul2_rn(G, bc(dx)), gdy = __fmul2_rn(G, dy);
                const float2 dGx = __ffma2_rn(gdx, bc(g0.z), __fmul2_rn(gdy, bc(g0.w)));      // -dG/ddelx
                const float2 dGy = __ffma2_rn(gdy, bc(g1.x), __fmul2_rn(gdx, bc(g0.w)));      // -dG/ddely
                v0 = hsum(__fmul2_rn(__fmul2_rn(dLdG, dGx), bc(nddelx)));
                v1 = hsum(__fmul2_rn(__fmul2_rn(dLdG, dGy), bc(nddely)));
                const float2 hg = __fmul2_rn(dLdG, bc(-0.5f));
                const float2 X = __fmul2_rn(hg, gdx);
                v2 = hsum(__fmul2_rn(X, bc(dx)));
                v3 = hsum(__fmul2_rn(X, dy));
                v4 = hsum(__fmul2_rn(__fmul2_rn(hg, gdy), dy));
                v5 = hsum(__fmul2_rn(G, dL));
            }
            float* const a = acc + (size_t)__shfl_sync(FULL, id_c, j) * ACC_STRIDE;
            if (__popc(act_mask) <= direct_max) {
                if (act) {      // three vector reductions on the 48-byte slot instead of ten scalar ones
                    red_add_v4(a, v0, v1, v2, v3); red_add_v4(a + 4, v4, v5, v6, v7); red_add_v2(a + 8, v8, v9);
                }
                continue;
            }
            const float a0_ = bfly_pair(v0, v1, 16, b4), a1_ = bfly_pair(v2, v3, 16, b4), a2_ = bfly_pair(v4, v5, 16, b4);
            const float a3_ = bfly_pair(v6, v7, 16, b4), a4_ = bfly_pair(v8, v9, 16, b4);
            const float c0 = bfly_pair(a0_, a1_, 8, b3), c1 = bfly_pair(a2_, a3_, 8, b3), c2 = bfly_one(a4_, 8);
            const float e0 = bfly_pair(bfly_pair(c0, c1, 4, b2), bfly_one(c2, 4), 2, b1);
            const float f0 = bfly_one(e0, 1);
            if (writer && f0 != 0.0f) atomicAdd(a + slot, f0);
        }
        __syncwarp();                       // every lane is done with this buffer before the next stage overwrites it
        have_c = have_n; pos_c = pos_n; id_c = id_n; buf ^= 1;
    }
}

// Visits with at most this many contributing pixels skip the warp reduction (tunable: GSR_DIRECT_MAX).
// Round 1, scalar atomics (the four-warp kernel): C2 1.244 ms (0) -> 1.192 ms (8); the 18-term surfel variant only lost
// (default 0). Round 2b, the one-warp kernels commit a direct visit with THREE vector reductions (red.global.add.v4.f32,
// SASS REDG.E.ADD.F32x4: the 48-byte slot is 16-byte aligned) instead of ten scalar ones, which moves the break-even
// far up: C2 0.647 ms (8, scalar) -> 0.626 (8) -> 0.605 (16) -> 0.590 (24) -> **0.586 (28)** -> 0.587 (31) -> 0.742 (32:
// full-warp visits are the common case and 96 vector atomics each is too many); 2DGS, five reductions on the 80-byte
// slot: C4 1.239 ms (0) -> 1.116 (8) -> 1.065 (16) -> **1.058 (20)** -> 1.081 (24) -> 1.299 (31).
int blend_direct_max(int dflt) {
    static int env = -2;
    if (env == -2) { const char* e = getenv("GSR_DIRECT_MAX"); env = e ? atoi(e) : -1; if (env > 32) env = 32; }
    return env >= 0 ? env : dflt;
}

cudaError_t launch_render_backward(int W, int H, int gx, int gy, const GeomState& g, const ImageState& img,
                                   const BinningState& bin, const float* bg, const float* out_alpha,
                                   const float* dL_dpix, const float* dL_ddepth, const float* dL_dalpha,
                                   float* acc, cudaStream_t st) {
    static int per_warp = -1;           // GSR_BWD_WARP=0: the four-warp CTA with barriers (round 1/2a)
    if (per_warp < 0) { const char* e = getenv("GSR_BWD_WARP"); per_warp = e ? (atoi(e) ? 1 : 0) : 1; }
    if (per_warp)
        render_backward_warp_kernel<<<gx * gy * 4, 32, 0, st>>>(W, H, gx, img.ranges, img.tile_order, bin.point_list, g.rec, bg, out_alpha,
                                                                img.n_contrib, dL_dpix, dL_ddepth, dL_dalpha, acc, blend_direct_max(28),
                                                                reinterpret_cast<const uint32_t*>(bin.contrib));
    else
    render_backward_kernel<<<gx * gy, BLEND_THREADS, 0, st>>>(W, H, gx, img.ranges, img.tile_order, bin.point_list, g.rec, bg, out_alpha,
                                                    img.n_contrib, dL_dpix, dL_ddepth, dL_dalpha, acc, blend_direct_max(8),
                                                              reinterpret_cast<const uint32_t*>(bin.contrib));
    return cudaGetLastError();
}

}  // namespace gsr
