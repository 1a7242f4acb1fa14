// surfel_render_bwd.cu -- back-to-front re-traversal of each tile's list for the 2DGS surfel path.
//
// Replaces renderCUDA (reference diff-surfel-rasterization/cuda_rasterizer/backward.cu:145-446).
// Same structure as render_bwd.cu: traversal starts at the tile's largest n_contrib, records are
// staged with cp.async, a warp visits only the entries the forward recorded as contributing to its 8x4 pixel block,
// and the up-to-16 per-pair gradient terms (colour 3, normal 3, transMat 9 or mean2D 2 + transMat 3,
// opacity 1) are reduced over the warp with a multi-value butterfly (20 shuffles for 18 values) and
// committed as one red.global.add.f32 per term per (warp, surfel) into an 80-byte accumulator slot.
#include "surfel.cuh"
#include "blend.cuh"

namespace gsr {

int blend_direct_max(int dflt);     // render_bwd.cu
constexpr float kNearB = 0.2f, kFarB = 100.0f;
constexpr int SURFEL_SMEM_B = 2 * BATCH * (int)sizeof(SurfelRec) + 8 * 8 * 4 + 8 * 4;

// One butterfly stage over mask m: N values -> (N+1)/2 values. Pair (2i, 2i+1): lanes with the bit
// clear end up with value 2i summed over {lane, lane^m}, lanes with the bit set with value 2i+1.
template <int N>
__device__ __forceinline__ void bfly_stage(const float (&in)[N], float (&out)[(N + 1) / 2], int m, bool hi) {
#pragma unroll
    for (int i = 0; i < N / 2; ++i) {
        const float send = hi ? in[2 * i] : in[2 * i + 1];
        const float keep = hi ? in[2 * i + 1] : in[2 * i];
        out[i] = keep + __shfl_xor_sync(0xffffffffu, send, m);
    }
    if (N & 1) out[N / 2] = in[N - 1] + __shfl_xor_sync(0xffffffffu, in[N - 1], m);
}

// Which of the 18 input values ends up in this lane after the five stages, and whether this lane
// is the designated writer for it (values that ride an odd slot are duplicated in both partners).
__device__ __forceinline__ void bfly_owner18(int lane, int& slot, bool& writer) {
    const int sizes[5] = { 18, 9, 5, 3, 2 };
    const int masks[5] = { 16, 8, 4, 2, 1 };
    int idx = 0; writer = true;
#pragma unroll
    for (int s = 4; s >= 0; --s) {
        const bool hi = (lane & masks[s]) != 0;
        if (idx < sizes[s] / 2) idx = 2 * idx + (hi ? 1 : 0);
        else { idx = sizes[s] - 1; if (hi) writer = false; }
    }
    slot = idx;
}

__global__ void __launch_bounds__(256, 3)      // 80 registers, 48 B of spill: 3 CTAs/SM. Measured at C4: 2 CTAs/SM (105 registers, no spill) 1.63 ms, 3 CTAs/SM 1.50 ms, 4 CTAs/SM 2.60 ms (round 1)
surfel_render_backward_kernel(int W, int H, int gx,
                              const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order, const uint32_t* __restrict__ point_list,
                              const SurfelRec* __restrict__ recs, const float* __restrict__ bg,
                              const uint32_t* __restrict__ n_contrib, const uint32_t* __restrict__ median_contrib,
                              const float* __restrict__ final_Ts, const float* __restrict__ in_M1,
                              const float* __restrict__ in_M2,
                              const float* __restrict__ dL_dpixels, const float* __restrict__ dL_dothers,
                              float* __restrict__ acc, int direct_max, const uint32_t* __restrict__ contrib) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SurfelRec (*s_rec)[BATCH] = reinterpret_cast<SurfelRec (*)[BATCH]>(smem_raw);
    uint32_t (*s_bal)[8] = reinterpret_cast<uint32_t (*)[8]>(smem_raw + 2 * BATCH * sizeof(SurfelRec));
    uint32_t* s_max = reinterpret_cast<uint32_t*>(smem_raw + 2 * BATCH * sizeof(SurfelRec) + 8 * 8 * 4);

    const int tile = (int)tile_order[blockIdx.x];      // longest lists first (binning.cu)
    const int tile_x = (tile % gx) * TILE, tile_y = (tile / gx) * TILE;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int bx = (wid & 1) * 8, by = (wid >> 1) * 4;
    const uint32_t pxi = tile_x + bx + (lane & 7), pyi = tile_y + by + (lane >> 3);
    const bool inside = pxi < (uint32_t)W && pyi < (uint32_t)H;
    const float pxf = (float)pxi, pyf = (float)pyi;
    const uint32_t pix_id = pyi * W + pxi;
    const size_t hw = (size_t)H * W;

    const uint2 range = ranges[tile];
    const int total = (int)(range.y - range.x);

    const uint32_t last_contributor = inside ? n_contrib[pix_id] : 0u;
    const uint32_t median_contributor = inside ? median_contrib[pix_id] : 0u;
    const float T_final = inside ? final_Ts[pix_id] : 0.0f;
    const float final_D = inside ? in_M1[pix_id] : 0.0f;
    const float final_D2 = inside ? in_M2[pix_id] : 0.0f;
    const float final_A = 1 - T_final;
    float dpix0 = 0.f, dpix1 = 0.f, dpix2 = 0.f;
    float dL_ddepth = 0.f, dL_daccum = 0.f, dL_dreg = 0.f, dn0 = 0.f, dn1 = 0.f, dn2 = 0.f, dL_dmedian_depth = 0.f;
    if (inside) {
        dpix0 = dL_dpixels[pix_id]; dpix1 = dL_dpixels[hw + pix_id]; dpix2 = dL_dpixels[2 * hw + pix_id];
        dL_ddepth = dL_dothers[pix_id];                  // DEPTH_OFFSET
        dL_daccum = dL_dothers[hw + pix_id];             // ALPHA_OFFSET
        dn0 = dL_dothers[2 * hw + pix_id]; dn1 = dL_dothers[3 * hw + pix_id]; dn2 = dL_dothers[4 * hw + pix_id];
        dL_dmedian_depth = dL_dothers[5 * hw + pix_id];  // MIDDEPTH_OFFSET
        dL_dreg = dL_dothers[6 * hw + pix_id];           // DISTORTION_OFFSET
    }
    float bg_dot_dpixel = 0;
    bg_dot_dpixel += __ldg(bg) * dpix0; bg_dot_dpixel += __ldg(bg + 1) * dpix1; bg_dot_dpixel += __ldg(bg + 2) * dpix2;

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

    // Running state per pixel. The reference carries (last_alpha, last_value, accum_rec) per blended quantity and forms
    // accum_rec = last_alpha * last_value + (1 - last_alpha) * accum_rec at the NEXT contributor (backward.cu:334-372);
    // B below is that value formed one step earlier, B <- B + alpha * (value - B): one register per quantity instead of
    // two plus the shared last_alpha (render_bwd.cu does the same for 3DGS).
    float T = T_final;
    float B0 = 0.f, B1 = 0.f, B2 = 0.f, Bd = 0.f, Ba = 0.f, Bn0 = 0.f, Bn1 = 0.f, Bn2 = 0.f;
    float last_dL_dT = 0.f;

    int slot; bool writer;
    bfly_owner18(lane, slot, writer);
    const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4, h2 = lane & 2, h1 = lane & 1;

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

    const uint32_t* my_contrib = contrib + contrib_base(range.x, tile) * 8;
    const int r_start = (int)((max_contrib - 1) / BATCH);
    issue(r_start);
    for (int r = r_start; r >= 0; --r) {
        __syncthreads();
        if (r > 0) { issue(r - 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
        __syncthreads();

        const SurfelRec* buf = s_rec[r & 1];
        // which entries to visit: the forward recorded, per warp block, the entries that blended into some pixel of the
        // block (surfel_render_fwd.cu); everything else contributes nothing here either
        if (tid < 64) {
            const int w = tid >> 3, v = tid & 7;
            s_bal[w][v] = (r * BATCH + v * 32 < total) ? __ldg(my_contrib + (size_t)(r * 8 + v) * 8 + w) : 0u;
        }
        __syncthreads();

        if ((uint32_t)(r * BATCH) >= warp_max) continue;
        for (int v = 7; v >= 0; --v) {
            uint32_t bits = s_bal[wid][v];
            while (bits) {
                const int jb = 31 - __clz(bits);
                bits &= ~(1u << jb);
                const int j = v * 32 + jb;
                const uint32_t pos = (uint32_t)(r * BATCH + j);      // 0-based list position == reference `contributor`
                if (pos >= warp_max) continue;

                const float4* q = reinterpret_cast<const float4*>(&buf[j]);
                const float4 q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3];
                const float3 Tu = make_float3(q0.z, q0.w, q1.x), Tv = make_float3(q1.y, q1.z, q1.w);
                const float3 Tw = make_float3(q2.x, q2.y, q2.z);
                // backward.cu:289-322
                const float3 k = make_float3(pxf * Tw.x - Tu.x, pxf * Tw.y - Tu.y, pxf * Tw.z - Tu.z);
                const float3 l = make_float3(pyf * Tw.x - Tv.x, pyf * Tw.y - Tv.y, pyf * Tw.z - Tv.z);
                const float3 p = make_float3(k.y * l.z - k.z * l.y, k.z * l.x - k.x * l.z, k.x * l.y - k.y * l.x);
                bool act = (pos < last_contributor) && !(p.z == 0.0f);
                float sx = 0.f, sy = 0.f, rho3d = 0.f, rho2d = 0.f, c_d = 0.f, G = 0.f, alpha = 0.f;
                const float dx = q0.x - pxf, dy = q0.y - pyf;
                if (act) {
                    sx = p.x / p.z; sy = p.y / p.z;
                    rho3d = (sx * sx + sy * sy);
                    rho2d = 2.0f * (dx * dx + dy * dy);
                    const float rho = fminf(rho3d, rho2d);
                    c_d = (sx * Tw.x + sy * Tw.y) + Tw.z;
                    const float power = -0.5f * rho;
                    act = !(c_d < kNearB) && !(power > 0.0f);
                    if (act) {
                        G = expf(power);
                        alpha = fminf(0.99f, q3.z * G);
                        act = !(alpha < 1.0f / 255.0f);
                    }
                }
                const uint32_t act_mask = __ballot_sync(0xffffffffu, act);
                if (act_mask == 0) continue;

                float val[18];
#pragma unroll
                for (int i = 0; i < 18; ++i) val[i] = 0.0f;
                if (act) {
                    const float4 q4 = q[4];
                    const float nrm0 = q2.w, nrm1 = q3.x, nrm2 = q3.y, opa = q3.z;
                    const float col0 = q3.w, col1 = q4.x, col2 = q4.y;
                    // Gradient-only quotients use the bare MUFU.RCP (rcp.approx.ftz, relative error 2^-23 against a tolerance of
                    // 1e-4; every operand is a normal number: 1 - alpha in [0.01, 1], c_d >= 0.2). The two divisions above that
                    // form sx, sy stay IEEE: they decide branches and must reproduce the forward bit for bit.
                    const float inv_1ma = rcp_approx(1.f - alpha);
                    T = T * inv_1ma;
                    const float dchannel_dcolor = alpha * T;
                    float dL_dalpha = 0.0f;
                    const float e0 = col0 - B0, e1 = col1 - B1, e2 = col2 - B2;
                    dL_dalpha += e0 * dpix0; B0 += alpha * e0;
                    val[SACC_R] = dchannel_dcolor * dpix0;
                    dL_dalpha += e1 * dpix1; B1 += alpha * e1;
                    val[SACC_G] = dchannel_dcolor * dpix1;
                    dL_dalpha += e2 * dpix2; B2 += alpha * e2;
                    val[SACC_B] = dchannel_dcolor * dpix2;

                    float dL_dz = 0.0f;
                    float dL_dweight = 0;
                    const float inv_cd = rcp_approx(c_d);
                    const float m_d = kFarB / (kFarB - kNearB) * (1 - kNearB * inv_cd);
                    const float dmd_dd = ((kFarB * kNearB) / (kFarB - kNearB)) * (inv_cd * inv_cd);
                    if (pos == median_contributor - 1u) dL_dz += dL_dmedian_depth;
                    dL_dweight += (final_D2 + m_d * m_d * final_A - 2 * m_d * final_D) * dL_dreg;
                    dL_dalpha += dL_dweight - last_dL_dT;
                    last_dL_dT = dL_dweight * alpha + (1 - alpha) * last_dL_dT;
                    const float dL_dmd = 2.0f * (T * alpha) * (m_d * final_A - final_D) * dL_dreg;
                    dL_dz += dL_dmd * dmd_dd;

                    const float ed = c_d - Bd, ea = 1.0f - Ba;
                    dL_dalpha += ed * dL_ddepth; Bd += alpha * ed;
                    dL_dalpha += ea * dL_daccum; Ba += alpha * ea;

                    const float n0 = nrm0 - Bn0, n1 = nrm1 - Bn1, n2 = nrm2 - Bn2;
                    dL_dalpha += n0 * dn0; Bn0 += alpha * n0;
                    val[SACC_NX] = alpha * T * dn0;
                    dL_dalpha += n1 * dn1; Bn1 += alpha * n1;
                    val[SACC_NY] = alpha * T * dn1;
                    dL_dalpha += n2 * dn2; Bn2 += alpha * n2;
                    val[SACC_NZ] = alpha * T * dn2;

                    dL_dalpha *= T;
                    dL_dalpha += (-T_final * inv_1ma) * bg_dot_dpixel;

                    const float dL_dG = opa * dL_dalpha;
                    dL_dz += alpha * T * dL_ddepth;

                    if (rho3d <= rho2d) {
                        // backward.cu:399-429
                        const float dLdsx = dL_dG * -G * sx + dL_dz * Tw.x;
                        const float dLdsy = dL_dG * -G * sy + dL_dz * Tw.y;
                        const float inv_pz = rcp_approx(p.z);
                        const float dsx_pz = dLdsx * inv_pz;
                        const float dsy_pz = dLdsy * inv_pz;
                        const float3 dL_dp = make_float3(dsx_pz, dsy_pz, -(dsx_pz * sx + dsy_pz * sy));
                        const float3 dL_dk = make_float3(l.y * dL_dp.z - l.z * dL_dp.y, l.z * dL_dp.x - l.x * dL_dp.z,
                                                         l.x * dL_dp.y - l.y * dL_dp.x);
                        const float3 dL_dl = make_float3(dL_dp.y * k.z - dL_dp.z * k.y, dL_dp.z * k.x - dL_dp.x * k.z,
                                                         dL_dp.x * k.y - dL_dp.y * k.x);
                        val[0] = -dL_dk.x; val[1] = -dL_dk.y; val[2] = -dL_dk.z;
                        val[3] = -dL_dl.x; val[4] = -dL_dl.y; val[5] = -dL_dl.z;
                        val[6] = pxf * dL_dk.x + pyf * dL_dl.x + dL_dz * sx;
                        val[7] = pxf * dL_dk.y + pyf * dL_dl.y + dL_dz * sy;
                        val[8] = pxf * dL_dk.z + pyf * dL_dl.z + dL_dz * 1.0f;
                    } else {
                        // backward.cu:430-440
                        const float dG_ddelx = -G * 2.0f * dx;
                        const float dG_ddely = -G * 2.0f * dy;
                        val[SACC_MX] = dL_dG * dG_ddelx;
                        val[SACC_MY] = dL_dG * dG_ddely;
                        val[6] = sx * dL_dz;
                        val[7] = sy * dL_dz;
                        val[8] = dL_dz;
                    }
                    val[SACC_OP] = G * dL_dalpha;
                }
                if (__popc(act_mask) <= direct_max) {       // splat edges: direct predicated REDs, no warp reduction
                    if (act) {
                        float* a = acc + (size_t)__ldg(point_list + range.x + pos) * SACC_STRIDE;
#pragma unroll
                        for (int i = 0; i < 18; ++i) if (val[i] != 0.0f) atomicAdd(a + i, val[i]);
                    }
                    continue;
                }
                float s1[9], s2[5], s3[3], s4[2], s5[1];
                bfly_stage<18>(val, s1, 16, h16);
                bfly_stage<9>(s1, s2, 8, h8);
                bfly_stage<5>(s2, s3, 4, h4);
                bfly_stage<3>(s3, s4, 2, h2);
                bfly_stage<2>(s4, s5, 1, h1);
                if (writer && s5[0] != 0.0f) {
                    const uint32_t id = __ldg(point_list + range.x + pos);
                    atomicAdd(acc + (size_t)id * SACC_STRIDE + slot, s5[0]);
                }
            }
        }
    }
}

// ---- one warp per CTA (round 2b; see render_bwd.cu, render_backward_warp_kernel) ---------------------------------------
// A CTA is one warp and one 8x4 pixel block: it walks its own contribution words back to front, packs the flagged
// entries 32 at a time (one per lane: list position -> surfel id -> six 16-byte cp.async of the record into the warp's
// own buffer) and blends a chunk while the next one is in flight. No CTA barriers; the visit arithmetic is the one above.
constexpr int SWCHUNK = 32;
__global__ void __launch_bounds__(32, 24)      // 80 registers, 8 B of spill. Measured at C4: 20 CTAs/SM (92 registers, no spill) 1.128 ms, 24 -> 1.060 ms, 28 (72 registers, 40 B spill) 1.102 ms
surfel_render_backward_warp_kernel(int W, int H, int gx,
                                   const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order, const uint32_t* __restrict__ point_list,
                                   const SurfelRec* __restrict__ recs, const float* __restrict__ bg,
                                   const uint32_t* __restrict__ n_contrib, const uint32_t* __restrict__ median_contrib,
                                   const float* __restrict__ final_Ts, const float* __restrict__ in_M1,
                                   const float* __restrict__ in_M2,
                                   const float* __restrict__ dL_dpixels, const float* __restrict__ dL_dothers,
                                   float* __restrict__ acc, int direct_max, const uint32_t* __restrict__ contrib) {
    __shared__ __align__(16) SurfelRec s_rec[2][SWCHUNK];
    constexpr uint32_t FULL = 0xffffffffu;
    const int tile = (int)tile_order[blockIdx.x >> 3];      // longest lists first (binning.cu)
    const int wid = blockIdx.x & 7, lane = threadIdx.x;
    const int tile_x = (tile % gx) * TILE, tile_y = (tile / gx) * TILE;
    const int bx = (wid & 1) * 8, by = (wid >> 1) * 4;
    const uint32_t pxi = tile_x + bx + (lane & 7), pyi = tile_y + by + (lane >> 3);
    const bool inside = pxi < (uint32_t)W && pyi < (uint32_t)H;
    const float pxf = (float)pxi, pyf = (float)pyi;
    const uint32_t pix_id = pyi * W + pxi;
    const size_t hw = (size_t)H * W;

    const uint2 range = ranges[tile];
    const int total = (int)(range.y - range.x);
    const uint32_t last_contributor = inside ? n_contrib[pix_id] : 0u;
    const uint32_t warp_max = min(__reduce_max_sync(FULL, last_contributor), (uint32_t)total);
    if (warp_max == 0) return;
    const uint32_t* my_contrib = contrib + contrib_base(range.x, tile) * 8 + wid;      // word k covers entries [32k, 32k+32)
    const uint32_t* ids = point_list + range.x;

    const uint32_t median_contributor = inside ? median_contrib[pix_id] : 0u;
    const float T_final = inside ? final_Ts[pix_id] : 0.0f;
    const float final_D = inside ? in_M1[pix_id] : 0.0f;
    const float final_D2 = inside ? in_M2[pix_id] : 0.0f;
    const float final_A = 1 - T_final;
    float dpix0 = 0.f, dpix1 = 0.f, dpix2 = 0.f;
    float dL_ddepth = 0.f, dL_daccum = 0.f, dL_dreg = 0.f, dn0 = 0.f, dn1 = 0.f, dn2 = 0.f, dL_dmedian_depth = 0.f;
    if (inside) {
        dpix0 = dL_dpixels[pix_id]; dpix1 = dL_dpixels[hw + pix_id]; dpix2 = dL_dpixels[2 * hw + pix_id];
        dL_ddepth = dL_dothers[pix_id];                  // DEPTH_OFFSET
        dL_daccum = dL_dothers[hw + pix_id];             // ALPHA_OFFSET
        dn0 = dL_dothers[2 * hw + pix_id]; dn1 = dL_dothers[3 * hw + pix_id]; dn2 = dL_dothers[4 * hw + pix_id];
        dL_dmedian_depth = dL_dothers[5 * hw + pix_id];  // MIDDEPTH_OFFSET
        dL_dreg = dL_dothers[6 * hw + pix_id];           // DISTORTION_OFFSET
    }
    float bg_dot_dpixel = 0;
    bg_dot_dpixel += __ldg(bg) * dpix0; bg_dot_dpixel += __ldg(bg + 1) * dpix1; bg_dot_dpixel += __ldg(bg + 2) * dpix2;

    float T = T_final;
    float B0 = 0.f, B1 = 0.f, B2 = 0.f, Bd = 0.f, Ba = 0.f, Bn0 = 0.f, Bn1 = 0.f, Bn2 = 0.f;
    float last_dL_dT = 0.f;

    int slot; bool writer;
    bfly_owner18(lane, slot, writer);
    const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4, h2 = lane & 2, h1 = lane & 1;

    auto nth_from_top = [](uint32_t bits, int n) {      // position of the (n+1)-th highest set bit
        int b = 0;
#pragma unroll
        for (int s = 16; s; s >>= 1) if (__popc(bits >> (b + s)) > n) b += s;
        return b;
    };
    int kw = (int)((warp_max - 1) >> 5), kbase = kw;      // current contribution word
    auto load_words = [&]() { const int kk = kbase - lane; return kk >= 0 ? __ldg(my_contrib + (size_t)kk * 8) : 0u; };
    uint32_t words = load_words();
    uint32_t bits = __shfl_sync(FULL, words, 0);
    if (((warp_max - 1) & 31u) != 31u) bits &= (2u << ((warp_max - 1) & 31u)) - 1u;
    auto fill = [&](uint32_t& mypos) {
        int have = 0;
        mypos = 0u;
        while (have < SWCHUNK) {
            if (bits == 0u) {
                if (kw == 0) break;
                --kw;
                if (kbase - kw == 32) { kbase = kw; words = load_words(); }
                bits = __shfl_sync(FULL, words, kbase - kw);
                continue;
            }
            const int n = __popc(bits), take = min(n, SWCHUNK - have), mine = lane - have;
            if (mine >= 0 && mine < take) mypos = (uint32_t)(kw * 32 + nth_from_top(bits, mine));
            if (take == n) bits = 0u;
            else bits &= (1u << nth_from_top(bits, take - 1)) - 1u;
            have += take;
        }
        return have;
    };
    auto stage = [&](int b, int have, uint32_t mypos, uint32_t& myid) {
        if (lane < have) {
            myid = __ldg(ids + mypos);
            const char* src = reinterpret_cast<const char*>(recs + myid);
            char* dst = reinterpret_cast<char*>(&s_rec[b][lane]);
#pragma unroll
            for (int q = 0; q < 6; ++q) cp_async16(dst + 16 * q, src + 16 * q);
        }
        cp_async_commit();
    };

    uint32_t pos_c, id_c = 0u, pos_n, id_n = 0u;
    int have_c = fill(pos_c), buf = 0;
    stage(0, have_c, pos_c, id_c);
    while (have_c > 0) {
        const int have_n = fill(pos_n);
        stage(buf ^ 1, have_n, pos_n, id_n);
        cp_async_wait<1>();
        __syncwarp();
        for (int j = 0; j < have_c; ++j) {
            const uint32_t pos = __shfl_sync(FULL, pos_c, j);      // 0-based list position == reference `contributor`
            const float4* q = reinterpret_cast<const float4*>(&s_rec[buf][j]);
            const float4 q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3];
            const float3 Tu = make_float3(q0.z, q0.w, q1.x), Tv = make_float3(q1.y, q1.z, q1.w);
            const float3 Tw = make_float3(q2.x, q2.y, q2.z);
            // backward.cu:289-322
            const float3 k = make_float3(pxf * Tw.x - Tu.x, pxf * Tw.y - Tu.y, pxf * Tw.z - Tu.z);
            const float3 l = make_float3(pyf * Tw.x - Tv.x, pyf * Tw.y - Tv.y, pyf * Tw.z - Tv.z);
            const float3 p = make_float3(k.y * l.z - k.z * l.y, k.z * l.x - k.x * l.z, k.x * l.y - k.y * l.x);
            bool act = (pos < last_contributor) && !(p.z == 0.0f);
            float sx = 0.f, sy = 0.f, rho3d = 0.f, rho2d = 0.f, c_d = 0.f, G = 0.f, alpha = 0.f;
            const float dx = q0.x - pxf, dy = q0.y - pyf;
            if (act) {
                sx = p.x / p.z; sy = p.y / p.z;
                rho3d = (sx * sx + sy * sy);
                rho2d = 2.0f * (dx * dx + dy * dy);
                const float rho = fminf(rho3d, rho2d);
                c_d = (sx * Tw.x + sy * Tw.y) + Tw.z;
                const float power = -0.5f * rho;
                act = !(c_d < kNearB) && !(power > 0.0f);
                if (act) {
                    G = expf(power);
                    alpha = fminf(0.99f, q3.z * G);
                    act = !(alpha < 1.0f / 255.0f);
                }
            }
            const uint32_t act_mask = __ballot_sync(0xffffffffu, act);
            if (act_mask == 0) continue;

            float val[18];
#pragma unroll
            for (int i = 0; i < 18; ++i) val[i] = 0.0f;
            if (act) {
                const float4 q4 = q[4];
                const float nrm0 = q2.w, nrm1 = q3.x, nrm2 = q3.y, opa = q3.z;
                const float col0 = q3.w, col1 = q4.x, col2 = q4.y;
                // Gradient-only quotients use the bare MUFU.RCP (rcp.approx.ftz, relative error 2^-23 against a tolerance of
                // 1e-4; every operand is a normal number: 1 - alpha in [0.01, 1], c_d >= 0.2). The two divisions above that
                // form sx, sy stay IEEE: they decide branches and must reproduce the forward bit for bit.
                const float inv_1ma = rcp_approx(1.f - alpha);
                T = T * inv_1ma;
                const float dchannel_dcolor = alpha * T;
                float dL_dalpha = 0.0f;
                const float e0 = col0 - B0, e1 = col1 - B1, e2 = col2 - B2;
                dL_dalpha += e0 * dpix0; B0 += alpha * e0;
                val[SACC_R] = dchannel_dcolor * dpix0;
                dL_dalpha += e1 * dpix1; B1 += alpha * e1;
                val[SACC_G] = dchannel_dcolor * dpix1;
                dL_dalpha += e2 * dpix2; B2 += alpha * e2;
                val[SACC_B] = dchannel_dcolor * dpix2;

                float dL_dz = 0.0f;
                float dL_dweight = 0;
                const float inv_cd = rcp_approx(c_d);
                const float m_d = kFarB / (kFarB - kNearB) * (1 - kNearB * inv_cd);
                const float dmd_dd = ((kFarB * kNearB) / (kFarB - kNearB)) * (inv_cd * inv_cd);
                if (pos == median_contributor - 1u) dL_dz += dL_dmedian_depth;
                dL_dweight += (final_D2 + m_d * m_d * final_A - 2 * m_d * final_D) * dL_dreg;
                dL_dalpha += dL_dweight - last_dL_dT;
                last_dL_dT = dL_dweight * alpha + (1 - alpha) * last_dL_dT;
                const float dL_dmd = 2.0f * (T * alpha) * (m_d * final_A - final_D) * dL_dreg;
                dL_dz += dL_dmd * dmd_dd;

                const float ed = c_d - Bd, ea = 1.0f - Ba;
                dL_dalpha += ed * dL_ddepth; Bd += alpha * ed;
                dL_dalpha += ea * dL_daccum; Ba += alpha * ea;

                const float n0 = nrm0 - Bn0, n1 = nrm1 - Bn1, n2 = nrm2 - Bn2;
                dL_dalpha += n0 * dn0; Bn0 += alpha * n0;
                val[SACC_NX] = alpha * T * dn0;
                dL_dalpha += n1 * dn1; Bn1 += alpha * n1;
                val[SACC_NY] = alpha * T * dn1;
                dL_dalpha += n2 * dn2; Bn2 += alpha * n2;
                val[SACC_NZ] = alpha * T * dn2;

                dL_dalpha *= T;
                dL_dalpha += (-T_final * inv_1ma) * bg_dot_dpixel;

                const float dL_dG = opa * dL_dalpha;
                dL_dz += alpha * T * dL_ddepth;

                if (rho3d <= rho2d) {
                    // backward.cu:399-429
                    const float dLdsx = dL_dG * -G * sx + dL_dz * Tw.x;
                    const float dLdsy = dL_dG * -G * sy + dL_dz * Tw.y;
                    const float inv_pz = rcp_approx(p.z);
                    const float dsx_pz = dLdsx * inv_pz;
                    const float dsy_pz = dLdsy * inv_pz;
                    const float3 dL_dp = make_float3(dsx_pz, dsy_pz, -(dsx_pz * sx + dsy_pz * sy));
                    const float3 dL_dk = make_float3(l.y * dL_dp.z - l.z * dL_dp.y, l.z * dL_dp.x - l.x * dL_dp.z,
                                                     l.x * dL_dp.y - l.y * dL_dp.x);
                    const float3 dL_dl = make_float3(dL_dp.y * k.z - dL_dp.z * k.y, dL_dp.z * k.x - dL_dp.x * k.z,
                                                     dL_dp.x * k.y - dL_dp.y * k.x);
                    val[0] = -dL_dk.x; val[1] = -dL_dk.y; val[2] = -dL_dk.z;
                    val[3] = -dL_dl.x; val[4] = -dL_dl.y; val[5] = -dL_dl.z;
                    val[6] = pxf * dL_dk.x + pyf * dL_dl.x + dL_dz * sx;
                    val[7] = pxf * dL_dk.y + pyf * dL_dl.y + dL_dz * sy;
                    val[8] = pxf * dL_dk.z + pyf * dL_dl.z + dL_dz * 1.0f;
                } else {
                    // backward.cu:430-440
                    const float dG_ddelx = -G * 2.0f * dx;
                    const float dG_ddely = -G * 2.0f * dy;
                    val[SACC_MX] = dL_dG * dG_ddelx;
                    val[SACC_MY] = dL_dG * dG_ddely;
                    val[6] = sx * dL_dz;
                    val[7] = sy * dL_dz;
                    val[8] = dL_dz;
                }
                val[SACC_OP] = G * dL_dalpha;
            }
            float* const a = acc + (size_t)__shfl_sync(FULL, id_c, j) * SACC_STRIDE;
            if (__popc(act_mask) <= direct_max) {       // splat edges: direct predicated REDs, no warp reduction
                if (act) {      // vector reductions on the 80-byte slot (16-byte aligned)
                    red_add_v4(a, val[0], val[1], val[2], val[3]); red_add_v4(a + 4, val[4], val[5], val[6], val[7]);
                    red_add_v4(a + 8, val[8], val[9], val[10], val[11]); red_add_v4(a + 12, val[12], val[13], val[14], val[15]);
                    red_add_v2(a + 16, val[16], val[17]);
                }
                continue;
            }
            float s1[9], s2[5], s3[3], s4[2], s5[1];
            bfly_stage<18>(val, s1, 16, h16);
            bfly_stage<9>(s1, s2, 8, h8);
            bfly_stage<5>(s2, s3, 4, h4);
            bfly_stage<3>(s3, s4, 2, h2);
            bfly_stage<2>(s4, s5, 1, h1);
            if (writer && s5[0] != 0.0f) atomicAdd(a + slot, s5[0]);
        }
        __syncwarp();
        have_c = have_n; pos_c = pos_n; id_c = id_n; buf ^= 1;
    }
}

cudaError_t launch_surfel_render_backward(int W, int H, int gx, int gy, const SurfelGeom& g, const SurfelImage& img,
                                          const BinningState& bin, const float* bg, const float* dL_dpix,
                                          const float* dL_dothers, float* acc, cudaStream_t st) {
    static int per_warp = -1;           // GSR_BWD_WARP=0: the eight-warp CTA with barriers (round 1/2a)
    if (per_warp < 0) { const char* e = getenv("GSR_BWD_WARP"); per_warp = e ? (atoi(e) ? 1 : 0) : 1; }
    if (per_warp) {
        surfel_render_backward_warp_kernel<<<gx * gy * 8, 32, 0, st>>>(W, H, gx, img.ranges, img.tile_order, bin.point_list, g.rec, bg,
                                                                     img.n_contrib, img.median_contrib, img.final_T,
                                                                     img.M1, img.M2, dL_dpix, dL_dothers, acc, blend_direct_max(20),
                                                                     reinterpret_cast<const uint32_t*>(bin.contrib));
        return cudaGetLastError();
    }
    cudaError_t e = cudaFuncSetAttribute(surfel_render_backward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SURFEL_SMEM_B);
    if (e != cudaSuccess) return e;
    surfel_render_backward_kernel<<<gx * gy, 256, SURFEL_SMEM_B, st>>>(W, H, gx, img.ranges, img.tile_order, bin.point_list, g.rec, bg,
                                                                      img.n_contrib, img.median_contrib, img.final_T,
                                                                      img.M1, img.M2, dL_dpix, dL_dothers, acc, blend_direct_max(0),
                                                                      reinterpret_cast<const uint32_t*>(bin.contrib));
    return cudaGetLastError();
}

}  // namespace gsr
