// preprocess_bwd.cu -- per-Gaussian backward: conic -> cov3D/mean, projection and depth -> mean,
// SH backward, cov3D -> scale/rotation, in ONE pass that also writes every output row (zeros for
// invisible Gaussians), so no gradient tensor needs a memset.
//
// Replaces computeCov2DCUDA (reference backward.cu:144-274), preprocessCUDA (:347-412),
// computeColorFromSH backward (:20-139) and computeCov3D backward (:278-341), plus the ten
// torch::zeros of RasterizeGaussiansBackwardCUDA (rasterize_points.cu:158-167).
#include "common.cuh"
#include "sh.cuh"

namespace gsr {

static_assert(SH_TILE == 128, "api.cu / surfel_api.cu align the row-range launches (launch_in_row_chunks) to 128 rows");

// M3: 3 * SH coefficients per Gaussian (0: colours were precomputed). RAW: model-space variant -- SH rows split into
// features_dc / features_rest, and the gradients are taken through sigmoid / exp / normalize to the stored parameters.
template <int M3, bool RAW, bool V4 = false>
__global__ void __launch_bounds__(SH_TILE, 8)
preprocess_backward_kernel(BwdArgs a) {
    static_assert(!V4 || (M3 == 48 && !RAW), "the 16-byte tile is for whole 48-float rows");
    constexpr int TILE_FLOATS = M3 ? SH_TILE * (V4 ? 48 : ShTile<M3>::STRIDE) : 4;
    __shared__ __align__(16) float s_sh[TILE_FLOATS];
    __shared__ uint8_t s_vis[SH_TILE];
    const size_t row0 = (size_t)a.row_begin + (size_t)blockIdx.x * SH_TILE;     // this launch covers rows [row_begin, row_end)
    const int nrows = min(SH_TILE, a.row_end - (int)row0);
    const int idx = (int)row0 + threadIdx.x;
    const bool valid = idx < a.row_end;
    const bool vis = valid && a.radii[idx] > 0;
    if constexpr (M3 != 0) {
        s_vis[threadIdx.x] = vis;
        __syncthreads();
        // asynchronous (LDGSTS): the whole tile is in flight while the per-Gaussian inputs below are fetched and
        // the covariance part is computed; waited for just before the SH backward
        if constexpr (V4)  sh48_load_async(a.shs, row0, nrows, s_vis, reinterpret_cast<float4*>(s_sh));
        else if constexpr (RAW) sh_tile_load_async_split<M3>(a.shs, a.shs_rest, row0, nrows, s_vis, s_sh);
        else               sh_tile_load_async<M3>(a.shs, row0, nrows, s_vis, s_sh);
    }

    float dcov[6] = { 0, 0, 0, 0, 0, 0 };
    V3 dmean = { 0, 0, 0 }, dm2d = { 0, 0, 0 }, dcol = { 0, 0, 0 }, dscale = { 0, 0, 0 }, m = { 0, 0, 0 };
    float4 drot = { 0, 0, 0, 0 };
    float dop = 0.0f;
    float* const my_sh = s_sh + threadIdx.x * (V4 ? 48 : ShTile<M3>::STRIDE);      // coefficients in, gradients out (in place)

    if (vis) {
        const float* ac = a.acc + (size_t)idx * ACC_STRIDE;
        const float4 q0 = __ldg(reinterpret_cast<const float4*>(ac));
        const float4 q1 = __ldg(reinterpret_cast<const float4*>(ac) + 1);
        const float4 q2 = __ldg(reinterpret_cast<const float4*>(ac) + 2);
        dm2d = { q0.x, q0.y, 0.0f };
        const float dconx = q0.z, dcony = q0.w, dconw = q1.x;
        dop = q1.y;
        dcol = { q1.z, q1.w, q2.x };
        const float ddepth = q2.y;

        m = ld3(a.means3D + 3 * idx);
        const float* v = a.view;
        const float* p = a.proj;
        const float* c = a.cov3D + 6 * (size_t)idx;
        const float c0 = __ldg(c), c1 = __ldg(c + 1), c2 = __ldg(c + 2), c3 = __ldg(c + 3), c4 = __ldg(c + 4), c5 = __ldg(c + 5);

        // ---- conic -> cov2D -> cov3D, T, J, t (backward.cu:144-274)
        float tx = v[0] * m.x + v[4] * m.y + v[8] * m.z + v[12];
        float ty = v[1] * m.x + v[5] * m.y + v[9] * m.z + v[13];
        const float tz = v[2] * m.x + v[6] * m.y + v[10] * m.z + v[14];
        const float limx = 1.3f * a.tanfovx, limy = 1.3f * a.tanfovy;
        const float txtz = tx / tz, tytz = ty / tz;
        tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
        ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
        const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
        const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
        const float J00 = a.focal_x / tz, J02 = -(a.focal_x * tx) / (tz * tz);
        const float J11 = a.focal_y / tz, J12 = -(a.focal_y * ty) / (tz * tz);
        const float T00 = v[0] * J00 + v[2] * J02, T01 = v[4] * J00 + v[6] * J02, T02 = v[8] * J00 + v[10] * J02;
        const float T10 = v[1] * J11 + v[2] * J12, T11 = v[5] * J11 + v[6] * J12, T12 = v[9] * J11 + v[10] * J12;
        const float X0 = T00 * c0 + T01 * c1 + T02 * c2, X1 = T00 * c1 + T01 * c3 + T02 * c4, X2 = T00 * c2 + T01 * c4 + T02 * c5;
        const float Y0 = T10 * c0 + T11 * c1 + T12 * c2, Y1 = T10 * c1 + T11 * c3 + T12 * c4, Y2 = T10 * c2 + T11 * c4 + T12 * c5;
        const float ca = T00 * X0 + T01 * X1 + T02 * X2 + 0.3f;
        const float cb = T00 * Y0 + T01 * Y1 + T02 * Y2;
        const float cc = T10 * Y0 + T11 * Y1 + T12 * Y2 + 0.3f;
        const float denom = ca * cc - cb * cb;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        float da = 0.f, db = 0.f, dc = 0.f;
        if (denom2inv != 0.f) {
            da = denom2inv * (-cc * cc * dconx + 2 * cb * cc * dcony + (denom - ca * cc) * dconw);
            dc = denom2inv * (-ca * ca * dconw + 2 * ca * cb * dcony + (denom - ca * cc) * dconx);
            db = denom2inv * 2 * (cb * cc * dconx - (denom + 2 * cb * cb) * dcony + ca * cb * dconw);
            dcov[0] = T00 * T00 * da + T00 * T10 * db + T10 * T10 * dc;
            dcov[3] = T01 * T01 * da + T01 * T11 * db + T11 * T11 * dc;
            dcov[5] = T02 * T02 * da + T02 * T12 * db + T12 * T12 * dc;
            dcov[1] = 2 * T00 * T01 * da + (T00 * T11 + T01 * T10) * db + 2 * T10 * T11 * dc;
            dcov[2] = 2 * T00 * T02 * da + (T00 * T12 + T02 * T10) * db + 2 * T10 * T12 * dc;
            dcov[4] = 2 * T02 * T01 * da + (T01 * T12 + T02 * T11) * db + 2 * T11 * T12 * dc;
        }
        const float dT00 = 2 * X0 * da + Y0 * db, dT01 = 2 * X1 * da + Y1 * db, dT02 = 2 * X2 * da + Y2 * db;
        const float dT10 = 2 * Y0 * dc + X0 * db, dT11 = 2 * Y1 * dc + X1 * db, dT12 = 2 * Y2 * dc + X2 * db;
        const float dJ00 = v[0] * dT00 + v[4] * dT01 + v[8] * dT02;
        const float dJ02 = v[2] * dT00 + v[6] * dT01 + v[10] * dT02;
        const float dJ11 = v[1] * dT10 + v[5] * dT11 + v[9] * dT12;
        const float dJ12 = v[2] * dT10 + v[6] * dT11 + v[10] * dT12;
        const float itz = 1.f / tz, itz2 = itz * itz, itz3 = itz2 * itz;
        const float dtx = x_grad_mul * -a.focal_x * itz2 * dJ02;
        const float dty = y_grad_mul * -a.focal_y * itz2 * dJ12;
        const float dtz = -a.focal_x * itz2 * dJ00 - a.focal_y * itz2 * dJ11
                        + (2 * a.focal_x * tx) * itz3 * dJ02 + (2 * a.focal_y * ty) * itz3 * dJ12;
        dmean = { v[0] * dtx + v[1] * dty + v[2] * dtz, v[4] * dtx + v[5] * dty + v[6] * dtz, v[8] * dtx + v[9] * dty + v[10] * dtz };

        // ---- screen-space mean and depth -> mean (backward.cu:371-403)
        const float hw_ = p[3] * m.x + p[7] * m.y + p[11] * m.z + p[15];
        const float m_w = 1.0f / (hw_ + 0.0000001f);
        const float mul1 = (p[0] * m.x + p[4] * m.y + p[8] * m.z + p[12]) * m_w * m_w;
        const float mul2 = (p[1] * m.x + p[5] * m.y + p[9] * m.z + p[13]) * m_w * m_w;
        dmean.x += (p[0] * m_w - p[3] * mul1) * dm2d.x + (p[1] * m_w - p[3] * mul2) * dm2d.y;
        dmean.y += (p[4] * m_w - p[7] * mul1) * dm2d.x + (p[5] * m_w - p[7] * mul2) * dm2d.y;
        dmean.z += (p[8] * m_w - p[11] * mul1) * dm2d.x + (p[9] * m_w - p[11] * mul2) * dm2d.y;
        const float mul3 = v[2] * m.x + v[6] * m.y + v[10] * m.z + v[14];
        dmean.x += (v[2] - v[3] * mul3) * ddepth;
        dmean.y += (v[6] - v[7] * mul3) * ddepth;
        dmean.z += (v[10] - v[11] * mul3) * ddepth;

        // ---- cov3D -> scale, rotation (backward.cu:278-341)
        if (!a.cov_precomp) {
            const float* qp = a.rotations + 4 * (size_t)idx;
            float4 q = make_float4(__ldg(qp), __ldg(qp + 1), __ldg(qp + 2), __ldg(qp + 3));
            V3 sc = ld3(a.scales + 3 * idx);
            float qn = 1.0f;
            if constexpr (RAW) {
                qn = quat_norm(q); q = act_rotation(q);
                sc = { act_scale(sc.x), act_scale(sc.y), act_scale(sc.z) };
            }
            const float r = q.x, x = q.y, y = q.z, z = q.w;
            const V3 s = a.scale_modifier * sc;
            // rows of R-hat (R-hat(row,col) = glm R[col][row])
            const V3 R0 = { 1.f - 2.f * (y * y + z * z), 2.f * (x * y + r * z), 2.f * (x * z - r * y) };
            const V3 R1 = { 2.f * (x * y - r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z + r * x) };
            const V3 R2 = { 2.f * (x * z + r * y), 2.f * (y * z - r * x), 1.f - 2.f * (x * x + y * y) };
            const V3 S0 = { dcov[0], 0.5f * dcov[1], 0.5f * dcov[2] };
            const V3 S1 = { 0.5f * dcov[1], dcov[3], 0.5f * dcov[4] };
            const V3 S2 = { 0.5f * dcov[2], 0.5f * dcov[4], dcov[5] };
            auto rowmul = [&](V3 rr, float sc) {   // 2 * sc * rr * dSigma
                V3 o = { rr.x * S0.x + rr.y * S1.x + rr.z * S2.x, rr.x * S0.y + rr.y * S1.y + rr.z * S2.y,
                         rr.x * S0.z + rr.y * S1.z + rr.z * S2.z };
                return (2.0f * sc) * o;
            };
            const V3 dM0 = rowmul(R0, s.x), dM1 = rowmul(R1, s.y), dM2 = rowmul(R2, s.z);
            dscale = { dot(R0, dM0), dot(R1, dM1), dot(R2, dM2) };
            const V3 N0 = s.x * dM0, N1 = s.y * dM1, N2 = s.z * dM2;
            drot.x = 2 * z * (N0.y - N1.x) + 2 * y * (N2.x - N0.z) + 2 * x * (N1.z - N2.y);
            drot.y = 2 * y * (N1.x + N0.y) + 2 * z * (N2.x + N0.z) + 2 * r * (N1.z - N2.y) - 4 * x * (N2.z + N1.y);
            drot.z = 2 * x * (N1.x + N0.y) + 2 * r * (N2.x - N0.z) + 2 * z * (N1.z + N2.y) - 4 * y * (N2.z + N0.x);
            drot.w = 2 * r * (N0.y - N1.x) + 2 * x * (N2.x + N0.z) + 2 * y * (N1.z + N2.y) - 4 * z * (N1.y + N0.x);
            if constexpr (RAW) {
                // exp: d/draw = d/dscale * scale;  normalize: d/draw = (g - q (q.g)) / |raw|
                dscale = { dscale.x * sc.x, dscale.y * sc.y, dscale.z * sc.z };
                const float qg = q.x * drot.x + q.y * drot.y + q.z * drot.z + q.w * drot.w;
                const float in = 1.0f / qn;
                drot = make_float4((drot.x - q.x * qg) * in, (drot.y - q.y * qg) * in, (drot.z - q.z * qg) * in, (drot.w - q.w * qg) * in);
            }
        }
        if constexpr (RAW) {      // sigmoid: d/draw = d/dopacity * s (1 - s)
            const float so = act_opacity(__ldg(a.opacities + idx));
            dop = dop * ((1.0f - so) * so);
        }
    }
    if constexpr (M3 != 0) {
        // ---- colour -> SH (+ mean)
        const uint8_t cl = vis ? a.clamped[idx] : 0;
        cp_async_wait<0>();
        __syncthreads();
        if (vis) {
            const V3 dRGB = { (cl & 1) ? 0.f : dcol.x, (cl & 2) ? 0.f : dcol.y, (cl & 4) ? 0.f : dcol.z };
            if constexpr (V4) dmean = dmean + sh48_backward(a.D, reinterpret_cast<float4*>(my_sh), threadIdx.x, m, ld3(a.campos), dRGB);
            else              dmean = dmean + sh_backward(a.D, a.M, my_sh, m, ld3(a.campos), dRGB, my_sh);
        }
        __syncthreads();
        if constexpr (V4) { if (a.dL_dsh) sh48_store(a.dL_dsh, row0, nrows, s_vis, reinterpret_cast<const float4*>(s_sh)); }
        else if constexpr (RAW) sh_tile_store_split<M3>(a.dL_dsh, a.dL_dsh_rest, row0, nrows, s_vis, s_sh);
        else if (a.dL_dsh) sh_tile_store<M3>(a.dL_dsh, row0, nrows, s_vis, s_sh);    // NULL: the caller rebuilds it (sh_exchange.cu)
    }
    if (!valid) return;
    bool head_sent = false;
    if constexpr (!RAW && M3 != 0) {
        if (a.peer_head) {      // multi-GPU: the row goes to this rank's replica of the symmetric head (common.cuh, phase 0)
            const float hv[PEER_HEAD_STRIDE] = { dmean.x, dmean.y, dmean.z, dop, dscale.x, dscale.y, dscale.z, drot.x,
                                                 drot.y, drot.z, drot.w, 0.0f };
            peer_head_store(a.peer_head, idx, hv);
            head_sent = true;
        }
    }

    st3(a.dL_dmeans2D + 3 * idx, dm2d);
    if constexpr (!RAW) {       // no model parameter behind these two in the model-space variant
        st3(a.dL_dcolors + 3 * idx, dcol);
#pragma unroll
        for (int k = 0; k < 6; ++k) a.dL_dcov3D[6 * (size_t)idx + k] = dcov[k];
    }
    if (head_sent) return;
    st3(a.dL_dmeans3D + 3 * idx, dmean);
    a.dL_dopacity[idx] = dop;
    st3(a.dL_dscales + 3 * idx, dscale);
    float* dq = a.dL_drotations + 4 * (size_t)idx;     // scalar: output may sit at any 4-byte offset
    dq[0] = drot.x; dq[1] = drot.y; dq[2] = drot.z; dq[3] = drot.w;
}

cudaError_t launch_preprocess_backward(const BwdArgs& a, cudaStream_t st) {
    const int blocks = (a.row_end - a.row_begin + SH_TILE - 1) / SH_TILE;
    if (blocks <= 0) return cudaSuccess;
    if (a.raw) {
        switch (3 * a.M) {
        case 3:  preprocess_backward_kernel<3, true><<<blocks, SH_TILE, 0, st>>>(a); break;
        case 12: preprocess_backward_kernel<12, true><<<blocks, SH_TILE, 0, st>>>(a); break;
        case 27: preprocess_backward_kernel<27, true><<<blocks, SH_TILE, 0, st>>>(a); break;
        case 48: preprocess_backward_kernel<48, true><<<blocks, SH_TILE, 0, st>>>(a); break;
        default: return cudaErrorInvalidValue;
        }
        return cudaGetLastError();
    }
    switch (a.shs ? 3 * a.M : 0) {
    case 0:  preprocess_backward_kernel<0, false><<<blocks, SH_TILE, 0, st>>>(a); break;
    case 3:  preprocess_backward_kernel<3, false><<<blocks, SH_TILE, 0, st>>>(a); break;
    case 12: preprocess_backward_kernel<12, false><<<blocks, SH_TILE, 0, st>>>(a); break;
    case 27: preprocess_backward_kernel<27, false><<<blocks, SH_TILE, 0, st>>>(a); break;
    case 48:      // whole rows at 16-byte granularity when the two bases allow it (sh.cuh)
        if ((((uintptr_t)a.shs | (uintptr_t)a.dL_dsh) & 15) == 0) preprocess_backward_kernel<48, false, true><<<blocks, SH_TILE, 0, st>>>(a);
        else preprocess_backward_kernel<48, false><<<blocks, SH_TILE, 0, st>>>(a);
        break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

}  // namespace gsr
