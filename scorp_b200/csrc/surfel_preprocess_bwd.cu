// surfel_preprocess_bwd.cu -- per-surfel backward of the 2DGS path: transMat/normal gradients ->
// mean, scale, rotation; SH backward; densification hack on dL_dmean2D; writes every output row.
//
// Replaces preprocessCUDA (reference diff-surfel-rasterization/cuda_rasterizer/backward.cu:582-636),
// compute_transmat_aabb (:449-579), quat_to_rotmat_vjp (auxiliary.h:238-282), the SH backward
// (:20-139, shared with 3DGS in sh.cuh) and the nine torch::zeros of the glue
// (rasterize_points.cu:192-200). Reference quirks kept on purpose: the backward rebuilds T from the
// UNMODIFIED scale (scale_to_mat(scale, 1.0f), :498) and recovers W, H as int(focal*tanfov*2) (:613-614).
#include "surfel.cuh"
#include "sh.cuh"

namespace gsr {

// M3: 3 * SH coefficients per surfel (0: colours were precomputed). RAW: model-space variant (split SH rows,
// gradients taken through sigmoid / exp / normalize to the stored parameters), as in preprocess_bwd.cu.
template <int M3, bool RAW, bool V4 = false>
__global__ void __launch_bounds__(SH_TILE, 8)
surfel_preprocess_backward_kernel(SurfelBwdArgs a) {
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
        // asynchronous (LDGSTS): in flight while the per-surfel inputs are fetched and the transMat part is computed
        if constexpr (V4)  sh48_load_async(a.shs, row0, nrows, s_vis, reinterpret_cast<float4*>(s_sh));
        else if constexpr (RAW) sh_tile_load_async_split<M3>(a.shs, a.shs_rest, row0, nrows, s_vis, s_sh);
        else               sh_tile_load_async<M3>(a.shs, row0, nrows, s_vis, s_sh);
    }

    float dT[3][3] = { { 0, 0, 0 }, { 0, 0, 0 }, { 0, 0, 0 } };   // column-major like glm: dT[j] = gradient of row j of transMat
    V3 dmean = { 0, 0, 0 }, dm2d = { 0, 0, 0 }, dcol = { 0, 0, 0 }, m = { 0, 0, 0 };
    float dscale0 = 0.f, dscale1 = 0.f, dop = 0.f;
    float4 drot = { 0, 0, 0, 0 };
    float* const my_sh = s_sh + threadIdx.x * (V4 ? 48 : ShTile<M3>::STRIDE);      // coefficients in, gradients out (in place)

    if (vis) {
        const float* ac = a.acc + (size_t)idx * SACC_STRIDE;
        float acv[20];
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(ac) + k);
            acv[4 * k] = v.x; acv[4 * k + 1] = v.y; acv[4 * k + 2] = v.z; acv[4 * k + 3] = v.w;
        }
#pragma unroll
        for (int j = 0; j < 3; ++j)
#pragma unroll
            for (int i = 0; i < 3; ++i) dT[j][i] = acv[3 * j + i];
        const float dmx = acv[SACC_MX], dmy = acv[SACC_MY];
        dop = acv[SACC_OP];
        dcol = { acv[SACC_R], acv[SACC_G], acv[SACC_B] };
        const V3 dnormal = { acv[SACC_NX], acv[SACC_NY], acv[SACC_NZ] };

        const int W = int(a.focal_x * a.tanfovx * 2);
        const int H = int(a.focal_y * a.tanfovy * 2);
        const float4* rq = reinterpret_cast<const float4*>(a.rec + idx);
        const float4 r0 = __ldg(rq), r1 = __ldg(rq + 1), r2 = __ldg(rq + 2);
        const float Tfwd[9] = { r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y, r2.z };   // forward transMat
        m = ld3(a.means3D + 3 * idx);
        const float* v = a.view;
        const float* pr = a.proj;

        float T[3][3], Pm[3][4], R[3][3];
        float q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f, sc0 = 0.f, sc1 = 0.f, qn = 1.0f;
        V3 normal = { 0, 0, 0 };
        if (a.T_precomp) {
#pragma unroll
            for (int j = 0; j < 3; ++j)
#pragma unroll
                for (int i = 0; i < 3; ++i) T[j][i] = Tfwd[3 * j + i];
        } else {
            const float* qp = a.rotations + 4 * (size_t)idx;
            q0 = __ldg(qp); q1 = __ldg(qp + 1); q2 = __ldg(qp + 2); q3 = __ldg(qp + 3);
            sc0 = __ldg(a.scales + 2 * idx); sc1 = __ldg(a.scales + 2 * idx + 1);
            if constexpr (RAW) {
                const float4 qr = make_float4(q0, q1, q2, q3);
                qn = quat_norm(qr);
                const float4 qa = act_rotation(qr);
                q0 = qa.x; q1 = qa.y; q2 = qa.z; q3 = qa.w;
                sc0 = act_scale(sc0); sc1 = act_scale(sc1);
            }
            const float s = rsqrtf(q3 * q3 + q0 * q0 + q1 * q1 + q2 * q2);
            const float w = q0 * s, x = q1 * s, y = q2 * s, z = q3 * s;
            R[0][0] = 1.f - 2.f * (y * y + z * z); R[0][1] = 2.f * (x * y + w * z); R[0][2] = 2.f * (x * z - w * y);
            R[1][0] = 2.f * (x * y - w * z); R[1][1] = 1.f - 2.f * (x * x + z * z); R[1][2] = 2.f * (y * z + w * x);
            R[2][0] = 2.f * (x * z + w * y); R[2][1] = 2.f * (y * z - w * x); R[2][2] = 1.f - 2.f * (x * x + y * y);
            const float s00 = 1.0f * sc0, s11 = 1.0f * sc1;
            float L[3][3];
#pragma unroll
            for (int i = 0; i < 3; ++i) { L[0][i] = R[0][i] * s00; L[1][i] = R[1][i] * s11; L[2][i] = R[2][i]; }
            const float w2 = (float)(float(W) / 2.0), wm = (float)(float(W - 1) / 2.0);
            const float h2 = (float)(float(H) / 2.0), hm = (float)(float(H - 1) / 2.0);
            // P = world2ndc * ndc2pix (mat3x4): Pm[j][i], i = 0..3
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                Pm[0][i] = pr[4 * i] * w2 + pr[4 * i + 3] * wm;
                Pm[1][i] = pr[4 * i + 1] * h2 + pr[4 * i + 3] * hm;
                Pm[2][i] = pr[4 * i + 3];
            }
            // T = transpose(M) * P: T[j][r] = sum_k Mt[k][r] * P[j][k], Mt[k] = (L0[k], L1[k], p[k]), Mt[3] = (0,0,1)
            const float Mt[4][3] = { { L[0][0], L[1][0], m.x }, { L[0][1], L[1][1], m.y }, { L[0][2], L[1][2], m.z }, { 0.f, 0.f, 1.f } };
#pragma unroll
            for (int j = 0; j < 3; ++j)
#pragma unroll
                for (int r = 0; r < 3; ++r)
                    T[j][r] = Mt[0][r] * Pm[j][0] + Mt[1][r] * Pm[j][1] + Mt[2][r] * Pm[j][2] + Mt[3][r] * Pm[j][3];
            normal = { v[0] * L[2][0] + v[4] * L[2][1] + v[8] * L[2][2], v[1] * L[2][0] + v[5] * L[2][1] + v[9] * L[2][2],
                       v[2] * L[2][0] + v[6] * L[2][1] + v[10] * L[2][2] };
        }

        // raw blend accumulation; the reference keeps this in global memory unless (precomp && dmean2D != 0)
        float dT_out[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) dT_out[k] = acv[k];

        if (dmx != 0 || dmy != 0) {     // AABB-centre VJP (backward.cu:521-536)
            const float t0 = 9.0f, t1 = 9.0f, t2 = -1.0f;
            const float d = t0 * (T[2][0] * T[2][0]) + t1 * (T[2][1] * T[2][1]) + t2 * (T[2][2] * T[2][2]);
            const float inv = 1.0f / d;
            const float f[3] = { t0 * inv, t1 * inv, t2 * inv };
            const float tv[3] = { t0, t1, t2 };
            float dT3[3], dLdf[3];
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                dT[0][i] += dmx * f[i] * T[2][i];
                dT[1][i] += dmy * f[i] * T[2][i];
                dT3[i] = dmx * f[i] * T[0][i] + dmy * f[i] * T[1][i];
                dLdf[i] = dmx * T[0][i] * T[2][i] + dmy * T[1][i] * T[2][i];
            }
            const float dL_dd = (float)((dLdf[0] * f[0] + dLdf[1] * f[1] + dLdf[2] * f[2]) * (-1.0 / d));   // double, as the reference
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                dT3[i] += dL_dd * (tv[i] * T[2][i] * 2.0f);
                dT[2][i] += dT3[i];
            }
            if (a.T_precomp) {
#pragma unroll
                for (int j = 0; j < 3; ++j)
#pragma unroll
                    for (int i = 0; i < 3; ++i) dT_out[3 * j + i] = dT[j][i];
            }
        }

        if (!a.T_precomp) {
            // dL_dM = P * transpose(dL_dT): dM[j][i] = sum_k P[k][i] * dT[k][j]   (i = 0..3)
            float dM[3][4];
#pragma unroll
            for (int j = 0; j < 3; ++j)
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    dM[j][i] = Pm[0][i] * dT[0][j] + Pm[1][i] * dT[1][j] + Pm[2][i] * dT[2][j];
            V3 dtn = { v[0] * dnormal.x + v[1] * dnormal.y + v[2] * dnormal.z,
                       v[4] * dnormal.x + v[5] * dnormal.y + v[6] * dnormal.z,
                       v[8] * dnormal.x + v[9] * dnormal.y + v[10] * dnormal.z };
            const V3 pv = { v[0] * m.x + v[4] * m.y + v[8] * m.z + v[12], v[1] * m.x + v[5] * m.y + v[9] * m.z + v[13],
                            v[2] * m.x + v[6] * m.y + v[10] * m.z + v[14] };
            const float cosv = -(pv.x * normal.x + pv.y * normal.y + pv.z * normal.z);
            const float mult = cosv > 0 ? 1.0f : -1.0f;
            dtn = mult * dtn;
            const V3 dRS0 = { dM[0][0], dM[0][1], dM[0][2] }, dRS1 = { dM[1][0], dM[1][1], dM[1][2] };
            // v_R columns (glm v_R[c][r])
            const float vR[3][3] = { { dRS0.x * sc0, dRS0.y * sc0, dRS0.z * sc0 }, { dRS1.x * sc1, dRS1.y * sc1, dRS1.z * sc1 },
                                     { dtn.x, dtn.y, dtn.z } };
            const float s = rsqrtf(q3 * q3 + q0 * q0 + q1 * q1 + q2 * q2);
            const float w = q0 * s, x = q1 * s, y = q2 * s, z = q3 * s;
            drot.x = 2.f * (x * (vR[1][2] - vR[2][1]) + y * (vR[2][0] - vR[0][2]) + z * (vR[0][1] - vR[1][0]));
            drot.y = 2.f * (-2.f * x * (vR[1][1] + vR[2][2]) + y * (vR[0][1] + vR[1][0]) + z * (vR[0][2] + vR[2][0]) + w * (vR[1][2] - vR[2][1]));
            drot.z = 2.f * (x * (vR[0][1] + vR[1][0]) - 2.f * y * (vR[0][0] + vR[2][2]) + z * (vR[1][2] + vR[2][1]) + w * (vR[2][0] - vR[0][2]));
            drot.w = 2.f * (x * (vR[0][2] + vR[2][0]) + y * (vR[1][2] + vR[2][1]) - 2.f * z * (vR[0][0] + vR[1][1]) + w * (vR[0][1] - vR[1][0]));
            dscale0 = dRS0.x * R[0][0] + dRS0.y * R[0][1] + dRS0.z * R[0][2];
            dscale1 = dRS1.x * R[1][0] + dRS1.y * R[1][1] + dRS1.z * R[1][2];
            dmean = { dM[2][0], dM[2][1], dM[2][2] };
            if constexpr (RAW) {      // exp and normalize (preprocess_bwd.cu)
                dscale0 *= sc0; dscale1 *= sc1;
                const float qg = q0 * drot.x + q1 * drot.y + q2 * drot.z + q3 * drot.w;
                const float in = 1.0f / qn;
                drot = make_float4((drot.x - q0 * qg) * in, (drot.y - q1 * qg) * in, (drot.z - q2 * qg) * in, (drot.w - q3 * qg) * in);
            }
        }
        if constexpr (RAW) {          // sigmoid
            const float so = act_opacity(__ldg(a.opacities + idx));
            dop = dop * ((1.0f - so) * so);
        }

        // densification hack (backward.cu:632-635)
        const float depth = Tfwd[8];
        dm2d.x = (float)(dT_out[2] * depth * 0.5 * float(W));
        dm2d.y = (float)(dT_out[5] * depth * 0.5 * float(H));
        dm2d.z = 0.0f;
        if constexpr (!RAW) {
#pragma unroll
            for (int k = 0; k < 9; ++k) a.dL_dtransMat[9 * (size_t)idx + k] = dT_out[k];
        }
    } else if (valid) {
        if constexpr (!RAW) {
#pragma unroll
            for (int k = 0; k < 9; ++k) a.dL_dtransMat[9 * (size_t)idx + k] = 0.0f;
        }
    }
    if constexpr (M3 != 0) {
        // ---- colour -> SH (+ mean), once the tile has landed
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
            const float hv[PEER_HEAD_STRIDE] = { dmean.x, dmean.y, dmean.z, dop, dscale0, dscale1, drot.x, drot.y,
                                                 drot.z, drot.w, 0.0f, 0.0f };
            peer_head_store(a.peer_head, idx, hv);
            head_sent = true;
        }
    }

    st3(a.dL_dmeans2D + 3 * idx, dm2d);
    if constexpr (!RAW) st3(a.dL_dcolors + 3 * idx, dcol);
    if (head_sent) return;
    st3(a.dL_dmeans3D + 3 * idx, dmean);
    a.dL_dopacity[idx] = dop;
    a.dL_dscales[2 * (size_t)idx] = dscale0; a.dL_dscales[2 * (size_t)idx + 1] = dscale1;
    float* dq = a.dL_drotations + 4 * (size_t)idx;
    dq[0] = drot.x; dq[1] = drot.y; dq[2] = drot.z; dq[3] = drot.w;
}

cudaError_t launch_surfel_preprocess_backward(const SurfelBwdArgs& a, cudaStream_t st) {
    const int blocks = (a.row_end - a.row_begin + SH_TILE - 1) / SH_TILE;
    if (blocks <= 0) return cudaSuccess;
    if (a.raw) {
        switch (3 * a.M) {
        case 3:  surfel_preprocess_backward_kernel<3, true><<<blocks, SH_TILE, 0, st>>>(a); break;
        case 12: surfel_preprocess_backward_kernel<12, true><<<blocks, SH_TILE, 0, st>>>(a); break;
        case 27: surfel_preprocess_backward_kernel<27, true><<<blocks, SH_TILE, 0, st>>>(a); break;
        case 48: surfel_preprocess_backward_kernel<48, true><<<blocks, SH_TILE, 0, st>>>(a); break;
        default: return cudaErrorInvalidValue;
        }
        return cudaGetLastError();
    }
    switch (a.shs ? 3 * a.M : 0) {
    case 0:  surfel_preprocess_backward_kernel<0, false><<<blocks, SH_TILE, 0, st>>>(a); break;
    case 3:  surfel_preprocess_backward_kernel<3, false><<<blocks, SH_TILE, 0, st>>>(a); break;
    case 12: surfel_preprocess_backward_kernel<12, false><<<blocks, SH_TILE, 0, st>>>(a); break;
    case 27: surfel_preprocess_backward_kernel<27, false><<<blocks, SH_TILE, 0, st>>>(a); break;
    case 48:      // whole rows at 16-byte granularity when the two bases allow it (sh.cuh)
        if ((((uintptr_t)a.shs | (uintptr_t)a.dL_dsh) & 15) == 0) surfel_preprocess_backward_kernel<48, false, true><<<blocks, SH_TILE, 0, st>>>(a);
        else surfel_preprocess_backward_kernel<48, false><<<blocks, SH_TILE, 0, st>>>(a);
        break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

}  // namespace gsr
