// surfel_preprocess.cu -- per-surfel forward stage of the 2DGS path: cull, splat-to-pixel matrix,
// AABB centre/extent, radius, tile rect, SH -> RGB, per-tile histogram.
//
// Replaces preprocessCUDA (reference diff-surfel-rasterization/cuda_rasterizer/forward.cu:149-251),
// compute_transmat (:75-115), compute_aabb (:119-145), quat_to_rotmat / scale_to_mat
// (auxiliary.h:213-235,285-292) and in_frustum (auxiliary.h:183-210).
//
// The matrix algebra is written with the reference's expression trees (glm's operator* formulas
// spelled out) and left to nvcc's default contraction, the same compiler decisions the reference
// build gets; the depth key uses the explicit sequence proven bit-exact on the 3DGS path.
#include "surfel.cuh"
#include <cstdlib>
#include "sh.cuh"
#include <cfloat>

namespace gsr {

struct M3 { float c[3][3]; };   // column-major like glm::mat3: c[col][row]

// auxiliary.h:213-235 (normalising; the w component of the stored quaternion is field x)
__device__ __forceinline__ M3 surfel_quat_to_rotmat(float q0, float q1, float q2, float q3) {
    const float s = rsqrtf(q3 * q3 + q0 * q0 + q1 * q1 + q2 * q2);
    const float w = q0 * s, x = q1 * s, y = q2 * s, z = q3 * s;
    M3 R;
    R.c[0][0] = 1.f - 2.f * (y * y + z * z);
    R.c[0][1] = 2.f * (x * y + w * z);
    R.c[0][2] = 2.f * (x * z - w * y);
    R.c[1][0] = 2.f * (x * y - w * z);
    R.c[1][1] = 1.f - 2.f * (x * x + z * z);
    R.c[1][2] = 2.f * (y * z + w * x);
    R.c[2][0] = 2.f * (x * z + w * y);
    R.c[2][1] = 2.f * (y * z - w * x);
    R.c[2][2] = 1.f - 2.f * (x * x + y * y);
    return R;
}

// forward.cu:75-115. T (glm column-major: T.c[j] = j-th stored row Tu/Tv/Tw), normal in view space.
__device__ __forceinline__ void surfel_transmat(float px, float py, float pz, float sx, float sy, float mod,
                                                float q0, float q1, float q2, float q3,
                                                const float* __restrict__ proj, const float* __restrict__ view,
                                                int W, int H, M3& T, float3& normal) {
    const M3 R = surfel_quat_to_rotmat(q0, q1, q2, q3);
    const float s00 = mod * sx, s11 = mod * sy;
    // L = R * S with S = diag(s00, s11, 1): glm mat3*mat3, zero terms kept as written
    float L[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        L[0][i] = R.c[0][i] * s00 + R.c[1][i] * 0.0f + R.c[2][i] * 0.0f;
        L[1][i] = R.c[0][i] * 0.0f + R.c[1][i] * s11 + R.c[2][i] * 0.0f;
        L[2][i] = R.c[0][i] * 0.0f + R.c[1][i] * 0.0f + R.c[2][i] * 1.0f;
    }
    // m1 = transpose(splat2world) (4 columns of 3): m1[c] = (L0[c], L1[c], p[c]); m1[3] = (0, 0, 1)
    const float m1[4][3] = { { L[0][0], L[1][0], px }, { L[0][1], L[1][1], py }, { L[0][2], L[1][2], pz }, { 0.0f, 0.0f, 1.0f } };
    // A = m1 * world2ndc (mat4x3): A[c][r] = sum_k m1[k][r] * proj[c + 4k]
    float A[4][3];
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int r = 0; r < 3; ++r)
            A[c][r] = m1[0][r] * proj[c] + m1[1][r] * proj[c + 4] + m1[2][r] * proj[c + 8] + m1[3][r] * proj[c + 12];
    // T = A * ndc2pix (mat3x4 columns (W/2,0,0,(W-1)/2), (0,H/2,0,(H-1)/2), (0,0,0,1))
    const float w2 = (float)(float(W) / 2.0), wm = (float)(float(W - 1) / 2.0);
    const float h2 = (float)(float(H) / 2.0), hm = (float)(float(H - 1) / 2.0);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        T.c[0][i] = A[0][i] * w2 + A[1][i] * 0.0f + A[2][i] * 0.0f + A[3][i] * wm;
        T.c[1][i] = A[0][i] * 0.0f + A[1][i] * h2 + A[2][i] * 0.0f + A[3][i] * hm;
        T.c[2][i] = A[0][i] * 0.0f + A[1][i] * 0.0f + A[2][i] * 0.0f + A[3][i] * 1.0f;
    }
    // normal = transformVec4x3(L[2], viewmatrix)
    normal.x = view[0] * L[2][0] + view[4] * L[2][1] + view[8] * L[2][2];
    normal.y = view[1] * L[2][0] + view[5] * L[2][1] + view[9] * L[2][2];
    normal.z = view[2] * L[2][0] + view[6] * L[2][1] + view[10] * L[2][2];
}

// forward.cu:119-145. Explicit rounding sequence = what the reference build executes (nvcc 12.9, default
// -fmad: NVVM contracts the glm dot products, ptxas additionally fuses d's last term and reassociates
// p*p - (q9 - w) into fma(p, p, w - q9)); verified bit-exact against the reference on the B200.
// glm::max(a, b) compiles to FMNMX (NaN -> 1e-4).
__device__ __forceinline__ bool surfel_aabb(const M3& T, float2& center, float2& extent) {
    const float T20 = T.c[2][0], T21 = T.c[2][1], T22 = T.c[2][2];
    const float d = fmar(-T22, T22, fmar(mulr(T20, T20), 9.0f, mulr(mulr(T21, T21), 9.0f)));
    if (d == 0.0f) return false;
    const float inv = __frcp_rn(d);
    const float f9 = mulr(inv, 9.0f);
    float p[2], h0[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const float a0 = T.c[k][0], a1 = T.c[k][1], a2 = T.c[k][2];
        p[k] = fmar(-mulr(a2, T22), inv, fmar(f9, mulr(a0, T20), mulr(f9, mulr(a1, T21))));
        const float q9 = fmar(f9, mulr(a0, a0), mulr(f9, mulr(a1, a1)));
        h0[k] = fmar(p[k], p[k], fmar(mulr(a2, a2), inv, -q9));
    }
    center = make_float2(p[0], p[1]);
    extent = make_float2(__fsqrt_rn(fmaxf(h0[0], 1e-4f)), __fsqrt_rn(fmaxf(h0[1], 1e-4f)));
    return true;
}

// Conservative pixel-space box outside which this surfel's alpha is below 1/255 for every pixel:
// union of the low-pass disc (rho2d <= c2) and the projected ellipse (rho3d <= c2), inflated for
// the rounding of this computation and of the per-pixel evaluation. Unbounded when the projection
// is not a bounded ellipse. (No reference counterpart: a skipped pixel is one the reference skips.)
__device__ __forceinline__ float4 surfel_cull_box(const M3& T, float cx, float cy, float opacity) {
    const float BIG = 3.0e38f;
    if (opacity < 1.0f / 255.0f) return make_float4(BIG, -BIG, BIG, -BIG);       // alpha <= opacity: never contributes
    const float c2 = 2.0f * logf(255.0f * opacity) * 1.002f + 0.004f;
    if (!(c2 > 0.0f) || !(c2 < 1.0e4f)) return make_float4(-BIG, BIG, -BIG, BIG);
    const float a = T.c[2][0] * T.c[2][0] + T.c[2][1] * T.c[2][1], w2 = T.c[2][2] * T.c[2][2];
    const float d = c2 * a - w2;
    if (!(d < -1.0e-3f * w2)) return make_float4(-BIG, BIG, -BIG, BIG);          // crosses (or nearly) the camera plane
    const float f0 = c2 / d, f2 = -1.0f / d;
    float lo[2], hi[2];
    const float ctr[2] = { cx, cy };
    const float r2 = sqrtf(0.5f * c2) + 0.01f;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const float p = f0 * (T.c[k][0] * T.c[2][0] + T.c[k][1] * T.c[2][1]) + f2 * (T.c[k][2] * T.c[2][2]);
        const float q = f0 * (T.c[k][0] * T.c[k][0] + T.c[k][1] * T.c[k][1]) + f2 * (T.c[k][2] * T.c[k][2]);
        const float h2 = p * p - q;
        const float err = 4.0e-6f * (p * p + fabsf(q));
        const float h = sqrtf(fmaxf(h2, 0.0f) + err) + 0.02f + 1.0e-5f * fabsf(p);
        lo[k] = fminf(p - h, ctr[k] - r2);
        hi[k] = fmaxf(p + h, ctr[k] + r2);
        if (!(h < BIG) || !(fabsf(p) < BIG)) { lo[k] = -BIG; hi[k] = BIG; }
    }
    return make_float4(lo[0], hi[0], lo[1], hi[1]);
}

template <bool VEC, bool RAW>
__global__ void __launch_bounds__(256, 4)
surfel_preprocess_kernel(SurfelFwdArgs a, SurfelGeom g, int32_t* __restrict__ radii) {
    const int idx = blockIdx.x * 256 + threadIdx.x;
    uint32_t tiles = 0, rx0 = 0, ry0 = 0, rw = 0;
    if (idx < a.P) {
        int my_radius = 0;
        do {
            const float px = __ldg(a.means3D + 3 * idx), py = __ldg(a.means3D + 3 * idx + 1), pz = __ldg(a.means3D + 3 * idx + 2);
            const float depth = xform_row(a.view, 2, px, py, pz);
            if (depth <= 0.2f) {
                if (a.prefiltered) __trap();
                break;
            }
            const float vx = xform_row(a.view, 0, px, py, pz), vy = xform_row(a.view, 1, px, py, pz);
            M3 T; float3 normal;
            if (a.transMat_precomp == nullptr) {
                const float* qp = a.rotations + 4 * (size_t)idx;
                float4 q = make_float4(__ldg(qp), __ldg(qp + 1), __ldg(qp + 2), __ldg(qp + 3));
                float s0 = __ldg(a.scales + 2 * idx), s1 = __ldg(a.scales + 2 * idx + 1);
                if constexpr (RAW) { q = act_rotation(q); s0 = act_scale(s0); s1 = act_scale(s1); }
                surfel_transmat(px, py, pz, s0, s1, a.scale_modifier, q.x, q.y, q.z, q.w, a.proj, a.view, a.W, a.H, T, normal);
            } else {
                const float* tp = a.transMat_precomp + 9 * (size_t)idx;
#pragma unroll
                for (int j = 0; j < 3; ++j)
#pragma unroll
                    for (int i = 0; i < 3; ++i) T.c[j][i] = __ldg(tp + 3 * j + i);
                normal = make_float3(0.0f, 0.0f, 1.0f);
            }
            // DUAL_VISIABLE (forward.cu:209-214)
            const float cosv = -(vx * normal.x + vy * normal.y + depth * normal.z);
            if (cosv == 0) break;
            const float mult = cosv > 0 ? 1.0f : -1.0f;
            normal.x = mult * normal.x; normal.y = mult * normal.y; normal.z = mult * normal.z;

            float2 center, extent;
            if (!surfel_aabb(T, center, extent)) break;
            const float radius = ceilf(fmaxf(fmaxf(extent.x, extent.y), 3.0f * 0.707106f));
            const int rad = (int)radius;
            uint32_t x0, y0, x1, y1;
            tile_rect(center.x, center.y, rad, a.gx, a.gy, x0, y0, x1, y1);
            const uint32_t cnt = (x1 - x0) * (y1 - y0);
            if (cnt == 0) break;

            float cr, cg, cb;
            if (a.colors_precomp) {
                cr = __ldg(a.colors_precomp + 3 * idx); cg = __ldg(a.colors_precomp + 3 * idx + 1); cb = __ldg(a.colors_precomp + 3 * idx + 2);
            } else {
                ShReader<VEC> sh;
                const float* base = RAW ? a.shs + (size_t)idx * 3 : a.shs + (size_t)idx * a.M * 3;
                sh.base = base;
                sh.rest = RAW ? a.shs_rest + (size_t)idx * (a.M - 1) * 3 : base + 3;
                if constexpr (VEC) {
#pragma unroll
                    for (int k = 0; k < 12; k += 2) {
                        if (a.sh_wide) ldg256(base + 4 * k, sh.v[k], sh.v[k + 1]);       // (kernel-uniform)
                        else { sh.v[k] = __ldg(reinterpret_cast<const float4*>(base) + k); sh.v[k + 1] = __ldg(reinterpret_cast<const float4*>(base) + k + 1); }
                    }
                }
                const float dx = subr(px, __ldg(a.campos)), dy = subr(py, __ldg(a.campos + 1)), dz = subr(pz, __ldg(a.campos + 2));
                const float len = __fsqrt_rn(fmar(dz, dz, fmar(dx, dx, mulr(dy, dy))));
                Sh3 col = eval_sh<VEC>(a.D, sh, divr(dx, len), divr(dy, len), divr(dz, len));
                g.clamped[idx] = (col.r < 0.0f ? 1 : 0) | (col.g < 0.0f ? 2 : 0) | (col.b < 0.0f ? 4 : 0);
                cr = (col.r < 0.0f) ? 0.0f : col.r; cg = (col.g < 0.0f) ? 0.0f : col.g; cb = (col.b < 0.0f) ? 0.0f : col.b;
            }
            const float op = RAW ? act_opacity(__ldg(a.opacities + idx)) : __ldg(a.opacities + idx);
            const float4 box = surfel_cull_box(T, center.x, center.y, op);
            float4* dst = reinterpret_cast<float4*>(g.rec + idx);
            dst[0] = make_float4(center.x, center.y, T.c[0][0], T.c[0][1]);
            dst[1] = make_float4(T.c[0][2], T.c[1][0], T.c[1][1], T.c[1][2]);
            dst[2] = make_float4(T.c[2][0], T.c[2][1], T.c[2][2], normal.x);
            dst[3] = make_float4(normal.y, normal.z, op, cr);
            dst[4] = make_float4(cg, cb, depth, 0.0f);
            dst[5] = box;

            my_radius = rad;
            tiles = cnt; rx0 = x0; ry0 = y0; rw = x1 - x0;
        } while (false);
        radii[idx] = my_radius;
        g.tiles_touched[idx] = tiles;
    }
    tile_histogram(g.tile_count, a.gx, rx0, ry0, rw, tiles);      // per-tile instance counts, warp-cooperative
    const uint32_t wsum = __reduce_add_sync(0xffffffffu, tiles);
    if ((threadIdx.x & 31) == 0 && wsum) atomicAdd(g.counters, wsum);
}

cudaError_t launch_surfel_preprocess(const SurfelFwdArgs& a_in, const SurfelGeom& g, int32_t* radii, cudaStream_t st) {
    SurfelFwdArgs a = a_in;
    const int blocks = (a.P + 255) / 256;
    const bool vec = a.shs && a.M == 16 && ((reinterpret_cast<uintptr_t>(a.shs) & 15) == 0);
    a.sh_wide = vec && ((reinterpret_cast<uintptr_t>(a.shs) & 31) == 0) && !getenv("GSR_SH_LOAD128");
    if (a.raw)    surfel_preprocess_kernel<false, true><<<blocks, 256, 0, st>>>(a, g, radii);
    else if (vec) surfel_preprocess_kernel<true, false><<<blocks, 256, 0, st>>>(a, g, radii);
    else          surfel_preprocess_kernel<false, false><<<blocks, 256, 0, st>>>(a, g, radii);
    return cudaGetLastError();
}

}  // namespace gsr
