// preprocess.cu -- per-Gaussian forward stage: cull, project, cov3D -> cov2D -> conic, radius,
// tile rect, SH -> RGB, plus the per-tile histogram that replaces the reference's scan.
//
// Replaces preprocessCUDA (reference diff-gaussian-rasterization/cuda_rasterizer/forward.cu:156-256),
// computeCov3D (:118-152), computeCov2D (:74-113), computeColorFromSH (:20-71), in_frustum
// (auxiliary.h:139-164) and checkFrustum (rasterizer_impl.cu:54-66).
//
// Bit-exactness: every float that decides an integer (depth key, radius, tile rect) is produced
// with an explicit rounding sequence equal to what nvcc emits for the reference; conic and rgb
// follow the same sequences so the blend inputs are bit-identical too.
#include "common.cuh"
#include "sh.cuh"
#include <cfloat>
#include <cstdlib>

namespace gsr {

// Sigma = (S R)^T (S R) from scale*mod and the un-normalised quaternion (forward.cu:118-152).
__device__ __forceinline__ void cov3d_from_scale_rot(float sx, float sy, float sz, float mod,
                                                    float r, float x, float y, float z, float* c) {
    const float xz = mulr(x, z), rx = mulr(r, x), rz = mulr(r, z);
    const float yy = mulr(y, y), zz = mulr(z, z);
    const float xz_p_ry = fmar(r, y, xz), xz_m_ry = fmar(-r, y, xz);
    const float yz_m_rx = fmar(y, z, -rx), yz_p_rx = fmar(y, z, rx);
    const float xy_m_rz = fmar(x, y, -rz), xy_p_rz = fmar(x, y, rz);
    const float xx_yy = fmar(x, x, yy), xx_zz = fmar(x, x, zz), yy_zz = addr(yy, zz);
    const float R00 = subr(1.0f, addr(yy_zz, yy_zz));
    const float R11 = subr(1.0f, addr(xx_zz, xx_zz));
    const float R22 = subr(1.0f, addr(xx_yy, xx_yy));
    const float s0 = mulr(mod, sx), s1 = mulr(mod, sy), s2 = mulr(mod, sz);
    // M = S * R (glm column-major): column j = (s0*R[j][0], s1*R[j][1], s2*R[j][2])
    const float m00 = mulr(s0, R00),                  m01 = mulr(s1, addr(xy_m_rz, xy_m_rz)), m02 = mulr(s2, addr(xz_p_ry, xz_p_ry));
    const float m10 = mulr(s0, addr(xy_p_rz, xy_p_rz)), m11 = mulr(s1, R11),                  m12 = mulr(s2, addr(yz_m_rx, yz_m_rx));
    const float m20 = mulr(s0, addr(xz_m_ry, xz_m_ry)), m21 = mulr(s1, addr(yz_p_rx, yz_p_rx)), m22 = mulr(s2, R22);
    c[0] = fmar(m02, m02, fmar(m00, m00, mulr(m01, m01)));
    c[1] = fmar(m12, m02, fmar(m10, m00, mulr(m11, m01)));
    c[2] = fmar(m22, m02, fmar(m20, m00, mulr(m21, m01)));
    c[3] = fmar(m12, m12, fmar(m10, m10, mulr(m11, m11)));
    c[4] = fmar(m22, m12, fmar(m20, m10, mulr(m21, m11)));
    c[5] = fmar(m22, m22, fmar(m20, m20, mulr(m21, m21)));
}

// EWA projection of Sigma to the screen (forward.cu:74-113). Returns (a, b, c) with the 0.3 low-pass.
__device__ __forceinline__ float3 cov2d_from_cov3d(float tx, float ty, float tz, float focal_x, float focal_y,
                                                  float tanfovx, float tanfovy, const float* c,
                                                  const float* __restrict__ v) {
    const float limx = mulr(tanfovx, 1.3f), limy = mulr(tanfovy, 1.3f);
    const float cx = fminf(limx, fmaxf(-limx, divr(tx, tz)));
    const float cy = fminf(limy, fmaxf(-limy, divr(ty, tz)));
    const float ntz = -tz, tz2 = mulr(tz, tz);
    const float J00 = divr(focal_x, tz);
    const float J02 = divr(mulr(focal_x, mulr(cx, ntz)), tz2);
    const float J11 = divr(focal_y, tz);
    const float J12 = divr(mulr(focal_y, mulr(cy, ntz)), tz2);
    const float v0 = v[0], v1 = v[1], v2 = v[2], v4 = v[4], v5 = v[5], v6 = v[6], v8 = v[8], v9 = v[9], v10 = v[10];
    // T = W * J, zero products kept literally as the reference build keeps them
    const float T00 = fmar(v2, J02, fmar(v0, J00, mulr(v1, 0.0f)));
    const float T01 = fmar(v6, J02, fmar(v4, J00, mulr(v5, 0.0f)));
    const float T02 = fmar(J02, v10, fmar(v8, J00, mulr(v9, 0.0f)));
    const float T10 = fmar(v2, J12, fmar(v0, 0.0f, mulr(J11, v1)));
    const float T11 = fmar(v6, J12, fmar(v4, 0.0f, mulr(J11, v5)));
    const float T12 = fmar(J12, v10, fmar(v8, 0.0f, mulr(J11, v9)));
    const float X0 = fmar(T02, c[2], fmar(T00, c[0], mulr(T01, c[1])));
    const float Y0 = fmar(T12, c[2], fmar(T10, c[0], mulr(T11, c[1])));
    const float X1 = fmar(T02, c[4], fmar(T00, c[1], mulr(T01, c[3])));
    const float Y1 = fmar(T12, c[4], fmar(T10, c[1], mulr(T11, c[3])));
    const float X2 = fmar(T02, c[5], fmar(T00, c[2], mulr(T01, c[4])));
    const float Y2 = fmar(T12, c[5], fmar(T10, c[2], mulr(T11, c[4])));
    const float cov00 = fmar(T02, X2, fmar(T00, X0, mulr(T01, X1)));
    const float cov01 = fmar(T02, Y2, fmar(T00, Y0, mulr(T01, Y1)));
    const float cov11 = fmar(T12, Y2, fmar(T10, Y0, mulr(T11, Y1)));
    return make_float3(addr(cov00, 0.3f), cov01, addr(cov11, 0.3f));
}

// (Tried in round 2: the CTA's SH rows staged through shared memory with unit-stride 16-byte cp.async, as the backward
// does (sh.cuh, sh48_*), either right after the depth test or once the rectangle is known: C3 0.36 -> 0.44 ms both ways.
// 48 KB of shared memory per CTA leaves ~60 KB of L1 for the strided 4-byte reads of means / scales / rotations and the
// 48-byte record stores, which is what this kernel leans on; the float4 row reads below stay.)
template <bool VEC, bool RAW>
__global__ void __launch_bounds__(256, 4)
preprocess_kernel(FwdArgs a, GeomState g, int32_t* __restrict__ radii) {
    const int idx = blockIdx.x * 256 + threadIdx.x;
    // batched views (gsr_rasterize_forward_views): blockIdx.y = view; its camera, its P records and its tile counters.
    // A single-view launch has gridDim.y == 1 and shifts nothing.
    const int view_id = blockIdx.y;
    if (view_id) {
        a.view += 16 * view_id; a.proj += 16 * view_id; a.campos += 3 * view_id;
        g.rec += (size_t)view_id * a.P; g.clamped += (size_t)view_id * a.P; g.tiles_touched += (size_t)view_id * a.P;
        g.tile_count += tc((size_t)view_id * a.gx * a.gy);
        radii += (size_t)view_id * a.P;
    }
    uint32_t tiles = 0, rx0 = 0, ry0 = 0, rw = 0;
    if (idx < a.P) {
        int my_radius = 0;
        do {
            const float px = __ldg(a.means3D + 3 * idx), py = __ldg(a.means3D + 3 * idx + 1), pz = __ldg(a.means3D + 3 * idx + 2);
            const float depth = xform_row(a.view, 2, px, py, pz);
            if (depth <= 0.2f) {                          // in_frustum, auxiliary.h:154 (NaN passes, as there)
                if (a.prefiltered) __trap();
                break;
            }
            const float hx = xform_row(a.proj, 0, px, py, pz);
            const float hy = xform_row(a.proj, 1, px, py, pz);
            const float hw = xform_row(a.proj, 3, px, py, pz);
            const float p_w = __frcp_rn(addr(hw, 0.0000001f));
            const float projx = mulr(hx, p_w), projy = mulr(hy, p_w);

            float c3[6];
            if (a.cov3D_precomp) {
#pragma unroll
                for (int k = 0; k < 6; ++k) c3[k] = __ldg(a.cov3D_precomp + 6 * idx + k);
            } else {
                // scalar loads: the caller's tensor may be a view at any 4-byte offset
                const float* qp = a.rotations + 4 * (size_t)idx;
                float4 q = make_float4(__ldg(qp), __ldg(qp + 1), __ldg(qp + 2), __ldg(qp + 3));
                float sx = __ldg(a.scales + 3 * idx), sy = __ldg(a.scales + 3 * idx + 1), sz = __ldg(a.scales + 3 * idx + 2);
                if constexpr (RAW) { q = act_rotation(q); sx = act_scale(sx); sy = act_scale(sy); sz = act_scale(sz); }
                cov3d_from_scale_rot(sx, sy, sz, a.scale_modifier, q.x, q.y, q.z, q.w, c3);
                if (view_id == 0) {      // the same for every view
#pragma unroll
                    for (int k = 0; k < 6; ++k) g.cov3D[6 * idx + k] = c3[k];
                }
            }
            const float tx = xform_row(a.view, 0, px, py, pz), ty = xform_row(a.view, 1, px, py, pz);
            const float3 cov = cov2d_from_cov3d(tx, ty, depth, a.focal_x, a.focal_y, a.tanfovx, a.tanfovy, c3, a.view);

            const float det = fmar(cov.x, cov.z, -mulr(cov.y, cov.y));
            if (det == 0.0f) break;
            const float det_inv = __frcp_rn(det);
            const float ca = mulr(cov.z, det_inv), cb = mulr(det_inv, -cov.y), cc = mulr(cov.x, det_inv);
            const float mid = mulr(addr(cov.x, cov.z), 0.5f);
            const float disc = __fsqrt_rn(fmaxf(fmar(mid, mid, -det), 0.1f));
            const float lam = fmaxf(addr(mid, disc), subr(mid, disc));
            const float rad_f = ceilf(mulr(__fsqrt_rn(lam), 3.0f));
            const float ix = ndc_to_pix(projx, a.W), iy = ndc_to_pix(projy, a.H);
            const int rad = (int)rad_f;
            uint32_t x0, y0, x1, y1;
            tile_rect(ix, iy, rad, a.gx, a.gy, x0, y0, x1, y1);
            const uint32_t cnt = (x1 - x0) * (y1 - y0);
            if (cnt == 0) break;

            GaussRec rec;
            if (a.colors_precomp) {
                rec.r = __ldg(a.colors_precomp + 3 * idx); rec.g = __ldg(a.colors_precomp + 3 * idx + 1);
                rec.b = __ldg(a.colors_precomp + 3 * idx + 2);
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
                const uint8_t cl = (col.r < 0.0f ? 1 : 0) | (col.g < 0.0f ? 2 : 0) | (col.b < 0.0f ? 4 : 0);
                g.clamped[idx] = cl;
                rec.r = (col.r < 0.0f) ? 0.0f : col.r; rec.g = (col.g < 0.0f) ? 0.0f : col.g; rec.b = (col.b < 0.0f) ? 0.0f : col.b;
            }
            const float op = RAW ? act_opacity(__ldg(a.opacities + idx)) : __ldg(a.opacities + idx);
            rec.x = ix; rec.y = iy; rec.ca = ca; rec.cb = cb; rec.cc = cc; rec.opacity = op; rec.depth = depth;
            // alpha = min(0.99, op*exp(power)) < 1/255 is implied by power < thr (0.2 % margin).
            rec.thr = (op > 0.0f) ? (-logf(255.0f * op) - 0.002f) : ((op <= 0.0f) ? FLT_MAX : -FLT_MAX);
            rec.pad = 0.0f;
            float4* dst = reinterpret_cast<float4*>(g.rec + idx);
            dst[0] = make_float4(rec.x, rec.y, rec.ca, rec.cb);
            dst[1] = make_float4(rec.cc, rec.opacity, rec.depth, rec.thr);
            dst[2] = make_float4(rec.r, rec.g, rec.b, 0.0f);

            my_radius = rad;
            tiles = cnt; rx0 = x0; ry0 = y0; rw = x1 - x0;
        } while (false);
        radii[idx] = my_radius;
        g.tiles_touched[idx] = tiles;
    }
    // per-tile instance histogram (replaces the reference's scan over tiles_touched + duplicateWithKeys counting):
    // the warp's 32 rectangles are walked 32 tiles at a time (common.cuh, warp_rects_packed)
    if (!a.skip_histogram) tile_histogram(g.tile_count, a.gx, rx0, ry0, rw, tiles);
    // total instance count: one atomic per warp
    const uint32_t wsum = __reduce_add_sync(0xffffffffu, tiles);
    if ((threadIdx.x & 31) == 0 && wsum) atomicAdd(g.counters, wsum);
}

cudaError_t launch_preprocess(const FwdArgs& a_in, const GeomState& g, int32_t* radii, cudaStream_t st, int views) {
    FwdArgs a = a_in;
    const dim3 blocks((a.P + 255) / 256, views);
    const bool vec = a.shs && a.M == 16 && ((reinterpret_cast<uintptr_t>(a.shs) & 15) == 0);
    a.sh_wide = vec && ((reinterpret_cast<uintptr_t>(a.shs) & 31) == 0) && !getenv("GSR_SH_LOAD128");
    if (a.raw)    preprocess_kernel<false, true><<<blocks, 256, 0, st>>>(a, g, radii);
    else if (vec) preprocess_kernel<true, false><<<blocks, 256, 0, st>>>(a, g, radii);
    else          preprocess_kernel<false, false><<<blocks, 256, 0, st>>>(a, g, radii);
    return cudaGetLastError();
}

// markVisible (rasterizer_impl.cu:54-66,141-153): present = view-space z > 0.2
__global__ void __launch_bounds__(256)
mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ view, uint8_t* __restrict__ present) {
    const int idx = blockIdx.x * 256 + threadIdx.x;
    if (idx >= P) return;
    const float z = xform_row(view, 2, __ldg(means3D + 3 * idx), __ldg(means3D + 3 * idx + 1), __ldg(means3D + 3 * idx + 2));
    present[idx] = (z <= 0.2f) ? 0 : 1;
}
cudaError_t launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t st) {
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, means3D, view, present);
    return cudaGetLastError();
}

}  // namespace gsr
