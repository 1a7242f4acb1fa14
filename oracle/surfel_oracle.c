/* surfel_oracle.c -- TEST INFRASTRUCTURE (checker): plain-C CPU restatement of the reference 2DGS
 * surfel rasterizer (forward + backward). Never linked into, imported by or called from scorp_b200/.
 * Callers: tests/, __graft_entry__.smoke(), bench.py's cpu_baseline leg.
 *
 * Follows DSR = /root/reference/submodules/diff-surfel-rasterization/cuda_rasterizer:
 *   auxiliary.h:213-235 (quat_to_rotmat), :285-292 (scale_to_mat), :238-282 (quat_to_rotmat_vjp)
 *   forward.cu:75-115 (compute_transmat), :119-145 (compute_aabb), :149-251 (preprocess), :258-448 (blend)
 *   rasterizer_impl.cu:70-138 (keys, ranges; identical to DGR), :197-327 (orchestration)
 *   backward.cu:145-446 (blend bwd), :449-579 (compute_transmat_aabb), :582-636 (preprocess bwd)
 * Parity status: PINNED by tests/golden/s*.npz (outputs of the reference CUDA build on a B200).
 * Gradient sums of the blend are accumulated in double (the reference's float atomicAdd order is
 * non-deterministic; a double sum is the value its runs scatter around).
 */
#include "oracle_common.h"
#include <stdio.h>

static const float near_n = 0.2f, far_n = 100.0f, FilterSize = 0.707106f, FilterInvSquare = 2.0f;   /* auxiliary.h:37-40 */

typedef struct { float m[3][3]; } m3;      /* glm layout m[col][row] */

typedef struct SurfelOracle {
    int P, W, H, gx, gy, tiles;
    int64_t N;
    float *depths, *means2D, *transMat, *normal_opacity, *rgb;
    uint8_t* clamped;
    uint32_t *tiles_touched, *point_offsets;
    uint64_t* keys;
    uint32_t* point_list;
    uint32_t* ranges;
    uint32_t *n_contrib, *median_contrib;     /* per pixel (the reference's n_contrib[0], [1]) */
    float *final_T, *M1, *M2;                 /* accum_alpha[0..2] */
    int* radii;
} SurfelOracle;

static uint32_t f2u_(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

/* auxiliary.h:213-235 */
static m3 quat_to_rotmat(const float* q) {
    float s = 1.0f / sqrtf(q[3] * q[3] + q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
    float w = q[0] * s, x = q[1] * s, y = q[2] * s, z = q[3] * s;
    m3 R;
    R.m[0][0] = 1.f - 2.f * (y * y + z * z); R.m[0][1] = 2.f * (x * y + w * z); R.m[0][2] = 2.f * (x * z - w * y);
    R.m[1][0] = 2.f * (x * y - w * z); R.m[1][1] = 1.f - 2.f * (x * x + z * z); R.m[1][2] = 2.f * (y * z + w * x);
    R.m[2][0] = 2.f * (x * z + w * y); R.m[2][1] = 2.f * (y * z - w * x); R.m[2][2] = 1.f - 2.f * (x * x + y * y);
    return R;
}

/* world2ndc * ndc2pix as a 3x4 (P[j][i], i = 0..3), forward.cu:96-109 / backward.cu:503-516 */
static void proj_to_pix(const float* proj, int W, int H, float Pm[3][4]) {
    const float w2 = (float)((float)W / 2.0), wm = (float)((float)(W - 1) / 2.0);
    const float h2 = (float)((float)H / 2.0), hm = (float)((float)(H - 1) / 2.0);
    for (int i = 0; i < 4; ++i) {
        Pm[0][i] = proj[4 * i] * w2 + proj[4 * i + 3] * wm;
        Pm[1][i] = proj[4 * i + 1] * h2 + proj[4 * i + 3] * hm;
        Pm[2][i] = proj[4 * i + 3];
    }
}

/* forward.cu:75-115. The forward associates (splat2world^T * world2ndc) * ndc2pix. */
static void compute_transmat(vec3 p, const float* scale, float mod, const float* rot, const float* proj,
                             const float* view, int W, int H, m3* T, vec3* normal, m3* R_out) {
    m3 R = quat_to_rotmat(rot);
    float L[3][3];
    for (int i = 0; i < 3; ++i) { L[0][i] = R.m[0][i] * (mod * scale[0]); L[1][i] = R.m[1][i] * (mod * scale[1]); L[2][i] = R.m[2][i]; }
    const float m1[4][3] = { { L[0][0], L[1][0], p.x }, { L[0][1], L[1][1], p.y }, { L[0][2], L[1][2], p.z }, { 0, 0, 1 } };
    float A[4][3];
    for (int c = 0; c < 4; ++c) for (int r = 0; r < 3; ++r)
        A[c][r] = m1[0][r] * proj[c] + m1[1][r] * proj[c + 4] + m1[2][r] * proj[c + 8] + m1[3][r] * proj[c + 12];
    const float w2 = (float)((float)W / 2.0), wm = (float)((float)(W - 1) / 2.0);
    const float h2 = (float)((float)H / 2.0), hm = (float)((float)(H - 1) / 2.0);
    for (int i = 0; i < 3; ++i) {
        T->m[0][i] = A[0][i] * w2 + A[3][i] * wm;
        T->m[1][i] = A[1][i] * h2 + A[3][i] * hm;
        T->m[2][i] = A[3][i];
    }
    *normal = v3(view[0] * L[2][0] + view[4] * L[2][1] + view[8] * L[2][2],
                 view[1] * L[2][0] + view[5] * L[2][1] + view[9] * L[2][2],
                 view[2] * L[2][0] + view[6] * L[2][1] + view[10] * L[2][2]);
    if (R_out) *R_out = R;
}

/* forward.cu:119-145 */
static int compute_aabb(const m3* T, float cutoff, float* center, float* extent) {
    const float t[3] = { cutoff * cutoff, cutoff * cutoff, -1.0f };
    float d = t[0] * (T->m[2][0] * T->m[2][0]) + t[1] * (T->m[2][1] * T->m[2][1]) + t[2] * (T->m[2][2] * T->m[2][2]);
    if (d == 0.0f) return 0;
    float f[3] = { (1 / d) * t[0], (1 / d) * t[1], (1 / d) * t[2] };
    for (int k = 0; k < 2; ++k) {
        float p = f[0] * (T->m[k][0] * T->m[2][0]) + f[1] * (T->m[k][1] * T->m[2][1]) + f[2] * (T->m[k][2] * T->m[2][2]);
        float q = f[0] * (T->m[k][0] * T->m[k][0]) + f[1] * (T->m[k][1] * T->m[k][1]) + f[2] * (T->m[k][2] * T->m[k][2]);
        float h0 = p * p - q;
        center[k] = p;
        extent[k] = sqrtf(1e-4f < h0 ? h0 : 1e-4f);
    }
    return 1;
}

void surfel_oracle_free(SurfelOracle* s) {
    if (!s) return;
    free(s->depths); free(s->means2D); free(s->transMat); free(s->normal_opacity); free(s->rgb); free(s->clamped);
    free(s->tiles_touched); free(s->point_offsets); free(s->keys); free(s->point_list); free(s->ranges);
    free(s->n_contrib); free(s->median_contrib); free(s->final_T); free(s->M1); free(s->M2); free(s->radii);
    free(s);
}

/* forward.cu:258-448, one tile */
static void surfel_render_tile(const SurfelOracle* s, int tile, const float* features, const float* transMats,
                               const float* bg, float* out_color, float* out_others) {
    const int W = s->W, H = s->H;
    const size_t hw = (size_t)H * W;
    const int tx = tile % s->gx, ty = tile / s->gx;
    const uint32_t r0 = s->ranges[2 * tile], r1 = s->ranges[2 * tile + 1];
    for (int ly = 0; ly < BLOCK_Y; ++ly) for (int lx = 0; lx < BLOCK_X; ++lx) {
        const int px = tx * BLOCK_X + lx, py = ty * BLOCK_Y + ly;
        if (px >= W || py >= H) continue;
        const uint32_t pix_id = (uint32_t)W * py + px;
        const float pixfx = (float)px, pixfy = (float)py;
        float T = 1.0f, C[3] = { 0, 0, 0 }, Nn[3] = { 0, 0, 0 }, D = 0, M1 = 0, M2 = 0, distortion = 0, median_depth = 0;
        uint32_t contributor = 0, last_contributor = 0, median_contributor = 0;
        for (uint32_t k = r0; k < r1; ++k) {
            contributor++;
            const uint32_t id = s->point_list[k];
            const float* Tm = transMats + 9 * (size_t)id;
            const float Tu[3] = { Tm[0], Tm[1], Tm[2] }, Tv[3] = { Tm[3], Tm[4], Tm[5] }, Tw[3] = { Tm[6], Tm[7], Tm[8] };
            const float kx = pixfx * Tw[0] - Tu[0], ky = pixfx * Tw[1] - Tu[1], kz = pixfx * Tw[2] - Tu[2];
            const float lx_ = pixfy * Tw[0] - Tv[0], ly_ = pixfy * Tw[1] - Tv[1], lz_ = pixfy * Tw[2] - Tv[2];
            const float p0 = ky * lz_ - kz * ly_, p1 = kz * lx_ - kx * lz_, p2 = kx * ly_ - ky * lx_;
            if (p2 == 0.0f) continue;
            const float sx = p0 / p2, sy = p1 / p2;
            const float rho3d = sx * sx + sy * sy;
            const float dx = s->means2D[2 * id] - pixfx, dy = s->means2D[2 * id + 1] - pixfy;
            const float rho2d = FilterInvSquare * (dx * dx + dy * dy);
            const float rho = fminf(rho3d, rho2d);
            const float depth = (sx * Tw[0] + sy * Tw[1]) + Tw[2];
            if (depth < near_n) continue;
            const float* no = s->normal_opacity + 4 * (size_t)id;
            const float power = -0.5f * rho;
            if (power > 0.0f) continue;
            const float alpha = fminf(0.99f, no[3] * expf(power));
            if (alpha < 1.0f / 255.0f) continue;
            const float test_T = T * (1 - alpha);
            if (test_T < 0.0001f) break;                               /* done = true (:381-385) */
            const float w = alpha * T;
            const float A = 1 - T;
            const float m = far_n / (far_n - near_n) * (1 - near_n / depth);
            distortion += (m * m * A + M2 - 2 * m * M1) * w;
            D += depth * w; M1 += m * w; M2 += m * m * w;
            if (T > 0.5f) { median_depth = depth; median_contributor = contributor; }
            for (int ch = 0; ch < 3; ++ch) Nn[ch] += no[ch] * w;
            for (int ch = 0; ch < 3; ++ch) C[ch] += features[id * 3 + ch] * w;
            T = test_T;
            last_contributor = contributor;
        }
        s->final_T[pix_id] = T; s->n_contrib[pix_id] = last_contributor; s->median_contrib[pix_id] = median_contributor;
        s->M1[pix_id] = M1; s->M2[pix_id] = M2;
        for (int ch = 0; ch < 3; ++ch) out_color[ch * hw + pix_id] = C[ch] + T * bg[ch];
        out_others[0 * hw + pix_id] = D;
        out_others[1 * hw + pix_id] = 1 - T;
        for (int ch = 0; ch < 3; ++ch) out_others[(2 + ch) * hw + pix_id] = Nn[ch];
        out_others[5 * hw + pix_id] = median_depth;
        out_others[6 * hw + pix_id] = distortion;
    }
}

/* Rasterizer::forward, rasterizer_impl.cu:197-327 */
SurfelOracle* surfel_oracle_forward(int P, int D, int M, const float* bg, int W, int H,
                                    const float* means3D, const float* shs, const float* colors_precomp,
                                    const float* opacities, const float* scales, float scale_modifier,
                                    const float* rotations, const float* transMat_precomp,
                                    const float* view, const float* proj, const float* campos,
                                    float tan_fovx, float tan_fovy,
                                    float* out_color, float* out_others, int* radii_out) {
    (void)tan_fovx; (void)tan_fovy;
    if (!shs && !colors_precomp) return NULL;
    SurfelOracle* s = calloc(1, sizeof(SurfelOracle));
    s->P = P; s->W = W; s->H = H; s->gx = (W + BLOCK_X - 1) / BLOCK_X; s->gy = (H + BLOCK_Y - 1) / BLOCK_Y;
    s->tiles = s->gx * s->gy;
    const size_t Pn = P > 0 ? P : 1, hw = (size_t)W * H;
    s->depths = calloc(Pn, 4); s->means2D = calloc(Pn, 8); s->transMat = calloc(Pn, 36); s->normal_opacity = calloc(Pn, 16);
    s->rgb = calloc(Pn, 12); s->clamped = calloc(Pn, 3); s->tiles_touched = calloc(Pn, 4); s->point_offsets = calloc(Pn, 4);
    s->radii = calloc(Pn, 4); s->ranges = calloc((size_t)s->tiles, 8);
    s->n_contrib = calloc(hw, 4); s->median_contrib = calloc(hw, 4); s->final_T = calloc(hw, 4); s->M1 = calloc(hw, 4); s->M2 = calloc(hw, 4);
    /* preprocessCUDA, forward.cu:149-251 */
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; ++idx) {
        vec3 p = v3(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]);
        vec3 p_view = transformPoint4x3(p, view);
        if (p_view.z <= 0.2f) continue;
        m3 T; vec3 normal;
        if (!transMat_precomp) {
            compute_transmat(p, scales + 2 * (size_t)idx, scale_modifier, rotations + 4 * (size_t)idx, proj, view, W, H, &T, &normal, NULL);
            for (int j = 0; j < 3; ++j) for (int i = 0; i < 3; ++i) s->transMat[9 * (size_t)idx + 3 * j + i] = T.m[j][i];
        } else {
            for (int j = 0; j < 3; ++j) for (int i = 0; i < 3; ++i) T.m[j][i] = transMat_precomp[9 * (size_t)idx + 3 * j + i];
            normal = v3(0, 0, 1);
        }
        float cosv = -(p_view.x * normal.x + p_view.y * normal.y + p_view.z * normal.z);
        if (cosv == 0) continue;
        float mult = cosv > 0 ? 1.0f : -1.0f;
        normal = v3scale(mult, normal);
        const float cutoff = 3.0f;
        float center[2], extent[2];
        if (!compute_aabb(&T, cutoff, center, extent)) continue;
        float radius = ceilf(fmaxf(fmaxf(extent[0], extent[1]), cutoff * FilterSize));
        uint32_t x0, y0, x1, y1;
        getRect(center[0], center[1], (int)radius, s->gx, s->gy, &x0, &y0, &x1, &y1);
        if ((x1 - x0) * (y1 - y0) == 0) continue;
        if (!colors_precomp) {
            vec3 c = orc_computeColorFromSH(idx, D, M, means3D, campos, shs, s->clamped);
            s->rgb[3 * idx] = c.x; s->rgb[3 * idx + 1] = c.y; s->rgb[3 * idx + 2] = c.z;
        }
        s->depths[idx] = p_view.z;
        s->radii[idx] = (int)radius;
        s->means2D[2 * idx] = center[0]; s->means2D[2 * idx + 1] = center[1];
        float* no = s->normal_opacity + 4 * (size_t)idx;
        no[0] = normal.x; no[1] = normal.y; no[2] = normal.z; no[3] = opacities[idx];
        s->tiles_touched[idx] = (y1 - y0) * (x1 - x0);
    }
    uint32_t run = 0;
    for (int i = 0; i < P; ++i) { run += s->tiles_touched[i]; s->point_offsets[i] = run; }
    s->N = run;
    size_t Nn = run ? run : 1;
    s->keys = malloc(Nn * 8); s->point_list = malloc(Nn * 4);
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; ++idx) {
        if (s->radii[idx] <= 0) continue;
        uint32_t off = idx == 0 ? 0 : s->point_offsets[idx - 1];
        uint32_t x0, y0, x1, y1;
        getRect(s->means2D[2 * idx], s->means2D[2 * idx + 1], s->radii[idx], s->gx, s->gy, &x0, &y0, &x1, &y1);
        for (uint32_t y = y0; y < y1; ++y) for (uint32_t x = x0; x < x1; ++x) {
            uint64_t key = (uint64_t)(y * (uint32_t)s->gx + x);
            key <<= 32; key |= f2u_(s->depths[idx]);
            s->keys[off] = key; s->point_list[off] = (uint32_t)idx; off++;
        }
    }
    sort_pairs_stable(s->keys, s->point_list, (size_t)run, 32 + (int)getHigherMsb((uint32_t)s->tiles));
    for (uint32_t i = 0; i < run; ++i) {
        uint32_t cur = (uint32_t)(s->keys[i] >> 32);
        if (i == 0) s->ranges[2 * cur] = 0;
        else { uint32_t prev = (uint32_t)(s->keys[i - 1] >> 32);
               if (cur != prev) { s->ranges[2 * prev + 1] = i; s->ranges[2 * cur] = i; } }
        if (i == run - 1) s->ranges[2 * cur + 1] = run;
    }
    const float* features = colors_precomp ? colors_precomp : s->rgb;
    const float* tm = transMat_precomp ? transMat_precomp : s->transMat;
#pragma omp parallel for schedule(dynamic, 4)
    for (int t = 0; t < s->tiles; ++t) surfel_render_tile(s, t, features, tm, bg, out_color, out_others);
    if (radii_out) memcpy(radii_out, s->radii, (size_t)P * 4);
    return s;
}

int64_t surfel_oracle_num_rendered(const SurfelOracle* s) { return s->N; }
/* which: 0 depths 1 means2D 2 transMat 3 normal_opacity 4 rgb 5 clamped 6 tiles_touched 7 point_offsets
 *        8 keys 9 point_list 10 ranges 11 n_contrib 12 median_contrib 13 final_T 14 M1 15 M2 */
const void* surfel_oracle_array(const SurfelOracle* s, int which) {
    switch (which) {
    case 0: return s->depths; case 1: return s->means2D; case 2: return s->transMat; case 3: return s->normal_opacity;
    case 4: return s->rgb; case 5: return s->clamped; case 6: return s->tiles_touched; case 7: return s->point_offsets;
    case 8: return s->keys; case 9: return s->point_list; case 10: return s->ranges; case 11: return s->n_contrib;
    case 12: return s->median_contrib; case 13: return s->final_T; case 14: return s->M1; case 15: return s->M2;
    default: return NULL;
    }
}

/* ---------------------------------------------------------------------------------------------
 * Backward. Per-surfel sums from the blend are accumulated in double; acc layout (18 per surfel):
 * transMat[9], mean2D.x, mean2D.y, opacity, color[3], normal[3].
 * ------------------------------------------------------------------------------------------- */
static inline void dadd_(double* p, double v) {
#pragma omp atomic
    *p += v;
}

/* backward.cu:145-446, one tile */
static void surfel_render_tile_backward(const SurfelOracle* s, int tile, const float* bg, const float* colors,
                                        const float* transMats, const float* dL_dpixels, const float* dL_depths,
                                        double* acc) {
    const int W = s->W, H = s->H;
    const size_t hw = (size_t)H * W;
    const int tx = tile % s->gx, ty = tile / s->gx;
    const uint32_t r0 = s->ranges[2 * tile], r1 = s->ranges[2 * tile + 1];
    for (int ly = 0; ly < BLOCK_Y; ++ly) for (int lx = 0; lx < BLOCK_X; ++lx) {
        const int px = tx * BLOCK_X + lx, py = ty * BLOCK_Y + ly;
        if (px >= W || py >= H) continue;
        const uint32_t pix_id = (uint32_t)W * py + px;
        const float pixfx = (float)px, pixfy = (float)py;
        const float T_final = s->final_T[pix_id];
        float T = T_final;
        uint32_t contributor = r1 - r0;
        const uint32_t last_contributor = s->n_contrib[pix_id];
        const uint32_t median_contributor = s->median_contrib[pix_id];
        float accum_rec[3] = { 0, 0, 0 }, dL_dpixel[3];
        for (int i = 0; i < 3; ++i) dL_dpixel[i] = dL_dpixels[i * hw + pix_id];
        const float dL_ddepth = dL_depths[0 * hw + pix_id], dL_daccum = dL_depths[1 * hw + pix_id];
        const float dL_dreg = dL_depths[6 * hw + pix_id], dL_dmedian_depth = dL_depths[5 * hw + pix_id];
        float dL_dnormal2D[3];
        for (int i = 0; i < 3; ++i) dL_dnormal2D[i] = dL_depths[(2 + i) * hw + pix_id];
        float last_depth = 0, last_normal[3] = { 0, 0, 0 }, accum_depth_rec = 0, accum_alpha_rec = 0, accum_normal_rec[3] = { 0, 0, 0 };
        const float final_D = s->M1[pix_id], final_D2 = s->M2[pix_id], final_A = 1 - T_final;
        float last_dL_dT = 0, last_alpha = 0, last_color[3] = { 0, 0, 0 };
        for (uint32_t k = r1; k-- > r0;) {
            contributor--;
            if (contributor >= last_contributor) continue;
            const uint32_t id = s->point_list[k];
            const float* Tm = transMats + 9 * (size_t)id;
            const float Tu[3] = { Tm[0], Tm[1], Tm[2] }, Tv[3] = { Tm[3], Tm[4], Tm[5] }, Tw[3] = { Tm[6], Tm[7], Tm[8] };
            const float kv[3] = { pixfx * Tw[0] - Tu[0], pixfx * Tw[1] - Tu[1], pixfx * Tw[2] - Tu[2] };
            const float lv[3] = { pixfy * Tw[0] - Tv[0], pixfy * Tw[1] - Tv[1], pixfy * Tw[2] - Tv[2] };
            const float p[3] = { kv[1] * lv[2] - kv[2] * lv[1], kv[2] * lv[0] - kv[0] * lv[2], kv[0] * lv[1] - kv[1] * lv[0] };
            if (p[2] == 0.0f) continue;
            const float sx = p[0] / p[2], sy = p[1] / p[2];
            const float rho3d = sx * sx + sy * sy;
            const float dx = s->means2D[2 * id] - pixfx, dy = s->means2D[2 * id + 1] - pixfy;
            const float rho2d = FilterInvSquare * (dx * dx + dy * dy);
            const float rho = fminf(rho3d, rho2d);
            const float c_d = (sx * Tw[0] + sy * Tw[1]) + Tw[2];
            if (c_d < near_n) continue;
            const float* no = s->normal_opacity + 4 * (size_t)id;
            const float power = -0.5f * rho;
            if (power > 0.0f) continue;
            const float G = expf(power);
            const float alpha = fminf(0.99f, no[3] * G);
            if (alpha < 1.0f / 255.0f) continue;
            T = T / (1.f - alpha);
            const float dchannel_dcolor = alpha * T;
            double* a = acc + 18 * (size_t)id;
            float dL_dalpha = 0.0f;
            for (int ch = 0; ch < 3; ++ch) {
                const float c = colors[id * 3 + ch];
                accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                last_color[ch] = c;
                dL_dalpha += (c - accum_rec[ch]) * dL_dpixel[ch];
                dadd_(a + 12 + ch, dchannel_dcolor * dL_dpixel[ch]);
            }
            float dL_dz = 0.0f, dL_dweight = 0;
            const float m_d = far_n / (far_n - near_n) * (1 - near_n / c_d);
            const float dmd_dd = (far_n * near_n) / ((far_n - near_n) * c_d * c_d);
            if (contributor == median_contributor - 1u) dL_dz += dL_dmedian_depth;
            dL_dweight += (final_D2 + m_d * m_d * final_A - 2 * m_d * final_D) * dL_dreg;
            dL_dalpha += dL_dweight - last_dL_dT;
            last_dL_dT = dL_dweight * alpha + (1 - alpha) * last_dL_dT;
            const float dL_dmd = 2.0f * (T * alpha) * (m_d * final_A - final_D) * dL_dreg;
            dL_dz += dL_dmd * dmd_dd;
            accum_depth_rec = last_alpha * last_depth + (1.f - last_alpha) * accum_depth_rec;
            last_depth = c_d;
            dL_dalpha += (c_d - accum_depth_rec) * dL_ddepth;
            accum_alpha_rec = last_alpha * 1.0f + (1.f - last_alpha) * accum_alpha_rec;
            dL_dalpha += (1 - accum_alpha_rec) * dL_daccum;
            for (int ch = 0; ch < 3; ++ch) {
                accum_normal_rec[ch] = last_alpha * last_normal[ch] + (1.f - last_alpha) * accum_normal_rec[ch];
                last_normal[ch] = no[ch];
                dL_dalpha += (no[ch] - accum_normal_rec[ch]) * dL_dnormal2D[ch];
                dadd_(a + 15 + ch, alpha * T * dL_dnormal2D[ch]);
            }
            dL_dalpha *= T;
            last_alpha = alpha;
            float bg_dot_dpixel = 0;
            for (int i = 0; i < 3; ++i) bg_dot_dpixel += bg[i] * dL_dpixel[i];
            dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot_dpixel;
            const float dL_dG = no[3] * dL_dalpha;
            dL_dz += alpha * T * dL_ddepth;
            if (rho3d <= rho2d) {
                const float dLdsx = dL_dG * -G * sx + dL_dz * Tw[0], dLdsy = dL_dG * -G * sy + dL_dz * Tw[1];
                const float dsx_pz = dLdsx / p[2], dsy_pz = dLdsy / p[2];
                const float dp[3] = { dsx_pz, dsy_pz, -(dsx_pz * sx + dsy_pz * sy) };
                const float dk[3] = { lv[1] * dp[2] - lv[2] * dp[1], lv[2] * dp[0] - lv[0] * dp[2], lv[0] * dp[1] - lv[1] * dp[0] };
                const float dl[3] = { dp[1] * kv[2] - dp[2] * kv[1], dp[2] * kv[0] - dp[0] * kv[2], dp[0] * kv[1] - dp[1] * kv[0] };
                const float dz_dTw[3] = { sx, sy, 1.0f };
                for (int i = 0; i < 3; ++i) {
                    dadd_(a + 0 + i, -dk[i]);
                    dadd_(a + 3 + i, -dl[i]);
                    dadd_(a + 6 + i, pixfx * dk[i] + pixfy * dl[i] + dL_dz * dz_dTw[i]);
                }
            } else {
                const float dG_ddelx = -G * FilterInvSquare * dx, dG_ddely = -G * FilterInvSquare * dy;
                dadd_(a + 9, dL_dG * dG_ddelx);
                dadd_(a + 10, dL_dG * dG_ddely);
                dadd_(a + 6, sx * dL_dz); dadd_(a + 7, sy * dL_dz); dadd_(a + 8, dL_dz);
            }
            dadd_(a + 11, G * dL_dalpha);
        }
    }
}

/* compute_transmat_aabb, backward.cu:449-579. dL_dTs (9 per surfel) is read and, in the precomputed
 * case with a non-zero dL_dmean2D, overwritten, exactly like the reference's global array. */
static void compute_transmat_aabb(int idx, const float* Ts_precomp, const float* p_origs, const float* scales,
                                  const float* rots, const float* proj, const float* view, int W, int H,
                                  const float* dL_dnormals, const float* dL_dmean2Ds, float* dL_dTs,
                                  float* dL_dmeans, float* dL_dscales, float* dL_drots) {
    m3 T, R; vec3 normal = v3(0, 0, 0), p_orig = v3(0, 0, 0);
    float Pm[3][4]; const float* rot = NULL; float scale[2] = { 0, 0 };
    memset(&R, 0, sizeof(R)); memset(Pm, 0, sizeof(Pm));
    if (Ts_precomp) {
        for (int j = 0; j < 3; ++j) for (int i = 0; i < 3; ++i) T.m[j][i] = Ts_precomp[9 * (size_t)idx + 3 * j + i];
    } else {
        p_orig = v3(p_origs[3 * idx], p_origs[3 * idx + 1], p_origs[3 * idx + 2]);
        rot = rots + 4 * (size_t)idx; scale[0] = scales[2 * idx]; scale[1] = scales[2 * idx + 1];
        R = quat_to_rotmat(rot);
        float L[3][3];
        for (int i = 0; i < 3; ++i) { L[0][i] = R.m[0][i] * (1.0f * scale[0]); L[1][i] = R.m[1][i] * (1.0f * scale[1]); L[2][i] = R.m[2][i]; }
        proj_to_pix(proj, W, H, Pm);
        const float Mt[4][3] = { { L[0][0], L[1][0], p_orig.x }, { L[0][1], L[1][1], p_orig.y }, { L[0][2], L[1][2], p_orig.z }, { 0, 0, 1 } };
        for (int j = 0; j < 3; ++j) for (int r = 0; r < 3; ++r)
            T.m[j][r] = Mt[0][r] * Pm[j][0] + Mt[1][r] * Pm[j][1] + Mt[2][r] * Pm[j][2] + Mt[3][r] * Pm[j][3];
        normal = v3(view[0] * L[2][0] + view[4] * L[2][1] + view[8] * L[2][2],
                    view[1] * L[2][0] + view[5] * L[2][1] + view[9] * L[2][2],
                    view[2] * L[2][0] + view[6] * L[2][1] + view[10] * L[2][2]);
    }
    float dT[3][3];
    for (int j = 0; j < 3; ++j) for (int i = 0; i < 3; ++i) dT[j][i] = dL_dTs[9 * (size_t)idx + 3 * j + i];
    const float dmx = dL_dmean2Ds[3 * idx], dmy = dL_dmean2Ds[3 * idx + 1];
    if (dmx != 0 || dmy != 0) {
        const float t[3] = { 9.0f, 9.0f, -1.0f };
        float d = t[0] * (T.m[2][0] * T.m[2][0]) + t[1] * (T.m[2][1] * T.m[2][1]) + t[2] * (T.m[2][2] * T.m[2][2]);
        float f[3] = { t[0] * (1.0f / d), t[1] * (1.0f / d), t[2] * (1.0f / d) };
        float dT3[3], dLdf[3];
        for (int i = 0; i < 3; ++i) {
            dT[0][i] += dmx * f[i] * T.m[2][i];
            dT[1][i] += dmy * f[i] * T.m[2][i];
            dT3[i] = dmx * f[i] * T.m[0][i] + dmy * f[i] * T.m[1][i];
            dLdf[i] = dmx * T.m[0][i] * T.m[2][i] + dmy * T.m[1][i] * T.m[2][i];
        }
        float dL_dd = (float)((dLdf[0] * f[0] + dLdf[1] * f[1] + dLdf[2] * f[2]) * (-1.0 / d));
        for (int i = 0; i < 3; ++i) { dT3[i] += dL_dd * (t[i] * T.m[2][i] * 2.0f); dT[2][i] += dT3[i]; }
        if (Ts_precomp) {
            for (int j = 0; j < 3; ++j) for (int i = 0; i < 3; ++i) dL_dTs[9 * (size_t)idx + 3 * j + i] = dT[j][i];
            return;
        }
    }
    if (Ts_precomp) return;
    float dM[3][4];
    for (int j = 0; j < 3; ++j) for (int i = 0; i < 4; ++i)
        dM[j][i] = Pm[0][i] * dT[0][j] + Pm[1][i] * dT[1][j] + Pm[2][i] * dT[2][j];
    vec3 dn = v3(dL_dnormals[3 * idx], dL_dnormals[3 * idx + 1], dL_dnormals[3 * idx + 2]);
    vec3 dtn = transformVec4x3Transpose(dn, view);
    vec3 p_view = transformPoint4x3(p_orig, view);
    float cosv = -(p_view.x * normal.x + p_view.y * normal.y + p_view.z * normal.z);
    float mult = cosv > 0 ? 1.0f : -1.0f;
    dtn = v3scale(mult, dtn);
    const float vR[3][3] = { { dM[0][0] * scale[0], dM[0][1] * scale[0], dM[0][2] * scale[0] },
                             { dM[1][0] * scale[1], dM[1][1] * scale[1], dM[1][2] * scale[1] },
                             { dtn.x, dtn.y, dtn.z } };
    /* quat_to_rotmat_vjp, auxiliary.h:238-282 */
    float s = 1.0f / sqrtf(rot[3] * rot[3] + rot[0] * rot[0] + rot[1] * rot[1] + rot[2] * rot[2]);
    float w = rot[0] * s, x = rot[1] * s, y = rot[2] * s, z = rot[3] * s;
    float* dq = dL_drots + 4 * (size_t)idx;
    dq[0] = 2.f * (x * (vR[1][2] - vR[2][1]) + y * (vR[2][0] - vR[0][2]) + z * (vR[0][1] - vR[1][0]));
    dq[1] = 2.f * (-2.f * x * (vR[1][1] + vR[2][2]) + y * (vR[0][1] + vR[1][0]) + z * (vR[0][2] + vR[2][0]) + w * (vR[1][2] - vR[2][1]));
    dq[2] = 2.f * (x * (vR[0][1] + vR[1][0]) - 2.f * y * (vR[0][0] + vR[2][2]) + z * (vR[1][2] + vR[2][1]) + w * (vR[2][0] - vR[0][2]));
    dq[3] = 2.f * (x * (vR[0][2] + vR[2][0]) + y * (vR[1][2] + vR[2][1]) - 2.f * z * (vR[0][0] + vR[1][1]) + w * (vR[0][1] - vR[1][0]));
    dL_dscales[2 * idx] = dM[0][0] * R.m[0][0] + dM[0][1] * R.m[0][1] + dM[0][2] * R.m[0][2];
    dL_dscales[2 * idx + 1] = dM[1][0] * R.m[1][0] + dM[1][1] * R.m[1][1] + dM[1][2] * R.m[1][2];
    dL_dmeans[3 * idx] = dM[2][0]; dL_dmeans[3 * idx + 1] = dM[2][1]; dL_dmeans[3 * idx + 2] = dM[2][2];
}

/* Rasterizer::backward, rasterizer_impl.cu:331-448, on top of the nine zero-filled tensors of the glue */
void surfel_oracle_backward(const SurfelOracle* s, int D, int M, const float* bg,
                            const float* means3D, const float* shs, const float* colors_precomp,
                            const float* scales, float scale_modifier, const float* rotations,
                            const float* transMat_precomp, const float* view, const float* proj, const float* campos,
                            float tan_fovx, float tan_fovy, const float* dL_dpix, const float* dL_depths,
                            float* dL_dmean2D /*P*3*/, float* dL_dnormal /*P*3*/, float* dL_dopacity, float* dL_dcolor,
                            float* dL_dmean3D, float* dL_dtransMat /*P*9*/, float* dL_dsh, float* dL_dscale /*P*2*/,
                            float* dL_drot) {
    (void)scale_modifier;
    const int P = s->P, W = s->W, H = s->H;
    const float focal_y = H / (2.0f * tan_fovy), focal_x = W / (2.0f * tan_fovx);
    memset(dL_dmean2D, 0, (size_t)P * 12); memset(dL_dnormal, 0, (size_t)P * 12); memset(dL_dopacity, 0, (size_t)P * 4);
    memset(dL_dcolor, 0, (size_t)P * 12); memset(dL_dmean3D, 0, (size_t)P * 12); memset(dL_dtransMat, 0, (size_t)P * 36);
    if (dL_dsh && M) memset(dL_dsh, 0, (size_t)P * M * 12);
    memset(dL_dscale, 0, (size_t)P * 8); memset(dL_drot, 0, (size_t)P * 16);
    double* acc = (double*)calloc((size_t)(P > 0 ? P : 1) * 18, sizeof(double));
    const float* color_ptr = colors_precomp ? colors_precomp : s->rgb;
    const float* tm = transMat_precomp ? transMat_precomp : s->transMat;
#pragma omp parallel for schedule(dynamic, 4)
    for (int t = 0; t < s->tiles; ++t) surfel_render_tile_backward(s, t, bg, color_ptr, tm, dL_dpix, dL_depths, acc);
    for (int i = 0; i < P; ++i) {
        const double* a = acc + 18 * (size_t)i;
        for (int k = 0; k < 9; ++k) dL_dtransMat[9 * (size_t)i + k] = (float)a[k];
        dL_dmean2D[3 * i] = (float)a[9]; dL_dmean2D[3 * i + 1] = (float)a[10];
        dL_dopacity[i] = (float)a[11];
        for (int k = 0; k < 3; ++k) { dL_dcolor[3 * i + k] = (float)a[12 + k]; dL_dnormal[3 * i + k] = (float)a[15 + k]; }
    }
    free(acc);
    /* preprocessCUDA backward, backward.cu:582-636 */
    const int Wb = (int)(focal_x * tan_fovx * 2), Hb = (int)(focal_y * tan_fovy * 2);
    const float* Ts_precomp = scales ? NULL : tm;
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; ++idx) {
        if (!(s->radii[idx] > 0)) continue;
        compute_transmat_aabb(idx, Ts_precomp, means3D, scales, rotations, proj, view, Wb, Hb, dL_dnormal, dL_dmean2D,
                              dL_dtransMat, dL_dmean3D, dL_dscale, dL_drot);
        if (shs) orc_sh_backward(idx, D, M, means3D, campos, shs, s->clamped, dL_dcolor, dL_dmean3D, dL_dsh);
        float depth = tm[9 * (size_t)idx + 8];
        dL_dmean2D[3 * idx] = (float)(dL_dtransMat[9 * (size_t)idx + 2] * depth * 0.5 * (float)Wb);
        dL_dmean2D[3 * idx + 1] = (float)(dL_dtransMat[9 * (size_t)idx + 5] * depth * 0.5 * (float)Hb);
    }
}
