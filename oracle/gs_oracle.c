/* gs_oracle.c -- TEST INFRASTRUCTURE (checker): plain-C CPU restatement of the reference 3DGS
 * rasterizer (forward + backward). Never linked into, imported by or called from scorp_b200/.
 * Callers: tests/, __graft_entry__.smoke(), bench.py's cpu_baseline leg.
 *
 * Follows DGR = /root/reference/submodules/diff-gaussian-rasterization/cuda_rasterizer:
 *   forward.cu:20-71 (SH), :74-113 (cov2D), :118-152 (cov3D), :156-256 (preprocess), :263-381 (blend)
 *   rasterizer_impl.cu:70-138 (keys, ranges), :197-339 (orchestration)
 *   backward.cu:20-139 (SH bwd), :144-274 (cov2D bwd), :278-341 (cov3D bwd), :347-412 (preprocess bwd),
 *               :417-601 (blend bwd)
 * Parity status: PINNED by tests/golden/ (outputs of the reference CUDA build on a B200).
 * Gradient sums are accumulated in double (the reference's float atomicAdd order is
 * non-deterministic; a double sum is the value its runs scatter around).
 */
#include "oracle_common.h"
#include <stdio.h>

/* 3x3 matrices in glm's layout: m[col][row] */
typedef struct { float m[3][3]; } mat3;

static mat3 mat3_mul(mat3 a, mat3 b) {   /* glm: (a*b)[j][i] = sum_k a[k][i] * b[j][k] */
    mat3 r;
    for (int j = 0; j < 3; ++j)
        for (int i = 0; i < 3; ++i)
            r.m[j][i] = a.m[0][i] * b.m[j][0] + a.m[1][i] * b.m[j][1] + a.m[2][i] * b.m[j][2];
    return r;
}
static mat3 mat3_transpose(mat3 a) {
    mat3 r;
    for (int j = 0; j < 3; ++j) for (int i = 0; i < 3; ++i) r.m[j][i] = a.m[i][j];
    return r;
}

typedef struct GsOracle {
    int P, W, H, gx, gy, tiles;
    int64_t N;
    float *depths, *means2D, *cov3D, *conic_opacity, *rgb;
    uint8_t* clamped;          /* 3 per Gaussian */
    uint32_t *tiles_touched, *point_offsets;
    uint64_t* keys;            /* sorted (tile << 32 | depth bits) */
    uint32_t* point_list;      /* sorted Gaussian indices */
    uint32_t* ranges;          /* 2 per tile */
    uint32_t* n_contrib;       /* per pixel */
    int* radii;
} GsOracle;

/* forward.cu:20-71 */
vec3 orc_computeColorFromSH(int idx, int deg, int max_coeffs, const float* means, const float* campos,
                               const float* shs, uint8_t* clamped) {
    vec3 pos = v3(means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]);
    vec3 dir = v3(pos.x - campos[0], pos.y - campos[1], pos.z - campos[2]);
    float len = sqrtf(v3dot(dir, dir));
    dir = v3(dir.x / len, dir.y / len, dir.z / len);
    const vec3* sh = (const vec3*)shs + (size_t)idx * max_coeffs;
    vec3 result = v3scale(SH_C0, sh[0]);
    if (deg > 0) {
        float x = dir.x, y = dir.y, z = dir.z;
        result = v3add(v3add(v3add(result, v3scale(-SH_C1 * y, sh[1])), v3scale(SH_C1 * z, sh[2])), v3scale(-SH_C1 * x, sh[3]));
        if (deg > 1) {
            float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            result = v3add(result, v3scale(SH_C2[0] * xy, sh[4]));
            result = v3add(result, v3scale(SH_C2[1] * yz, sh[5]));
            result = v3add(result, v3scale(SH_C2[2] * (2.0f * zz - xx - yy), sh[6]));
            result = v3add(result, v3scale(SH_C2[3] * xz, sh[7]));
            result = v3add(result, v3scale(SH_C2[4] * (xx - yy), sh[8]));
            if (deg > 2) {
                result = v3add(result, v3scale(SH_C3[0] * y * (3.0f * xx - yy), sh[9]));
                result = v3add(result, v3scale(SH_C3[1] * xy * z, sh[10]));
                result = v3add(result, v3scale(SH_C3[2] * y * (4.0f * zz - xx - yy), sh[11]));
                result = v3add(result, v3scale(SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy), sh[12]));
                result = v3add(result, v3scale(SH_C3[4] * x * (4.0f * zz - xx - yy), sh[13]));
                result = v3add(result, v3scale(SH_C3[5] * z * (xx - yy), sh[14]));
                result = v3add(result, v3scale(SH_C3[6] * x * (xx - 3.0f * yy), sh[15]));
            }
        }
    }
    result.x += 0.5f; result.y += 0.5f; result.z += 0.5f;
    clamped[3 * idx + 0] = result.x < 0; clamped[3 * idx + 1] = result.y < 0; clamped[3 * idx + 2] = result.z < 0;
    return v3(result.x > 0.0f ? result.x : 0.0f, result.y > 0.0f ? result.y : 0.0f, result.z > 0.0f ? result.z : 0.0f);
}

/* forward.cu:118-152 */
static void computeCov3D(const float* scale, float mod, const float* rot, float* cov3D) {
    mat3 S; memset(&S, 0, sizeof(S));
    S.m[0][0] = mod * scale[0]; S.m[1][1] = mod * scale[1]; S.m[2][2] = mod * scale[2];
    float r = rot[0], x = rot[1], y = rot[2], z = rot[3];      /* not normalised (:127) */
    mat3 R = { { { 1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y) },
                 { 2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x) },
                 { 2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y) } } };
    mat3 M = mat3_mul(S, R);
    mat3 Sigma = mat3_mul(mat3_transpose(M), M);
    cov3D[0] = Sigma.m[0][0]; cov3D[1] = Sigma.m[0][1]; cov3D[2] = Sigma.m[0][2];
    cov3D[3] = Sigma.m[1][1]; cov3D[4] = Sigma.m[1][2]; cov3D[5] = Sigma.m[2][2];
}

/* shared by forward.cu:74-113 and backward.cu:160-197: T = W*J and the clamped view-space mean */
static mat3 ewa_T(vec3 mean, float focal_x, float focal_y, float tan_fovx, float tan_fovy, const float* view,
                  vec3* t_out, float* x_grad_mul, float* y_grad_mul) {
    vec3 t = transformPoint4x3(mean, view);
    const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
    const float txtz = t.x / t.z, tytz = t.y / t.z;
    t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
    t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;
    if (x_grad_mul) *x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
    if (y_grad_mul) *y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
    mat3 J = { { { focal_x / t.z, 0.0f, -(focal_x * t.x) / (t.z * t.z) },
                 { 0.0f, focal_y / t.z, -(focal_y * t.y) / (t.z * t.z) },
                 { 0, 0, 0 } } };
    mat3 Wm = { { { view[0], view[4], view[8] }, { view[1], view[5], view[9] }, { view[2], view[6], view[10] } } };
    if (t_out) *t_out = t;
    return mat3_mul(Wm, J);
}
static mat3 vrk_of(const float* c) {
    mat3 V = { { { c[0], c[1], c[2] }, { c[1], c[3], c[4] }, { c[2], c[4], c[5] } } };
    return V;
}
/* forward.cu:74-113 */
static vec3 computeCov2D(vec3 mean, float focal_x, float focal_y, float tan_fovx, float tan_fovy,
                         const float* cov3D, const float* view) {
    mat3 T = ewa_T(mean, focal_x, focal_y, tan_fovx, tan_fovy, view, NULL, NULL, NULL);
    mat3 cov = mat3_mul(mat3_mul(mat3_transpose(T), mat3_transpose(vrk_of(cov3D))), T);
    cov.m[0][0] += 0.3f; cov.m[1][1] += 0.3f;
    return v3(cov.m[0][0], cov.m[0][1], cov.m[1][1]);
}

static uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

void gs_oracle_free(GsOracle* s) {
    if (!s) return;
    free(s->depths); free(s->means2D); free(s->cov3D); free(s->conic_opacity); free(s->rgb); free(s->clamped);
    free(s->tiles_touched); free(s->point_offsets); free(s->keys); free(s->point_list); free(s->ranges);
    free(s->n_contrib); free(s->radii); free(s);
}

/* forward.cu:263-381, one tile; pixels are visited sequentially (their results are independent) */
static void render_tile(const GsOracle* s, int tile, const float* features, const float* bg,
                        float* out_color, float* out_depth, float* out_alpha) {
    const int W = s->W, H = s->H;
    const int tx = tile % s->gx, ty = tile / s->gx;
    const uint32_t r0 = s->ranges[2 * tile], r1 = s->ranges[2 * tile + 1];
    for (int ly = 0; ly < BLOCK_Y; ++ly) for (int lx = 0; lx < BLOCK_X; ++lx) {
        const int px = tx * BLOCK_X + lx, py = ty * BLOCK_Y + ly;
        if (px >= W || py >= H) continue;
        const uint32_t pix_id = (uint32_t)W * py + px;
        const float pixfx = (float)px, pixfy = (float)py;
        float T = 1.0f, C[NUM_CHANNELS] = { 0, 0, 0 }, weight = 0, D = 0;
        uint32_t contributor = 0, last_contributor = 0;
        for (uint32_t k = r0; k < r1; ++k) {
            contributor++;
            const uint32_t id = s->point_list[k];
            const float dx = s->means2D[2 * id] - pixfx, dy = s->means2D[2 * id + 1] - pixfy;
            const float* co = s->conic_opacity + 4 * (size_t)id;
            const float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
            if (power > 0.0f) continue;
            const float alpha = fminf(0.99f, co[3] * expf(power));
            if (alpha < 1.0f / 255.0f) continue;
            const float test_T = T * (1 - alpha);
            if (test_T < 0.0001f) break;                        /* done = true (:349-353) */
            for (int ch = 0; ch < NUM_CHANNELS; ++ch) C[ch] += features[id * NUM_CHANNELS + ch] * alpha * T;
            weight += alpha * T;
            D += s->depths[id] * alpha * T;
            T = test_T;
            last_contributor = contributor;
        }
        s->n_contrib[pix_id] = last_contributor;
        for (int ch = 0; ch < NUM_CHANNELS; ++ch) out_color[(size_t)ch * H * W + pix_id] = C[ch] + T * bg[ch];
        out_alpha[pix_id] = weight;
        out_depth[pix_id] = D;
    }
}

/* rasterizer_impl.cu:197-339. Returns a state object holding every intermediate (NULL on error). */
GsOracle* gs_oracle_forward(int P, int D, int M, const float* bg, int W, int H,
                            const float* means3D, const float* shs, const float* colors_precomp,
                            const float* opacities, const float* scales, float scale_modifier,
                            const float* rotations, const float* cov3D_precomp,
                            const float* view, const float* proj, const float* campos,
                            float tan_fovx, float tan_fovy,
                            float* out_color, float* out_depth, float* out_alpha, int* radii_out) {
    if (colors_precomp == NULL && shs == NULL) return NULL;     /* rasterizer_impl.cu:243-246 */
    GsOracle* s = (GsOracle*)calloc(1, sizeof(GsOracle));
    const float focal_y = H / (2.0f * tan_fovy), focal_x = W / (2.0f * tan_fovx);   /* :224-225 */
    s->P = P; s->W = W; s->H = H;
    s->gx = (W + BLOCK_X - 1) / BLOCK_X; s->gy = (H + BLOCK_Y - 1) / BLOCK_Y; s->tiles = s->gx * s->gy;
    size_t Pn = P > 0 ? (size_t)P : 1;
    s->depths = calloc(Pn, 4); s->means2D = calloc(Pn, 8); s->cov3D = calloc(Pn, 24);
    s->conic_opacity = calloc(Pn, 16); s->rgb = calloc(Pn, 12); s->clamped = calloc(Pn, 3);
    s->tiles_touched = calloc(Pn, 4); s->point_offsets = calloc(Pn, 4); s->radii = calloc(Pn, 4);
    s->ranges = calloc((size_t)s->tiles, 8); s->n_contrib = calloc((size_t)W * H, 4);

    /* preprocessCUDA, forward.cu:156-256 */
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; ++idx) {
        vec3 p_orig = v3(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]);
        vec3 p_view = transformPoint4x3(p_orig, view);
        if (p_view.z <= 0.2f) continue;                          /* auxiliary.h:154 */
        vec4 p_hom = transformPoint4x4(p_orig, proj);
        float p_w = 1.0f / (p_hom.w + 0.0000001f);
        vec3 p_proj = v3(p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w);
        const float* cov3D;
        if (cov3D_precomp) cov3D = cov3D_precomp + 6 * (size_t)idx;
        else { computeCov3D(scales + 3 * (size_t)idx, scale_modifier, rotations + 4 * (size_t)idx, s->cov3D + 6 * (size_t)idx);
               cov3D = s->cov3D + 6 * (size_t)idx; }
        vec3 cov = computeCov2D(p_orig, focal_x, focal_y, tan_fovx, tan_fovy, cov3D, view);
        float det = (cov.x * cov.z - cov.y * cov.y);
        if (det == 0.0f) continue;
        float det_inv = 1.f / det;
        vec3 conic = v3(cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv);
        float mid = 0.5f * (cov.x + cov.z);
        float lambda1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
        float lambda2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
        float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
        float pix_x = ndc2Pix(p_proj.x, W), pix_y = ndc2Pix(p_proj.y, H);
        uint32_t x0, y0, x1, y1;
        getRect(pix_x, pix_y, (int)my_radius, s->gx, s->gy, &x0, &y0, &x1, &y1);
        if ((x1 - x0) * (y1 - y0) == 0) continue;
        if (colors_precomp == NULL) {
            vec3 c = orc_computeColorFromSH(idx, D, M, means3D, campos, shs, s->clamped);
            s->rgb[3 * idx] = c.x; s->rgb[3 * idx + 1] = c.y; s->rgb[3 * idx + 2] = c.z;
        }
        s->depths[idx] = p_view.z;
        s->radii[idx] = (int)my_radius;
        s->means2D[2 * idx] = pix_x; s->means2D[2 * idx + 1] = pix_y;
        float* co = s->conic_opacity + 4 * (size_t)idx;
        co[0] = conic.x; co[1] = conic.y; co[2] = conic.z; co[3] = opacities[idx];
        s->tiles_touched[idx] = (y1 - y0) * (x1 - x0);
    }
    /* InclusiveSum, rasterizer_impl.cu:278-282 */
    uint32_t run = 0;
    for (int i = 0; i < P; ++i) { run += s->tiles_touched[i]; s->point_offsets[i] = run; }
    s->N = run;
    size_t Nn = run ? run : 1;
    s->keys = malloc(Nn * 8); s->point_list = malloc(Nn * 4);
    /* duplicateWithKeys, rasterizer_impl.cu:70-111 */
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; ++idx) {
        if (s->radii[idx] <= 0) continue;
        uint32_t off = idx == 0 ? 0 : s->point_offsets[idx - 1];
        uint32_t x0, y0, x1, y1;
        getRect(s->means2D[2 * idx], s->means2D[2 * idx + 1], s->radii[idx], s->gx, s->gy, &x0, &y0, &x1, &y1);
        for (uint32_t y = y0; y < y1; ++y) for (uint32_t x = x0; x < x1; ++x) {
            uint64_t key = (uint64_t)(y * (uint32_t)s->gx + x);
            key <<= 32; key |= f2u(s->depths[idx]);
            s->keys[off] = key; s->point_list[off] = (uint32_t)idx; off++;
        }
    }
    /* SortPairs on bits [0, 32+bit), rasterizer_impl.cu:299-309 */
    sort_pairs_stable(s->keys, s->point_list, (size_t)run, 32 + (int)getHigherMsb((uint32_t)s->tiles));
    /* identifyTileRanges, rasterizer_impl.cu:116-138 (ranges zeroed at :311) */
    for (uint32_t i = 0; i < run; ++i) {
        uint32_t cur = (uint32_t)(s->keys[i] >> 32);
        if (i == 0) s->ranges[2 * cur] = 0;
        else { uint32_t prev = (uint32_t)(s->keys[i - 1] >> 32);
               if (cur != prev) { s->ranges[2 * prev + 1] = i; s->ranges[2 * cur] = i; } }
        if (i == run - 1) s->ranges[2 * cur + 1] = run;
    }
    /* renderCUDA */
    const float* features = colors_precomp ? colors_precomp : s->rgb;   /* :322 */
#pragma omp parallel for schedule(dynamic, 4)
    for (int t = 0; t < s->tiles; ++t) render_tile(s, t, features, bg, out_color, out_depth, out_alpha);
    if (radii_out) memcpy(radii_out, s->radii, (size_t)P * 4);
    return s;
}

int64_t gs_oracle_num_rendered(const GsOracle* s) { return s->N; }
/* which: 0 depths 1 means2D 2 cov3D 3 conic_opacity 4 rgb 5 clamped 6 tiles_touched 7 point_offsets
 *        8 keys 9 point_list 10 ranges 11 n_contrib */
const void* gs_oracle_array(const GsOracle* s, int which) {
    switch (which) {
    case 0: return s->depths; case 1: return s->means2D; case 2: return s->cov3D; case 3: return s->conic_opacity;
    case 4: return s->rgb; case 5: return s->clamped; case 6: return s->tiles_touched; case 7: return s->point_offsets;
    case 8: return s->keys; case 9: return s->point_list; case 10: return s->ranges; case 11: return s->n_contrib;
    default: return NULL;
    }
}

/* ---------------------------------------------------------------------------------------------
 * Backward. Per-Gaussian sums from the blend are accumulated in double (acc: 10 per Gaussian:
 * mean2D.x, mean2D.y, conic.x, conic.y, conic.w, opacity, color[3], depth).
 * ------------------------------------------------------------------------------------------- */
static inline void dadd(double* p, double v) {
#pragma omp atomic
    *p += v;
}

/* backward.cu:417-601, one tile */
static void render_tile_backward(const GsOracle* s, int tile, const float* bg, const float* colors,
                                 const float* alphas, const float* dL_dpixels, const float* dL_dpixel_depths,
                                 const float* dL_dalphas, double* acc) {
    const int W = s->W, H = s->H;
    const int tx = tile % s->gx, ty = tile / s->gx;
    const uint32_t r0 = s->ranges[2 * tile], r1 = s->ranges[2 * tile + 1];
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
    for (int ly = 0; ly < BLOCK_Y; ++ly) for (int lx = 0; lx < BLOCK_X; ++lx) {
        const int px = tx * BLOCK_X + lx, py = ty * BLOCK_Y + ly;
        if (px >= W || py >= H) continue;
        const uint32_t pix_id = (uint32_t)W * py + px;
        const float pixfx = (float)px, pixfy = (float)py;
        const float T_final = 1 - alphas[pix_id];                       /* :463 */
        float T = T_final;
        uint32_t contributor = r1 - r0;
        const uint32_t last_contributor = s->n_contrib[pix_id];
        float accum_rec[3] = { 0, 0, 0 }, dL_dpixel[3], accum_depth_rec = 0, accum_alpha_rec = 0;
        for (int i = 0; i < 3; ++i) dL_dpixel[i] = dL_dpixels[(size_t)i * H * W + pix_id];
        const float dL_dpixel_depth = dL_dpixel_depths[pix_id], dL_dalpha = dL_dalphas[pix_id];
        float last_alpha = 0, last_color[3] = { 0, 0, 0 }, last_depth = 0;
        for (uint32_t k = r1; k-- > r0;) {                              /* back to front (:502) */
            contributor--;
            if (contributor >= last_contributor) continue;
            const uint32_t id = s->point_list[k];
            const float dx = s->means2D[2 * id] - pixfx, dy = s->means2D[2 * id + 1] - pixfy;
            const float* co = s->conic_opacity + 4 * (size_t)id;
            const float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
            if (power > 0.0f) continue;
            const float G = expf(power);
            const float alpha = fminf(0.99f, co[3] * G);
            if (alpha < 1.0f / 255.0f) continue;
            T = T / (1.f - alpha);
            const float dchannel_dcolor = alpha * T;
            double* a = acc + 10 * (size_t)id;
            float dL_dopa = 0.0f;
            for (int ch = 0; ch < 3; ++ch) {
                const float c = colors[id * 3 + ch];
                accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                last_color[ch] = c;
                dL_dopa += (c - accum_rec[ch]) * dL_dpixel[ch];
                dadd(a + 6 + ch, dchannel_dcolor * dL_dpixel[ch]);
            }
            const float c_d = s->depths[id];
            accum_depth_rec = last_alpha * last_depth + (1.f - last_alpha) * accum_depth_rec;
            last_depth = c_d;
            dL_dopa += (c_d - accum_depth_rec) * dL_dpixel_depth;
            dadd(a + 9, dchannel_dcolor * dL_dpixel_depth);
            accum_alpha_rec = last_alpha + (1.f - last_alpha) * accum_alpha_rec;
            dL_dopa += (1 - accum_alpha_rec) * dL_dalpha;
            dL_dopa *= T;
            last_alpha = alpha;
            float bg_dot_dpixel = 0;
            for (int i = 0; i < 3; ++i) bg_dot_dpixel += bg[i] * dL_dpixel[i];
            dL_dopa += (-T_final / (1.f - alpha)) * bg_dot_dpixel;
            const float dL_dG = co[3] * dL_dopa;
            const float gdx = G * dx, gdy = G * dy;
            const float dG_ddelx = -gdx * co[0] - gdy * co[1];
            const float dG_ddely = -gdy * co[2] - gdx * co[1];
            dadd(a + 0, dL_dG * dG_ddelx * ddelx_dx);
            dadd(a + 1, dL_dG * dG_ddely * ddely_dy);
            dadd(a + 2, -0.5f * gdx * dx * dL_dG);
            dadd(a + 3, -0.5f * gdx * dy * dL_dG);
            dadd(a + 4, -0.5f * gdy * dy * dL_dG);
            dadd(a + 5, G * dL_dopa);
        }
    }
}

/* backward.cu:20-139 */
void orc_sh_backward(int idx, int deg, int max_coeffs, const float* means, const float* campos, const float* shs,
                        const uint8_t* clamped, const float* dL_dcolor, float* dL_dmeans, float* dL_dshs) {
    vec3 pos = v3(means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]);
    vec3 dir_orig = v3(pos.x - campos[0], pos.y - campos[1], pos.z - campos[2]);
    float len = sqrtf(v3dot(dir_orig, dir_orig));
    vec3 dir = v3(dir_orig.x / len, dir_orig.y / len, dir_orig.z / len);
    const vec3* sh = (const vec3*)shs + (size_t)idx * max_coeffs;
    vec3 dL_dRGB = v3(dL_dcolor[3 * idx], dL_dcolor[3 * idx + 1], dL_dcolor[3 * idx + 2]);
    dL_dRGB.x *= clamped[3 * idx + 0] ? 0 : 1; dL_dRGB.y *= clamped[3 * idx + 1] ? 0 : 1; dL_dRGB.z *= clamped[3 * idx + 2] ? 0 : 1;
    vec3 dRGBdx = v3(0, 0, 0), dRGBdy = v3(0, 0, 0), dRGBdz = v3(0, 0, 0);
    float x = dir.x, y = dir.y, z = dir.z;
    vec3* dL_dsh = (vec3*)dL_dshs + (size_t)idx * max_coeffs;
    dL_dsh[0] = v3scale(SH_C0, dL_dRGB);
    if (deg > 0) {
        dL_dsh[1] = v3scale(-SH_C1 * y, dL_dRGB); dL_dsh[2] = v3scale(SH_C1 * z, dL_dRGB); dL_dsh[3] = v3scale(-SH_C1 * x, dL_dRGB);
        dRGBdx = v3scale(-SH_C1, sh[3]); dRGBdy = v3scale(-SH_C1, sh[1]); dRGBdz = v3scale(SH_C1, sh[2]);
        if (deg > 1) {
            float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            dL_dsh[4] = v3scale(SH_C2[0] * xy, dL_dRGB); dL_dsh[5] = v3scale(SH_C2[1] * yz, dL_dRGB);
            dL_dsh[6] = v3scale(SH_C2[2] * (2.f * zz - xx - yy), dL_dRGB);
            dL_dsh[7] = v3scale(SH_C2[3] * xz, dL_dRGB); dL_dsh[8] = v3scale(SH_C2[4] * (xx - yy), dL_dRGB);
            dRGBdx = v3add(dRGBdx, v3add(v3add(v3add(v3scale(SH_C2[0] * y, sh[4]), v3scale(SH_C2[2] * 2.f * -x, sh[6])),
                                               v3scale(SH_C2[3] * z, sh[7])), v3scale(SH_C2[4] * 2.f * x, sh[8])));
            dRGBdy = v3add(dRGBdy, v3add(v3add(v3add(v3scale(SH_C2[0] * x, sh[4]), v3scale(SH_C2[1] * z, sh[5])),
                                               v3scale(SH_C2[2] * 2.f * -y, sh[6])), v3scale(SH_C2[4] * 2.f * -y, sh[8])));
            dRGBdz = v3add(dRGBdz, v3add(v3add(v3scale(SH_C2[1] * y, sh[5]), v3scale(SH_C2[2] * 2.f * 2.f * z, sh[6])),
                                         v3scale(SH_C2[3] * x, sh[7])));
            if (deg > 2) {
                dL_dsh[9] = v3scale(SH_C3[0] * y * (3.f * xx - yy), dL_dRGB);
                dL_dsh[10] = v3scale(SH_C3[1] * xy * z, dL_dRGB);
                dL_dsh[11] = v3scale(SH_C3[2] * y * (4.f * zz - xx - yy), dL_dRGB);
                dL_dsh[12] = v3scale(SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy), dL_dRGB);
                dL_dsh[13] = v3scale(SH_C3[4] * x * (4.f * zz - xx - yy), dL_dRGB);
                dL_dsh[14] = v3scale(SH_C3[5] * z * (xx - yy), dL_dRGB);
                dL_dsh[15] = v3scale(SH_C3[6] * x * (xx - 3.f * yy), dL_dRGB);
                vec3 ax = v3(0, 0, 0), ay = v3(0, 0, 0), az = v3(0, 0, 0);
                ax = v3add(ax, v3scale(3.f * 2.f * xy, v3scale(SH_C3[0], sh[9])));
                ax = v3add(ax, v3scale(yz, v3scale(SH_C3[1], sh[10])));
                ax = v3add(ax, v3scale(xy, v3scale(-2.f, v3scale(SH_C3[2], sh[11]))));
                ax = v3add(ax, v3scale(xz, v3scale(2.f, v3scale(-3.f, v3scale(SH_C3[3], sh[12])))));
                ax = v3add(ax, v3scale(-3.f * xx + 4.f * zz - yy, v3scale(SH_C3[4], sh[13])));
                ax = v3add(ax, v3scale(xz, v3scale(2.f, v3scale(SH_C3[5], sh[14]))));
                ax = v3add(ax, v3scale(xx - yy, v3scale(3.f, v3scale(SH_C3[6], sh[15]))));
                ay = v3add(ay, v3scale(xx - yy, v3scale(3.f, v3scale(SH_C3[0], sh[9]))));
                ay = v3add(ay, v3scale(xz, v3scale(SH_C3[1], sh[10])));
                ay = v3add(ay, v3scale(-3.f * yy + 4.f * zz - xx, v3scale(SH_C3[2], sh[11])));
                ay = v3add(ay, v3scale(yz, v3scale(2.f, v3scale(-3.f, v3scale(SH_C3[3], sh[12])))));
                ay = v3add(ay, v3scale(xy, v3scale(-2.f, v3scale(SH_C3[4], sh[13]))));
                ay = v3add(ay, v3scale(yz, v3scale(-2.f, v3scale(SH_C3[5], sh[14]))));
                ay = v3add(ay, v3scale(xy, v3scale(2.f, v3scale(-3.f, v3scale(SH_C3[6], sh[15])))));
                az = v3add(az, v3scale(xy, v3scale(SH_C3[1], sh[10])));
                az = v3add(az, v3scale(yz, v3scale(2.f, v3scale(4.f, v3scale(SH_C3[2], sh[11])))));
                az = v3add(az, v3scale(2.f * zz - xx - yy, v3scale(3.f, v3scale(SH_C3[3], sh[12]))));
                az = v3add(az, v3scale(xz, v3scale(2.f, v3scale(4.f, v3scale(SH_C3[4], sh[13])))));
                az = v3add(az, v3scale(xx - yy, v3scale(SH_C3[5], sh[14])));
                dRGBdx = v3add(dRGBdx, ax); dRGBdy = v3add(dRGBdy, ay); dRGBdz = v3add(dRGBdz, az);
            }
        }
    }
    vec3 dL_ddir = v3(v3dot(dRGBdx, dL_dRGB), v3dot(dRGBdy, dL_dRGB), v3dot(dRGBdz, dL_dRGB));
    vec3 dm = dnormvdv(dir_orig, dL_ddir);
    dL_dmeans[3 * idx] += dm.x; dL_dmeans[3 * idx + 1] += dm.y; dL_dmeans[3 * idx + 2] += dm.z;
}

/* backward.cu:144-274 */
static void cov2d_backward(int idx, const float* means, const float* cov3Ds, float h_x, float h_y, float tan_fovx,
                           float tan_fovy, const float* view, const float* dL_dconics /*4 per*/,
                           float* dL_dmeans, float* dL_dcov) {
    const float* cov3D = cov3Ds + 6 * (size_t)idx;
    vec3 mean = v3(means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]);
    vec3 dL_dconic = v3(dL_dconics[4 * idx], dL_dconics[4 * idx + 1], dL_dconics[4 * idx + 3]);
    vec3 t; float x_grad_mul, y_grad_mul;
    mat3 T = ewa_T(mean, h_x, h_y, tan_fovx, tan_fovy, view, &t, &x_grad_mul, &y_grad_mul);
    mat3 Vrk = vrk_of(cov3D);
    mat3 Wm = { { { view[0], view[4], view[8] }, { view[1], view[5], view[9] }, { view[2], view[6], view[10] } } };
    mat3 cov2D = mat3_mul(mat3_mul(mat3_transpose(T), mat3_transpose(Vrk)), T);
    float a = cov2D.m[0][0] += 0.3f, b = cov2D.m[0][1], c = cov2D.m[1][1] += 0.3f;
    float denom = a * c - b * b;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    float* dc = dL_dcov + 6 * (size_t)idx;
#define TT(i, j) T.m[i][j]
    if (denom2inv != 0) {
        dL_da = denom2inv * (-c * c * dL_dconic.x + 2 * b * c * dL_dconic.y + (denom - a * c) * dL_dconic.z);
        dL_dc = denom2inv * (-a * a * dL_dconic.z + 2 * a * b * dL_dconic.y + (denom - a * c) * dL_dconic.x);
        dL_db = denom2inv * 2 * (b * c * dL_dconic.x - (denom + 2 * b * b) * dL_dconic.y + a * b * dL_dconic.z);
        dc[0] = (TT(0,0) * TT(0,0) * dL_da + TT(0,0) * TT(1,0) * dL_db + TT(1,0) * TT(1,0) * dL_dc);
        dc[3] = (TT(0,1) * TT(0,1) * dL_da + TT(0,1) * TT(1,1) * dL_db + TT(1,1) * TT(1,1) * dL_dc);
        dc[5] = (TT(0,2) * TT(0,2) * dL_da + TT(0,2) * TT(1,2) * dL_db + TT(1,2) * TT(1,2) * dL_dc);
        dc[1] = 2 * TT(0,0) * TT(0,1) * dL_da + (TT(0,0) * TT(1,1) + TT(0,1) * TT(1,0)) * dL_db + 2 * TT(1,0) * TT(1,1) * dL_dc;
        dc[2] = 2 * TT(0,0) * TT(0,2) * dL_da + (TT(0,0) * TT(1,2) + TT(0,2) * TT(1,0)) * dL_db + 2 * TT(1,0) * TT(1,2) * dL_dc;
        dc[4] = 2 * TT(0,2) * TT(0,1) * dL_da + (TT(0,1) * TT(1,2) + TT(0,2) * TT(1,1)) * dL_db + 2 * TT(1,1) * TT(1,2) * dL_dc;
    } else {
        for (int i = 0; i < 6; ++i) dc[i] = 0;
    }
#define VV(i, j) Vrk.m[i][j]
    float dL_dT00 = 2 * (TT(0,0) * VV(0,0) + TT(0,1) * VV(0,1) + TT(0,2) * VV(0,2)) * dL_da + (TT(1,0) * VV(0,0) + TT(1,1) * VV(0,1) + TT(1,2) * VV(0,2)) * dL_db;
    float dL_dT01 = 2 * (TT(0,0) * VV(1,0) + TT(0,1) * VV(1,1) + TT(0,2) * VV(1,2)) * dL_da + (TT(1,0) * VV(1,0) + TT(1,1) * VV(1,1) + TT(1,2) * VV(1,2)) * dL_db;
    float dL_dT02 = 2 * (TT(0,0) * VV(2,0) + TT(0,1) * VV(2,1) + TT(0,2) * VV(2,2)) * dL_da + (TT(1,0) * VV(2,0) + TT(1,1) * VV(2,1) + TT(1,2) * VV(2,2)) * dL_db;
    float dL_dT10 = 2 * (TT(1,0) * VV(0,0) + TT(1,1) * VV(0,1) + TT(1,2) * VV(0,2)) * dL_dc + (TT(0,0) * VV(0,0) + TT(0,1) * VV(0,1) + TT(0,2) * VV(0,2)) * dL_db;
    float dL_dT11 = 2 * (TT(1,0) * VV(1,0) + TT(1,1) * VV(1,1) + TT(1,2) * VV(1,2)) * dL_dc + (TT(0,0) * VV(1,0) + TT(0,1) * VV(1,1) + TT(0,2) * VV(1,2)) * dL_db;
    float dL_dT12 = 2 * (TT(1,0) * VV(2,0) + TT(1,1) * VV(2,1) + TT(1,2) * VV(2,2)) * dL_dc + (TT(0,0) * VV(2,0) + TT(0,1) * VV(2,1) + TT(0,2) * VV(2,2)) * dL_db;
#define WW(i, j) Wm.m[i][j]
    float dL_dJ00 = WW(0,0) * dL_dT00 + WW(0,1) * dL_dT01 + WW(0,2) * dL_dT02;
    float dL_dJ02 = WW(2,0) * dL_dT00 + WW(2,1) * dL_dT01 + WW(2,2) * dL_dT02;
    float dL_dJ11 = WW(1,0) * dL_dT10 + WW(1,1) * dL_dT11 + WW(1,2) * dL_dT12;
    float dL_dJ12 = WW(2,0) * dL_dT10 + WW(2,1) * dL_dT11 + WW(2,2) * dL_dT12;
#undef TT
#undef VV
#undef WW
    float tz = 1.f / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
    float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
    float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
    float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 + (2 * h_y * t.y) * tz3 * dL_dJ12;
    vec3 dm = transformVec4x3Transpose(v3(dL_dtx, dL_dty, dL_dtz), view);
    dL_dmeans[3 * idx] = dm.x; dL_dmeans[3 * idx + 1] = dm.y; dL_dmeans[3 * idx + 2] = dm.z;   /* assignment (:273) */
}

/* backward.cu:278-341 */
static void cov3d_backward(int idx, const float* scale, float mod, const float* rot, const float* dL_dcov3Ds,
                           float* dL_dscales, float* dL_drots) {
    float r = rot[0], x = rot[1], y = rot[2], z = rot[3];
    mat3 R = { { { 1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y) },
                 { 2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x) },
                 { 2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y) } } };
    mat3 S; memset(&S, 0, sizeof(S));
    vec3 s = v3(mod * scale[0], mod * scale[1], mod * scale[2]);
    S.m[0][0] = s.x; S.m[1][1] = s.y; S.m[2][2] = s.z;
    mat3 M = mat3_mul(S, R);
    const float* d = dL_dcov3Ds + 6 * (size_t)idx;
    mat3 dL_dSigma = { { { d[0], 0.5f * d[1], 0.5f * d[2] }, { 0.5f * d[1], d[3], 0.5f * d[4] }, { 0.5f * d[2], 0.5f * d[4], d[5] } } };
    mat3 M2; for (int j = 0; j < 3; ++j) for (int i = 0; i < 3; ++i) M2.m[j][i] = 2.0f * M.m[j][i];
    mat3 dL_dM = mat3_mul(M2, dL_dSigma);
    mat3 Rt = mat3_transpose(R), dL_dMt = mat3_transpose(dL_dM);
    float* ds = dL_dscales + 3 * (size_t)idx;
    for (int k = 0; k < 3; ++k)
        ds[k] = Rt.m[k][0] * dL_dMt.m[k][0] + Rt.m[k][1] * dL_dMt.m[k][1] + Rt.m[k][2] * dL_dMt.m[k][2];
    for (int i = 0; i < 3; ++i) { dL_dMt.m[0][i] *= s.x; dL_dMt.m[1][i] *= s.y; dL_dMt.m[2][i] *= s.z; }
#define D(i, j) dL_dMt.m[i][j]
    float* q = dL_drots + 4 * (size_t)idx;
    q[0] = 2 * z * (D(0,1) - D(1,0)) + 2 * y * (D(2,0) - D(0,2)) + 2 * x * (D(1,2) - D(2,1));
    q[1] = 2 * y * (D(1,0) + D(0,1)) + 2 * z * (D(2,0) + D(0,2)) + 2 * r * (D(1,2) - D(2,1)) - 4 * x * (D(2,2) + D(1,1));
    q[2] = 2 * x * (D(1,0) + D(0,1)) + 2 * r * (D(2,0) - D(0,2)) + 2 * z * (D(1,2) + D(2,1)) - 4 * y * (D(2,2) + D(0,0));
    q[3] = 2 * r * (D(0,1) - D(1,0)) + 2 * x * (D(2,0) + D(0,2)) + 2 * y * (D(1,2) + D(2,1)) - 4 * z * (D(1,1) + D(0,0));
#undef D
}

/* rasterizer_impl.cu:343-446 + RasterizeGaussiansBackwardCUDA's zero fill (rasterize_points.cu:158-167).
 * All outputs are written in full. dL_dconic is [P*4] (x, y, unused, w), dL_ddepths [P]. */
void gs_oracle_backward(const GsOracle* s, int D, int M, const float* bg,
                        const float* means3D, const float* shs, const float* colors_precomp,
                        const float* alphas, const float* scales, float scale_modifier,
                        const float* rotations, const float* cov3D_precomp,
                        const float* view, const float* proj, const float* campos,
                        float tan_fovx, float tan_fovy,
                        const float* dL_dpix, const float* dL_dpix_depth, const float* dL_dalphas,
                        float* dL_dmean2D /*P*3*/, float* dL_dconic /*P*4*/, float* dL_dopacity, float* dL_dcolor /*P*3*/,
                        float* dL_ddepth, float* dL_dmean3D, float* dL_dcov3D /*P*6*/, float* dL_dsh,
                        float* dL_dscale, float* dL_drot) {
    const int P = s->P, W = s->W, H = s->H;
    const float focal_y = H / (2.0f * tan_fovy), focal_x = W / (2.0f * tan_fovx);
    memset(dL_dmean2D, 0, (size_t)P * 12); memset(dL_dconic, 0, (size_t)P * 16); memset(dL_dopacity, 0, (size_t)P * 4);
    memset(dL_dcolor, 0, (size_t)P * 12); memset(dL_ddepth, 0, (size_t)P * 4); memset(dL_dmean3D, 0, (size_t)P * 12);
    memset(dL_dcov3D, 0, (size_t)P * 24); if (dL_dsh && M) memset(dL_dsh, 0, (size_t)P * M * 12);
    memset(dL_dscale, 0, (size_t)P * 12); memset(dL_drot, 0, (size_t)P * 16);
    double* acc = (double*)calloc((size_t)(P > 0 ? P : 1) * 10, sizeof(double));
    const float* color_ptr = colors_precomp ? colors_precomp : s->rgb;      /* rasterizer_impl.cu:396 */
#pragma omp parallel for schedule(dynamic, 4)
    for (int t = 0; t < s->tiles; ++t)
        render_tile_backward(s, t, bg, color_ptr, alphas, dL_dpix, dL_dpix_depth, dL_dalphas, acc);
    for (int i = 0; i < P; ++i) {
        const double* a = acc + 10 * (size_t)i;
        dL_dmean2D[3 * i] = (float)a[0]; dL_dmean2D[3 * i + 1] = (float)a[1];
        dL_dconic[4 * i] = (float)a[2]; dL_dconic[4 * i + 1] = (float)a[3]; dL_dconic[4 * i + 3] = (float)a[4];
        dL_dopacity[i] = (float)a[5];
        dL_dcolor[3 * i] = (float)a[6]; dL_dcolor[3 * i + 1] = (float)a[7]; dL_dcolor[3 * i + 2] = (float)a[8];
        dL_ddepth[i] = (float)a[9];
    }
    free(acc);
    const float* cov3D_ptr = cov3D_precomp ? cov3D_precomp : s->cov3D;      /* rasterizer_impl.cu:423 */
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; ++idx) {
        if (!(s->radii[idx] > 0)) continue;
        cov2d_backward(idx, means3D, cov3D_ptr, focal_x, focal_y, tan_fovx, tan_fovy, view, dL_dconic, dL_dmean3D, dL_dcov3D);
        /* preprocessCUDA backward, backward.cu:347-412 */
        vec3 m = v3(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]);
        vec4 m_hom = transformPoint4x4(m, proj);
        float m_w = 1.0f / (m_hom.w + 0.0000001f);
        float mul1 = (proj[0] * m.x + proj[4] * m.y + proj[8] * m.z + proj[12]) * m_w * m_w;
        float mul2 = (proj[1] * m.x + proj[5] * m.y + proj[9] * m.z + proj[13]) * m_w * m_w;
        const float gx_ = dL_dmean2D[3 * idx], gy_ = dL_dmean2D[3 * idx + 1];
        dL_dmean3D[3 * idx + 0] += (proj[0] * m_w - proj[3] * mul1) * gx_ + (proj[1] * m_w - proj[3] * mul2) * gy_;
        dL_dmean3D[3 * idx + 1] += (proj[4] * m_w - proj[7] * mul1) * gx_ + (proj[5] * m_w - proj[7] * mul2) * gy_;
        dL_dmean3D[3 * idx + 2] += (proj[8] * m_w - proj[11] * mul1) * gx_ + (proj[9] * m_w - proj[11] * mul2) * gy_;
        float mul3 = view[2] * m.x + view[6] * m.y + view[10] * m.z + view[14];
        dL_dmean3D[3 * idx + 0] += (view[2] - view[3] * mul3) * dL_ddepth[idx];
        dL_dmean3D[3 * idx + 1] += (view[6] - view[7] * mul3) * dL_ddepth[idx];
        dL_dmean3D[3 * idx + 2] += (view[10] - view[11] * mul3) * dL_ddepth[idx];
        if (shs) orc_sh_backward(idx, D, M, means3D, campos, shs, s->clamped, dL_dcolor, dL_dmean3D, dL_dsh);
        if (scales) cov3d_backward(idx, scales + 3 * (size_t)idx, scale_modifier, rotations + 4 * (size_t)idx, dL_dcov3D, dL_dscale, dL_drot);
    }
}

/* checkFrustum / markVisible, rasterizer_impl.cu:54-66,141-153 */
void gs_oracle_mark_visible(int P, const float* means3D, const float* view, const float* proj, uint8_t* present) {
    (void)proj;
    for (int i = 0; i < P; ++i) {
        vec3 pv = transformPoint4x3(v3(means3D[3 * i], means3D[3 * i + 1], means3D[3 * i + 2]), view);
        present[i] = pv.z <= 0.2f ? 0 : 1;
    }
}
