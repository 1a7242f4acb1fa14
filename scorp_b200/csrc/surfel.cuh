// surfel.cuh -- declarations of the 2DGS surfel path (diff_surfel_rasterization replacement).
//
// Reference: submodules/diff-surfel-rasterization/cuda_rasterizer/ (forward.cu, backward.cu,
// rasterizer_impl.cu). Same tiling, binning and traversal as the 3DGS path; the per-Gaussian
// record carries the 3x3 splat-to-pixel matrix instead of a conic.
#pragma once
#include "common.cuh"

namespace gsr {

// Per-surfel render record, 96 bytes (six 16-byte quads), what the blend kernels stage.
//   q0 = (x, y, Tu.x, Tu.y)   q1 = (Tu.z, Tv.x, Tv.y, Tv.z)   q2 = (Tw.x, Tw.y, Tw.z, n.x)
//   q3 = (n.y, n.z, opacity, r)   q4 = (g, b, depth, pad)   q5 = (bx0, bx1, by0, by1)
// (x, y) = AABB centre used by the low-pass filter (points_xy_image); Tu/Tv/Tw = rows of transMat
// (forward.cu:195-198); n = view-space normal, flipped towards the camera (:209-214); depth =
// view-space z (sort key); b* = conservative pixel-space box outside which alpha < 1/255.
struct __align__(16) SurfelRec {
    float x, y;
    float T[9];
    float nx, ny, nz, opacity;
    float r, g, b;
    float depth, pad;
    float bx0, bx1, by0, by1;
};
static_assert(sizeof(SurfelRec) == 96, "SurfelRec must be 96 bytes");
constexpr int SURFEL_DEPTH_OFF = 18 * 4;

struct SurfelGeom {
    SurfelRec* rec;
    uint8_t*   clamped;
    uint32_t*  tiles_touched;
    uint32_t*  tile_count;
    uint32_t*  counters;
};
struct SurfelImage {
    uint2*    ranges;
    uint32_t* n_contrib;        // last contributor (1-based list position), forward.cu:428
    uint32_t* median_contrib;   // contributor of the median depth, 0 if none (:432)
    float*    final_T;          // (:427)
    float*    M1;               // sum m*w  (:433)
    float*    M2;               // sum m^2*w (:434)
    uint32_t* tile_order;       // tiles by decreasing list length (binning.cu)
};

inline size_t carve_surfel_geom(char* base, int P, int tiles, SurfelGeom* g, size_t* off = nullptr) {
    size_t o = 0, k = 0;
    auto take = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes); if (off) off[k] = at; ++k; return at; };
    size_t a0 = take((size_t)P * sizeof(SurfelRec));
    size_t a1 = take((size_t)P);
    size_t a2 = take((size_t)P * 4);
    size_t a3 = take((size_t)tiles * 4 * TC_STRIDE);
    size_t a4 = take(GEOM_COUNTERS * 4);
    if (off) off[5] = o;
    if (g) {
        g->rec = (SurfelRec*)(base + a0); g->clamped = (uint8_t*)(base + a1);
        g->tiles_touched = (uint32_t*)(base + a2); g->tile_count = (uint32_t*)(base + a3);
        g->counters = (uint32_t*)(base + a4);
    }
    return o;
}
inline size_t carve_surfel_image(char* base, int W, int H, SurfelImage* s, size_t* off = nullptr) {
    const int tiles = ((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE);
    const size_t hw = (size_t)W * H;
    size_t o = 0;
    size_t a[6];
    a[0] = o; o = align_up(o + (size_t)tiles * sizeof(uint2));
    for (int k = 1; k < 6; ++k) { a[k] = o; o = align_up(o + hw * 4); }
    const size_t a_order = o; o = align_up(o + (size_t)tiles * 4);
    if (off) { off[0] = a[0]; off[1] = a[1]; off[2] = a[2]; off[3] = a[3]; }   // ranges, n_contrib, median, final_T (M1, M2 follow at equal stride)
    if (s) {
        s->ranges = (uint2*)(base + a[0]); s->n_contrib = (uint32_t*)(base + a[1]);
        s->median_contrib = (uint32_t*)(base + a[2]); s->final_T = (float*)(base + a[3]);
        s->M1 = (float*)(base + a[4]); s->M2 = (float*)(base + a[5]);
        s->tile_order = (uint32_t*)(base + a_order);
    }
    return o;
}

// accumulator slots written by the surfel blend backward (20 floats = 80 B per surfel)
enum { SACC_T = 0 /* 9: dL_dtransMat */, SACC_MX = 9, SACC_MY = 10, SACC_OP = 11, SACC_R = 12, SACC_G = 13,
       SACC_B = 14, SACC_NX = 15, SACC_NY = 16, SACC_NZ = 17, SACC_USED = 18, SACC_STRIDE = 20 };

struct SurfelFwdArgs {
    int P, M, D, W, H, gx, gy;
    float scale_modifier;
    const float *means3D, *shs, *colors_precomp, *opacities, *scales, *rotations, *transMat_precomp;
    const float *view, *proj, *campos;
    bool prefiltered;
    // model-space variant (gsr_surfel_forward_raw): see FwdArgs in common.cuh
    const float* shs_rest = nullptr;
    bool raw = false;
    bool sh_wide = false;              // SH rows 32-byte aligned: read with 32-byte loads (set by launch_surfel_preprocess)
};
cudaError_t launch_surfel_preprocess(const SurfelFwdArgs& a, const SurfelGeom& g, int32_t* radii, cudaStream_t st);
cudaError_t launch_surfel_render_forward(int W, int H, int gx, int gy, const SurfelGeom& g, const SurfelImage& img,
                                         const BinningState& bin, const float* bg, float* out_color, float* out_others,
                                         cudaStream_t st);
cudaError_t launch_surfel_render_backward(int W, int H, int gx, int gy, const SurfelGeom& g, const SurfelImage& img,
                                          const BinningState& bin, const float* bg, const float* dL_dpix,
                                          const float* dL_dothers, float* acc, cudaStream_t st);
struct SurfelBwdArgs {
    int P, M, D, W, H;
    float tanfovx, tanfovy, focal_x, focal_y;
    const float *means3D, *shs, *scales, *rotations;
    bool T_precomp, colors_precomp;
    const float *view, *proj, *campos;
    const int32_t* radii;
    const SurfelRec* rec;
    const uint8_t* clamped;
    const float* acc;
    float *dL_dmeans2D, *dL_dcolors, *dL_dopacity, *dL_dmeans3D, *dL_dtransMat, *dL_dsh, *dL_dscales, *dL_drotations;
    // model-space variant (gsr_surfel_backward_raw)
    const float* shs_rest = nullptr;
    const float* opacities = nullptr;
    float* dL_dsh_rest = nullptr;
    bool raw = false;
    int row_begin = 0, row_end = 0;     // as BwdArgs
    float* peer_head = nullptr;    // as BwdArgs::peer_head: [P][12] = (means3D xyz, opacity | scales xy, rot rx | rot yz, 0, 0)
};
cudaError_t launch_surfel_preprocess_backward(const SurfelBwdArgs& a, cudaStream_t st);

// simple_knn
cudaError_t launch_knn(int P, const float* xyz, float* out, void* tmp, size_t tmp_bytes, cudaStream_t st);
size_t knn_workspace_bytes(int P);

}  // namespace gsr
