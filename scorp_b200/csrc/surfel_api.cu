// surfel_api.cu -- extern "C" entry points of the 2DGS surfel path and of simple_knn.
// Orchestration mirrors CudaRasterizer::Rasterizer::forward/backward of the surfel rasterizer
// (reference diff-surfel-rasterization/cuda_rasterizer/rasterizer_impl.cu:197-327, :331-448).
#include "surfel.cuh"

namespace gsr { void prof_mark(int i, cudaStream_t st); }
using namespace gsr;

static bool surfel_sh_ok(int M, int D) {
    return (M == 1 || M == 4 || M == 9 || M == 16) && D >= 0 && D <= 3 && (D + 1) * (D + 1) <= M;
}

extern "C" {

void gsr_surfel_geom_layout(int P, int tiles, size_t off[6], size_t* total) { *total = carve_surfel_geom(nullptr, P, tiles, nullptr, off); }
void gsr_surfel_image_layout(int W, int H, size_t off[4], size_t* total) { *total = carve_surfel_image(nullptr, W, H, nullptr, off); }
size_t gsr_surfel_backward_workspace_bytes(int P) { return align_up((size_t)(P > 0 ? P : 1) * SACC_STRIDE * sizeof(float)); }

static int surfel_forward_impl(
    const GsrSettings* s, int P, int M,
    const float* means3D, const float* shs, const float* shs_rest, bool raw, const float* colors_precomp,
    const float* opacities, const float* scales, const float* rotations, const float* transMat_precomp,
    float* out_color, float* out_others, int32_t* radii,
    gsr_alloc_fn alloc, void* alloc_user, int* num_rendered, void* stream, const NoSync* nosync = nullptr)
{
    cudaStream_t st = (cudaStream_t)stream;
    StageRanges nvtx_guard;
    if (!s || !alloc || !num_rendered) { set_error("null settings/allocator/num_rendered"); return GSR_EINVAL; }
    *num_rendered = 0;
    const int W = s->image_width, H = s->image_height;
    if (P < 0 || W <= 0 || H <= 0) { set_error("bad sizes P=%d W=%d H=%d", P, W, H); return GSR_EINVAL; }
    if (!out_color || !out_others) { set_error("null output image"); return GSR_EINVAL; }
    const size_t hw = (size_t)W * H;
    int rc;
    if (P == 0) {   // rasterize_points.cu:86: zero outputs, nothing else
        if ((rc = check_cuda(cudaMemsetAsync(out_color, 0, 3 * hw * 4, st), "memset color"))) return rc;
        return check_cuda(cudaMemsetAsync(out_others, 0, 7 * hw * 4, st), "memset others");
    }
    if (!means3D || !opacities || !radii || !s->bg || !s->viewmatrix || !s->projmatrix || !s->campos) {
        set_error("null required input"); return GSR_EINVAL;
    }
    if ((shs == nullptr) == (colors_precomp == nullptr)) { set_error("provide exactly one of shs / colors_precomp"); return GSR_EINVAL; }
    if (((scales == nullptr) || (rotations == nullptr)) == (transMat_precomp == nullptr)) {
        set_error("provide exactly one of scales+rotations / transMat_precomp"); return GSR_EINVAL;
    }
    if (shs && !surfel_sh_ok(M, s->sh_degree)) { set_error("unsupported SH layout: M=%d degree=%d", M, s->sh_degree); return GSR_EUNSUPPORTED; }
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE, tiles = gx * gy;

    SurfelGeom g; SurfelImage img; BinningState bin;
    char* gbase = (char*)alloc(alloc_user, GSR_BUF_GEOM, carve_surfel_geom(nullptr, P, tiles, nullptr));
    char* ibase = (char*)alloc(alloc_user, GSR_BUF_IMAGE, carve_surfel_image(nullptr, W, H, nullptr));
    if (!gbase || !ibase) { set_error("allocator returned NULL"); return GSR_EALLOC; }
    carve_surfel_geom(gbase, P, tiles, &g);
    carve_surfel_image(ibase, W, H, &img);
    const size_t zero_bytes = (size_t)((char*)g.counters - (char*)g.tile_count) + GEOM_COUNTERS * 4;
    if ((rc = check_cuda(cudaMemsetAsync(g.tile_count, 0, zero_bytes, st), "memset tile counters"))) return rc;

    prof_mark(0, st);
    SurfelFwdArgs a;
    a.P = P; a.M = shs ? M : 0; a.D = s->sh_degree; a.W = W; a.H = H; a.gx = gx; a.gy = gy;
    a.scale_modifier = s->scale_modifier;
    a.means3D = means3D; a.shs = shs; a.colors_precomp = colors_precomp; a.opacities = opacities;
    a.scales = scales; a.rotations = rotations; a.transMat_precomp = transMat_precomp;
    a.view = s->viewmatrix; a.proj = s->projmatrix; a.campos = s->campos;
    a.prefiltered = s->prefiltered != 0;
    a.shs_rest = shs_rest; a.raw = raw;
    if ((rc = check_cuda(launch_surfel_preprocess(a, g, radii, st), "surfel preprocess launch"))) return rc;
    if ((rc = stage_check(s, st, "surfel preprocess"))) return rc;
    prof_mark(1, st);

    // the scan over tiles needs nothing from the host, so it runs ahead of the readback and adds the longest
    // tile segment (counters[1]) to it
    if (nosync && (nosync->capacity < 1 || !nosync->status)) { set_error("no-sync forward needs capacity >= 1 and a pinned status slot"); return GSR_EINVAL; }
    if ((rc = check_cuda(launch_tile_scan(gx * gy, g.tile_count, img.ranges, g.counters, img.tile_order, st,
                                          nosync ? (uint32_t)nosync->capacity : 0xffffffffu), "tile scan launch"))) return rc;
    int N, max_len = 0;
    if (nosync) {       // see gsr_rasterize_forward_nosync (api.cu)
        if ((rc = check_cuda(cudaMemcpyAsync(nosync->status, g.counters, 12, cudaMemcpyDeviceToHost, st), "copy status"))) return rc;
        N = nosync->capacity;
    } else {
        uint32_t n_stack[2] = {0, 0};
        uint32_t* n_host = readback_pad() ? readback_pad() : n_stack;
        if ((rc = check_cuda(cudaMemcpyAsync(n_host, g.counters, 8, cudaMemcpyDeviceToHost, st), "copy num_rendered"))) return rc;
        if ((rc = check_cuda(cudaStreamSynchronize(st), "sync num_rendered"))) return rc;
        N = (int)n_host[0]; max_len = (int)n_host[1];
    }
    *num_rendered = N;

    char* bbase = (char*)alloc(alloc_user, GSR_BUF_BINNING, carve_binning(nullptr, N, nullptr, nullptr, gx * gy));
    if (!bbase) { set_error("allocator returned NULL (binning)"); return GSR_EALLOC; }
    carve_binning(bbase, N, &bin, nullptr, gx * gy);
    if ((rc = check_cuda(launch_binning(P, gx, gy, N, max_len, g.rec, (int)sizeof(SurfelRec), SURFEL_DEPTH_OFF, g.tile_count, radii,
                                        img.ranges, bin, st, 1, nosync ? g.counters : nullptr, img.tile_order), "binning launch"))) return rc;
    if ((rc = stage_check(s, st, "binning"))) return rc;
    prof_mark(2, st);
    if ((rc = check_cuda(cudaMemsetAsync(bin.contrib, 0, contrib_bytes(N, gx * gy), st), "memset contribution masks"))) return rc;
    if ((rc = check_cuda(launch_surfel_render_forward(W, H, gx, gy, g, img, bin, s->bg, out_color, out_others, st),
                         "surfel render forward launch"))) return rc;
    prof_mark(3, st);
    return stage_check(s, st, "surfel render forward");
}

int gsr_surfel_forward(
    const GsrSettings* s, int P, int M,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* opacities, const float* scales, const float* rotations, const float* transMat_precomp,
    float* out_color, float* out_others, int32_t* radii,
    gsr_alloc_fn alloc, void* alloc_user, int* num_rendered, void* stream)
{
    return surfel_forward_impl(s, P, M, means3D, shs, nullptr, false, colors_precomp, opacities, scales, rotations,
                               transMat_precomp, out_color, out_others, radii, alloc, alloc_user, num_rendered, stream);
}

int gsr_surfel_forward_nosync(
    const GsrSettings* s, int P, int M,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* opacities, const float* scales, const float* rotations, const float* transMat_precomp,
    float* out_color, float* out_others, int32_t* radii,
    gsr_alloc_fn alloc, void* alloc_user, int capacity, uint32_t* status_pinned, void* stream)
{
    const NoSync ns{capacity, status_pinned};
    int n = 0;
    return surfel_forward_impl(s, P, M, means3D, shs, nullptr, false, colors_precomp, opacities, scales, rotations,
                               transMat_precomp, out_color, out_others, radii, alloc, alloc_user, &n, stream, &ns);
}

static bool model_params_ok(int P, int M, const GsrModelParams* mp) {
    return mp && (P <= 0 || (mp->xyz && mp->features_dc && mp->opacity && mp->scaling && mp->rotation && (M <= 1 || mp->features_rest)));
}

int gsr_surfel_forward_raw(
    const GsrSettings* s, int P, int M, const GsrModelParams* mp,
    float* out_color, float* out_others, int32_t* radii,
    gsr_alloc_fn alloc, void* alloc_user, int* num_rendered, void* stream)
{
    if (!model_params_ok(P, M, mp)) { set_error("null model parameter"); return GSR_EINVAL; }
    if (M < 1) { set_error("model-space entry needs SH coefficients (M >= 1)"); return GSR_EUNSUPPORTED; }
    return surfel_forward_impl(s, P, M, mp->xyz, mp->features_dc, mp->features_rest, true, nullptr, mp->opacity, mp->scaling,
                               mp->rotation, nullptr, out_color, out_others, radii, alloc, alloc_user, num_rendered, stream);
}

struct SurfelRawBwd { const float* shs_rest; const float* opacities; float* dL_dsh_rest; };

static int surfel_backward_impl(
    const GsrSettings* s, int P, int M, int R,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* scales, const float* rotations, const float* transMat_precomp,
    const int32_t* radii,
    const void* geom, const void* binning, const void* image,
    const float* dL_dout_color, const float* dL_dout_others,
    float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D,
    float* dL_dtransMat, float* dL_dsh, float* dL_dscales, float* dL_drotations,
    void* workspace, size_t workspace_bytes, void* stream, const SurfelRawBwd* raw);

int gsr_surfel_backward(
    const GsrSettings* s, int P, int M, int R,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* scales, const float* rotations, const float* transMat_precomp,
    const int32_t* radii,
    const void* geom, const void* binning, const void* image,
    const float* dL_dout_color, const float* dL_dout_others,
    float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D,
    float* dL_dtransMat, float* dL_dsh, float* dL_dscales, float* dL_drotations,
    void* workspace, size_t workspace_bytes, void* stream)
{
    return surfel_backward_impl(s, P, M, R, means3D, shs, colors_precomp, scales, rotations, transMat_precomp, radii, geom,
                                binning, image, dL_dout_color, dL_dout_others, dL_dmeans2D, dL_dcolors, dL_dopacity,
                                dL_dmeans3D, dL_dtransMat, dL_dsh, dL_dscales, dL_drotations, workspace, workspace_bytes,
                                stream, nullptr);
}

int gsr_surfel_backward_raw(
    const GsrSettings* s, int P, int M, int R, const GsrModelParams* mp,
    const int32_t* radii, const void* geom, const void* binning, const void* image,
    const float* dL_dout_color, const float* dL_dout_others,
    float* dL_dmeans2D, const GsrModelGrads* mg, void* workspace, size_t workspace_bytes, void* stream)
{
    if (!model_params_ok(P, M, mp) || !mg) { set_error("null model parameters/gradients"); return GSR_EINVAL; }
    if (P > 0 && (!mg->xyz || !mg->features_dc || !mg->opacity || !mg->scaling || !mg->rotation || (M > 1 && !mg->features_rest))) {
        set_error("null model gradient"); return GSR_EINVAL;
    }
    if (M < 1) { set_error("model-space entry needs SH coefficients (M >= 1)"); return GSR_EUNSUPPORTED; }
    const SurfelRawBwd raw = { mp->features_rest, mp->opacity, mg->features_rest };
    // dL_dcolors / dL_dtransMat have no model parameter behind them: not produced (any non-NULL pointer passes the checks)
    return surfel_backward_impl(s, P, M, R, mp->xyz, mp->features_dc, nullptr, mp->scaling, mp->rotation, nullptr, radii, geom,
                                binning, image, dL_dout_color, dL_dout_others, dL_dmeans2D, mg->xyz, mg->opacity, mg->xyz,
                                mg->xyz, mg->features_dc, mg->scaling, mg->rotation, workspace, workspace_bytes, stream, &raw);
}

static int surfel_backward_impl(
    const GsrSettings* s, int P, int M, int R,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* scales, const float* rotations, const float* transMat_precomp,
    const int32_t* radii,
    const void* geom, const void* binning, const void* image,
    const float* dL_dout_color, const float* dL_dout_others,
    float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D,
    float* dL_dtransMat, float* dL_dsh, float* dL_dscales, float* dL_drotations,
    void* workspace, size_t workspace_bytes, void* stream, const SurfelRawBwd* raw)
{
    cudaStream_t st = (cudaStream_t)stream;
    StageRanges nvtx_guard;
    const EarlyRgb e = take_early_rgb();         // consumed by THIS call, also when it returns early (see api.cu)
    const PeerHead peer_head = take_peer_head();
    if (!s) { set_error("null settings"); return GSR_EINVAL; }
    if (P == 0) return GSR_OK;
    const int W = s->image_width, H = s->image_height;
    if (!geom || !binning || !image || !radii || !means3D || !dL_dout_color || !dL_dout_others ||
        !dL_dmeans2D || !dL_dcolors || !dL_dopacity || !dL_dmeans3D || !dL_dtransMat || !dL_dscales || !dL_drotations || !workspace) {
        set_error("null required pointer in surfel backward"); return GSR_EINVAL;
    }
    if (!transMat_precomp && (!scales || !rotations)) { set_error("scales/rotations missing"); return GSR_EINVAL; }
    if (workspace_bytes < gsr_surfel_backward_workspace_bytes(P)) { set_error("workspace too small"); return GSR_EINVAL; }
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE, tiles = gx * gy;
    SurfelGeom g; SurfelImage img; BinningState bin;
    carve_surfel_geom((char*)geom, P, tiles, &g);
    carve_surfel_image((char*)image, W, H, &img);
    carve_binning((char*)binning, R, &bin, nullptr, gx * gy);

    float* acc = (float*)workspace;
    int rc;
    prof_mark(4, st);
    if ((rc = check_cuda(cudaMemsetAsync(acc, 0, (size_t)P * SACC_STRIDE * sizeof(float), st), "memset accumulators"))) return rc;
    if (R > 0) {
        if ((rc = check_cuda(launch_surfel_render_backward(W, H, gx, gy, g, img, bin, s->bg, dL_dout_color, dL_dout_others,
                                                           acc, st), "surfel render backward launch"))) return rc;
        if ((rc = stage_check(s, st, "surfel render backward"))) return rc;
    }
    prof_mark(5, st);
    {   // multi-GPU: see gsr_rasterize_backward
        if (e.armed && shs) {
            if (e.peers.n > 0 || e.peers.mc_drgb) {
                if ((rc = check_cuda(launch_clamped_rgb_push(P, radii, g.clamped, acc, SACC_STRIDE, SACC_R, e.peers, st),
                                     "peer colour gradient push"))) return rc;
            } else if ((rc = check_cuda(launch_clamped_rgb_from_acc(P, radii, g.clamped, acc, SACC_STRIDE, SACC_R, e.drgb, st),
                                        "early colour gradient launch"))) return rc;
            if (e.ready) e.ready(e.user);
        }
    }
    SurfelBwdArgs b;
    b.P = P; b.M = shs ? M : 0; b.D = s->sh_degree; b.W = W; b.H = H;
    b.tanfovx = s->tanfovx; b.tanfovy = s->tanfovy;
    b.focal_y = H / (2.0f * s->tanfovy); b.focal_x = W / (2.0f * s->tanfovx);     // rasterizer_impl.cu:374-375
    b.means3D = means3D; b.shs = shs; b.scales = scales; b.rotations = rotations;
    b.T_precomp = transMat_precomp != nullptr; b.colors_precomp = colors_precomp != nullptr;
    b.view = s->viewmatrix; b.proj = s->projmatrix; b.campos = s->campos;
    b.radii = radii; b.rec = g.rec; b.clamped = g.clamped; b.acc = acc;
    b.dL_dmeans2D = dL_dmeans2D; b.dL_dcolors = dL_dcolors; b.dL_dopacity = dL_dopacity; b.dL_dmeans3D = dL_dmeans3D;
    b.dL_dtransMat = dL_dtransMat; b.dL_dsh = shs ? dL_dsh : nullptr; b.dL_dscales = dL_dscales; b.dL_drotations = dL_drotations;
    if (raw) { b.raw = true; b.shs_rest = raw->shs_rest; b.opacities = raw->opacities; b.dL_dsh_rest = raw->dL_dsh_rest; }
    b.peer_head = (shs && !raw) ? peer_head.head : nullptr;
    PeerHead ph = peer_head; ph.head = b.peer_head;             // multi-GPU: row-range launches, see gsr_rasterize_backward
    if ((rc = check_cuda(launch_in_row_chunks(P, 128, ph, [&](int r0, int r1) {
            b.row_begin = r0; b.row_end = r1; return launch_surfel_preprocess_backward(b, st); }), "surfel preprocess backward launch"))) return rc;
    prof_mark(6, st);
    return stage_check(s, st, "surfel preprocess backward");
}

size_t gsr_knn_workspace_bytes(int P) { return knn_workspace_bytes(P); }
int gsr_knn_dist2(int P, const float* xyz, float* out, void* tmp, size_t tmp_bytes, void* stream) {
    if (P == 0) return GSR_OK;
    if (P < 0 || !xyz || !out || !tmp) { set_error("bad argument to gsr_knn_dist2"); return GSR_EINVAL; }
    if (tmp_bytes < knn_workspace_bytes(P)) { set_error("knn workspace too small"); return GSR_EINVAL; }
    return check_cuda(launch_knn(P, xyz, out, tmp, tmp_bytes, (cudaStream_t)stream), "knn launch");
}

}  // extern "C"
