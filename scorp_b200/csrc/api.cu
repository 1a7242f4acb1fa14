// api.cu -- extern "C" entry points of libgsr_b200.so (3DGS path, markVisible, layouts).
// Orchestration mirrors CudaRasterizer::Rasterizer::forward/backward
// (reference rasterizer_impl.cu:197-339, :343-446) with the stages of this library.
#include "common.cuh"
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <nvtx3/nvToolsExt.h>      // header-only; the ranges cost nothing unless a tool (nsys, ncu --nvtx) is attached

namespace gsr {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
}
int check_cuda(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return GSR_OK;
    set_error("%s: %s", what, cudaGetErrorString(e));
    return GSR_ECUDA;
}
// debug mode: synchronise and surface asynchronous errors after each stage (auxiliary.h:166-173)
int stage_check(const GsrSettings* s, cudaStream_t st, const char* what) {
    if (!s->debug) return GSR_OK;
    return check_cuda(cudaStreamSynchronize(st), what);
}

// ---- optional per-stage timing (bench.py's live roofline numbers) --------------------------
// When enabled, an event is recorded on the caller's stream after every stage; gsr_profile_read
// synchronises on the last event of each pass and returns the elapsed times.
constexpr int PROF_RING = 64, PROF_EV = 7;
struct StageProf {
    bool on = false, made = false;
    int device = -1;                          // device the events were created on
    int n_fwd = 0, n_bwd = 0;                 // passes recorded since enable (ring of PROF_RING)
    cudaEvent_t ev[PROF_RING][PROF_EV] = {};
};
static StageProf g_prof;
// events: 0 start fwd, 1 after preprocess, 2 after binning, 3 after blend fwd,
//         4 start bwd, 5 after blend bwd, 6 after preprocess bwd
// NVTX range per stage (SURVEY.md 5 "tracing"), only while gsr_profile_enable(1) is in force: prof_mark(i) closes the
// range of the stage that just ended and opens the next one; StageRanges closes whatever is open when the entry point
// returns (also on an error path).
static thread_local bool g_nvtx_open = false;
static void nvtx_stage(const char* name) {
    if (g_nvtx_open) { nvtxRangePop(); g_nvtx_open = false; }
    if (name) { nvtxRangePushA(name); g_nvtx_open = true; }
}
StageRanges::~StageRanges() { nvtx_stage(nullptr); }
void prof_mark(int i, cudaStream_t st) {
    if (!g_prof.on) { return; }
    static const char* const kNext[PROF_EV] = { "gsr:preprocess", "gsr:binning", "gsr:blend_forward", nullptr,
                                                "gsr:blend_backward", "gsr:preprocess_backward", nullptr };
    nvtx_stage(kNext[i]);
    // the events belong to the device that is current when profiling is first used; a later call on another device
    // re-creates them there (profiling is a single-device diagnostic: bench.py enables it on the rank's own GPU)
    int dev = 0;
    cudaGetDevice(&dev);
    if (g_prof.made && dev != g_prof.device) {
        for (auto& r : g_prof.ev) for (auto& e : r) cudaEventDestroy(e);
        g_prof.made = false; g_prof.n_fwd = g_prof.n_bwd = 0;
    }
    if (!g_prof.made) { for (auto& r : g_prof.ev) for (auto& e : r) cudaEventCreate(&e); g_prof.made = true; g_prof.device = dev; }
    const int slot = (i < 4 ? g_prof.n_fwd : g_prof.n_bwd) % PROF_RING;
    cudaEventRecord(g_prof.ev[slot][i], st);
    if (i == 3) ++g_prof.n_fwd;
    if (i == 6) ++g_prof.n_bwd;
}

static bool sh_count_ok(int M) { return M == 0 || M == 1 || M == 4 || M == 9 || M == 16; }

// pinned landing pad for the one device->host readback of a forward (a pageable destination makes
// cudaMemcpyAsync stage through a driver bounce buffer); one per host thread, never freed
uint32_t* readback_pad() {
    static thread_local uint32_t* pad = nullptr;
    if (!pad && cudaHostAlloc((void**)&pad, 64, cudaHostAllocDefault) != cudaSuccess) { pad = nullptr; cudaGetLastError(); }
    return pad;
}

thread_local EarlyRgb g_early = { nullptr, nullptr, nullptr, false, {} };
thread_local PeerHead g_peer_head;
PeerHead take_peer_head() { const PeerHead h = g_peer_head; g_peer_head = PeerHead{}; return h; }
EarlyRgb take_early_rgb() { const EarlyRgb e = g_early; g_early.armed = false; return e; }

}  // namespace gsr

using namespace gsr;

extern "C" {

const char* gsr_last_error(void) { return g_err; }
int gsr_version(void) { return 100; }

void gsr_profile_enable(int on) { g_prof.on = on != 0; g_prof.n_fwd = g_prof.n_bwd = 0; }
int gsr_profile_read(float* ms) {
    if (!ms) { set_error("null ms"); return GSR_EINVAL; }
    for (int i = 0; i < GSR_PROF_STAGES; ++i) ms[i] = -1.0f;
    if (!g_prof.made) return GSR_OK;
    const int nf = g_prof.n_fwd < PROF_RING ? g_prof.n_fwd : PROF_RING;
    const int nb = g_prof.n_bwd < PROF_RING ? g_prof.n_bwd : PROF_RING;
    float acc[GSR_PROF_STAGES] = {0, 0, 0, 0, 0};
    for (int k = 0; k < nf; ++k) {
        if (cudaEventSynchronize(g_prof.ev[k][3]) != cudaSuccess) { set_error("event sync"); return GSR_ECUDA; }
        for (int i = 0; i < 3; ++i) { float t = 0; cudaEventElapsedTime(&t, g_prof.ev[k][i], g_prof.ev[k][i + 1]); acc[i] += t; }
    }
    for (int k = 0; k < nb; ++k) {
        if (cudaEventSynchronize(g_prof.ev[k][6]) != cudaSuccess) { set_error("event sync"); return GSR_ECUDA; }
        float t = 0;
        cudaEventElapsedTime(&t, g_prof.ev[k][4], g_prof.ev[k][5]); acc[3] += t;
        cudaEventElapsedTime(&t, g_prof.ev[k][5], g_prof.ev[k][6]); acc[4] += t;
    }
    for (int i = 0; i < 3; ++i) if (nf) ms[i] = acc[i] / nf;
    for (int i = 3; i < 5; ++i) if (nb) ms[i] = acc[i] / nb;
    return GSR_OK;
}

void gsr_geom_layout(int P, int tiles, size_t off[6], size_t* total) { *total = carve_geom(nullptr, P, tiles, nullptr, off); }
void gsr_image_layout(int W, int H, size_t off[2], size_t* total) { *total = carve_image(nullptr, W, H, nullptr, off); }
void gsr_binning_layout(int N, size_t off[3], size_t* total) { *total = carve_binning(nullptr, N, nullptr, off); }

size_t gsr_backward_workspace_bytes(int P) { return align_up((size_t)(P > 0 ? P : 1) * ACC_STRIDE * sizeof(float)); }

// shared by gsr_rasterize_forward and gsr_rasterize_forward_raw (raw: shs = features_dc, shs_rest = features_rest,
// opacities / scales / rotations are the stored parameters)
static int forward_impl(
    const GsrSettings* s, int P, int M,
    const float* means3D, const float* shs, const float* shs_rest, bool raw, const float* colors_precomp,
    const float* opacities, const float* scales, const float* rotations, const float* cov3D_precomp,
    float* out_color, float* out_depth, float* out_alpha, int32_t* radii,
    gsr_alloc_fn alloc, void* alloc_user, int* num_rendered, void* stream, int views = 1, const NoSync* nosync = nullptr)
{
    cudaStream_t st = (cudaStream_t)stream;
    StageRanges nvtx_guard;
    if (!s || !alloc || !num_rendered) { set_error("null settings/allocator/num_rendered"); return GSR_EINVAL; }
    if (nosync && (nosync->capacity < 1 || !nosync->status)) { set_error("no-sync forward needs capacity >= 1 and a pinned status slot"); return GSR_EINVAL; }
    if (views < 1 || (long long)views * P > 0x7fffffffLL) { set_error("bad view count %d for P=%d", views, P); return GSR_EINVAL; }
    *num_rendered = 0;
    const int W = s->image_width, H = s->image_height;
    if (P < 0 || W <= 0 || H <= 0) { set_error("bad sizes P=%d W=%d H=%d", P, W, H); return GSR_EINVAL; }
    if (!out_color || !out_depth || !out_alpha) { set_error("null output image"); return GSR_EINVAL; }
    const size_t hw = (size_t)W * H * views;
    int rc;
    if (P == 0) {   // rasterize_points.cu:83: zero outputs, nothing else
        if ((rc = check_cuda(cudaMemsetAsync(out_color, 0, 3 * hw * 4, st), "memset color"))) return rc;
        if ((rc = check_cuda(cudaMemsetAsync(out_depth, 0, hw * 4, st), "memset depth"))) return rc;
        return check_cuda(cudaMemsetAsync(out_alpha, 0, hw * 4, st), "memset alpha");
    }
    if (!means3D || !opacities || !radii || !s->bg || !s->viewmatrix || !s->projmatrix || !s->campos) {
        set_error("null required input"); return GSR_EINVAL;
    }
    if ((shs == nullptr) == (colors_precomp == nullptr)) { set_error("provide exactly one of shs / colors_precomp"); return GSR_EINVAL; }
    if (((scales == nullptr) || (rotations == nullptr)) == (cov3D_precomp == nullptr)) {
        set_error("provide exactly one of scales+rotations / cov3D_precomp"); return GSR_EINVAL;
    }
    if (shs && (!sh_count_ok(M) || M == 0 || (s->sh_degree + 1) * (s->sh_degree + 1) > M || s->sh_degree < 0 || s->sh_degree > 3)) {
        set_error("unsupported SH layout: M=%d degree=%d", M, s->sh_degree); return GSR_EUNSUPPORTED;
    }
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE, tiles = gx * gy * views;
    const int PV = P * views;          // records: view v owns [v*P, (v+1)*P)

    GeomState g; ImageState img; BinningState bin;
    char* gbase = (char*)alloc(alloc_user, GSR_BUF_GEOM, carve_geom(nullptr, PV, tiles, nullptr));
    char* ibase = (char*)alloc(alloc_user, GSR_BUF_IMAGE, carve_image(nullptr, W, H, nullptr, nullptr, views));
    if (!gbase || !ibase) { set_error("allocator returned NULL"); return GSR_EALLOC; }
    carve_geom(gbase, PV, tiles, &g);
    carve_image(ibase, W, H, &img, nullptr, views);

    // tile_count and counters are adjacent at the end of the geometry buffer: one memset
    const size_t zero_bytes = (size_t)((char*)g.counters - (char*)g.tile_count) + GEOM_COUNTERS * 4;
    if ((rc = check_cuda(cudaMemsetAsync(g.tile_count, 0, zero_bytes, st), "memset tile counters"))) return rc;

    prof_mark(0, st);
    FwdArgs a;
    a.P = P; a.M = shs ? M : 0; a.D = s->sh_degree; a.W = W; a.H = H; a.gx = gx; a.gy = gy;
    a.tanfovx = s->tanfovx; a.tanfovy = s->tanfovy;
    a.focal_y = H / (2.0f * s->tanfovy); a.focal_x = W / (2.0f * s->tanfovx);     // rasterizer_impl.cu:224-225
    a.scale_modifier = s->scale_modifier;
    a.means3D = means3D; a.shs = shs; a.colors_precomp = colors_precomp; a.opacities = opacities;
    a.scales = scales; a.rotations = rotations; a.cov3D_precomp = cov3D_precomp;
    a.view = s->viewmatrix; a.proj = s->projmatrix; a.campos = s->campos; a.bg = s->bg;
    a.prefiltered = s->prefiltered != 0;
    a.shs_rest = shs_rest; a.raw = raw;
    a.skip_histogram = views_aggregated(views, gx * gy, PV / views);
    if ((rc = check_cuda(launch_preprocess(a, g, radii, st, views), "preprocess launch"))) return rc;
    if (a.skip_histogram && (rc = check_cuda(launch_views_histogram(PV, views, gx, gy, g.rec, (int)sizeof(GaussRec), radii, g.tile_count, st),
                                             "views histogram launch"))) return rc;
    if ((rc = stage_check(s, st, "preprocess"))) return rc;
    prof_mark(1, st);

    // number of instances -> host (the one blocking point, as rasterizer_impl.cu:282)
    // the scan over tiles needs nothing from the host, so it runs ahead of the readback and adds the longest
    // tile segment (counters[1]) to it
    if ((rc = check_cuda(launch_tile_scan(tiles, g.tile_count, img.ranges, g.counters, img.tile_order, st,
                                          nosync ? (uint32_t)nosync->capacity : 0xffffffffu), "tile scan launch"))) return rc;
    int N, max_len = 0;
    if (nosync) {
        // no host round trip (rasterizer_impl.cu:282 has one): the caller sized the binning buffer; N, the longest segment
        // and the overflow flag travel to its pinned slot behind the kernels and are read after ITS next synchronisation
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

    char* bbase = (char*)alloc(alloc_user, GSR_BUF_BINNING, carve_binning(nullptr, N, nullptr, nullptr, tiles));
    if (!bbase) { set_error("allocator returned NULL (binning)"); return GSR_EALLOC; }
    carve_binning(bbase, N, &bin, nullptr, tiles);

    if ((rc = check_cuda(launch_binning(PV, gx, gy, N, max_len, g.rec, (int)sizeof(GaussRec), 24, g.tile_count, radii, img.ranges, bin, st, views,
                                        nosync ? g.counters : nullptr, img.tile_order), "binning launch"))) return rc;
    if ((rc = stage_check(s, st, "binning"))) return rc;
    prof_mark(2, st);
    if ((rc = check_cuda(cudaMemsetAsync(bin.contrib, 0, contrib_slots(N, tiles) * sizeof(uint4), st), "memset contribution masks"))) return rc;
    if ((rc = check_cuda(launch_render_forward(W, H, gx, gy, g, img, bin, s->bg, out_color, out_depth, out_alpha, st, views),
                         "render forward launch"))) return rc;
    prof_mark(3, st);
    return stage_check(s, st, "render forward");
}

int gsr_rasterize_forward(
    const GsrSettings* s, int P, int M,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* opacities, const float* scales, const float* rotations, const float* cov3D_precomp,
    float* out_color, float* out_depth, float* out_alpha, int32_t* radii,
    gsr_alloc_fn alloc, void* alloc_user, int* num_rendered, void* stream)
{
    return forward_impl(s, P, M, means3D, shs, nullptr, false, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                        out_color, out_depth, out_alpha, radii, alloc, alloc_user, num_rendered, stream);
}

int gsr_rasterize_forward_nosync(
    const GsrSettings* s, int P, int M,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* opacities, const float* scales, const float* rotations, const float* cov3D_precomp,
    float* out_color, float* out_depth, float* out_alpha, int32_t* radii,
    gsr_alloc_fn alloc, void* alloc_user, int capacity, uint32_t* status_pinned, void* stream)
{
    const NoSync ns{capacity, status_pinned};
    int n = 0;
    return forward_impl(s, P, M, means3D, shs, nullptr, false, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                        out_color, out_depth, out_alpha, radii, alloc, alloc_user, &n, stream, 1, &ns);
}

int gsr_rasterize_forward_raw(
    const GsrSettings* s, int P, int M, const GsrModelParams* mp,
    float* out_color, float* out_depth, float* out_alpha, int32_t* radii,
    gsr_alloc_fn alloc, void* alloc_user, int* num_rendered, void* stream)
{
    if (!mp) { set_error("null model parameters"); return GSR_EINVAL; }
    if (P > 0 && (!mp->features_dc || !mp->opacity || !mp->scaling || !mp->rotation || (M > 1 && !mp->features_rest))) {
        set_error("null model parameter"); return GSR_EINVAL;
    }
    if (M < 1) { set_error("model-space entry needs SH coefficients (M >= 1)"); return GSR_EUNSUPPORTED; }
    return forward_impl(s, P, M, mp->xyz, mp->features_dc, mp->features_rest, true, nullptr, mp->opacity, mp->scaling,
                        mp->rotation, nullptr, out_color, out_depth, out_alpha, radii, alloc, alloc_user, num_rendered, stream);
}

int gsr_rasterize_forward_views(
    const GsrSettings* s, int V, int P, int M,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* opacities, const float* scales, const float* rotations, const float* cov3D_precomp,
    float* out_color, float* out_depth, float* out_alpha, int32_t* radii,
    gsr_alloc_fn alloc, void* alloc_user, int* num_rendered, void* stream)
{
    return forward_impl(s, P, M, means3D, shs, nullptr, false, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                        out_color, out_depth, out_alpha, radii, alloc, alloc_user, num_rendered, stream, V);
}

struct RawBwd { const float* shs_rest; const float* opacities; float* dL_dsh_rest; };

static int backward_impl(
    const GsrSettings* s, int P, int M, int R,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* scales, const float* rotations, const float* cov3D_precomp,
    const int32_t* radii,
    const void* geom, const void* binning, const void* image,
    const float* out_alpha,
    const float* dL_dout_color, const float* dL_dout_depth, const float* dL_dout_alpha,
    float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D,
    float* dL_dcov3D, float* dL_dsh, float* dL_dscales, float* dL_drotations,
    void* workspace, size_t workspace_bytes, void* stream, const RawBwd* raw);

int gsr_rasterize_backward(
    const GsrSettings* s, int P, int M, int R,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* scales, const float* rotations, const float* cov3D_precomp,
    const int32_t* radii,
    const void* geom, const void* binning, const void* image,
    const float* out_alpha,
    const float* dL_dout_color, const float* dL_dout_depth, const float* dL_dout_alpha,
    float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D,
    float* dL_dcov3D, float* dL_dsh, float* dL_dscales, float* dL_drotations,
    void* workspace, size_t workspace_bytes, void* stream)
{
    return backward_impl(s, P, M, R, means3D, shs, colors_precomp, scales, rotations, cov3D_precomp, radii, geom, binning,
                         image, out_alpha, dL_dout_color, dL_dout_depth, dL_dout_alpha, dL_dmeans2D, dL_dcolors, dL_dopacity,
                         dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations, workspace, workspace_bytes, stream, nullptr);
}

int gsr_rasterize_backward_raw(
    const GsrSettings* s, int P, int M, int R, const GsrModelParams* mp,
    const int32_t* radii, const void* geom, const void* binning, const void* image, const float* out_alpha,
    const float* dL_dout_color, const float* dL_dout_depth, const float* dL_dout_alpha,
    float* dL_dmeans2D, const GsrModelGrads* mg, void* workspace, size_t workspace_bytes, void* stream)
{
    if (!mp || !mg) { set_error("null model parameters/gradients"); return GSR_EINVAL; }
    if (P > 0 && (!mp->features_dc || !mp->opacity || !mp->scaling || !mp->rotation || (M > 1 && !mp->features_rest) ||
                  !mg->features_dc || !mg->opacity || !mg->scaling || !mg->rotation || !mg->xyz || (M > 1 && !mg->features_rest))) {
        set_error("null model parameter or gradient"); return GSR_EINVAL;
    }
    if (M < 1) { set_error("model-space entry needs SH coefficients (M >= 1)"); return GSR_EUNSUPPORTED; }
    const RawBwd raw = { mp->features_rest, mp->opacity, mg->features_rest };
    // dL_dcolors / dL_dcov3D have no model parameter behind them: not produced (any non-NULL pointer passes the checks)
    return backward_impl(s, P, M, R, mp->xyz, mp->features_dc, nullptr, mp->scaling, mp->rotation, nullptr, radii, geom, binning,
                         image, out_alpha, dL_dout_color, dL_dout_depth, dL_dout_alpha, dL_dmeans2D, mg->xyz, mg->opacity,
                         mg->xyz, mg->xyz, mg->features_dc, mg->scaling, mg->rotation, workspace, workspace_bytes, stream, &raw);
}

static int backward_impl(
    const GsrSettings* s, int P, int M, int R,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* scales, const float* rotations, const float* cov3D_precomp,
    const int32_t* radii,
    const void* geom, const void* binning, const void* image,
    const float* out_alpha,
    const float* dL_dout_color, const float* dL_dout_depth, const float* dL_dout_alpha,
    float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D,
    float* dL_dcov3D, float* dL_dsh, float* dL_dscales, float* dL_drotations,
    void* workspace, size_t workspace_bytes, void* stream, const RawBwd* raw)
{
    cudaStream_t st = (cudaStream_t)stream;
    // whatever was armed for "the next backward of this thread" is consumed by THIS call, also when it returns early:
    // the armed state holds raw pointers into a caller-owned gradient arena and into peer buffers
    StageRanges nvtx_guard;
    const EarlyRgb e = take_early_rgb();
    const PeerHead peer_head = take_peer_head();
    if (!s) { set_error("null settings"); return GSR_EINVAL; }
    if (P == 0) return GSR_OK;
    const int W = s->image_width, H = s->image_height;
    if (!geom || !binning || !image || !radii || !means3D || !out_alpha || !dL_dout_color || !dL_dout_depth || !dL_dout_alpha ||
        !dL_dmeans2D || !dL_dcolors || !dL_dopacity || !dL_dmeans3D || !dL_dcov3D || !dL_dscales || !dL_drotations || !workspace) {
        set_error("null required pointer in backward"); return GSR_EINVAL;
    }
    if (workspace_bytes < gsr_backward_workspace_bytes(P)) { set_error("workspace too small"); return GSR_EINVAL; }
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE, tiles = gx * gy;
    GeomState g; ImageState img; BinningState bin;
    carve_geom((char*)geom, P, tiles, &g);
    carve_image((char*)image, W, H, &img);
    carve_binning((char*)binning, R, &bin, nullptr, gx * gy);

    float* acc = (float*)workspace;
    int rc;
    prof_mark(4, st);
    if ((rc = check_cuda(cudaMemsetAsync(acc, 0, (size_t)P * ACC_STRIDE * sizeof(float), st), "memset accumulators"))) return rc;
    if (R > 0) {
        if ((rc = check_cuda(launch_render_backward(W, H, gx, gy, g, img, bin, s->bg, out_alpha, dL_dout_color,
                                                    dL_dout_depth, dL_dout_alpha, acc, st), "render backward launch"))) return rc;
        if ((rc = stage_check(s, st, "render backward"))) return rc;
    }
    prof_mark(5, st);
    {   // multi-GPU: the colour gradient is final here; hand it out so its all-gather overlaps the next kernel
        if (e.armed && shs) {
            if (e.peers.n > 0 || e.peers.mc_drgb) {
                if ((rc = check_cuda(launch_clamped_rgb_push(P, radii, g.clamped, acc, ACC_STRIDE, ACC_R, e.peers, st),
                                     "peer colour gradient push"))) return rc;
            } else if ((rc = check_cuda(launch_clamped_rgb_from_acc(P, radii, g.clamped, acc, ACC_STRIDE, ACC_R, e.drgb, st),
                                        "early colour gradient launch"))) return rc;
            if (e.ready) e.ready(e.user);
        }
    }
    BwdArgs b;
    b.P = P; b.M = shs ? M : 0; b.D = s->sh_degree; b.W = W; b.H = H;
    b.tanfovx = s->tanfovx; b.tanfovy = s->tanfovy;
    b.focal_y = H / (2.0f * s->tanfovy); b.focal_x = W / (2.0f * s->tanfovx);
    b.scale_modifier = s->scale_modifier;
    b.means3D = means3D; b.shs = shs; b.colors_precomp = colors_precomp; b.scales = scales; b.rotations = rotations;
    b.cov_precomp = cov3D_precomp != nullptr;
    b.cov3D = cov3D_precomp ? cov3D_precomp : g.cov3D;
    b.view = s->viewmatrix; b.proj = s->projmatrix; b.campos = s->campos;
    b.radii = radii; b.clamped = g.clamped; b.acc = acc;
    b.dL_dmeans2D = dL_dmeans2D; b.dL_dcolors = dL_dcolors; b.dL_dopacity = dL_dopacity; b.dL_dmeans3D = dL_dmeans3D;
    b.dL_dcov3D = dL_dcov3D; b.dL_dsh = shs ? dL_dsh : nullptr; b.dL_dscales = dL_dscales; b.dL_drotations = dL_drotations;
    if (raw) { b.raw = true; b.shs_rest = raw->shs_rest; b.opacities = raw->opacities; b.dL_dsh_rest = raw->dL_dsh_rest; }
    b.peer_head = (shs && !raw) ? peer_head.head : nullptr;
    // multi-GPU: with an armed head the kernel runs as several launches over consecutive row ranges and the caller is told
    // after each one, so the cross-rank reduction of a range overlaps the computation of the next (gsr_backward_peer_head)
    PeerHead ph = peer_head; ph.head = b.peer_head;
    if ((rc = check_cuda(launch_in_row_chunks(P, 128, ph, [&](int r0, int r1) {
            b.row_begin = r0; b.row_end = r1; return launch_preprocess_backward(b, st); }), "preprocess backward launch"))) return rc;
    prof_mark(6, st);
    return stage_check(s, st, "preprocess backward");
}

int gsr_backward_early_rgb(float* drgb, void (*ready)(void*), void* user) {
    if (!drgb) { gsr::g_early = gsr::EarlyRgb{nullptr, nullptr, nullptr, false, {}}; return GSR_OK; }
    gsr::g_early = gsr::EarlyRgb{drgb, ready, user, true, {}};
    return GSR_OK;
}

int gsr_backward_peer_head(float* symmetric_head, int chunks, void (*chunk_done)(void*, int, int, int), void* user) {
    if ((uintptr_t)symmetric_head & 15) { set_error("gsr_backward_peer_head: the head must be 16-byte aligned"); return GSR_EINVAL; }
    if (symmetric_head && (chunks < 1 || chunks > 64)) { set_error("gsr_backward_peer_head: 1..64 chunks"); return GSR_EINVAL; }
    gsr::PeerHead h; h.head = symmetric_head; h.chunks = chunks; h.chunk_done = chunk_done; h.user = user;
    gsr::g_peer_head = symmetric_head ? h : gsr::PeerHead{};
    return GSR_OK;
}

int gsr_peer_consume(const float* symmetric_src, float* dst, int P, int row0, int rows, int surfel, void* stream) {
    if (P == 0 || rows == 0) return GSR_OK;
    if (P < 0 || row0 < 0 || rows < 0 || row0 + rows > P || !symmetric_src || !dst) { set_error("gsr_peer_consume: bad argument"); return GSR_EINVAL; }
    return check_cuda(gsr::launch_peer_consume(symmetric_src, dst, P, row0, rows, surfel != 0, (cudaStream_t)stream), "peer consume launch");
}

int gsr_backward_early_rgb_peers(int n_dst, float* const* drgb_dst, float* const* campos_dst, const float* campos_src,
                                 void* side_stream, void (*ready)(void*), void* user) {
    if (n_dst <= 0) { gsr::g_early = gsr::EarlyRgb{nullptr, nullptr, nullptr, false, {}}; return GSR_OK; }
    if (n_dst > gsr::PEER_MAX || !drgb_dst || !campos_dst || !campos_src || !side_stream) {
        set_error("gsr_backward_early_rgb_peers: bad argument (at most %d destinations, a side stream is required)", gsr::PEER_MAX);
        return GSR_EINVAL;
    }
    gsr::EarlyRgb e{nullptr, ready, user, true, {}};
    e.peers.n = n_dst; e.peers.campos_src = campos_src; e.peers.side = (cudaStream_t)side_stream;
    for (int k = 0; k < n_dst; ++k) {
        if (!drgb_dst[k] || !campos_dst[k]) { set_error("gsr_backward_early_rgb_peers: null destination %d", k); return GSR_EINVAL; }
        e.peers.drgb[k] = drgb_dst[k]; e.peers.campos[k] = campos_dst[k];
    }
    gsr::g_early = e;
    return GSR_OK;
}

int gsr_backward_early_rgb_multicast(float* mc_drgb_slot, float* mc_campos_slot, const float* campos_src,
                                     void* side_stream, void (*ready)(void*), void* user) {
    if (!mc_drgb_slot) { gsr::g_early = gsr::EarlyRgb{nullptr, nullptr, nullptr, false, {}}; return GSR_OK; }
    if (!mc_campos_slot || !campos_src || !side_stream || ((uintptr_t)mc_drgb_slot & 3) || ((uintptr_t)mc_campos_slot & 3)) {
        set_error("gsr_backward_early_rgb_multicast: bad argument (4-byte aligned multicast slots and a side stream are required)");
        return GSR_EINVAL;
    }
    gsr::EarlyRgb e{nullptr, ready, user, true, {}};
    e.peers.n = 0; e.peers.campos_src = campos_src; e.peers.side = (cudaStream_t)side_stream;
    e.peers.mc_drgb = mc_drgb_slot; e.peers.mc_campos = mc_campos_slot;
    gsr::g_early = e;
    return GSR_OK;
}

int gsr_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, void* stream)
{
    (void)projmatrix;   // the reference computes p_proj and never uses it (auxiliary.h:147-154)
    if (P == 0) return GSR_OK;
    if (P < 0 || !means3D || !viewmatrix || !present) { set_error("bad argument to gsr_mark_visible"); return GSR_EINVAL; }
    return check_cuda(launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream), "mark_visible launch");
}

}  // extern "C"
