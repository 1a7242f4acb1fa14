/*
 * gsr_b200.h -- C ABI of the B200-native Gaussian-splatting rasterizer (libgsr_b200.so).
 *
 * This is the drop-in boundary for SCORP's rasterizer hot path. Every entry point takes
 * plain device pointers, sizes and a CUDA stream handle (passed as void*); there are no
 * torch or C++ types in any signature. All pointers are DEVICE pointers owned by the
 * caller unless marked "host". The library allocates no device memory: scratch that must
 * survive until the backward pass is obtained through a caller-supplied allocator callback
 * (the same shape as the reference's std::function<char*(size_t)> resize lambdas,
 * reference submodules/diff-gaussian-rasterization/rasterize_points.cu:27-33,75-80).
 *
 * Return value: 0 on success, a negative GSR_E* code otherwise; gsr_last_error() returns a
 * host string describing the last failure on the calling thread. No exception crosses the ABI.
 *
 * Reference interfaces replaced (paths relative to /root/reference/submodules):
 *   gsr_rasterize_forward   <- CudaRasterizer::Rasterizer::forward
 *                              diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h:29-55
 *                              (called from RasterizeGaussiansCUDA, rasterize_points.cu:36-119)
 *   gsr_rasterize_backward  <- CudaRasterizer::Rasterizer::backward, rasterizer.h:57-92
 *                              (RasterizeGaussiansBackwardCUDA, rasterize_points.cu:122-208)
 *   gsr_mark_visible        <- CudaRasterizer::Rasterizer::markVisible, rasterizer.h:22-27
 *                              (markVisible, rasterize_points.cu:210-229)
 *   gsr_surfel_forward      <- diff-surfel-rasterization/cuda_rasterizer/rasterizer.h:29-54
 *   gsr_surfel_backward     <- diff-surfel-rasterization/cuda_rasterizer/rasterizer.h:56-88
 *   gsr_knn_dist2           <- SimpleKNN::knn, simple-knn/simple_knn.h:15-19
 *                              (distCUDA2, simple-knn/spatial.cu:15-26)
 */
#ifndef GSR_B200_H_INCLUDED
#define GSR_B200_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSR_OK            0
#define GSR_EINVAL       -1   /* bad argument (null pointer, negative size, ...)        */
#define GSR_ECUDA        -2   /* a CUDA runtime call or kernel launch failed            */
#define GSR_EALLOC       -3   /* the allocator callback returned NULL / too little      */
#define GSR_EUNSUPPORTED -4   /* e.g. SH coefficient count other than 0,1,4,9,16        */

/* Scratch buffers requested through the allocator callback. */
#define GSR_BUF_GEOM    0     /* per-Gaussian state, kept for backward                  */
#define GSR_BUF_BINNING 1     /* per-instance sorted lists, kept for backward           */
#define GSR_BUF_IMAGE   2     /* per-pixel / per-tile state, kept for backward          */

/* Must return a device pointer to at least `bytes` bytes, 256-byte aligned, that stays
 * valid until the matching backward has run. Called on the calling host thread. */
typedef void* (*gsr_alloc_fn)(void* user, int which, size_t bytes);

/* Mirrors GaussianRasterizationSettings (diff_gaussian_rasterization/__init__.py:160-172).
 * bg[3], viewmatrix[16], projmatrix[16], campos[3] are DEVICE pointers (the reference keeps
 * them as CUDA tensors); matrices are the transposed, column-major-indexed 4x4s of
 * gs3dgs/scene/cameras.py:82-97. */
typedef struct GsrSettings {
    int32_t image_height;
    int32_t image_width;
    float   tanfovx;
    float   tanfovy;
    float   scale_modifier;
    int32_t sh_degree;
    int32_t prefiltered;
    int32_t debug;          /* !=0: synchronise + check after every stage (auxiliary.h:166-173) */
    const float* bg;
    const float* viewmatrix;
    const float* projmatrix;
    const float* campos;
} GsrSettings;

const char* gsr_last_error(void);
int         gsr_version(void);

/* Optional per-stage device timing (no reference counterpart; the reference only has the
 * caller-side torch.cuda.Event pair of train_3dgs.py:48-49). When enabled, CUDA events are
 * recorded on the caller's stream between stages; gsr_profile_read waits for the recorded
 * passes (up to 64 since enable) and fills ms[GSR_PROF_STAGES] with their mean: 0 preprocess fwd, 1 binning, 2 blend fwd, 3 blend bwd (incl.
 * accumulator memset), 4 preprocess bwd; -1 where nothing was recorded. */
#define GSR_PROF_STAGES 5
void gsr_profile_enable(int on);
int  gsr_profile_read(float* ms);

/* ---- 3DGS (diff_gaussian_rasterization) ------------------------------------------------
 * P Gaussians, M SH coefficients per Gaussian (0 when colors_precomp is given).
 * Optional inputs are NULL when absent (the reference's 0-element tensor sentinel):
 *   exactly one of shs[P*M*3] / colors_precomp[P*3];
 *   exactly one of (scales[P*3], rotations[P*4]) / cov3D_precomp[P*6].
 * Outputs: out_color[3*H*W], out_depth[H*W], out_alpha[H*W], radii[P] (int32).
 * *num_rendered (host) receives the number of (Gaussian, tile) instances. */
int gsr_rasterize_forward(
    const GsrSettings* s, int P, int M,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* opacities, const float* scales, const float* rotations,
    const float* cov3D_precomp,
    float* out_color, float* out_depth, float* out_alpha, int32_t* radii,
    gsr_alloc_fn alloc, void* alloc_user, int* num_rendered, void* stream);

/* Workspace for the backward pass (per-Gaussian gradient accumulators); the caller passes a
 * device buffer of this many bytes. It need not be zeroed. */
size_t gsr_backward_workspace_bytes(int P);

/* R = num_rendered from the forward. geom/binning/image are the buffers the forward obtained
 * through `alloc` (read-only here: backward may run several times per forward).
 * Every dL_d* output is fully written (zeros for invisible Gaussians): no pre-zeroing needed.
 * Shapes: dL_dmeans2D[P*3], dL_dcolors[P*3], dL_dopacity[P], dL_dmeans3D[P*3],
 * dL_dcov3D[P*6], dL_dsh[P*M*3], dL_dscales[P*3], dL_drotations[P*4]. */
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
    void* workspace, size_t workspace_bytes, void* stream);

/* ---- model-space entry points (caller-side parameter staging) --------
 * No reference counterpart at the C++ level: these fold what the reference's caller runs as separate torch kernels
 * right before the boundary -- GaussianModel.get_features = torch.cat(features_dc, features_rest)
 * (gs3dgs/scene/gaussian_model.py:139-142), get_opacity = sigmoid, get_scaling = exp, get_rotation = normalize
 * (:31-46,126-146), all called from gaussian_renderer.render() (gs3dgs/gaussian_renderer/__init__.py:67-90) -- and
 * their autograd mirrors, into the preprocess kernels. Inputs are the tensors GaussianModel STORES:
 *   xyz[P*3], features_dc[P*1*3], features_rest[P*(M-1)*3] (NULL when M == 1), opacity[P] (pre-sigmoid),
 *   scaling[P*3] (pre-exp), rotation[P*4] (un-normalised quaternion).
 * The result equals gsr_rasterize_forward on the activated, concatenated tensors. The stock entry points above are
 * unchanged; this pair is opt-in (scorp_b200/model_space.py). */
typedef struct GsrModelParams {
    const float* xyz;
    const float* features_dc;
    const float* features_rest;
    const float* opacity;
    const float* scaling;
    const float* rotation;
} GsrModelParams;
typedef struct GsrModelGrads {      /* same shapes as GsrModelParams; every row is written */
    float* xyz;
    float* features_dc;
    float* features_rest;
    float* opacity;
    float* scaling;
    float* rotation;
} GsrModelGrads;
int gsr_rasterize_forward_raw(
    const GsrSettings* s, int P, int M, const GsrModelParams* params,
    float* out_color, float* out_depth, float* out_alpha, int32_t* radii,
    gsr_alloc_fn alloc, void* alloc_user, int* num_rendered, void* stream);
/* Gradients with respect to the STORED parameters (through the activations); dL_dmeans2D[P*3] is the screen-space
 * gradient the densification statistics use (train_3dgs.py:180-183). Workspace as for gsr_rasterize_backward. */
int gsr_rasterize_backward_raw(
    const GsrSettings* s, int P, int M, int R, const GsrModelParams* params,
    const int32_t* radii, const void* geom, const void* binning, const void* image, const float* out_alpha,
    const float* dL_dout_color, const float* dL_dout_depth, const float* dL_dout_alpha,
    float* dL_dmeans2D, const GsrModelGrads* grads, void* workspace, size_t workspace_bytes, void* stream);

/* The same pair for the 2DGS surfel rasterizer (gs2dgs/scene/gaussian_model.py stores the same six tensors; scaling
 * is [P*2]). Outputs as gsr_surfel_forward / gsr_surfel_backward; dL_dtransMat and dL_dcolors are not produced. */
int gsr_surfel_forward_raw(
    const GsrSettings* s, int P, int M, const GsrModelParams* params,
    float* out_color, float* out_others, int32_t* radii,
    gsr_alloc_fn alloc, void* alloc_user, int* num_rendered, void* stream);
int gsr_surfel_backward_raw(
    const GsrSettings* s, int P, int M, int R, const GsrModelParams* params,
    const int32_t* radii, const void* geom, const void* binning, const void* image,
    const float* dL_dout_color, const float* dL_dout_others,
    float* dL_dmeans2D, const GsrModelGrads* grads, void* workspace, size_t workspace_bytes, void* stream);

/* present[P] (1 byte each) = view-space z > 0.2 */
int gsr_mark_visible(int P, const float* means3D, const float* viewmatrix,
                     const float* projmatrix, uint8_t* present, void* stream);

/* ---- introspection of the private scratch layout (tests, golden-vector comparison) ------
 * Byte offsets from the (256-byte aligned) base of each buffer. Index meaning:
 * geom:    0 rec (P x 48 B: xy, conic, opacity, depth, power threshold, rgb), 1 cov3D (6P f32),
 *          2 clamped (P u8, bit c = channel c clamped), 3 tiles_touched (P u32),
 *          4 tile_count (tiles u32), 5 counters (4 u32: [0] = num_rendered)
 * image:   0 ranges (tiles x uint2), 1 n_contrib (W*H u32)
 * binning: 0 point_list (N u32), 1 point_list_keys (N u64), 2 sort scratch (N u64) */
void gsr_geom_layout(int P, int tiles, size_t off[6], size_t* total);
void gsr_image_layout(int W, int H, size_t off[2], size_t* total);
void gsr_binning_layout(int N, size_t off[3], size_t* total);

/* ---- 2DGS surfels (diff_surfel_rasterization) ------------------------------------------
 * scales[P*2]; transMat_precomp[P*9] replaces cov3D_precomp; out_others[7*H*W] with channels
 * depth 0, alpha 1, normal 2-4, median depth 5, distortion 6 (auxiliary.h:23-27). */
int gsr_surfel_forward(
    const GsrSettings* s, int P, int M,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* opacities, const float* scales, const float* rotations,
    const float* transMat_precomp,
    float* out_color, float* out_others, int32_t* radii,
    gsr_alloc_fn alloc, void* alloc_user, int* num_rendered, void* stream);

size_t gsr_surfel_backward_workspace_bytes(int P);

int gsr_surfel_backward(
    const GsrSettings* s, int P, int M, int R,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* scales, const float* rotations, const float* transMat_precomp,
    const int32_t* radii,
    const void* geom, const void* binning, const void* image,
    const float* dL_dout_color, const float* dL_dout_others,
    float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D,
    float* dL_dtransMat, float* dL_dsh, float* dL_dscales, float* dL_drotations,
    void* workspace, size_t workspace_bytes, void* stream);

void gsr_surfel_geom_layout(int P, int tiles, size_t off[6], size_t* total);
void gsr_surfel_image_layout(int W, int H, size_t off[4], size_t* total);

/* ---- forward without the host round trip ------------------------------------------------------------------------
 * The reference reads the number of (Gaussian, tile) instances back to the host in the middle of every forward because it
 * sizes the binning buffer from it (cuda_rasterizer/rasterizer_impl.cu:278-286: cudaMemcpy of the scan's last element);
 * gsr_rasterize_forward / gsr_surfel_forward keep that behaviour. These variants take the size from the CALLER instead:
 * `capacity` = how many instances the binning buffer may hold (e.g. the largest count seen so far plus a margin). Nothing
 * in the call synchronises, so the host runs ahead of the GPU and the call can be captured in a CUDA graph.
 * status_pinned: page-locked host uint32[4]; {num_rendered, longest tile list, overflow flag, -} arrive there behind the
 * kernels and are valid after the caller's next synchronisation of `stream`. If num_rendered > capacity the overflow flag
 * is 1, NOTHING was binned or blended (the outputs hold the background) and the call must be repeated with a larger
 * capacity -- the caller checks the flag at the point where it synchronises anyway (loss.item() in train_3dgs.py:158).
 * The binning buffer is requested with the size for `capacity`; pass R = capacity to the backward call. */
int gsr_rasterize_forward_nosync(
    const GsrSettings* s, int P, int M,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* opacities, const float* scales, const float* rotations, const float* cov3D_precomp,
    float* out_color, float* out_depth, float* out_alpha, int32_t* radii,
    gsr_alloc_fn alloc, void* alloc_user, int capacity, uint32_t* status_pinned, void* stream);
int gsr_surfel_forward_nosync(
    const GsrSettings* s, int P, int M,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* opacities, const float* scales, const float* rotations, const float* transMat_precomp,
    float* out_color, float* out_others, int32_t* radii,
    gsr_alloc_fn alloc, void* alloc_user, int capacity, uint32_t* status_pinned, void* stream);

/* ---- simple_knn.distCUDA2 ---------------------------------------------------------------
 * out[P] = mean squared distance to the 3 nearest other points. tmp: device scratch of
 * gsr_knn_workspace_bytes(P) bytes. */
size_t gsr_knn_workspace_bytes(int P);
int gsr_knn_dist2(int P, const float* xyz, float* out, void* tmp, size_t tmp_bytes, void* stream);
/* Allocation-free variant of gsr_knn_dist2. NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 * Same result as gsr_knn_dist2, bit for bit, with the scratch in a static __device__ array inside the library instead of
 * `tmp`: the caller passes only xyz[P*3] and out[P]; the call performs no cudaMalloc / cudaFree and calls no allocating
 * library routine. Limit: P <= GSR_KNN_STATIC_MAX_POINTS (gsr_knn_static_max_points()); a larger P is rejected with
 * GSR_EINVAL and a message naming the limit. The array (about 6.3 MB) is held by every GPU that loads the library until
 * the process exits. All calls on a device share it: the call re-initialises what must start clean, so results do not
 * depend on the previous call, but calls that OVERLAP on the GPU (two streams, two host threads) race -- serialise them. */
#define GSR_KNN_STATIC_MAX_POINTS 262144
int gsr_knn_static_max_points(void);
int gsr_knn_dist2_static(int P, const float* xyz, float* out, void* stream);

/* ---- view-sharded multi-GPU helpers (no reference counterpart; the reference is single-GPU) ----
 * The SH gradient of ONE view is an outer product, dL_dsh[k] = b_k(dir) * dL_dRGB with dir = normalize(mean - campos)
 * and dL_dRGB zeroed on the channels the forward clamped (diff-gaussian-rasterization/cuda_rasterizer/backward.cu:20-141).
 * Ranks can therefore exchange 3 floats per Gaussian per view (all-gather) instead of all-reducing 3M floats, and
 * rebuild the sum over views locally:
 *   gsr_clamped_color_grad : drgb[P][3] = dL_dcolors with the forward's clamp bits applied (geom_buffer of that view's
 *                            forward; surfel != 0 for a gsr_surfel_forward buffer)
 *   gsr_sh_grad_from_views : dL_dsh[P][M][3] = sum over v < n_views, in order, of b(normalize(means3D - campos[v])) (x)
 *                            drgb[v][P][3]
 * gsr_rasterize_backward / gsr_surfel_backward accept dL_dsh == NULL with shs given and then skip the SH rows. */
int gsr_clamped_color_grad(int P, int W, int H, int surfel, const void* geom_buffer, const float* dL_dcolors,
                           float* drgb, void* stream);
/* Arms the NEXT gsr_rasterize_backward / gsr_surfel_backward call made by this thread: right after its blend kernel
 * (i.e. before the preprocess-backward kernel is enqueued) it writes the clamped colour gradient to drgb[P][3] and
 * calls ready(user) on the host, so the caller can start the all-gather of drgb underneath the rest of the backward.
 * drgb == NULL disarms. Only takes effect when the backward is given SH coefficients. */
int gsr_backward_early_rgb(float* drgb, void (*ready)(void* user), void* user);
/* Peer-memory variant: instead of writing drgb locally for a later all-gather, the next backward writes it straight
 * into every rank's gather buffer through NVLink peer pointers -- drgb_dst[k] / campos_dst[k] (host arrays of n_dst
 * device pointers, n_dst <= 16) are THIS rank's slot in rank k's buffer ([P*3] floats and [3] floats; the local rank
 * included) -- from a kernel on side_stream that the library orders after its blend kernel, so the stores travel
 * underneath the preprocess-backward kernel. ready(user) is called once that kernel is enqueued; the caller then
 * synchronises the ranks (scorp_b200/multiview.py:PeerExchange uses torch symmetric memory for pointers and barrier).
 * n_dst <= 0 disarms. */
int gsr_backward_early_rgb_peers(int n_dst, float* const* drgb_dst, float* const* campos_dst, const float* campos_src,
                                 void* side_stream, void (*ready)(void* user), void* user);
/* Multicast variant of gsr_backward_early_rgb_peers: mc_drgb_slot / mc_campos_slot are the MULTICAST (NVLS) addresses of
 * this rank's slot ([P*3] floats, [3] floats) in a symmetric gather buffer; the push kernel issues one multimem.st per 16
 * bytes and the NVSwitch replicates it into every rank's buffer, so the SMs and the outbound link carry the colour
 * gradient once instead of once per rank. NULL disarms. */
int gsr_backward_early_rgb_multicast(float* mc_drgb_slot, float* mc_campos_slot, const float* campos_src,
                                     void* side_stream, void (*ready)(void* user), void* user);
/* Sum of the four non-SH parameter gradients over ranks, done in the NVSwitch (NVLS) in two phases so that the bytes a GPU
 * moves do not grow with the number of ranks. `head` is a symmetric buffer (same offset on every rank, with a multicast
 * mapping) of P rows of 12 floats (means3D xyz, opacity | scales, rotations, padding), 16-byte aligned.
 *   gsr_backward_peer_head(local replica, chunks, chunk_done, user)  arms the NEXT gsr_rasterize_backward /
 *       gsr_surfel_backward of this thread (SH colours only): its preprocess-backward kernel stores every Gaussian's row
 *       (zeros when invisible) into the local replica; dL_dmeans3D / dL_dopacity / dL_dscales / dL_drotations are then NOT
 *       written by that call. The kernel runs as `chunks` launches over consecutive row ranges and chunk_done(user, k,
 *       row0, rows) is called on the host right after launch k is enqueued, so the caller can reduce range k on another
 *       stream while range k+1 is being computed. NULL disarms.
 *   -- cross-rank barrier --
 *   gsr_peer_reduce_shard(multicast address, row0, rows, rank, world)  rank r sums its 1/world share of rows
 *       [row0, row0 + rows) over all replicas with multimem.ld_reduce and stores the sums into all replicas with multimem.st.
 *   -- cross-rank barrier --
 *   gsr_peer_consume(local replica, dst, P, row0, rows)  un-interleaves the summed rows into dst = the planar head
 *       [means3D 3P | opacity P | scales 3P (2P surfels) | rotations 4P]; gsr_peer_finish does that for all P rows inside
 *       the SH rebuild kernel (gsr_sh_grad_from_views with the same first nine arguments), one launch for both. */
int gsr_backward_peer_head(float* symmetric_head, int chunks, void (*chunk_done)(void* user, int k, int row0, int rows), void* user);
int gsr_peer_reduce_shard(float* multicast_head, int row0, int rows, int rank, int world, void* stream);
int gsr_peer_consume(const float* symmetric_src, float* dst, int P, int row0, int rows, int surfel, void* stream);
int gsr_peer_finish(int P, int M, int D, int n_views, const float* means3D, const float* campos, const float* drgb,
                    float* dL_dsh, const float* head_rows, float* head_dst, int surfel, void* stream);
int gsr_sh_grad_from_views(int P, int M, int D, int n_views, const float* means3D, const float* campos,
                           const float* drgb, float* dL_dsh, void* stream);

/* ---- batched multi-view forward ----------------------------------------
 * One call renders the SAME Gaussians from V cameras of identical resolution and intrinsics: the pose sweep of
 * align_3dgs_clpe_9dof.py:146-160 and utils/mask.py:120-122, which the reference runs as a Python loop of V forward
 * calls (V host synchronisations on num_rendered, 3V allocations). Here the V views go through ONE preprocess launch
 * (grid.y = view), ONE counting sort over V * tiles tile lists (the view index takes the place of the key's high
 * bits) and ONE blend launch over V * tiles CTAs, with a single readback.
 * s->viewmatrix / s->projmatrix / s->campos point to V consecutive matrices / positions ([V*16], [V*16], [V*3]);
 * tanfov, resolution, bg, scale_modifier and sh_degree are shared. Outputs: out_color[V*3*H*W], out_depth[V*H*W],
 * out_alpha[V*H*W], radii[V*P]; view v's slice is bit-identical to gsr_rasterize_forward with camera v.
 * *num_rendered is the total over the views. Forward only: the scratch buffers have the batched layout and are not
 * accepted by gsr_rasterize_backward. */
int gsr_rasterize_forward_views(
    const GsrSettings* s, int V, int P, int M,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* opacities, const float* scales, const float* rotations,
    const float* cov3D_precomp,
    float* out_color, float* out_depth, float* out_alpha, int32_t* radii,
    gsr_alloc_fn alloc, void* alloc_user, int* num_rendered, void* stream);

/* ---- photometric loss (the step right after the boundary) ----------------
 * Replaces l1_loss + ssim of gs3dgs/utils/loss_utils.py:17-18,45-73 (five grouped 11x11 F.conv2d calls and ~20
 * elementwise torch kernels per forward, mirrored by autograd) as used by train_3dgs.py:106-107:
 *     loss = (1 - lambda_dssim) * mean|img - gt| + lambda_dssim * (1 - mean(ssim_map(img, gt)))
 * img, gt: [C*H*W] fp32 (C planes, F.conv2d zero padding 5, window gaussian(11, 1.5)).
 * forward writes out3 (device) = { loss, l1, ssim } and fills the workspace (three derivative maps + per-CTA
 * partial sums; gsr_l1_ssim_workspace_bytes) that backward reads. Deterministic (no float atomics).
 * backward writes dL_dimg[C*H*W] = upstream * (w_l1 * d l1/d img + w_ssim * d ssim/d img); upstream is a DEVICE
 * scalar (NULL = 1). For the fused training loss pass w_l1 = 1 - lambda, w_ssim = -lambda; (1, 0) and (0, 1) give the
 * gradients of l1_loss and ssim alone. gt receives no gradient (it never does in the reference's loops). */
size_t gsr_l1_ssim_workspace_bytes(int C, int H, int W);
int gsr_l1_ssim_forward(int C, int H, int W, const float* img, const float* gt, float lambda_dssim,
                        float* out3, void* workspace, size_t workspace_bytes, void* stream);
int gsr_l1_ssim_backward(int C, int H, int W, const float* img, const float* gt, const float* upstream,
                         float w_l1, float w_ssim, const void* workspace, float* dL_dimg, void* stream);

/* ---- optimizer step and densification statistics ------------------------
 * gsr_adam_step replaces gaussians.optimizer.step() -- torch.optim.Adam(l, lr=0.0, eps=1e-15) over the six parameter
 * groups (gs3dgs/scene/gaussian_model.py:192-206, train_3dgs.py:192), i.e. torch's seven-pass multi-tensor update --
 * with one launch over all groups: param, exp_avg, exp_avg_sq updated in place from grad, bit-identical to torch
 * (amsgrad, weight_decay and maximize are off in the reference and not built). `groups` is a HOST array of at most 8
 * entries; n = number of floats; lr per group; step = the 1-based step count after the increment, as torch uses it. */
typedef struct GsrAdamGroup {
    float*       param;
    const float* grad;
    float*       exp_avg;
    float*       exp_avg_sq;
    int64_t      n;
    double       lr;
} GsrAdamGroup;
int gsr_adam_step(int n_groups, const GsrAdamGroup* groups, double beta1, double beta2, double eps, int step, void* stream);
/* For every Gaussian with radii > 0 (train_3dgs.py:180-181, gaussian_model.py:603-605):
 *   max_radii2D = max(max_radii2D, radii); xyz_gradient_accum += |dL_dmeans2D[:, :2]|; denom += 1.
 * dL_dmeans2D has grad_stride floats per Gaussian (3); max_radii2D may be NULL. No host synchronisation (the
 * reference's boolean-mask indexing runs nonzero()). */
int gsr_densification_stats(int P, const int32_t* radii, const float* dL_dmeans2D, int grad_stride,
                            float* xyz_gradient_accum, float* denom, float* max_radii2D, void* stream);

/* GaussianModel.reset_opacity (gaussian_model.py:253-256; train_3dgs.py:187-188) in place:
 *   opacity = inverse_sigmoid(min(sigmoid(opacity), cap)), cap = 0.01, and -- replace_tensor_to_optimizer, :410-423 --
 *   the opacity group's Adam moments zeroed (either may be NULL). */
int gsr_reset_opacity(int P, float* opacity, float* exp_avg, float* exp_avg_sq, float cap, void* stream);

/* --- densify_and_prune ------------------------------------------------------------------------------------------
 * Replaces GaussianModel.densify_and_prune (gs3dgs/scene/gaussian_model.py:572-586, gs2dgs/scene/gaussian_model.py:
 * 478-492) with everything below it: densify_and_clone (:559-570), densify_and_split (:502-527),
 * densification_postfix (:482-500), cat_tensors_to_optimizer (:460-480), prune_points (:443-458). The reference's four
 * rounds of mask indexing and torch.cat become one per-Gaussian decision (plan) and one gather of every surviving row
 * of the parameters and Adam moments (apply). Result rows, in the reference's order: surviving originals, surviving
 * clones, then the n_split copies of the surviving split children (copy-major).
 * Thresholds are the reference's Python floats rounded to float:
 *   grad_threshold = max_grad, dense_extent = percent_dense * extent, min_opacity, big_extent = 0.1 * extent,
 *   prune_big = bool(max_screen_size), max_screen_size (only its sign matters: max_radii2D has just been zeroed,
 *   :498), n_split = N (2). */
typedef struct GsrDensifyCfg {
    float grad_threshold;
    float dense_extent;
    float min_opacity;
    float big_extent;
    float max_screen_size;
    int   prune_big;
    int   n_split;
} GsrDensifyCfg;
size_t gsr_densify_workspace_bytes(int P, int n_split);
/* Plan. scaling [P][scale_dim] (3: 3DGS, 2: 2DGS) and opacity [P] are the STORED (pre-activation) tensors;
 * xyz_gradient_accum, denom [P]. Writes the decision into workspace and returns, after one stream synchronisation,
 * counts = { surviving originals, surviving clones, surviving split parents (each yields n_split rows),
 *            split-selected Gaussians before the prune (the reference draws n_split * counts[3] normal samples) }. */
int gsr_densify_plan(int P, int scale_dim, const float* scaling, const float* opacity, const float* xyz_gradient_accum,
                     const float* denom, const GsrDensifyCfg* cfg, void* workspace, int32_t counts[4], void* stream);
#define GSR_DENSIFY_COPY    0   /* rows are copied                                                   */
#define GSR_DENSIFY_XYZ     1   /* children: R(rotation) . (normal * exp(scaling)) + xyz     (:511-515) */
#define GSR_DENSIFY_SCALING 2   /* children: log(exp(scaling) / (0.8 * n_split))             (:516)     */
typedef struct GsrDensifyGroup {
    const float* param;         /* [P][width]                                         */
    const float* exp_avg;       /* [P][width] or NULL (no optimizer state yet)        */
    const float* exp_avg_sq;
    float*       out_param;     /* [rows][width], rows = counts[0] + counts[1] + n_split * counts[2] */
    float*       out_exp_avg;   /* [rows][width] or NULL; new rows are zero           */
    float*       out_exp_avg_sq;
    int          width;
    int          kind;
} GsrDensifyGroup;
/* Apply the plan held in workspace: up to 8 groups in one gather launch. rotation [P][4] and scaling [P][scale_dim]
 * are the stored tensors, normals [n_split * counts[3]][3] standard normal samples (what torch.normal draws before it
 * scales by std); needed only when a GSR_DENSIFY_XYZ group is present and counts[2] > 0. No synchronisation. */
int gsr_densify_apply(int P, int scale_dim, const GsrDensifyCfg* cfg, void* workspace, const int32_t counts[4],
                      int n_groups, const GsrDensifyGroup* groups, const float* rotation, const float* scaling,
                      const float* normals, void* stream);

/* --- PLY vertex rows ---------------------------------------------------------------------------------------------
 * The tensor shuffling of GaussianModel.save_ply / load_ply (gs3dgs/scene/gaussian_model.py:220-251, 287-332; same
 * pair in gs2dgs/scene/gaussian_model.py) as one kernel each, so the model crosses PCIe in a single copy. A vertex row
 * is  x y z nx ny nz f_dc_0..2 f_rest_0..3R-1 opacity scale_0..S-1 rot_0..3  (float32, normals zero, SH channel-major:
 * f_rest_{c*R+k} = features_rest[i][k][c], :238-239). The PLY header is written / parsed by the host layer
 * (scorp_b200/ply_io.py). */
int gsr_ply_row_floats(int sh_rest, int scale_dim);
/* rows [P][gsr_ply_row_floats(sh_rest, scale_dim)] (device) <- the stored model tensors (sh_rest = R coefficients
 * per channel besides DC; features_rest may be NULL when R == 0). */
int gsr_ply_pack(int P, int sh_rest, int scale_dim, const GsrModelParams* model, float* rows, void* stream);
/* The inverse for any property order: rows [P][row_floats] (device, float32 properties only) and a HOST array
 * columns[3 + 3 + 3R + 1 + S + 4] giving, for every model float in the order xyz, features_dc[c], features_rest[k][c],
 * opacity, scaling, rotation, the column of the row it is read from. `out` tensors have GsrModelParams' shapes. */
int gsr_ply_unpack(int P, int row_floats, const float* rows, const int32_t* columns, int sh_rest, int scale_dim,
                   const GsrModelGrads* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GSR_B200_H_INCLUDED */
