// common.cuh -- shared declarations for the sm_100a Gaussian-splatting kernels.
//
// Conventions (reference: submodules/diff-gaussian-rasterization/cuda_rasterizer/):
//  * 16x16 pixel tiles (config.h:15-17); one CTA of 256 threads per tile in the blend kernels.
//  * viewmatrix/projmatrix are indexed column-major, m[4*col+row] (auxiliary.h:58-87).
//  * Integer-deciding arithmetic (depth key, radius, tile rect) is written with explicit
//    __fmul_rn/__fadd_rn/__fmaf_rn so the rounding sequence is fixed by the source and equals
//    the sequence nvcc 12.9 emits for the reference (default -fmad contraction); see DESIGN.md.
#pragma once
#include <cstdint>
#include <cstddef>
#include <cuda_runtime.h>
#include "../../include/gsr_b200.h"

namespace gsr {

constexpr int TILE = 16;            // BLOCK_X == BLOCK_Y == 16 (config.h:16-17)
constexpr int TILE_PIX = TILE * TILE;
constexpr int BATCH = 256;          // Gaussians staged per round (== BLOCK_SIZE in the reference)
// Per-tile instance counters (histogram in preprocess, cursor in scatter) sit TC_STRIDE words apart, one per 32-byte
// sector: packed 4 bytes apart, the counters of a whole view share a few hundred cache lines, and atomics to one line
// serialise in L2. Measured on B200 (binning stage, C3 / C2 / C1): packed 0.919 / 0.153 / 0.060 ms, one per 32-byte sector
// 0.850 / 0.140 / 0.054 ms, one per 128-byte line 0.867 / 0.145 / 0.055 ms; preprocess (histogram atomics) gains too
// (C1: 0.020 -> 0.013 ms). Index with tc(tile).
constexpr int TC_STRIDE = 8;
__host__ __device__ inline size_t tc(size_t tile) { return tile * TC_STRIDE; }
constexpr int GEOM_COUNTERS = 8;    // uint32 slots in GeomState::counters / SurfelGeom::counters (binning.cu: tile_scan_kernel)

// Per-Gaussian render record, 48 bytes, 16-byte aligned: what the blend kernels stage.
struct __align__(16) GaussRec {
    float x, y;          // pixel-space mean           (points_xy_image)
    float ca, cb;        // conic.x, conic.y           (conic_opacity.xy)
    float cc, opacity;   // conic.z, opacity           (conic_opacity.zw)
    float depth;         // view-space z               (depths)
    float thr;           // conservative power threshold: power < thr  =>  alpha < 1/255
    float r, g, b;       // colour                     (rgb / colors_precomp)
    float pad;
};
static_assert(sizeof(GaussRec) == 48, "GaussRec must be 48 bytes");

// Private scratch layouts (see gsr_*_layout in include/gsr_b200.h).
struct GeomState {
    GaussRec* rec;
    float*    cov3D;
    uint8_t*  clamped;
    uint32_t* tiles_touched;
    uint32_t* tile_count;
    uint32_t* counters;
};
struct ImageState {
    uint2*    ranges;
    uint32_t* n_contrib;
    uint32_t* tile_order;     // tiles by decreasing list length: CTA b of a blend kernel takes tile_order[b]
};
struct BinningState {
    uint32_t* point_list;
    uint64_t* point_list_keys;
    uint64_t* scratch;
    uint4*    contrib;        // forward -> backward: per 32 list entries of a tile, one word per warp block with a bit
                              // for every entry that contributed to some pixel of that block (3DGS blends)
};
// first contrib slot of a tile whose list starts at `first`: slots of different tiles never overlap because
// floor((x+len)/32) + 1 >= floor(x/32) + ceil(len/32)
__host__ __device__ inline size_t contrib_base(uint32_t first, int tile) { return (size_t)(first >> 5) + (size_t)tile; }

inline size_t align_up(size_t v, size_t a = 256) { return (v + a - 1) / a * a; }

inline size_t carve_geom(char* base, int P, int tiles, GeomState* g, size_t* off = nullptr) {
    size_t o = 0, k = 0;
    auto take = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes); if (off) off[k] = at; ++k; return at; };
    size_t a0 = take((size_t)P * sizeof(GaussRec));
    size_t a1 = take((size_t)P * 6 * sizeof(float));
    size_t a2 = take((size_t)P);
    size_t a3 = take((size_t)P * 4);
    size_t a4 = take((size_t)tiles * 4 * TC_STRIDE);
    size_t a5 = take(GEOM_COUNTERS * 4);
    if (g) {
        g->rec = (GaussRec*)(base + a0); g->cov3D = (float*)(base + a1);
        g->clamped = (uint8_t*)(base + a2); g->tiles_touched = (uint32_t*)(base + a3);
        g->tile_count = (uint32_t*)(base + a4); g->counters = (uint32_t*)(base + a5);
    }
    return o;
}
inline size_t carve_image(char* base, int W, int H, ImageState* s, size_t* off = nullptr, int views = 1) {
    int tiles = ((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE) * views;
    size_t o = 0;
    size_t a0 = o; o = align_up(o + (size_t)tiles * sizeof(uint2));
    size_t a1 = o; o = align_up(o + (size_t)W * H * 4 * views);
    size_t a2 = o; o = align_up(o + (size_t)tiles * 4);
    if (off) { off[0] = a0; off[1] = a1; }
    if (s) { s->ranges = (uint2*)(base + a0); s->n_contrib = (uint32_t*)(base + a1); s->tile_order = (uint32_t*)(base + a2); }
    return o;
}
inline size_t contrib_slots(int N, int tiles) { return (size_t)(N > 0 ? N : 0) / 32 + (size_t)tiles + 2; }
// one slot = 32 list entries of a tile; 3DGS tiles have 4 warp blocks (4 words per slot), surfel tiles 8 (8 words per slot)
constexpr int CONTRIB_WORDS_MAX = 8;
inline size_t contrib_bytes(int N, int tiles) { return contrib_slots(N, tiles) * CONTRIB_WORDS_MAX * sizeof(uint32_t); }
inline size_t carve_binning(char* base, int N, BinningState* b, size_t* off = nullptr, int tiles = 0) {
    size_t n = (size_t)(N > 0 ? N : 1);
    size_t o = 0;
    size_t a0 = o; o = align_up(o + n * 4);
    size_t a1 = o; o = align_up(o + n * 8);
    size_t a2 = o; o = align_up(o + n * 8);
    size_t a3 = o; o = align_up(o + contrib_bytes(N, tiles));                     // last: the public offsets do not move
    if (off) { off[0] = a0; off[1] = a1; off[2] = a2; }
    if (b) { b->point_list = (uint32_t*)(base + a0); b->point_list_keys = (uint64_t*)(base + a1);
             b->scratch = (uint64_t*)(base + a2); b->contrib = (uint4*)(base + a3); }
    return o;
}

// gsr_backward_early_rgb (api.cu): armed per thread for the next backward call
constexpr int PEER_MAX = 16;
struct PeerDst {                    // gsr_backward_early_rgb_peers: one destination per rank (this rank's slot in that rank's buffer)
    int n = 0;
    float* drgb[PEER_MAX] = {};
    float* campos[PEER_MAX] = {};
    const float* campos_src = nullptr;
    cudaStream_t side = nullptr;    // the push kernel runs here, ordered after the blend backward by an event
    // gsr_backward_early_rgb_multicast: ONE multicast (NVLS) address of this rank's slot instead of n peer pointers; the
    // switch replicates every store into all replicas (n = 0 then)
    float* mc_drgb = nullptr;
    float* mc_campos = nullptr;
};
struct EarlyRgb { float* drgb; void (*ready)(void*); void* user; bool armed; PeerDst peers; };
EarlyRgb take_early_rgb();          // returns the armed request (armed == false if none) and disarms it
cudaError_t launch_clamped_rgb_from_acc(int P, const int32_t* radii, const uint8_t* clamped, const float* acc, int stride,
                                        int rgb_off, float* drgb, cudaStream_t st);
// head rows [P][12] -> the planar arena head [means3D 3P | opacity P | scales 3P (2P) | rotations 4P]
cudaError_t launch_peer_consume(const float* src, float* dst, int P, int row0, int rows, bool surfel, cudaStream_t st);
// the same values written straight into every rank's gather buffer through peer pointers (all-gather by stores)
cudaError_t launch_clamped_rgb_push(int P, const int32_t* radii, const uint8_t* clamped, const float* acc, int stride,
                                    int rgb_off, const PeerDst& dst, cudaStream_t main);

// gsr_backward_peer_head (api.cu): armed per thread for the next backward call; take_peer_head returns it and disarms
struct PeerHead {
    float* head = nullptr;              // local replica of the symmetric head rows (NULL: not armed)
    int chunks = 1;                     // the preprocess-backward runs as this many launches over consecutive row ranges ...
    void (*chunk_done)(void*, int, int, int) = nullptr;     // ... and this is called (user, k, row0, rows) after each is enqueued
    void* user = nullptr;
};
PeerHead take_peer_head();
// shared by api.cu / surfel_api.cu: runs launch(row_begin, row_end) over the armed chunking (one launch when not armed)
template <class Launch>
inline cudaError_t launch_in_row_chunks(int P, int tile, const PeerHead& ph, Launch&& launch) {
    const int chunks = (ph.head && ph.chunks > 1) ? ph.chunks : 1;
    const int per = ((P + chunks - 1) / chunks + tile - 1) / tile * tile;      // chunk starts stay multiples of the kernel's tile
    for (int k = 0, r0 = 0; r0 < P; ++k, r0 += per) {
        const int r1 = r0 + per < P ? r0 + per : P;
        cudaError_t e = launch(r0, r1);
        if (e != cudaSuccess) return e;
        if (ph.head && ph.chunk_done) ph.chunk_done(ph.user, k, r0, r1 - r0);
    }
    return cudaSuccess;
}
uint32_t* readback_pad();           // api.cu: pinned 64-byte buffer of the calling host thread (NULL if it cannot be had)

// gsr_*_forward_nosync: the binning buffer is sized by the caller (capacity instances) instead of by a device->host
// readback of N; status = pinned host uint32[4] that receives {N, longest segment, overflow flag, 0} asynchronously
struct NoSync { int capacity; uint32_t* status; };

struct StageRanges { ~StageRanges(); };   // api.cu: closes the NVTX stage range an entry point may leave open

// error plumbing (api.cu)
void set_error(const char* fmt, ...);
int  check_cuda(cudaError_t e, const char* what);
int  stage_check(const GsrSettings* s, cudaStream_t st, const char* what);

#ifdef __CUDACC__
// ---- exact-rounding helpers -------------------------------------------------------------
__device__ __forceinline__ float mulr(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float addr(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float subr(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fmar(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ float divr(float a, float b) { return __fdiv_rn(a, b); }

// ---- parameter activations of GaussianModel (gs3dgs/scene/gaussian_model.py:31-46), fused into the model-space
// entry points. Written as the float sequences torch's CUDA kernels execute: sigmoid = 1 / (1 + expf(-x)) with an IEEE
// division, exp = expf, normalize = x / max(sqrt((x0^2 + x2^2) + (x1^2 + x3^2)), 1e-12) (squares rounded; this pairing is what torch's
// reduction kernel does for a [P,4] row: two lanes, each summing a stride-2 pair -- identified on a B200 by
// tools/gpu_diag_norm.py, 0 of 200000 rows differ).
__device__ __forceinline__ float act_opacity(float x) { return divr(1.0f, addr(1.0f, expf(-x))); }
__device__ __forceinline__ float act_scale(float x) { return expf(x); }
__device__ __forceinline__ float quat_norm(float4 q) {
    return fmaxf(__fsqrt_rn(addr(addr(mulr(q.x, q.x), mulr(q.z, q.z)), addr(mulr(q.y, q.y), mulr(q.w, q.w)))), 1e-12f);
}
__device__ __forceinline__ float4 act_rotation(float4 q) {
    const float n = quat_norm(q);
    return make_float4(divr(q.x, n), divr(q.y, n), divr(q.z, n), divr(q.w, n));
}

// row r of (matrix * [p,1]) with the reference's rounding sequence
// (transformPoint4x3/4x4, auxiliary.h:58-80: FMUL(y*m4) FFMA(x,m0) FFMA(z,m8) FADD(m12)).
__device__ __forceinline__ float xform_row(const float* __restrict__ m, int r, float x, float y, float z) {
    float t = mulr(y, m[4 + r]);
    t = fmar(x, m[r], t);
    t = fmar(z, m[8 + r], t);
    return addr(t, m[12 + r]);
}

// ndc2Pix (auxiliary.h:41-44): evaluated in double, DFMA-contracted as nvcc does.
__device__ __forceinline__ float ndc_to_pix(float v, int S) {
    double d = __dadd_rn((double)v, 1.0);
    d = __fma_rn(d, (double)S, -1.0);
    d = __dmul_rn(d, 0.5);
    return (float)d;
}

// getRect (auxiliary.h:46-56). Float divisions by 16 are exact multiplications by 1/16.
__device__ __forceinline__ void tile_rect(float px, float py, int radius, int gx, int gy,
                                          uint32_t& x0, uint32_t& y0, uint32_t& x1, uint32_t& y1) {
    const float rf = (float)radius;
    x0 = min((uint32_t)gx, (uint32_t)max(0, (int)mulr(subr(px, rf), 0.0625f)));
    y0 = min((uint32_t)gy, (uint32_t)max(0, (int)mulr(subr(py, rf), 0.0625f)));
    x1 = min((uint32_t)gx, (uint32_t)max(0, (int)mulr(addr(addr(addr(px, rf), 16.0f), -1.0f), 0.0625f)));
    y1 = min((uint32_t)gy, (uint32_t)max(0, (int)mulr(addr(addr(addr(py, rf), 16.0f), -1.0f), 0.0625f)));
}

// Warp-wide walk over the tiles of 32 rectangles. Every lane brings one rectangle -- t0 = index of its first tile,
// w = width in tiles, cnt = w * h tiles (0: none) -- and, if PAY, a 64-bit payload; f(tile, payload of the rectangle's
// lane) is called once per covered tile. The tiles of all 32 rectangles are laid end to end (warp prefix sum over cnt) and
// taken 32 at a time, one per lane: a lane finds the rectangle its slot belongs to with a 5-step binary search over the
// prefix sums (shuffles), so every round has 32 independent atomics in flight whatever the rectangle sizes are.
// (Round 1/2a: rectangles of <= 4 tiles by their own lane, each larger one spread over the warp in a round of its own:
// at C3, ~14 dependent rounds per warp and 45 % of preprocess_kernel's instructions, most of them the i / w, i % w pair;
// now ceil(sum cnt / 32) = 4.3 rounds and a float reciprocal.) Must be called by all 32 lanes.
template <bool PAY, class F>
__device__ __forceinline__ void warp_rects_packed(int gx, uint32_t t0, uint32_t w, uint32_t cnt, uint64_t pay, F&& f) {
    constexpr uint32_t FULL = 0xffffffffu;
    const uint32_t lane = threadIdx.x & 31;
    uint32_t incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t n = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += n; }
    const uint32_t total = __shfl_sync(FULL, incl, 31);
    const uint32_t excl = incl - cnt;
    const bool fast = total < (1u << 20);       // (i + 0.5) / w in float is then exact enough to floor (error < 0.25 / w)
    for (uint32_t base = 0; base < total; base += 32) {
        const uint32_t j = base + lane;
        uint32_t own = 0, oe = 0;               // the last lane whose first slot is <= j (an empty lane shares its slot with the next)
#pragma unroll
        for (int s = 16; s; s >>= 1) {
            const uint32_t e = __shfl_sync(FULL, excl, own + s);
            if (e <= j) { own += s; oe = e; }
        }
        const uint32_t ot0 = __shfl_sync(FULL, t0, own), ow = __shfl_sync(FULL, w, own);
        uint64_t opay = 0;
        if constexpr (PAY) opay = __shfl_sync(FULL, pay, own);
        if (j < total) {
            const uint32_t i = j - oe;
            const uint32_t row = fast ? (uint32_t)__fdividef((float)i + 0.5f, (float)ow) : i / ow;
            f(ot0 + row * (uint32_t)gx + (i - row * ow), opay);
        }
    }
}
// one is added to tile_count for every tile of every lane's rectangle (origin x0,y0, width w, cnt tiles)
__device__ __forceinline__ void tile_histogram(uint32_t* tile_count, int gx, uint32_t x0, uint32_t y0, uint32_t w, uint32_t cnt) {
    warp_rects_packed<false>(gx, y0 * (uint32_t)gx + x0, w, cnt, 0ull, [&](uint32_t t, uint64_t) { atomicAdd(tile_count + tc(t), 1u); });
}

// ---- view-sharded multi-GPU: the non-SH parameter gradients are summed over ranks in the NVSwitch (NVLS) -------------
// Two-phase (reduce-scatter + all-gather) so the bytes a GPU moves do not grow with the number of ranks:
//   phase 0  the preprocess-backward kernel stores its 12 head floats per Gaussian into its LOCAL replica of a
//            symmetric buffer [P][12] (peer_head_store; every row, zeros for invisible Gaussians: nothing to clear);
//   barrier  (symmetric-memory signal pads, scorp_b200/multiview.py)
//   phase 1  rank r reads rows [r*P/n, (r+1)*P/n) with multimem.ld_reduce -- the switch pulls the row from every
//            replica and returns the sum -- and writes the sum back with multimem.st, which the switch stores into
//            every replica (peer_reduce_shard_kernel, sh_exchange.cu). Per GPU: 48 B x P out, 48 B x P in, for any n.
//   barrier, then phase 2 reads the local replica (peer_finish_kernel).
// (Round 1 added every row into every replica with multimem.red from inside the preprocess-backward kernel: an
// all-reduce made of n multicast reductions, n x 48 B x P_vis of ingress per GPU -- 0.28 ms of the step at 8 GPUs.)
constexpr int PEER_HEAD_STRIDE = 12;        // floats per Gaussian in the head rows (three 16-byte groups)
__device__ __forceinline__ float4 mc_ld_reduce4(const float4* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void mc_st1(float* mc, float v) {
    asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" :: "l"(mc), "f"(v) : "memory");
}
__device__ __forceinline__ void mc_st4(float4* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                 :: "l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// one Gaussian's head row into the local replica (three 16-byte stores)
__device__ __forceinline__ void peer_head_store(float* head, int idx, const float (&v)[PEER_HEAD_STRIDE]) {
    float4* row = reinterpret_cast<float4*>(head + (size_t)idx * PEER_HEAD_STRIDE);
    row[0] = make_float4(v[0], v[1], v[2], v[3]);
    row[1] = make_float4(v[4], v[5], v[6], v[7]);
    row[2] = make_float4(v[8], v[9], v[10], v[11]);
}

// vector reductions (sm_90+): one REDG.E.ADD.F32x4 / F32x2 adds 4 / 2 consecutive floats at a 16- / 8-byte aligned address
__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" :: "l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void red_add_v2(float* p, float a, float b) {
    asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" :: "l"(p), "f"(a), "f"(b) : "memory");
}
// cp.async 16-byte global->shared (LDGSTS), L2-only caching: the records are read once per CTA.
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" :: "r"(d), "l"(gsrc));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" :: "n"(N)); }
#endif  // __CUDACC__

// kernel launchers (one per .cu)
struct FwdArgs {
    int P, M, D, W, H, gx, gy;
    float tanfovx, tanfovy, focal_x, focal_y, scale_modifier;
    const float *means3D, *shs, *colors_precomp, *opacities, *scales, *rotations, *cov3D_precomp;
    const float *view, *proj, *campos, *bg;
    bool prefiltered;
    // model-space variant (gsr_rasterize_forward_raw): shs = features_dc[P,1,3], shs_rest = features_rest[P,M-1,3],
    // opacities / scales / rotations hold the stored (pre-activation) parameters
    const float* shs_rest = nullptr;
    bool raw = false;
    bool sh_wide = false;              // SH rows 32-byte aligned: read with 32-byte loads (set by launch_preprocess)
    bool skip_histogram = false;       // batched views: launch_views_histogram counts the instances per tile instead
};
cudaError_t launch_preprocess(const FwdArgs& a, const GeomState& g, int32_t* radii, cudaStream_t st, int views = 1);
cudaError_t launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t st);
// counters = GeomState::counters ([0] N in, [1] longest segment out, [2..4] see binning.cu); capacity: instances the
// binning buffer holds (no-sync forward), beyond it the overflow flag is raised and every range is left empty
cudaError_t launch_tile_scan(int tiles, uint32_t* tile_count, uint2* ranges, uint32_t* counters, uint32_t* order,
                             cudaStream_t st, uint32_t capacity = 0xffffffffu);
// views > 1: P is the TOTAL record count (views * Gaussians), record v*Pv + i belongs to view v and its tiles are
// numbered v * gx*gy + local tile
cudaError_t launch_binning(int P, int gx, int gy, int N, int max_len, const void* recs, int rec_stride, int depth_off,
                           uint32_t* tile_count, const int32_t* radii, const uint2* ranges, const BinningState& bin,
                           cudaStream_t st, int views = 1, const uint32_t* counters = nullptr, const uint32_t* order = nullptr);
// batched views with a small tile grid: preprocess leaves the per-tile histogram to launch_views_histogram (atomics
// aggregated per CTA in shared memory, binning.cu 3b) and launch_binning scatters the same way
bool views_aggregated(int views, int tiles_per_view, int per_view);
cudaError_t launch_views_histogram(int P_total, int views, int gx, int gy, const void* recs, int rec_stride, const int32_t* radii,
                                   uint32_t* tile_count, cudaStream_t st);
cudaError_t launch_render_forward(int W, int H, int gx, int gy, const GeomState& g, const ImageState& img,
                                  const BinningState& bin, const float* bg,
                                  float* out_color, float* out_depth, float* out_alpha, cudaStream_t st, int views = 1);
cudaError_t launch_render_backward(int W, int H, int gx, int gy, const GeomState& g, const ImageState& img,
                                   const BinningState& bin, const float* bg, const float* out_alpha,
                                   const float* dL_dpix, const float* dL_ddepth, const float* dL_dalpha,
                                   float* acc, cudaStream_t st);
struct BwdArgs {
    int P, M, D, W, H;
    float tanfovx, tanfovy, focal_x, focal_y, scale_modifier;
    const float *means3D, *shs, *colors_precomp, *scales, *rotations, *cov3D;  // cov3D: precomp or geom
    bool cov_precomp;
    const float *view, *proj, *campos;
    const int32_t* radii;
    const uint8_t* clamped;
    const float* acc;  // 12 floats per Gaussian from the blend backward
    float *dL_dmeans2D, *dL_dcolors, *dL_dopacity, *dL_dmeans3D, *dL_dcov3D, *dL_dsh, *dL_dscales, *dL_drotations;
    // model-space variant (gsr_rasterize_backward_raw): split SH rows, gradients through the activations
    const float* shs_rest = nullptr;
    const float* opacities = nullptr;      // raw opacity (only read when raw)
    float* dL_dsh_rest = nullptr;
    bool raw = false;
    // gsr_backward_peer_head: this rank's replica of the symmetric head [P][12] = (means3D xyz, opacity | scales xyz,
    // rot r | rot xyz, 0); the four parameter gradients are stored there (every row) instead of into the arena
    float* peer_head = nullptr;
    int row_begin = 0, row_end = 0;     // Gaussians this launch covers (row_begin a multiple of SH_TILE); set by the caller
};
cudaError_t launch_preprocess_backward(const BwdArgs& a, cudaStream_t st);

// accumulator slots written by the blend backward (12 floats = 48 B per Gaussian)
enum { ACC_MX = 0, ACC_MY = 1, ACC_CA = 2, ACC_CB = 3, ACC_CC = 4, ACC_OP = 5,
       ACC_R = 6, ACC_G = 7, ACC_B = 8, ACC_D = 9, ACC_STRIDE = 12 };

}  // namespace gsr
