// sh_exchange.cu -- view-sharded multi-GPU helper (no reference counterpart).
//
// The SH gradient a view produces for a Gaussian is an outer product: dL/dsh[k] = b_k(dir) * dL/dRGB, with
// dir = normalize(mean - campos) and dL/dRGB zeroed on the channels the forward clamped (backward.cu:20-141).
// Summing the SH gradients of n views across ranks therefore does not need the 3M floats per Gaussian per rank
// an all-reduce moves: every rank all-GATHERS 3 floats per Gaussian per view (plus the camera positions) and
// rebuilds the sum locally, in view order (deterministic, unlike a ring reduction):
//   gsr_clamped_color_grad  : dL_dcolors + the forward's clamp bits -> the 3 floats to exchange
//   gsr_sh_grad_from_views  : sum_v b(dir_v) (x) drgb_v -> dL_dsh[P][M][3]
// With n views in flight the exchange per Gaussian is 12 n bytes instead of 192 (M = 16): worth it up to ~8-12
// views per step; scorp_b200/multiview.py picks. The backward is then called with dL_dsh = NULL and does not
// write the SH rows at all.
#include <cstdlib>
#include "sh.cuh"
#include "surfel.cuh"

namespace gsr {

__global__ void __launch_bounds__(256)
clamped_color_grad_kernel(int P, const uint8_t* __restrict__ clamped, const float* __restrict__ dcol, float* __restrict__ drgb) {
    const int i = blockIdx.x * 256 + threadIdx.x;        // one float per thread: unit-stride in and out
    if (i >= 3 * P) return;
    const int g = i / 3, c = i - 3 * g;
    drgb[i] = ((clamped[g] >> c) & 1) ? 0.0f : dcol[i];
}

// the same three floats, straight from the blend backward's accumulator slots (available BEFORE the
// preprocess-backward kernel runs, so the all-gather can start underneath it; gsr_backward_early_rgb)
__global__ void __launch_bounds__(256)
clamped_rgb_from_acc_kernel(int P, const int32_t* __restrict__ radii, const uint8_t* __restrict__ clamped,
                            const float* __restrict__ acc, int stride, int rgb_off, float* __restrict__ drgb) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= 3 * P) return;
    const int g = i / 3, c = i - 3 * g;
    float v = 0.0f;
    if (radii[g] > 0 && !((clamped[g] >> c) & 1)) v = acc[(size_t)g * stride + rgb_off + c];
    drgb[i] = v;
}
cudaError_t launch_clamped_rgb_from_acc(int P, const int32_t* radii, const uint8_t* clamped, const float* acc, int stride,
                                        int rgb_off, float* drgb, cudaStream_t st) {
    if (P > 0) clamped_rgb_from_acc_kernel<<<(3 * P + 255) / 256, 256, 0, st>>>(P, radii, clamped, acc, stride, rgb_off, drgb);
    return cudaGetLastError();
}

// Peer variant: the three floats go straight into every rank's gather buffer through NVLink peer pointers (this
// rank's slot in each buffer, the local one included), i.e. the all-gather is done by the producing kernel's stores
// instead of a collective afterwards. Runs on a side stream ordered after the blend backward, so the stores travel
// underneath the preprocess-backward kernel of the main stream. Block 0 also publishes the camera position.
__global__ void __launch_bounds__(256)
clamped_rgb_push_kernel(int P, const int32_t* __restrict__ radii, const uint8_t* __restrict__ clamped,
                        const float* __restrict__ acc, int stride, int rgb_off, PeerDst d) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (blockIdx.x == 0 && threadIdx.x < 3) {
        const float c = d.campos_src[threadIdx.x];
        for (int k = 0; k < d.n; ++k) d.campos[k][threadIdx.x] = c;
    }
    if (i >= 3 * P) return;
    const int g = i / 3, c = i - 3 * g;
    float v = 0.0f;
    if (radii[g] > 0 && !((clamped[g] >> c) & 1)) v = acc[(size_t)g * stride + rgb_off + c];
#pragma unroll 4
    for (int k = 0; k < d.n; ++k) d.drgb[k][i] = v;
}
// Multicast variant: one multimem.st per 16 bytes, replicated into every rank's gather buffer by the switch -- the SMs
// issue 1/n of the stores and the link carries the 12 B x P once outbound instead of n times. The slot starts at an
// arbitrary float offset of the symmetric buffer (rank * 3P), so the floats before the first and after the last
// 16-byte-aligned group go out as scalar multimem.st.
__global__ void __launch_bounds__(256)
clamped_rgb_push_mc_kernel(int P, const int32_t* __restrict__ radii, const uint8_t* __restrict__ clamped,
                           const float* __restrict__ acc, int stride, int rgb_off, float* __restrict__ mc_drgb,
                           float* __restrict__ mc_campos, const float* __restrict__ campos_src) {
    if (blockIdx.x == 0 && threadIdx.x < 3) mc_st1(mc_campos + threadIdx.x, campos_src[threadIdx.x]);
    const int n = 3 * P;
    auto value = [&](int i) {
        const int g = i / 3, c = i - 3 * g;
        return (radii[g] > 0 && !((clamped[g] >> c) & 1)) ? acc[(size_t)g * stride + rgb_off + c] : 0.0f;
    };
    const int lead = (int)((16 - ((uintptr_t)mc_drgb & 15)) & 15) / 4;     // floats before the first aligned group (0..3)
    const int q = blockIdx.x * 256 + threadIdx.x;                          // aligned group index
    const int i0 = lead + 4 * q;
    if (q == 0) for (int i = 0; i < lead && i < n; ++i) mc_st1(mc_drgb + i, value(i));
    if (i0 + 3 < n) {
        mc_st4(reinterpret_cast<float4*>(mc_drgb + i0), make_float4(value(i0), value(i0 + 1), value(i0 + 2), value(i0 + 3)));
    } else {
        for (int i = i0; i < n; ++i) mc_st1(mc_drgb + i, value(i));        // the last, partial group
    }
}

cudaError_t launch_clamped_rgb_push(int P, const int32_t* radii, const uint8_t* clamped, const float* acc, int stride,
                                    int rgb_off, const PeerDst& dst, cudaStream_t main) {
    // one event per (host thread, device): an event belongs to the device that was current when it was created
    static thread_local cudaEvent_t evs[64] = {};
    cudaError_t e;
    int dev = 0;
    if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    cudaEvent_t& ev = evs[dev];
    if (!ev && (e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming)) != cudaSuccess) return e;
    if ((e = cudaEventRecord(ev, main)) != cudaSuccess) return e;
    if ((e = cudaStreamWaitEvent(dst.side, ev, 0)) != cudaSuccess) return e;
    if (P > 0 && dst.mc_drgb) {
        const int groups = (3 * P + 3) / 4 + 1;
        clamped_rgb_push_mc_kernel<<<(groups + 255) / 256, 256, 0, dst.side>>>(P, radii, clamped, acc, stride, rgb_off, dst.mc_drgb,
                                                                               dst.mc_campos, dst.campos_src);
    } else if (P > 0) {
        clamped_rgb_push_kernel<<<(3 * P + 255) / 256, 256, 0, dst.side>>>(P, radii, clamped, acc, stride, rgb_off, dst);
    }
    return cudaGetLastError();
}

// head row (common.cuh: PEER_HEAD_STRIDE floats) -> the planar arena head [means3D 3P | opacity P | scales 3P (2P) | rotations 4P]
__device__ __forceinline__ void head_row_to_planar(const float* __restrict__ src, float* __restrict__ dst, int P, int g, bool surfel) {
    const float4* row = reinterpret_cast<const float4*>(src + (size_t)g * PEER_HEAD_STRIDE);
    const float4 a = row[0], b = row[1], c = row[2];
    const size_t Pz = (size_t)P;
    dst[3 * (size_t)g] = a.x; dst[3 * (size_t)g + 1] = a.y; dst[3 * (size_t)g + 2] = a.z;
    dst[3 * Pz + g] = a.w;
    if (surfel) {       // scales 2, rotations 4
        float* sc = dst + 4 * Pz + 2 * (size_t)g; sc[0] = b.x; sc[1] = b.y;
        float* rq = dst + 6 * Pz + 4 * (size_t)g; rq[0] = b.z; rq[1] = b.w; rq[2] = c.x; rq[3] = c.y;
    } else {            // scales 3, rotations 4
        float* sc = dst + 4 * Pz + 3 * (size_t)g; sc[0] = b.x; sc[1] = b.y; sc[2] = b.z;
        float* rq = dst + 7 * Pz + 4 * (size_t)g; rq[0] = b.w; rq[1] = c.x; rq[2] = c.y; rq[3] = c.z;
    }
}
__global__ void __launch_bounds__(256)
peer_consume_kernel(const float* __restrict__ src, float* __restrict__ dst, int P, int row0, int rows, bool surfel) {
    const int g = blockIdx.x * 256 + threadIdx.x;
    if (g < rows) head_row_to_planar(src, dst, P, row0 + g, surfel);
}
cudaError_t launch_peer_consume(const float* src, float* dst, int P, int row0, int rows, bool surfel, cudaStream_t st) {
    peer_consume_kernel<<<(rows + 255) / 256, 256, 0, st>>>(src, dst, P, row0, rows, surfel);
    return cudaGetLastError();
}

// Phase 1 of the head reduce (common.cuh): this rank's share of the 16-byte groups, [q0, q1), summed over all replicas by
// the switch (multimem.ld_reduce) and stored back into all replicas (multimem.st). A ld_reduce is a round trip through
// the switch (microseconds), so the kernel is bound by bytes in flight: every thread issues U independent 16-byte
// reductions (consecutive lanes on consecutive groups: a warp instruction covers 512 contiguous bytes) before the
// first dependent store. Measured at 2 GPUs, 24 MB share: U = 1 0.150 ms.
template <int U>
__global__ void __launch_bounds__(256)
peer_reduce_shard_kernel(float4* __restrict__ mc, size_t q0, size_t q1) {
    const size_t step = (size_t)gridDim.x * 256 * U;
    for (size_t base = q0 + (size_t)blockIdx.x * 256 * U + threadIdx.x; base < q1; base += step) {
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) if (base + (size_t)u * 256 < q1) v[u] = mc_ld_reduce4(mc + base + (size_t)u * 256);
#pragma unroll
        for (int u = 0; u < U; ++u) if (base + (size_t)u * 256 < q1) mc_st4(mc + base + (size_t)u * 256, v[u]);
    }
}

// head != NULL: also hands over the switch-reduced head rows of the same Gaussians (phase 2 of the head reduce).
template <int M3>
__global__ void __launch_bounds__(SH_TILE)
sh_grad_from_views_kernel(int P, int D, int n_views, const float* __restrict__ means3D,
                          const float* __restrict__ campos, const float* __restrict__ drgb, float* __restrict__ dsh,
                          const float* __restrict__ head, float* __restrict__ head_dst, bool surfel) {
    constexpr int M = M3 / 3;
    __shared__ __align__(16) float s_sh[SH_TILE * ShTile<M3>::STRIDE];
    __shared__ uint8_t s_all[SH_TILE];
    const size_t row0 = (size_t)blockIdx.x * SH_TILE;
    const int nrows = min(SH_TILE, P - (int)row0);
    const int g = (int)row0 + threadIdx.x;
    float acc[M3];
#pragma unroll
    for (int i = 0; i < M3; ++i) acc[i] = 0.0f;
    if (g < P) {
        if (head) head_row_to_planar(head, head_dst, P, g, surfel);
        const V3 m = ld3(means3D + 3 * (size_t)g);
        // views in batches of four: the batch's colour gradients are loaded before any arithmetic, so four independent
        // loads are in flight per thread instead of one (the loop was latency-bound: 9 us per extra view at 1 M Gaussians)
        for (int v0 = 0; v0 < n_views; v0 += 4) {
            V3 dv[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int v = v0 + u;
                dv[u] = (v < n_views) ? ld3(drgb + ((size_t)v * P + g) * 3) : V3{0.0f, 0.0f, 0.0f};
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const V3 d = dv[u];
                if (d.x == 0.0f && d.y == 0.0f && d.z == 0.0f) continue;      // not visible in that view (or fully clamped)
                const int v = v0 + u;
                const V3 dir = { m.x - __ldg(campos + 3 * v), m.y - __ldg(campos + 3 * v + 1), m.z - __ldg(campos + 3 * v + 2) };
                const float inv_len = 1.0f / sqrtf(dot(dir, dir));
                float basis[16];
                const int used = sh_basis(D, dir.x * inv_len, dir.y * inv_len, dir.z * inv_len, basis);
#pragma unroll
                for (int k = 0; k < M; ++k) {
                    if (k < used) {
                        acc[3 * k] = fmaf(basis[k], d.x, acc[3 * k]);
                        acc[3 * k + 1] = fmaf(basis[k], d.y, acc[3 * k + 1]);
                        acc[3 * k + 2] = fmaf(basis[k], d.z, acc[3 * k + 2]);
                    }
                }
            }
        }
    }
    s_all[threadIdx.x] = 1;
    if constexpr (M3 == 48) {
        if ((reinterpret_cast<uintptr_t>(dsh) & 15) == 0) {      // whole rows as swizzled 16-byte chunks (sh.cuh, sh48_*)
            float4* s4 = reinterpret_cast<float4*>(s_sh);
            const int rot = threadIdx.x % 12;
#pragma unroll
            for (int q = 0; q < 12; ++q) {
                int p = q + rot;
                p -= (p >= 12) ? 12 : 0;
                s4[threadIdx.x * 12 + p] = make_float4(acc[4 * q], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3]);
            }
            __syncthreads();
            sh48_store(dsh, row0, nrows, s_all, s4);
            return;                                              // (CTA-uniform)
        }
    }
    float* mine = s_sh + threadIdx.x * ShTile<M3>::STRIDE;
#pragma unroll
    for (int i = 0; i < M3; ++i) mine[i] = acc[i];
    __syncthreads();
    sh_tile_store<M3>(dsh, row0, nrows, s_all, s_sh);
}

}  // namespace gsr

using namespace gsr;

extern "C" int gsr_clamped_color_grad(int P, int W, int H, int surfel, const void* geom_buffer, const float* dL_dcolors,
                                      float* drgb, void* stream) {
    if (P < 0 || W <= 0 || H <= 0) { set_error("bad sizes"); return GSR_EINVAL; }
    if (P == 0) return GSR_OK;
    if (!geom_buffer || !dL_dcolors || !drgb) { set_error("null pointer in gsr_clamped_color_grad"); return GSR_EINVAL; }
    const int tiles = ((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE);
    const uint8_t* clamped;
    if (surfel) { SurfelGeom g; carve_surfel_geom((char*)geom_buffer, P, tiles, &g); clamped = g.clamped; }
    else { GeomState g; carve_geom((char*)geom_buffer, P, tiles, &g); clamped = g.clamped; }
    cudaStream_t st = (cudaStream_t)stream;
    clamped_color_grad_kernel<<<(3 * P + 255) / 256, 256, 0, st>>>(P, clamped, dL_dcolors, drgb);
    return check_cuda(cudaGetLastError(), "clamped colour gradient launch");
}

static int sh_from_views_impl(int P, int M, int D, int n_views, const float* means3D, const float* campos,
                              const float* drgb, float* dL_dsh, const float* head, float* head_dst, bool surfel, void* stream) {
    if (P < 0 || M < 1 || M > 16 || D < 0 || D > 3 || (D + 1) * (D + 1) > M || n_views < 1) {
        set_error("bad sizes in gsr_sh_grad_from_views"); return GSR_EINVAL;
    }
    if (P == 0) return GSR_OK;
    if (!means3D || !campos || !drgb || !dL_dsh) { set_error("null pointer in gsr_sh_grad_from_views"); return GSR_EINVAL; }
    cudaStream_t st = (cudaStream_t)stream;
    const int blocks = (P + SH_TILE - 1) / SH_TILE;
    switch (3 * M) {
    case 3:  sh_grad_from_views_kernel<3><<<blocks, SH_TILE, 0, st>>>(P, D, n_views, means3D, campos, drgb, dL_dsh, head, head_dst, surfel); break;
    case 12: sh_grad_from_views_kernel<12><<<blocks, SH_TILE, 0, st>>>(P, D, n_views, means3D, campos, drgb, dL_dsh, head, head_dst, surfel); break;
    case 27: sh_grad_from_views_kernel<27><<<blocks, SH_TILE, 0, st>>>(P, D, n_views, means3D, campos, drgb, dL_dsh, head, head_dst, surfel); break;
    case 48: sh_grad_from_views_kernel<48><<<blocks, SH_TILE, 0, st>>>(P, D, n_views, means3D, campos, drgb, dL_dsh, head, head_dst, surfel); break;
    default: set_error("gsr_sh_grad_from_views: M must be 1, 4, 9 or 16"); return GSR_EINVAL;
    }
    return check_cuda(cudaGetLastError(), "SH gradient from views launch");
}

extern "C" int gsr_sh_grad_from_views(int P, int M, int D, int n_views, const float* means3D, const float* campos,
                                      const float* drgb, float* dL_dsh, void* stream) {
    return sh_from_views_impl(P, M, D, n_views, means3D, campos, drgb, dL_dsh, nullptr, nullptr, false, stream);
}

extern "C" int gsr_peer_finish(int P, int M, int D, int n_views, const float* means3D, const float* campos, const float* drgb,
                               float* dL_dsh, const float* head_rows, float* head_dst, int surfel, void* stream) {
    if (P > 0 && (!head_rows || !head_dst)) { set_error("gsr_peer_finish: null head"); return GSR_EINVAL; }
    return sh_from_views_impl(P, M, D, n_views, means3D, campos, drgb, dL_dsh, head_rows, head_dst, surfel != 0, stream);
}

extern "C" int gsr_peer_reduce_shard(float* multicast_head, int row0, int rows, int rank, int world, void* stream) {
    if (rows == 0) return GSR_OK;
    if (row0 < 0 || rows < 0 || !multicast_head || world < 1 || rank < 0 || rank >= world || ((uintptr_t)multicast_head & 15)) {
        set_error("gsr_peer_reduce_shard: bad argument"); return GSR_EINVAL;
    }
    // this rank's share of the 16-byte groups of rows [row0, row0 + rows)
    const size_t g0 = (size_t)row0 * (PEER_HEAD_STRIDE / 4), groups = (size_t)rows * (PEER_HEAD_STRIDE / 4);
    const size_t per = (groups + world - 1) / world;
    const size_t q0 = g0 + per * (size_t)rank, q1 = (per * (size_t)(rank + 1) < groups ? g0 + per * (size_t)(rank + 1) : g0 + groups);
    if (q0 >= q1) return GSR_OK;
    constexpr int U = 8;
    const size_t want = (q1 - q0 + 256 * U - 1) / (256 * U);
    // grid-stride, one CTA of 256 threads per SM, 8 groups per thread: 4.8 MB in flight (the link's bandwidth x the
    // round trip through the switch), and a footprint small enough to co-reside with the kernel it is pipelined with
    static int cap = -1;             // CTAs (tunable for the measurement in profiles/: GSR_REDUCE_BLOCKS)
    if (cap < 0) { const char* e = getenv("GSR_REDUCE_BLOCKS"); cap = e ? atoi(e) : 148; if (cap < 1) cap = 148; }
    const int blocks = (int)(want < (size_t)cap ? want : (size_t)cap);
    peer_reduce_shard_kernel<U><<<blocks, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<float4*>(multicast_head), q0, q1);
    return check_cuda(cudaGetLastError(), "peer head reduce launch");
}
