// densify.cu -- densify_and_prune (last item) as one plan and one gather.
//
// Replaces GaussianModel.densify_and_prune (reference gs3dgs/scene/gaussian_model.py:572-586; the 2DGS copy is
// gs2dgs/scene/gaussian_model.py:478-492) and what it calls: densify_and_clone (:559-570), densify_and_split
// (:502-527), densification_postfix (:482-500), cat_tensors_to_optimizer (:460-480), prune_points / _prune_optimizer
// (:425-458). The reference runs it as four rounds of boolean-mask indexing and torch.cat over the six parameter
// tensors and their twelve Adam moment tensors (clone -> append, split -> append, prune the split parents, prune by
// opacity / size): every round re-allocates and re-copies the whole model, and every mask indexing synchronises with
// the host. Here the four rounds collapse into a per-Gaussian decision and one compaction:
//
//   plan   one pass over (scaling, opacity, accum, denom): which originals survive, which are cloned, which are
//          split and whether their children survive the final prune; CTA counts; one small scan; one readback of the
//          four totals (the caller needs them to size the output tensors).
//   apply  the destination row of every survivor (CTA-local scan + CTA offsets), then ONE gather over all eighteen
//          arrays: read each surviving row once, write it once.
//
// Row order of the result is the reference's: surviving originals in index order, then surviving clones in index
// order, then the N copies of the surviving split children, copy-major (tensor.repeat(N, 1) tiles the whole block).
// Float sequences follow the torch kernels the reference calls, so every decision and every copied / recomputed value
// is bit-identical, except the child positions, where torch.bmm's accumulation order is cuBLAS's (tests allow 1 ulp
// scale differences there): grads = accum / denom with NaN -> 0; torch.norm over one element = sqrt(g * g);
// exp / sigmoid as in common.cuh; tensor / python_scalar is a multiplication by the float reciprocal
// (BinaryDivTrueKernel.cu); Python-float thresholds are rounded to float once on the host.
#include "common.cuh"

namespace gsr {

constexpr int DN_THREADS = 256;
constexpr int DN_ITEMS = 4;
constexpr int DN_TILE = DN_THREADS * DN_ITEMS;      // Gaussians per CTA
constexpr int DN_MAX_GROUPS = 8;

enum : uint8_t { DN_KEEP = 1, DN_CLONE = 2, DN_CHILD = 4, DN_SPLIT = 8 };

struct DensifyWs {           // carved from the caller's workspace
    uint8_t* cls;            // [P]
    int32_t* cta;            // [4][n_cta + 1]  per-CTA counts, then exclusive offsets (last = total)
    int32_t* src;            // [out_rows]      source Gaussian of every output row
    int32_t* zrow;           // [n_split * n_child] row of the normal sample of every child
    size_t bytes;
};
static inline size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }
static DensifyWs carve_densify(char* base, int P, int n_split) {
    DensifyWs w; size_t off = 0;
    const int n_cta = (P + DN_TILE - 1) / DN_TILE;
    w.cls = (uint8_t*)(base + off); off += al256((size_t)P);
    w.cta = (int32_t*)(base + off); off += al256((size_t)4 * (n_cta + 1) * 4);
    w.src = (int32_t*)(base + off); off += al256((size_t)(2 + n_split) * P * 4);
    w.zrow = (int32_t*)(base + off); off += al256((size_t)n_split * P * 4);
    w.bytes = off;
    return w;
}

struct DensifyCfgDev {
    float grad_thr, dense_ext, min_opacity, big_ext, max_screen; int prune_big; int n_split; float inv_split;
};

// the per-Gaussian decision of the whole densify_and_prune sequence
template <int SD>
__device__ __forceinline__ uint8_t classify_one(const float* __restrict__ scaling, const float* __restrict__ opacity,
                                                const float* __restrict__ accum, const float* __restrict__ denom,
                                                int i, const DensifyCfgDev& c) {
    float g = divr(__ldg(accum + i), __ldg(denom + i));            // :573-574
    if (isnan(g)) g = 0.0f;
    float maxs = -INFINITY, cmax = -INFINITY;
#pragma unroll
    for (int k = 0; k < SD; ++k) {
        const float s = act_scale(__ldg(scaling + (size_t)i * SD + k));
        maxs = fmaxf(maxs, s);
        // child scale: log(exp(s) / (0.8 N)), seen again through exp by the final prune (:516, :580-583)
        cmax = fmaxf(cmax, act_scale(logf(mulr(s, c.inv_split))));
    }
    const bool clone = (__fsqrt_rn(mulr(g, g)) >= c.grad_thr) && (maxs <= c.dense_ext);     // :561-563
    const bool split = (g >= c.grad_thr) && (maxs > c.dense_ext);                           // :507-509
    const bool low = act_opacity(__ldg(opacity + i)) < c.min_opacity;                         // :579
    // max_radii2D was just zeroed by densification_postfix (:498), so "max_radii2D > max_screen_size" is 0 > size
    const bool vs = c.prune_big && (0.0f > c.max_screen);
    const bool prune = low || vs || (c.prune_big && maxs > c.big_ext);
    const bool prune_child = low || vs || (c.prune_big && cmax > c.big_ext);
    uint8_t f = 0;
    if (!split && !prune) f |= DN_KEEP;
    if (clone && !prune) f |= DN_CLONE;
    if (split) f |= DN_SPLIT;
    if (split && !prune_child) f |= DN_CHILD;
    return f;
}

__device__ __forceinline__ int4 flags4(uint8_t f) {
    return make_int4((f & DN_KEEP) != 0, (f & DN_CLONE) != 0, (f & DN_CHILD) != 0, (f & DN_SPLIT) != 0);
}
__device__ __forceinline__ int4 add4(int4 a, int4 b) { return make_int4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ int4 shfl_up4(int4 v, int d) {
    return make_int4(__shfl_up_sync(0xffffffffu, v.x, d), __shfl_up_sync(0xffffffffu, v.y, d),
                     __shfl_up_sync(0xffffffffu, v.z, d), __shfl_up_sync(0xffffffffu, v.w, d));
}

// exclusive scan of one int4 per thread over the CTA; returns the thread's exclusive prefix, total in *total
__device__ __forceinline__ int4 cta_exscan4(int4 v, int4* total, int4* s_warp /*[DN_THREADS/32]*/) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int4 inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const int4 o = shfl_up4(inc, d); if (lane >= d) inc = add4(inc, o); }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    int4 base = make_int4(0, 0, 0, 0), tot = make_int4(0, 0, 0, 0);
#pragma unroll
    for (int w = 0; w < DN_THREADS / 32; ++w) { const int4 t = s_warp[w]; if (w < warp) base = add4(base, t); tot = add4(tot, t); }
    *total = tot;
    return make_int4(base.x + inc.x - v.x, base.y + inc.y - v.y, base.z + inc.z - v.z, base.w + inc.w - v.w);
}

template <int SD>
__global__ void __launch_bounds__(DN_THREADS)
densify_classify_kernel(int P, int n_cta, const float* __restrict__ scaling, const float* __restrict__ opacity,
                        const float* __restrict__ accum, const float* __restrict__ denom, DensifyCfgDev c,
                        uint8_t* __restrict__ cls, int32_t* __restrict__ cta) {
    __shared__ int4 s_warp[DN_THREADS / 32];
    const int i0 = blockIdx.x * DN_TILE + threadIdx.x * DN_ITEMS;
    int4 mine = make_int4(0, 0, 0, 0);
    uint32_t packed = 0;
#pragma unroll
    for (int j = 0; j < DN_ITEMS; ++j) {
        uint8_t f = 0;
        if (i0 + j < P) f = classify_one<SD>(scaling, opacity, accum, denom, i0 + j, c);
        packed |= (uint32_t)f << (8 * j);
        mine = add4(mine, flags4(f));
    }
    if (i0 + DN_ITEMS <= P) *reinterpret_cast<uint32_t*>(cls + i0) = packed;      // cls is 256-byte aligned, i0 % 4 == 0
    else for (int j = 0; j < DN_ITEMS && i0 + j < P; ++j) cls[i0 + j] = (uint8_t)(packed >> (8 * j));
    int4 tot;
    cta_exscan4(mine, &tot, s_warp);
    if (threadIdx.x == 0) {
        const int stride = n_cta + 1;
        cta[0 * stride + blockIdx.x] = tot.x; cta[1 * stride + blockIdx.x] = tot.y;
        cta[2 * stride + blockIdx.x] = tot.z; cta[3 * stride + blockIdx.x] = tot.w;
    }
}

// one CTA: the four rows of CTA counts become exclusive offsets, totals in the last column and in totals[4]
__global__ void __launch_bounds__(DN_THREADS)
densify_scan_kernel(int n_cta, int32_t* __restrict__ cta, int32_t* __restrict__ totals) {
    __shared__ int4 s_warp[DN_THREADS / 32];
    const int stride = n_cta + 1;
    int4 carry = make_int4(0, 0, 0, 0);
    for (int b0 = 0; b0 < n_cta; b0 += DN_THREADS) {
        const int b = b0 + threadIdx.x;
        int4 v = make_int4(0, 0, 0, 0);
        if (b < n_cta) v = make_int4(cta[b], cta[stride + b], cta[2 * stride + b], cta[3 * stride + b]);
        int4 tot;
        const int4 ex = add4(cta_exscan4(v, &tot, s_warp), carry);
        if (b < n_cta) { cta[b] = ex.x; cta[stride + b] = ex.y; cta[2 * stride + b] = ex.z; cta[3 * stride + b] = ex.w; }
        carry = add4(carry, tot);
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        cta[n_cta] = carry.x; cta[stride + n_cta] = carry.y; cta[2 * stride + n_cta] = carry.z; cta[3 * stride + n_cta] = carry.w;
        totals[0] = carry.x; totals[1] = carry.y; totals[2] = carry.z; totals[3] = carry.w;
    }
}

// destination rows: src[row] = source Gaussian, zrow[child row] = row of its normal sample
__global__ void __launch_bounds__(DN_THREADS)
densify_index_kernel(int P, int n_cta, int n_split, const uint8_t* __restrict__ cls, const int32_t* __restrict__ cta,
                     int32_t* __restrict__ src, int32_t* __restrict__ zrow) {
    __shared__ int4 s_warp[DN_THREADS / 32];
    const int stride = n_cta + 1;
    const int n_keep = cta[n_cta], n_clone = cta[stride + n_cta], n_child = cta[2 * stride + n_cta], n_sel = cta[3 * stride + n_cta];
    const int i0 = blockIdx.x * DN_TILE + threadIdx.x * DN_ITEMS;
    uint8_t f[DN_ITEMS];
    int4 mine = make_int4(0, 0, 0, 0);
#pragma unroll
    for (int j = 0; j < DN_ITEMS; ++j) { f[j] = (i0 + j < P) ? cls[i0 + j] : 0; mine = add4(mine, flags4(f[j])); }
    int4 tot;
    int4 at = cta_exscan4(mine, &tot, s_warp);
    at = add4(at, make_int4(cta[blockIdx.x], cta[stride + blockIdx.x], cta[2 * stride + blockIdx.x], cta[3 * stride + blockIdx.x]));
#pragma unroll
    for (int j = 0; j < DN_ITEMS; ++j) {
        const int i = i0 + j;
        if (f[j] & DN_KEEP) src[at.x++] = i;
        if (f[j] & DN_CLONE) src[n_keep + at.y++] = i;
        if (f[j] & DN_CHILD) {
            for (int k = 0; k < n_split; ++k) {
                src[n_keep + n_clone + k * n_child + at.z] = i;
                zrow[k * n_child + at.z] = k * n_sel + at.w;       // the reference draws N * n_sel samples, copy-major
            }
            ++at.z;
        }
        if (f[j] & DN_SPLIT) ++at.w;
    }
}

struct DensifyGroupDev {
    const float* param; const float* m; const float* v;
    float* out_param; float* out_m; float* out_v;
    int width; int kind;           // GSR_DENSIFY_COPY / _XYZ / _SCALING
    long long first_block;         // prefix sum of blocks over the groups
};
struct DensifyGatherArgs {
    DensifyGroupDev g[DN_MAX_GROUPS];
    const int32_t* src; const int32_t* zrow;
    const float* rotation; const float* scaling; const float* normals;
    int n_keep, n_clone, out_rows, scale_dim, n_groups;
    float inv_split;
};

// row k of build_rotation(q) (reference gs3dgs/utils/general_utils.py:93-115), one torch op per rounding
__device__ __forceinline__ float3 rotation_row(const float* __restrict__ rot, int k) {
    const float a = __ldg(rot), b = __ldg(rot + 1), c = __ldg(rot + 2), d = __ldg(rot + 3);
    const float n = __fsqrt_rn(addr(addr(addr(mulr(a, a), mulr(b, b)), mulr(c, c)), mulr(d, d)));
    const float r = divr(a, n), x = divr(b, n), y = divr(c, n), z = divr(d, n);
    if (k == 0) return make_float3(subr(1.0f, mulr(2.0f, addr(mulr(y, y), mulr(z, z)))), mulr(2.0f, subr(mulr(x, y), mulr(r, z))),
                                   mulr(2.0f, addr(mulr(x, z), mulr(r, y))));
    if (k == 1) return make_float3(mulr(2.0f, addr(mulr(x, y), mulr(r, z))), subr(1.0f, mulr(2.0f, addr(mulr(x, x), mulr(z, z)))),
                                   mulr(2.0f, subr(mulr(y, z), mulr(r, x))));
    return make_float3(mulr(2.0f, subr(mulr(x, z), mulr(r, y))), mulr(2.0f, addr(mulr(y, z), mulr(r, x))),
                       subr(1.0f, mulr(2.0f, addr(mulr(x, x), mulr(y, y)))));
}

// the recomputed values of a split child (xyz and scaling groups only)
struct ChildSrc { const int32_t* zrow; const float* rotation; const float* scaling; const float* normals; int scale_dim; float inv_split; };
__device__ __noinline__ float child_value(ChildSrc a, int kind, float val, int s, int k, int child_row) {
    if (kind == GSR_DENSIFY_SCALING) return logf(mulr(act_scale(val), a.inv_split));          // :516
    const int zr = __ldg(a.zrow + child_row);
    float smp[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {                            // torch.normal(mean = 0, std): z * std + 0  (:511-513)
        const float sd = j < a.scale_dim ? act_scale(__ldg(a.scaling + (size_t)s * a.scale_dim + j)) : 0.0f;
        smp[j] = addr(mulr(__ldg(a.normals + (size_t)zr * 3 + j), sd), 0.0f);
    }
    const float3 R = rotation_row(a.rotation + (size_t)s * 4, k);
    return addr(fmar(R.z, smp[2], fmar(R.y, smp[1], mulr(R.x, smp[0]))), val);                // bmm row + xyz (:515)
}

// one launch over all groups (each owns a contiguous range of blocks); DN_GATHER floats of one output tensor per
// thread, all loads issued before the stores. Element counts fit 32 bits (checked by the host).
constexpr int DN_GATHER = 4;
__global__ void __launch_bounds__(256)
densify_gather_kernel(DensifyGatherArgs a) {
    int gi = 0;
#pragma unroll
    for (int k = 1; k < DN_MAX_GROUPS; ++k)
        if (k < a.n_groups && (long long)blockIdx.x >= a.g[k].first_block) gi = k;
    const DensifyGroupDev& G = a.g[gi];
    const uint32_t w = (uint32_t)G.width;
    const uint32_t total = (uint32_t)a.out_rows * w;
    const uint32_t e0 = (uint32_t)(blockIdx.x - G.first_block) * (256u * DN_GATHER) + threadIdx.x;
    const bool has_m = G.out_m != nullptr && G.m != nullptr, has_v = G.out_v != nullptr && G.v != nullptr;
    uint32_t r[DN_GATHER], k[DN_GATHER]; int s[DN_GATHER];
#pragma unroll
    for (int j = 0; j < DN_GATHER; ++j) {
        const uint32_t e = e0 + 256u * j;
        r[j] = e / w; k[j] = e - r[j] * w;
        s[j] = e < total ? __ldg(a.src + r[j]) : 0;
    }
    float p[DN_GATHER], m[DN_GATHER], v[DN_GATHER];
#pragma unroll
    for (int j = 0; j < DN_GATHER; ++j) {
        const bool in = e0 + 256u * j < total;
        const bool kept = in && r[j] < (uint32_t)a.n_keep;   // surviving original: parameter and both moments move
        const size_t from = (size_t)s[j] * w + k[j];
        p[j] = in ? __ldg(G.param + from) : 0.0f;
        m[j] = (kept && has_m) ? __ldg(G.m + from) : 0.0f;   // new rows: zeros_like(extension), cat_tensors_to_optimizer
        v[j] = (kept && has_v) ? __ldg(G.v + from) : 0.0f;
    }
#pragma unroll
    for (int j = 0; j < DN_GATHER; ++j) {
        const uint32_t e = e0 + 256u * j;
        if (e >= total) continue;
        if (G.kind != GSR_DENSIFY_COPY && r[j] >= (uint32_t)(a.n_keep + a.n_clone))
            p[j] = child_value(ChildSrc{ a.zrow, a.rotation, a.scaling, a.normals, a.scale_dim, a.inv_split }, G.kind, p[j],
                               s[j], (int)k[j], (int)r[j] - a.n_keep - a.n_clone);
        G.out_param[e] = p[j];
        if (G.out_m) G.out_m[e] = m[j];
        if (G.out_v) G.out_v[e] = v[j];
    }
}

}  // namespace gsr

using namespace gsr;

static bool cfg_to_dev(const GsrDensifyCfg* c, DensifyCfgDev* d) {
    if (!c || c->n_split < 1 || c->n_split > 8) return false;
    d->grad_thr = c->grad_threshold; d->dense_ext = c->dense_extent; d->min_opacity = c->min_opacity;
    d->big_ext = c->big_extent; d->max_screen = c->max_screen_size; d->prune_big = c->prune_big; d->n_split = c->n_split;
    // tensor / (0.8 * N): the divisor is a Python float, cast to float, and the CUDA kernel multiplies by its reciprocal
    d->inv_split = 1.0f / (float)(0.8 * (double)c->n_split);
    return true;
}

extern "C" {

size_t gsr_densify_workspace_bytes(int P, int n_split)
{
    if (P <= 0 || n_split < 1) return 0;
    return carve_densify(nullptr, P, n_split).bytes;
}

int gsr_densify_plan(int P, int scale_dim, const float* scaling, const float* opacity, const float* xyz_gradient_accum,
                     const float* denom, const GsrDensifyCfg* cfg, void* workspace, int32_t counts[4], void* stream)
{
    if (!counts) { set_error("gsr_densify_plan: null counts"); return GSR_EINVAL; }
    counts[0] = counts[1] = counts[2] = counts[3] = 0;
    if (P == 0) return GSR_OK;
    DensifyCfgDev c;
    if (P < 0 || !scaling || !opacity || !xyz_gradient_accum || !denom || !workspace || !cfg_to_dev(cfg, &c)) {
        set_error("bad argument to gsr_densify_plan"); return GSR_EINVAL;
    }
    if (scale_dim != 2 && scale_dim != 3) { set_error("gsr_densify_plan: scale_dim %d (2 or 3)", scale_dim); return GSR_EUNSUPPORTED; }
    if ((long long)(2 + c.n_split) * P > 0x7fffffffLL) { set_error("gsr_densify_plan: too many rows"); return GSR_EUNSUPPORTED; }
    const cudaStream_t st = (cudaStream_t)stream;
    const DensifyWs w = carve_densify((char*)workspace, P, c.n_split);
    const int n_cta = (P + DN_TILE - 1) / DN_TILE;
    if (scale_dim == 3) densify_classify_kernel<3><<<n_cta, DN_THREADS, 0, st>>>(P, n_cta, scaling, opacity, xyz_gradient_accum, denom, c, w.cls, w.cta);
    else                densify_classify_kernel<2><<<n_cta, DN_THREADS, 0, st>>>(P, n_cta, scaling, opacity, xyz_gradient_accum, denom, c, w.cls, w.cta);
    int32_t* totals = w.zrow;                   // four ints of a buffer the index kernel only writes later
    densify_scan_kernel<<<1, DN_THREADS, 0, st>>>(n_cta, w.cta, totals);
    int rc;
    if ((rc = check_cuda(cudaGetLastError(), "densify plan launch"))) return rc;
    int32_t stack[4];
    int32_t* host = readback_pad() ? (int32_t*)readback_pad() : stack;
    if ((rc = check_cuda(cudaMemcpyAsync(host, totals, 16, cudaMemcpyDeviceToHost, st), "copy densify counts"))) return rc;
    if ((rc = check_cuda(cudaStreamSynchronize(st), "sync densify counts"))) return rc;
    for (int k = 0; k < 4; ++k) counts[k] = host[k];
    return GSR_OK;
}

int gsr_densify_apply(int P, int scale_dim, const GsrDensifyCfg* cfg, void* workspace, const int32_t counts[4],
                      int n_groups, const GsrDensifyGroup* groups, const float* rotation, const float* scaling,
                      const float* normals, void* stream)
{
    DensifyCfgDev c;
    if (P < 0 || !counts || !cfg_to_dev(cfg, &c) || n_groups < 0 || n_groups > DN_MAX_GROUPS || (n_groups && !groups)) {
        set_error("bad argument to gsr_densify_apply"); return GSR_EINVAL;
    }
    const int n_keep = counts[0], n_clone = counts[1], n_child = counts[2], n_sel = counts[3];
    if (n_keep < 0 || n_clone < 0 || n_child < 0 || n_sel < n_child || n_keep > P || n_clone > P || n_sel > P) {
        set_error("gsr_densify_apply: counts are not from gsr_densify_plan"); return GSR_EINVAL;
    }
    const long long out_rows = (long long)n_keep + n_clone + (long long)c.n_split * n_child;
    if (P == 0 || out_rows == 0 || n_groups == 0) return GSR_OK;
    if (!workspace) { set_error("gsr_densify_apply: null workspace"); return GSR_EINVAL; }
    if (scale_dim != 2 && scale_dim != 3) { set_error("gsr_densify_apply: scale_dim %d (2 or 3)", scale_dim); return GSR_EUNSUPPORTED; }
    const cudaStream_t st = (cudaStream_t)stream;
    const DensifyWs w = carve_densify((char*)workspace, P, c.n_split);
    const int n_cta = (P + DN_TILE - 1) / DN_TILE;

    DensifyGatherArgs a;
    long long blocks = 0; bool need_xyz = false;
    for (int k = 0; k < n_groups; ++k) {
        const GsrDensifyGroup& s = groups[k];
        if (!s.param || !s.out_param || s.width < 1) { set_error("gsr_densify_apply: group %d has a null tensor or width < 1", k); return GSR_EINVAL; }
        if (s.kind < GSR_DENSIFY_COPY || s.kind > GSR_DENSIFY_SCALING) { set_error("gsr_densify_apply: group %d kind %d", k, s.kind); return GSR_EINVAL; }
        if (s.kind == GSR_DENSIFY_XYZ && s.width != 3) { set_error("gsr_densify_apply: the xyz group must be 3 wide"); return GSR_EINVAL; }
        a.g[k] = { s.param, s.exp_avg, s.exp_avg_sq, s.out_param, s.out_exp_avg, s.out_exp_avg_sq, s.width, s.kind, blocks };
        if (out_rows * s.width > 0xffffffffLL) { set_error("gsr_densify_apply: group %d has more than 2^32 elements", k); return GSR_EUNSUPPORTED; }
        blocks += (out_rows * s.width + 256 * DN_GATHER - 1) / (256 * DN_GATHER);
        need_xyz |= s.kind == GSR_DENSIFY_XYZ;
    }
    if (need_xyz && n_child > 0 && (!rotation || !scaling || !normals)) {
        set_error("gsr_densify_apply: rotation, scaling and normals are needed to place split children"); return GSR_EINVAL;
    }
    a.src = w.src; a.zrow = w.zrow; a.rotation = rotation; a.scaling = scaling; a.normals = normals;
    a.n_keep = n_keep; a.n_clone = n_clone; a.out_rows = (int)out_rows; a.scale_dim = scale_dim; a.n_groups = n_groups;
    a.inv_split = c.inv_split;

    densify_index_kernel<<<n_cta, DN_THREADS, 0, st>>>(P, n_cta, c.n_split, w.cls, w.cta, w.src, w.zrow);
    if (blocks > 0x7fffffffLL) { set_error("gsr_densify_apply: too many elements"); return GSR_EUNSUPPORTED; }
    densify_gather_kernel<<<(unsigned)blocks, 256, 0, st>>>(a);
    return check_cuda(cudaGetLastError(), "densify apply launch");
}

}  // extern "C"
