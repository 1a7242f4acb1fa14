// optim.cu -- the two per-iteration steps that follow the rasterizer backward in the training loops
// (SURVEY.md 8(f) rank 4): the Adam update of the six parameter groups and the densification statistics.
//
// Replaces
//   gaussians.optimizer.step()  -- torch.optim.Adam(l, lr=0.0, eps=1e-15) over xyz, f_dc, f_rest, opacity, scaling,
//     rotation (reference gs3dgs/scene/gaussian_model.py:192-206, train_3dgs.py:192). torch's default CUDA path is
//     the multi-tensor "foreach" implementation: lerp_, mul_, addcmul_, sqrt, div_, add_, addcdiv_ -- seven passes,
//     72 bytes of traffic per parameter float. adam_kernel does the same arithmetic in ONE pass (read param, grad,
//     both moments; write param and both moments: 28 bytes per float) over all groups in one launch.
//   gaussians.max_radii2D[vis] = max(...); gaussians.add_densification_stats(viewspace_point_tensor, vis)
//     (train_3dgs.py:180-181, gaussian_model.py:603-605): boolean-mask indexing, i.e. nonzero() with a host
//     synchronisation, gathers, a norm and three scatters; densify_stats_kernel is one pass over the Gaussians.
//
// The float sequences are those of torch's functors, so the result is bit-identical to torch.optim.Adam (checked on
// the GPU by tests/test_gpu_optim.py): every scalar (1 - beta1, beta2, 1 - beta2, sqrt(bias_correction2), eps,
// -lr / bias_correction1) is formed in double on the host, as Python does, and rounded to float once.
#include "common.cuh"
#include <cmath>

namespace gsr {

constexpr int ADAM_MAX_GROUPS = 8;
constexpr int ADAM_THREADS = 256;
constexpr int ADAM_VEC = 4;                    // floats per thread per chunk step

struct AdamGroupDev {
    float* param; const float* grad; float* exp_avg; float* exp_avg_sq;
    long long n;               // floats
    long long first_chunk;     // prefix sum of chunks (one chunk = ADAM_THREADS * ADAM_VEC floats)
    float neg_step_size;       // -(lr / bias_correction1)
    int vec_ok;                // all four pointers 16-byte aligned
};
struct AdamArgs {
    AdamGroupDev g[ADAM_MAX_GROUPS];
    int n_groups;
    float w1;                  // 1 - beta1            (lerp weight)
    float beta2, w2;           // beta2, 1 - beta2
    float bc2_sqrt;            // sqrt(1 - beta2^step)
    float eps;
};

// one element, torch's functor sequence:
//   exp_avg.lerp_(grad, w1)                    -> fma(w1, grad - m, m)                  (|w1| < 0.5 branch of lerp)
//   exp_avg_sq.mul_(beta2).addcmul_(g, g, w2)  -> fma(w2, g * g, v * beta2)              (DeviceAddCmulCdiv.cuh: explicit fma)
//   denom = (sqrt(v) / bc2_sqrt) + eps ;  param.addcdiv_(m, denom, value = neg_step) -> fma(neg_step, m / denom, p)
__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const AdamArgs& a, float neg_step) {
    m = (a.w1 < 0.5f) ? fmar(a.w1, subr(g, m), m) : subr(g, mulr(subr(g, m), subr(1.0f, a.w1)));
    v = fmar(a.w2, mulr(g, g), mulr(v, a.beta2));
    const float denom = addr(divr(__fsqrt_rn(v), a.bc2_sqrt), a.eps);
    p = fmar(neg_step, divr(m, denom), p);
}

__global__ void __launch_bounds__(ADAM_THREADS)
adam_kernel(AdamArgs a) {
    // which group does this CTA's chunk belong to (<= 8 groups: linear search over the prefix sums)
    const long long chunk = blockIdx.x;
    int gi = 0;
#pragma unroll
    for (int k = 1; k < ADAM_MAX_GROUPS; ++k)
        if (k < a.n_groups && chunk >= a.g[k].first_chunk) gi = k;
    const AdamGroupDev& G = a.g[gi];
    const long long base = (chunk - G.first_chunk) * (ADAM_THREADS * ADAM_VEC) + (long long)threadIdx.x * ADAM_VEC;
    if (base >= G.n) return;
    const float ns = G.neg_step_size;
    if (G.vec_ok && base + ADAM_VEC <= G.n) {
        float4 p = *reinterpret_cast<const float4*>(G.param + base);
        const float4 g = __ldg(reinterpret_cast<const float4*>(G.grad + base));
        float4 m = *reinterpret_cast<const float4*>(G.exp_avg + base);
        float4 v = *reinterpret_cast<const float4*>(G.exp_avg_sq + base);
        adam_one(p.x, g.x, m.x, v.x, a, ns); adam_one(p.y, g.y, m.y, v.y, a, ns);
        adam_one(p.z, g.z, m.z, v.z, a, ns); adam_one(p.w, g.w, m.w, v.w, a, ns);
        *reinterpret_cast<float4*>(G.param + base) = p;
        *reinterpret_cast<float4*>(G.exp_avg + base) = m;
        *reinterpret_cast<float4*>(G.exp_avg_sq + base) = v;
    } else {
        for (int e = 0; e < ADAM_VEC && base + e < G.n; ++e) {
            float p = G.param[base + e], m = G.exp_avg[base + e], v = G.exp_avg_sq[base + e];
            adam_one(p, __ldg(G.grad + base + e), m, v, a, ns);
            G.param[base + e] = p; G.exp_avg[base + e] = m; G.exp_avg_sq[base + e] = v;
        }
    }
}

// train_3dgs.py:180-181 + gaussian_model.py:603-605 for the Gaussians with radii > 0:
//   max_radii2D = max(max_radii2D, float(radii));  xyz_gradient_accum += sqrt(gx^2 + gy^2);  denom += 1
// (torch.norm over two elements: both squares rounded, one add, IEEE sqrt).
__global__ void __launch_bounds__(256)
densify_stats_kernel(int P, const int32_t* __restrict__ radii, const float* __restrict__ grad2d, int grad_stride,
                     float* __restrict__ accum, float* __restrict__ denom, float* __restrict__ max_radii) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= P) return;
    const int r = __ldg(radii + i);
    if (r <= 0) return;
    const float gx = __ldg(grad2d + (size_t)i * grad_stride), gy = __ldg(grad2d + (size_t)i * grad_stride + 1);
    accum[i] = addr(accum[i], __fsqrt_rn(addr(mulr(gx, gx), mulr(gy, gy))));
    denom[i] = addr(denom[i], 1.0f);
    if (max_radii) max_radii[i] = fmaxf(max_radii[i], (float)r);
}

// GaussianModel.reset_opacity (gaussian_model.py:253-256) + replace_tensor_to_optimizer (:410-423):
//   opacity = inverse_sigmoid(min(sigmoid(opacity), 0.01)) = log(x / (1 - x));  exp_avg = exp_avg_sq = 0
// (torch: sigmoid, ones_like * 0.01, min, rsub, div, log -- one rounding each).
__global__ void __launch_bounds__(256)
reset_opacity_kernel(int n, float* __restrict__ opacity, float* __restrict__ exp_avg, float* __restrict__ exp_avg_sq, float cap) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const float a = act_opacity(opacity[i]);
    const float x = (a != a) ? a : fminf(a, cap);                     // torch.min propagates NaN
    opacity[i] = logf(divr(x, subr(1.0f, x)));
    if (exp_avg) exp_avg[i] = 0.0f;
    if (exp_avg_sq) exp_avg_sq[i] = 0.0f;
}

}  // namespace gsr

using namespace gsr;

extern "C" {

int gsr_adam_step(int n_groups, const GsrAdamGroup* groups, double beta1, double beta2, double eps, int step, void* stream)
{
    if (n_groups < 0 || n_groups > ADAM_MAX_GROUPS) { set_error("gsr_adam_step: %d groups (max %d)", n_groups, ADAM_MAX_GROUPS); return GSR_EINVAL; }
    if (n_groups == 0) return GSR_OK;
    if (!groups || step < 1) { set_error("gsr_adam_step: null groups or step < 1"); return GSR_EINVAL; }
    AdamArgs a;
    a.n_groups = 0;
    // scalars as torch.optim.adam._multi_tensor_adam forms them (Python floats = doubles), rounded to float once
    const double bc1 = 1.0 - std::pow(beta1, (double)step);
    const double bc2 = 1.0 - std::pow(beta2, (double)step);
    a.w1 = (float)(1.0 - beta1); a.beta2 = (float)beta2; a.w2 = (float)(1.0 - beta2);
    a.bc2_sqrt = (float)std::sqrt(bc2); a.eps = (float)eps;
    long long chunks = 0;
    for (int k = 0; k < n_groups; ++k) {
        const GsrAdamGroup& s = groups[k];
        if (s.n < 0) { set_error("gsr_adam_step: negative size in group %d", k); return GSR_EINVAL; }
        if (s.n == 0) continue;
        if (!s.param || !s.grad || !s.exp_avg || !s.exp_avg_sq) { set_error("gsr_adam_step: null pointer in group %d", k); return GSR_EINVAL; }
        AdamGroupDev& d = a.g[a.n_groups++];
        d.param = s.param; d.grad = s.grad; d.exp_avg = s.exp_avg; d.exp_avg_sq = s.exp_avg_sq; d.n = s.n;
        d.first_chunk = chunks;
        d.neg_step_size = (float)((s.lr / bc1) * -1.0);
        d.vec_ok = (((uintptr_t)s.param | (uintptr_t)s.grad | (uintptr_t)s.exp_avg | (uintptr_t)s.exp_avg_sq) & 15) == 0;
        chunks += (s.n + ADAM_THREADS * ADAM_VEC - 1) / (ADAM_THREADS * ADAM_VEC);
    }
    if (chunks == 0) return GSR_OK;
    if (chunks > 0x7fffffffLL) { set_error("gsr_adam_step: too many elements"); return GSR_EUNSUPPORTED; }
    adam_kernel<<<(unsigned)chunks, ADAM_THREADS, 0, (cudaStream_t)stream>>>(a);
    return check_cuda(cudaGetLastError(), "adam launch");
}

int gsr_reset_opacity(int P, float* opacity, float* exp_avg, float* exp_avg_sq, float cap, void* stream)
{
    if (P == 0) return GSR_OK;
    if (P < 0 || !opacity) { set_error("bad argument to gsr_reset_opacity"); return GSR_EINVAL; }
    reset_opacity_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream>>>(P, opacity, exp_avg, exp_avg_sq, cap);
    return check_cuda(cudaGetLastError(), "reset opacity launch");
}

int gsr_densification_stats(int P, const int32_t* radii, const float* dL_dmeans2D, int grad_stride,
                            float* xyz_gradient_accum, float* denom, float* max_radii2D, void* stream)
{
    if (P == 0) return GSR_OK;
    if (P < 0 || !radii || !dL_dmeans2D || !xyz_gradient_accum || !denom || grad_stride < 2) {
        set_error("bad argument to gsr_densification_stats"); return GSR_EINVAL;
    }
    densify_stats_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream>>>(P, radii, dL_dmeans2D, grad_stride,
                                                                            xyz_gradient_accum, denom, max_radii2D);
    return check_cuda(cudaGetLastError(), "densification stats launch");
}

}  // extern "C"
