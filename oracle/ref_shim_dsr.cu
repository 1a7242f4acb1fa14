// oracle/ref_shim_dsr.cu -- TEST INFRASTRUCTURE, not product code.
//
// C-ABI wrapper (written for this repo) around the UNMODIFIED reference 2DGS surfel
// rasterizer library, compiled from
// /root/reference/submodules/diff-surfel-rasterization/cuda_rasterizer/ in place
// (reference rasterizer.h:19-83, scratch layout rasterizer_impl.h:21-72).
// Nothing here is linked into, imported by or called from scorp_b200/.
#include <cstddef>
#include <cstdint>
#include <functional>
#include <stdexcept>
#include <cuda_runtime.h>
#include "cuda_rasterizer/rasterizer.h"
#include "cuda_rasterizer/rasterizer_impl.h"

extern "C" {

typedef char* (*ref_alloc_fn)(void* ctx, int which, size_t bytes);

int ref_dsr_forward(ref_alloc_fn alloc, void* ctx,
                    int P, int D, int M, const float* background, int width, int height,
                    const float* means3D, const float* shs, const float* colors_precomp,
                    const float* opacities, const float* scales, float scale_modifier,
                    const float* rotations, const float* transMat_precomp,
                    const float* viewmatrix, const float* projmatrix, const float* cam_pos,
                    float tan_fovx, float tan_fovy, int prefiltered,
                    float* out_color, float* out_others, int* radii, int debug)
{
    try {
        return CudaRasterizer::Rasterizer::forward(
            [=](size_t n) { return alloc(ctx, 0, n); },
            [=](size_t n) { return alloc(ctx, 1, n); },
            [=](size_t n) { return alloc(ctx, 2, n); },
            P, D, M, background, width, height, means3D, shs, colors_precomp, opacities,
            scales, scale_modifier, rotations, transMat_precomp, viewmatrix, projmatrix, cam_pos,
            tan_fovx, tan_fovy, prefiltered != 0, out_color, out_others, radii, debug != 0);
    } catch (const std::exception&) {
        return -1;
    }
}

int ref_dsr_backward(int P, int D, int M, int R, const float* background, int width, int height,
                     const float* means3D, const float* shs, const float* colors_precomp,
                     const float* scales, float scale_modifier, const float* rotations,
                     const float* transMat_precomp, const float* viewmatrix,
                     const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy,
                     const int* radii, char* geom_buffer, char* binning_buffer, char* image_buffer,
                     const float* dL_dpix, const float* dL_depths,
                     float* dL_dmean2D, float* dL_dnormal, float* dL_dopacity, float* dL_dcolor,
                     float* dL_dmean3D, float* dL_dtransMat, float* dL_dsh, float* dL_dscale,
                     float* dL_drot, int debug)
{
    try {
        CudaRasterizer::Rasterizer::backward(
            P, D, M, R, background, width, height, means3D, shs, colors_precomp, scales,
            scale_modifier, rotations, transMat_precomp, viewmatrix, projmatrix, campos, tan_fovx,
            tan_fovy, radii, geom_buffer, binning_buffer, image_buffer, dL_dpix, dL_depths,
            dL_dmean2D, dL_dnormal, dL_dopacity, dL_dcolor, dL_dmean3D, dL_dtransMat, dL_dsh,
            dL_dscale, dL_drot, debug != 0);
        return 0;
    } catch (const std::exception&) {
        return -1;
    }
}

// geom[10]: depths, clamped, internal_radii, means2D, transMat, normal_opacity, rgb,
//           tiles_touched, scanning_space, point_offsets   (rasterizer_impl.cu:155-170)
void ref_dsr_geom_layout(size_t P, size_t* off, size_t* total)
{
    char* chunk = nullptr;
    auto g = CudaRasterizer::GeometryState::fromChunk(chunk, P);
    off[0] = (size_t)g.depths; off[1] = (size_t)g.clamped; off[2] = (size_t)g.internal_radii;
    off[3] = (size_t)g.means2D; off[4] = (size_t)g.transMat; off[5] = (size_t)g.normal_opacity;
    off[6] = (size_t)g.rgb; off[7] = (size_t)g.tiles_touched; off[8] = (size_t)g.scanning_space;
    off[9] = (size_t)g.point_offsets;
    *total = (size_t)chunk;
}
// img[3]: accum_alpha (3 planes), n_contrib (2 planes), ranges
void ref_dsr_image_layout(size_t N, size_t* off, size_t* total)
{
    char* chunk = nullptr;
    auto s = CudaRasterizer::ImageState::fromChunk(chunk, N);
    off[0] = (size_t)s.accum_alpha; off[1] = (size_t)s.n_contrib; off[2] = (size_t)s.ranges;
    *total = (size_t)chunk;
}
void ref_dsr_binning_layout(size_t N, size_t* off, size_t* total)
{
    char* chunk = nullptr;
    auto b = CudaRasterizer::BinningState::fromChunk(chunk, N);
    off[0] = (size_t)b.point_list; off[1] = (size_t)b.point_list_unsorted;
    off[2] = (size_t)b.point_list_keys; off[3] = (size_t)b.point_list_keys_unsorted;
    off[4] = (size_t)b.list_sorting_space;
    *total = (size_t)chunk;
}

}  // extern "C"
