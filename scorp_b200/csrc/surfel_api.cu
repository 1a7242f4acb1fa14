// surfel_api.cu -- extern "C" entry points of the 2DGS surfel path and simple_knn.
// PLACEHOLDER until the surfel/knn kernels land: every call fails loudly (no fallback).
#include "common.cuh"

using namespace gsr;

extern "C" {

int gsr_surfel_forward(const GsrSettings*, int, int, const float*, const float*, const float*, const float*,
                       const float*, const float*, const float*, float*, float*, int32_t*, gsr_alloc_fn, void*,
                       int*, void*) {
    set_error("gsr_surfel_forward: not built in this revision");
    return GSR_EUNSUPPORTED;
}
size_t gsr_surfel_backward_workspace_bytes(int P) { return align_up((size_t)(P > 0 ? P : 1) * 20 * sizeof(float)); }
int gsr_surfel_backward(const GsrSettings*, int, int, int, const float*, const float*, const float*, const float*,
                        const float*, const float*, const int32_t*, const void*, const void*, const void*,
                        const float*, const float*, float*, float*, float*, float*, float*, float*, float*, float*,
                        void*, size_t, void*) {
    set_error("gsr_surfel_backward: not built in this revision");
    return GSR_EUNSUPPORTED;
}
void gsr_surfel_geom_layout(int, int, size_t off[6], size_t* total) { for (int i = 0; i < 6; ++i) off[i] = 0; *total = 0; }
void gsr_surfel_image_layout(int, int, size_t off[4], size_t* total) { for (int i = 0; i < 4; ++i) off[i] = 0; *total = 0; }

size_t gsr_knn_workspace_bytes(int P) { return align_up((size_t)(P > 0 ? P : 1) * 32); }
int gsr_knn_dist2(int, const float*, float*, void*, size_t, void*) {
    set_error("gsr_knn_dist2: not built in this revision");
    return GSR_EUNSUPPORTED;
}

}  // extern "C"
