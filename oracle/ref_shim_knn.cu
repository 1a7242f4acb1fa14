// oracle/ref_shim_knn.cu -- TEST INFRASTRUCTURE, not product code.
// C-ABI wrapper around the UNMODIFIED reference simple-knn (SimpleKNN::knn,
// /root/reference/submodules/simple-knn/simple_knn.h:15-19), compiled in place.
#include <cuda_runtime.h>
#include "simple_knn.h"
extern "C" void ref_knn_dist2(int P, float* points_xyz, float* mean_dists)
{
    SimpleKNN::knn(P, (float3*)points_xyz, mean_dists);
}
