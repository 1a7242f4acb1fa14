// ply.cu -- the model's on-disk format (last item): PLY vertex rows packed / unpacked on the GPU.
//
// Replaces the tensor shuffling of GaussianModel.save_ply / load_ply (reference gs3dgs/scene/gaussian_model.py:220-251,
// 287-332; gs2dgs/scene/gaussian_model.py has the same pair with two scales). The reference copies the six tensors
// to the host one by one, transposes the SH tensors there, concatenates, and fills a numpy structured array through
// `list(map(tuple, attributes))` -- a Python loop over every Gaussian -- and on load fills one numpy column per
// property. Here the vertex rows are built in HBM by one kernel and leave in ONE device->host copy (and the reverse on
// load); the host only writes / parses the PLY header.
//
// Row layout written by save_ply (:220-232), all float32:
//   x y z | nx ny nz (zeros) | f_dc_0..2 | f_rest_0..3R-1 | opacity | scale_0..S-1 | rot_0..3
// with the SH tensors channel-major: f_dc_c = features_dc[i][0][c], f_rest_{c*R+k} = features_rest[i][k][c]
// (transpose(1, 2).flatten(start_dim=1), :238-239; load_ply's reshape (P, 3, R) + transpose is the inverse, :309-310).
#include "common.cuh"

namespace gsr {

constexpr int PLY_MAX_COLS = 128;
struct PlyCols { int16_t col[PLY_MAX_COLS]; };     // destination float (model order) -> column of the vertex row

// model order of the destination floats: xyz(3) f_dc[c](3) f_rest[k][c](3R) opacity(1) scaling(S) rotation(4)
__device__ __forceinline__ float* model_slot(float* xyz, float* dc, float* rest, float* op, float* sc, float* rot,
                                             int R, int S, size_t i, int d) {
    if (d < 3) return xyz + i * 3 + d;
    d -= 3;
    if (d < 3) return dc + i * 3 + d;
    d -= 3;
    if (d < 3 * R) return rest + i * (size_t)(3 * R) + d;
    d -= 3 * R;
    if (d < 1) return op + i;
    d -= 1;
    if (d < S) return sc + i * (size_t)S + d;
    d -= S;
    return rot + i * 4 + d;
}

// one float of one vertex row per thread
__global__ void __launch_bounds__(256)
ply_pack_kernel(int P, int R, int S, GsrModelParams m, float* __restrict__ rows) {
    const int RF = 6 + 3 + 3 * R + 1 + S + 4;
    const long long e = (long long)blockIdx.x * 256 + threadIdx.x;
    if (e >= (long long)P * RF) return;
    const size_t i = (size_t)(e / RF);
    const int c = (int)(e - (long long)i * RF);
    float v;
    if (c < 3) v = __ldg(m.xyz + i * 3 + c);
    else if (c < 6) v = 0.0f;                                            // normals = zeros_like(xyz), :236
    else if (c < 9) v = __ldg(m.features_dc + i * 3 + (c - 6));
    else if (c < 9 + 3 * R) { const int q = c - 9, ch = q / R, k = q - ch * R; v = __ldg(m.features_rest + i * (size_t)(3 * R) + k * 3 + ch); }
    else if (c < 10 + 3 * R) v = __ldg(m.opacity + i);
    else if (c < 10 + 3 * R + S) v = __ldg(m.scaling + i * (size_t)S + (c - 10 - 3 * R));
    else v = __ldg(m.rotation + i * 4 + (c - 10 - 3 * R - S));
    rows[e] = v;
}

// one destination float per thread; the row is read through the column map, so any property order loads
__global__ void __launch_bounds__(256)
ply_unpack_kernel(int P, int row_floats, int R, int S, const float* __restrict__ rows, PlyCols cols,
                  float* xyz, float* dc, float* rest, float* op, float* sc, float* rot) {
    const int DF = 3 + 3 + 3 * R + 1 + S + 4;
    const long long e = (long long)blockIdx.x * 256 + threadIdx.x;
    if (e >= (long long)P * DF) return;
    const size_t i = (size_t)(e / DF);
    const int d = (int)(e - (long long)i * DF);
    *model_slot(xyz, dc, rest, op, sc, rot, R, S, i, d) = __ldg(rows + i * (size_t)row_floats + cols.col[d]);
}

}  // namespace gsr

using namespace gsr;

extern "C" {

int gsr_ply_row_floats(int sh_rest, int scale_dim) { return 6 + 3 + 3 * sh_rest + 1 + scale_dim + 4; }

int gsr_ply_pack(int P, int sh_rest, int scale_dim, const GsrModelParams* model, float* rows, void* stream)
{
    if (P == 0) return GSR_OK;
    if (P < 0 || sh_rest < 0 || scale_dim < 1 || !model || !rows || !model->xyz || !model->features_dc || !model->opacity ||
        !model->scaling || !model->rotation || (sh_rest > 0 && !model->features_rest)) {
        set_error("bad argument to gsr_ply_pack"); return GSR_EINVAL;
    }
    const long long n = (long long)P * gsr_ply_row_floats(sh_rest, scale_dim);
    ply_pack_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(P, sh_rest, scale_dim, *model, rows);
    return check_cuda(cudaGetLastError(), "ply pack launch");
}

int gsr_ply_unpack(int P, int row_floats, const float* rows, const int32_t* columns, int sh_rest, int scale_dim,
                   const GsrModelGrads* out, void* stream)
{
    if (P == 0) return GSR_OK;
    const int DF = 3 + 3 + 3 * sh_rest + 1 + scale_dim + 4;
    if (P < 0 || sh_rest < 0 || scale_dim < 1 || row_floats < 1 || !rows || !columns || !out || !out->xyz || !out->features_dc ||
        !out->opacity || !out->scaling || !out->rotation || (sh_rest > 0 && !out->features_rest)) {
        set_error("bad argument to gsr_ply_unpack"); return GSR_EINVAL;
    }
    if (DF > PLY_MAX_COLS) { set_error("gsr_ply_unpack: %d model floats per Gaussian (max %d)", DF, PLY_MAX_COLS); return GSR_EUNSUPPORTED; }
    if (row_floats > 32767) { set_error("gsr_ply_unpack: row too wide"); return GSR_EUNSUPPORTED; }
    PlyCols cols;
    for (int d = 0; d < DF; ++d) {
        if (columns[d] < 0 || columns[d] >= row_floats) { set_error("gsr_ply_unpack: column %d of destination %d is outside the row", columns[d], d); return GSR_EINVAL; }
        cols.col[d] = (int16_t)columns[d];
    }
    const long long n = (long long)P * DF;
    ply_unpack_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        P, row_floats, sh_rest, scale_dim, rows, cols, out->xyz, out->features_dc, out->features_rest, out->opacity,
        out->scaling, out->rotation);
    return check_cuda(cudaGetLastError(), "ply unpack launch");
}

}  // extern "C"
