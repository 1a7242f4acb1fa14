// knn.cu -- simple_knn.distCUDA2: mean squared distance to the 3 nearest other points.
//
// Replaces SimpleKNN::knn (reference submodules/simple-knn/simple_knn.cu:186-221) with its helpers
// coord2Morton (:46-71), boxMinMax (:79-118), distBoxPoint (:120-130), updateKBest<3> (:132-146)
// and boxMeanDist (:148-184). The reference search is EXACT (boxes are only skipped when they
// cannot improve the current 3rd-best), so any exact 3-NN gives the same three distances; the
// distance itself is evaluated with the reference's expression, so the output is bit-identical.
//
// Design: instead of Morton sort + 1024-point boxes scanned linearly by every point
// (O(P * P/1024) box tests), points are counting-sorted into a uniform grid of ~2-4 points per cell
// (histogram -> scan -> scatter, the same shape as the rasterizer's tile binning) and every point
// searches cubic shells of cells around its own cell until the shell's inner face is farther than
// its current 3rd-best. Threads run in cell order, so a warp's neighbourhoods overlap in L1/L2.
#include "surfel.cuh"
#include <cfloat>
#include <cmath>

namespace gsr {

struct KnnGrid {
    float minx, miny, minz, cs, inv_cs;
    int nx, ny, nz;
};

__device__ __forceinline__ uint32_t f2ord(float f) {
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
static inline float ord2f(uint32_t o) {
    const uint32_t u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
    float f; memcpy(&f, &u, 4); return f;
}

// bbox[0..2] = min (ordered-uint encoding, initialised to 0xffffffff), bbox[3..5] = max (initialised to 0)
__global__ void __launch_bounds__(256)
knn_bbox_kernel(int P, const float* __restrict__ xyz, uint32_t* __restrict__ bbox) {
    float lo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, hi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
    for (int i = blockIdx.x * 256 + threadIdx.x; i < P; i += gridDim.x * 256) {
#pragma unroll
        for (int k = 0; k < 3; ++k) { const float v = __ldg(xyz + 3 * (size_t)i + k); lo[k] = fminf(lo[k], v); hi[k] = fmaxf(hi[k], v); }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) { atomicMin(bbox + k, f2ord(lo[k])); atomicMax(bbox + 3 + k, f2ord(hi[k])); }
    }
}

__device__ __forceinline__ int3 knn_cell_of(const KnnGrid& g, float x, float y, float z) {
    int cx = (int)((x - g.minx) * g.inv_cs), cy = (int)((y - g.miny) * g.inv_cs), cz = (int)((z - g.minz) * g.inv_cs);
    cx = min(max(cx, 0), g.nx - 1); cy = min(max(cy, 0), g.ny - 1); cz = min(max(cz, 0), g.nz - 1);
    return make_int3(cx, cy, cz);
}

__global__ void __launch_bounds__(256)
knn_count_kernel(int P, KnnGrid g, const float* __restrict__ xyz, uint32_t* __restrict__ cell_of, uint32_t* __restrict__ count) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= P) return;
    const int3 c = knn_cell_of(g, __ldg(xyz + 3 * (size_t)i), __ldg(xyz + 3 * (size_t)i + 1), __ldg(xyz + 3 * (size_t)i + 2));
    const uint32_t id = ((uint32_t)c.z * g.ny + c.y) * g.nx + c.x;
    cell_of[i] = id;
    atomicAdd(count + id, 1u);
}

// exclusive scan of count[0..n) into start[0..n], start[n] = total. One CTA; thread t owns a contiguous chunk.
__global__ void __launch_bounds__(1024)
knn_scan_kernel(int n, const uint32_t* __restrict__ count, uint32_t* __restrict__ start) {
    __shared__ uint32_t warp_sums[32];
    const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
    const int chunk = (n + 1023) / 1024;
    const int b = min(n, t * chunk), e = min(n, b + chunk);
    uint32_t s = 0;
    for (int i = b; i < e; ++i) s += count[i];
    uint32_t v = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t x = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += x; }
    if (lane == 31) warp_sums[wid] = v;
    __syncthreads();
    if (wid == 0) {
        uint32_t w = warp_sums[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t x = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += x; }
        warp_sums[lane] = w;
    }
    __syncthreads();
    uint32_t run = v - s + (wid ? warp_sums[wid - 1] : 0u);
    for (int i = b; i < e; ++i) { start[i] = run; run += count[i]; }
    if (t == 1023) start[n] = run;
}

__global__ void __launch_bounds__(256)
knn_scatter_kernel(int P, const float* __restrict__ xyz, const uint32_t* __restrict__ cell_of, uint32_t* __restrict__ count,
                   const uint32_t* __restrict__ start, float4* __restrict__ sorted) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= P) return;
    const uint32_t id = cell_of[i];
    const uint32_t slot = atomicSub(count + id, 1u) - 1u;
    sorted[start[id] + slot] = make_float4(__ldg(xyz + 3 * (size_t)i), __ldg(xyz + 3 * (size_t)i + 1), __ldg(xyz + 3 * (size_t)i + 2),
                                           __uint_as_float((uint32_t)i));
}

// updateKBest<3> (simple_knn.cu:132-146), distance in the reference's expression
__device__ __forceinline__ void knn_update3(float qx, float qy, float qz, float4 o, float& b0, float& b1, float& b2) {
    const float dx = o.x - qx, dy = o.y - qy, dz = o.z - qz;
    float dist = dx * dx + dy * dy + dz * dz;
    if (b0 > dist) { const float t = b0; b0 = dist; dist = t; }
    if (b1 > dist) { const float t = b1; b1 = dist; dist = t; }
    if (b2 > dist) { b2 = dist; }
}

__global__ void __launch_bounds__(128)
knn_query_kernel(int P, KnnGrid g, const float4* __restrict__ sorted, const uint32_t* __restrict__ start, float* __restrict__ out) {
    const int i = blockIdx.x * 128 + threadIdx.x;
    if (i >= P) return;
    const float4 q = sorted[i];
    const int3 c = knn_cell_of(g, q.x, q.y, q.z);
    float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;
    const int rmax = max(g.nx, max(g.ny, g.nz));
    for (int r = 0; r <= rmax; ++r) {
        const int z0 = max(c.z - r, 0), z1 = min(c.z + r, g.nz - 1);
        const int y0 = max(c.y - r, 0), y1 = min(c.y + r, g.ny - 1);
        const int x0 = max(c.x - r, 0), x1 = min(c.x + r, g.nx - 1);
        for (int z = z0; z <= z1; ++z) {
            const bool zface = (z == c.z - r) || (z == c.z + r);
            for (int y = y0; y <= y1; ++y) {
                const bool yzface = zface || (y == c.y - r) || (y == c.y + r);
                const uint32_t row = ((uint32_t)z * g.ny + y) * g.nx;
                if (yzface) {       // the whole x-run belongs to the shell: one contiguous range of the sorted array
                    const uint32_t s = __ldg(start + row + x0), e = __ldg(start + row + x1 + 1);
                    for (uint32_t k = s; k < e; ++k) if ((int)k != i) knn_update3(q.x, q.y, q.z, sorted[k], b0, b1, b2);
                } else {            // only the two end cells
                    if (c.x - r >= 0) {
                        const uint32_t s = __ldg(start + row + x0), e = __ldg(start + row + x0 + 1);
                        for (uint32_t k = s; k < e; ++k) if ((int)k != i) knn_update3(q.x, q.y, q.z, sorted[k], b0, b1, b2);
                    }
                    if (c.x + r <= g.nx - 1 && r > 0) {
                        const uint32_t s = __ldg(start + row + x1), e = __ldg(start + row + x1 + 1);
                        for (uint32_t k = s; k < e; ++k) if ((int)k != i) knn_update3(q.x, q.y, q.z, sorted[k], b0, b1, b2);
                    }
                }
            }
        }
        // every unvisited point lies outside the cube of cells [c-r, c+r]; faces beyond the grid bound nothing
        float face = FLT_MAX;
        const float m = 1.0e-3f * g.cs;                   // covers the rounding of the cell assignment
        const float rx = q.x - g.minx, ry = q.y - g.miny, rz = q.z - g.minz;
        if (c.x - r > 0)        face = fminf(face, rx - (c.x - r) * g.cs - m);
        if (c.x + r < g.nx - 1) face = fminf(face, (c.x + r + 1) * g.cs - rx - m);
        if (c.y - r > 0)        face = fminf(face, ry - (c.y - r) * g.cs - m);
        if (c.y + r < g.ny - 1) face = fminf(face, (c.y + r + 1) * g.cs - ry - m);
        if (c.z - r > 0)        face = fminf(face, rz - (c.z - r) * g.cs - m);
        if (c.z + r < g.nz - 1) face = fminf(face, (c.z + r + 1) * g.cs - rz - m);
        if (face == FLT_MAX) break;                       // the cube covers the whole grid
        face = fmaxf(face, 0.0f);
        if (b2 <= face * face * 0.9999f) break;
    }
    out[__float_as_uint(q.w)] = (b0 + b1 + b2) / 3.0f;
}

// workspace: [0,256) bbox + pad | cell_of P u32 | sorted P float4 | count (P/2 + 1024) u32 | start (P/2 + 1025) u32
static inline size_t knn_max_cells(int P) { return (size_t)P / 2 + 1024; }
size_t knn_workspace_bytes(int P) {
    const size_t p = (size_t)(P > 0 ? P : 1);
    return 256 + align_up(p * 4) + align_up(p * 16) + align_up(knn_max_cells(P) * 4) + align_up((knn_max_cells(P) + 1) * 4);
}

cudaError_t launch_knn(int P, const float* xyz, float* out, void* tmp, size_t, cudaStream_t st) {
    char* base = (char*)tmp;
    const size_t p = (size_t)P;
    uint32_t* bbox = (uint32_t*)base;
    uint32_t* cell_of = (uint32_t*)(base + 256);
    float4* sorted = (float4*)(base + 256 + align_up(p * 4));
    uint32_t* count = (uint32_t*)((char*)sorted + align_up(p * 16));
    uint32_t* start = (uint32_t*)((char*)count + align_up(knn_max_cells(P) * 4));
    cudaError_t e;
    if ((e = cudaMemsetAsync(bbox, 0xff, 12, st)) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(bbox + 3, 0x00, 12, st)) != cudaSuccess) return e;
    knn_bbox_kernel<<<min((P + 255) / 256, 148 * 8), 256, 0, st>>>(P, xyz, bbox);
    uint32_t hb[6];
    if ((e = cudaMemcpyAsync(hb, bbox, 24, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
    if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return e;
    float lo[3], ext[3], maxext = 0.0f;
    for (int k = 0; k < 3; ++k) { lo[k] = ord2f(hb[k]); ext[k] = ord2f(hb[3 + k]) - lo[k]; if (!(ext[k] >= 0.0f) || !std::isfinite(ext[k])) ext[k] = 0.0f; maxext = std::fmax(maxext, ext[k]); }
    KnnGrid g;
    g.minx = lo[0]; g.miny = lo[1]; g.minz = lo[2];
    const size_t budget = knn_max_cells(P) - 512;
    if (maxext <= 0.0f) { g.cs = 1.0f; g.nx = g.ny = g.nz = 1; }
    else {
        float cs = maxext / 1024.0f;
        for (;;) {
            const size_t nx = (size_t)(ext[0] / cs) + 1, ny = (size_t)(ext[1] / cs) + 1, nz = (size_t)(ext[2] / cs) + 1;
            if (nx * ny * nz <= budget && nx <= 1024 + 1 && ny <= 1024 + 1 && nz <= 1024 + 1) { g.nx = (int)nx; g.ny = (int)ny; g.nz = (int)nz; break; }
            cs *= 1.2599211f;
        }
        g.cs = cs;
    }
    g.inv_cs = 1.0f / g.cs;
    const int cells = g.nx * g.ny * g.nz;
    if ((e = cudaMemsetAsync(count, 0, (size_t)cells * 4, st)) != cudaSuccess) return e;
    knn_count_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, g, xyz, cell_of, count);
    knn_scan_kernel<<<1, 1024, 0, st>>>(cells, count, start);
    knn_scatter_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, xyz, cell_of, count, start, sorted);
    knn_query_kernel<<<(P + 127) / 128, 128, 0, st>>>(P, g, sorted, start, out);
    return cudaGetLastError();
}

}  // namespace gsr

// ---- gsr_knn_dist2_static: the same computation with NO caller scratch and NO allocation ---------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA-graph capture regions, tight serving loops, harnesses that forbid it). The workspace launch_knn needs lives in a
// file-scope __device__ array sized for GSR_KNN_STATIC_MAX_POINTS points (about 6.3 MB, held by every GPU that loads the
// library until the process exits). In host code `g_knn_static_scratch` is only a host-side stand-in for the symbol, not
// a device address: the address a kernel may use comes from cudaGetSymbolAddress, queried once per device with that
// device current, below. Every call shares the array; launch_knn re-initialises the parts that must start clean (the
// bounding box and the per-cell counters are memset, every other sub-array is written before it is read), so the
// previous call's contents do not matter -- but two calls that overlap on the GPU (different streams, or different host
// threads) race on it.
#include <mutex>

namespace {

constexpr int KNN_STATIC_MAX_P = GSR_KNN_STATIC_MAX_POINTS;
constexpr size_t cexpr_align(size_t v) { return (v + 255) / 256 * 256; }
constexpr size_t cexpr_cells(size_t P) { return P / 2 + 1024; }
// == knn_workspace_bytes(KNN_STATIC_MAX_P), restated as a constant expression (checked at run time below)
constexpr size_t KNN_STATIC_BYTES = 256 + cexpr_align((size_t)KNN_STATIC_MAX_P * 4) + cexpr_align((size_t)KNN_STATIC_MAX_P * 16) +
                                    cexpr_align(cexpr_cells(KNN_STATIC_MAX_P) * 4) + cexpr_align((cexpr_cells(KNN_STATIC_MAX_P) + 1) * 4);

static __device__ __align__(256) unsigned char g_knn_static_scratch[KNN_STATIC_BYTES];

constexpr int MAX_DEVICES = 64;
std::mutex g_knn_static_mu;
void* g_knn_static_ptr[MAX_DEVICES] = {};

// device address of the scratch array on the CURRENT device (cached per device); nullptr + error text on failure
void* knn_static_scratch_ptr() {
    int dev = -1;
    if (gsr::check_cuda(cudaGetDevice(&dev), "cudaGetDevice")) return nullptr;
    if (dev < 0 || dev >= MAX_DEVICES) { gsr::set_error("gsr_knn_dist2_static: device ordinal %d not supported", dev); return nullptr; }
    std::lock_guard<std::mutex> lock(g_knn_static_mu);
    if (!g_knn_static_ptr[dev]) {
        void* p = nullptr;
        if (gsr::check_cuda(cudaGetSymbolAddress(&p, g_knn_static_scratch), "cudaGetSymbolAddress(knn scratch)")) return nullptr;
        g_knn_static_ptr[dev] = p;
    }
    return g_knn_static_ptr[dev];
}

}  // namespace

extern "C" int gsr_knn_static_max_points(void) { return KNN_STATIC_MAX_P; }

extern "C" int gsr_knn_dist2_static(int P, const float* xyz, float* out, void* stream) {
    if (P == 0) return GSR_OK;
    if (P < 0 || !xyz || !out) { gsr::set_error("bad argument to gsr_knn_dist2_static"); return GSR_EINVAL; }
    if (P > KNN_STATIC_MAX_P) {
        gsr::set_error("gsr_knn_dist2_static: %d points exceed the built-in scratch (at most %d); use gsr_knn_dist2 with caller scratch",
                       P, KNN_STATIC_MAX_P);
        return GSR_EINVAL;
    }
    if (gsr::knn_workspace_bytes(P) > KNN_STATIC_BYTES) { gsr::set_error("gsr_knn_dist2_static: scratch size mismatch"); return GSR_EINVAL; }
    void* tmp = knn_static_scratch_ptr();
    if (!tmp) return GSR_ECUDA;
    return gsr::check_cuda(gsr::launch_knn(P, xyz, out, tmp, KNN_STATIC_BYTES, (cudaStream_t)stream), "knn launch (static scratch)");
}

