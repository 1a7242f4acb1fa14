// binning.cu -- (Gaussian, tile) instance lists sorted by (tile, depth, Gaussian index).
//
// Replaces cub::DeviceScan::InclusiveSum + duplicateWithKeys + cub::DeviceRadixSort::SortPairs
// + identifyTileRanges (reference rasterizer_impl.cu:70-138,278-318). The reference sorts
// N keys (tile<<32 | depth bits) with a stable LSD radix sort whose input is in Gaussian-index
// order, so the result within a tile is ordered by (depth bits, Gaussian index). Here:
//   1. preprocess already counted instances per tile (atomic histogram);
//   2. tile_scan: exclusive scan of the counts -> ranges[tile] (empty tiles stay (0,0));
//   3. scatter: every instance goes straight to its tile's segment as (depth bits<<32 | index);
//   4. tile_sort: one CTA sorts each segment on that unique 64-bit key (bitonic in shared
//      memory; global-memory merge steps for segments longer than the shared buffer).
// The output lists are bit-identical to the reference's and cost ~20 B/instance of traffic
// instead of ~150 B/instance for a 6-pass radix sort.
#include "common.cuh"

namespace gsr {

// ---- 2. exclusive scan over tiles (single CTA, tiles <= a few 10^4) ------------------------
__global__ void __launch_bounds__(1024)
tile_scan_kernel(int tiles, const uint32_t* __restrict__ count, uint2* __restrict__ ranges) {
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry_s;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < tiles; base += 1024) {
        const int t = base + threadIdx.x;
        const uint32_t c = (t < tiles) ? count[t] : 0u;
        uint32_t v = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t n = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += n; }
        if (lane == 31) warp_sums[wid] = v;
        __syncthreads();
        if (wid == 0) {
            uint32_t w = warp_sums[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { uint32_t n = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += n; }
            warp_sums[lane] = w;
        }
        __syncthreads();
        const uint32_t carry = carry_s;
        const uint32_t incl = carry + v + (wid ? warp_sums[wid - 1] : 0u);
        if (t < tiles) ranges[t] = c ? make_uint2(incl - c, incl) : make_uint2(0u, 0u);
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = incl;
        __syncthreads();
    }
}

// ---- 3. scatter instances into tile segments ----------------------------------------------
// tile_count is consumed as a cursor (atomicSub), leaving it zero. One lane per Gaussian computes the
// tile rectangle; rectangles of up to SMALL_RECT tiles are written by their own lane, larger ones are
// spread over the whole warp (32 tiles per round), so one screen-filling splat does not serialise
// hundreds of returning atomics behind a single thread.
constexpr uint32_t SMALL_RECT = 4;

__global__ void __launch_bounds__(256)
scatter_kernel(int P, int gx, int gy, const char* __restrict__ recs, int rec_stride, int depth_off,
               const int32_t* __restrict__ radii,
               uint32_t* __restrict__ tile_count, const uint2* __restrict__ ranges, uint64_t* __restrict__ keys) {
    const int idx = blockIdx.x * 256 + threadIdx.x;
    const int lane = threadIdx.x & 31;
    uint32_t x0 = 0, y0 = 0, w = 0, cnt = 0;
    uint64_t key = 0;
    if (idx < P) {
        const int rad = radii[idx];
        if (rad > 0) {
            // every record type starts with the pixel-space centre (x, y); the depth sits at depth_off bytes
            const char* rp = recs + (size_t)idx * rec_stride;
            const float2 a = __ldg(reinterpret_cast<const float2*>(rp));
            const float depth = __ldg(reinterpret_cast<const float*>(rp + depth_off));
            uint32_t x1, y1;
            tile_rect(a.x, a.y, rad, gx, gy, x0, y0, x1, y1);
            w = x1 - x0;
            cnt = w * (y1 - y0);
            key = ((uint64_t)__float_as_uint(depth) << 32) | (uint32_t)idx;
        }
    }
    auto put = [&](uint32_t t, uint64_t k) {
        const uint32_t slot = atomicSub(tile_count + t, 1u) - 1u;
        keys[ranges[t].x + slot] = k;
    };
    if (cnt > 0 && cnt <= SMALL_RECT) {
        for (uint32_t i = 0; i < cnt; ++i) put((y0 + i / w) * gx + x0 + i % w, key);
    }
    uint32_t big = __ballot_sync(0xffffffffu, cnt > SMALL_RECT);
    while (big) {
        const int src = __ffs(big) - 1;
        big &= big - 1;
        const uint32_t bx0 = __shfl_sync(0xffffffffu, x0, src), by0 = __shfl_sync(0xffffffffu, y0, src);
        const uint32_t bw = __shfl_sync(0xffffffffu, w, src), bcnt = __shfl_sync(0xffffffffu, cnt, src);
        const uint64_t bkey = __shfl_sync(0xffffffffu, key, src);
        for (uint32_t i = lane; i < bcnt; i += 32) put((by0 + i / bw) * gx + bx0 + i % bw, bkey);
    }
}

// ---- 4. per-tile sort ---------------------------------------------------------------------
// Bitonic network in its all-ascending form (first step of every merge pairs i with its mirror
// image, the remaining steps are half-cleaners). Every comparator puts the smaller key at the
// lower index, so a length L that is not a power of two is handled by treating indices >= L
// as +inf: comparators that touch them are no-ops and are skipped.
constexpr int SORT_CAP = 4096;        // u64 elements held in shared memory (32 KB)
constexpr int SORT_THREADS = 256;

__device__ __forceinline__ void cmpx(uint64_t* a, int lo, int hi) {
    const uint64_t x = a[lo], y = a[hi];
    if (x > y) { a[lo] = y; a[hi] = x; }
}
// merge step 1 for blocks of size k over n/2 comparators
__device__ __forceinline__ void mirror_step(uint64_t* a, int L, int n, int k) {
    const int h = k >> 1;
    for (int i = threadIdx.x; i < (n >> 1); i += SORT_THREADS) {
        const int blk = i / h, t = i - blk * h;
        const int lo = blk * k + t, hi = blk * k + k - 1 - t;
        if (hi < L) cmpx(a, lo, hi);
    }
    __syncthreads();
}
__device__ __forceinline__ void half_step(uint64_t* a, int L, int n, int j) {
    for (int i = threadIdx.x; i < (n >> 1); i += SORT_THREADS) {
        const int lo = ((i & ~(j - 1)) << 1) | (i & (j - 1));
        const int hi = lo | j;
        if (hi < L) cmpx(a, lo, hi);
    }
    __syncthreads();
}
__device__ void sort_smem(uint64_t* s, int L) {
    int n = 1; while (n < L) n <<= 1;
    for (int k = 2; k <= n; k <<= 1) {
        mirror_step(s, L, n, k);
        for (int j = k >> 2; j > 0; j >>= 1) half_step(s, L, n, j);
    }
}

__global__ void __launch_bounds__(SORT_THREADS)
tile_sort_kernel(int tiles, const uint2* __restrict__ ranges, uint64_t* __restrict__ keys,
                 uint32_t* __restrict__ point_list, uint64_t* __restrict__ point_list_keys) {
    __shared__ uint64_t s[SORT_CAP];
    const int tile = blockIdx.x;
    const uint2 r = ranges[tile];
    const int L = (int)(r.y - r.x);
    if (L <= 0) return;
    uint64_t* seg = keys + r.x;
    const uint64_t tile_hi = (uint64_t)tile << 32;
    if (L <= SORT_CAP) {
        for (int i = threadIdx.x; i < L; i += SORT_THREADS) s[i] = seg[i];
        __syncthreads();
        sort_smem(s, L);
        for (int i = threadIdx.x; i < L; i += SORT_THREADS) {
            const uint64_t k = s[i];
            point_list[r.x + i] = (uint32_t)k;
            point_list_keys[r.x + i] = tile_hi | (k >> 32);
        }
        return;
    }
    // Long segment: chunks sorted in shared memory, merge steps with stride >= SORT_CAP in
    // global (L2-resident) memory, the tail of every merge again per chunk in shared memory.
    int n = SORT_CAP; while (n < L) n <<= 1;
    for (int c0 = 0; c0 < L; c0 += SORT_CAP) {
        const int Lc = min(SORT_CAP, L - c0);
        for (int i = threadIdx.x; i < Lc; i += SORT_THREADS) s[i] = seg[c0 + i];
        __syncthreads();
        sort_smem(s, Lc);
        for (int i = threadIdx.x; i < Lc; i += SORT_THREADS) seg[c0 + i] = s[i];
        __syncthreads();
    }
    for (int k = 2 * SORT_CAP; k <= n; k <<= 1) {
        mirror_step(seg, L, n, k);
        for (int j = k >> 2; j >= SORT_CAP; j >>= 1) half_step(seg, L, n, j);
        for (int c0 = 0; c0 < L; c0 += SORT_CAP) {
            const int Lc = min(SORT_CAP, L - c0);
            for (int i = threadIdx.x; i < Lc; i += SORT_THREADS) s[i] = seg[c0 + i];
            __syncthreads();
            for (int j = SORT_CAP >> 1; j > 0; j >>= 1) half_step(s, Lc, SORT_CAP, j);
            for (int i = threadIdx.x; i < Lc; i += SORT_THREADS) seg[c0 + i] = s[i];
            __syncthreads();
        }
    }
    for (int i = threadIdx.x; i < L; i += SORT_THREADS) {
        const uint64_t k = seg[i];
        point_list[r.x + i] = (uint32_t)k;
        point_list_keys[r.x + i] = tile_hi | (k >> 32);
    }
}

cudaError_t launch_binning(int P, int gx, int gy, int N, const void* recs, int rec_stride, int depth_off,
                           uint32_t* tile_count, const int32_t* radii, uint2* ranges, const BinningState& bin,
                           cudaStream_t st) {
    const int tiles = gx * gy;
    tile_scan_kernel<<<1, 1024, 0, st>>>(tiles, tile_count, ranges);
    if (N > 0) {
        scatter_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, gx, gy, (const char*)recs, rec_stride, depth_off, radii,
                                                        tile_count, ranges, bin.scratch);
        tile_sort_kernel<<<tiles, SORT_THREADS, 0, st>>>(tiles, ranges, bin.scratch, bin.point_list, bin.point_list_keys);
    }
    return cudaGetLastError();
}

}  // namespace gsr
