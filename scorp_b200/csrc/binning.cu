// binning.cu -- (Gaussian, tile) instance lists sorted by (tile, depth, Gaussian index).
//
// Replaces cub::DeviceScan::InclusiveSum + duplicateWithKeys + cub::DeviceRadixSort::SortPairs
// + identifyTileRanges (reference rasterizer_impl.cu:70-138,278-318). The reference sorts
// N keys (tile<<32 | depth bits) with a stable LSD radix sort whose input is in Gaussian-index
// order, so the result within a tile is ordered by (depth bits, Gaussian index). Here:
//   1. preprocess already counted instances per tile (atomic histogram);
//   2. tile_scan: exclusive scan of the counts -> ranges[tile] (empty tiles stay (0,0)), the longest
//      segment, read back together with N so the host launches only the sort kernels that have work, and
//      in place of each count the absolute END of the tile's segment (the scatter's cursor);
//   3. scatter: every instance goes straight to its tile's segment as (depth bits<<32 | index), the
//      rectangles of a warp walked 32 tiles per round (common.cuh, warp_rects_packed);
//   4. tile_sort: one CTA sorts each segment on that unique 64-bit key: up to 256 entries a register-blocked
//      bitonic network in shared memory; longer ones a bucket pass over the tile's depth range with the
//      16-bit two-pass and the 32-bit four-pass radix sorts behind it for clustered depths (three size
//      classes); global-memory merge steps for segments longer than 8192.
// The output lists are bit-identical to the reference's and cost ~20 B/instance of traffic
// instead of ~150 B/instance for a 6-pass radix sort.
#include "common.cuh"
#include <cstdlib>

namespace gsr {

// ---- 2. exclusive scan over tiles (single CTA, tiles <= a few 10^4) ------------------------
constexpr int ORDER_BINS = 1024;      // == threads of the scan CTA
constexpr int SORT_S_CAP_ = 1024;     // == SORT_S_CAP below (segments up to this length go to the small sort kernel)
__device__ __forceinline__ int order_bin(uint32_t len) { return (int)min(len >> 3, (uint32_t)(ORDER_BINS - 1)); }
// counters (GeomState::counters): [0] N (summed by the preprocess kernel), [1] longest segment, and for the no-sync
// forward (gsr_rasterize_forward_nosync) [2] overflow flag (N > capacity: every range is (0,0), nothing is binned or
// blended), [3] / [4] how many entries at the head of `order` can need the mid / long sort kernel.
__global__ void __launch_bounds__(1024)
tile_scan_kernel(int tiles, uint32_t* __restrict__ count, uint2* __restrict__ ranges, uint32_t* __restrict__ counters,
                 uint32_t* __restrict__ order, uint32_t capacity) {
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry_s;
    __shared__ uint32_t bins[ORDER_BINS];
    uint32_t* const max_len = counters + 1;
    const bool overflow = counters[0] > capacity;
    bins[threadIdx.x] = 0;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    // (one block scan over per-thread runs of K consecutive tiles was tried instead of this loop: 11 -> 13.7 us at C2,
    // 18 -> 27 us at C3 under ncu; the strided second read of the counters costs more than the barriers)
    uint32_t longest = 0;
    for (int base = 0; base < tiles; base += 1024) {
        const int t = base + threadIdx.x;
        const uint32_t c = (t < tiles) ? count[tc(t)] : 0u;
        longest = max(longest, c);
        if (t < tiles) atomicAdd(&bins[order_bin(overflow ? 0u : c)], 1u);      // the class the order loop below will find in ranges[]
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
        if (t < tiles) {
            ranges[t] = (c && !overflow) ? make_uint2(incl - c, incl) : make_uint2(0u, 0u);
            count[tc(t)] = incl;         // from here on the scatter's cursor: absolute END of the tile's segment, counted down
        }
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = incl;
        __syncthreads();
    }
    longest = __reduce_max_sync(0xffffffffu, longest);      // the host picks the sort kernels by the longest segment
    if (lane == 0 && longest) atomicMax(max_len, longest);

    // Launch order for the blend kernels: longest segments first (counting sort on the quantised length), so
    // the crowded tiles of an object-centred view start at once instead of landing in the last wave.
    {
        const uint32_t c = bins[ORDER_BINS - 1 - threadIdx.x];          // thread i owns the i-th longest class
        uint32_t v = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t n = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += n; }
        __syncthreads();
        if (lane == 31) warp_sums[wid] = v;
        __syncthreads();
        if (wid == 0) {
            uint32_t w = warp_sums[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { uint32_t n = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += n; }
            warp_sums[lane] = w;
        }
        __syncthreads();
        bins[ORDER_BINS - 1 - threadIdx.x] = v - c + (wid ? warp_sums[wid - 1] : 0u);   // first slot of the class
        __syncthreads();
        if (threadIdx.x == 0) {
            // classes are 8 lengths wide: every segment of >= 1024 entries sits before the first slot of class 127,
            // every segment of >= 8184 in class 1023, which ends where class 1022 starts
            counters[2] = overflow ? 1u : 0u;
            counters[3] = overflow ? 0u : bins[(SORT_S_CAP_ >> 3) - 1];
            counters[4] = overflow ? 0u : bins[ORDER_BINS - 2];
        }
        __syncthreads();
        for (int t = threadIdx.x; t < tiles; t += 1024) {
            const uint2 r = ranges[t];           // (count[] holds the cursors by now)
            order[atomicAdd(&bins[order_bin(r.y - r.x)], 1u)] = (uint32_t)t;
        }
    }
}

// ---- 3. scatter instances into tile segments ----------------------------------------------
// tile_count is consumed as a cursor: tile_scan replaced each count by the absolute end of the tile's segment, and
// every instance takes the slot below it (atomicSub), so no range lookup is needed. One lane per Gaussian computes the
// tile rectangle; the warp then walks the tiles of its 32 rectangles 32 at a time (warp_rects_packed), so a
// screen-filling splat does not serialise hundreds of returning atomics behind a single thread and a warp of
// medium splats does not take one round trip per splat.
__global__ void __launch_bounds__(256)
scatter_kernel(int P, int per_view, int gx, int gy, const char* __restrict__ recs, int rec_stride, int depth_off,
               const int32_t* __restrict__ radii,
               uint32_t* __restrict__ tile_count, const uint2* __restrict__ ranges, uint64_t* __restrict__ keys,
               const uint32_t* __restrict__ overflow) {
    if (overflow && *overflow) return;        // no-sync forward whose instances exceed the caller's capacity: nothing is binned
    const int idx = blockIdx.x * 256 + threadIdx.x;
    uint32_t x0 = 0, y0 = 0, w = 0, cnt = 0;
    uint32_t tbase = 0;                       // first tile of this record's view (0 for a single view)
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
            tbase = (uint32_t)(idx / per_view) * (uint32_t)(gx * gy);
        }
    }
    // tiles of the warp's 32 rectangles, 32 per round (common.cuh): cursor atomic, then the dependent store
    warp_rects_packed<true>(gx, tbase + y0 * (uint32_t)gx + x0, w, cnt, key, [&](uint32_t t, uint64_t k) {
        keys[atomicSub(tile_count + tc(t), 1u) - 1u] = k;        // tile_scan left the END of the tile's segment there
    });
}

// ---- 3b. batched views: the same two jobs with the atomics aggregated per CTA ---------------------------------------
// A pose sweep (gsr_rasterize_forward_views, C5: 128 views x 200 k Gaussians, 1024 tiles per view) makes 10^8 instances
// land on the 1024 counters of their view: ~800 global atomics per address from preprocess (histogram) and again from
// scatter (cursor). Gaussians arrive in no spatial order, so a CTA of 256 of them cannot combine anything; here a CTA
// takes VIEWS_CHUNK Gaussians of ONE view, counts in a shared-memory copy of that view's tile grid and issues one global
// atomic per non-empty tile: ~17x fewer at C5. The order inside a tile segment is arbitrary either way (the sort fixes it).
// ONE large view was tried the same way (GSR_AGGREGATE=1, chunks of P/296 Gaussians, tile grids up to AGG_MAX_TILES):
// C3 binning 0.84 -> 1.39 ms, C2 0.137 -> 0.158. With ~8 k tiles a chunk of 20 k Gaussians combines only ~10 instances
// per tile and pays for two passes over the records and two sweeps over the grid; left as an experiment switch.
constexpr int VIEWS_CHUNK = 8192;
constexpr int AGG_THREADS = 256;
constexpr int AGG_MAX_TILES = 24576;           // two uint32 per tile in shared memory (192 KB; 8 K x 3 K pixels)

__device__ __forceinline__ void rect_of_record(const char* __restrict__ recs, int rec_stride, const int32_t* __restrict__ radii,
                                               size_t idx, bool in_range, int gx, int gy, uint32_t& x0, uint32_t& y0, uint32_t& w,
                                               uint32_t& cnt) {
    x0 = y0 = w = cnt = 0;
    if (!in_range) return;
    const int rad = radii[idx];
    if (rad <= 0) return;
    const float2 a = __ldg(reinterpret_cast<const float2*>(recs + idx * rec_stride));
    uint32_t x1, y1;
    tile_rect(a.x, a.y, rad, gx, gy, x0, y0, x1, y1);
    w = x1 - x0;
    cnt = w * (y1 - y0);
}
__global__ void __launch_bounds__(AGG_THREADS)
views_histogram_kernel(int per_view, int chunk, int gx, int gy, const char* __restrict__ recs, int rec_stride,
                       const int32_t* __restrict__ radii, uint32_t* __restrict__ tile_count) {
    extern __shared__ uint32_t s_cnt[];
    const int tiles = gx * gy, view = blockIdx.y;
    for (int t = threadIdx.x; t < tiles; t += AGG_THREADS) s_cnt[t] = 0u;
    __syncthreads();
    const int first = blockIdx.x * chunk, last = min(first + chunk, per_view);
    for (int base = first; base < last; base += AGG_THREADS) {            // CTA-uniform trip count
        const int i = base + threadIdx.x;
        uint32_t x0, y0, w, cnt;
        rect_of_record(recs, rec_stride, radii, (size_t)view * per_view + i, i < last, gx, gy, x0, y0, w, cnt);
        warp_rects_packed<false>(gx, y0 * (uint32_t)gx + x0, w, cnt, 0ull, [&](uint32_t t, uint64_t) { atomicAdd(&s_cnt[t], 1u); });
    }
    __syncthreads();
    uint32_t* out = tile_count + tc((size_t)view * tiles);
    for (int t = threadIdx.x; t < tiles; t += AGG_THREADS) { const uint32_t c = s_cnt[t]; if (c) atomicAdd(out + tc(t), c); }
}
__global__ void __launch_bounds__(AGG_THREADS)
views_scatter_kernel(int per_view, int chunk, int gx, int gy, const char* __restrict__ recs, int rec_stride, int depth_off,
                     const int32_t* __restrict__ radii, uint32_t* __restrict__ tile_count, const uint2* __restrict__ ranges,
                     uint64_t* __restrict__ keys, const uint32_t* __restrict__ overflow) {
    if (overflow && *overflow) return;
    extern __shared__ uint32_t s_cnt[];                           // [tiles] counts, then cursors; [tiles] first slot
    const int tiles = gx * gy, view = blockIdx.y;
    uint32_t* s_base = s_cnt + tiles;
    for (int t = threadIdx.x; t < tiles; t += AGG_THREADS) s_cnt[t] = 0u;
    __syncthreads();
    const int first = blockIdx.x * chunk, last = min(first + chunk, per_view);
    for (int base = first; base < last; base += AGG_THREADS) {
        const int i = base + threadIdx.x;
        uint32_t x0, y0, w, cnt;
        rect_of_record(recs, rec_stride, radii, (size_t)view * per_view + i, i < last, gx, gy, x0, y0, w, cnt);
        warp_rects_packed<false>(gx, y0 * (uint32_t)gx + x0, w, cnt, 0ull, [&](uint32_t t, uint64_t) { atomicAdd(&s_cnt[t], 1u); });
    }
    __syncthreads();
    // one cursor atomic per non-empty tile reserves this CTA's block of the tile's segment
    const size_t vt = (size_t)view * tiles;
    for (int t = threadIdx.x; t < tiles; t += AGG_THREADS) {
        const uint32_t c = s_cnt[t];
        if (c) s_base[t] = atomicSub(tile_count + tc(vt + t), c) - c;
        s_cnt[t] = 0u;
    }
    __syncthreads();
    for (int base = first; base < last; base += AGG_THREADS) {
        const int i = base + threadIdx.x;
        const size_t idx = (size_t)view * per_view + i;
        uint32_t x0, y0, w, cnt;
        rect_of_record(recs, rec_stride, radii, idx, i < last, gx, gy, x0, y0, w, cnt);
        uint64_t key = 0;
        if (cnt) key = ((uint64_t)__float_as_uint(__ldg(reinterpret_cast<const float*>(recs + idx * rec_stride + depth_off))) << 32) | (uint32_t)idx;
        warp_rects_packed<true>(gx, y0 * (uint32_t)gx + x0, w, cnt, key, [&](uint32_t t, uint64_t k) { keys[s_base[t] + atomicAdd(&s_cnt[t], 1u)] = k; });
    }
}

// ---- 4. per-tile sort ---------------------------------------------------------------------
// Bitonic network in its all-ascending form (first step of every merge pairs i with its mirror
// image, the remaining steps are half-cleaners). Every comparator puts the smaller key at the
// lower index, so a length L that is not a power of two is handled by treating indices >= L
// as +inf: such keys never move down, comparators that touch only them are skipped.
//
// Register blocking: a thread owns the 2^C elements that interact during C consecutive steps of
// the network (C <= 3), so the m steps of the merge of size 2^m cost ceil(m/3) passes over shared
// memory instead of m, and the merges of size 2, 4, 8 happen entirely in registers on the way in
// from global memory. The last pass of the last merge writes the two output lists directly.
constexpr uint64_t KEY_INF = ~0ull;
__device__ __forceinline__ int sphys(int i) { return i + (i >> 3); }      // 8 B pad per 8 keys: stride-8 owners hit distinct banks
constexpr size_t sphys_host(size_t n) { return n + (n >> 3); }
__device__ __forceinline__ void cx(uint64_t& a, uint64_t& b) {
    const uint64_t lo = a < b ? a : b, hi = a < b ? b : a;
    a = lo; b = hi;
}
// C steps on E = 2^C registers held in ascending address order: [mirror,] then half-cleaners.
template <int C, bool MIRROR>
__device__ __forceinline__ void net_group(uint64_t (&r)[1 << C]) {
    constexpr int E = 1 << C;
    if constexpr (MIRROR) {
#pragma unroll
        for (int e = 0; e < E / 2; ++e) cx(r[e], r[E - 1 - e]);
    }
#pragma unroll
    for (int j = MIRROR ? E / 4 : E / 2; j > 0; j >>= 1) {
#pragma unroll
        for (int e = 0; e < E; ++e)
            if (!(e & j)) cx(r[e], r[e | j]);
    }
}
__device__ __forceinline__ void net_sort8(uint64_t (&r)[8]) {
    cx(r[0], r[1]); cx(r[2], r[3]); cx(r[4], r[5]); cx(r[6], r[7]);                      // merges of 2
    cx(r[0], r[3]); cx(r[1], r[2]); cx(r[4], r[7]); cx(r[5], r[6]);                      // merges of 4
    cx(r[0], r[1]); cx(r[2], r[3]); cx(r[4], r[5]); cx(r[6], r[7]);
    net_group<3, true>(r);                                                               // merge of 8
}

struct SortOut { uint32_t* list; uint64_t* keys; uint64_t tile_hi; };

// One pass = C network steps of the merge of size k over the n-element (padded) segment.
// MIRROR: the group starts with the mirror step of that merge; otherwise `stride` is the distance
// between a thread's elements (top comparator distance = stride << (C-1)). EMIT: results go to the output lists.
template <int C, bool MIRROR, bool EMIT, int THREADS>
__device__ __forceinline__ void sort_pass(uint64_t* s, int L, int n, int k, int stride, const SortOut& o) {
    constexpr int E = 1 << C;
    for (int i = threadIdx.x; i < (n >> C); i += THREADS) {
        int addr[E];
        if constexpr (MIRROR) {
            const int st = k >> C;                      // stride inside each half
            const int blk = i / st, t = i - blk * st;
#pragma unroll
            for (int e = 0; e < E / 2; ++e) {
                addr[e] = blk * k + t + e * st;
                addr[E - 1 - e] = blk * k + k - 1 - t - e * st;
            }
        } else {
            const int low = i & (stride - 1);
            const int base = low | ((i - low) << C);
#pragma unroll
            for (int e = 0; e < E; ++e) addr[e] = base + e * stride;
        }
        if (addr[0] >= L) continue;                     // nothing but padding
        uint64_t r[E];
#pragma unroll
        for (int e = 0; e < E; ++e) r[e] = addr[e] < L ? s[sphys(addr[e])] : KEY_INF;
        net_group<C, MIRROR>(r);
#pragma unroll
        for (int e = 0; e < E; ++e) {
            if (addr[e] < L) {
                if constexpr (EMIT) { o.list[addr[e]] = (uint32_t)r[e]; o.keys[addr[e]] = o.tile_hi | (r[e] >> 32); }
                else s[sphys(addr[e])] = r[e];
            }
        }
    }
}

// Segments with lo_len < L <= CAP. Dynamic shared memory: sphys(CAP) keys.
// (every return below is taken by the whole CTA: L is CTA-uniform)
template <int THREADS, int CAP>
__device__ __forceinline__ void tile_sort_one(uint64_t* s, int tile, int lo_len, const uint2* __restrict__ ranges,
                                              const uint64_t* __restrict__ keys, uint32_t* __restrict__ point_list,
                                              uint64_t* __restrict__ point_list_keys) {
    const uint2 rg = ranges[tile];
    const int L = (int)(rg.y - rg.x);
    if (L <= lo_len || L > CAP) return;
    const uint64_t* seg = keys + rg.x;
    const SortOut o{point_list + rg.x, point_list_keys + rg.x, (uint64_t)tile << 32};
    int n = 8; while (n < L) n <<= 1;
    for (int i = threadIdx.x; i < (n >> 3); i += THREADS) {
        const int base = i << 3;
        if (base >= L) continue;
        uint64_t r[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) r[e] = base + e < L ? seg[base + e] : KEY_INF;
        net_sort8(r);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            if (base + e < L) {
                if (n == 8) { o.list[base + e] = (uint32_t)r[e]; o.keys[base + e] = o.tile_hi | (r[e] >> 32); }
                else s[sphys(base + e)] = r[e];
            }
        }
    }
    if (n == 8) return;
    __syncthreads();
    int m = 4;
    for (int k = 16; k <= n; k <<= 1, ++m) {
        const int c0 = (m % 3) ? (m % 3) : 3;            // steps in the leading (mirror) group; the rest are triples
        if (c0 == 1)      sort_pass<1, true, false, THREADS>(s, L, n, k, 0, o);
        else if (c0 == 2) sort_pass<2, true, false, THREADS>(s, L, n, k, 0, o);
        else              sort_pass<3, true, false, THREADS>(s, L, n, k, 0, o);
        __syncthreads();
        for (int d = k >> (c0 + 1); d >= 4; d >>= 3) {   // d = top comparator distance of the triple
            const int stride = d >> 2;
            if (stride == 1 && k == n) { sort_pass<3, false, true, THREADS>(s, L, n, k, 1, o); return; }
            sort_pass<3, false, false, THREADS>(s, L, n, k, stride, o);
            __syncthreads();
        }
    }
}
template <int THREADS, int CAP>
__global__ void __launch_bounds__(THREADS)
tile_sort_kernel(int lo_len, const uint2* __restrict__ ranges, const uint64_t* __restrict__ keys,
                 uint32_t* __restrict__ point_list, uint64_t* __restrict__ point_list_keys) {
    extern __shared__ uint64_t s[];
    tile_sort_one<THREADS, CAP>(s, blockIdx.x, lo_len, ranges, keys, point_list, point_list_keys);
}
// No-sync forward: the host does not know the segment lengths, so a fixed grid walks the first *n_work entries of
// `order` (tiles by decreasing length; tile_scan_kernel counted how many of them can belong to this size class).
template <int THREADS, int CAP>
__global__ void __launch_bounds__(THREADS)
tile_sort_queued_kernel(int lo_len, const uint32_t* __restrict__ n_work, const uint32_t* __restrict__ order,
                        const uint2* __restrict__ ranges, const uint64_t* __restrict__ keys,
                        uint32_t* __restrict__ point_list, uint64_t* __restrict__ point_list_keys) {
    extern __shared__ uint64_t s[];
    const uint32_t n = *n_work;
    for (uint32_t k = blockIdx.x; k < n; k += gridDim.x) {
        tile_sort_one<THREADS, CAP>(s, (int)order[k], lo_len, ranges, keys, point_list, point_list_keys);
        __syncthreads();
    }
}

// ---- 4b. per-tile radix sort (segments of more than RADIX_MIN entries) ----------------------------------------------
// The 64-bit keys of a segment are unique (depth bits << 32 | Gaussian index) and the reference's order is (depth bits,
// index); the segment arrives in arbitrary order (scatter_kernel's cursor is an atomic).
//
// A stable 8-bit pass ranks with match.any (same-digit lanes of a warp find each other; a warp-private counter per digit
// makes the rank stable) -- and match.any is what binds: ncu shows the ADU pipe at 71 % with the issue slots at 31 % when
// all 32 depth bits are sorted that way in four passes (profiles/r02_ncu_C3_kernels.md). So the sort is done on a 16-bit
// key instead: u = floor((depth - dmin) * 65535 / (dmax - dmin)), a monotone map of the depth over the TILE's own depth
// range (clustered depths still spread over the 65536 values). LSD needs stability only from the second pass on, so the
// low byte is ranked with plain shared-memory atomics and only the high byte with match.any: one match pass, not four.
// A last step puts every run of equal u in (depth bits, index) order by counting inside the run (runs are a few entries
// long). A tile whose entries all share one depth, or that holds a run longer than RUN_MAX, is sorted by the full
// four-pass sort on the depth bits (followed by the same ranking step, which then orders equal depths by index).
constexpr int RADIX_K = 8;                     // positions per thread
__device__ int g_bucket_first = 1;             // GSR_BUCKET_SORT=0 (experiments): skip the bucket pass
__device__ __forceinline__ bool bucket_first() { return g_bucket_first != 0; }
constexpr int RUN_MAX = 48;
constexpr uint32_t BUCKET_MAX = 16;          // the bucket pass gives up on a tile with a fuller bucket than this
template <int THREADS> struct RadixSmem {
    static constexpr int CAP = THREADS * RADIX_K, W = THREADS / 32;
    uint2 a[CAP], b[CAP];                      // .x = Gaussian index, .y = depth bits
    union {
        uint16_t cnt[W][256];                  // radix passes: per-warp digit counters
        uint32_t bin[CAP];                     // bucket pass: one counter, then first slot, per bucket
    };
    uint32_t base[256];
    uint32_t total[256];
    uint32_t skip[4];                          // per pass: every key has the same digit (the pass is the identity)
    uint32_t dmin, dmax, long_run;
};
// exclusive scan of sm.total[256] -> sm.base[256], by warp 0 (8 digits per lane)
template <int THREADS>
__device__ __forceinline__ void radix_scan_totals(RadixSmem<THREADS>& sm, int lane, int w) {
    if (w == 0) {
        uint32_t t[8], sum = 0;
#pragma unroll
        for (int e = 0; e < 8; ++e) { t[e] = sm.total[lane * 8 + e]; sum += t[e]; }
        uint32_t incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t n = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += n; }
        uint32_t run = incl - sum;
#pragma unroll
        for (int e = 0; e < 8; ++e) { sm.base[lane * 8 + e] = run; run += t[e]; }
    }
}
// One STABLE pass on digit(key) from src to dst. Entered and left with the CTA in step (barrier inside and at the end).
// Returns false when every key had the same digit (nothing was moved).
template <int THREADS, class Digit>
__device__ __forceinline__ bool radix_pass_stable(RadixSmem<THREADS>& sm, const uint2* src, uint2* dst, int L, int pass, Digit digit) {
    constexpr int W = THREADS / 32;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    const int p0 = w * 32 * RADIX_K + lane;      // this thread's position in round 0
    for (int i = tid; i < W * 128; i += THREADS) reinterpret_cast<uint32_t*>(&sm.cnt[0][0])[i] = 0u;
    __syncthreads();                            // counters cleared; src complete
    uint2 key[RADIX_K];
    uint32_t rank[RADIX_K];
#pragma unroll
    for (int r = 0; r < RADIX_K; ++r) {
        const int p = p0 + r * 32;
        if (p - lane >= L) { rank[r] = 0; key[r] = make_uint2(0u, 0u); continue; }      // warp-uniform: nothing left for this warp
        const bool valid = p < L;
        if (valid) key[r] = src[p];
        const uint32_t dig = valid ? digit(key[r]) : 256u;
        const uint32_t peers = __match_any_sync(0xffffffffu, dig);
        const int leader = __ffs(peers) - 1;
        uint32_t prior = 0;
        if (lane == leader && valid) {
            prior = sm.cnt[w][dig];
            sm.cnt[w][dig] = (uint16_t)(prior + __popc(peers));
        }
        prior = __shfl_sync(0xffffffffu, prior, leader);
        rank[r] = prior + __popc(peers & lt);
        __syncwarp();                            // the next round reads the counters this one wrote
    }
    __syncthreads();
    // counters -> exclusive prefix over the warps, per digit; totals per digit
    for (int d = tid; d < 256; d += THREADS) {
        uint32_t run = 0;
#pragma unroll 4
        for (int ww = 0; ww < W; ++ww) { const uint32_t c = sm.cnt[ww][d]; sm.cnt[ww][d] = (uint16_t)run; run += c; }
        sm.total[d] = run;
        if (run == (uint32_t)L) sm.skip[pass] = 1u;
    }
    __syncthreads();
    if (sm.skip[pass]) return false;             // CTA-uniform; nobody reads the counters any more
    radix_scan_totals<THREADS>(sm, lane, w);
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RADIX_K; ++r) {
        const int p = p0 + r * 32;
        if (p < L) {
            const uint32_t dig = digit(key[r]);
            dst[sm.base[dig] + sm.cnt[w][dig] + rank[r]] = key[r];
        }
    }
    __syncthreads();                             // scatter done: counters and bases may be overwritten, dst is complete
    return true;
}
// One pass on digit(key) whose order inside a digit is arbitrary (first pass of an LSD sort): slots from shared atomics.
template <int THREADS, class Digit>
__device__ __forceinline__ void radix_pass_unstable(RadixSmem<THREADS>& sm, const uint2* src, uint2* dst, int L, Digit digit) {
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    for (int d = tid; d < 256; d += THREADS) sm.total[d] = 0u;
    __syncthreads();
    uint2 key[RADIX_K];
    uint32_t slot[RADIX_K];
#pragma unroll
    for (int r = 0; r < RADIX_K; ++r) {
        const int p = tid + r * THREADS;         // any assignment of positions to threads will do here
        slot[r] = 0;
        if (p < L) { key[r] = src[p]; slot[r] = atomicAdd(&sm.total[digit(key[r])], 1u); }
    }
    __syncthreads();
    radix_scan_totals<THREADS>(sm, lane, w);
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RADIX_K; ++r) {
        const int p = tid + r * THREADS;
        if (p < L) dst[sm.base[digit(key[r])] + slot[r]] = key[r];
    }
    __syncthreads();
}
// Bucket pass (round 2b): ONE counting pass over CAP buckets of the tile's own depth range, u = floor((depth - dmin) *
// (CAP - 1) / (dmax - dmin)), slots inside a bucket handed out by shared-memory atomics in arrival order; since u is a
// monotone map of the depth, every key's final position is its bucket's first slot plus the number of keys in the bucket
// that are below it in (depth bits, index) order -- buckets hold one or two keys when the depths spread over the range.
// No match.any, no second pass. Returns false (nothing written) when a bucket holds more than BUCKET_MAX keys: clustered
// depths, left to the 16-bit two-pass sort below. Keys are read from sm.a, which is left intact.
template <int THREADS>
__device__ __forceinline__ bool tile_bucket_sort(RadixSmem<THREADS>& sm, int L, float dmin, float range, uint64_t tile_hi,
                                                 uint32_t* __restrict__ out_list, uint64_t* __restrict__ out_keys) {
    constexpr int CAP = THREADS * RADIX_K, W = THREADS / 32;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const float scale = __fdividef((float)(CAP - 1), range);
    auto bucket = [&](const uint2& k) { return min((uint32_t)(CAP - 1), (uint32_t)(__fsub_rn(__uint_as_float(k.y), dmin) * scale)); };
    uint4* bin4 = reinterpret_cast<uint4*>(sm.bin);
    bin4[2 * tid] = make_uint4(0u, 0u, 0u, 0u); bin4[2 * tid + 1] = make_uint4(0u, 0u, 0u, 0u);     // RADIX_K == 8 buckets per thread
    __syncthreads();
    uint2 key[RADIX_K];
    uint32_t slot[RADIX_K];
#pragma unroll
    for (int r = 0; r < RADIX_K; ++r) {
        const int p = tid + r * THREADS;
        slot[r] = 0;
        if (p < L) { key[r] = sm.a[p]; slot[r] = atomicAdd(&sm.bin[bucket(key[r])], 1u); }
    }
    __syncthreads();
    {   // counts -> first slots (exclusive scan over the CAP buckets, 8 consecutive ones per thread); fullest bucket
        const uint4 c0 = bin4[2 * tid], c1 = bin4[2 * tid + 1];
        const uint32_t sum = c0.x + c0.y + c0.z + c0.w + c1.x + c1.y + c1.z + c1.w;
        const uint32_t most = max(max(max(c0.x, c0.y), max(c0.z, c0.w)), max(max(c1.x, c1.y), max(c1.z, c1.w)));
        if (most > BUCKET_MAX) sm.long_run = 1u;
        uint32_t incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t n = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += n; }
        if (lane == 31) sm.total[w] = incl;
        __syncthreads();
        if (w == 0) {
            const uint32_t t = lane < W ? sm.total[lane] : 0u;
            uint32_t ti = t;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t n = __shfl_up_sync(0xffffffffu, ti, o); if (lane >= o) ti += n; }
            if (lane < W) sm.total[lane] = ti - t;
        }
        __syncthreads();
        if (sm.long_run) return false;                 // CTA-uniform
        uint32_t run = sm.total[w] + incl - sum;
        uint4 s0, s1;
        s0.x = run; run += c0.x; s0.y = run; run += c0.y; s0.z = run; run += c0.z; s0.w = run; run += c0.w;
        s1.x = run; run += c1.x; s1.y = run; run += c1.y; s1.z = run; run += c1.z; s1.w = run;
        bin4[2 * tid] = s0; bin4[2 * tid + 1] = s1;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RADIX_K; ++r) {
        const int p = tid + r * THREADS;
        if (p < L) sm.b[sm.bin[bucket(key[r])] + slot[r]] = key[r];
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RADIX_K; ++r) {
        const int p = tid + r * THREADS;
        if (p >= L) continue;
        const uint2 k = sm.b[p];
        const uint32_t u = bucket(k);
        const int s0 = (int)sm.bin[u], e0 = (u + 1 < (uint32_t)CAP) ? (int)sm.bin[u + 1] : L;
        int at = p;
        if (e0 - s0 > 1) {
            int below = 0;
            for (int q = s0; q < e0; ++q) { const uint2 o = sm.b[q]; below += (o.y < k.y) || (o.y == k.y && o.x < k.x); }
            at = s0 + below;
        }
        out_list[at] = k.x; out_keys[at] = tile_hi | k.y;
    }
    return true;
}
template <int THREADS>
__device__ __forceinline__ void tile_radix_one(RadixSmem<THREADS>& sm, int tile, int lo_len, const uint2* __restrict__ ranges,
                                               const uint64_t* __restrict__ keys, uint32_t* __restrict__ point_list,
                                               uint64_t* __restrict__ point_list_keys) {
    constexpr int CAP = THREADS * RADIX_K;
    const uint2 rg = ranges[tile];
    const int L = (int)(rg.y - rg.x);
    if (L <= lo_len || L > CAP) return;         // CTA-uniform
    const int tid = threadIdx.x;
    const uint64_t* seg = keys + rg.x;
    if (tid < 4) sm.skip[tid] = 0u;
    if (tid == 0) { sm.dmin = 0xffffffffu; sm.dmax = 0u; sm.long_run = 0u; }
    __syncthreads();
    {   // load; depth range of the tile (depths are positive floats: their bit patterns order like the values)
        uint32_t lo = 0xffffffffu, hi = 0u;
        for (int p = tid; p < L; p += THREADS) {
            const uint64_t k = seg[p];
            const uint32_t d = (uint32_t)(k >> 32);
            sm.a[p] = make_uint2((uint32_t)k, d);
            lo = min(lo, d); hi = max(hi, d);
        }
        lo = __reduce_min_sync(0xffffffffu, lo); hi = __reduce_max_sync(0xffffffffu, hi);
        if ((tid & 31) == 0) { atomicMin(&sm.dmin, lo); atomicMax(&sm.dmax, hi); }
    }
    __syncthreads();
    const uint32_t dmin_b = sm.dmin, dmax_b = sm.dmax;
    const float dmin = __uint_as_float(dmin_b);
    const float scale = __fdividef(65535.0f, __uint_as_float(dmax_b) - dmin);
    auto mapped = [&](const uint2& k) { return min(65535u, (uint32_t)(__fsub_rn(__uint_as_float(k.y), dmin) * scale)); };
    uint2* src = sm.a;
    uint2* dst = sm.b;
    bool full = (dmin_b == dmax_b);             // one depth for the whole tile: order by index through the full sort
    const uint64_t tile_hi = (uint64_t)tile << 32;
    uint32_t* out_list = point_list + rg.x;
    uint64_t* out_keys = point_list_keys + rg.x;
    if (!full && bucket_first()) {
        if (tile_bucket_sort<THREADS>(sm, L, dmin, __uint_as_float(dmax_b) - dmin, tile_hi, out_list, out_keys)) return;
        __syncthreads();                        // every thread has read the flag
        if (tid == 0) sm.long_run = 0u;         // (its next use is behind the barriers of the radix passes)
    }
    if (!full) {
        radix_pass_unstable<THREADS>(sm, src, dst, L, [&](const uint2& k) { return mapped(k) & 255u; });
        { uint2* t = src; src = dst; dst = t; }
        if (radix_pass_stable<THREADS>(sm, src, dst, L, 0, [&](const uint2& k) { return mapped(k) >> 8; })) { uint2* t = src; src = dst; dst = t; }
    }
    for (int attempt = 0; attempt < 2; ++attempt) {
        if (full) {
            // four stable 8-bit passes on the depth bits, from the segment as it arrived
            if (attempt) {
                for (int p = tid; p < L; p += THREADS) { const uint64_t k = seg[p]; sm.a[p] = make_uint2((uint32_t)k, (uint32_t)(k >> 32)); }
                if (tid < 4) sm.skip[tid] = 0u;
                src = sm.a; dst = sm.b;
                __syncthreads();
            }
#pragma unroll 1
            for (int pass = 0; pass < 4; ++pass) {
                const int shift = 8 * pass;
                if (radix_pass_stable<THREADS>(sm, src, dst, L, pass, [&](const uint2& k) { return (k.y >> shift) & 255u; })) {
                    uint2* t = src; src = dst; dst = t;
                }
            }
        }
        // final positions: runs of equal sort key are put in (depth bits, index) order by counting inside the run
        int at[RADIX_K];
        bool too_long = false;
#pragma unroll
        for (int r = 0; r < RADIX_K; ++r) {
            const int p = tid + r * THREADS;
            at[r] = p;
            if (p >= L) continue;
            const uint2 k = src[p];
            const uint32_t mine = full ? k.y : mapped(k);
            auto same = [&](int q) { const uint2 o = src[q]; return (full ? o.y : mapped(o)) == mine; };
            if ((p > 0 && same(p - 1)) || (p + 1 < L && same(p + 1))) {
                // (the mapped sort gives up on a run of more than ~RUN_MAX entries; exact ties of the full sort may be as
                // long as the list -- a whole plane at one depth -- and are counted to the end)
                int s0 = p, n = 0, below = 0;
                bool give_up = false;
                while (s0 > 0 && same(s0 - 1)) { --s0; if (!full && ++n > RUN_MAX) { give_up = true; break; } }
                if (!give_up) {
                    for (int q = s0; q < L && same(q); ++q) {
                        const uint2 o = src[q];
                        below += (o.y < k.y) || (o.y == k.y && o.x < k.x);
                        if (!full && ++n > 2 * RUN_MAX) { give_up = true; break; }
                    }
                }
                too_long = too_long || give_up;
                at[r] = s0 + below;
            }
        }
        if (!full) {
            if (too_long) sm.long_run = 1u;
            __syncthreads();
            if (sm.long_run) { full = true; continue; }            // CTA-uniform: redo this tile with the full sort
        }
#pragma unroll
        for (int r = 0; r < RADIX_K; ++r) {
            const int p = tid + r * THREADS;
            if (p < L) { const uint2 k = src[p]; out_list[at[r]] = k.x; out_keys[at[r]] = tile_hi | k.y; }
        }
        break;
    }
}
template <int THREADS>
__global__ void __launch_bounds__(THREADS, THREADS <= 512 ? 1024 / THREADS : 1)      // at most 64 registers
tile_radix_kernel(int lo_len, const uint2* __restrict__ ranges, const uint64_t* __restrict__ keys,
                  uint32_t* __restrict__ point_list, uint64_t* __restrict__ point_list_keys) {
    extern __shared__ __align__(16) unsigned char radix_smem[];
    tile_radix_one<THREADS>(*reinterpret_cast<RadixSmem<THREADS>*>(radix_smem), blockIdx.x, lo_len, ranges, keys, point_list,
                            point_list_keys);
}
template <int THREADS>
__global__ void __launch_bounds__(THREADS, THREADS <= 512 ? 1024 / THREADS : 1)
tile_radix_queued_kernel(int lo_len, const uint32_t* __restrict__ n_work, const uint32_t* __restrict__ order,
                         const uint2* __restrict__ ranges, const uint64_t* __restrict__ keys,
                         uint32_t* __restrict__ point_list, uint64_t* __restrict__ point_list_keys) {
    extern __shared__ __align__(16) unsigned char radix_smem[];
    const uint32_t n = *n_work;
    for (uint32_t k = blockIdx.x; k < n; k += gridDim.x) {
        tile_radix_one<THREADS>(*reinterpret_cast<RadixSmem<THREADS>*>(radix_smem), (int)order[k], lo_len, ranges, keys,
                                point_list, point_list_keys);
        __syncthreads();
    }
}

// Segments of up to SORT_S_CAP entries: the network for the short ones, the radix sort above radix_min.
template <int THREADS, int CAP>
__global__ void __launch_bounds__(THREADS)
tile_sort_small_kernel(int radix_min, const uint2* __restrict__ ranges, const uint64_t* __restrict__ keys,
                       uint32_t* __restrict__ point_list, uint64_t* __restrict__ point_list_keys) {
    static_assert(CAP == THREADS * RADIX_K, "one radix CTA covers the size class");
    extern __shared__ __align__(16) unsigned char small_smem[];
    const uint2 rg = ranges[blockIdx.x];
    const int L = (int)(rg.y - rg.x);
    if (L <= radix_min)
        tile_sort_one<THREADS, CAP>(reinterpret_cast<uint64_t*>(small_smem), blockIdx.x, 0, ranges, keys, point_list, point_list_keys);
    else
        tile_radix_one<THREADS>(*reinterpret_cast<RadixSmem<THREADS>*>(small_smem), blockIdx.x, radix_min, ranges, keys, point_list,
                                point_list_keys);
}

// Segments longer than LONG_MIN (rare: > 8192 splats on one tile): chunks sorted in shared memory, merge
// steps with stride >= LONG_CAP in global (L2-resident) memory, the tail of every merge per chunk in shared memory.
constexpr int LONG_CAP = 4096;
constexpr int LONG_THREADS = 1024;
__device__ __forceinline__ void cmpx(uint64_t* a, int lo, int hi) {
    const uint64_t x = a[lo], y = a[hi];
    if (x > y) { a[lo] = y; a[hi] = x; }
}
__device__ __forceinline__ void mirror_step(uint64_t* a, int L, int n, int k) {
    const int h = k >> 1;
    for (int i = threadIdx.x; i < (n >> 1); i += LONG_THREADS) {
        const int blk = i / h, t = i - blk * h;
        const int lo = blk * k + t, hi = blk * k + k - 1 - t;
        if (hi < L) cmpx(a, lo, hi);
    }
    __syncthreads();
}
__device__ __forceinline__ void half_step(uint64_t* a, int L, int n, int j) {
    for (int i = threadIdx.x; i < (n >> 1); i += LONG_THREADS) {
        const int lo = ((i & ~(j - 1)) << 1) | (i & (j - 1));
        const int hi = lo | j;
        if (hi < L) cmpx(a, lo, hi);
    }
    __syncthreads();
}
__device__ __forceinline__ void tile_sort_long_one(uint64_t* s, int tile, int lo_len, const uint2* __restrict__ ranges,
                                                   uint64_t* __restrict__ keys, uint32_t* __restrict__ point_list,
                                                   uint64_t* __restrict__ point_list_keys) {
    const uint2 r = ranges[tile];
    const int L = (int)(r.y - r.x);
    if (L <= lo_len) return;
    uint64_t* seg = keys + r.x;
    const uint64_t tile_hi = (uint64_t)tile << 32;
    int n = LONG_CAP; while (n < L) n <<= 1;
    for (int c0 = 0; c0 < L; c0 += LONG_CAP) {
        const int Lc = min(LONG_CAP, L - c0);
        for (int i = threadIdx.x; i < Lc; i += LONG_THREADS) s[i] = seg[c0 + i];
        __syncthreads();
        for (int k = 2; k <= LONG_CAP; k <<= 1) {
            mirror_step(s, Lc, LONG_CAP, k);
            for (int j = k >> 2; j > 0; j >>= 1) half_step(s, Lc, LONG_CAP, j);
        }
        for (int i = threadIdx.x; i < Lc; i += LONG_THREADS) seg[c0 + i] = s[i];
        __syncthreads();
    }
    for (int k = 2 * LONG_CAP; k <= n; k <<= 1) {
        mirror_step(seg, L, n, k);
        for (int j = k >> 2; j >= LONG_CAP; j >>= 1) half_step(seg, L, n, j);
        for (int c0 = 0; c0 < L; c0 += LONG_CAP) {
            const int Lc = min(LONG_CAP, L - c0);
            for (int i = threadIdx.x; i < Lc; i += LONG_THREADS) s[i] = seg[c0 + i];
            __syncthreads();
            for (int j = LONG_CAP >> 1; j > 0; j >>= 1) half_step(s, Lc, LONG_CAP, j);
            for (int i = threadIdx.x; i < Lc; i += LONG_THREADS) seg[c0 + i] = s[i];
            __syncthreads();
        }
    }
    for (int i = threadIdx.x; i < L; i += LONG_THREADS) {
        const uint64_t k = seg[i];
        point_list[r.x + i] = (uint32_t)k;
        point_list_keys[r.x + i] = tile_hi | (k >> 32);
    }
}
__global__ void __launch_bounds__(LONG_THREADS)
tile_sort_long_kernel(int lo_len, const uint2* __restrict__ ranges, uint64_t* __restrict__ keys,
                      uint32_t* __restrict__ point_list, uint64_t* __restrict__ point_list_keys) {
    __shared__ uint64_t s[LONG_CAP];
    tile_sort_long_one(s, blockIdx.x, lo_len, ranges, keys, point_list, point_list_keys);
}
__global__ void __launch_bounds__(LONG_THREADS)
tile_sort_long_queued_kernel(int lo_len, const uint32_t* __restrict__ n_work, const uint32_t* __restrict__ order,
                             const uint2* __restrict__ ranges, uint64_t* __restrict__ keys,
                             uint32_t* __restrict__ point_list, uint64_t* __restrict__ point_list_keys) {
    __shared__ uint64_t s[LONG_CAP];
    const uint32_t n = *n_work;
    for (uint32_t k = blockIdx.x; k < n; k += gridDim.x) {
        tile_sort_long_one(s, (int)order[k], lo_len, ranges, keys, point_list, point_list_keys);
        __syncthreads();
    }
}

constexpr int SORT_S_CAP = 1024, SORT_S_THREADS = 128;       // most tiles
constexpr int SORT_M1_CAP = 4096, SORT_M1_THREADS = 512;     // crowded tiles: radix sort, three CTAs per SM
constexpr int SORT_M_CAP = 8192, SORT_M_THREADS = 1024;      // very crowded tiles: radix sort, one CTA of 32 warps per SM
static_assert(SORT_S_CAP == SORT_S_CAP_ && SORT_S_CAP == SORT_S_THREADS * RADIX_K && SORT_M1_CAP == SORT_M1_THREADS * RADIX_K &&
              SORT_M_CAP == SORT_M_THREADS * RADIX_K, "size classes");

cudaError_t launch_tile_scan(int tiles, uint32_t* tile_count, uint2* ranges, uint32_t* counters, uint32_t* order,
                             cudaStream_t st, uint32_t capacity) {
    tile_scan_kernel<<<1, 1024, 0, st>>>(tiles, tile_count, ranges, counters, order, capacity);
    return cudaGetLastError();
}

// Shorter segments than this go through the network instead of the radix sort (tunable: GSR_RADIX_MIN; a value >= 1024
// sends the whole small class through the network). Measured on B200, binning stage at C2: 1024 -> 0.200 ms, 512 -> 0.154,
// 256 -> 0.153, 128 -> 0.152, 64 -> 0.153; at C1: 512 -> 0.064, 256 -> 0.060.
static int radix_min() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("GSR_RADIX_MIN"); v = e ? atoi(e) : 256; if (v < 8) v = 8; if (v > SORT_S_CAP) v = SORT_S_CAP; }
    return v;
}
static size_t small_smem_bytes() {
    const size_t a = sphys_host(SORT_S_CAP) * 8, b = sizeof(RadixSmem<SORT_S_THREADS>);
    return a > b ? a : b;
}
// dynamic shared memory above 48 KB has to be requested per kernel and device
static cudaError_t apply_bucket_switch() {
    static bool done[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64 || done[dev]) return cudaSuccess;
    done[dev] = true;
    const char* v = getenv("GSR_BUCKET_SORT");
    if (!v) return cudaSuccess;
    const int on = atoi(v) ? 1 : 0;
    return cudaMemcpyToSymbol(g_bucket_first, &on, sizeof(int));
}
static cudaError_t set_radix_attrs() {
    static bool done[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64 || done[dev]) return cudaSuccess;
    const int m1 = (int)sizeof(RadixSmem<SORT_M1_THREADS>), m = (int)sizeof(RadixSmem<SORT_M_THREADS>);
    if ((e = cudaFuncSetAttribute(tile_radix_kernel<SORT_M1_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, m1)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(tile_radix_kernel<SORT_M_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, m)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(tile_radix_queued_kernel<SORT_M1_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, m1)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(tile_radix_queued_kernel<SORT_M_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, m)) != cudaSuccess) return e;
    done[dev] = true;
    return cudaSuccess;
}

// GSR_AGGREGATE: 0 never, 1 whenever the tile grid fits (experiment), unset: batched views with a small grid
bool views_aggregated(int views, int tiles_per_view, int per_view) {
    static int mode = -2;
    if (mode == -2) {
        const char* e = getenv("GSR_AGGREGATE");
        mode = e ? atoi(e) : -1;
        const char* o = getenv("GSR_VIEWS_NO_AGGREGATE");
        if (o && atoi(o)) mode = 0;
    }
    if (mode == 0 || tiles_per_view > AGG_MAX_TILES) return false;
    if (mode == 1) return true;
    (void)per_view;
    return views > 1 && tiles_per_view <= 6144;
}
static int agg_chunk(int views, int per_view) {
    if (views > 1) return VIEWS_CHUNK;
    static int forced = -1;
    if (forced < 0) { const char* e = getenv("GSR_AGG_CHUNK"); forced = e ? atoi(e) : 0; }
    if (forced > 0) return (forced + AGG_THREADS - 1) / AGG_THREADS * AGG_THREADS;
    int c = (per_view + 2 * 148 - 1) / (2 * 148);
    c = (c + AGG_THREADS - 1) / AGG_THREADS * AGG_THREADS;
    return c < VIEWS_CHUNK ? VIEWS_CHUNK : c;
}
static cudaError_t set_agg_attrs() {
    static bool done[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64 || done[dev]) return cudaSuccess;
    if ((e = cudaFuncSetAttribute(views_histogram_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AGG_MAX_TILES * 4)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(views_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AGG_MAX_TILES * 8)) != cudaSuccess) return e;
    done[dev] = true;
    return cudaSuccess;
}
cudaError_t launch_views_histogram(int P_total, int views, int gx, int gy, const void* recs, int rec_stride, const int32_t* radii,
                                   uint32_t* tile_count, cudaStream_t st) {
    const int per_view = P_total / views, chunk = agg_chunk(views, per_view);
    cudaError_t e = set_agg_attrs();
    if (e != cudaSuccess) return e;
    views_histogram_kernel<<<dim3((per_view + chunk - 1) / chunk, views), AGG_THREADS, (size_t)gx * gy * 4, st>>>(
        per_view, chunk, gx, gy, (const char*)recs, rec_stride, radii, tile_count);
    return cudaGetLastError();
}

cudaError_t launch_binning(int P, int gx, int gy, int N, int max_len, const void* recs, int rec_stride, int depth_off,
                           uint32_t* tile_count, const int32_t* radii, const uint2* ranges, const BinningState& bin,
                           cudaStream_t st, int views, const uint32_t* counters, const uint32_t* order) {
    const int tiles = gx * gy * views;
    if (N <= 0) return cudaSuccess;
    cudaError_t e;
    if ((e = apply_bucket_switch()) != cudaSuccess) return e;
    if (views_aggregated(views, gx * gy, P / views)) {
        const int per_view = P / views, chunk = agg_chunk(views, per_view);
        if ((e = set_agg_attrs()) != cudaSuccess) return e;
        views_scatter_kernel<<<dim3((per_view + chunk - 1) / chunk, views), AGG_THREADS, (size_t)gx * gy * 8, st>>>(
            per_view, chunk, gx, gy, (const char*)recs, rec_stride, depth_off, radii, tile_count, ranges, bin.scratch,
            counters ? counters + 2 : nullptr);
    } else
    scatter_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, P / views, gx, gy, (const char*)recs, rec_stride, depth_off, radii,
                                                    tile_count, ranges, bin.scratch, counters ? counters + 2 : nullptr);
    tile_sort_small_kernel<SORT_S_THREADS, SORT_S_CAP><<<tiles, SORT_S_THREADS, small_smem_bytes(), st>>>(
        radix_min(), ranges, bin.scratch, bin.point_list, bin.point_list_keys);
    const size_t m1 = sizeof(RadixSmem<SORT_M1_THREADS>), m = sizeof(RadixSmem<SORT_M_THREADS>);
    if (counters) {
        // no-sync forward: N is the caller's capacity and the host knows no segment length; fixed grids walk the head of
        // `order` (counters[3] tiles can have more than SORT_S_CAP entries, counters[4] more than SORT_M_CAP) and leave at
        // once when there is nothing for them
        if ((e = set_radix_attrs()) != cudaSuccess) return e;
        const int g1 = tiles < 148 * 3 ? tiles : 148 * 3, g2 = tiles < 148 ? tiles : 148, g3 = tiles < 148 * 2 ? tiles : 148 * 2;
        tile_radix_queued_kernel<SORT_M1_THREADS><<<g1, SORT_M1_THREADS, m1, st>>>(
            SORT_S_CAP, counters + 3, order, ranges, bin.scratch, bin.point_list, bin.point_list_keys);
        tile_radix_queued_kernel<SORT_M_THREADS><<<g2, SORT_M_THREADS, m, st>>>(
            SORT_M1_CAP, counters + 3, order, ranges, bin.scratch, bin.point_list, bin.point_list_keys);
        tile_sort_long_queued_kernel<<<g3, LONG_THREADS, 0, st>>>(SORT_M_CAP, counters + 4, order, ranges, bin.scratch,
                                                                  bin.point_list, bin.point_list_keys);
        return cudaGetLastError();
    }
    // the host read the longest segment back with N: only the kernels that have work are launched
    if (max_len > SORT_S_CAP) {
        if ((e = set_radix_attrs()) != cudaSuccess) return e;
        tile_radix_kernel<SORT_M1_THREADS><<<tiles, SORT_M1_THREADS, m1, st>>>(SORT_S_CAP, ranges, bin.scratch, bin.point_list,
                                                                                bin.point_list_keys);
    }
    if (max_len > SORT_M1_CAP)
        tile_radix_kernel<SORT_M_THREADS><<<tiles, SORT_M_THREADS, m, st>>>(SORT_M1_CAP, ranges, bin.scratch, bin.point_list,
                                                                            bin.point_list_keys);
    if (max_len > SORT_M_CAP)
        tile_sort_long_kernel<<<tiles, LONG_THREADS, 0, st>>>(SORT_M_CAP, ranges, bin.scratch, bin.point_list, bin.point_list_keys);
    return cudaGetLastError();
}

}  // namespace gsr
