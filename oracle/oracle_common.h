/* oracle_common.h -- TEST INFRASTRUCTURE (checker), never linked into or called by the product.
 *
 * Shared helpers of the plain-C CPU restatement of the reference rasterizers. Every function
 * cites the reference file:line it restates (paths relative to /root/reference/submodules/).
 * DGR = diff-gaussian-rasterization/cuda_rasterizer, DSR = diff-surfel-rasterization/cuda_rasterizer.
 *
 * Arithmetic is fp32 in the reference's expression order, compiled with -ffp-contract=off:
 * the CPU cannot reproduce nvcc's FMA contraction or CUDA's expf, so floats agree with the
 * reference CUDA build to a few ulp and integer results (radii, tile rects, sort order) agree
 * except at rounding boundaries. Bit-exactness is asserted only against the reference CUDA build
 * (oracle/_ref); this oracle is pinned by the golden vectors in tests/golden/, which were produced
 * by that build on a B200 (tools/make_golden.py).
 */
#ifndef ORACLE_COMMON_H
#define ORACLE_COMMON_H
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define BLOCK_X 16          /* DGR/DSR config.h:16-17 */
#define BLOCK_Y 16
#define BLOCK_SIZE (BLOCK_X * BLOCK_Y)
#define NUM_CHANNELS 3      /* config.h:15 */

/* DGR auxiliary.h:22-39 */
static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = { 1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                -1.0925484305920792f, 0.5462742152960396f };
static const float SH_C3[7] = { -0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                -0.5900435899266435f };

typedef struct { float x, y, z; } vec3;
typedef struct { float x, y, z, w; } vec4;

static inline vec3 v3(float x, float y, float z) { vec3 r = { x, y, z }; return r; }
static inline vec3 v3add(vec3 a, vec3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline vec3 v3scale(float s, vec3 a) { return v3(s * a.x, s * a.y, s * a.z); }
static inline float v3dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

/* DGR auxiliary.h:41-44 -- note the double-precision intermediate */
static inline float ndc2Pix(float v, int S) { return (float)(((v + 1.0) * S - 1.0) * 0.5); }

/* DGR auxiliary.h:46-56 */
static inline void getRect(float px, float py, int max_radius, int gx, int gy,
                           uint32_t* minx, uint32_t* miny, uint32_t* maxx, uint32_t* maxy) {
    int v;
    v = (int)((px - max_radius) / BLOCK_X); *minx = (uint32_t)(v < 0 ? 0 : (v > gx ? gx : v));
    v = (int)((py - max_radius) / BLOCK_Y); *miny = (uint32_t)(v < 0 ? 0 : (v > gy ? gy : v));
    v = (int)((px + max_radius + BLOCK_X - 1) / BLOCK_X); *maxx = (uint32_t)(v < 0 ? 0 : (v > gx ? gx : v));
    v = (int)((py + max_radius + BLOCK_Y - 1) / BLOCK_Y); *maxy = (uint32_t)(v < 0 ? 0 : (v > gy ? gy : v));
}

/* DGR auxiliary.h:58-66: column-major 4x4, first three rows */
static inline vec3 transformPoint4x3(vec3 p, const float* m) {
    return v3(m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
              m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
              m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]);
}
/* DGR auxiliary.h:68-78 */
static inline vec4 transformPoint4x4(vec3 p, const float* m) {
    vec4 r = { m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
               m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
               m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14],
               m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15] };
    return r;
}
/* DGR auxiliary.h:90-98 */
static inline vec3 transformVec4x3Transpose(vec3 p, const float* m) {
    return v3(m[0] * p.x + m[1] * p.y + m[2] * p.z,
              m[4] * p.x + m[5] * p.y + m[6] * p.z,
              m[8] * p.x + m[9] * p.y + m[10] * p.z);
}
/* DGR auxiliary.h:107-117 */
static inline vec3 dnormvdv(vec3 v, vec3 dv) {
    float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
    float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    vec3 d;
    d.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
    d.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
    d.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
    return d;
}

/* DGR rasterizer_impl.cu:35-50 */
static inline uint32_t getHigherMsb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

/* Stable ascending LSD radix sort of (u64 key, u32 value) pairs on key bits [0, end_bit):
 * the contract of cub::DeviceRadixSort::SortPairs as called at DGR rasterizer_impl.cu:301-309. */
static inline void sort_pairs_stable(uint64_t* keys, uint32_t* vals, size_t n, int end_bit) {
    if (n == 0) return;
    uint64_t* k2 = (uint64_t*)malloc(n * sizeof(uint64_t));
    uint32_t* v2 = (uint32_t*)malloc(n * sizeof(uint32_t));
    uint64_t *ka = keys, *kb = k2; uint32_t *va = vals, *vb = v2;
    for (int shift = 0; shift < end_bit; shift += 8) {
        int bits = end_bit - shift < 8 ? end_bit - shift : 8;
        size_t cnt[257]; memset(cnt, 0, sizeof(cnt));
        uint64_t mask = ((uint64_t)1 << bits) - 1;
        for (size_t i = 0; i < n; ++i) cnt[((ka[i] >> shift) & mask) + 1]++;
        for (int d = 0; d < 256; ++d) cnt[d + 1] += cnt[d];
        for (size_t i = 0; i < n; ++i) { size_t o = cnt[(ka[i] >> shift) & mask]++; kb[o] = ka[i]; vb[o] = va[i]; }
        uint64_t* tk = ka; ka = kb; kb = tk; uint32_t* tv = va; va = vb; vb = tv;
    }
    if (ka != keys) { memcpy(keys, ka, n * sizeof(uint64_t)); memcpy(vals, va, n * sizeof(uint32_t)); }
    free(k2); free(v2);
}


/* shared by the 3DGS and the surfel restatements (the two reference submodules carry identical
 * copies: forward.cu:20-71, backward.cu:20-139); defined in gs_oracle.c. clamped: 3 bytes per Gaussian. */
vec3 orc_computeColorFromSH(int idx, int deg, int max_coeffs, const float* means, const float* campos,
                            const float* shs, uint8_t* clamped);
void orc_sh_backward(int idx, int deg, int max_coeffs, const float* means, const float* campos, const float* shs,
                     const uint8_t* clamped, const float* dL_dcolor, float* dL_dmeans, float* dL_dshs);

#endif
