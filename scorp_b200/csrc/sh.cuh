// sh.cuh -- spherical-harmonics colour evaluation and its backward, shared by the 3DGS and the 2DGS
// surfel paths (the two reference submodules carry identical copies: diff-gaussian-rasterization
// forward.cu:20-71 / backward.cu:20-139 and diff-surfel-rasterization forward.cu:20-71 / backward.cu:20-139).
#pragma once
#include "common.cuh"

namespace gsr {

constexpr float kC0 = 0.28209479177387814f;
constexpr float kC1 = 0.4886025119029199f;

struct Sh3 { float r, g, b; };

// 32-byte read-only load (sm_100: LDG.E.256): a thread-per-Gaussian read of a 192-byte SH row touches one 32-byte sector
// per lane and instruction; with 16-byte loads every sector is requested twice (the second time out of L1, if it is still
// there) and the row costs 12 x 32 LSU wavefronts per warp, with 32-byte loads 6 x 32. Needs a 32-byte aligned address.
__device__ __forceinline__ void ldg256(const float* p, float4& lo, float4& hi) {
    asm("ld.global.nc.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
        : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w), "=f"(hi.x), "=f"(hi.y), "=f"(hi.z), "=f"(hi.w) : "l"(p));
}

// Loads coefficient k (3 floats) of one Gaussian. VEC: 48 floats were preloaded as float4s.
template <bool VEC>
struct ShReader {
    const float* base;     // scalar path: pointer to this Gaussian's coefficient 0 (features_dc in the split layout)
    const float* rest;     // scalar path: coefficients 1.. (base + 3, or this Gaussian's features_rest row)
    float4 v[VEC ? 12 : 1];
    __device__ __forceinline__ Sh3 get(int k) const {
        if constexpr (VEC) {
            const float* f = reinterpret_cast<const float*>(v);
            return { f[3 * k], f[3 * k + 1], f[3 * k + 2] };
        } else {
            const float* p = (k == 0) ? base : rest + 3 * (k - 1);
            return { __ldg(p), __ldg(p + 1), __ldg(p + 2) };
        }
    }
};

__device__ __forceinline__ void sh_axpy(Sh3& acc, float w, const Sh3& c) {
    acc.r = fmar(w, c.r, acc.r); acc.g = fmar(w, c.g, acc.g); acc.b = fmar(w, c.b, acc.b);
}

// SH -> RGB for degree D (forward.cu:20-71), same rounding sequence as the reference build.
template <bool VEC>
__device__ __forceinline__ Sh3 eval_sh(int D, const ShReader<VEC>& sh, float x, float y, float z) {
    Sh3 c0 = sh.get(0);
    Sh3 res = { mulr(c0.r, kC0), mulr(c0.g, kC0), mulr(c0.b, kC0) };
    if (D > 0) {
        sh_axpy(res, -mulr(y, kC1), sh.get(1));
        sh_axpy(res,  mulr(z, kC1), sh.get(2));
        sh_axpy(res, -mulr(x, kC1), sh.get(3));
        if (D > 1) {
            const float xx = mulr(x, x), yy = mulr(y, y), zz = mulr(z, z);
            const float xy = mulr(x, y), yz = mulr(y, z), xz = mulr(x, z);
            const float zz2 = addr(zz, zz);
            const float xx_yy = subr(xx, yy);
            sh_axpy(res, mulr(xy, 1.0925484305920792f), sh.get(4));
            sh_axpy(res, mulr(yz, -1.0925484305920792f), sh.get(5));
            sh_axpy(res, mulr(subr(subr(zz2, xx), yy), 0.31539156525252005f), sh.get(6));
            sh_axpy(res, mulr(xz, -1.0925484305920792f), sh.get(7));
            sh_axpy(res, mulr(xx_yy, 0.5462742152960396f), sh.get(8));
            if (D > 2) {
                const float t4z = subr(fmar(zz, 4.0f, -xx), yy);   // 4zz - xx - yy
                sh_axpy(res, mulr(mulr(y, -0.5900435899266435f), fmar(xx, 3.0f, -yy)), sh.get(9));
                sh_axpy(res, mulr(z, mulr(xy, 2.890611442640554f)), sh.get(10));
                sh_axpy(res, mulr(mulr(y, -0.4570457994644658f), t4z), sh.get(11));
                sh_axpy(res, mulr(mulr(z, 0.3731763325901154f), fmar(yy, -3.0f, fmar(xx, -3.0f, zz2))), sh.get(12));
                sh_axpy(res, mulr(mulr(x, -0.4570457994644658f), t4z), sh.get(13));
                sh_axpy(res, mulr(mulr(z, 1.445305721320277f), xx_yy), sh.get(14));
                sh_axpy(res, mulr(mulr(x, -0.5900435899266435f), fmar(yy, -3.0f, xx)), sh.get(15));
            }
        }
    }
    res.r = addr(res.r, 0.5f); res.g = addr(res.g, 0.5f); res.b = addr(res.b, 0.5f);
    return res;
}

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
__device__ __forceinline__ V3 operator*(float s, V3 a) { return { s * a.x, s * a.y, s * a.z }; }
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

__device__ __forceinline__ V3 ld3(const float* p) { return { __ldg(p), __ldg(p + 1), __ldg(p + 2) }; }
__device__ __forceinline__ V3 ld3p(const float* p) { return { p[0], p[1], p[2] }; }      // generic (shared or global)
__device__ __forceinline__ void st3(float* p, V3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }

// Real SH basis b_k(dir) up to degree D (dir normalised); returns how many entries are set. dL/dsh[k] = b_k * dL/dRGB
// (backward.cu:20-141), which is what makes the SH gradient of ONE view a rank-1 outer product.
__device__ __forceinline__ int sh_basis(int D, float x, float y, float z, float (&basis)[16]) {
    const float C1 = 0.4886025119029199f;
    const float C2[5] = { 1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f, 0.5462742152960396f };
    const float C3[7] = { -0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                          -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f };
    basis[0] = 0.28209479177387814f;
    if (D < 1) return 1;
    basis[1] = -C1 * y; basis[2] = C1 * z; basis[3] = -C1 * x;
    if (D < 2) return 4;
    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
    basis[4] = C2[0] * xy; basis[5] = C2[1] * yz; basis[6] = C2[2] * (2.f * zz - xx - yy);
    basis[7] = C2[3] * xz; basis[8] = C2[4] * (xx - yy);
    if (D < 3) return 9;
    basis[9] = C3[0] * y * (3.f * xx - yy); basis[10] = C3[1] * xy * z;
    basis[11] = C3[2] * y * (4.f * zz - xx - yy); basis[12] = C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy);
    basis[13] = C3[4] * x * (4.f * zz - xx - yy); basis[14] = C3[5] * z * (xx - yy);
    basis[15] = C3[6] * x * (xx - 3.f * yy);
    return 16;
}

// SH backward (backward.cu:20-139). Writes dL_dsh[0..M) (zeros beyond the active degree) and
// returns the colour-induced part of dL_dmean.
__device__ __forceinline__ V3 sh_backward(int D, int M, const float* sh, V3 pos, V3 campos,
                                          V3 dRGB, float* dsh) {
    const V3 dir_orig = { pos.x - campos.x, pos.y - campos.y, pos.z - campos.z };
    const float inv_len = 1.0f / sqrtf(dot(dir_orig, dir_orig));
    const float x = dir_orig.x * inv_len, y = dir_orig.y * inv_len, z = dir_orig.z * inv_len;
    const float C1 = 0.4886025119029199f;
    const float C2[5] = { 1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f, 0.5462742152960396f };
    const float C3[7] = { -0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                          -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f };
    V3 gx = { 0, 0, 0 }, gy = { 0, 0, 0 }, gz = { 0, 0, 0 };   // dRGB/dx, /dy, /dz
    float basis[16];
    const int used = sh_basis(D, x, y, z, basis);
    if (D > 0) {
        gx = -C1 * ld3p(sh + 9); gy = -C1 * ld3p(sh + 3); gz = C1 * ld3p(sh + 6);
        if (D > 1) {
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            const V3 s4 = ld3p(sh + 12), s5 = ld3p(sh + 15), s6 = ld3p(sh + 18), s7 = ld3p(sh + 21), s8 = ld3p(sh + 24);
            gx = gx + (C2[0] * y) * s4 + (C2[2] * 2.f * -x) * s6 + (C2[3] * z) * s7 + (C2[4] * 2.f * x) * s8;
            gy = gy + (C2[0] * x) * s4 + (C2[1] * z) * s5 + (C2[2] * 2.f * -y) * s6 + (C2[4] * 2.f * -y) * s8;
            gz = gz + (C2[1] * y) * s5 + (C2[2] * 2.f * 2.f * z) * s6 + (C2[3] * x) * s7;
            if (D > 2) {
                const V3 s9 = ld3p(sh + 27), s10 = ld3p(sh + 30), s11 = ld3p(sh + 33), s12 = ld3p(sh + 36),
                         s13 = ld3p(sh + 39), s14 = ld3p(sh + 42), s15 = ld3p(sh + 45);
                gx = gx + (C3[0] * 3.f * 2.f * xy) * s9 + (C3[1] * yz) * s10 + (C3[2] * -2.f * xy) * s11
                        + (C3[3] * -3.f * 2.f * xz) * s12 + (C3[4] * (-3.f * xx + 4.f * zz - yy)) * s13
                        + (C3[5] * 2.f * xz) * s14 + (C3[6] * 3.f * (xx - yy)) * s15;
                gy = gy + (C3[0] * 3.f * (xx - yy)) * s9 + (C3[1] * xz) * s10 + (C3[2] * (-3.f * yy + 4.f * zz - xx)) * s11
                        + (C3[3] * -3.f * 2.f * yz) * s12 + (C3[4] * -2.f * xy) * s13 + (C3[5] * -2.f * yz) * s14
                        + (C3[6] * -3.f * 2.f * xy) * s15;
                gz = gz + (C3[1] * xy) * s10 + (C3[2] * 4.f * 2.f * yz) * s11 + (C3[3] * 3.f * (2.f * zz - xx - yy)) * s12
                        + (C3[4] * 4.f * 2.f * xz) * s13 + (C3[5] * (xx - yy)) * s14;
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 16; ++k) {       // M <= 16; constant indices keep basis[] in registers
        if (k < M) {
            const float bk = (k < used) ? basis[k] : 0.0f;
            st3(dsh + 3 * k, bk * dRGB);
        }
    }
    // through the normalisation of the view direction (dnormvdv, auxiliary.h:107-117)
    const V3 ddir = { dot(gx, dRGB), dot(gy, dRGB), dot(gz, dRGB) };
    const V3 v = dir_orig;
    const float sum2 = dot(v, v);
    const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    V3 dmean;
    dmean.x = ((sum2 - v.x * v.x) * ddir.x - v.y * v.x * ddir.y - v.z * v.x * ddir.z) * invsum32;
    dmean.y = (-v.x * v.y * ddir.x + (sum2 - v.y * v.y) * ddir.y - v.z * v.y * ddir.z) * invsum32;
    dmean.z = (-v.x * v.z * ddir.x - v.y * v.z * ddir.y + (sum2 - v.z * v.z) * ddir.z) * invsum32;
    return dmean;
}


// ---- shared-memory staging of one CTA's SH rows (backward kernels) ------------------------------
// The SH coefficients and their gradients are 3M contiguous floats per Gaussian: a thread-per-Gaussian
// access touches 32 different sectors per instruction. The CTA instead moves its SH_TILE rows with
// unit-stride accesses (one 128-byte line per warp instruction) into a padded shared tile, computes
// in place, and writes the gradient rows back the same way. Rows of invisible Gaussians are neither
// read (their sectors are never requested) nor computed; they are written as zeros.
constexpr int SH_TILE = 128;
template <int M3> struct ShTile { static constexpr int STRIDE = M3 | 1; };     // odd stride: conflict-free rows

template <int M3>
__device__ __forceinline__ void sh_tile_load(const float* __restrict__ shs, size_t row0, int nrows,
                                             const uint8_t* s_vis, float* s_sh) {
    const float* src = shs + row0 * M3;
#pragma unroll 8       // eight independent loads in flight per thread
    for (int e = threadIdx.x; e < nrows * M3; e += SH_TILE) {
        const int g = e / M3, c = e - g * M3;
        if (s_vis[g]) s_sh[g * ShTile<M3>::STRIDE + c] = __ldg(src + e);
    }
}
// same tile, fetched with 4-byte cp.async (the odd row stride rules out wider ones): every element of the tile is
// in flight at once; the caller waits (cp_async_wait<0> + barrier) before reading s_sh
__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gsrc) {
    unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" :: "r"(d), "l"(gsrc));
}
template <int M3>
__device__ __forceinline__ void sh_tile_load_async(const float* __restrict__ shs, size_t row0, int nrows,
                                                   const uint8_t* s_vis, float* s_sh) {
    const float* src = shs + row0 * M3;
#pragma unroll 8
    for (int e = threadIdx.x; e < nrows * M3; e += SH_TILE) {
        const int g = e / M3, c = e - g * M3;
        if (s_vis[g]) cp_async4(s_sh + g * ShTile<M3>::STRIDE + c, src + e);
    }
    cp_async_commit();
}
// split layout (model-space entry points): column 0..2 of the tile from dc[P,3], the rest from rest[P,M3-3]
template <int M3>
__device__ __forceinline__ void sh_tile_load_async_split(const float* __restrict__ dc, const float* __restrict__ rest,
                                                         size_t row0, int nrows, const uint8_t* s_vis, float* s_sh) {
    const float* src0 = dc + row0 * 3;
    for (int e = threadIdx.x; e < nrows * 3; e += SH_TILE) {
        const int g = e / 3, c = e - g * 3;
        if (s_vis[g]) cp_async4(s_sh + g * ShTile<M3>::STRIDE + c, src0 + e);
    }
    if constexpr (M3 > 3) {
        constexpr int R3 = M3 - 3;
        const float* src = rest + row0 * R3;
#pragma unroll 8
        for (int e = threadIdx.x; e < nrows * R3; e += SH_TILE) {
            const int g = e / R3, c = e - g * R3;
            if (s_vis[g]) cp_async4(s_sh + g * ShTile<M3>::STRIDE + 3 + c, src + e);
        }
    }
    cp_async_commit();
}
template <int M3>
__device__ __forceinline__ void sh_tile_store_split(float* __restrict__ ddc, float* __restrict__ drest, size_t row0, int nrows,
                                                    const uint8_t* s_vis, const float* s_sh) {
    float* dst0 = ddc + row0 * 3;
    for (int e = threadIdx.x; e < nrows * 3; e += SH_TILE) {
        const int g = e / 3, c = e - g * 3;
        dst0[e] = s_vis[g] ? s_sh[g * ShTile<M3>::STRIDE + c] : 0.0f;
    }
    if constexpr (M3 > 3) {
        constexpr int R3 = M3 - 3;
        float* dst = drest + row0 * R3;
#pragma unroll 8
        for (int e = threadIdx.x; e < nrows * R3; e += SH_TILE) {
            const int g = e / R3, c = e - g * R3;
            dst[e] = s_vis[g] ? s_sh[g * ShTile<M3>::STRIDE + 3 + c] : 0.0f;
        }
    }
}
template <int M3>
__device__ __forceinline__ void sh_tile_store(float* __restrict__ dsh, size_t row0, int nrows,
                                              const uint8_t* s_vis, const float* s_sh) {
    float* dst = dsh + row0 * M3;
#pragma unroll 8
    for (int e = threadIdx.x; e < nrows * M3; e += SH_TILE) {
        const int g = e / M3, c = e - g * M3;
        dst[e] = s_vis[g] ? s_sh[g * ShTile<M3>::STRIDE + c] : 0.0f;
    }
}

// ---- 16-byte variant of the staging for M3 == 48 (degree-3 storage, the headline configs) -------------------------
// The 4-byte tile above spends ~24 instructions per Gaussian-float4 on index arithmetic (ncu, C3: the two copies were
// 55 % of the kernel's instructions). Here a row is 12 float4 chunks at stride 48 floats and chunk q of row g sits at
// chunk (q + g) mod 12 of that row: a quarter warp (the unit a 128-bit shared access is served in) reading "its own
// row, chunk q" then hits 8 different 16-byte bank groups, so both the unit-stride copies and the per-Gaussian
// accesses are 128-bit and conflict-free. Needs the global rows 16-byte aligned (the launcher checks the bases).
__device__ __forceinline__ int sh48_chunk(int g, int q) {      // q in [0, 12)
    int p = q + g % 12;
    p -= (p >= 12) ? 12 : 0;
    return g * 12 + p;
}
template <int THREADS = SH_TILE>
__device__ __forceinline__ void sh48_load_async(const float* __restrict__ shs, size_t row0, int nrows,
                                                const uint8_t* s_vis, float4* s4) {
    const float4* src = reinterpret_cast<const float4*>(shs + row0 * 48);
#pragma unroll 12
    for (int e = threadIdx.x; e < nrows * 12; e += THREADS) {
        const int g = e / 12, q = e - g * 12;
        if (s_vis[g]) cp_async16(s4 + sh48_chunk(g, q), src + e);
    }
    cp_async_commit();
}
__device__ __forceinline__ void sh48_store(float* __restrict__ dsh, size_t row0, int nrows,
                                           const uint8_t* s_vis, const float4* s4) {
    float4* dst = reinterpret_cast<float4*>(dsh + row0 * 48);
#pragma unroll 12
    for (int e = threadIdx.x; e < nrows * 12; e += SH_TILE) {
        const int g = e / 12, q = e - g * 12;
        dst[e] = s_vis[g] ? s4[sh48_chunk(g, q)] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
}
// sh_backward on one row of that tile, in place. Same mathematics, organised per stored float j = 3k + channel so that
// a chunk is read once, used and overwritten with its gradients: d RGB_c / d(x,y,z) accumulate w{x,y,z}[k] * sh[j].
__device__ __forceinline__ V3 sh48_backward(int D, float4* row4, int g, V3 pos, V3 campos, V3 dRGB) {
    const V3 dir_orig = { pos.x - campos.x, pos.y - campos.y, pos.z - campos.z };
    const float inv_len = 1.0f / sqrtf(dot(dir_orig, dir_orig));
    const float x = dir_orig.x * inv_len, y = dir_orig.y * inv_len, z = dir_orig.z * inv_len;
    const float C1 = 0.4886025119029199f;
    const float C2[5] = { 1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f, 0.5462742152960396f };
    const float C3[7] = { -0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                          -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f };
    float basis[16];
    const int used = sh_basis(D, x, y, z, basis);
#pragma unroll
    for (int k = 0; k < 16; ++k) basis[k] = (k < used) ? basis[k] : 0.0f;
    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
    const float d1 = D > 0 ? 1.f : 0.f, d2 = D > 1 ? 1.f : 0.f, d3 = D > 2 ? 1.f : 0.f;
    // weights of coefficient k in dRGB/dx, /dy, /dz (backward.cu:58-127); entries that are identically zero are skipped
    const float wx[16] = { 0.f, 0.f, 0.f, -C1 * d1,
                           d2 * C2[0] * y, 0.f, d2 * C2[2] * 2.f * -x, d2 * C2[3] * z, d2 * C2[4] * 2.f * x,
                           d3 * C3[0] * 3.f * 2.f * xy, d3 * C3[1] * yz, d3 * C3[2] * -2.f * xy, d3 * C3[3] * -3.f * 2.f * xz,
                           d3 * C3[4] * (-3.f * xx + 4.f * zz - yy), d3 * C3[5] * 2.f * xz, d3 * C3[6] * 3.f * (xx - yy) };
    const float wy[16] = { 0.f, -C1 * d1, 0.f, 0.f,
                           d2 * C2[0] * x, d2 * C2[1] * z, d2 * C2[2] * 2.f * -y, 0.f, d2 * C2[4] * 2.f * -y,
                           d3 * C3[0] * 3.f * (xx - yy), d3 * C3[1] * xz, d3 * C3[2] * (-3.f * yy + 4.f * zz - xx), d3 * C3[3] * -3.f * 2.f * yz,
                           d3 * C3[4] * -2.f * xy, d3 * C3[5] * -2.f * yz, d3 * C3[6] * -3.f * 2.f * xy };
    const float wz[16] = { 0.f, 0.f, C1 * d1, 0.f,
                           0.f, d2 * C2[1] * y, d2 * C2[2] * 2.f * 2.f * z, d2 * C2[3] * x, 0.f,
                           0.f, d3 * C3[1] * xy, d3 * C3[2] * 4.f * 2.f * yz, d3 * C3[3] * 3.f * (2.f * zz - xx - yy),
                           d3 * C3[4] * 4.f * 2.f * xz, d3 * C3[5] * (xx - yy), 0.f };
    constexpr unsigned NZX = 0xFFD8u, NZY = 0xFF72u, NZZ = 0x7CE4u;      // bit k: w?[k] is not identically zero
    float gx[3] = { 0, 0, 0 }, gy[3] = { 0, 0, 0 }, gz[3] = { 0, 0, 0 };
    const float dr[3] = { dRGB.x, dRGB.y, dRGB.z };
    const int rot = g % 12;
#pragma unroll
    for (int q = 0; q < 12; ++q) {
        int p = q + rot;
        p -= (p >= 12) ? 12 : 0;
        const float4 v = row4[p];
        const float f[4] = { v.x, v.y, v.z, v.w };
        float o[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int j = 4 * q + i, k = j / 3, c = j - 3 * k;
            if ((NZX >> k) & 1) gx[c] = fmaf(wx[k], f[i], gx[c]);
            if ((NZY >> k) & 1) gy[c] = fmaf(wy[k], f[i], gy[c]);
            if ((NZZ >> k) & 1) gz[c] = fmaf(wz[k], f[i], gz[c]);
            o[i] = basis[k] * dr[c];
        }
        row4[p] = make_float4(o[0], o[1], o[2], o[3]);
    }
    const V3 ddir = { gx[0] * dr[0] + gx[1] * dr[1] + gx[2] * dr[2], gy[0] * dr[0] + gy[1] * dr[1] + gy[2] * dr[2],
                      gz[0] * dr[0] + gz[1] * dr[1] + gz[2] * dr[2] };
    const V3 v = dir_orig;
    const float sum2 = dot(v, v);
    const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    V3 dmean;
    dmean.x = ((sum2 - v.x * v.x) * ddir.x - v.y * v.x * ddir.y - v.z * v.x * ddir.z) * invsum32;
    dmean.y = (-v.x * v.y * ddir.x + (sum2 - v.y * v.y) * ddir.y - v.z * v.y * ddir.z) * invsum32;
    dmean.z = (-v.x * v.z * ddir.x - v.y * v.z * ddir.y + (sum2 - v.z * v.z) * ddir.z) * invsum32;
    return dmean;
}

}  // namespace gsr
