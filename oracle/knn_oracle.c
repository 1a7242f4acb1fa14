/* knn_oracle.c -- TEST INFRASTRUCTURE (checker): plain-C CPU restatement of the reference simple-knn
 * (distCUDA2). Never linked into, imported by or called from scorp_b200/.
 *
 * Follows /root/reference/submodules/simple-knn/simple_knn.cu:
 *   :38-71 (prepMorton, coord2Morton), :79-118 (boxMinMax), :120-130 (distBoxPoint),
 *   :132-146 (updateKBest<3>), :148-184 (boxMeanDist), :186-221 (SimpleKNN::knn; note the reductions
 *   start from init = {0,0,0}, so the Morton frame always contains the origin).
 * Parity status: PINNED by tests/golden/k*.npz (outputs of the reference CUDA build on a B200); the
 * search is exact, so a brute-force 3-NN (knn_oracle_brute) must and does give the same distances.
 */
#include "oracle_common.h"
#include <float.h>

#define BOX_SIZE 1024       /* simple_knn.cu:23 */

static uint32_t prepMorton(uint32_t x) {
    x = (x | (x << 16)) & 0x030000FF;
    x = (x | (x << 8)) & 0x0300F00F;
    x = (x | (x << 4)) & 0x030C30C3;
    x = (x | (x << 2)) & 0x09249249;
    return x;
}
static uint32_t coord2Morton(const float* c, const float* minn, const float* maxx) {
    uint32_t x = prepMorton((uint32_t)(((c[0] - minn[0]) / (maxx[0] - minn[0])) * ((1 << 10) - 1)));
    uint32_t y = prepMorton((uint32_t)(((c[1] - minn[1]) / (maxx[1] - minn[1])) * ((1 << 10) - 1)));
    uint32_t z = prepMorton((uint32_t)(((c[2] - minn[2]) / (maxx[2] - minn[2])) * ((1 << 10) - 1)));
    return x | (y << 1) | (z << 2);
}
static float distBoxPoint(const float* bmin, const float* bmax, const float* p) {
    float diff[3] = { 0, 0, 0 };
    for (int k = 0; k < 3; ++k)
        if (p[k] < bmin[k] || p[k] > bmax[k]) diff[k] = fminf(fabsf(p[k] - bmin[k]), fabsf(p[k] - bmax[k]));
    return diff[0] * diff[0] + diff[1] * diff[1] + diff[2] * diff[2];
}
static void updateKBest3(const float* ref, const float* point, float* knn) {
    float d[3] = { point[0] - ref[0], point[1] - ref[1], point[2] - ref[2] };
    float dist = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
    for (int j = 0; j < 3; ++j)
        if (knn[j] > dist) { float t = knn[j]; knn[j] = dist; dist = t; }
}

/* SimpleKNN::knn */
void knn_oracle(int P, const float* points, float* meanDists) {
    if (P <= 0) return;
    float minn[3] = { 0, 0, 0 }, maxx[3] = { 0, 0, 0 };
    for (int i = 0; i < P; ++i) for (int k = 0; k < 3; ++k) {
        minn[k] = fminf(minn[k], points[3 * i + k]); maxx[k] = fmaxf(maxx[k], points[3 * i + k]);
    }
    uint64_t* keys = malloc((size_t)P * 8);
    uint32_t* indices = malloc((size_t)P * 4);
    for (int i = 0; i < P; ++i) { keys[i] = coord2Morton(points + 3 * i, minn, maxx); indices[i] = (uint32_t)i; }
    sort_pairs_stable(keys, indices, (size_t)P, 32);
    const int num_boxes = (P + BOX_SIZE - 1) / BOX_SIZE;
    float* boxes = malloc((size_t)num_boxes * 24);
    for (int b = 0; b < num_boxes; ++b) {
        float* bx = boxes + 6 * (size_t)b;
        for (int k = 0; k < 3; ++k) { bx[k] = FLT_MAX; bx[3 + k] = -FLT_MAX; }
        for (int i = b * BOX_SIZE; i < P && i < (b + 1) * BOX_SIZE; ++i) for (int k = 0; k < 3; ++k) {
            bx[k] = fminf(bx[k], points[3 * indices[i] + k]); bx[3 + k] = fmaxf(bx[3 + k], points[3 * indices[i] + k]);
        }
    }
#pragma omp parallel for schedule(dynamic, 256)
    for (int idx = 0; idx < P; ++idx) {
        const float* point = points + 3 * (size_t)indices[idx];
        float best[3] = { FLT_MAX, FLT_MAX, FLT_MAX };
        const int lo = idx - 3 > 0 ? idx - 3 : 0, hi = idx + 3 < P - 1 ? idx + 3 : P - 1;
        for (int i = lo; i <= hi; ++i) { if (i == idx) continue; updateKBest3(point, points + 3 * (size_t)indices[i], best); }
        const float reject = best[2];
        best[0] = best[1] = best[2] = FLT_MAX;
        for (int b = 0; b < num_boxes; ++b) {
            const float dist = distBoxPoint(boxes + 6 * (size_t)b, boxes + 6 * (size_t)b + 3, point);
            if (dist > reject || dist > best[2]) continue;
            for (int i = b * BOX_SIZE; i < P && i < (b + 1) * BOX_SIZE; ++i) {
                if (i == idx) continue;
                updateKBest3(point, points + 3 * (size_t)indices[i], best);
            }
        }
        meanDists[indices[idx]] = (best[0] + best[1] + best[2]) / 3.0f;
    }
    free(keys); free(indices); free(boxes);
}

/* exhaustive 3-NN with the same distance expression (cross-check of the search's exactness) */
void knn_oracle_brute(int P, const float* points, float* meanDists) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int i = 0; i < P; ++i) {
        float best[3] = { FLT_MAX, FLT_MAX, FLT_MAX };
        for (int j = 0; j < P; ++j) { if (j == i) continue; updateKBest3(points + 3 * (size_t)i, points + 3 * (size_t)j, best); }
        meanDists[i] = (best[0] + best[1] + best[2]) / 3.0f;
    }
}
