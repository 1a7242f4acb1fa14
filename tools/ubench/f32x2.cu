// Microbenchmark: issue rate of scalar FFMA vs packed FFMA2/FMUL2/FADD2 on sm_100a (decides whether the blend
// kernels gain from two pixels per thread). Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 f32x2.cu -o f32x2
#include <cstdio>
#include <cuda_runtime.h>
constexpr int IT = 4096, U = 8;
__global__ void k_ffma(float* out, float a, float b) {
    float x[U];
    for (int i = 0; i < U; ++i) x[i] = threadIdx.x + i;
    for (int it = 0; it < IT; ++it)
#pragma unroll
        for (int i = 0; i < U; ++i) x[i] = __fmaf_rn(x[i], a, b);
    float s = 0; for (int i = 0; i < U; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_ffma2(float* out, float a, float b) {
    float2 x[U];
    const float2 a2 = make_float2(a, a + 1.f), b2 = make_float2(b, b + 1.f);
    for (int i = 0; i < U; ++i) x[i] = make_float2(threadIdx.x + i, threadIdx.x - i);
    for (int it = 0; it < IT; ++it)
#pragma unroll
        for (int i = 0; i < U; ++i) x[i] = __ffma2_rn(x[i], a2, b2);
    float s = 0; for (int i = 0; i < U; ++i) s += x[i].x + x[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_fmul2(float* out, float a, float b) {
    float2 x[U];
    const float2 a2 = make_float2(a, a + 1e-7f);
    for (int i = 0; i < U; ++i) x[i] = make_float2(threadIdx.x + i, threadIdx.x - i);
    for (int it = 0; it < IT; ++it)
#pragma unroll
        for (int i = 0; i < U; ++i) x[i] = __fmul2_rn(x[i], a2);
    float s = 0; for (int i = 0; i < U; ++i) s += x[i].x + x[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_mix(float* out, float a, float b) {     // FFMA + alu-pipe op (FMNMX) interleaved
    float x[U];
    for (int i = 0; i < U; ++i) x[i] = threadIdx.x + i;
    for (int it = 0; it < IT; ++it)
#pragma unroll
        for (int i = 0; i < U; ++i) { x[i] = __fmaf_rn(x[i], a, b); x[i] = fminf(x[i], 1e30f); }
    float s = 0; for (int i = 0; i < U; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <class K> float run(K k, float* out) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<<<148 * 8, 256>>>(out, 1.0000001f, 1e-9f);
    cudaEventRecord(e0);
    for (int i = 0; i < 5; ++i) k<<<148 * 8, 256>>>(out, 1.0000001f, 1e-9f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms / 5;
}
int main() {
    float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    const double inst = 148.0 * 8 * 8 * IT * U;            // warp instructions in the main loop
    float t;
    t = run(k_ffma, out);  printf("FFMA   %.3f ms  %.1f Gwarp-inst/s  %.1f TFLOP/s\n", t, inst / t / 1e6, inst * 64 / t / 1e9);
    t = run(k_ffma2, out); printf("FFMA2  %.3f ms  %.1f Gwarp-inst/s  %.1f TFLOP/s\n", t, inst / t / 1e6, inst * 128 / t / 1e9);
    t = run(k_fmul2, out); printf("FMUL2  %.3f ms  %.1f Gwarp-inst/s\n", t, inst / t / 1e6);
    t = run(k_mix, out);   printf("FFMA+FMNMX %.3f ms  %.1f Gwarp-inst/s (both)\n", t, 2 * inst / t / 1e6);
    return 0;
}
