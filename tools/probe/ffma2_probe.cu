// FFMA (3-register form, outer-product pattern) vs packed FFMA2 (fma.rn.f32x2) issue rate on sm_100a.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_ffma(float* out, const float* in, int iters) {
  float acc[8][8], x[8], w[8];
  for (int i = 0; i < 8; ++i) { x[i] = in[threadIdx.x + i]; w[i] = in[threadIdx.x + 8 + i]; }
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 8; ++j) acc[i][j] = i + j;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(x[i], w[j], acc[i][j]);
#pragma unroll
    for (int i = 0; i < 8; ++i) { x[i] += 1e-9f; }
  }
  float s = 0; for (int i = 0; i < 8; ++i) for (int j = 0; j < 8; ++j) s += acc[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__device__ __forceinline__ unsigned long long pk(float a, float b) { float2 v = make_float2(a, b); return *reinterpret_cast<unsigned long long*>(&v); }
__global__ void k_ffma2(float* out, const float* in, int iters) {
  unsigned long long acc[8][4], xx[8], w2[4];
  float x[8];
  for (int i = 0; i < 8; ++i) x[i] = in[threadIdx.x + i];
  for (int j = 0; j < 4; ++j) w2[j] = pk(in[threadIdx.x + 8 + 2 * j], in[threadIdx.x + 9 + 2 * j]);
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) acc[i][j] = pk(i + 2 * j, i + 2 * j + 1);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) xx[i] = pk(x[i], x[i]);
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc[i][j]) : "l"(xx[i]), "l"(w2[j]));
#pragma unroll
    for (int i = 0; i < 8; ++i) { x[i] += 1e-9f; }
  }
  float s = 0; for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) { float2 v = *reinterpret_cast<float2*>(&acc[i][j]); s += v.x + v.y; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
  float *out, *in; cudaMalloc(&out, 148 * 8 * 1024 * 4); cudaMalloc(&in, 4096 * 4); cudaMemset(in, 0, 4096 * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  int iters = 5000;
  for (int thr = 128; thr <= 512; thr *= 2)
  for (int mode = 0; mode < 2; ++mode) for (int rep = 0; rep < 2; ++rep) {
    cudaEventRecord(e0);
    if (mode == 0) k_ffma<<<148 * 2, thr>>>(out, in, iters);
    else k_ffma2<<<148 * 2, thr>>>(out, in, iters);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double fma = 148.0 * 2 * thr * 64 * iters;
    printf("thr %d %s: %.3f ms  %.1f TFLOP/s  (%.1f FMA/clk/SM @1.9GHz)\n", thr, mode ? "ffma2" : "ffma ", ms, 2 * fma / ms / 1e9, fma / (ms * 1e-3) / 148 / 1.9e9);
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}
