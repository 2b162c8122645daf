// Latency microbenchmarks for the serial chain of the planar growth (scratch tool).
#include <cstdio>
#include "../../x-edr-pointcloud-segmentation_b200/csrc/pcseg_math.cuh"
using namespace pcseg;
__global__ void k(float* out, long long* cyc, int iters, float seedv) {
  float x = seedv; long long t0, t1; int s = 0;
  // 0: fdiv chain
  t0 = clock64(); for (int i = 0; i < iters; ++i) x = 1.0f / (x + 1.5f); t1 = clock64(); cyc[s++] = (t1 - t0) / iters;
  // 1: fsqrt chain
  t0 = clock64(); for (int i = 0; i < iters; ++i) x = sqrtf(x + 1.5f); t1 = clock64(); cyc[s++] = (t1 - t0) / iters;
  // 2: atan2 double->float chain
  t0 = clock64(); for (int i = 0; i < iters; ++i) x = cr_atan2f(x + 0.3f, 0.7f); t1 = clock64(); cyc[s++] = (t1 - t0) / iters;
  // 3: cos+sin separate
  t0 = clock64(); for (int i = 0; i < iters; ++i) x = cr_cosf(x) + cr_sinf(x) * 0.5f; t1 = clock64(); cyc[s++] = (t1 - t0) / iters;
  // 4: sincos double
  t0 = clock64(); for (int i = 0; i < iters; ++i) { double sv, cv; sincos((double)x, &sv, &cv); x = (float)cv + (float)sv * 0.5f; } t1 = clock64(); cyc[s++] = (t1 - t0) / iters;
  // 5: fadd chain
  t0 = clock64(); for (int i = 0; i < iters; ++i) x = x + 1.0001f; t1 = clock64(); cyc[s++] = (t1 - t0) / iters;
  // 6: dfma chain
  double d = x; t0 = clock64(); for (int i = 0; i < iters; ++i) d = d * 1.0000001 + 0.5; t1 = clock64(); cyc[s++] = (t1 - t0) / iters; x = (float)d * 1e-9f;
  // 7: est_compute chain
  Estimator e; est_clear(e);
  e.plane.n = V3{0.1f, 0.2f, 0.9f};
  t0 = clock64();
  for (int i = 0; i < iters; ++i) {
    e.acc[0] = 30.5f + x; e.acc[1] = 3.1f; e.acc[2] = 12.2f; e.acc[3] = 25.0f; e.acc[4] = 9.3f; e.acc[5] = 40.0f;
    e.csum = V3{10.f, 8.f, 30.f}; e.w = 30.f;
    est_compute(e);
    x = e.plane.d * 1e-6f;
  }
  t1 = clock64(); cyc[s++] = (t1 - t0) / iters;
  // 8: atan2f (fp32 CUDA) chain, 9: shfl chain, 10: ballot chain
  t0 = clock64(); for (int i = 0; i < iters; ++i) x = atan2f(x + 0.3f, 0.7f); t1 = clock64(); cyc[s++] = (t1 - t0) / iters;
  t0 = clock64(); for (int i = 0; i < iters; ++i) x = __shfl_sync(0xffffffffu, x, (i + 1) & 31) + 1.0f; t1 = clock64(); cyc[s++] = (t1 - t0) / iters;
  unsigned b = 0; t0 = clock64(); for (int i = 0; i < iters; ++i) b += __ballot_sync(0xffffffffu, (b + i) & 1); t1 = clock64(); cyc[s++] = (t1 - t0) / iters;
  __shared__ float sm[64]; sm[threadIdx.x] = x; __syncwarp();
  int idx = threadIdx.x; t0 = clock64(); for (int i = 0; i < iters; ++i) { idx = (int)sm[idx & 31] & 31; } t1 = clock64(); cyc[s++] = (t1 - t0) / iters;
  t0 = clock64(); for (int i = 0; i < iters; ++i) { __syncwarp(); } t1 = clock64(); cyc[s++] = (t1 - t0) / iters;
  out[threadIdx.x] = x + b + idx;
}
int main() {
  float* out; long long* cyc; cudaMalloc(&out, 128); cudaMalloc(&cyc, 16 * 8); cudaMemset(cyc, 0, 128);
  k<<<1, 32>>>(out, cyc, 2000, 0.37f); cudaDeviceSynchronize();
  k<<<1, 32>>>(out, cyc, 2000, 0.37f); cudaDeviceSynchronize();
  long long h[16]; cudaMemcpy(h, cyc, 128, cudaMemcpyDeviceToHost);
  const char* names[] = {"fdiv", "fsqrt", "atan2(double)", "cos+sin(double)", "sincos(double)", "fadd", "dfma", "est_compute", "atan2f", "shfl+add", "ballot+add", "lds-dep", "syncwarp"};
  for (int i = 0; i < 13; ++i) printf("%-18s %lld cycles\n", names[i], h[i]);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
