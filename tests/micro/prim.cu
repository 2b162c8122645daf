// Latency of synchronisation / shared-memory primitives on one warp (scratch tool).
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(long long* cyc, int iters, int* gout) {
  __shared__ volatile int sv[64];
  __shared__ int sn[64];
  __shared__ unsigned sh[256];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  long long t0, t1; int s = 0; int x = lane;
  sv[threadIdx.x & 63] = lane; sn[threadIdx.x & 63] = lane; sh[threadIdx.x] = 0xffffffffu;
  __syncthreads();
  if (warp == 0) {
    t0 = clock64(); for (int i = 0; i < iters; ++i) { x += (int)clock64(); } t1 = clock64(); cyc[s++] = (t1 - t0) / iters;           // 0 clock64
    t0 = clock64(); for (int i = 0; i < iters; ++i) { __threadfence_block(); x += i; } t1 = clock64(); cyc[s++] = (t1 - t0) / iters;  // 1 fence.block
    t0 = clock64(); for (int i = 0; i < iters; ++i) { x = sv[x & 31] + 1; } t1 = clock64(); cyc[s++] = (t1 - t0) / iters;             // 2 dependent volatile LDS
    t0 = clock64(); for (int i = 0; i < iters; ++i) { x = sn[x & 31] + 1; } t1 = clock64(); cyc[s++] = (t1 - t0) / iters;             // 3 dependent LDS
    t0 = clock64(); for (int i = 0; i < iters; ++i) { __syncwarp(); x += i; } t1 = clock64(); cyc[s++] = (t1 - t0) / iters;            // 4 syncwarp
    t0 = clock64(); for (int i = 0; i < iters; ++i) { x += __ballot_sync(0xffffffffu, (x + i) & 1); } t1 = clock64(); cyc[s++] = (t1 - t0) / iters;  // 5 ballot chain
    t0 = clock64(); for (int i = 0; i < iters; ++i) { x += atomicMin(&sh[(x + i) & 255], (unsigned)x); } t1 = clock64(); cyc[s++] = (t1 - t0) / iters; // 6 smem atomicMin w/ return
    t0 = clock64(); for (int i = 0; i < iters; ++i) { atomicMin(&sh[(lane + i) & 255], (unsigned)i); } t1 = clock64(); cyc[s++] = (t1 - t0) / iters; // 7 smem atomicMin no return (throughput)
    t0 = clock64(); for (int i = 0; i < iters; ++i) { sv[lane] = x; __threadfence_block(); x = sv[(lane + 1) & 31] + i; } t1 = clock64(); cyc[s++] = (t1 - t0) / iters; // 8 store+fence+load
    t0 = clock64(); for (int i = 0; i < iters; ++i) { if (lane == 0) { sv[32] = i; sv[33] = i; sv[34] = i; } __syncwarp(); x += sv[32]; } t1 = clock64(); cyc[s++] = (t1 - t0) / iters; // 9 divergent lane-0 store + syncwarp + load
    t0 = clock64(); for (int i = 0; i < iters; ++i) { x = __shfl_sync(0xffffffffu, x, (i + 1) & 31) + 1; } t1 = clock64(); cyc[s++] = (t1 - t0) / iters; // 10 shfl
    t0 = clock64(); for (int i = 0; i < iters; ++i) { x = __popc(x) + __ffs(x + i); } t1 = clock64(); cyc[s++] = (t1 - t0) / iters; // 11 popc+ffs
    t0 = clock64(); for (int i = 0; i < iters; ++i) { gout[lane] = x + i; x += 1; } t1 = clock64(); cyc[s++] = (t1 - t0) / iters; // 12 global store issue
    t0 = clock64(); for (int i = 0; i < iters; ++i) { gout[lane] = x + i; __threadfence_block(); x += 1; } t1 = clock64(); cyc[s++] = (t1 - t0) / iters; // 13 global store + fence.block
    t0 = clock64(); for (int i = 0; i < iters; ++i) { __nanosleep(40); x += 1; } t1 = clock64(); cyc[s++] = (t1 - t0) / iters; // 14 nanosleep(40)
  }
  __syncthreads();
  // 15: named barrier over 128 threads (warps 0-3)
  if (warp < 4) {
    t0 = clock64(); for (int i = 0; i < iters; ++i) { asm volatile("bar.sync 1, 128;" ::: "memory"); x += i; } t1 = clock64();
    if (threadIdx.x == 0) cyc[15] = (t1 - t0) / iters;
  }
  gout[threadIdx.x] = x;
}
int main() {
  long long* cyc; int* g; cudaMalloc(&cyc, 32 * 8); cudaMalloc(&g, 4096); cudaMemset(cyc, 0, 256);
  k<<<1, 256>>>(cyc, 2000, g); cudaDeviceSynchronize();
  k<<<1, 256>>>(cyc, 2000, g); cudaDeviceSynchronize();
  long long h[32]; cudaMemcpy(h, cyc, 256, cudaMemcpyDeviceToHost);
  const char* names[] = {"clock64", "fence.block", "dep volatile LDS", "dep LDS", "syncwarp", "ballot chain", "smem atomicMin ret", "smem atomicMin noret", "sts+fence+lds", "lane0 3xSTS+syncwarp+LDS", "shfl", "popc+ffs", "global store issue", "global store+fence.block", "nanosleep(40)", "bar.sync 128"};
  for (int i = 0; i < 16; ++i) printf("%-28s %lld cycles\n", names[i], h[i]);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
