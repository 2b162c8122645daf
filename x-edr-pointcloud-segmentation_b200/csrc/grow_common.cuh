// grow_common.cuh -- shared pieces of the region-growing kernels.
//
// The reference grows regions with ONE FIFO queue, sequentially
// (ps/segmentation.h:239-292).  Labels must be bit-exact, so the kernels
// emulate the FIFO deterministically: a chunk Q[head, head+m) of the queue is
// evaluated in parallel and all side effects are resolved BY QUEUE POSITION:
//   * an entry that is rejected at chunk position q writes a transient code
//     (kRejBase - q) so that accepted entries at positions p > q see the cell
//     as kUnlabeled again (it was reset at "time" q) while entries p < q still
//     see it as "in the queue" (kAlreadyInUnexaminedPointsQueue);
//   * a neighbour cell n is enqueued by the SMALLEST (position, slot) key among
//     the accepted entries that see it unlabeled (atomicMin on claim[n]); the
//     winners are appended to the queue in key order (ordered compaction).
#pragma once
#include <climits>
#include <cstdint>
#include "pcseg_math.cuh"

namespace pcseg {

constexpr int kUnlabeled = -1;                      // ps/segmentation.h:36
constexpr int kAlreadyInQueue = -3;                 // ps/segmentation.h:41
constexpr int kAlreadyExamedPoint = -4;             // ps/segmentation.h:43
constexpr int kRejBase = -1000;                     // transient, never escapes a kernel

constexpr int kErrRegionCap = 1;    // more accepted regions than meta slots
constexpr int kErrBoundaryCap = 2;  // boundary buffer overflow
constexpr int kErrTraceCheck = 4;   // reference CHECK_GE(new_direction_index, 0) would abort
constexpr int kErrWatchdog = 8;     // a growth loop exceeded its iteration bound (internal error)

struct FrameStats {
  int n_planar, n_cluster, planar_inliers, cluster_inliers;
  int boundary_len, error_flags, grow_steps, refits;
};

struct PlanarMeta {
  int inlier_begin, inlier_count;
  int boundary_begin, boundary_count;
  int seed_index, label_id;
  float plane[4];
  float centroid[3];
  float area;
  float est[10];
};

struct ClusterMeta {
  int inlier_begin, inlier_count, label_id;
};

// Exclusive block scan of one int per thread (all threads must call).
template <int BLOCK>
__device__ __forceinline__ int block_scan_excl(int v, int* total, int* s_warp /*[BLOCK/32 + 1]*/) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) s_warp[warp] = x;
  __syncthreads();
  if (warp == 0) {
    int w = (lane < BLOCK / 32) ? s_warp[lane] : 0;
    int ws = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int y = __shfl_up_sync(0xffffffffu, ws, o);
      if (lane >= o) ws += y;
    }
    if (lane < BLOCK / 32) s_warp[lane] = ws - w;  // exclusive warp offsets
    if (lane == BLOCK / 32 - 1) s_warp[BLOCK / 32] = ws;
  }
  __syncthreads();
  int res = s_warp[warp] + x - v;
  *total = s_warp[BLOCK / 32];
  __syncthreads();
  return res;
}

// Block-wide max of one int per thread (all threads must call).
template <int BLOCK>
__device__ __forceinline__ int block_max(int v, int* s_warp) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
  if (lane == 0) s_warp[warp] = v;
  __syncthreads();
  int r = s_warp[0];
#pragma unroll
  for (int i = 1; i < BLOCK / 32; ++i) r = max(r, s_warp[i]);
  __syncthreads();
  return r;
}

// cell -> (row, col) for cell < 2^24 without an integer division: float
// reciprocal estimate corrected by one step (exact for the stated range).
struct FastDiv {
  int d;
  float inv;
};
__host__ __device__ inline FastDiv make_fastdiv(int d) { return FastDiv{d, 1.0f / (float)d}; }
__device__ __forceinline__ void fast_divmod(const FastDiv& fd, int x, int* q, int* r) {
  int qq = __float2int_rz(((float)x + 0.5f) * fd.inv);
  int rr = x - qq * fd.d;
  if (rr < 0) { --qq; rr += fd.d; }
  else if (rr >= fd.d) { ++qq; rr -= fd.d; }
  *q = qq;
  *r = rr;
}

// Seed list access: seeds == nullptr means "all indices descending"
// (seeds[k] = N-1-k, ps/clustering_test.cc:57-59).
__device__ __forceinline__ int seed_at(const int* seeds, int N, int k) {
  return seeds ? seeds[k] : (N - 1 - k);
}

}  // namespace pcseg
