// seeds.cuh -- FindSeedPointsFromPlaneSupport on the GPU.
//
// Replaces ps/segmentation.h:190-230 with detail::GetNumPlaneSupportPoints
// (:54-73).  Quirk Q1 is reproduced literally: the reference passes (col,row)
// to AtUnsafe(row,col), so the element it reads for loop variables (col,row) is
// data[row * Rows + col]; the seed index it emits is col * Rows + row.  The
// std::multimap<int,int> ordering (ascending support count, insertion order
// within a count) is a stable counting sort, done by one warp per frame.
#pragma once
#include "grow_common.cuh"

namespace pcseg {

struct SeedArgs {
  const float* xyz;      // [F][N][3]
  const float* normals;  // [F][N][3]
  int rows, cols;
  int half;              // neighborhood_size / 2
  int nbhd;              // neighborhood_size
  float max_plane_distance;
  int min_support;
  int* counts;           // [F][N] scratch: support count or -1
  int* seeds;            // [F][stride]
  int stride;
  int* n_seeds;          // [F]
};

constexpr int kSeedMaxKeys = 1024;

__global__ void __launch_bounds__(256) seed_support_count_kernel(SeedArgs a) {
  const int N = a.rows * a.cols;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= N) return;
  const int f = blockIdx.y;
  const float* xyz = a.xyz + (size_t)f * N * 3;
  const float* nrm = a.normals + (size_t)f * N * 3;
  const int rows = a.rows, cols = a.cols;
  const int row = t % rows, col = t / rows;   // loop variables of the reference
  int result = -1;
  if (rows >= a.nbhd && cols >= a.nbhd) {     // :202-204
    const int ti = row * rows + col;          // AtUnsafe(col, row) -> data[row*Rows + col]
    const V3 point = ld3(xyz, ti);
    const V3 normal = ld3(nrm, ti);
    if (vfinite(normal) && vfinite(point)) {
      const Plane pl = make_plane(normal, point);
      const int min_row = max(0, row - a.half), max_row = min(rows - 1, row + a.half);
      const int min_col = max(0, col - a.half), max_col = min(cols - 1, col + a.half);
      int num = 0;
      for (int sc = min_col; sc <= max_col; ++sc)
        for (int sr = min_row; sr <= max_row; ++sr)
          if (plane_abs_distance(pl, ld3(xyz, sr * rows + sc)) < a.max_plane_distance) ++num;
      if (num >= a.min_support) result = num;
    }
  }
  a.counts[(size_t)f * N + t] = result;
}

// One warp per frame: stable counting sort of (count, index) by count.
__global__ void __launch_bounds__(32) seed_order_kernel(SeedArgs a) {
  const unsigned FULL = 0xffffffffu;
  const int f = blockIdx.x, lane = threadIdx.x;
  const int N = a.rows * a.cols;
  const int* counts = a.counts + (size_t)f * N;
  int* out = a.seeds + (size_t)f * a.stride;
  __shared__ int s_run[kSeedMaxKeys];
  for (int k = lane; k < kSeedMaxKeys; k += 32) s_run[k] = 0;
  __syncwarp();
  for (int i = lane; i < N; i += 32) {
    const int key = counts[i];
    if (key >= 0) atomicAdd(&s_run[key], 1);
  }
  __syncwarp();
  if (lane == 0) {
    int acc = 0;
    for (int k = 0; k < kSeedMaxKeys; ++k) {
      const int c = s_run[k];
      s_run[k] = acc;
      acc += c;
    }
    a.n_seeds[f] = acc;
  }
  __syncwarp();
  for (int i0 = 0; i0 < N; i0 += 32) {
    const int i = i0 + lane;
    const int key = (i < N) ? counts[i] : -1;
    const unsigned peers = __match_any_sync(FULL, key);
    if (key >= 0) {
      const int leader = __ffs(peers) - 1;
      const int rank = __popc(peers & ((1u << lane) - 1u));
      int basepos = 0;
      if (lane == leader) {
        basepos = s_run[key];
        s_run[key] = basepos + __popc(peers);
      }
      basepos = __shfl_sync(peers, basepos, leader);
      out[basepos + rank] = i;  // seed index = col * Rows + row = linear cell index
    }
    __syncwarp();
  }
}

}  // namespace pcseg
