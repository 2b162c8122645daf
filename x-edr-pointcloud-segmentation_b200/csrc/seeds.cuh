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

// One CTA of kSeedWarps warps per frame: stable counting sort of (count, index)
// by count.  Warp w owns the contiguous cell segment [w*seg, (w+1)*seg); the
// per-(warp, key) histogram scanned key-major / warp-minor gives every warp the
// start of its run inside each key's block, so the output order is "ascending
// key, ascending cell index within a key" (= std::multimap insertion order).
// Dynamic shared memory: kSeedWarps * nkeys ints, nkeys = nbhd*nbhd + 1.
constexpr int kSeedWarps = 32;

__global__ void __launch_bounds__(32 * kSeedWarps) seed_order_kernel(SeedArgs a, int nkeys) {
  extern __shared__ int s_hist[];  // [kSeedWarps][nkeys]
  __shared__ int s_keybase[kSeedMaxKeys];
  const unsigned FULL = 0xffffffffu;
  const int f = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int N = a.rows * a.cols;
  const int* counts = a.counts + (size_t)f * N;
  int* out = a.seeds + (size_t)f * a.stride;
  const int seg = ((N + kSeedWarps - 1) / kSeedWarps + 31) & ~31;
  const int c0 = min(warp * seg, N), c1 = min(c0 + seg, N);
  int* hist = s_hist + warp * nkeys;
  for (int k = threadIdx.x; k < kSeedWarps * nkeys; k += blockDim.x) s_hist[k] = 0;
  __syncthreads();
  for (int i = c0 + lane; i < c1; i += 32) {
    const int key = counts[i];
    if (key >= 0) atomicAdd(&hist[key], 1);
  }
  __syncthreads();
  // per-key totals, exclusive scan over keys by warp 0, then warp-minor offsets
  for (int k = threadIdx.x; k < nkeys; k += blockDim.x) {
    int t = 0;
    for (int w = 0; w < kSeedWarps; ++w) t += s_hist[w * nkeys + k];
    s_keybase[k] = t;
  }
  __syncthreads();
  if (warp == 0) {
    int carry = 0;
    for (int k0 = 0; k0 < nkeys; k0 += 32) {
      const int k = k0 + lane;
      const int v = (k < nkeys) ? s_keybase[k] : 0;
      int x = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(FULL, x, o);
        if (lane >= o) x += y;
      }
      if (k < nkeys) s_keybase[k] = carry + x - v;
      carry += __shfl_sync(FULL, x, 31);
    }
    if (lane == 0) a.n_seeds[f] = carry;
  }
  __syncthreads();
  for (int k = threadIdx.x; k < nkeys; k += blockDim.x) {
    int run = s_keybase[k];
    for (int w = 0; w < kSeedWarps; ++w) {
      const int c = s_hist[w * nkeys + k];
      s_hist[w * nkeys + k] = run;
      run += c;
    }
  }
  __syncthreads();
  for (int i0 = c0; i0 < c1; i0 += 32) {
    const int i = i0 + lane;
    const int key = (i < c1) ? counts[i] : -1;
    const unsigned peers = __match_any_sync(FULL, key);
    if (key >= 0) {
      const int leader = __ffs(peers) - 1;
      const int rank = __popc(peers & ((1u << lane) - 1u));
      int basepos = 0;
      if (lane == leader) {
        basepos = hist[key];
        hist[key] = basepos + __popc(peers);
      }
      basepos = __shfl_sync(peers, basepos, leader);
      out[basepos + rank] = i;  // seed index = col * Rows + row = linear cell index
    }
    __syncwarp();
  }
}

}  // namespace pcseg
