// normals.cuh -- ComputeNormalsOrganized on the GPU.
//
// Replaces ps/algorithms.h:106-257 (detail::FindNormalSupportNeighbors) and
// ps/algorithms.h:329-375 (ComputeNormalsOrganized) of the reference.
//
// One thread per point; a warp covers 32 consecutive ROWS of one column, which
// is the contiguous direction of the reference's column-major layout, so the
// centre loads and the normal stores are coalesced.  Support neighbours are
// found with the reference's unbounded axis walks; the <= 9 support points are
// accumulated sequentially in the reference's order (centre, up, left, down,
// right, upper-left, lower-left, upper-right, lower-right) and the plane comes
// from the restated Eigen computeDirect, so normals are bit-identical to the
// CPU path (up to the libm trig policy).
//
// Quirk Q3: the reference reuses one PlaneEstimator and
// does not Clear() it after an invalid plane, so the next supported point
// accumulates on top of stale sums.  The parallel pass records every point
// whose standalone plane is invalid; normals_fixup_kernel then replays those
// (rare) chains sequentially in the reference's scan order.
#pragma once
#include "pcseg_math.cuh"

namespace pcseg {

struct NormalsArgs {
  const float* xyz;     // [n_frames][N][3]
  float* normals;       // [n_frames][N][3]
  float* curvature;     // nullable, [n_frames][N]
  int rows, cols, n_frames;
  int r0, r1, c0, c1;   // [r0,r1) x [c0,c1)
  float min2, max2;     // squared distances (float products, as the reference)
  int diag;
  int min_support;
  float sx, sy, sz;     // sensor translation cast to float (ps/algorithms.h:355)
  int* inv_list;        // [n_frames][inv_cap]
  int* inv_count;       // [n_frames]
  int inv_cap;
};

__device__ __forceinline__ bool in_range(V3 c, V3 p, float min2, float max2) {
  if (!vfinite(p)) return false;
  float d2 = vsqn(vsub(c, p));
  return d2 >= min2 && d2 <= max2;
}

// One axis walk of FindNormalSupportNeighbors (ps/algorithms.h:137-205): the first k in
// [0, count) whose point q[k * stride3 .. +2] is finite with min2 <= |c - p|^2 <= max2,
// or -1.  The walk dominates the kernel (about 25 steps per direction at the default
// min_neighbor_distance on a 640x480 depth frame), so it is written for instruction count
// and memory-level parallelism: four candidates are loaded and evaluated per trip (the loads
// of a trip do not depend on each other) and the first eligible one is taken, which is
// exactly the reference's sequential first match.
// CHK = false drops the explicit allFinite() test: a non-finite candidate gives d2 = NaN or
// +Inf, which fails `d2 >= min2 && d2 <= max2` whenever max2 is finite (the caller
// instantiates CHK = true otherwise), so the outcome is identical.
template <bool CHK>
__device__ __forceinline__ bool walk_hit(const float* __restrict__ q, V3 c, float min2, float max2) {
  const V3 p{__ldg(q), __ldg(q + 1), __ldg(q + 2)};
  if (CHK && !vfinite(p)) return false;
  const float d2 = vsqn(vsub(c, p));
  return d2 >= min2 && d2 <= max2;
}
template <bool CHK>
__device__ __forceinline__ int walk_first(const float* __restrict__ q, long stride3, int count, V3 c,
                                          float min2, float max2) {
  int k = 0;
  for (; k + 4 <= count; k += 4, q += 4 * stride3) {
    const bool e0 = walk_hit<CHK>(q, c, min2, max2);
    const bool e1 = walk_hit<CHK>(q + stride3, c, min2, max2);
    const bool e2 = walk_hit<CHK>(q + 2 * stride3, c, min2, max2);
    const bool e3 = walk_hit<CHK>(q + 3 * stride3, c, min2, max2);
    if (e0 | e1 | e2 | e3) return k + (e0 ? 0 : e1 ? 1 : e2 ? 2 : 3);
  }
  for (; k < count; ++k, q += stride3)
    if (walk_hit<CHK>(q, c, min2, max2)) return k;
  return -1;
}

// Returns the number of support indices (centre included), fills idx[<=9].
template <bool CHK>
__device__ __forceinline__ int find_support(const float* __restrict__ xyz, int rows, int cols,
                                            int row, int col, V3 center, float min2, float max2,
                                            bool diag, int idx[9]) {
  int n = 0;
  int min_row = max(row - 1, 0);
  int max_row = min(row + 1, rows - 1);
  int min_col = max(col - 1, 0);
  int max_col = min(col + 1, cols - 1);
  const int ci = col * rows + row;
  const float* qc = xyz + 3 * (size_t)ci;
  idx[n++] = ci;
  // upper neighbour: index -= 1
  {
    const int k = walk_first<CHK>(qc - 3, -3, row, center, min2, max2);
    if (k >= 0) { idx[n++] = ci - 1 - k; min_row = row - 1 - k; }
  }
  // left neighbour: index -= rows
  {
    const int k = walk_first<CHK>(qc - 3 * (long)rows, -3 * (long)rows, col, center, min2, max2);
    if (k >= 0) { idx[n++] = ci - (1 + k) * rows; min_col = col - 1 - k; }
  }
  // lower neighbour
  {
    const int k = walk_first<CHK>(qc + 3, 3, rows - 1 - row, center, min2, max2);
    if (k >= 0) { idx[n++] = ci + 1 + k; max_row = row + 1 + k; }
  }
  // right neighbour
  {
    const int k = walk_first<CHK>(qc + 3 * (long)rows, 3 * (long)rows, cols - 1 - col, center, min2, max2);
    if (k >= 0) { idx[n++] = ci + (1 + k) * rows; max_col = col + 1 + k; }
  }
  if (!diag) return n;
  if (min_col != col) {
    if (min_row != row) { int i = min_col * rows + min_row; if (in_range(center, ld3(xyz, i), min2, max2)) idx[n++] = i; }
    if (max_row != row) { int i = min_col * rows + max_row; if (in_range(center, ld3(xyz, i), min2, max2)) idx[n++] = i; }
  }
  if (max_col != col) {
    if (min_row != row) { int i = max_col * rows + min_row; if (in_range(center, ld3(xyz, i), min2, max2)) idx[n++] = i; }
    if (max_row != row) { int i = max_col * rows + max_row; if (in_range(center, ld3(xyz, i), min2, max2)) idx[n++] = i; }
  }
  return n;
}

__device__ __forceinline__ void store_nan3(float* normals, float* curvature, int i) {
  const float kNaN = __int_as_float(0x7fc00000);
  normals[3 * (size_t)i] = kNaN;
  normals[3 * (size_t)i + 1] = kNaN;
  normals[3 * (size_t)i + 2] = kNaN;
  if (curvature) curvature[i] = kNaN;
}

#ifndef NORMALS_MIN_BLOCKS
#define NORMALS_MIN_BLOCKS 6  // 40 registers: measured 10.0 -> 9.0 ms on the C2 batch (more warps to cover the walk loads)
#endif
#ifndef NORMALS_BLOCK
#define NORMALS_BLOCK 128  // small CTAs: one fits beside two resident planar-growth CTAs (9 K registers left)
#endif
constexpr int kNormalsBlock = NORMALS_BLOCK;
template <bool CHK>
__global__ void __launch_bounds__(kNormalsBlock, NORMALS_MIN_BLOCKS * 256 / kNormalsBlock) normals_kernel(NormalsArgs a) {
  const int nr = a.r1 - a.r0;
  const int nc = a.c1 - a.c0;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nr * nc) return;
  const int f = blockIdx.y;
  const size_t N = (size_t)a.rows * a.cols;
  const float* xyz = a.xyz + f * N * 3;
  float* normals = a.normals + f * N * 3;
  float* curvature = a.curvature ? a.curvature + f * N : nullptr;
  const int row = a.r0 + t % nr;
  const int col = a.c0 + t / nr;
  const int ci = col * a.rows + row;
  const V3 point = ld3(xyz, ci);
  int idx[9];
  int n = 0;
  if (vfinite(point))
    n = find_support<CHK>(xyz, a.rows, a.cols, row, col, point, a.min2, a.max2, a.diag != 0, idx);
  if (n == 0 || n < a.min_support) {
    store_nan3(normals, curvature, ci);
    return;
  }
  Estimator est;
  est_clear(est);
  est_set_orientation(est, vsub(V3{a.sx, a.sy, a.sz}, point));
  est_add(est, point);
  for (int k = 1; k < n; ++k) est_add(est, ld3(xyz, idx[k]));
  est_compute(est);
  if (!est.valid) {
    store_nan3(normals, curvature, ci);
    int slot = atomicAdd(a.inv_count + f, 1);
    if (slot < a.inv_cap) a.inv_list[(size_t)f * a.inv_cap + slot] = ci;
    return;
  }
  normals[3 * (size_t)ci] = est.plane.n.x;
  normals[3 * (size_t)ci + 1] = est.plane.n.y;
  normals[3 * (size_t)ci + 2] = est.plane.n.z;
  if (curvature) curvature[ci] = est.curvature;
}

// Sequential replay of the estimator-state leak (quirk Q3) from cell `start`
// in the reference's scan order (col outer, row inner, inside the range).
// stop_on_valid: return after the first valid plane (the chain ends with the
// Clear() at ps/algorithms.h:363).  Returns the linear index of the last cell
// the chain consumed.
__device__ int normals_chain(const NormalsArgs& a, const float* xyz, float* normals,
                             float* curvature, int start, bool stop_on_valid) {
  Estimator est;
  est_clear(est);
  int row = start % a.rows, col = start / a.rows;
  int last = start;
  const V3 sensor{a.sx, a.sy, a.sz};
  for (; col < a.c1; ++col, row = a.r0) {
    for (; row < a.r1; ++row) {
      const int ci = col * a.rows + row;
      last = ci;
      const V3 point = ld3(xyz, ci);
      int idx[9];
      int n = 0;
      if (vfinite(point))
        n = find_support<true>(xyz, a.rows, a.cols, row, col, point, a.min2, a.max2, a.diag != 0, idx);
      if (n == 0 || n < a.min_support) {
        store_nan3(normals, curvature, ci);
        continue;
      }
      est_set_orientation(est, vsub(sensor, point));
      for (int k = 0; k < n; ++k) est_add(est, ld3(xyz, idx[k]));
      est_compute(est);
      if (!est.valid) {
        store_nan3(normals, curvature, ci);
        continue;
      }
      normals[3 * (size_t)ci] = est.plane.n.x;
      normals[3 * (size_t)ci + 1] = est.plane.n.y;
      normals[3 * (size_t)ci + 2] = est.plane.n.z;
      if (curvature) curvature[ci] = est.curvature;
      est_clear(est);
      if (stop_on_valid) return ci;
    }
  }
  return last;
}

__global__ void __launch_bounds__(256) normals_fixup_kernel(NormalsArgs a) {
  const int f = blockIdx.x;
  const int count = a.inv_count[f];
  if (count == 0) return;
  const size_t N = (size_t)a.rows * a.cols;
  const float* xyz = a.xyz + f * N * 3;
  float* normals = a.normals + f * N * 3;
  float* curvature = a.curvature ? a.curvature + f * N : nullptr;
  __shared__ int s_red[256];
  __shared__ int s_pos;
  if (count > a.inv_cap) {
    // List overflowed: replay the whole range sequentially (exact, slow).
    if (threadIdx.x == 0)
      normals_chain(a, xyz, normals, curvature, a.c0 * a.rows + a.r0, false);
    return;
  }
  const int* list = a.inv_list + (size_t)f * a.inv_cap;
  int pos = -1;
  while (true) {
    int best = 0x7fffffff;
    for (int i = threadIdx.x; i < count; i += blockDim.x) {
      int v = list[i];
      if (v > pos && v < best) best = v;
    }
    s_red[threadIdx.x] = best;
    __syncthreads();
    for (int s = blockDim.x / 2; s > 0; s >>= 1) {
      if (threadIdx.x < s) s_red[threadIdx.x] = min(s_red[threadIdx.x], s_red[threadIdx.x + s]);
      __syncthreads();
    }
    best = s_red[0];
    __syncthreads();
    if (best == 0x7fffffff) break;
    if (threadIdx.x == 0) s_pos = normals_chain(a, xyz, normals, curvature, best, true);
    __syncthreads();
    pos = s_pos;
    __syncthreads();
  }
}

}  // namespace pcseg
