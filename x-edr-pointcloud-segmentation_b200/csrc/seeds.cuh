// seeds.cuh -- the three seed finders on the GPU: FindSeedPointsFromPlaneSupport (below),
// FindSeedPointsFromAverageNormals and FindSeedPointsFromLastPlanarRegions (end of file).
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
  const int tt = blockIdx.x * blockDim.x + threadIdx.x;
  if (tt >= N) return;
  const int f = blockIdx.y;
  const float* xyz = a.xyz + (size_t)f * N * 3;
  const float* nrm = a.normals + (size_t)f * N * 3;
  const int rows = a.rows, cols = a.cols;
  // loop variables of the reference; a warp takes 32 consecutive `col` of one `row`: the elements
  // the reference actually reads (quirk Q1: data[support_row * Rows + support_col]) are then
  // consecutive in memory across the warp
  const int row = tt / cols, col = tt % cols;
  const int t = col * rows + row;
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


// ---------------------------------------------------------------------------
// FindSeedPointsFromAverageNormals, ps/segmentation.h:137-184 with
// detail::GetSquaredAverageNormalLengthAndUpdateSlidingWindow (:79-120).
// Quirk Q1 literally: normals.AtUnsafe(col - n, row + i - k) reads
// data[(row + i - k) * Rows + (col - n)], the seed index is
// LinearizeIndex(row, col) - k = col * Rows + row - k.
// The sliding row sums are order-sensitive fp32 (subtract the leaving normal,
// add the entering one, column after column), so one thread walks one `row` of
// the reference's outer loop through all columns with its n running sums; rows
// are independent.  Flags go to a (row, col)-ordered byte grid; the seeds are
// emitted in the reference's loop order (row outer, col inner) by an ordered
// compaction.
// ---------------------------------------------------------------------------
constexpr int kAvgMaxNbhd = 32;

struct AvgSeedArgs {
  const float* normals;  // [F][N][3]
  int rows, cols;
  int nbhd, half, min_valid;
  float sq_min_len;      // min_avg_normal_length^2
  unsigned char* flags;  // [F][flag_stride]: (row - half) * cols + col
  size_t flag_stride;
  int* seeds;            // [F][stride]
  int stride;
  int* n_seeds;          // [F]
};

__global__ void __launch_bounds__(64) seed_avg_normals_flag_kernel(AvgSeedArgs a) {
  const int f = blockIdx.y;
  const int rr = blockIdx.x * blockDim.x + threadIdx.x;  // row - half
  const int rows = a.rows, cols = a.cols, n = a.nbhd, half = a.half;
  const int row = rr + half;
  if (row >= rows - half) return;
  const float* nrm = a.normals + (size_t)f * rows * cols * 3;
  unsigned char* flags = a.flags + (size_t)f * a.flag_stride + (size_t)rr * cols;
  V3 avg_row[kAvgMaxNbhd];
  int nvalid_row[kAvgMaxNbhd];
  for (int i = 0; i < n; ++i) { avg_row[i] = V3{0.0f, 0.0f, 0.0f}; nvalid_row[i] = 0; }
  for (int col = 0; col < cols; ++col) {
    V3 avg{0.0f, 0.0f, 0.0f};
    int num_valid = 0;
    for (int i = 0; i < n; ++i) {
      const int q = (row + i - half) * rows;  // Q1: AtUnsafe(col', row + i - k) -> data[(row+i-k)*Rows + col']
      if (col >= n) {
        const V3 leaving = ld3(nrm, q + col - n);
        if (vfinite(leaving)) { avg_row[i] = vsub(avg_row[i], leaving); --nvalid_row[i]; }
      }
      const V3 entering = ld3(nrm, q + col);
      if (vfinite(entering)) { avg_row[i] = vadd(avg_row[i], entering); ++nvalid_row[i]; }
      const int new_num = num_valid + nvalid_row[i];
      // (int * Vector3f + int * Vector3f) / int, the ints converted to float (:113-115)
      const float wa = (float)num_valid, wb = (float)nvalid_row[i];
      avg = vdiv(vadd(vscale(wa, avg), vscale(wb, avg_row[i])), (float)new_num);
      num_valid = new_num;
    }
    const float sq = vsqn(avg);
    flags[col] = (col >= n - 1 && num_valid >= a.min_valid && sq >= a.sq_min_len) ? 1 : 0;
  }
}

// One CTA per frame: seeds in (row outer, col inner) order.
__global__ void __launch_bounds__(1024) seed_avg_normals_order_kernel(AvgSeedArgs a) {
  __shared__ int s_w[33];
  const int f = blockIdx.x;
  const int rows = a.rows, cols = a.cols, half = a.half;
  const int total = (rows - 2 * half) * cols;
  const unsigned char* flags = a.flags + (size_t)f * a.flag_stride;
  int* out = a.seeds + (size_t)f * a.stride;
  int written = 0;
  for (int c0 = 0; c0 < total; c0 += 1024) {
    const int k = c0 + threadIdx.x;
    const bool keep = k < total && flags[k] != 0;
    int tot;
    const int off = block_scan_excl<1024>(keep ? 1 : 0, &tot, s_w);
    if (keep) {
      const int row = k / cols + half, col = k % cols;
      out[written + off] = col * rows + row - half;  // LinearizeIndex(row, col) - kHalfNeighborhoodSize
    }
    written += tot;
  }
  if (threadIdx.x == 0) a.n_seeds[f] = written;
}

// ---------------------------------------------------------------------------
// FindSeedPointsFromLastPlanarRegions, ps/planar_region.h:478-519: per previous
// region (already in multimap order) the nearest current point whose normal
// agrees with the transformed region normal; strict "<" keeps the first
// minimum in index order, so the reduction key is (distance, index).  One CTA
// per previous region; the transformed seed point / normal
// (current_pose_prev * centroid, quaternion * normal) are computed on the device.
// ---------------------------------------------------------------------------
struct LastSeedArgs {
  const float* xyz;       // [n_points][3]
  const float* normals;   // [n_points][3]
  int n_points;
  const float* centroids; // [n_prev][3]
  const float* pnormals;  // [n_prev][3]
  float qx, qy, qz, qw, tx, ty, tz;
  float max_sq, cos_max;
  int* out;               // [n_prev]: index or -1
};

__global__ void __launch_bounds__(512) seed_last_regions_kernel(LastSeedArgs a) {
  __shared__ unsigned long long s_best[16];
  const int r = blockIdx.x;
  const V3 sp = vadd(qrotate(a.qx, a.qy, a.qz, a.qw, ld3(a.centroids, r)), V3{a.tx, a.ty, a.tz});
  const V3 sn = qrotate(a.qx, a.qy, a.qz, a.qw, ld3(a.pnormals, r));
  // squared distances are >= 0, so their bit patterns order like the values; key = (bits, index)
  unsigned long long best = ~0ull;
  for (int j = threadIdx.x; j < a.n_points; j += blockDim.x) {
    const float sq = vsqn(vsub(sp, ld3(a.xyz, j)));
    if (sq < a.max_sq && vdot(sn, ld3(a.normals, j)) > a.cos_max) {
      const unsigned long long key = ((unsigned long long)__float_as_uint(sq + 0.0f) << 32) | (unsigned)j;
      best = key < best ? key : best;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long v = __shfl_xor_sync(0xffffffffu, best, o);
    best = v < best ? v : best;
  }
  if ((threadIdx.x & 31) == 0) s_best[threadIdx.x >> 5] = best;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 16; ++w) best = s_best[w] < best ? s_best[w] : best;
    a.out[r] = best == ~0ull ? -1 : (int)(unsigned)best;
  }
}

}  // namespace pcseg
