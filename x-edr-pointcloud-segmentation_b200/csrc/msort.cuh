// msort.cuh -- the sort of SlidingMeanShift on the device (ps/mean_shift_segmentation.h:262-280).
//
// The reference sorts its (seed, intensity) records with std::sort, comparing intensities only.
// Intensities are window counts (integers 50..121, quirk Q7), so the tie groups are huge and the
// permutation is whatever the library's unstable algorithm leaves.  Region ids, inlier orders and
// the suppression of later modes depend on it, so the product must reproduce it exactly.
// For libstdc++ (the library of this build; GCC 13 bits/stl_algo.h:1918-1951) the algorithm is
//   __introsort_loop: while (size > 16) { median-of-3 of (first+1, mid, last-1) swapped to first;
//                     Hoare partition of (first, last) around *first (__unguarded_partition);
//                     recurse on [cut, last); continue with [first, cut) }
//   __final_insertion_sort: an insertion sort, i.e. THE stable sort of what the loop left.
// The partition is data-parallel once it is stated by ranks: with L_k the k-th element >= pivot
// from the left and R_k the k-th element <= pivot from the right (inside the partitioned range),
// the loop swaps L_k <-> R_k for k = 1 .. S, S = #{k : L_k < R_k}, and returns
// cut = min(L_(S+1), R_S) (the two scan pointers stop on each other's last swap otherwise).
// Sub-ranges are independent, so every recursion LEVEL is one pass over the array: pivots, two
// prefix counts, rank -> position tables, swaps, children.  ~log2(n/16) levels, then a stable
// counting sort on the 8-bit keys.  The heapsort fall-back of introsort (depth 2*log2 n) cannot be
// replayed this way; it is detected and reported (it needs adversarial input: Hoare partitions of
// tie-heavy data are balanced).  tests/test_msort.py checks the permutation against std::sort.
#pragma once
#include "grow_common.cuh"

namespace pcseg {

struct SortSeg {
  int lo, hi;         // range [lo, hi)
  int pivot;          // key at lo after the median swap
  int baseL, baseR;   // global prefix counts before the range
  int totalR;
  int swaps;          // S
  int cut;
  int child_left, child_right;  // ids in the next level's table, -1 = done (size <= 16)
};

constexpr int kSortThreshold = 16;   // _S_threshold
constexpr int kScanBlock = 1024;
constexpr int kScanItems = 4;        // elements per thread
constexpr int kScanTile = kScanBlock * kScanItems;

// ---- device-wide inclusive scan of 64-bit values (two packed 32-bit counters) ----
__device__ __forceinline__ unsigned long long warp_incl_scan64(unsigned long long x, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned long long y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  return x;
}

// pass 1: per-tile totals
__global__ void __launch_bounds__(kScanBlock) scan64_reduce_kernel(const unsigned long long* __restrict__ v, int n,
                                                                    unsigned long long* tile_sums) {
  __shared__ unsigned long long s_w[32];
  const int base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
  unsigned long long t = 0;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k)
    if (base + k < n) t += v[base + k];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  t = warp_incl_scan64(t, lane);
  if (lane == 31) s_w[warp] = t;
  __syncthreads();
  if (warp == 0) {
    unsigned long long w = s_w[lane];
    w = warp_incl_scan64(w, lane);
    if (lane == 31) tile_sums[blockIdx.x] = w;
  }
}
// pass 2: one CTA turns the tile totals into exclusive tile offsets
__global__ void __launch_bounds__(kScanBlock) scan64_tiles_kernel(unsigned long long* tile_sums, int n_tiles) {
  __shared__ unsigned long long s_w[32];
  __shared__ unsigned long long s_carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int b0 = 0; b0 < n_tiles; b0 += kScanBlock) {
    const int i = b0 + threadIdx.x;
    const unsigned long long v = i < n_tiles ? tile_sums[i] : 0ull;
    unsigned long long x = warp_incl_scan64(v, lane);
    if (lane == 31) s_w[warp] = x;
    __syncthreads();
    if (warp == 0) {
      unsigned long long w = s_w[lane];
      w = warp_incl_scan64(w, lane);
      s_w[lane] = w;
    }
    __syncthreads();
    const unsigned long long carry = s_carry;
    const unsigned long long incl = carry + (warp ? s_w[warp - 1] : 0ull) + x;
    if (i < n_tiles) tile_sums[i] = incl - v;  // exclusive
    __syncthreads();
    if (threadIdx.x == kScanBlock - 1) s_carry = incl;
    __syncthreads();
  }
}
// pass 3: inclusive scan inside each tile + tile offset
__global__ void __launch_bounds__(kScanBlock) scan64_apply_kernel(const unsigned long long* __restrict__ v, int n,
                                                                   const unsigned long long* __restrict__ tile_off,
                                                                   unsigned long long* out) {
  __shared__ unsigned long long s_w[32];
  const int base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
  unsigned long long loc[kScanItems];
  unsigned long long t = 0;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    loc[k] = (base + k < n) ? v[base + k] : 0ull;
    t += loc[k];
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned long long incl = warp_incl_scan64(t, lane);
  if (lane == 31) s_w[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    unsigned long long w = s_w[lane];
    w = warp_incl_scan64(w, lane);
    s_w[lane] = w;
  }
  __syncthreads();
  unsigned long long run = tile_off[blockIdx.x] + (warp ? s_w[warp - 1] : 0ull) + incl - t;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    run += loc[k];
    if (base + k < n) out[base + k] = run;
  }
}

// ---- seeds -> (key, cell) list in seed order -------------------------------------------------
// intensity[i] < 0: cell i is not a valid seed.  flag = (valid, key out of the 8-bit range)
__global__ void __launch_bounds__(256) msort_valid_flag_kernel(const float* __restrict__ intensity, int n,
                                                               unsigned long long* v, int* bad) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float x = intensity[i];
  const bool valid = x >= 0.0f;
  if (valid && !(x <= 255.0f && x == (float)(int)x)) atomicOr(bad, 1);
  v[i] = valid ? 1ull : 0ull;
}
__global__ void __launch_bounds__(256) msort_compact_kernel(const float* __restrict__ intensity, int n,
                                                            const unsigned long long* __restrict__ incl,
                                                            unsigned char* key, int* idx, int* seg) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float x = intensity[i];
  if (x >= 0.0f) {
    const int o = (int)incl[i] - 1;
    key[o] = (unsigned char)(int)x;
    idx[o] = i;
    seg[o] = 0;
  }
}

// ---- one introsort level -----------------------------------------------------------------------
// K1: __move_median_to_first(first, first + 1, mid, last - 1), stl_algo.h:84-108
__global__ void __launch_bounds__(256) msort_pivot_kernel(SortSeg* segs, int n_segs, unsigned char* key, int* idx) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_segs) return;
  const int lo = segs[s].lo, hi = segs[s].hi;
  const int a = lo + 1, b = lo + (hi - lo) / 2, c = hi - 1;
  const int ka = key[a], kb = key[b], kc = key[c];
  int m;
  if (ka < kb) m = (kb < kc) ? b : ((ka < kc) ? c : a);
  else if (ka < kc) m = a;
  else if (kb < kc) m = c;
  else m = b;
  const unsigned char k0 = key[lo], km = key[m];
  key[lo] = km; key[m] = k0;
  const int i0 = idx[lo], im = idx[m];
  idx[lo] = im; idx[m] = i0;
  segs[s].pivot = km;
}
// K2: stoppers of the two scans of __unguarded_partition(first + 1, last, first): low word = "not < pivot",
// high word = "not > pivot"
__global__ void __launch_bounds__(256) msort_flag_kernel(const SortSeg* __restrict__ segs, const int* __restrict__ seg,
                                                         const unsigned char* __restrict__ key, int n,
                                                         unsigned long long* v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int s = seg[i];
  unsigned long long f = 0ull;
  if (s >= 0 && i != segs[s].lo) {
    const int k = key[i], p = segs[s].pivot;
    f = (k >= p ? 1ull : 0ull) | (k <= p ? (1ull << 32) : 0ull);
  }
  v[i] = f;
}
// K4: rank -> position tables (1-based global inclusive ranks)
__global__ void __launch_bounds__(256) msort_pos_kernel(const unsigned long long* __restrict__ v,
                                                        const unsigned long long* __restrict__ incl, int n,
                                                        int* posL, int* posR) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long f = v[i], p = incl[i];
  if (f & 1ull) posL[(unsigned)p] = i;
  if (f >> 32) posR[(unsigned)(p >> 32)] = i;
}
// K5: number of swaps, cut, children
__global__ void __launch_bounds__(256) msort_cut_kernel(SortSeg* segs, int n_segs, const unsigned long long* __restrict__ incl,
                                                        const int* __restrict__ posL, const int* __restrict__ posR,
                                                        SortSeg* next, int* n_next, int next_cap, int* overflow) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_segs) return;
  SortSeg g = segs[s];
  const unsigned long long b = incl[g.lo], e = incl[g.hi - 1];
  g.baseL = (int)(unsigned)b; g.baseR = (int)(unsigned)(b >> 32);
  const int totalL = (int)(unsigned)e - g.baseL;
  g.totalR = (int)(unsigned)(e >> 32) - g.baseR;
  // L_k = posL[baseL + k], R_k = posR[baseR + totalR - k + 1]; L_k < R_k holds for k = 1 .. S
  int lo = 0, hi = min(totalL, g.totalR);
  while (lo < hi) {
    const int k = (lo + hi + 1) >> 1;
    if (posL[g.baseL + k] < posR[g.baseR + g.totalR - k + 1]) lo = k; else hi = k - 1;
  }
  g.swaps = lo;
  int cut = INT_MAX;
  if (lo < totalL) cut = posL[g.baseL + lo + 1];
  if (lo >= 1) cut = min(cut, posR[g.baseR + g.totalR - lo + 1]);
  g.cut = cut;
  g.child_left = -1; g.child_right = -1;
  // __introsort_loop(cut, last) and the loop's own continuation on [first, cut)
  if (g.hi - cut > kSortThreshold) {
    const int id = atomicAdd(n_next, 1);
    if (id < next_cap) { next[id].lo = cut; next[id].hi = g.hi; g.child_right = id; } else atomicOr(overflow, 2);
  }
  if (cut - g.lo > kSortThreshold) {
    const int id = atomicAdd(n_next, 1);
    if (id < next_cap) { next[id].lo = g.lo; next[id].hi = cut; g.child_left = id; } else atomicOr(overflow, 2);
  }
  segs[s] = g;
}
// K6: the swaps (each pair by the thread of its left element) and the elements' next segment
__global__ void __launch_bounds__(256) msort_swap_kernel(const SortSeg* __restrict__ segs, int* seg,
                                                         const unsigned long long* __restrict__ v,
                                                         const unsigned long long* __restrict__ incl,
                                                         const int* __restrict__ posR, int n, unsigned char* key, int* idx) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int s = seg[i];
  if (s < 0) return;
  const SortSeg g = segs[s];
  if (v[i] & 1ull) {
    const int k = (int)(unsigned)incl[i] - g.baseL;
    if (k <= g.swaps) {
      const int j = posR[g.baseR + g.totalR - k + 1];
      const unsigned char ki = key[i], kj = key[j];
      key[i] = kj; key[j] = ki;
      const int ii = idx[i], ij = idx[j];
      idx[i] = ij; idx[j] = ii;
    }
  }
  seg[i] = (i >= g.cut) ? g.child_right : g.child_left;
}

// ---- stable counting sort on 8-bit keys (= __final_insertion_sort; also sort_mode 1) -----------
constexpr int kCountChunk = 2048;   // elements per warp
constexpr int kCountWarps = 8;

// hist[key * n_chunks + chunk]
__global__ void __launch_bounds__(32 * kCountWarps) msort_hist_kernel(const unsigned char* __restrict__ key, int n,
                                                                      int n_chunks, unsigned long long* hist) {
  __shared__ int s_h[kCountWarps][256];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int chunk = blockIdx.x * kCountWarps + warp;
  for (int k = lane; k < 256; k += 32) s_h[warp][k] = 0;
  __syncwarp();
  if (chunk < n_chunks) {
    const int c0 = chunk * kCountChunk, c1 = min(c0 + kCountChunk, n);
    for (int i = c0 + lane; i < c1; i += 32) atomicAdd(&s_h[warp][key[i]], 1);
  }
  __syncwarp();
  if (chunk < n_chunks)
    for (int k = lane; k < 256; k += 32) hist[(size_t)k * n_chunks + chunk] = (unsigned long long)s_h[warp][k];
}
// base[key * n_chunks + chunk] = inclusive scan of hist; a chunk's run for `key` starts at incl - count
__global__ void __launch_bounds__(32 * kCountWarps) msort_scatter_kernel(const unsigned char* __restrict__ key,
                                                                         const int* __restrict__ idx, int n, int n_chunks,
                                                                         const unsigned long long* __restrict__ hist,
                                                                         const unsigned long long* __restrict__ incl,
                                                                         int* out) {
  __shared__ int s_b[kCountWarps][256];
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int chunk = blockIdx.x * kCountWarps + warp;
  if (chunk >= n_chunks) return;
  for (int k = lane; k < 256; k += 32) {
    const size_t c = (size_t)k * n_chunks + chunk;
    s_b[warp][k] = (int)(incl[c] - hist[c]);
  }
  __syncwarp();
  const int c0 = chunk * kCountChunk, c1 = min(c0 + kCountChunk, n);
  for (int i0 = c0; i0 < c1; i0 += 32) {
    const int i = i0 + lane;
    const int k = (i < c1) ? (int)key[i] : -1;
    const unsigned peers = __match_any_sync(FULL, k);
    if (k >= 0) {
      const int leader = __ffs(peers) - 1;
      const int rank = __popc(peers & ((1u << lane) - 1u));
      int basepos = 0;
      if (lane == leader) {
        basepos = s_b[warp][k];
        s_b[warp][k] = basepos + __popc(peers);
      }
      basepos = __shfl_sync(peers, basepos, leader);
      out[basepos + rank] = idx[i];
    }
    __syncwarp();
  }
}

}  // namespace pcseg
