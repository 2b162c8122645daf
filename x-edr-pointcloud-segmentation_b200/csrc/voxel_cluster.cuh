// voxel_cluster.cuh -- fixed-radius Euclidean clustering of an UNORGANIZED cloud via a voxel-hash
// grid (BASELINE.json config 4: "unorganized 2M-point synthetic lidar sweep
// via voxel-hash neighbourhoods + Euclidean clustering into ClusterRegions").
//
// The reference has no unorganized path, so the semantics are DEFINED here and
// restated on the CPU in oracle/pcseg_oracle.cc (oracle_cluster_unorganized):
//   * a point takes part iff labels[i] == kUnlabeled and its coordinates are finite;
//   * two such points are neighbours iff ||p - q||^2 < radius^2 in fp32 with the reduction order
//     of ClusterRegion::CheckInlierAndUpdate, d0^2 + (d1^2 + d2^2) (ps/cluster_region.h:94);
//   * clusters are the connected components of that graph (a true closure, unlike the FIFO rule
//     of quirk Q5); components with >= min_region_inliers points become regions, the others stay
//     kUnlabeled;
//   * regions are numbered initial_id_offset, +1, ... by ascending SMALLEST MEMBER INDEX, and a
//     region's inlier list is ascending by index (the deterministic order north_star asks for);
//   * per region: centroid (double sums in a fixed order, see vox_stats_kernel) and axis-aligned
//     extent -- north_star item (4).
// Method: the grid cell is 0.57 * radius, so that a cell's diagonal (0.987 radius) is below the
// radius: ALL points of a cell are mutual neighbours and are united with the cell's first point
// without a single distance test, and every neighbour of a point lies within +-2 cells.  (With
// voxel = radius a 0.5 m voxel of a dense lidar sweep holds thousands of points and the pair
// tests explode: 3 s for the 2 M-point sweep; this formulation: milliseconds.)  Cell keys go into
// an open-addressing hash table (one slot per occupied cell, CAS on the key); a scan over the
// slots turns the per-cell counts into a CSR of the cells' points.  One warp per occupied cell
// then visits the 62 cells of the upper half-space around it: same component already -> skip;
// else the pairs are tested 32 at a time until the FIRST pair below the radius unites the two
// components (exact: components do not depend on which edges are found).  Union-find with
// lock-free hooking towards the SMALLER index (the root of a set is its smallest member, which
// is exactly the ordering key) and path halving; one pointer-jumping pass flattens it.  Labels, offsets and the sorted inlier
// lists then come from scans and the stable counting sort of msort.cuh -- no atomics-ordered
// output anywhere, so the result does not depend on thread scheduling.
#pragma once
#include "grow_common.cuh"

namespace pcseg {

constexpr unsigned long long kVoxEmpty = 0xffffffffffffffffull;
constexpr int kVoxBias = 1 << 20;   // voxel coordinates are clamped to [-2^20, 2^20)

struct VoxArgs {
  const float* xyz;      // [n][3]
  int* labels;           // [n] in/out
  int n;
  float radius, r2, cell;    // cell = grid cell size (0.57 * radius)
  int min_inliers, id_offset;
  unsigned long long* keys;  // [cap] voxel key per slot
  int* head;                 // [cap] number of points of the slot's cell (0 = empty); then its fill cursor
  int cap_mask;              // cap - 1 (power of two)
  int* next;                 // [n] cell-major point list (CSR payload)
  int* slot_of;              // [n] slot of the point's cell, -1 = the point does not take part
  int* parent;               // [n] union-find
  int* count;                // [n] members per root
};

__device__ __forceinline__ unsigned long long vox_key(int vx, int vy, int vz) {
  return ((unsigned long long)(unsigned)(vx + kVoxBias) << 42) | ((unsigned long long)(unsigned)(vy + kVoxBias) << 21) |
         (unsigned long long)(unsigned)(vz + kVoxBias);
}
__device__ __forceinline__ unsigned vox_hash(unsigned long long k) {
  k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
  return (unsigned)k;
}
// voxel coordinate of one component: floor(v / radius), fp32 division, clamped
__device__ __forceinline__ int vox_coord(float v, float radius) {
  const float q = floorf(v / radius);
  return (int)fminf(fmaxf(q, (float)(-kVoxBias)), (float)(kVoxBias - 1));
}

// 1: every participating point finds (or claims) the slot of its voxel and links itself into the list
__global__ void __launch_bounds__(256) vox_insert_kernel(VoxArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  a.parent[i] = i;
  a.count[i] = 0;
  const V3 p = ld3(a.xyz, i);
  if (a.labels[i] != kUnlabeled || !vfinite(p)) { a.slot_of[i] = -1; return; }
  const unsigned long long key = vox_key(vox_coord(p.x, a.cell), vox_coord(p.y, a.cell), vox_coord(p.z, a.cell));
  unsigned s = vox_hash(key) & (unsigned)a.cap_mask;
  while (true) {
    const unsigned long long old = atomicCAS(&a.keys[s], kVoxEmpty, key);
    if (old == kVoxEmpty || old == key) break;
    s = (s + 1) & (unsigned)a.cap_mask;
  }
  a.slot_of[i] = (int)s;
  atomicAdd(&a.head[s], 1);
}

// per slot: (occupied, point count) packed for the scan: low word -> cell ordinal, high word -> CSR offset
__global__ void __launch_bounds__(256) vox_slotflag_kernel(VoxArgs a, int cap, unsigned long long* v) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= cap) return;
  const int c = a.head[s];
  v[s] = c > 0 ? (1ull | ((unsigned long long)(unsigned)c << 32)) : 0ull;
}
// cell records + cursors reset
__global__ void __launch_bounds__(256) vox_cells_kernel(VoxArgs a, int cap, const unsigned long long* __restrict__ incl,
                                                        int* cell_slot, int* cell_begin, int* cell_count) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= cap) return;
  const int c = a.head[s];
  if (c > 0) {
    const int o = (int)(incl[s] & 0xffffffffull) - 1;
    cell_slot[o] = s;
    cell_count[o] = c;
    cell_begin[o] = (int)(incl[s] >> 32) - c;
    a.head[s] = 0;   // becomes the fill cursor of vox_scatter_kernel
  }
}
// CSR payload: the points of every cell (order inside a cell is irrelevant)
__global__ void __launch_bounds__(256) vox_scatter_kernel(VoxArgs a, const unsigned long long* __restrict__ incl,
                                                          const int* __restrict__ cell_begin) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  const int s = a.slot_of[i];
  if (s < 0) return;
  const int o = (int)(incl[s] & 0xffffffffull) - 1;
  a.next[cell_begin[o] + atomicAdd(&a.head[s], 1)] = i;
}

__device__ __forceinline__ int vox_find(int* parent, int x) {
  // path halving; parents only ever decrease, so a stale read is still an ancestor
  while (true) {
    const int p = *(volatile int*)(parent + x);
    if (p == x) return x;
    const int g = *(volatile int*)(parent + p);
    if (g != p) atomicMin(&parent[x], g);
    x = p;
  }
}
__device__ __forceinline__ void vox_unite(int* parent, int x, int y) {
  while (true) {
    x = vox_find(parent, x);
    y = vox_find(parent, y);
    if (x == y) return;
    if (x < y) { const int t = x; x = y; y = t; }   // hook the larger root under the smaller one
    const int old = atomicMin(&parent[x], y);
    if (old == x) return;                            // x was still a root: linked
    x = old;                                         // somebody hooked x first: go on from there
  }
}

// 2a: all points of a cell are mutual neighbours (cell diagonal < radius): unite with the cell's first point
__global__ void __launch_bounds__(256) vox_union_cell_kernel(VoxArgs a, const unsigned long long* __restrict__ incl,
                                                             const int* __restrict__ cell_begin) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  const int s = a.slot_of[i];
  if (s < 0) return;
  const int rep = a.next[cell_begin[(int)(incl[s] & 0xffffffffull) - 1]];
  if (rep != i) vox_unite(a.parent, i, rep);
}

// 2b-0: axis-aligned box of the points of every occupied cell (one warp per cell), for the pruning below
__global__ void __launch_bounds__(256) vox_cellbox_kernel(VoxArgs a, const int* __restrict__ cell_begin,
                                                          const int* __restrict__ cell_count, int n_cells, float* box) {
  const unsigned FULL = 0xffffffffu;
  const int ca = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (ca >= n_cells) return;
  const int b = cell_begin[ca], n = cell_count[ca];
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int i = lane; i < n; i += 32) {
    const V3 p = ld3(a.xyz, a.next[b + i]);
    lo[0] = fminf(lo[0], p.x); lo[1] = fminf(lo[1], p.y); lo[2] = fminf(lo[2], p.z);
    hi[0] = fmaxf(hi[0], p.x); hi[1] = fmaxf(hi[1], p.y); hi[2] = fmaxf(hi[2], p.z);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      lo[k] = fminf(lo[k], __shfl_xor_sync(FULL, lo[k], o));
      hi[k] = fmaxf(hi[k], __shfl_xor_sync(FULL, hi[k], o));
    }
  }
  if (lane < 3) box[6 * ca + lane] = lo[lane];
  else if (lane < 6) box[6 * ca + lane] = hi[lane - 3];
}

// squared distance from p to the box [lo, hi] (0 inside)
__device__ __forceinline__ float vox_box_dist2(const V3& p, const float* lo, const float* hi) {
  const float dx = fmaxf(fmaxf(lo[0] - p.x, p.x - hi[0]), 0.0f);
  const float dy = fmaxf(fmaxf(lo[1] - p.y, p.y - hi[1]), 0.0f);
  const float dz = fmaxf(fmaxf(lo[2] - p.z, p.z - hi[2]), 0.0f);
  return dx * dx + dy * dy + dz * dz;
}

// 2b: one warp per occupied cell: the 62 cells of the upper half-space within +-2 (each unordered cell pair
// is visited once); two cells already in one component are skipped, else their point pairs are tested 32 at a
// time until the first pair below the radius unites the components.  PRUNING (results unchanged: a pair is
// only ever skipped when it cannot be below the radius, with a 1e-4 margin on r^2 against fp32 rounding):
// cell pairs whose point boxes are farther apart than the radius, and points of the cell farther than the
// radius from the other cell's box.  Without it every cell pair that is close but NOT connected (the
// ground under an object, neighbouring scan rings far out) costs all nA x nB tests: 47 of the 50 ms of C4.
// NEAR = true: the 13 face / edge / corner neighbours (launched first: they unite most of a surface, so
// that the second launch, the 49 cells at distance 2, mostly ends at the same-component test).
template <bool NEAR>
__global__ void __launch_bounds__(256) vox_union_pairs_kernel(VoxArgs a, const unsigned long long* __restrict__ incl,
                                                              const int* __restrict__ cell_slot, const int* __restrict__ cell_begin,
                                                              const int* __restrict__ cell_count, int n_cells,
                                                              const float* __restrict__ box) {
  const unsigned FULL = 0xffffffffu;
  const int ca = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (ca >= n_cells) return;
  const unsigned long long keyA = a.keys[cell_slot[ca]];
  const int vx = (int)((keyA >> 42) & 0x1fffff) - kVoxBias, vy = (int)((keyA >> 21) & 0x1fffff) - kVoxBias,
            vz = (int)(keyA & 0x1fffff) - kVoxBias;
  const int bA = cell_begin[ca], nA = cell_count[ca];
  const int repA = a.next[bA];
  const float r2m = a.r2 * 1.0001f;
  float loA[3], hiA[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) { loA[k] = box[6 * ca + k]; hiA[k] = box[6 * ca + 3 + k]; }
  for (int o = 63; o < 125; ++o) {   // offsets (dz, dy, dx) lexicographically after (0, 0, 0)
    const int dz = o / 25 - 2, dy = (o / 5) % 5 - 2, dx = o % 5 - 2;
    const bool near = dz >= -1 && dz <= 1 && dy >= -1 && dy <= 1 && dx >= -1 && dx <= 1;
    if (near != NEAR) continue;
    const int nx = vx + dx, ny = vy + dy, nz = vz + dz;
    if (nx < -kVoxBias || nx >= kVoxBias || ny < -kVoxBias || ny >= kVoxBias || nz < -kVoxBias || nz >= kVoxBias) continue;
    int sB = -1;
    if (lane == 0) {
      const unsigned long long key = vox_key(nx, ny, nz);
      unsigned s = vox_hash(key) & (unsigned)a.cap_mask;
      while (true) {
        const unsigned long long k = a.keys[s];
        if (k == key) { sB = (int)s; break; }
        if (k == kVoxEmpty) break;
        s = (s + 1) & (unsigned)a.cap_mask;
      }
    }
    sB = __shfl_sync(FULL, sB, 0);
    if (sB < 0) continue;
    const int cb = (int)(incl[sB] & 0xffffffffull) - 1;
    float loB[3], hiB[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) { loB[k] = box[6 * cb + k]; hiB[k] = box[6 * cb + 3 + k]; }
    {  // box - box distance
      float d2 = 0.0f;
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const float d = fmaxf(fmaxf(loA[k] - hiB[k], loB[k] - hiA[k]), 0.0f);
        d2 += d * d;
      }
      if (d2 > r2m) continue;
    }
    const int bB = cell_begin[cb], nB = cell_count[cb];
    int same = 0;
    if (lane == 0) same = vox_find(a.parent, repA) == vox_find(a.parent, a.next[bB]);
    if (__shfl_sync(FULL, same, 0)) continue;
    bool done = false;
    for (int i0 = 0; i0 < nA && !done; i0 += 32) {
      // the points of A that can reach B's box at all, 32 candidates at a time
      int pi_l = -1;
      V3 p_l{0.0f, 0.0f, 0.0f};
      if (i0 + lane < nA) {
        pi_l = a.next[bA + i0 + lane];
        p_l = ld3(a.xyz, pi_l);
        if (vox_box_dist2(p_l, loB, hiB) > r2m) pi_l = -1;
      }
      unsigned cand = __ballot_sync(FULL, pi_l >= 0);
      while (cand && !done) {
        const int src = __ffs(cand) - 1;
        cand &= cand - 1u;
        const int pi = __shfl_sync(FULL, pi_l, src);
        const V3 p{__shfl_sync(FULL, p_l.x, src), __shfl_sync(FULL, p_l.y, src), __shfl_sync(FULL, p_l.z, src)};
        for (int j0 = 0; j0 < nB && !done; j0 += 32) {
          const int j = j0 + lane;
          bool hit = false;
          int pj = -1;
          if (j < nB) {
            pj = a.next[bB + j];
            hit = vsqn(vsub(ld3(a.xyz, pj), p)) < a.r2;
          }
          const unsigned m = __ballot_sync(FULL, hit);
          if (m) {
            if (lane == __ffs(m) - 1) vox_unite(a.parent, pi, pj);
            done = true;
          }
        }
      }
    }
    __syncwarp();
  }
}

// 3: flatten + member counts
__global__ void __launch_bounds__(256) vox_flatten_kernel(VoxArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool part = i < a.n && a.slot_of[i] >= 0;
  int r = -1;
  if (part) {
    r = vox_find(a.parent, i);
    a.parent[i] = r;
  }
  // one atomic per (warp, root) instead of one per point: a few dozen roots take all the increments
  const unsigned peers = __match_any_sync(0xffffffffu, r);
  if (part && (int)(threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(&a.count[r], __popc(peers));
}

// 4: packed (is an accepted root, its member count) per point for the scan: low word -> region
// ordinal, high word -> inlier offset
__global__ void __launch_bounds__(256) vox_rootflag_kernel(VoxArgs a, unsigned long long* v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  const bool root = a.slot_of[i] >= 0 && a.parent[i] == i && a.count[i] >= a.min_inliers;
  v[i] = root ? (1ull | ((unsigned long long)(unsigned)a.count[i] << 32)) : 0ull;
}

// 5: labels, per-point region ordinal (n_regions = "none"), per-region records
__global__ void __launch_bounds__(256) vox_label_kernel(VoxArgs a, const unsigned long long* __restrict__ incl, int n_regions,
                                                        int* ord, int* reg_begin, int* reg_count, int* reg_root) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  int o = n_regions;
  if (a.slot_of[i] >= 0) {
    const int r = a.parent[i];
    if (a.count[r] >= a.min_inliers) {
      o = (int)(incl[r] & 0xffffffffull) - 1;
      a.labels[i] = a.id_offset + o;
      if (r == i) {
        reg_count[o] = a.count[r];
        reg_begin[o] = (int)(incl[r] >> 32) - a.count[r];
        reg_root[o] = r;
      }
    }
  }
  ord[i] = o;
}

// one byte of the region ordinal of the points in their current order (LSD radix passes of the
// stable counting sort of msort.cuh)
__global__ void __launch_bounds__(256) vox_digit_kernel(const int* __restrict__ ord, const int* __restrict__ idx, int n, int shift,
                                                        unsigned char* key) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) key[k] = (unsigned char)((ord[idx ? idx[k] : k] >> shift) & 255);
}
__global__ void __launch_bounds__(256) vox_iota_kernel(int* idx, int n) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) idx[k] = k;
}

// 6: centroid + extent of every region: one CTA of 256 threads per region.  Fixed summation order
// (restated by the oracle): thread t adds the elements t, t + 256, ... of the sorted inlier list in
// double; the total is threads 0 .. 255 added in that order; centroid = (float)(total / count).
constexpr int kVoxStatsBlock = 256;
__global__ void __launch_bounds__(kVoxStatsBlock) vox_stats_kernel(const float* __restrict__ xyz, const int* __restrict__ inliers,
                                                                   const int* __restrict__ reg_begin, const int* __restrict__ reg_count,
                                                                   int n_regions, float* centroid, float* bbox) {
  __shared__ double s_sum[kVoxStatsBlock][3];
  __shared__ float s_mn[kVoxStatsBlock / 32][3], s_mx[kVoxStatsBlock / 32][3];
  const int r = blockIdx.x;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  if (r >= n_regions) return;
  const int b = reg_begin[r], n = reg_count[r];
  double sx = 0.0, sy = 0.0, sz = 0.0;
  float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  for (int i = t; i < n; i += kVoxStatsBlock) {
    const V3 p = ld3(xyz, inliers[b + i]);
    sx += (double)p.x; sy += (double)p.y; sz += (double)p.z;
    mn[0] = fminf(mn[0], p.x); mn[1] = fminf(mn[1], p.y); mn[2] = fminf(mn[2], p.z);
    mx[0] = fmaxf(mx[0], p.x); mx[1] = fmaxf(mx[1], p.y); mx[2] = fmaxf(mx[2], p.z);
  }
  s_sum[t][0] = sx; s_sum[t][1] = sy; s_sum[t][2] = sz;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
      mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
    }
  if (lane == 0)
    for (int k = 0; k < 3; ++k) { s_mn[warp][k] = mn[k]; s_mx[warp][k] = mx[k]; }
  __syncthreads();
  if (t < 3) {   // one coordinate each, threads 0 .. 255 in order
    double tot = 0.0;
    for (int l = 0; l < kVoxStatsBlock; ++l) tot += s_sum[l][t];
    centroid[3 * r + t] = (float)(tot / (double)n);
    float lo = s_mn[0][t], hi = s_mx[0][t];
    for (int w = 1; w < kVoxStatsBlock / 32; ++w) { lo = fminf(lo, s_mn[w][t]); hi = fmaxf(hi, s_mx[w][t]); }
    bbox[6 * r + t] = lo; bbox[6 * r + 3 + t] = hi;
  }
}

}  // namespace pcseg
