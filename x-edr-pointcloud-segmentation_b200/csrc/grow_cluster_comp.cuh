// grow_cluster_comp.cuh -- SegmentRegions<ClusterRegion<>> of ONE large frame, one CTA per
// connected component (the fast path of pcseg_segment_cluster for single frames).
//
// A frame's cluster growth is one FIFO (grow_cluster.cuh) and its time is the number of FIFO
// levels times the latency of a level -- on a lidar range image (128 x 16384) ~24 000 levels of
// a few hundred entries: 106 ms on one CTA (column bands: cluster_band_kernel) against 52 ms on a
// host core.  Regions of DIFFERENT connected components never influence one another:
//   * components are those of the graph "cells within the search window of one another whose
//     points are closer than the threshold" over the cells that can take part at all (unlabeled,
//     finite) -- the very predicate of CheckInlierAndUpdate (ps/cluster_region.h:94), which is
//     symmetric in its two points bit for bit;
//   * a region only ever accepts a candidate that passes that predicate against a cell of the
//     region, so it stays inside its component; a candidate of ANOTHER component is enqueued by
//     the reference, fails its test and reverts to kUnlabeled (ps/cluster_region.h:105): it
//     occupies a queue position and changes nothing else.  Never enqueueing it leaves labels,
//     inlier order and regions as they are.  (Quirk Q5 -- tested only against the first source,
//     re-enqueued later -- stays in force INSIDE a component, where the FIFO is still emulated
//     entry by entry.)
//   * a component with fewer cells than min_region_inliers cannot yield a region (every attempt
//     fails and is reset, ps/segmentation.h:280-291): it is skipped.
//   * the only state one CTA could see of another is the transient mark "rejected in this step"
//     (grow_common.cuh) on a cell of the other component; the eligibility test therefore also asks
//     whose cell it is (the flattened union-find) in that -- rare -- case.
// So: union-find over the grid (hooking to the smaller index: the root is the component's first
// seed in the ascending order in which the all-descending seed list is consumed), the cells of a
// component are re-coded kCompUnlabeledBase + root in the label array, one CTA per component runs
// cluster_grow_frame on "its" code, and the regions of all components are numbered by ascending
// seed index -- the order in which the reference accepts them -- with their inlier lists packed
// in that order (host: a sort of a few hundred records; device: cluster_comp_pack_kernel).
#pragma once
#include "grow_cluster.cuh"
#include "voxel_cluster.cuh"   // vox_find / vox_unite

namespace pcseg {

// parent[i] = i for the cells that can take part, -1 for the others
__global__ void __launch_bounds__(256) cluster_comp_init_kernel(const float* __restrict__ xyz, const int* __restrict__ labels,
                                                                int n, int* parent, int* count, int* hi) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const bool part = labels[i] == kUnlabeled && vfinite(ld3(xyz, i));
  parent[i] = part ? i : -1;
  count[i] = 0;
  hi[i] = -1;
}

// edges to the window cells with a LARGER index (every unordered pair once)
__global__ void __launch_bounds__(256) cluster_comp_link_kernel(const float* __restrict__ xyz, int rows, int cols, int h,
                                                                float sq_thr, int* parent) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * cols || parent[i] < 0) return;
  const int row = i % rows, col = i / rows;
  const V3 p = ld3(xyz, i);
  for (int dc = 0; dc <= h; ++dc) {
    const int cc = col + dc;
    if (cc >= cols) break;
    for (int dr = (dc == 0 ? 1 : -h); dr <= h; ++dr) {
      const int rc = row + dr;
      if (rc < 0 || rc >= rows) continue;
      const int j = cc * rows + rc;
      if (*(volatile int*)(parent + j) < 0) continue;
      // (candidate - source).squaredNorm() < threshold, ps/cluster_region.h:94 (same value both ways)
      if (vsqn(vsub(ld3(xyz, j), p)) < sq_thr) vox_unite(parent, i, j);
    }
  }
}

// root of every cell -> its code in the label array, member count and largest member index per root
__global__ void __launch_bounds__(256) cluster_comp_flatten_kernel(int n, int* parent, int* labels, int* count, int* hi) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool part = i < n && parent[i] >= 0;
  int r = -1;
  if (part) {
    r = vox_find(parent, i);
    parent[i] = r;   // roots only ever point to themselves: safe while other threads still search
    labels[i] = kCompUnlabeledBase + r;
  }
  // one atomic pair per (warp, root): the lanes of a warp are consecutive cells, mostly of one component
  const unsigned peers = __match_any_sync(0xffffffffu, r);
  if (part && (int)(threadIdx.x & 31) == __ffs(peers) - 1) {
    atomicAdd(&count[r], __popc(peers));
    atomicMax(&hi[r], i - (int)(threadIdx.x & 31) + 31 - __clz(peers));   // the highest lane of the group
  }
}

// components that can yield a region: {root, count, hi} appended to list (order irrelevant, sorted on the host)
__global__ void __launch_bounds__(256) cluster_comp_list_kernel(int n, const int* __restrict__ parent, const int* __restrict__ count,
                                                                const int* __restrict__ hi, int min_inliers, int* list, int cap,
                                                                int* n_list) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || parent[i] < 0 || count[i] < min_inliers || count[i] == 0) return;   // count > 0 only at roots
  const int k = atomicAdd(n_list, 1);
  if (k < cap) { list[3 * k] = i; list[3 * k + 1] = count[i]; list[3 * k + 2] = hi[i]; }
}

struct ClusterCompTable {
  const int* root;         // [n_comp]
  const int* hi;
  const long long* q_off;  // queue offset, in pairs
  const int* q_cap;
  const int* inl_off;      // the component's slice of the frame's inlier buffer
  const int* m_off;        // offset into the temporary region table
  int* queue;
  ClusterMeta* meta;
  int* out;                // [n_comp][4]
  int rcap_total;
  const int* comp;         // [N] root of every participating cell (the flattened union-find)
};

template <int BLOCK, bool H1>
__global__ void __launch_bounds__(BLOCK, H1 ? 2 : 1) cluster_comp_kernel(ClusterArgs a, ClusterCompTable t) {
  __shared__ int s_warp[BLOCK / 32 + 1];
  const int k = blockIdx.x;
  ClusterBand band;
  band.on = 1;
  band.c_begin = 0; band.c_end = 0;
  band.queue = t.queue + 2 * t.q_off[k]; band.qcap = t.q_cap[k];
  band.inl_base = t.inl_off[k];
  band.meta = t.meta + t.m_off[k];
  band.rcap = (k + 1 < (int)gridDim.x ? t.m_off[k + 1] : t.rcap_total) - t.m_off[k];
  band.out = t.out + 4 * k;
  band.comp_on = 1;
  band.unl = kCompUnlabeledBase + t.root[k];
  band.idx_lo = t.root[k]; band.idx_hi = t.hi[k];   // idx_lo is the root: the component's smallest index
  band.comp = t.comp;
  cluster_grow_frame<BLOCK, H1>(a, 0, a.id_offset, s_warp, band);   // provisional ids: renumbered afterwards
}

// one region per block row: its inlier list moves to its final place (tmp), its cells get the final id
struct ClusterCompMove { int src, count, dst, label_id; };
__global__ void __launch_bounds__(256) cluster_comp_pack_kernel(ClusterArgs a, const ClusterCompMove* __restrict__ mv, int n_regions, int* tmp) {
  for (int r = blockIdx.x; r < n_regions; r += gridDim.x) {
    const ClusterCompMove m = mv[r];
    for (int i = threadIdx.x; i < m.count; i += blockDim.x) {
      const int n = a.inliers[m.src + i];
      tmp[m.dst + i] = n;
      a.labels[n] = m.label_id;
    }
    if (threadIdx.x == 0) a.meta[r] = ClusterMeta{m.dst, m.count, m.label_id};
  }
}
__global__ void cluster_comp_stats_kernel(ClusterArgs a, int n_regions, int n_inliers, int err, int steps) {
  a.stats[0].n_cluster = n_regions;
  a.stats[0].cluster_inliers = n_inliers;
  a.stats[0].error_flags |= err;
  a.stats[0].grow_steps += steps;
}
// what is still coded (cells of components that were skipped, cells no region took) or marked as
// examined by a failed region is unlabeled again, ps/segmentation.h:287-291
__global__ void __launch_bounds__(256) cluster_comp_restore_kernel(int* labels, int n) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int l = labels[i];
    if (l < kCompUnlabeledBase + (1 << 30) || l == kAlreadyExamedPoint) labels[i] = kUnlabeled;
  }
}

}  // namespace pcseg
