// meanshift_grow.cuh -- SlidingMeanShift sort + growth phase on the GPU.
//
// Replaces ps/mean_shift_segmentation.h:262-329 and
// internal::AddNeighborsAndGrowReigon (:152-199).  The sort itself stays the
// reference's std::sort call on (seed order, intensity) keys on the host side
// of the C ABI (its permutation is standard-library specific, quirk Q7); this
// file does everything that touches the cloud:
//   ms_intensity_kernel  valid seeds -> intensity (else -1), 4 B per cell, for the sort
//   ms_gather_kernel     sorted cell ids -> (mode xyz, centre cell) records
//   ms_grow_kernel       ONE CTA walks the sorted modes in order:
//     * a mode is skipped when it lies within 1 m^2 of an ACCEPTED earlier mode
//       (:312-322 evaluated lazily: the set of accepted regions before mode j is
//       the same, so the outcome is identical to the eager O(R*S) marking);
//     * the region of a surviving mode is grown with the reference's single FIFO
//       emulated in chunks: queue = [centre cell] + inlier list (every accepted
//       cell is pushed to both in the same order, :190-192); for a chunk of
//       entries every (entry, window slot) pair is tested by one thread, a cell
//       is claimed by the smallest passing (entry, slot) key (atomicMin) and the
//       winners are appended in key order -- exactly the order in which the
//       sequential loop would have labelled them (labels are written at once,
//       there is no "in queue" state, quirk Q15);
//     * < min_region_inliers: labels reset to kUnlabeled (:323-327);
//     * modes whose first step would find nothing (they sit on an area labelled by an earlier
//       region) are no-ops and are filtered out per batch of 1024, in parallel.
#pragma once
#include "meanshift.cuh"

namespace pcseg {

struct MsSeedRec {  // 16 B per sorted valid seed
  float mx, my, mz;
  int centre;       // LinearizeIndex(RowIndex(), ColIndex()) of the shifted index
};

struct MsRegionMeta {
  int inlier_begin, inlier_count, label_id;
  float seed[3];
};

__global__ void ms_intensity_kernel(const ModeRecord* modes, int n, float* intensity) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const ModeRecord r = modes[i];
  intensity[i] = (r.is_seed != 0.0f && r.valid != 0.0f) ? r.intensity : -1.0f;
}

__global__ void ms_gather_kernel(const ModeRecord* modes, const int* sorted_cells, int n_sorted, int rows,
                                 MsSeedRec* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_sorted) return;
  const ModeRecord r = modes[sorted_cells[i]];
  MsSeedRec s;
  s.mx = r.mx; s.my = r.my; s.mz = r.mz;
  // PointIndex::RowIndex/ColIndex = std::round (half away from zero), ps/mean_shift_segmentation.cc:32-34
  s.centre = (int)roundf(r.icol) * rows + (int)roundf(r.irow);
  out[i] = s;
}

struct MsGrowArgs {
  const float* xyz;
  int* labels;
  int* claim;            // N ints, all INT_MAX on entry and on exit
  const MsSeedRec* seeds;
  int n_seeds;
  int rows, cols;
  int half_window;       // config.half_search_window (growth only, quirk Q9)
  int min_inliers;
  int id_offset;
  int* inliers;          // N ints
  MsRegionMeta* meta;    // rcap records
  int rcap;
  int* out;              // [0] = number of regions, [1] = error flags, [2] = total inliers
};

constexpr int kMsGrowBlock = 1024;

__global__ void __launch_bounds__(kMsGrowBlock) ms_grow_kernel(MsGrowArgs a) {
  __shared__ int s_warp[kMsGrowBlock / 32 + 1];
  __shared__ int s_first;
  const int t = threadIdx.x;
  const int rows = a.rows, cols = a.cols;
  const int W = 2 * a.half_window + 1;
  const int WW = W * W;
  const int CH = kMsGrowBlock / WW;  // queue entries per chunk: one (entry, slot) pair per thread
  int n_regions = 0, base = 0;
  for (int pos = 0; pos < a.n_seeds; pos += kMsGrowBlock) {
    const int j = pos + t;
    MsSeedRec me{0.0f, 0.0f, 0.0f, 0};
    bool alive = j < a.n_seeds;
    if (alive) {
      me = a.seeds[j];
      for (int r = 0; r < n_regions && alive; ++r) {  // suppression by regions accepted before this batch
        const float* s = a.meta[r].seed;
        const V3 d{me.mx - s[0], me.my - s[1], me.mz - s[2]};
        if ((double)vsqn(d) < 1.0) alive = false;     // kSquaredCentroidDistanceThreshold, :316-318
      }
    }
    // A mode whose FIRST queue step (the window of its centre cell, :172-199) finds no cell that
    // is unlabeled, not NaN and within 1 m^2 of the mode grows an empty region: rejected, labels
    // untouched, nothing suppressed -- a no-op.  Labels only move kUnlabeled -> label between now
    // and the mode's turn (a rejected region restores what it took), so "empty now" implies
    // "empty then" and such modes are dropped here, in parallel.  Without this every mode that
    // lands on an already labelled area costs a serial CTA-wide round (millions on C5).
    if (alive) {
      const V3 mm{me.mx, me.my, me.mz};
      const int crow = me.centre % rows, ccol = me.centre / rows;
      bool any = false;
      for (int s = 0; s < WW && !any; ++s) {
        const int rc = crow + s % W - a.half_window, cc = ccol + s / W - a.half_window;
        if (cc >= 0 && cc < cols && rc >= 0 && rc < rows) {
          const int n = cc * rows + rc;
          if (a.labels[n] == kUnlabeled) {
            const V3 p = ld3(a.xyz, n);
            if (!vhasnan(p) && !((double)vsqn(vsub(p, mm)) > 1.0)) any = true;
          }
        }
      }
      if (!any) alive = false;
    }
    while (true) {
      // first surviving mode of the batch
      if (t == 0) s_first = INT_MAX;
      __syncthreads();
      if (alive) atomicMin(&s_first, t);
      __syncthreads();
      const int first = s_first;
      __syncthreads();
      if (first == INT_MAX) break;
      const MsSeedRec cur = a.seeds[pos + first];
      const V3 mode{cur.mx, cur.my, cur.mz};
      const int label_id = n_regions + a.id_offset;   // :288-289
      // ---- FIFO growth, :294-308 ----
      int count = 0;   // accepted cells = inlier list length
      int head = 0;    // queue position; entry 0 = centre cell, entry q > 0 = inliers[base + q - 1]
      while (head < count + 1) {
        const bool first_entry = head == 0;            // first_unexamined_index
        const int m = first_entry ? 1 : min(count + 1 - head, CH);
        const int e = t / WW, s = t - e * WW;
        int cell = -1;
        bool pass = false;
        if (e < m && t < CH * WW) {
          const int centre = (head + e == 0) ? cur.centre : a.inliers[base + head + e - 1];
          const int dc = s / W - a.half_window, dr = s % W - a.half_window;  // delta_col outer, delta_row inner
          const int crow = centre % rows, ccol = centre / rows;
          const int rc = crow + dr, cc = ccol + dc;
          if (!(dc == 0 && dr == 0 && !first_entry) && cc >= 0 && cc < cols && rc >= 0 && rc < rows) {
            const int n = cc * rows + rc;
            if (a.labels[n] == kUnlabeled) {
              const V3 p = ld3(a.xyz, n);
              if (!vhasnan(p)) {
                pass = true;
                if ((double)vsqn(vsub(p, mode)) > 1.0) {  // :185-187
                  if (first_entry) pass = false;
                  else if ((double)vsqn(vsub(p, ld3(a.xyz, centre))) > 0.2 * 0.2) pass = false;  // :188-192
                }
                if (pass) {
                  cell = n;
                  atomicMin(&a.claim[n], t);
                }
              }
            }
          }
        }
        __syncthreads();
        const bool win = pass && __ldcg(&a.claim[cell]) == t;  // claims are L2 atomics: bypass L1
        int total;
        const int off = block_scan_excl<kMsGrowBlock>(win ? 1 : 0, &total, s_warp);
        if (win) {
          a.labels[cell] = label_id;                 // :195-196
          a.inliers[base + count + off] = cell;      // :194, :197
          a.claim[cell] = INT_MAX;
        }
        count += total;
        head += m;
        __threadfence_block();
        __syncthreads();
      }
      // ---- CheckAndFinalize (ps/cluster_region.h:111-116) ----
      const bool accepted = count >= a.min_inliers;
      if (accepted) {
        if (n_regions < a.rcap) {
          if (t == 0) {
            MsRegionMeta mt;
            mt.inlier_begin = base; mt.inlier_count = count; mt.label_id = label_id;
            mt.seed[0] = mode.x; mt.seed[1] = mode.y; mt.seed[2] = mode.z;
            a.meta[n_regions] = mt;
          }
          // later modes of this batch within 1 m^2 of the new region are dead, :312-322
          if (alive && t > first) {
            const V3 d{me.mx - mode.x, me.my - mode.y, me.mz - mode.z};
            if ((double)vsqn(d) < 1.0) alive = false;
          }
          ++n_regions;
          base += count;
        } else if (t == 0) {
          atomicOr(&a.out[1], kErrRegionCap);
        }
      } else {
        for (int i = t; i < count; i += kMsGrowBlock) a.labels[a.inliers[base + i]] = kUnlabeled;  // :323-327
      }
      if (t == first) alive = false;
      __threadfence_block();
      __syncthreads();
    }
  }
  if (t == 0) {
    a.out[0] = n_regions;
    a.out[2] = base;
  }
}

}  // namespace pcseg
