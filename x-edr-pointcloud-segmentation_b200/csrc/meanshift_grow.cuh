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
//       region) are no-ops and are filtered out in parallel;
//   ms_filter_kernel     the same two "dead mode" tests (suppressed by an accepted region, empty
//       first step) for ALL remaining modes on the whole GPU.  Both are monotone (accepted
//       regions only accumulate; labels only move kUnlabeled -> label across region boundaries),
//       so a mode found dead now is dead at its turn.  The host alternates filter and growth
//       launches: the growth CTA stops after kMsRegionsPerRound accepted regions, the filter
//       clears the flags of the modes those regions killed, and the growth CTA then skips
//       dead batches with one flag read instead of a serial CTA-wide round per 1024 modes.
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
  int* out;              // [0] = number of regions, [1] = error flags, [2] = total inliers, [3] = resume position
  unsigned char* alive;  // n_seeds flags: 0 = known dead (ms_filter_kernel)
  int* first_alive;      // nullable: ms_filter_kernel lowers it to the first position it leaves alive
};

constexpr int kMsRegionsPerRound = 8;

// Dead-mode tests for the modes [pos0, n_seeds) against the regions [r0, r1) accepted since the
// last pass and the current labels.
__global__ void ms_filter_kernel(MsGrowArgs a, int pos0, int r0, int r1) {
  __shared__ float s_seed[3 * 64];
  const int rows = a.rows, cols = a.cols;
  const int W = 2 * a.half_window + 1;
  const int WW = W * W;
  for (long long j0 = (long long)pos0 + (long long)blockIdx.x * blockDim.x; j0 < a.n_seeds; j0 += (long long)gridDim.x * blockDim.x) {
    const long long jl = j0 + threadIdx.x;
    const int j = (int)(jl < a.n_seeds ? jl : a.n_seeds - 1);
    bool alive = jl < a.n_seeds && a.alive[j] != 0;
    const MsSeedRec me = alive ? a.seeds[j] : MsSeedRec{0.0f, 0.0f, 0.0f, 0};
    for (int rb = r0; rb < r1; rb += 64) {
      const int cnt = min(64, r1 - rb);
      __syncthreads();
      if (threadIdx.x < 3 * cnt) s_seed[threadIdx.x] = a.meta[rb + threadIdx.x / 3].seed[threadIdx.x % 3];
      __syncthreads();
      if (alive) {
        for (int r = 0; r < cnt && alive; ++r) {
          const V3 d{me.mx - s_seed[3 * r], me.my - s_seed[3 * r + 1], me.mz - s_seed[3 * r + 2]};
          if ((double)vsqn(d) < 1.0) alive = false;   // kSquaredCentroidDistanceThreshold, :316-318
        }
      }
    }
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
    if (jl < a.n_seeds && !alive) a.alive[j] = 0;
    if (a.first_alive) {   // first live position (lanes hold consecutive positions: the lowest live lane)
      const unsigned live = __ballot_sync(0xffffffffu, alive);
      if (live && (int)(threadIdx.x & 31) == __ffs(live) - 1 && j < *(volatile int*)a.first_alive) atomicMin(a.first_alive, j);
    }
  }
}

constexpr int kMsGrowBlock = 1024;
constexpr int kMsPairs = 1;   // (entry, slot) pairs per thread and chunk (4: measured slower, the pair chains serialise)

// pos_begin: first mode of this launch; r_filter: regions [0, r_filter) are already applied to
// the alive flags (only later ones are tested here); base_begin: inliers written so far.
__global__ void __launch_bounds__(kMsGrowBlock) ms_grow_kernel(MsGrowArgs a, int pos_begin, int r_filter, int base_begin) {
  __shared__ int s_warp[kMsGrowBlock / 32 + 1];
  __shared__ int s_first;
  const int t = threadIdx.x;
  const int rows = a.rows, cols = a.cols;
  const int W = 2 * a.half_window + 1;
  const int WW = W * W;
  const int CH = kMsGrowBlock * kMsPairs / WW;  // queue entries per chunk: kMsPairs (entry, slot) pairs per thread
  int n_regions = r_filter, base = base_begin;
  int pos = pos_begin;
  for (; pos < a.n_seeds; pos += kMsGrowBlock) {
    if (n_regions - r_filter >= kMsRegionsPerRound) break;   // let the filter catch up (resume at pos)
    const int j = pos + t;
    MsSeedRec me{0.0f, 0.0f, 0.0f, 0};
    bool alive = j < a.n_seeds && a.alive[j] != 0;
    if (!__syncthreads_or(alive)) continue;                  // a dead batch costs one flag read
    if (alive) {
      me = a.seeds[j];
      for (int r = r_filter; r < n_regions && alive; ++r) {  // suppression by regions accepted since the last filter pass
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
        // thread t tests the (entry, slot) pairs t * kMsPairs + u: contiguous keys per thread, so the
        // key order of the winners is thread order x local order
        int cellv[kMsPairs];
        bool passv[kMsPairs];
#pragma unroll
        for (int u = 0; u < kMsPairs; ++u) {
          const int key = t * kMsPairs + u;
          const int e = key / WW, s = key - e * WW;
          cellv[u] = -1;
          passv[u] = false;
          if (e < m) {
            const int centre = (head + e == 0) ? cur.centre : a.inliers[base + head + e - 1];
            const int dc = s / W - a.half_window, dr = s % W - a.half_window;  // delta_col outer, delta_row inner
            const int crow = centre % rows, ccol = centre / rows;
            const int rc = crow + dr, cc = ccol + dc;
            if (!(dc == 0 && dr == 0 && !first_entry) && cc >= 0 && cc < cols && rc >= 0 && rc < rows) {
              const int n = cc * rows + rc;
              if (a.labels[n] == kUnlabeled) {
                const V3 p = ld3(a.xyz, n);
                if (!vhasnan(p)) {
                  bool pass = true;
                  if ((double)vsqn(vsub(p, mode)) > 1.0) {  // :185-187
                    if (first_entry) pass = false;
                    else if ((double)vsqn(vsub(p, ld3(a.xyz, centre))) > 0.2 * 0.2) pass = false;  // :188-192
                  }
                  if (pass) {
                    passv[u] = true;
                    cellv[u] = n;
                    atomicMin(&a.claim[n], key);
                  }
                }
              }
            }
          }
        }
        __syncthreads();
        int nw = 0;
#pragma unroll
        for (int u = 0; u < kMsPairs; ++u) {
          passv[u] = passv[u] && __ldcg(&a.claim[cellv[u]]) == t * kMsPairs + u;  // claims are L2 atomics: bypass L1
          nw += passv[u] ? 1 : 0;
        }
        int total;
        int off = block_scan_excl<kMsGrowBlock>(nw, &total, s_warp);
#pragma unroll
        for (int u = 0; u < kMsPairs; ++u) {
          if (passv[u]) {
            a.labels[cellv[u]] = label_id;                 // :195-196
            a.inliers[base + count + off] = cellv[u];      // :194, :197
            a.claim[cellv[u]] = INT_MAX;
            ++off;
          }
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
    a.out[3] = pos < a.n_seeds ? pos : a.n_seeds;
  }
}

}  // namespace pcseg
