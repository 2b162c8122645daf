// meanshift_wave.cuh -- SlidingMeanShift growth phase with SEVERAL regions in flight.
//
// Replaces the serial walk of ps/mean_shift_segmentation.h:282-329 (one mode after the other, each
// region grown by internal::AddNeighborsAndGrowReigon :152-199) without changing its result.
// ms_grow_kernel (meanshift_grow.cuh) emulates that walk with ONE CTA, which one SM's load / atomic
// throughput bounds (C5: 400 regions x 20 k cells, 224 ms).  Here the next K live modes of the
// sorted list ("ranks" 0..K-1 of a WAVE) are grown at the same time, one CTA each, speculatively,
// with temporary labels kMsTmpBase + rank, and a resolver then walks the ranks IN ORDER and keeps
// exactly what the serial walk would have produced:
//
//   * A lower rank is IMMUNE to higher ranks: it treats a cell labelled by a higher rank as
//     unlabeled and takes it (claims carry the rank in their high bits; labels are written with an
//     unsigned atomicMin, so the lowest rank sticks).  Its growth is therefore the growth on the
//     labels left by the ranks below it only.
//   * A rank is VALID iff no cell in the windows of its queue entries (centre + inliers, the
//     entries' own cells included) carries the temporary label of a LOWER rank when the wave has
//     finished (ms_wave_check_kernel; a rank that meets such a label while growing stops at once).
//     Then it never read or wrote a cell a lower rank touched, i.e. it grew on exactly the labels
//     the serial walk would have shown it, provided every lower rank resolved to what is in the
//     labels -- which the resolver guarantees by stopping at the first rank it cannot resolve.
//   * Resolution in rank order: suppressed by a region accepted earlier in this wave (:312-322) ->
//     undone; valid -> accepted if count >= min_region_inliers, else undone (:323-327); invalid ->
//     undone, then the "dead mode" test of ms_filter_kernel on the labels as they now are: a dead
//     mode is a no-op in the serial walk, so the walk goes on; a live one has to be grown again and
//     becomes rank 0 of the next wave, and every higher rank is undone.  Rank 0 has no lower rank,
//     so every wave resolves at least one mode.  Done by four small kernels instead of one serial
//     CTA (its undo loops and dependent global reads took as long as half the growth):
//       ms_wave_plan_kernel    the walk itself on shared-memory copies, ASSUMING every invalid rank
//                              is dead: states, provisional acceptance order, ids, list offsets;
//       ms_wave_undo_kernel    all ranks that are not accepted, undone in parallel (undos only
//                              touch the rank's own temporary label, so they commute);
//       ms_wave_dead_kernel    the dead test of every invalid rank, one warp each, on those labels
//                              (cells of higher ranks count as unlabeled: they are not there yet in
//                              the serial walk); the lowest live one is the STOP rank;
//       ms_wave_commit_kernel  accepted ranks below the stop get their final ids and their inlier
//                              lists packed behind one another; accepted ranks behind it are undone.
//     The dead test of rank k is evaluated as if all lower invalid ranks were no-ops; if one of
//     them is not, the stop lies below k and k's outcome is discarded anyway.
//
// Regions that do not touch (the usual case: the growth rule stops at depth gaps, NaN borders and
// 0.2 m jumps) resolve a whole wave at once.  Caller labels must be below kMsTmpBase.
#pragma once
#include "meanshift_grow.cuh"

namespace pcseg {

constexpr int kMsTmpBase = 0x40000000;  // temporary label of rank k = kMsTmpBase + k
constexpr int kMsWaveMax = 1024;        // ranks per wave (claim key = rank * 1024 + thread < 2^20)

struct MsWaveArgs {
  MsGrowArgs g;
  int* slot_inliers;      // inlier list of each rank: rank 0 has room for every cell (slot0 ints), the
  size_t slot0;           // higher ranks for slot_cap cells each -- a rank that outgrows its slot stops
  int slot_cap;           // as if it had met a lower rank and is grown again as rank 0 of a later wave
  // wave state (device): see the indices below
  int* wave;
};
// layout of MsWaveArgs::wave (ints)
constexpr int kWvNCand = 0;                       // candidates of this wave
constexpr int kWvStop = 1;                        // first unresolved rank (== n_cand: all resolved)
constexpr int kWvNAcc = 2;                        // regions accepted in this wave if nothing stops it
constexpr int kWvFirstAlive = 3;                  // ms_filter_kernel: first live position
constexpr int kWvTotReg = 4;                      // regions / packed inliers after the whole wave
constexpr int kWvTotBase = 5;
constexpr int kWvPos = 8;                         // [kMsWaveMax] position of rank k in the sorted modes
constexpr int kWvCount = kWvPos + kMsWaveMax;     // [kMsWaveMax] inliers of rank k
constexpr int kWvConflict = kWvCount + kMsWaveMax;
constexpr int kWvState = kWvConflict + kMsWaveMax;     // kMsAccepted ...
constexpr int kWvRegBefore = kWvState + kMsWaveMax;    // regions accepted before rank k
constexpr int kWvBaseBefore = kWvRegBefore + kMsWaveMax;  // inliers packed before rank k
constexpr int kWvInts = kWvBaseBefore + kMsWaveMax;
enum { kMsAccepted = 0, kMsRejected = 1, kMsSuppressed = 2, kMsInvalid = 3, kMsOverCap = 4 };

__device__ __forceinline__ int* ms_slot(const MsWaveArgs& w, int k) {
  return w.slot_inliers + (k == 0 ? (size_t)0 : w.slot0 + (size_t)(k - 1) * (size_t)w.slot_cap);
}
__device__ __forceinline__ bool ms_is_lower_tmp(int lab, int mylab) { return lab >= kMsTmpBase && lab < mylab; }

// The first `want` live modes at positions >= pos0 (one CTA).
__global__ void __launch_bounds__(1024) ms_wave_select_kernel(MsWaveArgs w, int pos0, int want) {
  __shared__ int s_warp[33];
  const int t = threadIdx.x;
  int found = 0;
  const int fa = w.wave[kWvFirstAlive];   // everything below it is dead (ms_filter_kernel)
  for (int pos = max(pos0, min(fa, w.g.n_seeds)); pos < w.g.n_seeds && found < want; pos += 1024) {
    const int j = pos + t;
    const bool alive = j < w.g.n_seeds && w.g.alive[j] != 0;
    if (!__syncthreads_or(alive)) continue;
    int total;
    const int off = block_scan_excl<1024>(alive ? 1 : 0, &total, s_warp);
    if (alive && found + off < want) {
      w.wave[kWvPos + found + off] = j;
      w.wave[kWvCount + found + off] = 0;
      w.wave[kWvConflict + found + off] = 0;
    }
    found = min(found + total, want);
  }
  if (t == 0) {
    w.wave[kWvNCand] = found;
    w.wave[kWvStop] = found;
    w.wave[kWvNAcc] = 0;
  }
}

// One CTA per rank: the FIFO growth of ms_grow_kernel (one (entry, window slot) pair per thread,
// min-key claims, ordered append) on temporary labels, immune to higher ranks, stopping at the
// first sign of a lower rank.  Nothing is finalised here.
__global__ void __launch_bounds__(kMsGrowBlock) ms_wave_grow_kernel(MsWaveArgs w) {
  __shared__ int s_warp[kMsGrowBlock / 32 + 1];
  __shared__ int s_conflict;
  const MsGrowArgs& a = w.g;
  const int k = blockIdx.x;
  if (k >= w.wave[kWvNCand]) return;
  const int t = threadIdx.x;
  const int rows = a.rows, cols = a.cols;
  const int W = 2 * a.half_window + 1;
  const int WW = W * W;
  const int CH = kMsGrowBlock / WW;   // queue entries per chunk
  const int mylab = kMsTmpBase + k;
  const int ckey = k * kMsGrowBlock + t;   // claims of a lower rank win
  int* inl = ms_slot(w, k);
  const MsSeedRec cur = a.seeds[w.wave[kWvPos + k]];
  const V3 mode{cur.mx, cur.my, cur.mz};
  if (t == 0) s_conflict = 0;
  __syncthreads();
  int count = 0;   // accepted cells = inlier list length
  int head = 0;    // queue position; entry 0 = centre cell, entry q > 0 = inl[q - 1]
  bool stop = false;
  while (head < count + 1 && !stop) {
    const bool first_entry = head == 0;
    const int m = first_entry ? 1 : min(count + 1 - head, CH);
    const int e = t / WW, s = t - e * WW;
    int cell = -1;
    bool pass = false;
    if (e < m) {
      const int centre = (head + e == 0) ? cur.centre : inl[head + e - 1];
      const int dc = s / W - a.half_window, dr = s % W - a.half_window;  // delta_col outer, delta_row inner
      const int crow = centre % rows, ccol = centre / rows;
      const int rc = crow + dr, cc = ccol + dc;
      if (!(dc == 0 && dr == 0 && !first_entry) && cc >= 0 && cc < cols && rc >= 0 && rc < rows) {
        const int n = cc * rows + rc;
        const int lab = __ldcg(&a.labels[n]);   // labels move under L2 atomics of other CTAs
        if (ms_is_lower_tmp(lab, mylab)) s_conflict = 1;
        if (lab == kUnlabeled || lab > mylab) {  // a higher rank's cell counts as unlabeled
          const V3 p = ld3(a.xyz, n);
          if (!vhasnan(p)) {
            bool ok = true;
            if ((double)vsqn(vsub(p, mode)) > 1.0) {  // :185-187
              if (first_entry) ok = false;
              else if ((double)vsqn(vsub(p, ld3(a.xyz, centre))) > 0.2 * 0.2) ok = false;  // :188-192
            }
            if (ok) {
              pass = true;
              cell = n;
              atomicMin(&a.claim[n], ckey);
            }
          }
        }
      }
    }
    __syncthreads();
    pass = pass && __ldcg(&a.claim[cell < 0 ? 0 : cell]) == ckey;
    int total;
    const int off = block_scan_excl<kMsGrowBlock>(pass ? 1 : 0, &total, s_warp);
    const bool fits = k == 0 || count + total <= w.slot_cap;
    if (pass) {
      if (fits) {
        const unsigned old = atomicMin(reinterpret_cast<unsigned*>(&a.labels[cell]), (unsigned)mylab);  // :195-196
        if (ms_is_lower_tmp((int)old, mylab)) s_conflict = 1;
        inl[count + off] = cell;                                                                     // :194, :197
      } else {
        s_conflict = 1;   // the slot is full: give up here, nothing of this chunk is taken
      }
      atomicCAS(&a.claim[cell], ckey, INT_MAX);
    }
    if (!fits) total = 0;
    count += total;
    head += m;
    __threadfence_block();
    stop = __syncthreads_or(s_conflict != 0);
  }
  if (t == 0) {
    w.wave[kWvCount + k] = count;
    if (stop) w.wave[kWvConflict + k] = 1;
  }
}

// Validity of the ranks: any cell in the windows of rank k's entries with a lower rank's label?
// grid = (blocks over the (entry, slot) pairs, rank).
__global__ void __launch_bounds__(256) ms_wave_check_kernel(MsWaveArgs w) {
  const MsGrowArgs& a = w.g;
  const int k = blockIdx.y;
  if (k == 0 || k >= w.wave[kWvNCand] || w.wave[kWvConflict + k] != 0) return;
  const int rows = a.rows, cols = a.cols;
  const int W = 2 * a.half_window + 1;
  const int WW = W * W;
  const int mylab = kMsTmpBase + k;
  const int count = w.wave[kWvCount + k];
  const int* inl = ms_slot(w, k);
  const int centre0 = a.seeds[w.wave[kWvPos + k]].centre;
  const long long pairs = (long long)(count + 1) * WW;
  bool bad = false;
  for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < pairs && !bad; q += (long long)gridDim.x * blockDim.x) {
    const int e = (int)(q / WW), s = (int)(q - (long long)e * WW);
    const int centre = e == 0 ? centre0 : inl[e - 1];
    const int rc = centre % rows + s % W - a.half_window, cc = centre / rows + s / W - a.half_window;
    if (cc >= 0 && cc < cols && rc >= 0 && rc < rows) bad = ms_is_lower_tmp(a.labels[cc * rows + rc], mylab);
  }
  if (bad) w.wave[kWvConflict + k] = 1;
}


// The walk over the ranks on shared-memory copies (one CTA of kMsWaveMax threads), under the
// assumption that every invalid rank is dead.  n_regions0 / base0: accepted before this wave.
__global__ void __launch_bounds__(kMsWaveMax) ms_wave_plan_kernel(MsWaveArgs w, int n_regions0, int base0) {
  __shared__ float s_mode[3 * kMsWaveMax];   // mode of rank k
  __shared__ float s_acc[3 * kMsWaveMax];    // modes of the regions accepted so far in this wave
  __shared__ int s_cnt[kMsWaveMax], s_conf[kMsWaveMax], s_state[kMsWaveMax], s_reg[kMsWaveMax], s_base[kMsWaveMax];
  const MsGrowArgs& a = w.g;
  const int t = threadIdx.x;
  const int n_cand = w.wave[kWvNCand];
  if (t < n_cand) {
    const MsSeedRec r = a.seeds[w.wave[kWvPos + t]];
    s_mode[3 * t] = r.mx; s_mode[3 * t + 1] = r.my; s_mode[3 * t + 2] = r.mz;
    s_cnt[t] = w.wave[kWvCount + t];
    s_conf[t] = w.wave[kWvConflict + t];
  }
  __syncthreads();
  int n_regions = n_regions0, base = base0, n_acc = 0;
  for (int k = 0; k < n_cand; ++k) {
    // suppression by a region accepted earlier in this wave, :312-322 (earlier waves: ms_filter_kernel)
    bool sup = false;
    if (t < n_acc) {
      const V3 d{s_mode[3 * k] - s_acc[3 * t], s_mode[3 * k + 1] - s_acc[3 * t + 1], s_mode[3 * k + 2] - s_acc[3 * t + 2]};
      sup = (double)vsqn(d) < 1.0;
    }
    sup = __syncthreads_or(sup);
    int state;
    if (sup) state = kMsSuppressed;
    else if (s_conf[k]) state = kMsInvalid;
    else if (s_cnt[k] < a.min_inliers) state = kMsRejected;    // CheckAndFinalize, ps/cluster_region.h:111-116
    else if (n_regions >= a.rcap) state = kMsOverCap;
    else state = kMsAccepted;
    if (t == 0) { s_state[k] = state; s_reg[k] = n_regions; s_base[k] = base; }
    if (state == kMsAccepted) {
      if (t == 0) { s_acc[3 * n_acc] = s_mode[3 * k]; s_acc[3 * n_acc + 1] = s_mode[3 * k + 1]; s_acc[3 * n_acc + 2] = s_mode[3 * k + 2]; }
      ++n_acc; ++n_regions; base += s_cnt[k];
    }
    __syncthreads();
  }
  if (t < n_cand) {
    w.wave[kWvState + t] = s_state[t];
    w.wave[kWvRegBefore + t] = s_reg[t];
    w.wave[kWvBaseBefore + t] = s_base[t];
  }
  if (t == 0) {
    w.wave[kWvStop] = n_cand;
    w.wave[kWvNAcc] = n_acc;
    w.wave[kWvTotReg] = n_regions;
    w.wave[kWvTotBase] = base;
  }
}

// Undo of a rank: its cells that still carry ITS label become kUnlabeled (cells a lower rank took
// keep that rank's label).  grid = (blocks, rank).
__device__ __forceinline__ void ms_wave_undo(const MsWaveArgs& w, int k) {
  const int* inl = ms_slot(w, k);
  const int count = w.wave[kWvCount + k];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
    atomicCAS(&w.g.labels[inl[i]], kMsTmpBase + k, kUnlabeled);
}
__global__ void __launch_bounds__(256) ms_wave_undo_kernel(MsWaveArgs w) {
  const int k = blockIdx.y;
  if (k >= w.wave[kWvNCand] || w.wave[kWvState + k] == kMsAccepted) return;
  ms_wave_undo(w, k);
}

// Dead test of the invalid ranks (one warp each): the first queue step of the mode finds nothing.
__global__ void __launch_bounds__(256) ms_wave_dead_kernel(MsWaveArgs w) {
  const MsGrowArgs& a = w.g;
  const int k = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (k >= w.wave[kWvNCand] || w.wave[kWvState + k] != kMsInvalid) return;
  const int rows = a.rows, cols = a.cols;
  const int W = 2 * a.half_window + 1;
  const int WW = W * W;
  const MsSeedRec cur = a.seeds[w.wave[kWvPos + k]];
  const V3 mm{cur.mx, cur.my, cur.mz};
  const int crow = cur.centre % rows, ccol = cur.centre / rows;
  const int mylab = kMsTmpBase + k;
  bool live = false;
  for (int s = lane; s < WW && !live; s += 32) {
    const int rc = crow + s % W - a.half_window, cc = ccol + s / W - a.half_window;
    if (cc >= 0 && cc < cols && rc >= 0 && rc < rows) {
      const int n = cc * rows + rc;
      const int lab = a.labels[n];
      if (lab == kUnlabeled || lab > mylab) {   // higher ranks are not there yet in the serial walk
        const V3 p = ld3(a.xyz, n);
        if (!vhasnan(p) && !((double)vsqn(vsub(p, mm)) > 1.0)) live = true;
      }
    }
  }
  if (__any_sync(0xffffffffu, live) && lane == 0) atomicMin(&w.wave[kWvStop], k);   // has to be grown again
}

// Final ids + packed inlier lists of the accepted ranks below the stop; undo of those behind it;
// dead flags; totals.  grid = (blocks, rank).
__global__ void __launch_bounds__(256) ms_wave_commit_kernel(MsWaveArgs w) {
  const MsGrowArgs& a = w.g;
  const int k = blockIdx.y;
  const int n_cand = w.wave[kWvNCand];
  const int stop = w.wave[kWvStop];
  const bool lead = blockIdx.x == 0 && threadIdx.x == 0;
  if (k == 0 && lead) {
    a.out[0] = stop < n_cand ? w.wave[kWvRegBefore + stop] : w.wave[kWvTotReg];
    a.out[2] = stop < n_cand ? w.wave[kWvBaseBefore + stop] : w.wave[kWvTotBase];
    a.out[3] = stop < n_cand ? w.wave[kWvPos + stop] : (n_cand > 0 ? w.wave[kWvPos + n_cand - 1] + 1 : a.n_seeds);
  }
  if (k >= n_cand) return;
  const int state = w.wave[kWvState + k];
  if (k >= stop) {
    if (state == kMsAccepted) ms_wave_undo(w, k);
    return;
  }
  const int pos = w.wave[kWvPos + k];
  if (lead) {
    a.alive[pos] = 0;
    if (state == kMsOverCap) atomicOr(&a.out[1], kErrRegionCap);
  }
  if (state != kMsAccepted) return;
  const int count = w.wave[kWvCount + k], base = w.wave[kWvBaseBefore + k];
  const int label = w.wave[kWvRegBefore + k] + a.id_offset;   // :288-289
  if (lead) {
    const MsSeedRec cur = a.seeds[pos];
    MsRegionMeta mt;
    mt.inlier_begin = base; mt.inlier_count = count; mt.label_id = label;
    mt.seed[0] = cur.mx; mt.seed[1] = cur.my; mt.seed[2] = cur.mz;
    a.meta[w.wave[kWvRegBefore + k]] = mt;
  }
  const int* inl = ms_slot(w, k);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
    const int n = inl[i];
    a.labels[n] = label;
    a.inliers[base + i] = n;
  }
}

}  // namespace pcseg
