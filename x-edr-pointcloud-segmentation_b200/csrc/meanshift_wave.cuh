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
//   * Resolver, rank by rank (ms_wave_resolve_kernel, one CTA): suppressed by a region accepted
//     earlier in this wave (:312-322) -> undone; valid -> accepted if count >= min_region_inliers,
//     else undone (:323-327); invalid -> undone, then the "dead mode" test of ms_filter_kernel on the
//     labels as they now are: a dead mode is a no-op in the serial walk, so the walk goes on; a live
//     one has to be grown again and becomes rank 0 of the next wave, and every higher rank is undone.
//     Rank 0 has no lower rank, so every wave resolves at least one mode.
//   * ms_wave_finalize_kernel writes the final label ids (acceptance order = rank order) and packs
//     the inlier lists of the accepted regions behind one another.
//
// Regions that do not touch (the usual case: the growth rule stops at depth gaps, NaN borders and
// 0.2 m jumps) resolve a whole wave at once.  Caller labels must be below kMsTmpBase.
#pragma once
#include "meanshift_grow.cuh"

namespace pcseg {

constexpr int kMsTmpBase = 0x40000000;  // temporary label of rank k = kMsTmpBase + k
constexpr int kMsWaveMax = 256;         // ranks per wave (claim key = rank * 1024 + thread)

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
constexpr int kWvNAcc = 2;                        // regions accepted in this wave
constexpr int kWvPos = 8;                         // [kMsWaveMax] position of rank k in the sorted modes
constexpr int kWvCount = kWvPos + kMsWaveMax;     // [kMsWaveMax] inliers of rank k
constexpr int kWvConflict = kWvCount + kMsWaveMax;
constexpr int kWvAccRank = kWvConflict + kMsWaveMax;   // [kMsWaveMax] accepted: rank
constexpr int kWvAccBase = kWvAccRank + kMsWaveMax;    //   inlier_begin
constexpr int kWvAccLabel = kWvAccBase + kMsWaveMax;   //   final label id
constexpr int kWvInts = kWvAccLabel + kMsWaveMax;

__device__ __forceinline__ int* ms_slot(const MsWaveArgs& w, int k) {
  return w.slot_inliers + (k == 0 ? (size_t)0 : w.slot0 + (size_t)(k - 1) * (size_t)w.slot_cap);
}
__device__ __forceinline__ bool ms_is_lower_tmp(int lab, int mylab) { return lab >= kMsTmpBase && lab < mylab; }

// The first `want` live modes at positions >= pos0 (one CTA).
__global__ void __launch_bounds__(1024) ms_wave_select_kernel(MsWaveArgs w, int pos0, int want) {
  __shared__ int s_warp[33];
  const int t = threadIdx.x;
  int found = 0;
  for (int pos = pos0; pos < w.g.n_seeds && found < want; pos += 1024) {
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

// Undo of rank k by the whole CTA: its cells that still carry ITS label become kUnlabeled (cells a
// lower rank took keep that rank's label).
__device__ __forceinline__ void ms_wave_undo(const MsWaveArgs& w, int k, int count) {
  const int* inl = ms_slot(w, k);
  for (int i = threadIdx.x; i < count; i += blockDim.x) atomicCAS(&w.g.labels[inl[i]], kMsTmpBase + k, kUnlabeled);
}

// The serial walk over the ranks of the wave (one CTA).  n_regions0 / base0: regions accepted and
// inliers packed before this wave.  out[] as ms_grow_kernel.
__global__ void __launch_bounds__(1024) ms_wave_resolve_kernel(MsWaveArgs w, int n_regions0, int base0) {
  __shared__ float s_acc[3 * kMsWaveMax];   // modes of the regions accepted in this wave
  __shared__ int s_flag;
  const MsGrowArgs& a = w.g;
  const int t = threadIdx.x;
  const int rows = a.rows, cols = a.cols;
  const int W = 2 * a.half_window + 1;
  const int WW = W * W;
  const int n_cand = w.wave[kWvNCand];
  int n_regions = n_regions0, base = base0, n_acc = 0;
  int k = 0;
  for (; k < n_cand; ++k) {
    const int pos = w.wave[kWvPos + k];
    const int count = w.wave[kWvCount + k];
    const bool conflict = w.wave[kWvConflict + k] != 0;
    const MsSeedRec cur = a.seeds[pos];
    // suppression by a region accepted earlier in this wave, :312-322 (earlier waves: ms_filter_kernel)
    bool sup = false;
    if (t < n_acc) {
      const V3 d{cur.mx - s_acc[3 * t], cur.my - s_acc[3 * t + 1], cur.mz - s_acc[3 * t + 2]};
      sup = (double)vsqn(d) < 1.0;
    }
    if (__syncthreads_or(sup)) {
      ms_wave_undo(w, k, count);
      if (t == 0) a.alive[pos] = 0;
      __threadfence();
      __syncthreads();
      continue;
    }
    if (!conflict) {
      if (count >= a.min_inliers) {               // CheckAndFinalize, ps/cluster_region.h:111-116
        if (n_regions < a.rcap) {
          if (t == 0) {
            MsRegionMeta mt;
            mt.inlier_begin = base; mt.inlier_count = count; mt.label_id = n_regions + a.id_offset;  // :288-289
            mt.seed[0] = cur.mx; mt.seed[1] = cur.my; mt.seed[2] = cur.mz;
            a.meta[n_regions] = mt;
            w.wave[kWvAccRank + n_acc] = k;
            w.wave[kWvAccBase + n_acc] = base;
            w.wave[kWvAccLabel + n_acc] = mt.label_id;
            s_acc[3 * n_acc] = cur.mx; s_acc[3 * n_acc + 1] = cur.my; s_acc[3 * n_acc + 2] = cur.mz;
            a.alive[pos] = 0;
          }
          ++n_regions; ++n_acc; base += count;
        } else {
          if (t == 0) { atomicOr(&a.out[1], kErrRegionCap); a.alive[pos] = 0; }
          ms_wave_undo(w, k, count);
        }
      } else {
        ms_wave_undo(w, k, count);                // :323-327
        if (t == 0) a.alive[pos] = 0;
      }
      __threadfence();
      __syncthreads();
      continue;
    }
    // invalid: undo, then decide whether the mode is a no-op on the labels as they are now
    ms_wave_undo(w, k, count);
    if (t == 0) s_flag = 0;
    __threadfence();
    __syncthreads();
    {
      const V3 mm{cur.mx, cur.my, cur.mz};
      const int crow = cur.centre % rows, ccol = cur.centre / rows;
      const int mylab = kMsTmpBase + k;
      for (int s = t; s < WW; s += blockDim.x) {
        const int rc = crow + s % W - a.half_window, cc = ccol + s / W - a.half_window;
        if (cc >= 0 && cc < cols && rc >= 0 && rc < rows) {
          const int n = cc * rows + rc;
          const int lab = __ldcg(&a.labels[n]);
          if (lab == kUnlabeled || lab > mylab) {   // higher ranks are not there yet in the serial walk
            const V3 p = ld3(a.xyz, n);
            if (!vhasnan(p) && !((double)vsqn(vsub(p, mm)) > 1.0)) s_flag = 1;
          }
        }
      }
    }
    __syncthreads();
    const bool live = s_flag != 0;
    __syncthreads();
    if (live) break;                 // has to be grown again: rank 0 of the next wave
    if (t == 0) a.alive[pos] = 0;    // dead: a no-op of the serial walk
  }
  // ranks behind the stop are undone (rank `k` itself already is)
  for (int j = k + 1; j < n_cand; ++j) ms_wave_undo(w, j, w.wave[kWvCount + j]);
  if (t == 0) {
    w.wave[kWvStop] = k;
    w.wave[kWvNAcc] = n_acc;
    a.out[0] = n_regions;
    a.out[2] = base;
    a.out[3] = k < n_cand ? w.wave[kWvPos + k] : (n_cand > 0 ? w.wave[kWvPos + n_cand - 1] + 1 : a.n_seeds);
  }
}

// Final label ids and packed inlier lists of the regions accepted in the wave.  grid = (blocks, kMsWaveMax).
__global__ void __launch_bounds__(256) ms_wave_finalize_kernel(MsWaveArgs w) {
  const int q = blockIdx.y;
  if (q >= w.wave[kWvNAcc]) return;
  const int k = w.wave[kWvAccRank + q], base = w.wave[kWvAccBase + q], label = w.wave[kWvAccLabel + q];
  const int count = w.wave[kWvCount + k];
  const int* inl = ms_slot(w, k);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
    const int n = inl[i];
    w.g.labels[n] = label;
    w.g.inliers[base + i] = n;
  }
}

}  // namespace pcseg
