// grow_cluster.cuh -- SegmentRegions<ClusterRegion<>> on the GPU.
//
// Replaces ps/segmentation.h:239-292 (driver) with ps/cluster_region.h:73-150
// (Init / CheckInlierAndUpdate / AddNeighbors / CheckAndFinalize) as the model.
// One CTA per frame (frames are the unit of parallelism: one FIFO, shared
// labels and sequential seeds make a single frame inherently ordered, SURVEY.md
// H1).  Each FIFO chunk of up to BLOCK entries is evaluated in parallel and
// resolved by queue position (see grow_common.cuh), which reproduces quirk Q5
// exactly: a candidate is tested only against the source that first enqueued
// it, reverts to kUnlabeled on failure and can be re-enqueued by a later entry.
#pragma once
#include "grow_common.cuh"

namespace pcseg {

struct ClusterArgs {
  const float* xyz;   // [F][N][3]
  int* labels;        // [F][N]
  int rows, cols;
  const int* seeds;   // shared list or nullptr (all descending)
  int n_seeds;
  int min_inliers;
  float sq_thr;
  int half_window;    // <= 7 (224 window slots in four 64-bit masks)
  int id_offset;      // used when id_offset_from_planar == 0
  int id_offset_from_planar;  // 1: offset = stats[f].n_planar
  int* queue;         // [F][qcap][2]
  int qcap;
  int* claim;         // [F][N], INT_MAX when idle
  int* inliers;       // [F][N]
  ClusterMeta* meta;  // [F][rcap]
  int rcap;
  FrameStats* stats;  // [F]
};

// One CTA's share of a SINGLE large frame (cluster_band_kernel): a band of consecutive columns that no
// window of the growth can leave, because at least `half_window` columns without any unlabeled finite
// cell separate it from the next band (cells there are labelled / masked, or not finite -- a non-finite
// cell can be enqueued and is always rejected, so it carries nothing across).  The seed list "all
// indices descending" is consumed back to front, i.e. in ascending index order, so the reference
// finishes every region of a lower band before it touches a higher one: the bands are grown
// concurrently and only the region ids and the order of the inlier lists are assigned afterwards, by
// band order (cluster_band_merge_kernel).  on == 0: the whole frame (the batched path).
struct ClusterBand {
  int on;
  int c_begin, c_end;   // columns [c_begin, c_end)
  int* queue;           // [qcap][2], qcap a power of two >= the band's cell count
  int qcap;
  int inl_base;         // the band's slice of the frame's inlier buffer starts here (= c_begin * rows)
  ClusterMeta* meta;    // the band's slice of the temporary region table
  int rcap;
  int* out;             // {regions, inliers, error flags, steps}
  // component mode (cluster_comp_kernel): the CTA owns ONE connected component of the frame instead of a
  // band of columns.  Its cells carry the code `unl` (= kCompUnlabeledBase + root) where a cell of the
  // frame would carry kUnlabeled, so "unlabeled AND mine" is one comparison; seeds are the indices
  // [idx_lo, idx_hi] in ascending order; the seed of every region is returned in ClusterMeta::label_id.
  int comp_on;
  int unl;
  int idx_lo, idx_hi;
  const int* comp;      // root of every participating cell: tells whose transient "rejected in this step" mark a cell carries
};
constexpr int kCompUnlabeledBase = INT_MIN;   // + root index (< 2^30): far below every other label value

// All cluster regions of frame f, grown by the calling CTA (every thread of the
// CTA must call).  Used by cluster_grow_kernel and, fused behind a frame's planar
// stage, by planar_grow_kernel (a frame's stages are sequential, different
// frames are independent: fusing removes the batch-wide barrier between them).
// H1 = true: instantiation for half_search_window == 1 (the default; 8 neighbours, lean register
// allocation so that two CTAs share an SM); H1 = false: any window up to h = 7 (the slot masks are
// kClW 64-bit words; claim key = thread * 256 + slot).
template <int BLOCK, bool H1>
__device__ __forceinline__ void cluster_grow_frame(const ClusterArgs& a, int f, int id_offset, int* s_warp,
                                                   const ClusterBand band = ClusterBand{0, 0, 0, nullptr, 0, 0, nullptr, 0, nullptr, 0, kUnlabeled, 0, 0, nullptr}) {
  const int tid = threadIdx.x;
  const int rows = a.rows, cols = a.cols;
  const int N = rows * cols;
  const float* __restrict__ xyz = a.xyz + (size_t)f * N * 3;
  volatile int* labels = a.labels + (size_t)f * N;
  volatile int* queue = band.on ? band.queue : a.queue + (size_t)f * a.qcap * 2;
  const int qmask = (band.on ? band.qcap : a.qcap) - 1;
  int* claim = a.claim + (size_t)f * N;
  volatile int* vclaim = claim;
  int* inliers = a.inliers + (size_t)f * N;
  ClusterMeta* meta = band.on ? band.meta : a.meta + (size_t)f * a.rcap;
  const int rcap = band.on ? band.rcap : a.rcap;
  const int h = H1 ? 1 : a.half_window;
  const int side = 2 * h + 1;

  // seeds are consumed back to front; a band (all-descending seeds only: seeds[k] = N-1-k) owns the
  // positions of its own cells
  int seed_pos = band.comp_on ? N - band.idx_lo : band.on ? N - band.c_begin * rows : a.n_seeds;
  const int seed_lo = band.comp_on ? N - (band.idx_hi + 1) : band.on ? N - band.c_end * rows : 0;
  const int unl = band.unl;   // what an unlabeled cell (of this CTA's component) looks like
  int nreg = 0, inl_total = band.on ? band.inl_base : 0, err = 0, steps = 0;
  bool any_failed = false;

  while (true) {
    // ---- next usable seed (ps/segmentation.h:254-260).  A seed whose point is
    // not finite (or a non-positive threshold) fails its own self-test, adds no
    // inlier and leaves labels untouched, so skipping it here is exact.
    int found = -1;
    while (seed_pos > seed_lo && found < 0) {
      int cand = -1;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        int pos = seed_pos - 1 - (tid + k * BLOCK);
        if (pos >= seed_lo && cand < 0) {
          int s = seed_at(a.seeds, N, pos);
          if (labels[s] == unl) {
            V3 p = ld3(xyz, s);
            if (vfinite(p) && (0.0f < a.sq_thr)) cand = pos;
          }
        }
      }
      // threads with smaller tid hold larger positions within each k; take max
      found = block_max<BLOCK>(cand, s_warp);
      if (found < 0) seed_pos = max(seed_pos - 4 * BLOCK, seed_lo);
    }
    if (found < 0) break;
    seed_pos = found;
    const int seed = seed_at(a.seeds, N, found);
    const int label_id = nreg + id_offset;
    const int base = inl_total;
    int head = 0, tail = 1, ninl = 0;
    if (tid == 0) {
      queue[0] = seed;
      queue[1] = seed;
    }
    __syncthreads();

    while (head < tail) {
      ++steps;
      const int m = min(BLOCK, tail - head);
      const bool mine = tid < m;
      int src = -1, cand = -1;
      bool acc = false;
      if (mine) {
        const int qi = (head + tid) & qmask;
        src = queue[2 * qi];
        cand = queue[2 * qi + 1];
        // (candidate - source).squaredNorm() < threshold, ps/cluster_region.h:94
        acc = vsqn(vsub(ld3(xyz, cand), ld3(xyz, src))) < a.sq_thr;
      }
      int nacc;
      const int rank = block_scan_excl<BLOCK>(acc ? 1 : 0, &nacc, s_warp);
      if (mine) {
        if (acc) {
          labels[cand] = label_id;
          inliers[base + ninl + rank] = cand;
        } else {
          labels[cand] = kRejBase - tid;
        }
      }
      __syncthreads();
      // AddNeighbors, ps/cluster_region.h:122-150: delta_col outer, delta_row inner.
      // All label probes of an entry are issued before the first one is used (one memory
      // round trip instead of one per neighbour); same for the claim read-back below.
      constexpr int kClW = H1 ? 1 : 4;             // 64-bit words of the slot masks
      constexpr int kKeyStride = H1 ? 64 : 256;   // claim key = tid * kKeyStride + slot
      unsigned long long el[kClW];
#pragma unroll
      for (int w = 0; w < kClW; ++w) el[w] = 0ull;
      const int row = cand % rows, col = cand / rows;
      if (acc) {
        if (H1) {
          int L[8], nn[8];
#pragma unroll
          for (int s = 0; s < 8; ++s) {
            const int cell = s + (s >= 4 ? 1 : 0);
            const int cc = col + cell / 3 - 1, rc = row + cell % 3 - 1;
            const bool in = cc >= 0 && cc < cols && rc >= 0 && rc < rows;
            nn[s] = in ? cc * rows + rc : -1;
            L[s] = in ? labels[nn[s]] : 0;
          }
#pragma unroll
          for (int s = 0; s < 8; ++s) {
            if (nn[s] >= 0 && (L[s] == unl || (L[s] <= kRejBase && (kRejBase - L[s]) < tid && (!band.comp_on || band.comp[nn[s]] == band.idx_lo)))) {
              el[0] |= 1ull << s;
              atomicMin(claim + nn[s], tid * kKeyStride + s);
            }
          }
        } else {
          // window slots in groups of kG: the label probes of a group are in flight together
          constexpr int kG = 16;
          const int nslots = side * side - 1;
#pragma unroll
          for (int w = 0; w < kClW; ++w) {
            if (w * 64 < nslots) {
#pragma unroll 1
              for (int g0 = 0; g0 < 64; g0 += kG) {
                const int s0 = w * 64 + g0;
                if (s0 >= nslots) break;
                int L[kG], nn[kG];
#pragma unroll
                for (int k = 0; k < kG; ++k) {
                  const int sl = s0 + k;
                  const int cell = sl + (sl >= (side * side) / 2 ? 1 : 0);  // skip the centre
                  const int cc = col + cell / side - h, rc = row + cell % side - h;
                  const bool in = sl < nslots && cc >= 0 && cc < cols && rc >= 0 && rc < rows;
                  nn[k] = in ? cc * rows + rc : -1;
                  L[k] = in ? labels[nn[k]] : 0;
                }
#pragma unroll
                for (int k = 0; k < kG; ++k) {
                  if (nn[k] >= 0 && (L[k] == unl || (L[k] <= kRejBase && (kRejBase - L[k]) < tid && (!band.comp_on || band.comp[nn[k]] == band.idx_lo)))) {
                    el[w] |= 1ull << (g0 + k);
                    atomicMin(claim + nn[k], tid * kKeyStride + s0 + k);
                  }
                }
              }
            }
          }
        }
      }
      __syncthreads();
      if (mine && !acc) labels[cand] = unl;  // ps/cluster_region.h:105
      unsigned long long win[kClW];
#pragma unroll
      for (int w = 0; w < kClW; ++w) win[w] = 0ull;
      if (acc) {
        if (H1) {
          int cv[8];
#pragma unroll
          for (int s = 0; s < 8; ++s) {
            const int cell = s + (s >= 4 ? 1 : 0);
            const int n = (col + cell / 3 - 1) * rows + (row + cell % 3 - 1);
            cv[s] = ((el[0] >> s) & 1ull) ? vclaim[n] : -1;
          }
#pragma unroll
          for (int s = 0; s < 8; ++s)
            if (cv[s] == tid * kKeyStride + s) win[0] |= 1ull << s;
        } else {
#pragma unroll
          for (int w = 0; w < kClW; ++w) {
            unsigned long long e = el[w];
            while (e) {  // claim read-back, 8 claimed slots per round trip
              int ss[8], cv[8];
#pragma unroll
              for (int k = 0; k < 8; ++k) {
                ss[k] = -1;
                if (e) {
                  ss[k] = w * 64 + __ffsll((long long)e) - 1;
                  e &= e - 1;
                }
              }
#pragma unroll
              for (int k = 0; k < 8; ++k) {
                cv[k] = -1;
                if (ss[k] >= 0) {
                  const int cell = ss[k] + (ss[k] >= (side * side) / 2 ? 1 : 0);  // skip the centre
                  cv[k] = vclaim[(col + cell / side - h) * rows + (row + cell % side - h)];
                }
              }
#pragma unroll
              for (int k = 0; k < 8; ++k)
                if (ss[k] >= 0 && cv[k] == tid * kKeyStride + ss[k]) win[w] |= 1ull << (ss[k] - w * 64);
            }
          }
        }
      }
      int nwin;
      int nmine = 0;
#pragma unroll
      for (int w = 0; w < kClW; ++w) nmine += __popcll(win[w]);
      int off = block_scan_excl<BLOCK>(nmine, &nwin, s_warp);  // syncs inside
#pragma unroll
      for (int w = 0; w < kClW; ++w) {
        unsigned long long ww = win[w];
        while (ww) {
          const int s = w * 64 + __ffsll((long long)ww) - 1;
          ww &= ww - 1;
          const int cell = s + (s >= (side * side) / 2 ? 1 : 0);
          const int dc = cell / side - h, dr = cell % side - h;
          const int n = (col + dc) * rows + (row + dr);
          const int qi = (tail + off) & qmask;
          queue[2 * qi] = cand;
          queue[2 * qi + 1] = n;
          labels[n] = kAlreadyInQueue;
          claim[n] = INT_MAX;
          ++off;
        }
      }
      __syncthreads();
      head += m;
      tail += nwin;
      ninl += nacc;
    }

    // CheckAndFinalize, ps/cluster_region.h:111-116
    if (ninl >= a.min_inliers) {
      if (nreg < rcap) {
        if (tid == 0) meta[nreg] = ClusterMeta{base, ninl, band.comp_on ? seed : label_id};
        ++nreg;
        inl_total += ninl;
      } else {
        err |= kErrRegionCap;
        break;
      }
    } else {
      // ps/segmentation.h:280-282
      for (int i = tid; i < ninl; i += BLOCK) labels[inliers[base + i]] = kAlreadyExamedPoint;
      any_failed = any_failed || (ninl > 0);
      __syncthreads();
    }
  }
  // ps/segmentation.h:287-291
  __syncthreads();
  if (any_failed && !band.comp_on) {   // component mode: cluster_comp_restore_kernel does it for the whole frame
    const int i0 = band.on ? band.c_begin * rows : 0, i1 = band.on ? band.c_end * rows : N;
    for (int i = i0 + tid; i < i1; i += BLOCK)
      if (labels[i] == kAlreadyExamedPoint) labels[i] = kUnlabeled;
  }
  if (band.on) {
    if (tid == 0) { band.out[0] = nreg; band.out[1] = inl_total - band.inl_base; band.out[2] = err; band.out[3] = steps; }
    return;
  }
  if (tid == 0) {
    a.stats[f].n_cluster = nreg;
    a.stats[f].cluster_inliers = inl_total;
    a.stats[f].error_flags |= err;
    a.stats[f].grow_steps += steps;
  }
}

template <int BLOCK, bool H1>
__global__ void __launch_bounds__(BLOCK, H1 ? 2 : 1) cluster_grow_kernel(ClusterArgs a) {
  __shared__ int s_warp[BLOCK / 32 + 1];
  const int f = blockIdx.x;
  cluster_grow_frame<BLOCK, H1>(a, f, a.id_offset_from_planar ? a.stats[f].n_planar : a.id_offset, s_warp);
}

// ---- one large frame: bands of columns grown concurrently (see ClusterBand) ----
struct ClusterBandTable {
  const int* c_begin;      // [n_bands], bands in the order the reference reaches them (ascending columns for the
                           // back-to-front consumption of the all-descending seed list)
  const int* c_end;
  const long long* q_off;  // queue offset of the band, in (source, candidate) pairs
  const int* q_cap;
  const int* m_off;        // offset into the temporary region table
  int* queue;              // band queues
  ClusterMeta* meta;       // temporary region table
  int* out;                // [n_bands][4]
  int rcap_total;
};

// a column is active iff it holds a cell that can become an inlier: unlabeled and finite
__global__ void __launch_bounds__(256) cluster_colflag_kernel(const float* __restrict__ xyz, const int* __restrict__ labels,
                                                              int rows, int cols, int* flag) {
  const int col = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (col >= cols) return;
  bool any = false;
  for (int r = lane; r < rows && !any; r += 32) {
    const int i = col * rows + r;
    any = labels[i] == kUnlabeled && vfinite(ld3(xyz, i));
  }
  any = __any_sync(0xffffffffu, any);
  if (lane == 0) flag[col] = any ? 1 : 0;
}

template <int BLOCK, bool H1>
__global__ void __launch_bounds__(BLOCK, H1 ? 2 : 1) cluster_band_kernel(ClusterArgs a, ClusterBandTable t) {
  __shared__ int s_warp[BLOCK / 32 + 1];
  const int bnd = blockIdx.x;
  ClusterBand band;
  band.on = 1;
  band.c_begin = t.c_begin[bnd]; band.c_end = t.c_end[bnd];
  band.queue = t.queue + 2 * t.q_off[bnd]; band.qcap = t.q_cap[bnd];
  band.inl_base = band.c_begin * a.rows;
  band.meta = t.meta + t.m_off[bnd];
  band.rcap = (bnd + 1 < (int)gridDim.x ? t.m_off[bnd + 1] : t.rcap_total) - t.m_off[bnd];
  band.out = t.out + 4 * bnd;
  band.comp_on = 0; band.unl = kUnlabeled; band.idx_lo = 0; band.idx_hi = 0; band.comp = nullptr;
  cluster_grow_frame<BLOCK, H1>(a, 0, a.id_offset, s_warp, band);   // provisional ids: renumbered by the merge
}

// region ids in the reference's order (band order, then the order inside the band), final region table
// with the inlier lists packed in that order (cluster_band_compact_kernel moves them), frame statistics.
// region_base: [2][n_bands] = regions / inliers of the bands in front
__global__ void __launch_bounds__(256) cluster_band_merge_kernel(ClusterArgs a, ClusterBandTable t, int n_bands, int* region_base) {
  __shared__ int s_total[3];
  if (threadIdx.x == 0) {
    int nreg = 0, ninl = 0, err = 0, steps = 0;
    for (int b = 0; b < n_bands; ++b) {
      region_base[b] = nreg;
      region_base[n_bands + b] = ninl;
      nreg += t.out[4 * b]; ninl += t.out[4 * b + 1]; err |= t.out[4 * b + 2]; steps += t.out[4 * b + 3];
    }
    if (nreg > a.rcap) { err |= kErrRegionCap; nreg = a.rcap; }
    a.stats[0].n_cluster = nreg;
    a.stats[0].cluster_inliers = ninl;
    a.stats[0].error_flags |= err;
    a.stats[0].grow_steps += steps;
    s_total[0] = nreg;
  }
  __syncthreads();
  const int total = s_total[0];
  for (int b = 0; b < n_bands; ++b) {
    const int n = t.out[4 * b], g0 = region_base[b];
    for (int r = threadIdx.x; r < n; r += blockDim.x) {
      if (g0 + r >= total) continue;
      ClusterMeta m = t.meta[t.m_off[b] + r];
      m.label_id = a.id_offset + g0 + r;
      m.inlier_begin = region_base[n_bands + b] + (m.inlier_begin - t.c_begin[b] * a.rows);
      a.meta[g0 + r] = m;
    }
  }
}
// the bands' inlier slices, packed in band order into tmp; then back into the frame's inlier buffer
__global__ void __launch_bounds__(256) cluster_band_compact_kernel(ClusterArgs a, ClusterBandTable t, int n_bands,
                                                                   const int* region_base, int* tmp) {
  const int b = blockIdx.x;
  const int n = t.out[4 * b + 1], src = t.c_begin[b] * a.rows, dst = region_base[n_bands + b];
  for (int i = threadIdx.x; i < n; i += blockDim.x) tmp[dst + i] = a.inliers[src + i];
}
__global__ void __launch_bounds__(256) cluster_band_copyback_kernel(ClusterArgs a, const int* tmp) {
  const int n = a.stats[0].cluster_inliers;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) a.inliers[i] = tmp[i];
}
// labels of every inlier <- the region's final id
__global__ void __launch_bounds__(256) cluster_band_relabel_kernel(ClusterArgs a) {
  const int n = a.stats[0].n_cluster;
  for (int r = blockIdx.x; r < n; r += gridDim.x) {
    const ClusterMeta m = a.meta[r];
    for (int i = threadIdx.x; i < m.inlier_count; i += blockDim.x) a.labels[a.inliers[m.inlier_begin + i]] = m.label_id;
  }
}

}  // namespace pcseg
