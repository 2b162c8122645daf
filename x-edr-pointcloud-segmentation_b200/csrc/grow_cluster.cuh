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

// All cluster regions of frame f, grown by the calling CTA (every thread of the
// CTA must call).  Used by cluster_grow_kernel and, fused behind a frame's planar
// stage, by planar_grow_kernel (a frame's stages are sequential, different
// frames are independent: fusing removes the batch-wide barrier between them).
// H1 = true: instantiation for half_search_window == 1 (the default; 8 neighbours, lean register
// allocation so that two CTAs share an SM); H1 = false: any window up to h = 7 (the slot masks are
// kClW 64-bit words; claim key = thread * 256 + slot).
template <int BLOCK, bool H1>
__device__ __forceinline__ void cluster_grow_frame(const ClusterArgs& a, int f, int id_offset, int* s_warp) {
  const int tid = threadIdx.x;
  const int rows = a.rows, cols = a.cols;
  const int N = rows * cols;
  const float* __restrict__ xyz = a.xyz + (size_t)f * N * 3;
  volatile int* labels = a.labels + (size_t)f * N;
  volatile int* queue = a.queue + (size_t)f * a.qcap * 2;
  int* claim = a.claim + (size_t)f * N;
  volatile int* vclaim = claim;
  int* inliers = a.inliers + (size_t)f * N;
  ClusterMeta* meta = a.meta + (size_t)f * a.rcap;
  const int h = H1 ? 1 : a.half_window;
  const int side = 2 * h + 1;

  int seed_pos = a.n_seeds;  // seeds are consumed back to front
  int nreg = 0, inl_total = 0, err = 0, steps = 0;
  bool any_failed = false;

  while (true) {
    // ---- next usable seed (ps/segmentation.h:254-260).  A seed whose point is
    // not finite (or a non-positive threshold) fails its own self-test, adds no
    // inlier and leaves labels untouched, so skipping it here is exact.
    int found = -1;
    while (seed_pos > 0 && found < 0) {
      int cand = -1;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        int pos = seed_pos - 1 - (tid + k * BLOCK);
        if (pos >= 0 && cand < 0) {
          int s = seed_at(a.seeds, N, pos);
          if (labels[s] == kUnlabeled) {
            V3 p = ld3(xyz, s);
            if (vfinite(p) && (0.0f < a.sq_thr)) cand = pos;
          }
        }
      }
      // threads with smaller tid hold larger positions within each k; take max
      found = block_max<BLOCK>(cand, s_warp);
      if (found < 0) seed_pos = max(seed_pos - 4 * BLOCK, 0);
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
        const int qi = (head + tid) & (a.qcap - 1);
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
            if (nn[s] >= 0 && (L[s] == kUnlabeled || (L[s] <= kRejBase && (kRejBase - L[s]) < tid))) {
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
                  if (nn[k] >= 0 && (L[k] == kUnlabeled || (L[k] <= kRejBase && (kRejBase - L[k]) < tid))) {
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
      if (mine && !acc) labels[cand] = kUnlabeled;  // ps/cluster_region.h:105
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
          const int qi = (tail + off) & (a.qcap - 1);
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
      if (nreg < a.rcap) {
        if (tid == 0) meta[nreg] = ClusterMeta{base, ninl, label_id};
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
  if (any_failed) {
    for (int i = tid; i < N; i += BLOCK)
      if (labels[i] == kAlreadyExamedPoint) labels[i] = kUnlabeled;
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

}  // namespace pcseg
