// grow_cluster_wave.cuh -- SegmentRegions<ClusterRegion<>> for half_search_window == 1 with the
// frame's label state in shared memory (the fast path of the cluster stage).
//
// Same sequential semantics as ps/cluster_region.h:73-150 under the driver of
// ps/segmentation.h:239-292, quirk Q5 included (a candidate is tested only against the source that
// first enqueued it, reverts to kUnlabeled on failure and can be enqueued again by a later entry).
// Same window machinery as grow_planar_wave2.cuh, without the speculation (there is no model to
// re-fit): the FIFO is walked in windows of up to BLOCK consecutive (source, candidate) entries, one
// per thread.  Per window: entry from a shared-memory ring (candidate cell + both points),
// distance test, the 8 neighbour tests against the 1-bit "unlabeled" mask, min-key claims
// (atomicMin on a shared-memory hash of (cell, entry * 8 + slot)) -- all before the first barrier;
// then the scan of the accept ballots, labels and inlier list, claim read-back and the ordered
// append of the winners: three CTA barriers and no dependent global round trip per window (the
// xyz of an eligible neighbour is loaded before the claims and stored into the ring by the winner).
// cluster_grow_kernel (grow_cluster.cuh: labels, claims and queue through L2) remains the path for
// h > 1 and for clouds whose mask does not fit shared memory.
#pragma once
#include "grow_cluster.cuh"
#include "grow_planar_wave2.cuh"

namespace pcseg {

constexpr int kCwBlock = 256;      // entries per window = threads
constexpr int kCwHash = 4096;      // claim table slots (<= 8 * 256 claims per window)
constexpr int kCwRej = 512;        // rejected-cell table slots
constexpr int kCwKeyBits = 11;     // claim key = entry * 8 + slot < 2048
constexpr int kCwMaxCells = 1 << (32 - kCwKeyBits);

struct CwCtl {
  alignas(16) unsigned wacc[8];
  alignas(16) int wtot[8];
  volatile int unres[2];
  int nwin[2];   // winners of the window, by window parity
};

// dynamic shared memory: mask[mask_words] | q_tag[ring] | q_cell[ring] | q_xyz[ring * 6] (candidate, source) |
// htab[kCwHash] | rej[kCwRej]
__global__ void __launch_bounds__(kCwBlock, 2) cluster_wave_kernel(ClusterArgs a, int mask_words, int ring) {
  extern __shared__ __align__(16) unsigned char cw_smem[];
  constexpr int BLOCK = kCwBlock;
  const unsigned FULL = 0xffffffffu;
  const int f = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int rows = a.rows, cols = a.cols;
  const int N = rows * cols;
  const float* __restrict__ xyz = a.xyz + (size_t)f * N * 3;
  int* labels = a.labels + (size_t)f * N;
  int* gqueue = a.queue + (size_t)f * a.qcap * 2;
  int* inliers = a.inliers + (size_t)f * N;
  ClusterMeta* meta = a.meta + (size_t)f * a.rcap;
  const int id_offset = a.id_offset_from_planar ? a.stats[f].n_planar : a.id_offset;
  const float thr = a.sq_thr;
  const int R = ring;
  const FastDiv fdrows = make_fastdiv(rows);

  unsigned* mask = reinterpret_cast<unsigned*>(cw_smem);
  int* q_tag = reinterpret_cast<int*>(mask + mask_words);
  int* q_cell = q_tag + ring;
  float* q_xyz = reinterpret_cast<float*>(q_cell + ring);
  unsigned* htab = reinterpret_cast<unsigned*>(q_xyz + 6 * ring);
  unsigned* rej = htab + kCwHash;
  __shared__ CwCtl ctl;
  __shared__ int s_warp[BLOCK / 32 + 1];

  for (int i = tid; i < ring; i += BLOCK) q_tag[i] = -1;
  for (int i = tid; i < kCwHash; i += BLOCK) htab[i] = kHashEmpty;
  for (int i = tid; i < kCwRej; i += BLOCK) rej[i] = kHashEmpty;
  if (tid == 0) { ctl.unres[0] = 0; ctl.unres[1] = 0; }
  // mask bit <=> labels[cell] == kUnlabeled
  for (int b0 = warp * 32; b0 < mask_words * 32; b0 += BLOCK) {
    const int cell = b0 + lane;
    const bool u = cell < N && labels[cell] == kUnlabeled;
    const unsigned bits = __ballot_sync(FULL, u);
    if (lane == 0) mask[b0 >> 5] = bits;
  }
  __syncthreads();

  int seed_pos = a.n_seeds;  // seeds are consumed back to front
  int nreg = 0, inl_total = 0, err = 0, steps = 0, tag_base = 0;
  bool any_failed = false;

  while (true) {
    // ---- next usable seed (ps/segmentation.h:254-260); a seed whose point is not finite (or a
    // non-positive threshold) fails its own self-test and leaves labels untouched: skipping is exact
    int found = -1;
    while (seed_pos > 0 && found < 0) {
      int cnd = -1;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int pos = seed_pos - 1 - (tid + k * BLOCK);
        if (pos >= 0 && cnd < 0) {
          const int s = seed_at(a.seeds, N, pos);
          if (mask_test(mask, s) && vfinite(ld3(xyz, s)) && (0.0f < thr)) cnd = pos;
        }
      }
      found = block_max<BLOCK>(cnd, s_warp);
      if (found < 0) seed_pos = max(seed_pos - 4 * BLOCK, 0);
    }
    if (found < 0) break;
    seed_pos = found;
    const int seed = seed_at(a.seeds, N, found);
    const int label_id = nreg + id_offset;
    const int base = inl_total;
    int head = 0, tail = 1, ninl = 0, k = 0;
    if (tid == 0) {  // Init, ps/cluster_region.h:73-81: the pair (seed, seed); the seed is NOT marked as queued
      const V3 sp = ld3(xyz, seed);
      q_tag[0] = tag_base;
      q_cell[0] = seed;
      q_xyz[0] = sp.x; q_xyz[1] = sp.y; q_xyz[2] = sp.z; q_xyz[3] = sp.x; q_xyz[4] = sp.y; q_xyz[5] = sp.z;
      gqueue[0] = seed; gqueue[1] = seed;
    }
    __syncthreads();

    while (head < tail) {
      ++steps;
      const int m = min(BLOCK, tail - head);
      const int aw = (m + 31) >> 5;
      const bool active = warp < aw;
      const bool mine = tid < m;
      int cand = -1;
      V3 p{0.0f, 0.0f, 0.0f};
      bool acc = false;
      int rej_slot = -1;
      unsigned unres = 0u, tried = 0u, win = 0u;   // bit s: neighbour slot s
      int nb[8], probe[8];
      V3 nx[8];
#pragma unroll
      for (int s = 0; s < 8; ++s) { nb[s] = -1; probe[s] = 0; }
      if (mine) {
        const int pos = head + tid;
        const int slot = pos & (R - 1);
        V3 ps;
        if (q_tag[slot] == tag_base + pos) {
          cand = q_cell[slot];
          const float* q = q_xyz + 6 * slot;
          p = V3{q[0], q[1], q[2]};
          ps = V3{q[3], q[4], q[5]};
        } else {
          const int qi = pos & (a.qcap - 1);
          const int src = *(volatile int*)(gqueue + 2 * qi);
          cand = *(volatile int*)(gqueue + 2 * qi + 1);
          p = ld3(xyz, cand);
          ps = ld3(xyz, src);
        }
        // (candidate - source).squaredNorm() < threshold, ps/cluster_region.h:94
        acc = vsqn(vsub(p, ps)) < thr;
        if (acc) {
          // AddNeighbors, ps/cluster_region.h:122-150: delta_col outer, delta_row inner, centre skipped
          int row, col;
          fast_divmod(fdrows, cand, &col, &row);
#pragma unroll
          for (int s = 0; s < 8; ++s) {
            const int c9 = s + (s >= 4 ? 1 : 0);
            const int cc = col + c9 / 3 - 1, rc = row + c9 % 3 - 1;
            if (cc < 0 || cc >= cols || rc < 0 || rc >= rows) continue;
            nb[s] = cc * rows + rc;
            if (mask_test(mask, nb[s])) {
              unres |= 1u << s;
              nx[s] = ld3(xyz, nb[s]);  // in flight during the claims; the winner stores it into the ring
              probe[s] = (int)w2_hash(nb[s], kCwHash - 1);
              atomicMin(&htab[probe[s]], ((unsigned)nb[s] << kCwKeyBits) | (unsigned)(tid * 8 + s));
            } else {
              tried |= 1u << s;
            }
          }
        } else {  // rejected: unlabeled again for LATER entries only, ps/cluster_region.h:105
          unsigned h = w2_hash(cand, kCwRej - 1);
          const unsigned v = ((unsigned)cand << kCwKeyBits) | (unsigned)tid;
          while (atomicCAS(&rej[h], kHashEmpty, v) != kHashEmpty) h = (h + 1) & (kCwRej - 1);
          rej_slot = (int)h;
        }
      }
      if (active) {
        const unsigned b = __ballot_sync(FULL, acc);
        if (lane == 0) ctl.wacc[warp] = b;
      } else if (lane == 0) {
        ctl.wacc[warp] = 0u;
        ctl.wtot[warp] = 0;
      }
      __syncthreads();  // B1
      // ---- scan of the accept bits
      const uint4 wa0 = *reinterpret_cast<const uint4*>(&ctl.wacc[0]);
      const uint4 wa1 = *reinterpret_cast<const uint4*>(&ctl.wacc[4]);
      const unsigned wv[8] = {wa0.x, wa0.y, wa0.z, wa0.w, wa1.x, wa1.y, wa1.z, wa1.w};
      int excl = 0, na = 0;
#pragma unroll
      for (int w = 0; w < 8; ++w) {
        const int c = __popc(wv[w]);
        if (w < warp) excl += c;
        if (w == warp) excl += __popc(wv[w] & ((1u << lane) - 1u));
        na += c;
      }
      const bool any_rej = na < m;
      int nwin = 0;
      if (active) {
        if (acc) {
          labels[cand] = label_id;                 // ps/cluster_region.h:98-100
          inliers[base + ninl + excl] = cand;
          // only the seed can still have its bit set when it is accepted
          if (k == 0 && tid == 0) atomicAnd(&mask[cand >> 5], ~(1u << (cand & 31)));
        }
        // every mask test of the window is done (B1): its rejected cells are unlabeled again
        if (rej_slot >= 0) atomicOr(&mask[cand >> 5], 1u << (cand & 31));
        if (any_rej) {
          // a neighbour that failed the mask test may have been rejected EARLIER in this window
          if (acc) {
#pragma unroll
            for (int s = 0; s < 8; ++s) {
              if (!((tried >> s) & 1u)) continue;
              unsigned h = w2_hash(nb[s], kCwRej - 1);
              while (true) {
                const unsigned v = rej[h];
                if (v == kHashEmpty) break;
                if ((int)(v >> kCwKeyBits) == nb[s]) {
                  if ((int)(v & ((1u << kCwKeyBits) - 1u)) < tid) {
                    unres |= 1u << s;
                    nx[s] = ld3(xyz, nb[s]);
                    probe[s] = (int)w2_hash(nb[s], kCwHash - 1);
                    atomicMin(&htab[probe[s]], ((unsigned)nb[s] << kCwKeyBits) | (unsigned)(tid * 8 + s));
                  }
                  break;
                }
                h = (h + 1) & (kCwRej - 1);
              }
            }
          }
          w2_bar(2, aw * 32);  // B3: the late claims are in, every rejected-table lookup is done
          if (rej_slot >= 0) rej[rej_slot] = kHashEmpty;
        }
        // ---- claim resolution: the smallest (entry, slot) key per neighbour cell wins the enqueue
        int round = 0, woff = 0;
        while (true) {
          unsigned left = 0u;
#pragma unroll
          for (int s = 0; s < 8; ++s) {
            if ((unres >> s) & 1u) {
              const unsigned v = ((volatile unsigned*)htab)[probe[s]];
              if ((int)(v >> kCwKeyBits) == nb[s]) {
                if ((v & ((1u << kCwKeyBits) - 1u)) == (unsigned)(tid * 8 + s)) win |= 1u << s;
                unres &= ~(1u << s);
              } else {  // slot taken by a smaller cell: next slot, one more round
                probe[s] = (probe[s] + 1) & (kCwHash - 1);
                left |= 1u << s;
              }
            }
          }
          // winners in key order: thread-major, slot-minor
          int mywins = __popc(win);
          int incl = mywins;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const int y = __shfl_up_sync(FULL, incl, o);
            if (lane >= o) incl += y;
          }
          woff = incl - mywins;
          const bool wleft = __any_sync(FULL, left != 0u);
          if (lane == 31) ctl.wtot[warp] = incl;
          if (lane == 0 && wleft) ctl.unres[round & 1] = 1;
          w2_bar(2, aw * 32);  // B4
          const bool again = ctl.unres[round & 1] != 0;
          if (tid == 0) ctl.unres[(round + 1) & 1] = 0;
          if (!again) break;
          ++round;
#pragma unroll
          for (int s = 0; s < 8; ++s)
            if ((unres >> s) & 1u) atomicMin(&htab[probe[s]], ((unsigned)nb[s] << kCwKeyBits) | (unsigned)(tid * 8 + s));
          w2_bar(2, aw * 32);
        }
        if (tid == 0) { ctl.unres[0] = 0; ctl.unres[1] = 0; }
        {
          const int4 t0 = *reinterpret_cast<const int4*>(&ctl.wtot[0]);
          const int4 t1 = *reinterpret_cast<const int4*>(&ctl.wtot[4]);
          const int tv[8] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w};
#pragma unroll
          for (int w = 0; w < 8; ++w) {
            const int c = w < aw ? tv[w] : 0;
            if (w < warp) woff += c;
            nwin += c;
          }
        }
        // ---- ordered append of the winners: (source = this entry's candidate, candidate = neighbour)
        int qp = tail + woff;
#pragma unroll
        for (int s = 0; s < 8; ++s) {
          if ((win >> s) & 1u) {
            const int n = nb[s];
            htab[probe[s]] = kHashEmpty;
            atomicAnd(&mask[n >> 5], ~(1u << (n & 31)));  // kAlreadyInUnexaminedPointsQueue
            const int qi = qp & (a.qcap - 1);
            gqueue[2 * qi] = cand;
            gqueue[2 * qi + 1] = n;
            if (qp - head < R) {
              const int sl = qp & (R - 1);
              q_tag[sl] = tag_base + qp;
              q_cell[sl] = n;
              float* q = q_xyz + 6 * sl;
              q[0] = nx[s].x; q[1] = nx[s].y; q[2] = nx[s].z; q[3] = p.x; q[4] = p.y; q[5] = p.z;
            }
            ++qp;
          }
        }
        if (tid == 0) ctl.nwin[k & 1] = nwin;
      }
      __syncthreads();  // B5
      nwin = ctl.nwin[k & 1];
      head += m;
      tail += nwin;
      ninl += na;
      ++k;
    }
    tag_base += tail;

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
      // ps/segmentation.h:280-282 (parked cells are neither unlabeled nor queued: their mask bits stay clear)
      for (int i = tid; i < ninl; i += BLOCK) labels[inliers[base + i]] = kAlreadyExamedPoint;
      any_failed = any_failed || (ninl > 0);
    }
    __syncthreads();
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

}  // namespace pcseg
