// grow_planar_wave.cuh -- window-parallel, speculative FIFO growth of one planar
// region by the whole CTA (the fast path of planar_grow_kernel, MODE 2).
//
// Same sequential semantics as ps/planar_region.h:144-255 under the driver of
// ps/segmentation.h:265-275.  The reference pops one queue entry at a time and
// re-fits the plane at every `period`-th inlier; here a WINDOW of up to W
// consecutive queue entries (one per entry thread) is processed per step:
//
//   1  evaluate   every entry against the plane that is final at the window
//                 start (entries behind a re-fit inside the window are thereby
//                 evaluated with a stale plane: value speculation);
//   2  scan       accept ballots -> position of every inlier in the region's
//                 inlier list -> which re-fit (0 .. J) precedes every entry;
//   3a helper warp: PlaneEstimator::AddPoint of the window's inliers in FIFO
//                 order (SEQUENTIAL fp32 sums, lane j owns accumulator j), the
//                 sums posted at each of the J re-fit points; then all J
//                 eigen-solves at once, one per lane; then the eigenvector sign
//                 chain (ps/plane_estimator.cc:208-216) -> the J true planes;
//   3b entry warps, meanwhile: AddNeighbors of the accepted entries against the
//                 1-bit label mask, min-key claims, ordered append to the queue
//                 (global + shared-memory ring with cp.async prefetch of xyz) --
//                 nothing that cannot be overwritten;
//   4  verify     every entry behind a re-fit is evaluated again with its TRUE
//                 plane.  All accept bits equal (the usual case: a 30-inlier
//                 update of a plane fitted to thousands of points moves it by
//                 far less than the distance margin) -> 5; else the window is
//                 cut at the first differing entry t* and redone (entries before
//                 t* are exact by induction, entry 0 always is, so it terminates);
//   5  commit     labels, inlier list, mask updates, head/tail/plane.
//
// Chunking is invisible to the semantics; only the accept bits, the FIFO order
// of the appended neighbours and the summation order matter, and all three are
// resolved by queue position exactly as in grow_common.cuh.
#pragma once
#include "grow_planar.cuh"
#include "grow_planar_smem.cuh"

namespace pcseg {

constexpr int kWaveHash = 4096;   // claim table slots (<= 4 * W claims per window)
constexpr int kWaveRej = 1024;    // rejected-cell table slots (<= W rejected entries per window)
constexpr int kWaveMaxJ = 32;     // re-fits per window (one eigen-solve per helper lane)
constexpr int kWaveKeyBits = 11;  // claim key = entry * 4 + slot < 2048
constexpr int kWaveMaxCells = 1 << (32 - kWaveKeyBits);  // N must be smaller

struct WaveCtl {
  float plane[4];            // plane_ at the window start
  float prev_n[3];           // normal of plane_estimator_'s plane_ (sign reference)
  float centroid[3];
  float planes[kWaveMaxJ + 1][4];  // [j] = plane_ after the j-th re-fit of the window (j >= 1)
  float new_prev_n[3], new_centroid[3];
  float req[kWaveMaxJ][10];  // sums at the re-fit points
  float cur_s[10];           // sums after the window
  unsigned wacc[32];         // accept ballots per entry warp
  int wtot[32];              // winners per entry warp
  volatile int unres[2];
  int tstar;
#ifdef PCSEG_PROFILE
  long long prof[16];
#endif
};

struct WaveSmem {
  unsigned* mask;   // [mask_words]
  int* q_tag;       // [ring]
  int* q_cell;      // [ring]
  float* q_xyz;     // [ring * 3]
  float* stage;     // [W][10] AddPoint terms (xx xy xz yy yz zz x y z 1) of the window's inliers, FIFO order
  unsigned* htab;   // [kWaveHash]
  unsigned* rej;    // [kWaveRej]
  int ring;
};

// barrier of the window's active entry warps only
__device__ __forceinline__ void wave_bar_entry(int nthreads) {
  asm volatile("bar.sync 1, %0;" ::"r"(nthreads) : "memory");
}

// ComputePlane's sums-only part as a call: one copy of the code, its own register allocation
__device__ __noinline__ RawFit est_raw_call(const float* S) {
  return est_raw<true>(S, V3{S[6], S[7], S[8]}, S[9]);
}

__device__ __forceinline__ unsigned wave_hash(int cell, int size_mask) {
  return (((unsigned)cell * 2654435761u) >> 15) & (unsigned)size_mask;
}

// One region.  Every thread of the CTA calls it (BLOCK = W + 32: warps 0 .. W/32-1 hold one
// queue entry each, the last warp is the helper).  labels: the frame's label grid in HBM.
template <int BLOCK>
__device__ __forceinline__ void planar_grow_wave(const PlanarArgs& a, const float* __restrict__ xyz,
                                                 const float* __restrict__ nrm, int* labels, int* gqueue,
                                                 int* inliers, int seed, int label_id, int base, int tag_base,
                                                 const WaveSmem& sm, WaveCtl& ctl, GrowResult* out) {
  constexpr int W = BLOCK - 32;
  constexpr int NW = W / 32;
  const unsigned FULL = 0xffffffffu;
#ifdef PCSEG_WAVE_HELPER_FIRST
  // physical warp 0 is the helper; entry thread e is physical thread e + 32
  const int tid = threadIdx.x < 32 ? W + (int)threadIdx.x : (int)threadIdx.x - 32;
#else
  const int tid = threadIdx.x;
#endif
  const int lane = tid & 31, warp = tid >> 5;
  const bool helper = warp == NW;
  const int rows = a.rows, cols = a.cols;
  const int period = a.period;
  const float thr = a.max_plane_distance;
  const int R = sm.ring;
  const FastDiv fdrows = make_fastdiv(rows);
  // window width: at most kWaveMaxJ re-fits can fall inside one window
  const long long capl = (long long)kWaveMaxJ * period;
  const int wcap = capl < W ? (int)capl : W;

  int head = 0, tail = 1, ninl = 0, inper = 0, steps = 0, refits = 0, any_refit = 0;
  int force_lim = W;
  float accj = 0.0f;  // helper lane j < 10: accumulator j (acc[0..5], csum xyz, weight)

  if (tid == 0) {
    // Init, ps/planar_region.h:144-159
    const V3 sn = ld3(nrm, seed), sp = ld3(xyz, seed);
    const Plane pl = make_plane(sn, sp);
    ctl.plane[0] = pl.n.x; ctl.plane[1] = pl.n.y; ctl.plane[2] = pl.n.z; ctl.plane[3] = pl.d;
    ctl.prev_n[0] = sn.x; ctl.prev_n[1] = sn.y; ctl.prev_n[2] = sn.z;
    ctl.centroid[0] = 0.0f; ctl.centroid[1] = 0.0f; ctl.centroid[2] = 0.0f;
    ctl.unres[0] = 0; ctl.unres[1] = 0;
    ctl.tstar = INT_MAX;
    sm.q_tag[0] = tag_base;
    sm.q_cell[0] = seed;
    sm.q_xyz[0] = sp.x; sm.q_xyz[1] = sp.y; sm.q_xyz[2] = sp.z;
    gqueue[0] = seed;
  }
  __syncthreads();
#ifdef PCSEG_PROFILE
  // counters live in shared memory (in registers they change the code that is measured); slots
  // 8, 9 are written by the helper's lane 0, the others by entry thread 0
#undef PROF_ACC
#undef PROF_CNT
#define PROF_ACC(slot, t0, t1) do { if (lane == 0 && ((slot == 8 || slot == 9) ? helper : tid == 0)) ctl.prof[slot] += (t1) - (t0); } while (0)
#define PROF_CNT(slot, n) do { if (tid == 0) ctl.prof[slot] += (n); } while (0)
  if (tid < 16) ctl.prof[tid] = 0;
  __syncthreads();
#endif

  while (head < tail) {
    PROF_T(w0);
    // at most kWaveMaxJ re-fits per window: its inliers <= kWaveMaxJ * period - inper
    int m = min(min(wcap, tail - head), force_lim);
    if (capl - inper < m) m = (int)(capl - inper);
    const float acc_snap = accj;  // helper: sums before the window (restored on a mis-speculation)
    // ---- 1: evaluate with the plane of the window start (ps/planar_region.h:168)
    const Plane P{V3{ctl.plane[0], ctl.plane[1], ctl.plane[2]}, ctl.plane[3]};
    int cand = -1;
    V3 p{0.0f, 0.0f, 0.0f};
    bool acc = false;
    const int aw = (m + 31) >> 5;            // entry warps that hold entries of this window
    const bool active = warp < aw;           // (the others only meet the CTA barriers)
    const bool mine = !helper && tid < m;
    if (mine) {
      const int pos = head + tid;
      const int slot = pos & (R - 1);
      if (sm.q_tag[slot] == tag_base + pos) {
        cand = sm.q_cell[slot];
        p = V3{sm.q_xyz[3 * slot], sm.q_xyz[3 * slot + 1], sm.q_xyz[3 * slot + 2]};
      } else {
        cand = *(volatile int*)(gqueue + (pos & (a.qcap - 1)));
        p = ld3(xyz, cand);
      }
      acc = plane_abs_distance(P, p) < thr;
    }
    if (active) {
      const unsigned b = __ballot_sync(FULL, acc);
      if (lane == 0) ctl.wacc[warp] = b;
    }
    __syncthreads();  // #1
    PROF_T(w1);
    // ---- 2: scan
    int excl = 0, na = 0;
    for (int w = 0; w < aw; ++w) {
      const int c = __popc(ctl.wacc[w]);
      if (w < warp) excl += c;
      na += c;
    }
    if (active) excl += __popc(ctl.wacc[warp] & ((1u << lane) - 1u));
    const int J = (inper + na) / period;          // re-fits triggered inside the window
    const int jt = (inper + excl) / period;       // re-fits before this entry
    const bool any_rej = na < m;
    if (acc) {  // plane_estimator_.AddPoint terms (ps/plane_estimator.cc:55-65); summed by the helper
      float2* st = reinterpret_cast<float2*>(sm.stage + 10 * excl);
      st[0] = make_float2(p.x * p.x, p.x * p.y);
      st[1] = make_float2(p.x * p.z, p.y * p.y);
      st[2] = make_float2(p.y * p.z, p.z * p.z);
      st[3] = make_float2(p.x, p.y);
      st[4] = make_float2(p.z, 1.0f);
    }
    int rej_slot = -1;
    if (mine && !acc) {  // rejected: unlabeled again for LATER entries only (:181)
      unsigned h = wave_hash(cand, kWaveRej - 1);
      const unsigned v = ((unsigned)cand << kWaveKeyBits) | (unsigned)tid;
      while (atomicCAS(&sm.rej[h], kHashEmpty, v) != kHashEmpty) h = (h + 1) & (kWaveRej - 1);
      rej_slot = (int)h;
    }
    __syncthreads();  // #2
    PROF_T(w2);

    bool win[4] = {false, false, false, false};
    bool claimed[4] = {false, false, false, false};
    int nb[4] = {-1, -1, -1, -1};
    int probe[4] = {0, 0, 0, 0};
    int woff = 0;
    if (helper) {
      // ---- 3a: sums in FIFO order (ps/plane_estimator.cc:55-65), posted at the re-fit points
      const float* st = sm.stage + (lane < 10 ? lane : 0);
      int i = 0, jj = 0, nxt = period - inper;
      while (i < na) {
        const int end = min(na, i + nxt);
        const bool post = (end - i) == nxt;
        int k = i;
#pragma unroll 1
        for (; k + 10 <= end; k += 10) {
          float v[10];
#pragma unroll
          for (int u = 0; u < 10; ++u) v[u] = st[10 * (k + u)];
#pragma unroll
          for (int u = 0; u < 10; ++u) accj += v[u];
        }
#pragma unroll 1
        for (; k < end; ++k) accj += st[10 * k];
        if (post) {
          if (lane < 10) ctl.req[jj][lane] = accj;
          ++jj;
          nxt = period;
        }
        i = end;
      }
      if (lane < 10) ctl.cur_s[lane] = accj;
      __syncwarp();
      PROF_T(h1);
      PROF_ACC(8, w2, h1);
      if (J > 0) {
        // all re-fits of the window at once: ComputePlane's sums-only part, one per lane
        RawFit r;
        r.valid = true;
        r.ev = V3{0.0f, 0.0f, 0.0f};
        r.centroid = r.ev;
        if (lane < J) {
          r = est_raw_call(ctl.req[lane]);
        }
        // eigenvector sign / invalid-fit fallback depend on the previous plane (:208-226)
        const V3 pn0{ctl.prev_n[0], ctl.prev_n[1], ctl.prev_n[2]};
        const unsigned jm = J >= 32 ? FULL : ((1u << J) - 1u);
        const unsigned inval = __ballot_sync(FULL, !r.valid) & jm;
        V3 n;
        if (inval == 0u) {
          // n_j = (n_(j-1) . ev_j > 0) ? ev_j : -ev_j with n_(j-1) = s_(j-1) ev_(j-1): negation is exact,
          // so s_j = s_(j-1) * sign(d_j), d_j = ev_(j-1) . ev_j; d_j == 0 or NaN gives s_j = -1.
          V3 pr{__shfl_up_sync(FULL, r.ev.x, 1), __shfl_up_sync(FULL, r.ev.y, 1), __shfl_up_sync(FULL, r.ev.z, 1)};
          if (lane == 0) pr = pn0;
          const float d = vdot(pr, r.ev);
          const unsigned neg = __ballot_sync(FULL, d < 0.0f) & jm;
          const unsigned rst = __ballot_sync(FULL, !(d > 0.0f) && !(d < 0.0f)) & jm;
          const unsigned le = lane == 31 ? FULL : ((2u << lane) - 1u);
          const unsigned rs = rst & le;
          int flips;
          if (rs) {
            const int rp = 31 - __clz(rs);
            const unsigned above = rp == 31 ? 0u : ~((2u << rp) - 1u);
            flips = 1 + __popc(neg & le & above);
          } else {
            flips = __popc(neg & le);
          }
          n = (flips & 1) ? vneg(r.ev) : r.ev;
        } else {
          V3 pn = pn0;
          n = pn0;
          for (int j = 0; j < J; ++j) {
            const V3 ev{__shfl_sync(FULL, r.ev.x, j), __shfl_sync(FULL, r.ev.y, j), __shfl_sync(FULL, r.ev.z, j)};
            const bool valid = !((inval >> j) & 1u);
            if (valid) pn = (vdot(pn, ev) > 0.0f) ? ev : vneg(ev);
            if (lane == j) n = pn;
          }
        }
        if (lane < J) {
          const Plane pl = make_plane(n, r.centroid);
          float* dst = ctl.planes[lane + 1];
          dst[0] = pl.n.x; dst[1] = pl.n.y; dst[2] = pl.n.z; dst[3] = pl.d;
          if (lane == J - 1) {
            ctl.new_prev_n[0] = n.x; ctl.new_prev_n[1] = n.y; ctl.new_prev_n[2] = n.z;
            ctl.new_centroid[0] = r.centroid.x; ctl.new_centroid[1] = r.centroid.y; ctl.new_centroid[2] = r.centroid.z;
          }
        }
      }
      PROF_T(h2);
      PROF_ACC(9, h1, h2);
#ifdef PCSEG_PROFILE
      if (lane == 0 && J == 1) { ctl.prof[7] += h2 - h1; ctl.prof[5] += 1; }
#endif
    }
#ifdef PCSEG_WAVE_SERIAL
    __syncthreads();  // experiment: the entry warps expand only after the helper is done
#endif
    if (!helper && active) {
      // ---- 3b: AddNeighbors, ps/planar_region.h:229-255: (c-1,r), (c,r-1), (c,r+1), (c+1,r)
      bool unres[4] = {false, false, false, false};
      if (acc) {
        int row, col;
        fast_divmod(fdrows, cand, &col, &row);
        if (col > 0) nb[0] = cand - rows;
        if (row > 0) nb[1] = cand - 1;
        if (row < rows - 1) nb[2] = cand + 1;
        if (col < cols - 1) nb[3] = cand + rows;
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          if (nb[s] < 0) continue;
          bool el = mask_test(sm.mask, nb[s]);
          if (!el && any_rej) {  // rejected earlier in this window?
            unsigned h = wave_hash(nb[s], kWaveRej - 1);
            while (true) {
              const unsigned v = sm.rej[h];
              if (v == kHashEmpty) break;
              if ((int)(v >> kWaveKeyBits) == nb[s]) {
                el = (int)(v & ((1u << kWaveKeyBits) - 1u)) < tid;
                break;
              }
              h = (h + 1) & (kWaveRej - 1);
            }
          }
          unres[s] = el;
          claimed[s] = el;
          probe[s] = (int)wave_hash(nb[s], kWaveHash - 1);
        }
      }
      // the smallest (entry, slot) key per neighbour cell wins the enqueue
      int round = 0;
      while (true) {
#pragma unroll
        for (int s = 0; s < 4; ++s)
          if (unres[s]) atomicMin(&sm.htab[probe[s]], ((unsigned)nb[s] << kWaveKeyBits) | (unsigned)(tid * 4 + s));
        wave_bar_entry(aw * 32);
        bool left = false;
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          if (unres[s]) {
            const unsigned v = ((volatile unsigned*)sm.htab)[probe[s]];
            if ((int)(v >> kWaveKeyBits) == nb[s]) {
              win[s] = (v & ((1u << kWaveKeyBits) - 1u)) == (unsigned)(tid * 4 + s);
              unres[s] = false;
            } else {
              probe[s] = (probe[s] + 1) & (kWaveHash - 1);
              left = true;
            }
          }
        }
        const unsigned b0 = __ballot_sync(FULL, win[0]), b1 = __ballot_sync(FULL, win[1]);
        const unsigned b2 = __ballot_sync(FULL, win[2]), b3 = __ballot_sync(FULL, win[3]);
        const bool wleft = __any_sync(FULL, left);
        const unsigned lt = (1u << lane) - 1u;
        woff = __popc(b0 & lt) + __popc(b1 & lt) + __popc(b2 & lt) + __popc(b3 & lt);
        if (lane == 0) {
          ctl.wtot[warp] = __popc(b0) + __popc(b1) + __popc(b2) + __popc(b3);
          if (wleft) ctl.unres[round & 1] = 1;
        }
        wave_bar_entry(aw * 32);
        const bool again = ctl.unres[round & 1] != 0;
        if (tid == 0) ctl.unres[(round + 1) & 1] = 0;
        if (!again) break;
        ++round;
      }
      if (tid == 0) { ctl.unres[0] = 0; ctl.unres[1] = 0; }
      for (int w = 0; w < warp; ++w) woff += ctl.wtot[w];
      // speculative append: global queue + ring (tag, cell, xyz by cp.async)
      int qp = tail + woff;
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        if (win[s]) {
          const int n = nb[s];
          gqueue[qp & (a.qcap - 1)] = n;
          if (qp - head < R) {
            const int sl = qp & (R - 1);
            sm.q_tag[sl] = tag_base + qp;
            sm.q_cell[sl] = n;
            const float* src = xyz + 3 * (size_t)n;
            cp_async4(sm.q_xyz + 3 * sl, src);
            cp_async4(sm.q_xyz + 3 * sl + 1, src + 1);
            cp_async4(sm.q_xyz + 3 * sl + 2, src + 2);
          }
          ++qp;
        }
      }
      cp_async_commit();
      PROF_T(e1);
      PROF_ACC(10, w2, e1);
    }
    __syncthreads();  // #3
    PROF_T(w3);
    // ---- 4: verify the entries behind a re-fit with their true plane
    bool bad = false;
    if (mine && jt > 0) {
      const float* q = ctl.planes[jt];
      const Plane T{V3{q[0], q[1], q[2]}, q[3]};
      bad = (plane_abs_distance(T, p) < thr) != acc;
      if (bad) atomicMin(&ctl.tstar, tid);
    }
#ifdef PCSEG_FORCE_ROLLBACK
    if (mine && tid > 0 && tid == m - 1 && ((steps + head) % PCSEG_FORCE_ROLLBACK) == PCSEG_FORCE_ROLLBACK - 1 && force_lim == W) {
      bad = true;
      atomicMin(&ctl.tstar, tid);
    }
#endif
    // rejected cells become unlabeled (every mask test of the window is done; the winners clear
    // their bits after the next barrier)
    if (mine && !acc) atomicOr(&sm.mask[cand >> 5], 1u << (cand & 31));
    const int bad_any = __syncthreads_or(bad);  // #4
    PROF_T(w4);
    int nwin = 0;
    for (int w = 0; w < aw; ++w) nwin += ctl.wtot[w];
    if (bad_any) {
      // mis-speculation: undo the speculative state and redo the window up to the first
      // differing entry t* (entries before it are exact by induction; entry 0 always is)
      const int tstar = ctl.tstar;
      if (rej_slot >= 0) {
        atomicAnd(&sm.mask[cand >> 5], ~(1u << (cand & 31)));
        sm.rej[rej_slot] = kHashEmpty;
      }
      if (active) {
        cp_async_wait<0>();
        int qp = tail + woff;
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          if (claimed[s]) sm.htab[probe[s]] = kHashEmpty;
          if (win[s]) {
            if (qp - head < R) sm.q_tag[qp & (R - 1)] = -1;
            ++qp;
          }
        }
      }
      accj = acc_snap;
      force_lim = tstar;
      __syncthreads();
      if (tid == 0) ctl.tstar = INT_MAX;
      continue;
    }
    // ---- 5: commit
    if (active) {
      if (acc) {
        labels[cand] = label_id;                 // :177
        inliers[base + ninl + excl] = cand;      // :169
        // only the seed can still have its bit set when it is accepted
        if (head == 0 && tid == 0) atomicAnd(&sm.mask[cand >> 5], ~(1u << (cand & 31)));
      }
      if (rej_slot >= 0) sm.rej[rej_slot] = kHashEmpty;  // every lookup of the window is done (#3)
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        if (win[s]) {
          sm.htab[probe[s]] = kHashEmpty;
          atomicAnd(&sm.mask[nb[s] >> 5], ~(1u << (nb[s] & 31)));  // kAlreadyInUnexaminedPointsQueue
        }
      }
      cp_async_wait<0>();
    } else if (J > 0 && lane < 4) {
      ctl.plane[lane] = ctl.planes[J][lane];
      if (lane < 3) {
        ctl.prev_n[lane] = ctl.new_prev_n[lane];
        ctl.centroid[lane] = ctl.new_centroid[lane];
      }
    }
    head += m;
    tail += nwin;
    ninl += na;
    inper = (inper + na) % period;
    refits += J;
    any_refit |= J;
    ++steps;
    force_lim = W;
    __syncthreads();  // #5
    PROF_T(w5);
    PROF_ACC(0, w0, w1);
    PROF_ACC(1, w1, w2);
    PROF_ACC(2, w2, w3);
    PROF_ACC(3, w3, w4);
    PROF_ACC(4, w4, w5);
    PROF_CNT(6, 1);
    PROF_CNT(11, m);
    PROF_CNT(12, J);
  }
#ifdef PCSEG_PROFILE
  if (a.prof) {
    long long* pr = a.prof + (size_t)blockIdx.x * 16;
    __syncthreads();
    if (tid < 13) pr[tid] += ctl.prof[tid];
  }
#endif
  // ---- CheckAndFinalize first lines, ps/planar_region.h:190-196: Plane() / Centroid() of the
  // estimator (a ComputePlane unless the last inlier triggered one)
  if (helper) {
    Estimator est;
    est_clear(est);
#pragma unroll
    for (int i = 0; i < 6; ++i) est.acc[i] = ctl.cur_s[i];
    est.csum = V3{ctl.cur_s[6], ctl.cur_s[7], ctl.cur_s[8]};
    est.w = ctl.cur_s[9];
    const V3 prev_n{ctl.prev_n[0], ctl.prev_n[1], ctl.prev_n[2]};
    est.plane = Plane{prev_n, ctl.plane[3]};
    est.centroid = V3{ctl.centroid[0], ctl.centroid[1], ctl.centroid[2]};
    if (ninl >= a.min_inliers && inper != 0) {
      const RawFit r = est_raw_call(ctl.cur_s);
      est_apply(est, r);
    }
    else if (!any_refit) est.plane = Plane{prev_n, 0.0f};
    if (lane == 0) {
      out->ninl = ninl;
      out->tail = tail;
      out->steps = steps;
      out->refits = refits;
      out->plane[0] = est.plane.n.x; out->plane[1] = est.plane.n.y;
      out->plane[2] = est.plane.n.z; out->plane[3] = est.plane.d;
      out->centroid[0] = est.centroid.x; out->centroid[1] = est.centroid.y; out->centroid[2] = est.centroid.z;
      for (int i = 0; i < 6; ++i) out->est[i] = est.acc[i];
      out->est[6] = est.csum.x; out->est[7] = est.csum.y; out->est[8] = est.csum.z; out->est[9] = est.w;
    }
  }
}

}  // namespace pcseg
