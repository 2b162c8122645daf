// grow_planar_wave2.cuh -- window-parallel FIFO growth of one planar region with the
// plane re-fits OFF the critical path (the fast path of planar_grow_kernel, MODE 3).
//
// Same sequential semantics as ps/planar_region.h:144-255 under the driver of
// ps/segmentation.h:265-275.  Roles inside the frame's CTA (BLOCK = W + 64 threads):
//
//   entry warps (W threads, one queue entry each) walk the FIFO in WINDOWS of up to W
//     consecutive entries: evaluate all of them with the newest plane the re-fit warp has
//     published (usually one or two windows stale: value speculation), scan the accept bits
//     (inlier positions, re-fit points), stage the AddPoint terms for warp S, expand the
//     accepted entries (AddNeighbors against the 1-bit label mask, min-key claims, ordered
//     append to the queue) and COMMIT at once: labels, inlier list, mask.  Each thread keeps
//     its entry (cell, xyz, accept bit, number of re-fits in front of it) of the last two
//     windows in registers; when the planes of an older window are final every thread
//     evaluates its entry of that window again with the TRUE plane.  All accept bits equal
//     (almost always: one more 30-inlier update of a plane fitted to thousands of points
//     moves it far less than the distance margin) => the window is verified.  Otherwise the
//     unverified windows are undone in reverse order from their records (appended cells ->
//     unlabeled, rejected entries -> queued again, accepted entries -> kUnlabeled, tail,
//     sums) and the growth resumes at the failed window, cut at the first differing entry.
//   warp S  adds the staged terms in FIFO order (sequential fp32, lane j = accumulator j,
//     ps/plane_estimator.cc:55-65) and posts the sums at every re-fit point.
//   warp R  solves the posted re-fits, up to 32 at once (one per lane), applies the
//     eigenvector sign chain in order (ps/plane_estimator.cc:208-226) and publishes the
//     planes by version (= number of re-fits applied).
//
// Exactness: windows are verified in queue order, so when window j is checked every plane
// it is checked against derives from verified accept bits only (induction over the queue);
// a redone window starts from a verified plane and entry 0 never needs a re-fit of its own
// window, so the cut position strictly shrinks until the window passes.
#pragma once
#include "grow_planar.cuh"
#include "grow_planar_smem.cuh"

#ifndef PCSEG_W2_HELPER_NS
#define PCSEG_W2_HELPER_NS 20   // back-off of the idle warps S and R between polls
#endif

namespace pcseg {

#ifdef PCSEG_W2_SMALL
// footprint for three frames per SM (BLOCK 224 at <= 96 registers, ~72 KB of shared memory for C1)
constexpr int kW2Hash = 2048;    // claim table slots
constexpr int kW2Rej = 256;      // rejected-cell table slots (<= W rejected entries per window)
constexpr int kW2MaxJ = 15;      // re-fits per window
constexpr int kW2Ver = 64;       // plane / request ring (versions in flight <= 4 * kW2MaxJ + 1)
#else
constexpr int kW2Hash = 2048;    // claim table slots
constexpr int kW2Rej = 512;      // rejected-cell table slots
constexpr int kW2MaxJ = 16;      // re-fits per window
constexpr int kW2Ver = 128;      // plane / request ring (versions in flight <= 4 * kW2MaxJ + 1)
#endif
constexpr int kW2Rec = 4;        // window records (<= 3 in flight)
constexpr int kW2KeyBits = 10;   // claim key = entry * 4 + slot < 1024  (W <= 256)
constexpr int kW2MaxCells = 1 << (32 - kW2KeyBits);

struct W2Rec {
  int head, m, tailb, nwin, ninl0, inper0, g0, J;
  unsigned accb[8];   // accept ballots per entry warp
  float sums0[10];    // warp S: sums before the window
};

struct W2Ctl {
  // entry leader -> helpers
  volatile int e_pub;       // windows [0, e_pub) are staged and recorded
  volatile int rb_epoch;    // bumped on a roll-back to window rb_k / re-fit rb_g
  volatile int rb_k, rb_g;
  volatile int done;
  // S -> R, entry
  volatile int s_posted;    // re-fit requests [0, s_posted) are in req[]
  volatile int s_ninl;      // staged inliers consumed
  volatile int s_k;         // windows consumed
  volatile int s_ack;
  // R -> entry
  volatile int r_done;      // planes of versions [0, r_done] are final (given the speculation)
  volatile int r_ack;
  float planes[kW2Ver][4];  // plane_ after `version` re-fits
  float cents[kW2Ver][3];
  float req[kW2Ver][10];
  float cur_s[10];          // warp S: sums after the last consumed window
  W2Rec rec[kW2Rec];
  // entry-side scratch
  alignas(16) unsigned wacc[8];   // accept ballots per entry warp (read as two uint4)
  alignas(16) int wtot[8];        // winners per entry warp
  volatile int unres[2];
  int tstar;
  volatile int wd;          // a bounded wait of the leader ran out (internal error): everyone leaves
  int bcast[2][4];          // leader -> entry threads, by iteration parity: ready, nwin, plane version to use
#ifdef PCSEG_PROFILE
  long long prof[16];
#endif
};

struct W2Smem {
  unsigned* mask;   // [mask_words]
  int* q_tag;       // [ring]
  int* q_cell;      // [ring]
  float* q_xyz;     // [ring * 3]
  float* stage;     // [nstage][10] AddPoint terms of the inliers, ring by inlier ordinal
  unsigned* htab;   // [kW2Hash]
  unsigned* rej;    // [kW2Rej]
  int ring, nstage; // powers of two
};

__device__ __forceinline__ void w2_bar(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
// barrier + OR reduction of a predicate over the threads of the barrier
__device__ __forceinline__ bool w2_bar_or(int id, int nthreads, bool pred) {
  unsigned r;
  asm volatile(
      "{ .reg .pred p, q; setp.ne.u32 q, %1, 0; bar.red.or.pred p, %2, %3, q; selp.u32 %0, 1, 0, p; }"
      : "=r"(r) : "r"((unsigned)pred), "r"(id), "r"(nthreads) : "memory");
  return r != 0u;
}
__device__ __forceinline__ unsigned w2_hash(int cell, int size_mask) {
  return (((unsigned)cell * 2654435761u) >> 15) & (unsigned)size_mask;
}
__device__ __noinline__ RawFit w2_est_raw_call(const float* S) {
  return est_raw<true>(S, V3{S[6], S[7], S[8]}, S[9]);
}

// ---------------------------------------------------------------------------
// warp S
// ---------------------------------------------------------------------------
__device__ __noinline__ void w2_warp_S(W2Ctl& ctl, const float* stage, int nstage, int period) {
  const int lane = threadIdx.x & 31;
  const float* st = stage + (lane < 10 ? lane : 0);
  float accj = 0.0f;
  int k = 0, g = 0, ord = 0, epoch = 0;  // next window, next re-fit index, inliers consumed
  while (true) {
    const int e = ctl.rb_epoch;
    if (e != epoch) {
      __threadfence_block();
      const int rk = ctl.rb_k;
      if (k > rk) {
        const W2Rec& rr = ctl.rec[rk & (kW2Rec - 1)];
        k = rk;
        accj = lane < 10 ? rr.sums0[lane] : 0.0f;
        ord = rr.ninl0;
        g = rr.g0;
      }
      epoch = e;
      __syncwarp();
      if (lane == 0) {
        ctl.s_posted = g;
        ctl.s_ninl = ord;
        ctl.s_k = k;
        __threadfence_block();
        ctl.s_ack = e;
      }
      continue;
    }
    if (ctl.e_pub <= k) {
      if (ctl.done) break;
      __nanosleep(PCSEG_W2_HELPER_NS);
      continue;
    }
    __threadfence_block();
    W2Rec& rc = ctl.rec[k & (kW2Rec - 1)];
    const int inper0 = rc.inper0;
    int na = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) na += __popc(rc.accb[w]);
    if (lane < 10) rc.sums0[lane] = accj;
    int i = 0, nxt = period - inper0;
    while (i < na) {
      const int end = min(na, i + nxt);
      const bool post = (end - i) == nxt;
      int q = i;
#pragma unroll 1
      for (; q + 10 <= end; q += 10) {
        float v[10];
#pragma unroll
        for (int u = 0; u < 10; ++u) v[u] = st[10 * ((ord + q + u) & (nstage - 1))];
#pragma unroll
        for (int u = 0; u < 10; ++u) accj += v[u];
      }
#pragma unroll 1
      for (; q < end; ++q) accj += st[10 * ((ord + q) & (nstage - 1))];
      if (post) {
        if (lane < 10) ctl.req[g & (kW2Ver - 1)][lane] = accj;
        ++g;
        nxt = period;
        __syncwarp();
        if (lane == 0) {
          __threadfence_block();
          ctl.s_posted = g;
        }
      }
      i = end;
    }
    ord += na;
    ++k;
    if (lane < 10) ctl.cur_s[lane] = accj;
    __syncwarp();
    if (lane == 0) {
      __threadfence_block();
      ctl.s_ninl = ord;
      ctl.s_k = k;
    }
  }
}

// ---------------------------------------------------------------------------
// warp R
// ---------------------------------------------------------------------------
__device__ __noinline__ void w2_warp_R(W2Ctl& ctl) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  int g = 0, epoch = 0;  // re-fits [0, g) are applied: planes of versions [0, g] are published
  V3 pn{ctl.planes[0][0], ctl.planes[0][1], ctl.planes[0][2]};  // normal of plane_estimator_'s plane_
  while (true) {
    const int e = ctl.rb_epoch;
    if (e != epoch) {
      // warp S first: it withdraws the requests behind the roll-back point (s_posted) before its ack
      if (ctl.s_ack != e) {
        if (ctl.done) break;
        __nanosleep(PCSEG_W2_HELPER_NS);
        continue;
      }
      __threadfence_block();
      const int rg = ctl.rb_g;
      if (g > rg) {
        g = rg;
        const float* pl = ctl.planes[g & (kW2Ver - 1)];
        pn = V3{pl[0], pl[1], pl[2]};
      }
      epoch = e;
      if (lane == 0) {
        ctl.r_done = g;
        __threadfence_block();
        ctl.r_ack = e;
      }
      __syncwarp();
      continue;
    }
    const int posted = ctl.s_posted;
    if (posted <= g) {
      if (ctl.done) break;
      __nanosleep(PCSEG_W2_HELPER_NS);
      continue;
    }
    __threadfence_block();
    const int J = min(32, posted - g);
    RawFit r;
    r.valid = true;
    r.ev = V3{0.0f, 0.0f, 0.0f};
    r.centroid = r.ev;
    if (lane < J) r = w2_est_raw_call(ctl.req[(g + lane) & (kW2Ver - 1)]);
    const unsigned jm = J >= 32 ? FULL : ((1u << J) - 1u);
    const unsigned inval = __ballot_sync(FULL, !r.valid) & jm;
    V3 n;
    if (inval == 0u) {
      // n_j = (n_(j-1) . ev_j > 0) ? ev_j : -ev_j with n_(j-1) = s_(j-1) ev_(j-1): negation is exact,
      // so s_j = s_(j-1) * sign(d_j), d_j = ev_(j-1) . ev_j; d_j == 0 or NaN gives s_j = -1.
      V3 pr{__shfl_up_sync(FULL, r.ev.x, 1), __shfl_up_sync(FULL, r.ev.y, 1), __shfl_up_sync(FULL, r.ev.z, 1)};
      if (lane == 0) pr = pn;
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
      V3 q = pn;
      n = pn;
      for (int j = 0; j < J; ++j) {
        const V3 ev{__shfl_sync(FULL, r.ev.x, j), __shfl_sync(FULL, r.ev.y, j), __shfl_sync(FULL, r.ev.z, j)};
        if (!((inval >> j) & 1u)) q = (vdot(q, ev) > 0.0f) ? ev : vneg(ev);
        if (lane == j) n = q;
      }
    }
    if (lane < J) {
      const Plane pl = make_plane(n, r.centroid);
      const int v = (g + lane + 1) & (kW2Ver - 1);
      ctl.planes[v][0] = pl.n.x; ctl.planes[v][1] = pl.n.y; ctl.planes[v][2] = pl.n.z; ctl.planes[v][3] = pl.d;
      ctl.cents[v][0] = r.centroid.x; ctl.cents[v][1] = r.centroid.y; ctl.cents[v][2] = r.centroid.z;
    }
    pn = V3{__shfl_sync(FULL, n.x, J - 1), __shfl_sync(FULL, n.y, J - 1), __shfl_sync(FULL, n.z, J - 1)};
    g += J;
    __syncwarp();
    if (lane == 0) {
      __threadfence_block();
      if (ctl.rb_epoch == epoch) ctl.r_done = g;  // (a roll-back in flight re-publishes after its ack)
    }
  }
}

// per-thread copy of its entry in an unverified window
struct W2Saved {
  int cand;
  V3 p;
  int meta;   // bit 0: accepted, bit 1: entry exists, bits 2..: re-fits of the window in front of it
};

// ---------------------------------------------------------------------------
// One region.  Every thread of the CTA calls it (BLOCK = W + 64: warps 0 .. W/32-1 are the
// entry warps, then warp S, then warp R).  labels: the frame's label grid in HBM.
// ---------------------------------------------------------------------------
template <int BLOCK>
__device__ __forceinline__ void planar_grow_wave2(const PlanarArgs& a, const float* __restrict__ xyz,
                                                  const float* __restrict__ nrm, int* labels, int* gqueue,
                                                  int* inliers, int seed, int label_id, int base, int tag_base,
                                                  const W2Smem& sm, W2Ctl& ctl, GrowResult* out) {
  constexpr int W = BLOCK - 64;
  constexpr int NW = W / 32;
  static_assert(NW <= 8 && W * 4 <= (1 << kW2KeyBits), "entry warps / claim key bits");
  const unsigned FULL = 0xffffffffu;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int rows = a.rows, cols = a.cols;
  const int period = a.period;
  const float inv_period = 1.0f / (float)period;
  const float thr = a.max_plane_distance;
  const int R = sm.ring;
  const int NS = sm.nstage;
  const FastDiv fdrows = make_fastdiv(rows);
  const long long capl = (long long)kW2MaxJ * period;   // inliers of a window <= capl - inper

  if (tid == 0) {
    // Init, ps/planar_region.h:144-159
    const V3 sn = ld3(nrm, seed), sp = ld3(xyz, seed);
    const Plane pl = make_plane(sn, sp);
    ctl.planes[0][0] = pl.n.x; ctl.planes[0][1] = pl.n.y; ctl.planes[0][2] = pl.n.z; ctl.planes[0][3] = pl.d;
    ctl.cents[0][0] = 0.0f; ctl.cents[0][1] = 0.0f; ctl.cents[0][2] = 0.0f;
    ctl.e_pub = 0; ctl.rb_epoch = 0; ctl.rb_k = 0; ctl.rb_g = 0; ctl.done = 0;
    ctl.s_posted = 0; ctl.s_ninl = 0; ctl.s_k = 0; ctl.s_ack = 0;
    ctl.r_done = 0; ctl.r_ack = 0;
    ctl.unres[0] = 0; ctl.unres[1] = 0;
    ctl.tstar = INT_MAX;
    ctl.wd = 0;
    for (int i = 0; i < 4; ++i) { ctl.bcast[0][i] = 0; ctl.bcast[1][i] = 0; }
    for (int i = 0; i < 10; ++i) ctl.cur_s[i] = 0.0f;
    sm.q_tag[0] = tag_base;
    sm.q_cell[0] = seed;
    sm.q_xyz[0] = sp.x; sm.q_xyz[1] = sp.y; sm.q_xyz[2] = sp.z;
    gqueue[0] = seed;
#ifdef PCSEG_PROFILE
    for (int i = 0; i < 16; ++i) ctl.prof[i] = 0;
#endif
  }
  __syncthreads();

  int ninl = 0, inper = 0, gtot = 0, steps = 0, tail = 1;
  if (warp == NW) {
    w2_warp_S(ctl, sm.stage, NS, period);
  } else if (warp == NW + 1) {
    w2_warp_R(ctl);
  } else {
#ifdef PCSEG_PROFILE
#undef PROF_ACC
#undef PROF_CNT
#define PROF_ACC(slot, t0, t1) do { if (tid == 0) ctl.prof[slot] += (t1) - (t0); } while (0)
#define PROF_CNT(slot, n) do { if (tid == 0) ctl.prof[slot] += (n); } while (0)
#endif
    int head = 0, k = 0, kv = 0, force_lim = W, epoch = 0, it = 0;
    const int it_limit = (int)min((long long)INT_MAX - 1, 8LL * rows * cols + 4096);
    bool watchdog = false;
    int ready = 0, vcur = 0;
    int used1 = 0, used2 = 0, used3 = 0;   // plane version windows k-1, k-2, k-3 were evaluated with
    W2Saved s1, s2, s3;                    // this thread's entries of windows k-1, k-2, k-3
    s1.cand = s2.cand = s3.cand = -1;
    s1.meta = s2.meta = s3.meta = 0;
    s1.p = s2.p = s3.p = V3{0.0f, 0.0f, 0.0f};
#ifdef PCSEG_FORCE_ROLLBACK
    int forced_k = -1;
#endif
    while (true) {
      PROF_T(w0);
      // (a growth that does not end is a bug, never an input property: leave instead of hanging the GPU)
      if (++it > it_limit) { watchdog = true; break; }
      int* bc = ctl.bcast[it & 1];
      if (head == tail && k == kv) break;

      // ---- phase 1 (up to barrier B1): verification compare of the oldest open window (its planes
      // are final), and of the next window: entry, plane test, neighbour mask tests, first claims
      bool bad = false;
      if (ready) {
        const int age = k - kv;  // 1 .. 3
        const W2Saved sv = age == 1 ? s1 : (age == 2 ? s2 : s3);
        const int used = age == 1 ? used1 : (age == 2 ? used2 : used3);
        const W2Rec& rv = ctl.rec[kv & (kW2Rec - 1)];
        if (sv.meta & 2) {
          const int ver = rv.g0 + (sv.meta >> 2);
          if (ver != used) {
            const float* q = ctl.planes[ver & (kW2Ver - 1)];
            const Plane T{V3{q[0], q[1], q[2]}, q[3]};
            bad = (plane_abs_distance(T, sv.p) < thr) != (bool)(sv.meta & 1);
          }
#ifdef PCSEG_FORCE_ROLLBACK
          if (tid > 0 && tid == rv.m - 1 && kv != forced_k && (kv % PCSEG_FORCE_ROLLBACK) == PCSEG_FORCE_ROLLBACK - 1) bad = true;
#endif
          if (bad) atomicMin(&ctl.tstar, tid);
        }
      }
      int m = 0;
      if (head < tail && k - kv < 3 + (ready ? 1 : 0)) {
        m = min(min(W, tail - head), force_lim);
        if (capl - inper < m) m = (int)(capl - inper);  // at most kW2MaxJ re-fits inside a window
      }
      const int aw = (m + 31) >> 5;      // entry warps that hold entries of this window
      const bool active = warp < aw;
      const bool mine = tid < m;
      W2Saved cur;
      cur.cand = -1; cur.meta = 0; cur.p = V3{0.0f, 0.0f, 0.0f};
      bool acc = false;
      int rej_slot = -1;
      bool win[4] = {false, false, false, false};
      bool unres[4] = {false, false, false, false};
      bool tried[4] = {false, false, false, false};   // mask test done and negative
      int nb[4] = {-1, -1, -1, -1};
      int probe[4] = {0, 0, 0, 0};
      V3 nx[4];
      if (mine) {
        // evaluate with the newest published plane (ps/planar_region.h:168)
        const float* pc = ctl.planes[vcur & (kW2Ver - 1)];
        const Plane P{V3{pc[0], pc[1], pc[2]}, pc[3]};
        const int pos = head + tid;
        const int slot = pos & (R - 1);
        if (sm.q_tag[slot] == tag_base + pos) {
          cur.cand = sm.q_cell[slot];
          cur.p = V3{sm.q_xyz[3 * slot], sm.q_xyz[3 * slot + 1], sm.q_xyz[3 * slot + 2]};
        } else {
          cur.cand = *(volatile int*)(gqueue + (pos & (a.qcap - 1)));
          cur.p = ld3(xyz, cur.cand);
        }
        const int cand = cur.cand;
        acc = plane_abs_distance(P, cur.p) < thr;
        if (acc) {
          // AddNeighbors, ps/planar_region.h:229-255: (c-1,r), (c,r-1), (c,r+1), (c+1,r)
          int row, col;
          fast_divmod(fdrows, cand, &col, &row);
          if (col > 0) nb[0] = cand - rows;
          if (row > 0) nb[1] = cand - 1;
          if (row < rows - 1) nb[2] = cand + 1;
          if (col < cols - 1) nb[3] = cand + rows;
#pragma unroll
          for (int s = 0; s < 4; ++s) {
            if (nb[s] < 0) continue;
            if (mask_test(sm.mask, nb[s])) {
              unres[s] = true;
              nx[s] = ld3(xyz, nb[s]);  // in flight during the claims; the winner stores it into the ring
              probe[s] = (int)w2_hash(nb[s], kW2Hash - 1);
              // the smallest (entry, slot) key per neighbour cell wins the enqueue
              atomicMin(&sm.htab[probe[s]], ((unsigned)nb[s] << kW2KeyBits) | (unsigned)(tid * 4 + s));
            } else {
              tried[s] = true;
            }
          }
        } else {  // rejected: unlabeled again for LATER entries only (:181)
          unsigned h = w2_hash(cand, kW2Rej - 1);
          const unsigned v = ((unsigned)cand << kW2KeyBits) | (unsigned)tid;
          while (atomicCAS(&sm.rej[h], kHashEmpty, v) != kHashEmpty) h = (h + 1) & (kW2Rej - 1);
          rej_slot = (int)h;
        }
      }
      if (active) {
        const unsigned b = __ballot_sync(FULL, acc);
        if (lane == 0) ctl.wacc[warp] = b;
      }
      PROF_T(w1);
      PROF_ACC(1, w0, w1);
      const bool any_bad = w2_bar_or(1, W, bad);  // B1
      PROF_T(w2);
      PROF_ACC(0, w1, w2);
      if (any_bad) {
        // ---- mis-speculation in window kv.  Withdraw what phase 1 did for the next window, undo the
        // open windows k-1 .. kv in reverse order, resume at window kv cut at the first differing entry
        PROF_CNT(5, 1);
#pragma unroll
        for (int s = 0; s < 4; ++s)
          if (unres[s]) sm.htab[probe[s]] = kHashEmpty;
        if (rej_slot >= 0) sm.rej[rej_slot] = kHashEmpty;
        const int tstar = ctl.tstar;
#ifdef PCSEG_FORCE_ROLLBACK
        forced_k = kv;
#endif
        const int tail_old = tail;
        for (int r = k - 1; r >= kv; --r) {
          const int ag = k - r;
          const W2Saved su = ag == 1 ? s1 : (ag == 2 ? s2 : s3);
          const W2Rec& ru = ctl.rec[r & (kW2Rec - 1)];
          // cells this window appended: unlabeled again
          for (int i = tid; i < ru.nwin; i += W) {
            const int qp = ru.tailb + i;
            const int n = *(volatile int*)(gqueue + (qp & (a.qcap - 1)));
            atomicOr(&sm.mask[n >> 5], 1u << (n & 31));
          }
          w2_bar(1, W);
          // its rejected entries are queued again, its accepted entries lose their label
          if (su.meta & 2) {
            if (su.meta & 1) labels[su.cand] = kUnlabeled;
            else atomicAnd(&sm.mask[su.cand >> 5], ~(1u << (su.cand & 31)));
          }
          w2_bar(1, W);
        }
        const W2Rec& rj = ctl.rec[kv & (kW2Rec - 1)];
        head = rj.head; tail = rj.tailb; ninl = rj.ninl0; inper = rj.inper0; gtot = rj.g0;
        // ring slots of the dropped queue positions must never match again
        for (int qp = tail + tid; qp < tail_old; qp += W)
          if (sm.q_tag[qp & (R - 1)] == tag_base + qp) sm.q_tag[qp & (R - 1)] = -1;
        k = kv;
        force_lim = max(tstar, 1);   // entry 0 is evaluated with the exact plane in the redo
        ++epoch;
        if (tid == 0) {
          ctl.e_pub = kv;
          ctl.rb_k = kv;
          ctl.rb_g = gtot;
          __threadfence_block();
          ctl.rb_epoch = epoch;
          for (int spin = 0; ctl.s_ack != epoch || ctl.r_ack != epoch; ++spin) {
            if (spin > (1 << 22)) { ctl.wd = 1; break; }
            __nanosleep(20);
          }
          __threadfence_block();
          ctl.tstar = INT_MAX;
        }
        w2_bar(1, W);
        if (ctl.wd) { watchdog = true; break; }
        ready = 0;
        vcur = gtot;   // version g0 of the failed window: final and verified
        continue;
      }
      if (ready) ++kv;   // window kv is exact

      int na = 0, J = 0, nwin = 0;
      if (m > 0) {
        // ---- phase 2: scan of the accept bits, staging, claim resolution
        const uint4 wa0 = *reinterpret_cast<const uint4*>(&ctl.wacc[0]);
        const uint4 wa1 = *reinterpret_cast<const uint4*>(&ctl.wacc[4]);
        const unsigned wv[8] = {wa0.x, wa0.y, wa0.z, wa0.w, wa1.x, wa1.y, wa1.z, wa1.w};
        int excl = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) {
          const int c = w < aw ? __popc(wv[w]) : 0;
          if (w < warp) excl += c;
          if (w == warp) excl += __popc(wv[w] & ((1u << lane) - 1u));
          na += c;
        }
        J = __float2int_rz(((float)(inper + na) + 0.5f) * inv_period);               // re-fits inside the window
        const int jt = __float2int_rz(((float)(inper + excl) + 0.5f) * inv_period);  // re-fits before this entry
        cur.meta = (acc ? 1 : 0) | (mine ? 2 : 0) | (jt << 2);
        const bool any_rej = na < m;
        const int cand = cur.cand;
        if (acc) {  // plane_estimator_.AddPoint terms (ps/plane_estimator.cc:55-65); summed by warp S
          float2* st = reinterpret_cast<float2*>(sm.stage + 10 * ((ninl + excl) & (NS - 1)));
          const V3 p = cur.p;
          st[0] = make_float2(p.x * p.x, p.x * p.y);
          st[1] = make_float2(p.x * p.z, p.y * p.y);
          st[2] = make_float2(p.y * p.z, p.z * p.z);
          st[3] = make_float2(p.x, p.y);
          st[4] = make_float2(p.z, 1.0f);
          labels[cand] = label_id;                 // :177  (speculative until the window is verified)
          inliers[base + ninl + excl] = cand;      // :169
          // only the seed can still have its bit set when it is accepted
          if (k == 0 && tid == 0) atomicAnd(&sm.mask[cand >> 5], ~(1u << (cand & 31)));
        }
        if (tid < 8) {
          W2Rec& rc = ctl.rec[k & (kW2Rec - 1)];
          rc.accb[tid] = tid < aw ? wv[tid] : 0u;
          if (tid == 0) {
            rc.head = head; rc.m = m; rc.tailb = tail; rc.ninl0 = ninl; rc.inper0 = inper; rc.g0 = gtot; rc.J = J;
          }
        }
        PROF_T(x0);
        PROF_ACC(7, w2, x0);
        if (active) {
          // every mask test of the window is done (B1): its rejected cells are unlabeled again
          if (rej_slot >= 0) atomicOr(&sm.mask[cand >> 5], 1u << (cand & 31));
          if (any_rej) {
            // a neighbour that failed the mask test may have been rejected EARLIER in this window
            if (acc) {
#pragma unroll
              for (int s = 0; s < 4; ++s) {
                if (!tried[s]) continue;
                unsigned h = w2_hash(nb[s], kW2Rej - 1);
                while (true) {
                  const unsigned v = sm.rej[h];
                  if (v == kHashEmpty) break;
                  if ((int)(v >> kW2KeyBits) == nb[s]) {
                    if ((int)(v & ((1u << kW2KeyBits) - 1u)) < tid) {
                      unres[s] = true;
                      nx[s] = ld3(xyz, nb[s]);
                      probe[s] = (int)w2_hash(nb[s], kW2Hash - 1);
                      atomicMin(&sm.htab[probe[s]], ((unsigned)nb[s] << kW2KeyBits) | (unsigned)(tid * 4 + s));
                    }
                    break;
                  }
                  h = (h + 1) & (kW2Rej - 1);
                }
              }
            }
            w2_bar(2, aw * 32);  // B3: the late claims are in, every rejected-table lookup is done
            if (rej_slot >= 0) sm.rej[rej_slot] = kHashEmpty;
          }
          PROF_T(x1);
          PROF_ACC(8, x0, x1);
          int round = 0, woff = 0;
          while (true) {
            bool left = false;
#pragma unroll
            for (int s = 0; s < 4; ++s) {
              if (unres[s]) {
                const unsigned v = ((volatile unsigned*)sm.htab)[probe[s]];
                if ((int)(v >> kW2KeyBits) == nb[s]) {
                  win[s] = (v & ((1u << kW2KeyBits) - 1u)) == (unsigned)(tid * 4 + s);
                  unres[s] = false;
                } else {  // slot taken by a smaller cell: next slot, one more round
                  probe[s] = (probe[s] + 1) & (kW2Hash - 1);
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
            w2_bar(2, aw * 32);  // B4
            const bool again = ctl.unres[round & 1] != 0;
            if (tid == 0) ctl.unres[(round + 1) & 1] = 0;
            if (!again) break;
            ++round;
#pragma unroll
            for (int s = 0; s < 4; ++s)
              if (unres[s]) atomicMin(&sm.htab[probe[s]], ((unsigned)nb[s] << kW2KeyBits) | (unsigned)(tid * 4 + s));
            w2_bar(2, aw * 32);
          }
          if (tid == 0) {
            ctl.unres[0] = 0; ctl.unres[1] = 0;
            // the staged terms and the record are complete (B4): warp S may take the window
            __threadfence_block();
            ctl.e_pub = k + 1;
          }
          PROF_T(x2);
          PROF_ACC(9, x1, x2);
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
          // ---- phase 3: ordered append of the winners (undone from the record on a roll-back)
          int qp = tail + woff;
#pragma unroll
          for (int s = 0; s < 4; ++s) {
            if (win[s]) {
              const int n = nb[s];
              sm.htab[probe[s]] = kHashEmpty;
              atomicAnd(&sm.mask[n >> 5], ~(1u << (n & 31)));  // kAlreadyInUnexaminedPointsQueue
              gqueue[qp & (a.qcap - 1)] = n;
              if (qp - head < R) {
                const int sl = qp & (R - 1);
                sm.q_tag[sl] = tag_base + qp;
                sm.q_cell[sl] = n;
                sm.q_xyz[3 * sl] = nx[s].x; sm.q_xyz[3 * sl + 1] = nx[s].y; sm.q_xyz[3 * sl + 2] = nx[s].z;
              }
              ++qp;
            }
          }
          if (tid == 0) {
            ctl.rec[k & (kW2Rec - 1)].nwin = nwin;
            bc[1] = nwin;
          }
          PROF_T(x3);
          PROF_ACC(10, x2, x3);
        }
        PROF_CNT(6, 1);
        PROF_CNT(11, m);
        PROF_CNT(12, J);
      }
      // ---- leader: what the next iteration can do (planes of the oldest open window final?)
      const int k_next = k + (m > 0 ? 1 : 0);
      const int g_next = gtot + J;
      if (tid == 0) {
        PROF_T(p0);
        int rd = ctl.r_done;
        int rdy = 0;
        if (k_next > kv) {
          const W2Rec& rv = ctl.rec[kv & (kW2Rec - 1)];
          const int need = (kv == k) ? g_next : rv.g0 + rv.J;   // (record k's J is in a register here)
          // the queue is empty until the open windows are confirmed, or three are open: wait
          const bool must = (k_next - kv >= 3) || m == 0;
          if (must) {
            for (int spin = 0; (rd = ctl.r_done) < need; ++spin) {
              if (spin > (1 << 22)) { ctl.wd = 1; break; }
              __nanosleep(20);
            }
          }
          rdy = rd >= need;
        }
        // staged terms not yet consumed by warp S + one more window must fit the staging ring
        for (int spin = 0; ninl + na + W - ctl.s_ninl > NS; ++spin) {
          if (spin > (1 << 22)) { ctl.wd = 1; break; }
          __nanosleep(20);
        }
        __threadfence_block();
        bc[0] = rdy;
        bc[2] = rd;
        PROF_T(p1);
        PROF_ACC(3, p0, p1);
      }
      w2_bar(1, W);  // B5
      if (ctl.wd) { watchdog = true; break; }
      if (m > 0) {
        nwin = bc[1];
        s3 = s2; s2 = s1; s1 = cur;
        used3 = used2; used2 = used1; used1 = vcur;
        head += m;
        tail += nwin;
        ninl += na;
        inper = inper + na - J * period;
        gtot = g_next;
        k = k_next;
        ++steps;
        force_lim = W;
      }
      ready = bc[0];
      vcur = bc[2];
      PROF_T(w4);
      PROF_ACC(4, w0, w4);
    }
    // every window is verified; warp S has the final sums once it has consumed the last window
    if (tid == 0) {
      for (int spin = 0; !watchdog && ctl.s_k < k && spin < (1 << 22); ++spin) __nanosleep(20);
      __threadfence_block();
      ctl.done = 1;
    }
    if (watchdog) steps = -1;
  }
  __syncthreads();
  // ---- CheckAndFinalize first lines, ps/planar_region.h:190-196: Plane() / Centroid() of the
  // estimator (a ComputePlane unless the last inlier triggered one)
  if (tid == 0) {
    Estimator est;
    est_clear(est);
#pragma unroll
    for (int i = 0; i < 6; ++i) est.acc[i] = ctl.cur_s[i];
    est.csum = V3{ctl.cur_s[6], ctl.cur_s[7], ctl.cur_s[8]};
    est.w = ctl.cur_s[9];
    const float* pl = ctl.planes[gtot & (kW2Ver - 1)];
    const float* ce = ctl.cents[gtot & (kW2Ver - 1)];
    const V3 prev_n{pl[0], pl[1], pl[2]};
    est.plane = Plane{prev_n, pl[3]};
    est.centroid = V3{ce[0], ce[1], ce[2]};
    if (ninl >= a.min_inliers && inper != 0) {
      const RawFit r = w2_est_raw_call(ctl.cur_s);
      est_apply(est, r);
    } else if (gtot == 0) {
      est.plane = Plane{prev_n, 0.0f};
    }
    out->ninl = ninl;
    out->tail = tail;
    out->steps = steps;
    out->refits = gtot;
    out->plane[0] = est.plane.n.x; out->plane[1] = est.plane.n.y;
    out->plane[2] = est.plane.n.z; out->plane[3] = est.plane.d;
    out->centroid[0] = est.centroid.x; out->centroid[1] = est.centroid.y; out->centroid[2] = est.centroid.z;
    for (int i = 0; i < 6; ++i) out->est[i] = est.acc[i];
    out->est[6] = est.csum.x; out->est[7] = est.csum.y; out->est[8] = est.csum.z; out->est[9] = est.w;
  }
#ifdef PCSEG_PROFILE
  if (a.prof && tid < 13) a.prof[(size_t)blockIdx.x * 16 + tid] += ctl.prof[tid];
#endif
}

}  // namespace pcseg
