// grow_planar_pipe.cuh -- warp-specialised, speculative FIFO growth of one
// planar region (the fast path of planar_grow_kernel).
//
// Same sequential semantics as ps/planar_region.h:144-255 under the driver of
// ps/segmentation.h:265-275.  The serial chain of the reference is
//   evaluate candidates -> 30th inlier -> eigen-solve -> evaluate next ...
// and one eigen-solve is ~2000 dependent cycles on the GPU.  The chain is cut
// by value speculation and spread over the warps of the frame's CTA:
//
//   warp 0  "A"  walks the queue in chunks of <= 32 entries cut at the inlier
//                that triggers a re-fit, evaluates each chunk with the NEWEST
//                FINAL plane it has (possibly several re-fits stale), adds the
//                moments in FIFO order and posts the sums of every re-fit.
//                When the true plane of a chunk becomes final it re-evaluates
//                the chunk; identical accept bits => identical sums, the chunk
//                is VERIFIED (else A rolls back to it).  Only verified chunks
//                are handed on, so nothing speculative ever reaches memory.
//   warps 4-7 "B" expand verified chunks in order, one (entry, slot) pair per
//                thread: labels, inlier list, neighbour tests against the 1-bit
//                mask, min-key claims, ordered enqueue with cp.async prefetch of
//                the xyz into the ring.
//   6 warps  "R" each owns one re-fit slot: est_raw (centroid, covariance,
//                eigen-solve) of posted sums.  The only part of ComputePlane
//                that depends on the previous plane (eigenvector sign,
//                ps/plane_estimator.cc:208-216) is applied by A in order.
//
// Chunking is invisible to the semantics (the FIFO is popped entry by entry in
// the reference), so how far A runs ahead only changes speed, never results.
#pragma once
#include "grow_planar.cuh"
#include "grow_planar_smem.cuh"

namespace pcseg {

constexpr int kPipeDW = 6;    // re-fit warps / slots: warps 1, 2, 3, 9, 10, 11 of the 12-warp CTA
constexpr int kPipeNC = 16;   // chunk records in flight
constexpr int kPipeHash = kHashSize;

struct PipeCtl {
  // B -> A
  volatile int tail_pub;    // queue entries [0, tail_pub) are complete (cell, xyz)
  volatile int b_done;      // chunks fully expanded
  // A -> B
  volatile int a_verified;  // chunks [0, a_verified) are final
  volatile int done;        // region finished: all role loops exit
  // chunk records (written by A before a_verified passes them)
  volatile int cr_head[kPipeNC];
  volatile int cr_proc[kPipeNC];
  volatile unsigned cr_amask[kPipeNC];
  // A-private per chunk
  int cr_m[kPipeNC];
  unsigned cr_all[kPipeNC];
  int cr_vreq[kPipeNC], cr_used[kPipeNC], cr_inper[kPipeNC], cr_ninl[kPipeNC], cr_refit[kPipeNC];
  float cr_snap[kPipeNC][10];
  // re-fit slots
  volatile float req_s[kPipeDW][10];
  volatile int req_ticket[kPipeDW];
  volatile float res_v[kPipeDW][8];  // centroid xyz, ev xyz, smallest, trace
  volatile int res_valid[kPipeDW];
  volatile int res_ticket[kPipeDW];
  int req_chunk[kPipeDW];    // A-private: chunk whose end posted the request
  int req_expect[kPipeDW];   // A-private: ticket A waits for
  int ticket;                // persists across regions
};

using PipeSmem = PlanarSmem;

// Queue entry `pos` -> (cell, xyz); ring slot if its tag matches, else the global queue.
__device__ __forceinline__ void pipe_read_entry(const PipeSmem& sm, const int* gqueue, int qcap,
                                                const float* __restrict__ xyz, int tag_base, int pos,
                                                int* cell, V3* p) {
  const int slot = pos & (sm.ring - 1);
  if (((volatile int*)sm.q_tag)[slot] == tag_base + pos) {
    *cell = ((volatile int*)sm.q_cell)[slot];
    const volatile float* q = sm.q_xyz + 3 * slot;
    *p = V3{q[0], q[1], q[2]};
  } else {
    const int c = *(volatile const int*)(gqueue + (pos & (qcap - 1)));
    *cell = c;
    *p = ld3(xyz, c);
  }
}


// ---------------------------------------------------------------------------
// Warp A: speculative evaluation, ordered moment sums, verification, plane
// finalisation, termination.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void pipe_warp_A(const PlanarArgs& a, const float* __restrict__ xyz,
                                            const float* __restrict__ nrm, const int* gqueue, int seed,
                                            int tag_base, const PipeSmem& sm, PipeCtl& ctl,
                                            GrowResult* out) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int period = a.period;
  const float thr = a.max_plane_distance;
  const V3 seed_normal = ld3(nrm, seed);
  const V3 seed_point = ld3(xyz, seed);
  // Init, ps/planar_region.h:144-159
  Plane plane = make_plane(seed_normal, seed_point);
  V3 prev_n = seed_normal;          // normal of plane_estimator_'s plane_ (sign reference)
  V3 centroid{0.0f, 0.0f, 0.0f};
  float accj = 0.0f;                // lane j < 10 owns accumulator j: acc[0..5], csum xyz, weight
  int head = 0, ninl = 0, inper = 0;
  int c_spec = 0, c_ver = 0, n_posted = 0, n_final = 0;
  int tail_seen = 1, b_done_seen = 0;
  int ticket = ctl.ticket;
  int refits = 0, steps = 0;
#ifdef PCSEG_PROFILE
  long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#endif
  while (true) {
    bool progressed = false;
    PROF_T(t0);
    // (1) apply the next re-fit once the chunk that posted it is verified
    if (n_final < n_posted) {
      const int s = n_final % kPipeDW;
      if (ctl.req_chunk[s] < c_ver && ctl.res_ticket[s] == ctl.req_expect[s]) {
        __threadfence_block();
        centroid = V3{ctl.res_v[s][0], ctl.res_v[s][1], ctl.res_v[s][2]};
        const V3 ev{ctl.res_v[s][3], ctl.res_v[s][4], ctl.res_v[s][5]};
        V3 n = prev_n;  // invalid fit keeps the old normal, ps/plane_estimator.cc:219-226
        if (ctl.res_valid[s]) n = (vdot(prev_n, ev) > 0.0f) ? ev : vneg(ev);
        plane = make_plane(n, centroid);
        prev_n = n;
        ++n_final;
        ++refits;
        progressed = true;
      }
    }
    PROF_T(t1);
    // (2) verify the oldest unverified chunk if its true plane is final
    if (c_ver < c_spec) {
      const int r = c_ver % kPipeNC;
      if (ctl.cr_vreq[r] == n_final) {
        bool ok = true;
        if (ctl.cr_used[r] != n_final) {
          const int m = ctl.cr_m[r];
          bool acc = false;
          if (lane < m) {
            int cand;
            V3 p;
            pipe_read_entry(sm, gqueue, a.qcap, xyz, tag_base, ctl.cr_head[r] + lane, &cand, &p);
            acc = plane_abs_distance(plane, p) < thr;
          }
          ok = __ballot_sync(FULL, acc) == ctl.cr_all[r];
        }
        if (ok) {
          ++c_ver;
          if (lane == 0) {
            __threadfence_block();
            ctl.a_verified = c_ver;
          }
        } else {  // mis-speculation: resume from this chunk with the true plane
          PROF_CNT(5, 1);
          head = ctl.cr_head[r];
          ninl = ctl.cr_ninl[r];
          inper = ctl.cr_inper[r];
          if (lane < 10) accj = ctl.cr_snap[r][lane];
          c_spec = c_ver;
          n_posted = n_final;
        }
        progressed = true;
      }
    }
    PROF_T(t2);
    // (3) speculate on the next chunk with the newest final plane
    if (c_spec - b_done_seen >= kPipeNC) b_done_seen = ctl.b_done;
    if (c_spec - b_done_seen < kPipeNC && n_posted - n_final < kPipeDW) {
      if (head >= tail_seen) {
        tail_seen = ctl.tail_pub;
        __threadfence_block();
      }
      // A partial chunk is only worth evaluating when no evaluated chunk is still
      // waiting to be expanded (nothing more can arrive before A moves on).
      if (tail_seen - head < 32 && b_done_seen < c_spec) {
        b_done_seen = ctl.b_done;
        __threadfence_block();
        tail_seen = ctl.tail_pub;
      }
      if (tail_seen - head >= 32 || (head < tail_seen && b_done_seen >= c_spec)) {
        const int m = min(32, tail_seen - head);
        bool acc = false;
        V3 p{0.0f, 0.0f, 0.0f};
        if (lane < m) {
          int cand;
          pipe_read_entry(sm, gqueue, a.qcap, xyz, tag_base, head + lane, &cand, &p);
          acc = plane_abs_distance(plane, p) < thr;  // ps/planar_region.h:168
        }
        const unsigned all = __ballot_sync(FULL, acc);
        const int need = period - inper;
        const unsigned le = (lane == 31) ? FULL : ((2u << lane) - 1u);
        // lane holding the need-th accepted entry (if any): the re-fit cuts the chunk there
        const unsigned cutm = __ballot_sync(FULL, acc && (__popc(all & le) == need));
        const bool refit = cutm != 0u;
        const int processed = refit ? __ffs(cutm) : m;
        const unsigned pmask = processed >= 32 ? FULL : ((1u << processed) - 1u);
        const unsigned amask = all & pmask;
        const int r = c_spec % kPipeNC;
        if (lane < 10) ctl.cr_snap[r][lane] = accj;
        if (lane == 0) {
          ctl.cr_head[r] = head;
          ctl.cr_proc[r] = processed;
          ctl.cr_amask[r] = amask;
          ctl.cr_m[r] = m;
          ctl.cr_all[r] = all;
          ctl.cr_vreq[r] = n_posted;
          ctl.cr_used[r] = n_final;
          ctl.cr_inper[r] = inper;
          ctl.cr_ninl[r] = ninl;
        }
        if ((amask >> lane) & 1u) {  // plane_estimator_.AddPoint products, :170
          float* st = sm.stage + lane * 11;
          st[0] = p.x * p.x; st[1] = p.x * p.y; st[2] = p.x * p.z;
          st[3] = p.y * p.y; st[4] = p.y * p.z; st[5] = p.z * p.z;
          st[6] = p.x; st[7] = p.y; st[8] = p.z; st[9] = 1.0f;
        }
        __syncwarp();
        if (lane < 10) {
          // sequential fp32 sums in FIFO order; loads first (independent), adds after
          float v[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = sm.stage[i * 11 + lane];
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if ((amask >> i) & 1u) accj += v[i];
        }
        const int na = __popc(amask);
        ninl += na;
        inper += na;
        head += processed;
        if (refit) {  // plane_ = plane_estimator_.Plane(), :172-176 -- posted, applied in order later
          inper = 0;
          const int s = n_posted % kPipeDW;
          if (lane < 10) ctl.req_s[s][lane] = accj;
          ++ticket;
          __syncwarp();
          if (lane == 0) {
            ctl.req_chunk[s] = c_spec;
            ctl.req_expect[s] = ticket;
            __threadfence_block();
            ctl.req_ticket[s] = ticket;
          }
          ++n_posted;
        }
        ++c_spec;
        ++steps;
        __syncwarp();
        progressed = true;
      }
    }
    PROF_T(t3);
#ifdef PCSEG_PROFILE
    if (progressed) {
      PROF_ACC(0, t0, t1);
      PROF_ACC(1, t1, t2);
      PROF_ACC(2, t2, t3);
    } else {
      PROF_CNT(4, 1);
      if (n_posted - n_final >= kPipeDW) PROF_CNT(6, 1);
      else if (tail_seen - head < 32) PROF_CNT(7, 1);
    }
#endif
    // (4) nothing to do: refresh what the other warps published; finished?
    if (!progressed) {
      const int bd = ctl.b_done;
      __threadfence_block();
      const int tp = ctl.tail_pub;
      if (c_spec == c_ver && n_final == n_posted && bd == c_ver && head == tp) break;
      tail_seen = tp;
      b_done_seen = bd;
      PROF_T(t4);
      PROF_ACC(3, t0, t4);
    }
  }
  if (lane == 0) {
    ctl.ticket = ticket;
    __threadfence_block();
    ctl.done = 1;
  }
#ifdef PCSEG_PROFILE
  if (lane == 0 && a.prof)
    for (int i = 0; i < 8; ++i) a.prof[(size_t)blockIdx.x * 16 + i] += prof_acc[i];
#endif
  // CheckAndFinalize first lines, ps/planar_region.h:190-196
  Estimator est;
  est_clear(est);
#pragma unroll
  for (int i = 0; i < 6; ++i) est.acc[i] = __shfl_sync(FULL, accj, i);
  est.csum.x = __shfl_sync(FULL, accj, 6);
  est.csum.y = __shfl_sync(FULL, accj, 7);
  est.csum.z = __shfl_sync(FULL, accj, 8);
  est.w = __shfl_sync(FULL, accj, 9);
  est.plane = plane;
  est.plane.n = prev_n;
  est.centroid = centroid;
  if (ninl >= a.min_inliers && inper != 0) est_compute(est);
  else if (n_final == 0) est.plane = Plane{prev_n, 0.0f};
  if (lane == 0) {
    out->ninl = ninl;
    out->tail = head;
    out->steps = steps;
    out->refits = refits;
    out->plane[0] = est.plane.n.x; out->plane[1] = est.plane.n.y;
    out->plane[2] = est.plane.n.z; out->plane[3] = est.plane.d;
    out->centroid[0] = est.centroid.x; out->centroid[1] = est.centroid.y; out->centroid[2] = est.centroid.z;
    for (int i = 0; i < 6; ++i) out->est[i] = est.acc[i];
    out->est[6] = est.csum.x; out->est[7] = est.csum.y; out->est[8] = est.csum.z; out->est[9] = est.w;
  }
}

// ---------------------------------------------------------------------------
// Warp R: est_raw of posted sums (seqlock on the slot's ticket).
// ---------------------------------------------------------------------------
// `last` persists across the regions of a frame: on exit it equals the slot's
// newest ticket, so abandoned requests are never picked up later.
__device__ __forceinline__ void pipe_warp_R(PipeCtl& ctl, int s, int& last) {
  const int lane = threadIdx.x & 31;
  while (true) {
    const int t = ctl.req_ticket[s];
    if (t == last) {
      if (ctl.done) break;
      __nanosleep(40);  // keep the poll off the schedulers that A and B run on
      continue;
    }
    __threadfence_block();
    float S[10];
#pragma unroll
    for (int i = 0; i < 10; ++i) S[i] = ctl.req_s[s][i];
    __threadfence_block();
    if (ctl.req_ticket[s] != t) continue;  // re-posted while reading
    const RawFit r = est_raw(S, V3{S[6], S[7], S[8]}, S[9]);
    if (lane == 0) {
      ctl.res_v[s][0] = r.centroid.x; ctl.res_v[s][1] = r.centroid.y; ctl.res_v[s][2] = r.centroid.z;
      ctl.res_v[s][3] = r.ev.x; ctl.res_v[s][4] = r.ev.y; ctl.res_v[s][5] = r.ev.z;
      ctl.res_v[s][6] = r.smallest; ctl.res_v[s][7] = r.trace;
      ctl.res_valid[s] = r.valid ? 1 : 0;
      __threadfence_block();
      ctl.res_ticket[s] = t;
    }
    last = t;
  }
}

// ---------------------------------------------------------------------------
// Group B (4 warps, 128 threads): expansion of verified chunks -- labels,
// inlier list, AddNeighbors.  Thread t handles ONE (entry, slot) pair of the
// chunk, e = t >> 2, s = t & 3, so t itself is the (lane, slot) order key of
// grow_common.cuh: a neighbour cell is enqueued by its smallest claiming t and
// the winners are appended in t order.  The four warps meet at named barrier 1.
// ---------------------------------------------------------------------------
constexpr int kPipeBThreads = 128;

__device__ __forceinline__ void pipe_bar_B() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

struct PipeBShared {
  volatile int cmd;
  volatile unsigned wb[4];     // per-warp winner ballots
  volatile int unres[4];       // per-warp "some claim still probing"
  volatile int cand[32];       // chunk cells (rejection rule)
};

__device__ __forceinline__ void pipe_group_B(const PlanarArgs& a, const float* __restrict__ xyz, int* labels,
                                             int* gqueue, int* inliers, int label_id, int base, int tag_base,
                                             const PipeSmem& sm, PipeCtl& ctl, PipeBShared& bs, int t) {
  const unsigned FULL = 0xffffffffu;
  const int lane = t & 31, bw = t >> 5;
  const int e = t >> 2, s = t & 3;
  const int rows = a.rows, cols = a.cols;
  const int R = sm.ring;
  const FastDiv fdrows = make_fastdiv(rows);
  int k = 0, pub_k = 0, tail = 1, ninl = 0;
#ifdef PCSEG_PROFILE
  long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#endif
  while (true) {
    PROF_T(t0);
    if (t == 0) {
      int cmd;
      while (true) {
        if (ctl.a_verified > k) { cmd = 1; break; }
        if (pub_k < k) { cmd = 2; break; }  // idle with unpublished enqueues: flush
        if (ctl.done) { cmd = 0; break; }
      }
      __threadfence_block();
      bs.cmd = cmd;
    }
    pipe_bar_B();
    const int cmd = bs.cmd;
    if (cmd == 0) break;
    if (cmd == 2) {
      cp_async_wait<0>();
      pipe_bar_B();
      if (t == 0) {
        __threadfence_block();
        ctl.tail_pub = tail;
        __threadfence_block();
        ctl.b_done = k;
      }
      pub_k = k;
      continue;
    }
    PROF_T(t1);
    const int r = k % kPipeNC;
    const int head = ctl.cr_head[r];
    const int processed = ctl.cr_proc[r];
    const unsigned amask = ctl.cr_amask[r];
    const unsigned pmask = processed >= 32 ? FULL : ((1u << processed) - 1u);
    const unsigned rejmask = ~amask & pmask;
    const bool mine = e < processed;
    const bool acc = (amask >> e) & 1u;
    int cand = -1;
    if (mine) {
      const int pos = head + e;
      const int slot = pos & (R - 1);
      if (((volatile int*)sm.q_tag)[slot] == tag_base + pos) cand = ((volatile int*)sm.q_cell)[slot];
      else cand = *(volatile int*)(gqueue + (pos & (a.qcap - 1)));
    }
    // AddNeighbors, ps/planar_region.h:229-255: slot 0 (c-1,r), 1 (c,r-1), 2 (c,r+1), 3 (c+1,r)
    int nb = -1;
    if (acc) {
      int row, col;
      fast_divmod(fdrows, cand, &col, &row);
      if (s == 0) { if (col > 0) nb = cand - rows; }
      else if (s == 1) { if (row > 0) nb = cand - 1; }
      else if (s == 2) { if (row < rows - 1) nb = cand + 1; }
      else { if (col < cols - 1) nb = cand + rows; }
      if (s == 0) {
        labels[cand] = label_id;                                                  // :177
        inliers[base + ninl + __popc(amask & ((1u << e) - 1u))] = cand;           // :169
        // only the seed can still have its bit set when it is accepted
        if (k == 0) atomicAnd(&sm.mask[cand >> 5], ~(1u << (cand & 31)));
      }
    }
    bool el = nb >= 0 && mask_test(sm.mask, nb);
    // a cell rejected at entry q is kUnlabeled again for every later entry (:181)
    if (rejmask) {
      if (s == 0 && mine) bs.cand[e] = cand;
      pipe_bar_B();
      if (nb >= 0) {
        for (unsigned rm = rejmask & ((1u << e) - 1u); rm; rm &= rm - 1)
          el = el || (nb == bs.cand[__ffs(rm) - 1]);
      }
    }
    PROF_T(t2);
    // smallest key t per neighbour cell wins the enqueue
    bool win = false, unres = el, first = true;
    int probe = (int)(((unsigned)nb * 2654435761u) >> 21) & (kPipeHash - 1);
    while (true) {
      if (unres) atomicMin(&sm.htab[probe], ((unsigned)nb << 7) | (unsigned)t);
      pipe_bar_B();
      // every mask test of this chunk is done: rejected cells become unlabeled again
      // (ordered before the winners clear their bits, after the next barrier)
      if (first && mine && !acc && s == 0) atomicOr(&sm.mask[cand >> 5], 1u << (cand & 31));
      first = false;
      if (unres) {
        const unsigned v = ((volatile unsigned*)sm.htab)[probe];
        if ((int)(v >> 7) == nb) {
          win = (v & 127u) == (unsigned)t;
          unres = false;
        } else {
          probe = (probe + 1) & (kPipeHash - 1);
        }
      }
      const unsigned wbal = __ballot_sync(FULL, win);
      const unsigned ubal = __ballot_sync(FULL, unres);
      if (lane == 0) {
        bs.wb[bw] = wbal;
        bs.unres[bw] = ubal != 0u;
      }
      pipe_bar_B();
      if (!(bs.unres[0] | bs.unres[1] | bs.unres[2] | bs.unres[3])) break;
      pipe_bar_B();  // everyone has read the flags before they are rewritten
    }
    PROF_T(t3);
    const unsigned w0 = bs.wb[0], w1 = bs.wb[1], w2 = bs.wb[2], w3 = bs.wb[3];
    const int c0 = __popc(w0), c1 = __popc(w1), c2 = __popc(w2), c3 = __popc(w3);
    const unsigned myb = bw == 0 ? w0 : (bw == 1 ? w1 : (bw == 2 ? w2 : w3));
    const int off = (bw > 0 ? c0 : 0) + (bw > 1 ? c1 : 0) + (bw > 2 ? c2 : 0) + __popc(myb & ((1u << lane) - 1u));
    const int new_head = head + processed;
    if (win) {
      sm.htab[probe] = kHashEmpty;
      atomicAnd(&sm.mask[nb >> 5], ~(1u << (nb & 31)));  // kAlreadyInUnexaminedPointsQueue
      const int qp = tail + off;
      gqueue[qp & (a.qcap - 1)] = nb;
      if (qp - new_head < R) {
        const int sl = qp & (R - 1);
        sm.q_tag[sl] = tag_base + qp;
        sm.q_cell[sl] = nb;
        const float* src = xyz + 3 * (size_t)nb;
        cp_async4(sm.q_xyz + 3 * sl, src);
        cp_async4(sm.q_xyz + 3 * sl + 1, src + 1);
        cp_async4(sm.q_xyz + 3 * sl + 2, src + 2);
      }
    }
    cp_async_commit();
    const int tail_prev = tail;
    tail += c0 + c1 + c2 + c3;
    ninl += __popc(amask);
    ++k;
    // publish one chunk behind, so that the xyz copies stay off the critical path
    cp_async_wait<1>();
    pipe_bar_B();
    if (t == 0) {
      __threadfence_block();
      ctl.tail_pub = tail_prev;
      __threadfence_block();
      ctl.b_done = k - 1;
    }
    pub_k = k - 1;
    PROF_T(t4);
    PROF_ACC(0, t1, t2);
    PROF_ACC(1, t2, t3);
    PROF_ACC(2, t3, t4);
    PROF_ACC(3, t0, t1);
    PROF_CNT(4, 1);
  }
  cp_async_wait<0>();
#ifdef PCSEG_PROFILE
  if (t == 0 && a.prof)
    for (int i = 0; i < 8; ++i) a.prof[(size_t)blockIdx.x * 16 + 8 + i] += prof_acc[i];
#endif
}

}  // namespace pcseg
