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
//   warp 1  "B"  expands verified chunks in order: labels, inlier list,
//                neighbour tests against the 1-bit mask, min-(lane,slot) claims,
//                ordered enqueue with cp.async prefetch of the xyz into the ring.
//   warps 2..    "R" each owns one re-fit slot: est_raw (centroid, covariance,
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

constexpr int kPipeDW = 4;    // re-fit warps / slots: warps 2, 3, 6, 7 (not on A's / B's scheduler)
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
    const int c = *(volatile const int*)(gqueue + pos % qcap);
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
// Warp B: expansion of verified chunks (labels, inliers, AddNeighbors).
// ---------------------------------------------------------------------------
__device__ __forceinline__ void pipe_warp_B(const PlanarArgs& a, const float* __restrict__ xyz, int* labels,
                                            int* gqueue, int* inliers, int label_id, int base, int tag_base,
                                            const PipeSmem& sm, PipeCtl& ctl) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int rows = a.rows, cols = a.cols;
  const int R = sm.ring;
  const FastDiv fdrows = make_fastdiv(rows);
  int k = 0, tail = 1, ninl = 0;
#ifdef PCSEG_PROFILE
  long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#endif
  while (true) {
    PROF_T(t0);
    if (ctl.a_verified <= k) {
      if (ctl.done) break;
      PROF_T(tw);
      PROF_ACC(3, t0, tw);
      continue;
    }
    __threadfence_block();
    PROF_T(t1);
    const int r = k % kPipeNC;
    const int head = ctl.cr_head[r];
    const int processed = ctl.cr_proc[r];
    const unsigned amask = ctl.cr_amask[r];
    const bool mine = lane < processed;
    const bool acc = (amask >> lane) & 1u;
    const unsigned rejmask = __ballot_sync(FULL, mine && !acc);
    int cand = -1;
    if (mine) {
      const int pos = head + lane;
      const int slot = pos & (R - 1);
      if (((volatile int*)sm.q_tag)[slot] == tag_base + pos) cand = ((volatile int*)sm.q_cell)[slot];
      else cand = *(volatile int*)(gqueue + pos % a.qcap);
    }
    const int rank = __popc(amask & ((1u << lane) - 1u));
    // AddNeighbors, ps/planar_region.h:229-255: (c-1,r), (c,r-1), (c,r+1), (c+1,r)
    int nb[4] = {-1, -1, -1, -1};
    bool el[4] = {false, false, false, false};
    if (acc) {
      labels[cand] = label_id;             // :177
      inliers[base + ninl + rank] = cand;  // :169
      // only the seed can still have its bit set when it is accepted
      if (k == 0) atomicAnd(&sm.mask[cand >> 5], ~(1u << (cand & 31)));
      int row, col;
      fast_divmod(fdrows, cand, &col, &row);
      if (col > 0) nb[0] = cand - rows;
      if (row > 0) nb[1] = cand - 1;
      if (row < rows - 1) nb[2] = cand + 1;
      if (col < cols - 1) nb[3] = cand + rows;
    }
    __syncwarp();
    if (acc) {
#pragma unroll
      for (int s = 0; s < 4; ++s) el[s] = nb[s] >= 0 && mask_test(sm.mask, nb[s]);
    }
    // a cell rejected at lane q is kUnlabeled again for every later lane (:181)
    for (unsigned rm = rejmask; rm; rm &= rm - 1) {
      const int q = __ffs(rm) - 1;
      const int cq = __shfl_sync(FULL, cand, q);
      if (acc && lane > q) {
#pragma unroll
        for (int s = 0; s < 4; ++s) el[s] = el[s] || (nb[s] == cq);
      }
    }
    PROF_T(t2);
    // smallest (lane, slot) key per neighbour cell wins the enqueue
    bool win[4] = {false, false, false, false};
    int probe[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) probe[s] = (int)(((unsigned)nb[s] * 2654435761u) >> 21) & (kPipeHash - 1);
    bool unres[4] = {el[0], el[1], el[2], el[3]};
    while (__any_sync(FULL, unres[0] || unres[1] || unres[2] || unres[3])) {
#pragma unroll
      for (int s = 0; s < 4; ++s)
        if (unres[s]) atomicMin(&sm.htab[probe[s]], ((unsigned)nb[s] << 7) | (unsigned)(lane * 4 + s));
      __syncwarp();
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        if (unres[s]) {
          const unsigned v = ((volatile unsigned*)sm.htab)[probe[s]];
          if ((int)(v >> 7) == nb[s]) {
            win[s] = (v & 127u) == (unsigned)(lane * 4 + s);
            unres[s] = false;
          } else {
            probe[s] = (probe[s] + 1) & (kPipeHash - 1);
          }
        }
      }
      __syncwarp();
    }
    PROF_T(t3);
    if (mine && !acc) atomicOr(&sm.mask[cand >> 5], 1u << (cand & 31));
    __syncwarp();
    const unsigned b0 = __ballot_sync(FULL, win[0]), b1 = __ballot_sync(FULL, win[1]);
    const unsigned b2 = __ballot_sync(FULL, win[2]), b3 = __ballot_sync(FULL, win[3]);
    const unsigned lt = (1u << lane) - 1u;
    int off = __popc(b0 & lt) + __popc(b1 & lt) + __popc(b2 & lt) + __popc(b3 & lt);
    const int new_head = head + processed;
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      if (win[s]) {
        const int n = nb[s];
        sm.htab[probe[s]] = kHashEmpty;
        atomicAnd(&sm.mask[n >> 5], ~(1u << (n & 31)));  // kAlreadyInUnexaminedPointsQueue
        const int qp = tail + off;
        gqueue[qp % a.qcap] = n;
        if (qp - new_head < R) {
          const int sl = qp & (R - 1);
          sm.q_tag[sl] = tag_base + qp;
          sm.q_cell[sl] = n;
          const float* src = xyz + 3 * (size_t)n;
          cp_async4(sm.q_xyz + 3 * sl, src);
          cp_async4(sm.q_xyz + 3 * sl + 1, src + 1);
          cp_async4(sm.q_xyz + 3 * sl + 2, src + 2);
        }
        ++off;
      }
    }
    cp_async_commit();
    const int tail_prev = tail;
    tail += __popc(b0) + __popc(b1) + __popc(b2) + __popc(b3);
    ninl += __popc(amask);
    ++k;
    // publish: everything when A has nothing else to do, else lag one chunk
    // behind so that the xyz copies stay off the critical path
    const bool flush = ctl.a_verified == k;
    if (flush) cp_async_wait<0>();
    else cp_async_wait<1>();
    __syncwarp();
    if (lane == 0) {
      __threadfence_block();
      ctl.tail_pub = flush ? tail : tail_prev;
      __threadfence_block();
      ctl.b_done = k;
    }
    PROF_T(t4);
    PROF_ACC(0, t1, t2);
    PROF_ACC(1, t2, t3);
    PROF_ACC(2, t3, t4);
    PROF_CNT(4, 1);
  }
  cp_async_wait<0>();
#ifdef PCSEG_PROFILE
  if (lane == 0 && a.prof)
    for (int i = 0; i < 8; ++i) a.prof[(size_t)blockIdx.x * 16 + 8 + i] += prof_acc[i];
#endif
}
}  // namespace pcseg
