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
//                that triggers a re-fit and evaluates each chunk with the NEWEST
//                FINAL plane it has (possibly several re-fits stale).  When the
//                true plane of a chunk becomes final, A proves that no accept
//                bit can differ (plane change x point magnitude < the chunk's
//                distance margin, fp32 error included) or else re-evaluates the
//                chunk; identical accept bits => identical chunk, the chunk is
//                VERIFIED, else A rolls S and B back to it (stop-the-world: epoch
//                counter, acknowledged by both) and resumes with the true plane.
//   warp 1  "S"  adds the moments of the chunks' inliers in FIFO order
//                (sequential fp32) and posts the sums of every re-fit; keeps a
//                snapshot of the sums per chunk record for the roll-back.
//   warps 4-7 "B" expand the chunks A has written, verified or not, up to four
//                records (128 queue entries) per pass, one entry with its four
//                neighbour slots per thread: labels, inlier list, neighbour tests
//                against the 1-bit mask, min-key claims, ordered enqueue with
//                cp.async prefetch of the xyz into the ring.  The queue tail before
//                each record's enqueues is the undo log of a roll-back (appended
//                cells get their mask bit back, rejected entries lose theirs,
//                accepted entries become kUnlabeled again).
//   kPipeDW warps "R" (2, 3 and 8 ..): each owns one re-fit slot: est_raw (centroid,
//                covariance, eigen-solve) of posted sums.  The only part of
//                ComputePlane that depends on the previous plane (eigenvector sign,
//                ps/plane_estimator.cc:208-216) is applied by A in order.
//
// Chunking is invisible to the semantics (the FIFO is popped entry by entry in
// the reference), so how far A runs ahead only changes speed, never results.
#pragma once
#include "grow_planar.cuh"
#include "grow_planar_smem.cuh"

namespace pcseg {

// CTA of the fast path: warp 0 = A, 1 = S, 4-7 = B, re-fit warps 2, 3 and 8 .. (kPipeWarps - 1).
// Fewer warps leave registers on the SM for the kernels of other batches to run beside two
// resident frames; the default is the measured optimum on the C2 batch.
#ifndef PCSEG_PIPE_WARPS
#define PCSEG_PIPE_WARPS 11
#endif
constexpr int kPipeWarps = PCSEG_PIPE_WARPS;
constexpr int kPipeDW = kPipeWarps == 12 ? 5 : kPipeWarps - 6;  // re-fit warps / slots (12 warps: warp 8 idles)
static_assert(kPipeWarps >= 9 && kPipeWarps <= 12, "A + S + 4 B + 3..5 re-fit warps");
constexpr int kPipeNC = 16;   // chunk records in flight
constexpr int kPipePH = 8;    // plane history (> kPipeDW)
constexpr int kPipeHash = kHashSize;

struct PipeCtl {
  // B -> A
  volatile int tail_pub;    // queue entries [0, tail_pub) are complete (cell, xyz)
  volatile int b_done;      // chunks whose enqueues are all inside tail_pub
  volatile int b_ack;       // roll-back epoch acknowledged
  // S -> A
  volatile int s_done;      // chunks accumulated
  volatile int s_ack;       // roll-back epoch acknowledged
  volatile float cur_s[10]; // sums after chunk s_done - 1
  // A -> B, S
  volatile int a_spec;      // chunk records [0, a_spec) are written            (S, B)
  volatile int rb_epoch;    // bumped by A on a roll-back to chunk rb_chunk     (S, B)
  volatile int rb_chunk;
  volatile int done;        // region finished: all role loops exit
  // chunk records (written by A before a_spec passes them)
  volatile int cr_head[kPipeNC];
  volatile int cr_proc[kPipeNC];
  volatile unsigned cr_amask[kPipeNC];
  volatile int cr_post[kPipeNC];   // 0, or the ticket of the re-fit this chunk ends with
  volatile int cr_slot[kPipeNC];
  // A-private per chunk
  int cr_m[kPipeNC];
  unsigned cr_all[kPipeNC];
  int cr_vreq[kPipeNC], cr_used[kPipeNC], cr_inper[kPipeNC], cr_ninl[kPipeNC];
  float cr_margin[kPipeNC], cr_pmax[kPipeNC];
  float ph[kPipePH][4];            // plane of version v at v % kPipePH
  // S-private per chunk: sums before the chunk
  float cr_snap[kPipeNC][10];
  // B-private per chunk: queue tail before the chunk's enqueues (undo log)
  int cr_tail0[kPipeNC];
  // re-fit slots
  volatile float req_s[kPipeDW][10];
  volatile int req_ticket[kPipeDW];
  volatile float res_v[kPipeDW][8];  // centroid xyz, ev xyz, smallest, trace
  volatile int res_valid[kPipeDW];
  volatile int res_ticket[kPipeDW];
  int req_chunk[kPipeDW];    // A-private: chunk whose end posts the request
  int req_expect[kPipeDW];   // A-private: ticket A waits for
  int ticket;                // persists across regions
};

using PipeSmem = PlanarSmem;

// Back-off of an idle role: keeps its polling off the issue slots and the shared-memory
// pipe that the busy roles of both resident frames need.
#ifndef PIPE_IDLE_NS
#define PIPE_IDLE_NS 20
#endif
#define PIPE_IDLE() __nanosleep(PIPE_IDLE_NS)

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
// Warp A: speculative evaluation, verification, plane finalisation,
// termination.
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
  int head = 0, ninl = 0, inper = 0;
  int c_spec = 0, c_ver = 0, n_posted = 0, n_final = 0;
  int slot_post = 0, slot_fin = 0;  // n_posted % kPipeDW, n_final % kPipeDW
  int tail_seen = 1, b_done_seen = 0, s_done_seen = 0;
  int ticket = ctl.ticket;
  int epoch = 0;
  int refits = 0, steps = 0;
  if (lane == 0) {
    ctl.ph[0][0] = plane.n.x; ctl.ph[0][1] = plane.n.y; ctl.ph[0][2] = plane.n.z; ctl.ph[0][3] = plane.d;
  }
  __syncwarp();
#ifdef PCSEG_PROFILE
  long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#endif
  while (true) {
    bool progressed = false;
    PROF_T(t0);
    // (1) apply the next re-fit once the chunk that posts it is verified
    if (n_final < n_posted) {
      const int s = slot_fin;
      if (ctl.req_chunk[s] < c_ver && ctl.res_ticket[s] == ctl.req_expect[s]) {
        __threadfence_block();
        centroid = V3{ctl.res_v[s][0], ctl.res_v[s][1], ctl.res_v[s][2]};
        const V3 ev{ctl.res_v[s][3], ctl.res_v[s][4], ctl.res_v[s][5]};
        V3 n = prev_n;  // invalid fit keeps the old normal, ps/plane_estimator.cc:219-226
        if (ctl.res_valid[s]) n = (vdot(prev_n, ev) > 0.0f) ? ev : vneg(ev);
        plane = make_plane(n, centroid);
        prev_n = n;
        ++n_final;
        slot_fin = (slot_fin + 1 == kPipeDW) ? 0 : slot_fin + 1;
        ++refits;
        if (lane == 0) {
          float* h = ctl.ph[n_final % kPipePH];
          h[0] = plane.n.x; h[1] = plane.n.y; h[2] = plane.n.z; h[3] = plane.d;
        }
        __syncwarp();
        progressed = true;
      }
    }
    PROF_T(t1);
    // (2) verify the oldest unverified chunk if its true plane is final
    if (c_ver < c_spec) {
      const int r = c_ver % kPipeNC;
      if (ctl.cr_vreq[r] == n_final) {
        bool ok = true;
        const int used = ctl.cr_used[r];
        if (used != n_final) {
          // |dist'(p) - dist(p)| <= |dn|_1 * max|p_k| + |dd| for every entry; plus the fp32
          // rounding of both evaluations (<= ~8 roundings of magnitude |n|_1 max|p_k| + |d|).
          const float* h = ctl.ph[used % kPipePH];
          const float ux = h[0], uy = h[1], uz = h[2], ud = h[3];
          const float pmax = ctl.cr_pmax[r];
          const float dn = fabsf(plane.n.x - ux) + fabsf(plane.n.y - uy) + fabsf(plane.n.z - uz);
          const float dd = fabsf(plane.d - ud);
          const float nsum = fmaxf(fabsf(ux) + fabsf(uy) + fabsf(uz),
                                   fabsf(plane.n.x) + fabsf(plane.n.y) + fabsf(plane.n.z));
          const float dmax = fmaxf(fabsf(ud), fabsf(plane.d));
          const float bound = (dn * pmax + dd) * 1.01f + 2.0e-6f * (nsum * pmax + dmax);
          if (!(bound < ctl.cr_margin[r])) {  // inconclusive (or NaN): evaluate again
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
        }
#ifdef PCSEG_FORCE_ROLLBACK
        if (used != n_final && (c_ver % PCSEG_FORCE_ROLLBACK) == PCSEG_FORCE_ROLLBACK - 1) ok = false;
#endif
        if (ok) {
          ++c_ver;
        } else {  // mis-speculation: resume from this chunk with the true plane
          PROF_CNT(5, 1);
          head = ctl.cr_head[r];
          ninl = ctl.cr_ninl[r];
          inper = ctl.cr_inper[r];
          c_spec = c_ver;
          n_posted = n_final;
          slot_post = slot_fin;
          ++epoch;
          if (lane == 0) {
            ctl.a_spec = c_spec;
            ctl.rb_chunk = c_spec;
            __threadfence_block();
            ctl.rb_epoch = epoch;
          }
          while (ctl.s_ack != epoch || ctl.b_ack != epoch) {}
          __threadfence_block();
          s_done_seen = ctl.s_done;
          b_done_seen = ctl.b_done;
          tail_seen = ctl.tail_pub;
        }
        progressed = true;
      }
    }
    PROF_T(t2);
    // (3) speculate on the next chunk with the newest final plane
    if (c_spec - b_done_seen >= kPipeNC) b_done_seen = ctl.b_done;
    if (c_spec - s_done_seen >= kPipeNC) s_done_seen = ctl.s_done;
    if (c_spec - b_done_seen < kPipeNC && c_spec - s_done_seen < kPipeNC && c_spec - c_ver < kPipeNC &&
        n_posted - n_final < kPipeDW) {
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
        PROF_CNT(4, tail_seen - head);
        PROF_CNT(6, c_spec - c_ver);
        bool acc = false;
        unsigned mg = 0x7f800000u, pm = 0u;  // +inf margin / zero magnitude: entry can never flip
        if (lane < m) {
          int cand;
          V3 p;
          pipe_read_entry(sm, gqueue, a.qcap, xyz, tag_base, head + lane, &cand, &p);
          const float ad = plane_abs_distance(plane, p);
          acc = ad < thr;  // ps/planar_region.h:168
          if (vfinite(p)) {
            mg = __float_as_uint(fabsf(ad - thr));
            pm = __float_as_uint(fmaxf(fmaxf(fabsf(p.x), fabsf(p.y)), fabsf(p.z)));
          }
        }
        const unsigned all = __ballot_sync(FULL, acc);
        const float margin = __uint_as_float(__reduce_min_sync(FULL, mg));
        const float pmax = __uint_as_float(__reduce_max_sync(FULL, pm));
        const int need = period - inper;
        const unsigned le = (lane == 31) ? FULL : ((2u << lane) - 1u);
        // lane holding the need-th accepted entry (if any): the re-fit cuts the chunk there
        const unsigned cutm = __ballot_sync(FULL, acc && (__popc(all & le) == need));
        const bool refit = cutm != 0u;
        const int processed = refit ? __ffs(cutm) : m;
        const unsigned pmask = processed >= 32 ? FULL : ((1u << processed) - 1u);
        const unsigned amask = all & pmask;
        const int r = c_spec % kPipeNC;
        if (refit) ++ticket;
        if (lane == 0) {
          ctl.cr_head[r] = head;
          ctl.cr_proc[r] = processed;
          ctl.cr_amask[r] = amask;
          ctl.cr_post[r] = refit ? ticket : 0;
          ctl.cr_slot[r] = slot_post;
          ctl.cr_m[r] = m;
          ctl.cr_all[r] = all;
          ctl.cr_vreq[r] = n_posted;
          ctl.cr_used[r] = n_final;
          ctl.cr_inper[r] = inper;
          ctl.cr_ninl[r] = ninl;
          ctl.cr_margin[r] = margin;
          ctl.cr_pmax[r] = pmax;
          if (refit) {  // plane_ = plane_estimator_.Plane(), :172-176 -- posted by S, applied in order
            ctl.req_chunk[slot_post] = c_spec;
            ctl.req_expect[slot_post] = ticket;
          }
          __threadfence_block();
          ctl.a_spec = c_spec + 1;
        }
        const int na = __popc(amask);
        ninl += na;
        inper += na;
        head += processed;
        if (refit) {
          inper = 0;
          ++n_posted;
          slot_post = (slot_post + 1 == kPipeDW) ? 0 : slot_post + 1;
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
      if (tail_seen - head < 32) PROF_CNT(7, 1);
    }
#endif
    // (4) nothing to do: refresh what the other warps published; finished?
    if (!progressed) {
      const int bd = ctl.b_done;
      const int sd = ctl.s_done;
      __threadfence_block();
      const int tp = ctl.tail_pub;
      if (c_spec == c_ver && n_final == n_posted && bd == c_ver && sd == c_spec && head == tp) break;
      tail_seen = tp;
      b_done_seen = bd;
      s_done_seen = sd;
      PIPE_IDLE();
      PROF_T(t4);
      PROF_ACC(3, t0, t4);
    }
  }
  // final sums (S has accumulated every chunk)
  Estimator est;
  est_clear(est);
#pragma unroll
  for (int i = 0; i < 6; ++i) est.acc[i] = ctl.cur_s[i];
  est.csum = V3{ctl.cur_s[6], ctl.cur_s[7], ctl.cur_s[8]};
  est.w = ctl.cur_s[9];
  __syncwarp();
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
  est.plane = plane;
  est.plane.n = prev_n;
  est.centroid = centroid;
  if (ninl >= a.min_inliers && inper != 0) est_compute<true>(est);
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
// Warp S: PlaneEstimator::AddPoint of every inlier in FIFO order (sequential
// fp32 sums, ps/plane_estimator.cc:55-65) for the chunk records A has written,
// verified or not; posts the sums of a chunk that ends with a re-fit.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void pipe_warp_S(const PlanarArgs& a, const float* __restrict__ xyz,
                                            const int* gqueue, int tag_base, const PipeSmem& sm,
                                            PipeCtl& ctl) {
  const int lane = threadIdx.x & 31;
  float accj = 0.0f;  // lane j < 10 owns accumulator j: acc[0..5], csum xyz, weight
  int c_s = 0, my_epoch = 0;
  while (true) {
    const int e = ctl.rb_epoch;
    if (e != my_epoch) {  // A rolled back to chunk rb_chunk
      __threadfence_block();
      const int rc = ctl.rb_chunk;
      if (c_s > rc) {
        c_s = rc;
        if (lane < 10) accj = ctl.cr_snap[rc % kPipeNC][lane];
      }
      my_epoch = e;
      __syncwarp();
      if (lane == 0) {
        ctl.s_done = c_s;
        __threadfence_block();
        ctl.s_ack = e;
      }
      continue;
    }
    if (ctl.a_spec <= c_s) {
      if (ctl.done) break;
      PIPE_IDLE();
      continue;
    }
    __threadfence_block();
    const int r = c_s % kPipeNC;
    const int head = ctl.cr_head[r];
    const unsigned amask = ctl.cr_amask[r];
    const int post = ctl.cr_post[r];
    const int slot = ctl.cr_slot[r];
    if (lane < 10) ctl.cr_snap[r][lane] = accj;
    if ((amask >> lane) & 1u) {  // plane_estimator_.AddPoint products, ps/planar_region.h:170
      int cand;
      V3 p;
      pipe_read_entry(sm, gqueue, a.qcap, xyz, tag_base, head + lane, &cand, &p);
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
      ctl.cur_s[lane] = accj;
      if (post) ctl.req_s[slot][lane] = accj;
    }
    __syncwarp();
    ++c_s;
    if (lane == 0) {
      __threadfence_block();
      if (post) ctl.req_ticket[slot] = post;
      ctl.s_done = c_s;
    }
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
    const RawFit r = est_raw<true>(S, V3{S[6], S[7], S[8]}, S[9]);
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
// inlier list, AddNeighbors.  One pass takes up to kPipeBRec consecutive
// verified chunk records (<= 128 contiguous queue entries); thread e handles
// entry e and its four neighbour slots, so key = e * 4 + s is the (position,
// slot) order of grow_common.cuh: a neighbour cell is enqueued by its smallest
// claiming key and the winners are appended in key order.  The cost of a pass
// is latency (barriers, shared-memory round trips), nearly independent of the
// number of entries, so B catches up by taking more records whenever A is ahead.
// The four warps meet at named barrier 1.
// ---------------------------------------------------------------------------
constexpr int kPipeBThreads = 128;
constexpr int kPipeBRec = 4;

__device__ __forceinline__ void pipe_bar_B() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

struct PipeBShared {
  volatile int cmd;
  volatile unsigned wb[4][4];  // [warp][slot] winner ballots
  volatile int unres[4];       // per-warp "some claim still probing"
  volatile int cand[128];      // batch cells (rejection rule)
};

__device__ __forceinline__ void pipe_group_B(const PlanarArgs& a, const float* __restrict__ xyz, int* labels,
                                             int* gqueue, int* inliers, int label_id, int base, int tag_base,
                                             const PipeSmem& sm, PipeCtl& ctl, PipeBShared& bs, int t) {
  const unsigned FULL = 0xffffffffu;
  const int lane = t & 31, bw = t >> 5;
  const int e = t;
  const int rows = a.rows, cols = a.cols;
  const int R = sm.ring;
  const FastDiv fdrows = make_fastdiv(rows);
  int k = 0, pub_k = 0, tail = 1, ninl = 0, my_epoch = 0;
#ifdef PCSEG_PROFILE
  long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#endif
  while (true) {
    PROF_T(t0);
    if (t == 0) {
      int cmd;
      while (true) {
        if (ctl.rb_epoch != my_epoch) { cmd = -2; break; }  // roll-back: undo
        const int av = ctl.a_spec;
        if (av > k) { cmd = min(av - k, kPipeBRec); break; }
        if (pub_k < k) { cmd = -1; break; }  // idle with unpublished enqueues: flush
        if (ctl.done) { cmd = 0; break; }
        PIPE_IDLE();
      }
      __threadfence_block();
      bs.cmd = cmd;
    }
    pipe_bar_B();
    const int cmd = bs.cmd;
    if (cmd == 0) break;
    if (cmd == -2) {
      // A rolled back to chunk rb_chunk: undo the expansion of chunks k-1 .. rb_chunk in
      // reverse order.  Per chunk the inverse of "rejected entry -> bit set, winner -> bit
      // cleared + appended" is "appended cells -> bit set, then rejected entries -> bit
      // cleared"; accepted entries get their label back (still queued = kUnlabeled in HBM).
      cp_async_wait<0>();
      __threadfence_block();
      const int ep = ctl.rb_epoch;
      const int rc = ctl.rb_chunk;
      pipe_bar_B();
      int tail_hi = tail;
      for (int c = k - 1; c >= rc; --c) {
        const int r = c % kPipeNC;
        const int t0q = ctl.cr_tail0[r];
        for (int qp = t0q + t; qp < tail_hi; qp += kPipeBThreads) {
          const int n = *(volatile int*)(gqueue + (qp & (a.qcap - 1)));
          atomicOr(&sm.mask[n >> 5], 1u << (n & 31));
          // the slot must not match again if this position is re-appended while the ring is full
          if (((volatile int*)sm.q_tag)[qp & (R - 1)] == tag_base + qp) sm.q_tag[qp & (R - 1)] = -1;
        }
        pipe_bar_B();
        const int pj = ctl.cr_proc[r];
        const unsigned am = ctl.cr_amask[r];
        if (t < pj) {
          const int pos = ctl.cr_head[r] + t;
          const int slot = pos & (R - 1);
          int cell;
          if (((volatile int*)sm.q_tag)[slot] == tag_base + pos) cell = ((volatile int*)sm.q_cell)[slot];
          else cell = *(volatile int*)(gqueue + (pos & (a.qcap - 1)));
          if ((am >> t) & 1u) labels[cell] = kUnlabeled;
          else atomicAnd(&sm.mask[cell >> 5], ~(1u << (cell & 31)));
        }
        pipe_bar_B();
        ninl -= __popc(am);
        tail_hi = t0q;
      }
      if (k > rc) {
        tail = tail_hi;
        k = rc;
      }
      pub_k = k;
      my_epoch = ep;
      if (t == 0) {
        __threadfence_block();
        ctl.tail_pub = tail;
        ctl.b_done = k;
        __threadfence_block();
        ctl.b_ack = ep;
      }
      continue;
    }
    if (cmd < 0) {
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
    // the batch: records k .. k+cmd-1, entries [head, head + total)
    const int head = ctl.cr_head[k % kPipeNC];
    int total = 0, nacc_before = 0, nacc_total = 0, first_of = -1;
    bool mine = false, acc = false, any_rej = false;
#pragma unroll
    for (int j = 0; j < kPipeBRec; ++j) {
      if (j < cmd) {
        const int r = (k + j) % kPipeNC;
        const int pj = ctl.cr_proc[r];
        const unsigned am = ctl.cr_amask[r];
        const unsigned pm = pj >= 32 ? FULL : ((1u << pj) - 1u);
        any_rej = any_rej || ((~am & pm) != 0u);
        const int q = e - total;  // index inside record j
        if (q == 0) first_of = r;
        if (q >= 0 && q < pj) {
          mine = true;
          acc = (am >> q) & 1u;
          nacc_before = nacc_total + __popc(am & ((1u << q) - 1u));
        }
        total += pj;
        nacc_total += __popc(am);
      }
    }
    int cand = -1;
    if (mine) {
      const int pos = head + e;
      const int slot = pos & (R - 1);
      if (((volatile int*)sm.q_tag)[slot] == tag_base + pos) cand = ((volatile int*)sm.q_cell)[slot];
      else cand = *(volatile int*)(gqueue + (pos & (a.qcap - 1)));
    }
    // AddNeighbors, ps/planar_region.h:229-255: (c-1,r), (c,r-1), (c,r+1), (c+1,r)
    int nb[4] = {-1, -1, -1, -1};
    bool el[4] = {false, false, false, false};
    if (acc) {
      labels[cand] = label_id;                    // :177
      inliers[base + ninl + nacc_before] = cand;  // :169
      // only the seed can still have its bit set when it is accepted
      if (k == 0 && e == 0) atomicAnd(&sm.mask[cand >> 5], ~(1u << (cand & 31)));
      int row, col;
      fast_divmod(fdrows, cand, &col, &row);
      if (col > 0) nb[0] = cand - rows;
      if (row > 0) nb[1] = cand - 1;
      if (row < rows - 1) nb[2] = cand + 1;
      if (col < cols - 1) nb[3] = cand + rows;
#pragma unroll
      for (int s = 0; s < 4; ++s) el[s] = nb[s] >= 0 && mask_test(sm.mask, nb[s]);
    }
    // a cell rejected at entry q is kUnlabeled again for every later entry (:181)
    if (any_rej) {
      if (mine) bs.cand[e] = cand;
      pipe_bar_B();
      if (acc) {
        int off = 0;
#pragma unroll
        for (int j = 0; j < kPipeBRec; ++j) {
          if (j < cmd && off < e) {
            const int r = (k + j) % kPipeNC;
            const int pj = ctl.cr_proc[r];
            const unsigned pm = pj >= 32 ? FULL : ((1u << pj) - 1u);
            unsigned rm = ~ctl.cr_amask[r] & pm;
            const int lim = e - off;  // only entries before e
            if (lim < 32) rm &= (1u << lim) - 1u;
            for (; rm; rm &= rm - 1) {
              const int cq = bs.cand[off + __ffs(rm) - 1];
#pragma unroll
              for (int s = 0; s < 4; ++s) el[s] = el[s] || (nb[s] == cq);
            }
            off += pj;
          }
        }
      }
    }
    PROF_T(t2);
    // smallest key per neighbour cell wins the enqueue
    bool win[4] = {false, false, false, false};
    bool unres[4] = {el[0], el[1], el[2], el[3]};
    int probe[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) probe[s] = (int)(((unsigned)nb[s] * 2654435761u) >> 21) & (kPipeHash - 1);
    bool first = true;
    while (true) {
#pragma unroll
      for (int s = 0; s < 4; ++s)
        if (unres[s]) atomicMin(&sm.htab[probe[s]], ((unsigned)nb[s] << 9) | (unsigned)(e * 4 + s));
      pipe_bar_B();
      // every mask test of this batch is done: rejected cells become unlabeled again
      // (ordered before the winners clear their bits, after the next barrier)
      if (first && mine && !acc) atomicOr(&sm.mask[cand >> 5], 1u << (cand & 31));
      first = false;
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        if (unres[s]) {
          const unsigned v = ((volatile unsigned*)sm.htab)[probe[s]];
          if ((int)(v >> 9) == nb[s]) {
            win[s] = (v & 511u) == (unsigned)(e * 4 + s);
            unres[s] = false;
          } else {
            probe[s] = (probe[s] + 1) & (kPipeHash - 1);
          }
        }
      }
      const unsigned b0 = __ballot_sync(FULL, win[0]), b1 = __ballot_sync(FULL, win[1]);
      const unsigned b2 = __ballot_sync(FULL, win[2]), b3 = __ballot_sync(FULL, win[3]);
      const unsigned ub = __ballot_sync(FULL, unres[0] || unres[1] || unres[2] || unres[3]);
      if (lane == 0) {
        bs.wb[bw][0] = b0; bs.wb[bw][1] = b1; bs.wb[bw][2] = b2; bs.wb[bw][3] = b3;
        bs.unres[bw] = ub != 0u;
      }
      pipe_bar_B();
      if (!(bs.unres[0] | bs.unres[1] | bs.unres[2] | bs.unres[3])) break;
      pipe_bar_B();  // everyone has read the flags before they are rewritten
    }
    PROF_T(t3);
    int off = 0, nwin = 0;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const int c = __popc(bs.wb[w][0]) + __popc(bs.wb[w][1]) + __popc(bs.wb[w][2]) + __popc(bs.wb[w][3]);
      if (w < bw) off += c;
      nwin += c;
    }
    {
      const unsigned lt = (1u << lane) - 1u;
      off += __popc(bs.wb[bw][0] & lt) + __popc(bs.wb[bw][1] & lt) + __popc(bs.wb[bw][2] & lt) +
             __popc(bs.wb[bw][3] & lt);
    }
    const int new_head = head + total;
    if (first_of >= 0) ctl.cr_tail0[first_of] = tail + off;  // undo log: tail before this record's enqueues
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      if (win[s]) {
        const int n = nb[s];
        sm.htab[probe[s]] = kHashEmpty;
        atomicAnd(&sm.mask[n >> 5], ~(1u << (n & 31)));  // kAlreadyInUnexaminedPointsQueue
        const int qp = tail + off;
        gqueue[qp & (a.qcap - 1)] = n;
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
    const int k_prev = k;
    tail += nwin;
    ninl += nacc_total;
    k += cmd;
    // publish one pass behind, so that the xyz copies stay off the critical path
    cp_async_wait<1>();
    pipe_bar_B();
    if (t == 0) {
      __threadfence_block();
      ctl.tail_pub = tail_prev;
      __threadfence_block();
      ctl.b_done = k_prev;
    }
    pub_k = k_prev;
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
