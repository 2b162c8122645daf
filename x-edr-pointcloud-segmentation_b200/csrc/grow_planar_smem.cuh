// grow_planar_smem.cuh -- shared-memory resident FIFO growth of one planar
// region (the fast path of planar_grow_kernel).
//
// Same sequential semantics as planar_grow_warp (ps/planar_region.h:144-255 under
// the driver of ps/segmentation.h:265-275), but every structure on the serial
// chain lives in shared memory instead of L2:
//   * mask   : 1 bit per cell, set <=> labels[cell] == kUnlabeled (a cell that
//              is in the queue, claimed by a region or parked has the bit clear);
//   * ring   : the FIFO's next entries as (tag = queue position, cell, xyz); the
//              xyz of a newly enqueued cell is fetched with cp.async straight
//              into its slot, so the point gather is off the critical path.  The
//              global queue keeps every cell id; an entry whose slot was busy
//              (queue longer than the ring) is re-read from there (tag mismatch);
//   * htab   : open-addressing claim table, atomicMin on (cell << 7 | key), that
//              picks for every neighbour cell the smallest (lane, slot) key;
//   * stage  : per-lane moment products, so that lane j < 10 adds accumulator j
//              over the accepted lanes in FIFO order (sequential fp32 sums).
// Global memory only sees fire-and-forget stores (labels of accepted cells,
// inlier list, global queue).
#pragma once
#include "grow_planar.cuh"

namespace pcseg {

constexpr int kHashSize = 256;  // >= 2 * 128 claims per step
constexpr unsigned kHashEmpty = 0xffffffffu;

struct PlanarSmem {
  unsigned* mask;   // [mask_words]
  int* q_tag;       // [ring]
  int* q_cell;      // [ring]
  float* q_xyz;     // [ring * 3]
  float* stage;     // [32 * 11]
  unsigned* htab;   // [kHashSize]
  int ring;         // power of two
};

__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gsrc) {
  unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

__device__ __forceinline__ bool mask_test(const unsigned* mask, int cell) {
  return (mask[cell >> 5] >> (cell & 31)) & 1u;
}

__device__ __forceinline__ void planar_grow_warp_smem(const PlanarArgs& a, const float* __restrict__ xyz,
                                                      const float* __restrict__ nrm, int* labels, int* gqueue,
                                                      int* inliers, int seed, int label_id, int base,
                                                      int tag_base, const PlanarSmem& sm, GrowResult* out) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int rows = a.rows, cols = a.cols;
  const int R = sm.ring;
  const FastDiv fdrows = make_fastdiv(rows);
  const V3 seed_normal = ld3(nrm, seed);
  const V3 seed_point = ld3(xyz, seed);
  // Init, ps/planar_region.h:144-159
  Plane plane = make_plane(seed_normal, seed_point);
  Estimator est;
  est_clear(est);
  est_set_orientation(est, seed_normal);
  float accj = 0.0f;  // lane j < 10 owns accumulator j: acc[0..5], csum xyz, weight
  int head = 0, tail = 1, ninl = 0, steps = 0, refits = 0;
  int th2 = 0, th3 = 0;  // queue tail after the steps t-2 and t-3
  if (lane == 0) {
    sm.q_tag[0] = tag_base;
    sm.q_cell[0] = seed;
    sm.q_xyz[0] = seed_point.x; sm.q_xyz[1] = seed_point.y; sm.q_xyz[2] = seed_point.z;
    gqueue[0] = seed;
  }
  __syncwarp();
  const int period = a.period;
#ifdef PCSEG_PROFILE
  long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#endif
  while (head < tail) {
    ++steps;
    PROF_T(t0);
    const int m = min(32, tail - head);
    // cp.async groups are per thread: every lane waits for its own copies that can
    // cover the entries consumed now, then the warp synchronises.
    if (head + m <= th3) cp_async_wait<2>();
    else cp_async_wait<0>();
    __syncwarp();
    const int pos = head + lane;
    const int slot = pos & (R - 1);
    const bool have = lane < m;
    int cand = -1;
    V3 p{0.f, 0.f, 0.f};
    if (have) {
      if (sm.q_tag[slot] == tag_base + pos) {
        cand = sm.q_cell[slot];
        p = V3{sm.q_xyz[3 * slot], sm.q_xyz[3 * slot + 1], sm.q_xyz[3 * slot + 2]};
      } else {  // the ring was full when this entry was enqueued
        cand = *(volatile int*)(gqueue + pos % a.qcap);
        p = ld3(xyz, cand);
      }
    }
    bool acc = have && (plane_abs_distance(plane, p) < a.max_plane_distance);  // :168
    unsigned amask = __ballot_sync(FULL, acc);
    PROF_T(t1);
    PROF_ACC(0, t0, t1);
    const int need = period - (ninl % period);
    const unsigned le = (lane == 31) ? FULL : ((2u << lane) - 1u);
    // lane holding the need-th accepted entry (if any): the re-fit cuts the chunk there
    const unsigned cutm = __ballot_sync(FULL, acc && (__popc(amask & le) == need));
    int processed = m;
    const bool refit = cutm != 0u;
    if (refit) processed = __ffs(cutm);
    const unsigned pmask = processed >= 32 ? FULL : ((1u << processed) - 1u);
    amask &= pmask;
    const bool mine = lane < processed;
    acc = acc && mine;
    const unsigned rejmask = __ballot_sync(FULL, mine && !acc);
    const int rank = __popc(amask & ((1u << lane) - 1u));
    if (acc) {
      labels[cand] = label_id;             // :177
      inliers[base + ninl + rank] = cand;  // :169
      float* st = sm.stage + lane * 11;    // plane_estimator_.AddPoint products, :170
      st[0] = p.x * p.x; st[1] = p.x * p.y; st[2] = p.x * p.z;
      st[3] = p.y * p.y; st[4] = p.y * p.z; st[5] = p.z * p.z;
      st[6] = p.x; st[7] = p.y; st[8] = p.z; st[9] = 1.0f;
      // only the seed can still have its bit set when it is accepted
      if (steps == 1) atomicAnd(&sm.mask[cand >> 5], ~(1u << (cand & 31)));
    }
    __syncwarp();
    PROF_T(t1b);
    PROF_ACC(5, t1, t1b);
    if (lane < 10) {
      // sequential fp32 sums in FIFO order; loads first (independent), adds after
      float v[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = sm.stage[i * 11 + lane];
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if ((amask >> i) & 1u) accj += v[i];
    }
    PROF_T(t2);
    PROF_ACC(1, t1, t2);
    // AddNeighbors, :229-255: (c-1,r), (c,r-1), (c,r+1), (c+1,r)
    int nb[4] = {-1, -1, -1, -1};
    bool el[4] = {false, false, false, false};
    if (acc) {
      int row, col;
      fast_divmod(fdrows, cand, &col, &row);
      if (col > 0) nb[0] = cand - rows;
      if (row > 0) nb[1] = cand - 1;
      if (row < rows - 1) nb[2] = cand + 1;
      if (col < cols - 1) nb[3] = cand + rows;
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
    PROF_T(t2b);
    PROF_ACC(6, t2, t2b);
    // smallest (lane, slot) key per neighbour cell wins the enqueue
    bool win[4] = {false, false, false, false};
    int probe[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) probe[s] = (int)(((unsigned)nb[s] * 2654435761u) >> 24) & (kHashSize - 1);
    bool unres[4] = {el[0], el[1], el[2], el[3]};
    while (__any_sync(FULL, unres[0] || unres[1] || unres[2] || unres[3])) {
#pragma unroll
      for (int s = 0; s < 4; ++s)
        if (unres[s]) atomicMin(&sm.htab[probe[s]], ((unsigned)nb[s] << 7) | (unsigned)(lane * 4 + s));
      __syncwarp();
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        if (unres[s]) {
          const unsigned v = sm.htab[probe[s]];
          if ((int)(v >> 7) == nb[s]) {
            win[s] = (v & 127u) == (unsigned)(lane * 4 + s);
            unres[s] = false;
          } else {
            probe[s] = (probe[s] + 1) & (kHashSize - 1);
          }
        }
      }
      __syncwarp();
    }
    PROF_T(t3);
    PROF_ACC(2, t2, t3);
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
    th3 = th2;
    th2 = tail;
    tail += __popc(b0) + __popc(b1) + __popc(b2) + __popc(b3);
    head = new_head;
    ninl += __popc(amask);
    PROF_T(t4);
    PROF_ACC(3, t3, t4);
    if (refit) {
      // plane_ = plane_estimator_.Plane(), :172-176
#pragma unroll
      for (int i = 0; i < 6; ++i) est.acc[i] = __shfl_sync(FULL, accj, i);
      est.csum.x = __shfl_sync(FULL, accj, 6);
      est.csum.y = __shfl_sync(FULL, accj, 7);
      est.csum.z = __shfl_sync(FULL, accj, 8);
      est.w = __shfl_sync(FULL, accj, 9);
      est_compute(est);
      plane = est.plane;
      ++refits;
    }
    __syncwarp();
    PROF_T(t5);
    PROF_ACC(4, t4, t5);
  }
  cp_async_wait<0>();
#ifdef PCSEG_PROFILE
  if (lane == 0 && a.prof)
    for (int i = 0; i < 8; ++i) a.prof[(size_t)blockIdx.x * 8 + i] += prof_acc[i];
#endif
  // CheckAndFinalize first lines, :190-196
#pragma unroll
  for (int i = 0; i < 6; ++i) est.acc[i] = __shfl_sync(FULL, accj, i);
  est.csum.x = __shfl_sync(FULL, accj, 6);
  est.csum.y = __shfl_sync(FULL, accj, 7);
  est.csum.z = __shfl_sync(FULL, accj, 8);
  est.w = __shfl_sync(FULL, accj, 9);
  if (ninl >= a.min_inliers && (ninl % period) != 0) est_compute(est);
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

}  // namespace pcseg
