// grow_planar.cuh -- SegmentRegions<PlanarRegion<>> on the GPU.
//
// Replaces ps/segmentation.h:239-292 (driver) with ps/planar_region.h:144-255
// (Init / CheckInlierAndUpdate / AddNeighbors), :189-224 (CheckAndFinalize),
// :295-425 (FindLabeledRegionBoundary), ps/planar_region.cc:26-106 and
// ps/algorithms.h:264-292,526-549 as the model.
//
// One CTA per frame.  The plane is re-fitted every
// plane_model_reestimation_period-th inlier from SEQUENTIAL fp32 sums in FIFO
// order (ps/planar_region.h:168-176), so the growth is a serial chain of
// "evaluate <= 32 queue entries -> cut at the inlier that triggers the re-fit
// -> accumulate in order -> 3x3 eigen-solve".  That chain runs warp-
// synchronously in warp 0; the whole CTA is used for seed scanning, the
// boundary-start search, the hull gate, the discontinuity flags and label
// parking.  Latency-bound by construction: throughput comes
// from many frames in flight, one per CTA.
#pragma once
#include "grow_common.cuh"

namespace pcseg {

struct PlanarArgs {
  const float* xyz;      // [F][N][3]
  const float* normals;  // [F][N][3]
  int* labels;           // [F][N]
  int rows, cols;
  const int* seeds;      // nullptr = all descending; else shared or per-frame list
  int n_seeds;           // shared list length
  int seeds_stride;      // > 0: per-frame lists, seeds + f*stride, count n_seeds_dev[f]
  const int* n_seeds_dev;
  float max_plane_distance, min_region_area;
  int min_inliers, period;
  float disc_min_range, disc_max_range, disc_angle, disc_z_diff, disc_z_ratio;
  float qx, qy, qz, qw;
  int id_offset;
  int* queue;            // [F][2*qcap] (first qcap ints used; stride shared with the cluster stage)
  int qcap;
  int* claim;            // [F][N]
  int* inliers;          // [F][N]
  int* boundary;         // [F][bcap]
  unsigned char* bflags; // [F][bcap] discontinuity flag per boundary entry
  int bcap;
  PlanarMeta* meta;      // [F][rcap]
  int rcap;
  FrameStats* stats;     // [F]
  long long* prof;       // nullable: [F][16] clock64 totals / counters (diagnostics)
  int ring;              // smem ring entries (power of two) for the shared-memory path
  int mask_words;        // ceil(N / 32)
  int nstage;            // staged-inlier ring slots (power of two), grow_planar_wave2.cuh
};

#ifdef PCSEG_PROFILE
#define PROF_T(var) long long var = clock64()
#define PROF_ACC(slot, t0, t1) prof_acc[slot] += (t1) - (t0)
#define PROF_CNT(slot, n) prof_acc[slot] += (n)
#else
#define PROF_CNT(slot, n)
#define PROF_T(var)
#define PROF_ACC(slot, t0, t1)
#endif

struct GrowResult {  // warp 0 -> CTA broadcast
  int ninl;
  int steps, refits;
  int tail;  // number of queue positions used (smem ring tags stay unique across regions)
  float plane[4];
  float centroid[3];
  float est[10];
};

__device__ __constant__ int kDx8[8] = {-1, -1, 0, 1, 1, 1, 0, -1};  // ps/planar_region.cc:30-37
__device__ __constant__ int kDy8[8] = {0, -1, -1, -1, 0, 1, 1, 1};
__device__ __constant__ int kDx4[4] = {-1, 0, 1, 0};                // ps/planar_region.cc:39-42
__device__ __constant__ int kDy4[4] = {0, -1, 0, 1};

// FIFO growth of one region by warp 0 (all 32 lanes).  Returns via *out.
__device__ __forceinline__ void planar_grow_warp(const PlanarArgs& a, const float* __restrict__ xyz,
                                                 const float* __restrict__ nrm, volatile int* labels,
                                                 volatile int* queue, int* claim, int* inliers,
                                                 int seed, int label_id, int base, GrowResult* out) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int rows = a.rows, cols = a.cols;
  volatile int* vclaim = claim;
  const V3 seed_normal = ld3(nrm, seed);
  const V3 seed_point = ld3(xyz, seed);
  // Init, ps/planar_region.h:144-159
  Plane plane = make_plane(seed_normal, seed_point);
  Estimator est;
  est_clear(est);
  est_set_orientation(est, seed_normal);
  int head = 0, tail = 1, ninl = 0, steps = 0, refits = 0;
  if (lane == 0) queue[0] = seed;
  __syncwarp();
  const int period = a.period;
#ifdef PCSEG_PROFILE
  long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#endif
  while (head < tail) {
    ++steps;
    PROF_T(t0);
    const int m = min(32, tail - head);
    int cand = -1;
    V3 p{0.f, 0.f, 0.f};
    bool acc = false;
    if (lane < m) {
      cand = queue[(head + lane) & (a.qcap - 1)];
      p = ld3(xyz, cand);
      acc = plane_abs_distance(plane, p) < a.max_plane_distance;  // :168
    }
    unsigned amask = __ballot_sync(FULL, acc);
    PROF_T(t1);
    PROF_ACC(0, t0, t1);
    const int need = period - (ninl % period);
    int processed = m;
    bool refit = false;
    if (__popc(amask) >= need) {
      processed = (int)__fns(amask, 0, need) + 1;
      refit = true;
    }
    const unsigned pmask = processed >= 32 ? FULL : ((1u << processed) - 1u);
    amask &= pmask;
    const bool mine = lane < processed;
    acc = acc && mine;
    const int rank = __popc(amask & ((1u << lane) - 1u));
    if (mine) {
      if (acc) {
        labels[cand] = label_id;              // :177
        inliers[base + ninl + rank] = cand;   // :169
      } else {
        labels[cand] = kRejBase - lane;       // becomes kUnlabeled below (:181)
      }
    }
    // plane_estimator_.AddPoint in FIFO order, :170
    for (unsigned mm = amask; mm; mm &= mm - 1) {
      const int src = __ffs(mm) - 1;
      V3 q{__shfl_sync(FULL, p.x, src), __shfl_sync(FULL, p.y, src), __shfl_sync(FULL, p.z, src)};
      est_add(est, q);
    }
    __syncwarp();
    PROF_T(t2);
    PROF_ACC(1, t1, t2);
    // AddNeighbors, :229-255: (c-1,r), (c,r-1), (c,r+1), (c+1,r)
    int nb[4] = {-1, -1, -1, -1};
    bool el[4] = {false, false, false, false};
    if (acc) {
      const int row = cand % rows, col = cand / rows;
      if (col > 0) nb[0] = cand - rows;
      if (row > 0) nb[1] = cand - 1;
      if (row < rows - 1) nb[2] = cand + 1;
      if (col < cols - 1) nb[3] = cand + rows;
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        if (nb[s] >= 0) {
          const int L = labels[nb[s]];
          el[s] = (L == kUnlabeled) || (L <= kRejBase && (kRejBase - L) < lane);
        }
      }
#pragma unroll
      for (int s = 0; s < 4; ++s)
        if (el[s]) atomicMin(claim + nb[s], lane * 4 + s);
    }
    __syncwarp();
    PROF_T(t3);
    PROF_ACC(2, t2, t3);
    if (mine && !acc) labels[cand] = kUnlabeled;
    bool win[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) win[s] = el[s] && (vclaim[nb[s]] == lane * 4 + s);
    __syncwarp();
    const unsigned b0 = __ballot_sync(FULL, win[0]), b1 = __ballot_sync(FULL, win[1]);
    const unsigned b2 = __ballot_sync(FULL, win[2]), b3 = __ballot_sync(FULL, win[3]);
    const unsigned lt = (1u << lane) - 1u;
    int off = __popc(b0 & lt) + __popc(b1 & lt) + __popc(b2 & lt) + __popc(b3 & lt);
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      if (win[s]) {
        queue[(tail + off) & (a.qcap - 1)] = nb[s];
        labels[nb[s]] = kAlreadyInQueue;
        claim[nb[s]] = INT_MAX;
        ++off;
      }
    }
    tail += __popc(b0) + __popc(b1) + __popc(b2) + __popc(b3);
    head += processed;
    ninl += __popc(amask);
    PROF_T(t4);
    PROF_ACC(3, t3, t4);
    if (refit) {
      est_compute(est);   // plane_ = plane_estimator_.Plane(), :172-176
      plane = est.plane;
      ++refits;
    }
    __syncwarp();
    PROF_T(t5);
    PROF_ACC(4, t4, t5);
  }
#ifdef PCSEG_PROFILE
  if (lane == 0 && a.prof)
    for (int i = 0; i < 8; ++i) a.prof[(size_t)blockIdx.x * 16 + i] += prof_acc[i];
#endif
  // CheckAndFinalize first lines, :190-196
  if (ninl >= a.min_inliers && !est.up_to_date) est_compute(est);
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

// detail::FindInitialPredecessorDirection, ps/planar_region.cc:67-82
// Label reads here and in trace_boundary are plain (L1-cached) loads: the frame's
// labels are written only by this CTA and every write is ordered by a barrier.
__device__ __forceinline__ int initial_predecessor_dir(const int* labels, int rows, int cols,
                                                       int idx, int region_label) {
  const int x = idx / rows, y = idx % rows;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int nx = x + kDx8[i], ny = y + kDy8[i];
    if (nx >= 0 && nx < cols && ny >= 0 && ny < rows &&
        labels[idx + kDx8[i] * rows + kDy8[i]] != region_label)
      return i;
  }
  return -1;
}

struct TraceResult {
  int nb, min_x, max_x, min_y, max_y, err;
};

// FindLabeledRegionBoundary (Moore trace), ps/planar_region.h:295-425, executed
// by one thread.  Discontinuity flags are computed afterwards in parallel.
__device__ void trace_boundary(const int* labels, int rows, int cols, int start_idx,
                               int region_label, int* boundary, int cap, TraceResult* out) {
  int curr_idx = start_idx;
  int cx = start_idx / rows, cy = start_idx % rows;
  int dir = initial_predecessor_dir(labels, rows, cols, start_idx, region_label);
  int nb = 0, err = 0;
  int min_x = cols, max_x = 0, min_y = rows, max_y = 0;  // ps/planar_region.cc:93-96
  bool found_start = false;
  int first = -1;
  while (true) {
    // detail::FindNextBoundaryPointDirection, ps/planar_region.cc:47-65
    // bit i of `same` <=> neighbour i (table order) is in bounds and carries the region label
    unsigned same = 0u;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int nx = cx + kDx8[i], ny = cy + kDy8[i];
      if (nx >= 0 && nx < cols && ny >= 0 && ny < rows &&
          labels[curr_idx + kDx8[i] * rows + kDy8[i]] == region_label)
        same |= 1u << i;
    }
    // first direction in the order (dir+1 .. dir+8) mod 8: rotate so that bit 0 = dir+1
    int nd = -1;
    if (same) {
      const unsigned sh = (unsigned)(dir + 1) & 7u;
      const unsigned rot = ((same >> sh) | (same << (8u - sh))) & 0xffu;
      nd = (int)((sh + (unsigned)__ffs(rot) - 1u) & 7u);
    }
    if (nd < 0) { err = kErrTraceCheck; break; }  // CHECK_GE(new_direction_index, 0)
    dir = (nd + 4) & 7;
    curr_idx += kDx8[nd] * rows + kDy8[nd];
    cx += kDx8[nd];
    cy += kDy8[nd];
    if (found_start) {
      if (curr_idx == first) break;
      found_start = false;
    }
    if (nb >= cap) { err = kErrBoundaryCap; break; }
    boundary[nb] = curr_idx;
    if (nb == 0) first = curr_idx;
    ++nb;
    min_x = min(min_x, cx); max_x = max(max_x, cx);
    min_y = min(min_y, cy); max_y = max(max_y, cy);
    if (curr_idx == start_idx) found_start = true;
  }
  out->nb = nb; out->min_x = min_x; out->max_x = max_x;
  out->min_y = min_y; out->max_y = max_y; out->err = err;
}

// The same trace by the whole CTA: the serial walk (thread 0) reads "carries the region label" flags of a
// 64 x 64 cell tile centred on its position from shared memory; all threads reload the tile (coalesced along
// the rows) whenever the walk leaves the tile's interior.  Same steps, same output; the dependent label reads
// of the walk cost a shared-memory access instead of an L2 round trip each.
constexpr int kTraceTile = 64;
struct TraceState {
  int curr_idx, cx, cy, dir, nb, err, found_start, first, done;
  int min_x, max_x, min_y, max_y, ox, oy;
};
template <int BLOCK>
__device__ __forceinline__ void trace_boundary_tiled(const int* labels, int rows, int cols, int start_idx, int region_label,
                                                     int* boundary, int cap, TraceResult* out, unsigned char* tile /*[64*64]*/,
                                                     TraceState* st) {
  const int tid = threadIdx.x;
  if (tid == 0) {
    st->curr_idx = start_idx;
    st->cx = start_idx / rows; st->cy = start_idx % rows;
    st->dir = initial_predecessor_dir(labels, rows, cols, start_idx, region_label);
    st->nb = 0; st->err = 0; st->found_start = 0; st->first = -1; st->done = 0;
    st->min_x = cols; st->max_x = 0; st->min_y = rows; st->max_y = 0;  // ps/planar_region.cc:93-96
    st->ox = st->cx - kTraceTile / 2; st->oy = st->cy - kTraceTile / 2;
  }
  __syncthreads();
  while (true) {
    const int ox = st->ox, oy = st->oy;
    for (int t = tid; t < kTraceTile * kTraceTile; t += BLOCK) {
      const int x = ox + t / kTraceTile, y = oy + t % kTraceTile;
      tile[t] = (x >= 0 && x < cols && y >= 0 && y < rows && labels[x * rows + y] == region_label) ? 1 : 0;
    }
    __syncthreads();
    if (tid == 0) {
      int curr_idx = st->curr_idx, cx = st->cx, cy = st->cy, dir = st->dir, nb = st->nb, err = 0;
      bool found_start = st->found_start != 0;
      int first = st->first;
      int min_x = st->min_x, max_x = st->max_x, min_y = st->min_y, max_y = st->max_y;
      bool done = false;
      // all 8 neighbours of (cx, cy) are inside the tile while 1 <= cx - ox <= 62 (same for y)
      while (cx - ox >= 1 && cx - ox <= kTraceTile - 2 && cy - oy >= 1 && cy - oy <= kTraceTile - 2) {
        // detail::FindNextBoundaryPointDirection, ps/planar_region.cc:47-65 (out-of-bounds cells hold 0)
        const unsigned char* c = tile + (cx - ox) * kTraceTile + (cy - oy);
        unsigned same = 0u;
#pragma unroll
        for (int i = 0; i < 8; ++i) same |= (unsigned)c[kDx8[i] * kTraceTile + kDy8[i]] << i;
        int nd = -1;
        if (same) {
          const unsigned sh = (unsigned)(dir + 1) & 7u;
          const unsigned rot = ((same >> sh) | (same << (8u - sh))) & 0xffu;
          nd = (int)((sh + (unsigned)__ffs(rot) - 1u) & 7u);
        }
        if (nd < 0) { err = kErrTraceCheck; done = true; break; }  // CHECK_GE(new_direction_index, 0)
        dir = (nd + 4) & 7;
        curr_idx += kDx8[nd] * rows + kDy8[nd];
        cx += kDx8[nd];
        cy += kDy8[nd];
        if (found_start) {
          if (curr_idx == first) { done = true; break; }
          found_start = false;
        }
        if (nb >= cap) { err = kErrBoundaryCap; done = true; break; }
        boundary[nb] = curr_idx;
        if (nb == 0) first = curr_idx;
        ++nb;
        min_x = min(min_x, cx); max_x = max(max_x, cx);
        min_y = min(min_y, cy); max_y = max(max_y, cy);
        if (curr_idx == start_idx) found_start = true;
      }
      st->curr_idx = curr_idx; st->cx = cx; st->cy = cy; st->dir = dir; st->nb = nb; st->err = err;
      st->found_start = found_start ? 1 : 0; st->first = first; st->done = done ? 1 : 0;
      st->min_x = min_x; st->max_x = max_x; st->min_y = min_y; st->max_y = max_y;
      st->ox = cx - kTraceTile / 2; st->oy = cy - kTraceTile / 2;
    }
    __syncthreads();
    if (st->done) break;
  }
  if (tid == 0) {
    out->nb = st->nb; out->min_x = st->min_x; out->max_x = st->max_x;
    out->min_y = st->min_y; out->max_y = st->max_y; out->err = st->err;
  }
}

// Plane-frame (x, y) of a point as in PlanarConvexHull, ps/algorithms.h:527-538
// with eigenmath::PoseFromPlane restated (not in the reference tree):
// rotation = Quaternion::FromTwoVectors(UnitZ, normal), origin = -offset*normal.
struct PlaneFrame {
  float qx, qy, qz, qw;  // world_pose_plane rotation
  V3 ti;                 // translation of the inverse pose
};
__device__ __forceinline__ PlaneFrame make_plane_frame(const float plane[4]) {
  V3 n{plane[0], plane[1], plane[2]};
  V3 v0{0.0f, 0.0f, 1.0f};
  V3 v1 = n;
  float z = vsqn(n);
  if (z > 0.0f) v1 = vdiv(n, sqrtf(z));
  float c = vdot(v1, v0);
  PlaneFrame pf;
  if (c < -1.0f + FLT_EPSILON) {
    pf.qx = 1.0f; pf.qy = 0.0f; pf.qz = 0.0f; pf.qw = 0.0f;
  } else {
    V3 axis = vcross(v0, v1);
    float s = sqrtf((1.0f + c) * 2.0f);
    float invs = 1.0f / s;
    pf.qx = axis.x * invs; pf.qy = axis.y * invs; pf.qz = axis.z * invs; pf.qw = s * 0.5f;
  }
  V3 t = vscale(-plane[3], n);
  pf.ti = vneg(qrotate(-pf.qx, -pf.qy, -pf.qz, pf.qw, t));
  return pf;
}
__device__ __forceinline__ void plane_xy(const PlaneFrame& pf, V3 p, double* x, double* y) {
  V3 pp = vadd(qrotate(-pf.qx, -pf.qy, -pf.qz, pf.qw, p), pf.ti);
  *x = (double)pp.x;
  *y = (double)pp.y;
}

}  // namespace pcseg
