// planar_kernel.cuh -- the per-frame planar segmentation kernel (seed loop,
// FIFO growth, CheckAndFinalize).  See grow_planar.cuh / grow_planar_smem.cuh.
#pragma once
#include "grow_planar.cuh"
#include "grow_planar_smem.cuh"
#include "grow_planar_pipe.cuh"
#include "grow_planar_wave.cuh"
#include "grow_planar_wave2.cuh"
#include <type_traits>
#include "grow_cluster.cuh"

namespace pcseg {

// MODE = 3 : window-parallel growth with asynchronous re-fit warps and speculative commit
//            (grow_planar_wave2.cuh); BLOCK = W + 64; mask_words*4 + ring*20 + nstage*40 +
//            kW2Hash*4 + kW2Rej*4 bytes of dynamic shared memory.
// MODE = 2 : window-parallel speculative growth by the whole CTA (grow_planar_wave.cuh): labels
//            mirrored as a 1-bit/cell mask + FIFO ring + claim tables in dynamic shared memory
//            (mask_words*4 + ring*20 + W*40 + kWaveHash*4 + kWaveRej*4 bytes); BLOCK = W + 32.
// MODE = 1 : warp-specialised pipeline on chunks of 32 entries (grow_planar_pipe.cuh), same
//            mask + ring (mask_words*4 + ring*20 + 32*11*4 + kHashSize*4 bytes).
// MODE = 0 : everything through L2 (planar_grow_warp); any cloud size.
// FUSE = true : the frame's CTA goes on with SegmentRegions<ClusterRegion<>> (seeds all
//               descending or ca.seeds, id offset = the frame's plane count) as soon as its
//               own planar stage is done, instead of waiting for the slowest frame of the batch.
#ifdef PLANAR_MAXNREG
#define PLANAR_REG_ATTR __maxnreg__(PLANAR_MAXNREG)
#else
#ifndef PCSEG_W2_MINB
#define PCSEG_W2_MINB 2   // frames per SM the register budget of the MODE 3 kernel allows
#endif
#define PLANAR_REG_ATTR __launch_bounds__(MODE == 1 ? 384 : BLOCK, MODE == 3 ? PCSEG_W2_MINB : (MODE ? 2 : 1))  // pipeline: register budget of the 12-warp CTA (<= 85) for every warp count
#endif
template <int BLOCK, int MODE, bool FUSE>
__global__ void PLANAR_REG_ATTR planar_grow_kernel(PlanarArgs a, ClusterArgs ca) {
  extern __shared__ __align__(16) unsigned char dyn_smem[];
  const int f = blockIdx.x;
  const int tid = threadIdx.x;
  const int rows = a.rows, cols = a.cols;
  const int N = rows * cols;
  const float* __restrict__ xyz = a.xyz + (size_t)f * N * 3;
  const float* __restrict__ nrm = a.normals + (size_t)f * N * 3;
  volatile int* labels = a.labels + (size_t)f * N;
  volatile int* queue = a.queue + (size_t)f * a.qcap * 2;  // per-frame stride shared with the cluster stage ([qcap][2])
  int* claim = a.claim + (size_t)f * N;
  int* inliers = a.inliers + (size_t)f * N;
  int* boundary = a.boundary + (size_t)f * a.bcap;
  unsigned char* bflags = a.bflags + (size_t)f * a.bcap;
  PlanarMeta* meta = a.meta + (size_t)f * a.rcap;
  const int* seeds = nullptr;
  int n_seeds = N;
  if (a.seeds) {
    seeds = a.seeds_stride > 0 ? a.seeds + (size_t)f * a.seeds_stride : a.seeds;
    n_seeds = a.seeds_stride > 0 ? a.n_seeds_dev[f] : a.n_seeds;
  }

  __shared__ int s_warp[BLOCK / 32 + 1];
  __shared__ GrowResult s_res;
  __shared__ TraceResult s_trace;
  __shared__ float s_area;
  __shared__ float s_cross[3 * BLOCK];
  __shared__ double s_ext[4][BLOCK / 32];
  __shared__ unsigned char s_ttile[MODE >= 2 ? kTraceTile * kTraceTile : 4];
  __shared__ TraceState s_tstate;
  constexpr bool SMEM = MODE != 0;

  PlanarSmem sm;
  WaveSmem wsm;
  int tag_base = 0;
  int refit_last = 0;  // warp R: newest ticket seen on its slot
  W2Smem w2sm;
  __shared__ std::conditional_t<MODE == 3, W2Ctl, std::conditional_t<MODE == 2, WaveCtl, PipeCtl>> s_ctl;
  (void)w2sm;
  if constexpr (MODE == 3) {
    w2sm.mask = reinterpret_cast<unsigned*>(dyn_smem);
    w2sm.q_tag = reinterpret_cast<int*>(w2sm.mask + a.mask_words);
    w2sm.q_cell = w2sm.q_tag + a.ring;
    w2sm.q_xyz = reinterpret_cast<float*>(w2sm.q_cell + a.ring);
    w2sm.stage = w2sm.q_xyz + 3 * a.ring + ((a.mask_words + 5 * a.ring) & 1);  // 8-byte aligned
    w2sm.htab = reinterpret_cast<unsigned*>(w2sm.stage + 10 * a.nstage);
    w2sm.rej = w2sm.htab + kW2Hash;
    w2sm.ring = a.ring;
    w2sm.nstage = a.nstage;
    sm.mask = w2sm.mask;
    for (int i = tid; i < a.ring; i += BLOCK) w2sm.q_tag[i] = -1;
    for (int i = tid; i < kW2Hash; i += BLOCK) w2sm.htab[i] = kHashEmpty;
    for (int i = tid; i < kW2Rej; i += BLOCK) w2sm.rej[i] = kHashEmpty;
  }
  static_assert(MODE != 1 || BLOCK == 32 * kPipeWarps, "pipeline roles: warp 0 = A, 1 = S, 4-7 = B, 2-3 and 8.. = re-fit");
  static_assert(MODE != 2 || (BLOCK % 32 == 0 && (BLOCK - 32) * 4 <= (1 << kWaveKeyBits)), "wave: claim key bits");
  __shared__ std::conditional_t<MODE == 1, PipeBShared, int> s_bsh;
  (void)wsm; (void)refit_last; (void)s_bsh;
  if constexpr (MODE == 2) {
    wsm.mask = reinterpret_cast<unsigned*>(dyn_smem);
    wsm.q_tag = reinterpret_cast<int*>(wsm.mask + a.mask_words);
    wsm.q_cell = wsm.q_tag + a.ring;
    wsm.q_xyz = reinterpret_cast<float*>(wsm.q_cell + a.ring);
    wsm.stage = wsm.q_xyz + 3 * a.ring + ((a.mask_words + 5 * a.ring) & 1);  // 8-byte aligned
    wsm.htab = reinterpret_cast<unsigned*>(wsm.stage + 10 * (BLOCK - 32));
    wsm.rej = wsm.htab + kWaveHash;
    wsm.ring = a.ring;
    sm.mask = wsm.mask;
    for (int i = tid; i < a.ring; i += BLOCK) wsm.q_tag[i] = -1;
    for (int i = tid; i < kWaveHash; i += BLOCK) wsm.htab[i] = kHashEmpty;
    for (int i = tid; i < kWaveRej; i += BLOCK) wsm.rej[i] = kHashEmpty;
  }
  if constexpr (MODE == 1) {
    if (tid == 0) {
      s_ctl.ticket = 0;
      for (int i = 0; i < kPipeDW; ++i) { s_ctl.req_ticket[i] = 0; s_ctl.res_ticket[i] = 0; s_ctl.req_expect[i] = 0; }
    }
    sm.mask = reinterpret_cast<unsigned*>(dyn_smem);
    sm.q_tag = reinterpret_cast<int*>(sm.mask + a.mask_words);
    sm.q_cell = sm.q_tag + a.ring;
    sm.q_xyz = reinterpret_cast<float*>(sm.q_cell + a.ring);
    sm.stage = sm.q_xyz + 3 * a.ring;
    sm.htab = reinterpret_cast<unsigned*>(sm.stage + 32 * 11);
    sm.ring = a.ring;
    for (int i = tid; i < a.ring; i += BLOCK) sm.q_tag[i] = -1;
    for (int i = tid; i < kHashSize; i += BLOCK) sm.htab[i] = kHashEmpty;
  }
  if constexpr (SMEM) {
    // mask bit <=> labels[cell] == kUnlabeled
    const int lane = tid & 31;
    for (int b0 = (tid >> 5) * 32; b0 < a.mask_words * 32; b0 += BLOCK) {
      const int cell = b0 + lane;
      const bool u = cell < N && labels[cell] == kUnlabeled;
      const unsigned bits = __ballot_sync(0xffffffffu, u);
      if (lane == 0) sm.mask[b0 >> 5] = bits;
    }
    __syncthreads();
  }
  (void)tag_base;

  int seed_pos = n_seeds;
  int nreg = 0, inl_total = 0, b_total = 0, err = 0, steps = 0, refits = 0;
  bool any_failed = false;
  const float min_range_sq = a.disc_min_range * a.disc_min_range;  // eigenmath::Square
  const float max_range_sq = a.disc_max_range * a.disc_max_range;

#ifdef PCSEG_PROFILE
  long long kp_seed = 0, kp_fin = 0, kp_regions = 0, kp_tiny = 0;
  const long long kp_start = clock64();
#endif
  while (true) {
    PROF_T(k0);
    // ---- next usable seed (ps/segmentation.h:254-260).  A seed with a
    // non-finite point or normal fails its own plane test (NaN distance), adds
    // no inlier and leaves labels untouched: skipping it here is exact.
    int found = -1;
    while (seed_pos > 0 && found < 0) {
      int cand = -1;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        int pos = seed_pos - 1 - (tid + k * BLOCK);
        if (pos >= 0 && cand < 0) {
          int s = seed_at(seeds, N, pos);
          bool unl;
          if constexpr (SMEM) unl = mask_test(sm.mask, s);
          else unl = labels[s] == kUnlabeled;
          if (unl && vfinite(ld3(xyz, s)) && vfinite(ld3(nrm, s))) cand = pos;
        }
      }
      found = block_max<BLOCK>(cand, s_warp);
      if (found < 0) seed_pos = max(seed_pos - 4 * BLOCK, 0);
    }
    if (found < 0) break;
    PROF_T(k1);
#ifdef PCSEG_PROFILE
    kp_seed += k1 - k0;
    ++kp_regions;
#endif
    seed_pos = found;
    const int seed = seed_at(seeds, N, found);
    const int label_id = nreg + a.id_offset;
    const int base = inl_total;

    if constexpr (MODE == 3) {
      planar_grow_wave2<BLOCK>(a, xyz, nrm, a.labels + (size_t)f * N, a.queue + (size_t)f * a.qcap * 2, inliers, seed,
                               label_id, base, tag_base, w2sm, s_ctl, &s_res);
    } else if constexpr (MODE == 2) {
      planar_grow_wave<BLOCK>(a, xyz, nrm, a.labels + (size_t)f * N, a.queue + (size_t)f * a.qcap * 2, inliers, seed,
                              label_id, base, tag_base, wsm, s_ctl, &s_res);
    } else if constexpr (MODE == 1) {
      int* gqueue = a.queue + (size_t)f * a.qcap * 2;
      if (tid == 0) {
        const V3 sp = ld3(xyz, seed);
        sm.q_tag[0] = tag_base;
        sm.q_cell[0] = seed;
        sm.q_xyz[0] = sp.x; sm.q_xyz[1] = sp.y; sm.q_xyz[2] = sp.z;
        gqueue[0] = seed;
        s_ctl.tail_pub = 1;
        s_ctl.b_done = 0;
        s_ctl.b_ack = 0;
        s_ctl.a_spec = 0;
        s_ctl.s_done = 0;
        s_ctl.s_ack = 0;
        s_ctl.rb_epoch = 0;
        s_ctl.rb_chunk = 0;
        s_ctl.done = 0;
      }
      __syncthreads();
      const int warp = tid >> 5;
      if (warp == 0) pipe_warp_A(a, xyz, nrm, gqueue, seed, tag_base, sm, s_ctl, &s_res);
      else if (warp >= 4 && warp < 8)
        pipe_group_B(a, xyz, a.labels + (size_t)f * N, gqueue, inliers, label_id, base, tag_base, sm, s_ctl, s_bsh, tid - 128);
      else if (warp == 1) pipe_warp_S(a, xyz, gqueue, tag_base, sm, s_ctl);
      else if (kPipeWarps == 12) { if (warp != 8) pipe_warp_R(s_ctl, warp < 4 ? warp - 2 : warp - 7, refit_last); }
      else pipe_warp_R(s_ctl, warp < 4 ? warp - 2 : warp - 6, refit_last);
    } else {
      if (tid < 32) planar_grow_warp(a, xyz, nrm, labels, queue, claim, inliers, seed, label_id, base, &s_res);
    }
    __syncthreads();
    const int ninl = s_res.ninl;
    tag_base += s_res.tail;
    if (s_res.steps < 0) { err |= kErrWatchdog; break; }
    steps += s_res.steps;
    refits += s_res.refits;

    PROF_T(k2);
    // ---- CheckAndFinalize, ps/planar_region.h:189-224
    bool ok = ninl >= a.min_inliers;
    int nb = 0;
    if (ok) {
      // last inlier (FIFO order) with an in-bounds 8-neighbour of another label, :198-203
      int start_i = -1;
      for (int hi = ninl; hi > 0 && start_i < 0; hi -= BLOCK) {
        const int i = hi - 1 - tid;
        int q = -1;
        if (i >= 0 && initial_predecessor_dir(const_cast<const int*>(labels), rows, cols, inliers[base + i], label_id) >= 0) q = i;
        start_i = block_max<BLOCK>(q, s_warp);
      }
      if (start_i < 0) {
        ok = false;  // boundary_indices_ empty -> CheckMinRowsAndCols fails, :205-212
      } else {
        if constexpr (MODE >= 2) {
          trace_boundary_tiled<BLOCK>(const_cast<const int*>(labels), rows, cols, inliers[base + start_i], label_id,
                                      boundary + b_total, a.bcap - b_total, &s_trace, s_ttile, &s_tstate);
        } else if (tid == 0) {
          trace_boundary(const_cast<const int*>(labels), rows, cols, inliers[base + start_i], label_id, boundary + b_total,
                         a.bcap - b_total, &s_trace);
        }
        __syncthreads();
        nb = s_trace.nb;
        if (s_trace.err) {
          err |= s_trace.err;
          ok = false;
        }
        // detail::CheckMinRowsAndCols(boundary, rows, cols, 3, 3), ps/planar_region.cc:91-106
        if (!((s_trace.max_x - s_trace.min_x) > 3 && (s_trace.max_y - s_trace.min_y) > 3)) ok = false;
        if (nb == 0) ok = false;
      }
    }
    if (ok) {
      // PlanarConvexHull(...).Size() >= 3 (patch P1), ps/planar_region.h:214-219: the hull
      // of the plane-frame boundary points has >= 3 vertices iff they are not all collinear.
      const PlaneFrame pf = make_plane_frame(s_res.plane);
      const double kInf = 1.0e300;
      double ax = kInf, ay = kInf, bx = -kInf, by = -kInf;
      for (int k = tid; k < nb; k += BLOCK) {
        double x, y;
        plane_xy(pf, ld3(xyz, boundary[b_total + k]), &x, &y);
        if (x < ax || (x == ax && y < ay)) { ax = x; ay = y; }
        if (x > bx || (x == bx && y > by)) { bx = x; by = y; }
      }
      // lexicographic min / max over the CTA: warp shuffles, then one value per warp
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const double oax = __shfl_xor_sync(0xffffffffu, ax, o), oay = __shfl_xor_sync(0xffffffffu, ay, o);
        const double obx = __shfl_xor_sync(0xffffffffu, bx, o), oby = __shfl_xor_sync(0xffffffffu, by, o);
        if (oax < ax || (oax == ax && oay < ay)) { ax = oax; ay = oay; }
        if (obx > bx || (obx == bx && oby > by)) { bx = obx; by = oby; }
      }
      if ((tid & 31) == 0) {
        s_ext[0][tid >> 5] = ax; s_ext[1][tid >> 5] = ay; s_ext[2][tid >> 5] = bx; s_ext[3][tid >> 5] = by;
      }
      __syncthreads();
      ax = s_ext[0][0]; ay = s_ext[1][0]; bx = s_ext[2][0]; by = s_ext[3][0];
      for (int w = 1; w < BLOCK / 32; ++w) {
        const double oax = s_ext[0][w], oay = s_ext[1][w], obx = s_ext[2][w], oby = s_ext[3][w];
        if (oax < ax || (oax == ax && oay < ay)) { ax = oax; ay = oay; }
        if (obx > bx || (obx == bx && oby > by)) { bx = obx; by = oby; }
      }
      int off_line = 0;
      for (int k = tid; k < nb; k += BLOCK) {
        double x, y;
        plane_xy(pf, ld3(xyz, boundary[b_total + k]), &x, &y);
        const double cr = (bx - ax) * (y - ay) - (by - ay) * (x - ax);
        if (cr != 0.0) off_line = 1;
      }
      off_line = __syncthreads_or(off_line);
      if (!off_line) ok = false;
    }
    if (ok) {
      // PolygonArea, ps/algorithms.h:264-292: the cross products p_i x p_(i+1) are
      // computed by the whole CTA (batches of BLOCK, staged in shared memory so the
      // scattered point loads overlap); their SUM stays sequential fp32 in polygon order.
      {
        const int* b = boundary + b_total;
        V3 nsum{0.0f, 0.0f, 0.0f};
        for (int i0 = 0; i0 < nb; i0 += BLOCK) {
          const int i = i0 + tid;
          if (i < nb) {
            const V3 pa = ld3(xyz, b[i]);
            const V3 pb = ld3(xyz, b[(i + 1 == nb) ? 0 : i + 1]);
            const V3 c = vcross(pa, pb);
            s_cross[3 * tid] = c.x; s_cross[3 * tid + 1] = c.y; s_cross[3 * tid + 2] = c.z;
          }
          __syncthreads();
          if (tid == 0) {
            const int cnt = min(BLOCK, nb - i0);
            for (int k = 0; k < cnt; ++k)
              nsum = vadd(nsum, V3{s_cross[3 * k], s_cross[3 * k + 1], s_cross[3 * k + 2]});
          }
          __syncthreads();
        }
        if (tid == 0) s_area = 0.5f * sqrtf(vsqn(nsum));
      }
      __syncthreads();
      ok = s_area >= a.min_region_area;
    }
    if (ok && nreg >= a.rcap) {
      err |= kErrRegionCap;
      break;
    }
    if (ok) {
      // discontinuity side output, ps/planar_region.h:360-417 (per boundary entry; the
      // host turns the flags into the reference's set)
      for (int k = tid; k < nb; k += BLOCK) {
        const int idx = boundary[b_total + k];
        const V3 pc = ld3(xyz, idx);
        unsigned char flag = 0;
        const float r2 = vsqn(pc);
        const int x = idx / rows, y = idx % rows;
        if (r2 > min_range_sq && r2 < max_range_sq && x > 0 && x < cols - 1 && y > 0 && y < rows - 1) {
#pragma unroll
          for (int d = 0; d < 4; ++d) {
            const int ni = idx + kDx4[d] * rows + kDy4[d];
            const V3 pn = ld3(xyz, ni);
            if ((pc.z != pc.z) || (pn.z != pn.z)) continue;
            if (labels[ni] != label_id) continue;
            const V3 delta = qrotate(a.qx, a.qy, a.qz, a.qw, vsub(pc, pn));
            const float cosang = vdot(ld3(nrm, idx), ld3(nrm, ni));
            const float deg = (float)((double)cr_acosf(cosang) * (180.0 / 3.14159265358979323846));
            if (fabsf(deg) < a.disc_angle && fabsf(delta.z) < a.disc_z_diff) continue;
            if (fabsf(delta.z) / sqrtf(vsqn(delta)) < a.disc_z_ratio) continue;
            flag = 1;
          }
        }
        bflags[b_total + k] = flag;
      }
      if (tid == 0) {
        PlanarMeta m;
        m.inlier_begin = base; m.inlier_count = ninl;
        m.boundary_begin = b_total; m.boundary_count = nb;
        m.seed_index = seed; m.label_id = label_id;
        for (int i = 0; i < 4; ++i) m.plane[i] = s_res.plane[i];
        for (int i = 0; i < 3; ++i) m.centroid[i] = s_res.centroid[i];
        m.area = s_area;
        for (int i = 0; i < 10; ++i) m.est[i] = s_res.est[i];
        meta[nreg] = m;
      }
      ++nreg;
      inl_total += ninl;
      b_total += nb;
    } else {
      // ps/segmentation.h:280-282
      for (int i = tid; i < ninl; i += BLOCK) labels[inliers[base + i]] = kAlreadyExamedPoint;
      any_failed = any_failed || (ninl > 0);
    }
    __syncthreads();
#ifdef PCSEG_PROFILE
    { long long k3 = clock64(); kp_fin += k3 - k2; if (ninl < a.min_inliers) ++kp_tiny; }
#endif
    if (err & kErrBoundaryCap) break;
  }
  __syncthreads();
  // ps/segmentation.h:287-291
  if (any_failed) {
    for (int i = tid; i < N; i += BLOCK)
      if (labels[i] == kAlreadyExamedPoint) labels[i] = kUnlabeled;
  }
#ifdef PCSEG_PROFILE
  if (tid == 0 && a.prof) {
    a.prof[(size_t)f * 16 + 13] += kp_seed;
    a.prof[(size_t)f * 16 + 14] += kp_fin;
    a.prof[(size_t)f * 16 + 15] = clock64() - kp_start;  // planar stage of this frame, cycles
    (void)kp_regions; (void)kp_tiny;
  }
  const long long kp_planar_end = clock64();
#endif
  if (tid == 0) {
    a.stats[f].n_planar = nreg;
    a.stats[f].planar_inliers = inl_total;
    a.stats[f].boundary_len = b_total;
    a.stats[f].error_flags |= err;
    a.stats[f].grow_steps += steps;
    a.stats[f].refits += refits;
  }
  if constexpr (FUSE) {
    __syncthreads();  // the frame's labels (written by this CTA) are final
    if (ca.half_window == 1) cluster_grow_frame<BLOCK, true>(ca, f, nreg, s_warp);
    else cluster_grow_frame<BLOCK, false>(ca, f, nreg, s_warp);
#ifdef PCSEG_PROFILE
    if (tid == 0 && a.prof) a.prof[(size_t)f * 16 + 7] = clock64() - kp_planar_end;  // fused cluster stage, cycles
#endif
  }
}

}  // namespace pcseg
