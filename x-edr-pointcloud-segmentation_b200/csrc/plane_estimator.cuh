// plane_estimator.cuh -- mobility::PlaneEstimator (ps/plane_estimator.h:37-120,
// ps/plane_estimator.cc:46-229) as stand-alone device entry points.
//
//   fit_planes_csr_kernel      north_star item (2): least-squares plane fits of many regions at
//                              once.  One warp per region; the region's points are gathered 32
//                              at a time (coalesced index loads, independent point loads), their
//                              moment products staged in shared memory, and lane j < 10 adds
//                              accumulator j in the REFERENCE'S INSERTION ORDER (AddPoints over
//                              the index list, ps/plane_estimator.h:97-103: sequential fp32 sums,
//                              order-sensitive by the authors' own note, plane_estimator_test.cc:34-36);
//                              one 3x3 symmetric eigen-solve per region (ComputePlane, :184-229).
//   estimator_replay_kernel    the whole class, op by op, on one thread: weighted AddPoint (:67-77),
//                              SetNormalOrientation (:113-126), Clear (:46-53), Merge (:128-140),
//                              Transform (:142-182), lazy accessors.  Used for parity against the
//                              reference's golden vectors (plane_estimator_test.cc:40-254) and by
//                              the host-side PlaneEstimator mirror.
//   estimator_transform_kernel / estimator_merge_kernel
//                              Transform / Merge of many accumulator states (10 floats each) at once.
#pragma once
#include "pcseg_math.cuh"

namespace pcseg {

// Accumulator state as exported everywhere in this library: acc[6], sum xyz, weight.
struct EstState {
  float v[10];
};

__device__ __forceinline__ void est_add_weighted(Estimator& e, V3 p, float w) {  // .cc:67-77
  e.acc[0] += p.x * p.x * w;
  e.acc[1] += p.x * p.y * w;
  e.acc[2] += p.x * p.z * w;
  e.acc[3] += p.y * p.y * w;
  e.acc[4] += p.y * p.z * w;
  e.acc[5] += p.z * p.z * w;
  e.csum = vadd(e.csum, V3{p.x * w, p.y * w, p.z * w});
  e.w += w;
  e.up_to_date = false;
}

__device__ __forceinline__ void est_merge(Estimator& e, const Estimator& o) {  // .cc:128-133
  for (int i = 0; i < 6; ++i) e.acc[i] += o.acc[i];
  e.csum = vadd(e.csum, o.csum);
  e.w += o.w;
  e.up_to_date = false;
}

// Transform(new_pose_current), .cc:142-182.  pose: quaternion xyzw + translation, doubles
// (eigenmath::Pose3d): rotationMatrix() in double then cast<float>() (:159-160); the pose cast to
// float before it moves the centroid (:164).  Caller guarantees w > 0 (CHECK_GT, :143).
__device__ __forceinline__ void est_transform(Estimator& e, const double* pose) {
  if (!e.up_to_date) e.centroid = vdiv(e.csum, e.w);
  const float a0 = e.acc[0] / e.w, a1 = e.acc[1] / e.w, a2 = e.acc[2] / e.w;
  const float a3 = e.acc[3] / e.w, a4 = e.acc[4] / e.w, a5 = e.acc[5] / e.w;
  const V3 c = e.centroid;
  float cov[3][3];
  cov[0][0] = a0 - c.x * c.x;
  cov[1][0] = a1 - c.x * c.y;
  cov[2][0] = a2 - c.x * c.z;
  cov[1][1] = a3 - c.y * c.y;
  cov[2][1] = a4 - c.y * c.z;
  cov[2][2] = a5 - c.z * c.z;
  cov[0][1] = cov[1][0]; cov[0][2] = cov[2][0]; cov[1][2] = cov[2][1];
  const double qx = pose[0], qy = pose[1], qz = pose[2], qw = pose[3];
  // Quaterniond::toRotationMatrix; explicit _rn intrinsics: no double FMA contraction
  const double tx = __dmul_rn(2.0, qx), ty = __dmul_rn(2.0, qy), tz = __dmul_rn(2.0, qz);
  const double twx = __dmul_rn(tx, qw), twy = __dmul_rn(ty, qw), twz = __dmul_rn(tz, qw);
  const double txx = __dmul_rn(tx, qx), txy = __dmul_rn(ty, qx), txz = __dmul_rn(tz, qx);
  const double tyy = __dmul_rn(ty, qy), tyz = __dmul_rn(tz, qy), tzz = __dmul_rn(tz, qz);
  float rot[3][3];
  rot[0][0] = (float)__dsub_rn(1.0, __dadd_rn(tyy, tzz)); rot[0][1] = (float)__dsub_rn(txy, twz); rot[0][2] = (float)__dadd_rn(txz, twy);
  rot[1][0] = (float)__dadd_rn(txy, twz); rot[1][1] = (float)__dsub_rn(1.0, __dadd_rn(txx, tzz)); rot[1][2] = (float)__dsub_rn(tyz, twx);
  rot[2][0] = (float)__dsub_rn(txz, twy); rot[2][1] = (float)__dadd_rn(tyz, twx); rot[2][2] = (float)__dsub_rn(1.0, __dadd_rn(txx, tyy));
  // rotation * covariance * rotation^T: coefficient products reduced e0 + (e1 + e2)
  float rc[3][3], out[3][3];
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int k = 0; k < 3; ++k) rc[r][k] = rot[r][0] * cov[0][k] + (rot[r][1] * cov[1][k] + rot[r][2] * cov[2][k]);
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int k = 0; k < 3; ++k) out[r][k] = rc[r][0] * rot[k][0] + (rc[r][1] * rot[k][1] + rc[r][2] * rot[k][2]);
  const V3 nc = vadd(qrotate((float)qx, (float)qy, (float)qz, (float)qw, c),
                     V3{(float)pose[4], (float)pose[5], (float)pose[6]});
  e.acc[0] = out[0][0] + nc.x * nc.x;
  e.acc[1] = out[1][0] + nc.x * nc.y;
  e.acc[2] = out[2][0] + nc.x * nc.z;
  e.acc[3] = out[1][1] + nc.y * nc.y;
  e.acc[4] = out[2][1] + nc.y * nc.z;
  e.acc[5] = out[2][2] + nc.z * nc.z;
  for (int i = 0; i < 6; ++i) e.acc[i] *= e.w;
  e.csum = V3{nc.x * e.w, nc.y * e.w, nc.z * e.w};
  e.up_to_date = false;
}

// Query row shared with the oracle: {valid, curvature, centroid xyz, normal xyz, offset, weight}.
// With zero weight nothing is computed (Plane() right after Clear(), quirk Q14; ComputePlane
// would CHECK-abort, which the replay reports as status -3 before it gets here).
__device__ __forceinline__ void est_query(Estimator& e, float* o) {
  if (!(e.w > 0.0f)) {
    o[0] = e.valid ? 1.0f : 0.0f;
    o[1] = 0.0f; o[2] = 0.0f; o[3] = 0.0f; o[4] = 0.0f;
  } else {
    if (!e.up_to_date) est_compute(e);
    o[0] = e.valid ? 1.0f : 0.0f;
    o[1] = e.curvature;
    o[2] = e.centroid.x; o[3] = e.centroid.y; o[4] = e.centroid.z;
  }
  o[5] = e.plane.n.x; o[6] = e.plane.n.y; o[7] = e.plane.n.z; o[8] = e.plane.d;
  o[9] = e.w;
}

struct ReplayArgs {
  const int* ops;       // [n_ops]
  const float* args;    // [n_ops][4]
  int n_ops;
  const double* poses;  // [n_poses][7]
  int n_poses;
  float* out;           // [cap][10]
  int cap;
  int* result;          // [1]: number of queries, or < 0 (bad op: -2, reference CHECK: -3, capacity: -1)
};

__global__ void estimator_replay_kernel(ReplayArgs a) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  Estimator slots[4];
  for (int i = 0; i < 4; ++i) est_clear(slots[i]);
  int cur = 0, nq = 0, status = 0;
  for (int i = 0; i < a.n_ops && status == 0; ++i) {
    Estimator& e = slots[cur];
    const V3 p{a.args[4 * i], a.args[4 * i + 1], a.args[4 * i + 2]};
    const float w = a.args[4 * i + 3];
    switch (a.ops[i]) {
      case 0: est_add(e, p); break;
      case 1: est_add_weighted(e, p, w); break;
      case 2:
        if (!e.up_to_date && !(e.w > 0.0f)) { status = -3; break; }
        est_set_orientation(e, p);
        break;
      case 3:
        if (nq >= a.cap) { status = -1; break; }
        if (!e.up_to_date && !(e.w > 0.0f)) { status = -3; break; }
        est_query(e, a.out + 10 * nq);
        ++nq;
        break;
      case 4: est_clear(e); break;
      case 5: {
        const int k = (int)p.x;
        if (k < 0 || k >= 4) status = -2; else cur = k;
        break;
      }
      case 6: {
        const int k = (int)p.x;
        if (k < 0 || k >= 4 || k == cur) status = -2; else est_merge(e, slots[k]);
        break;
      }
      case 7: {
        const int pi = (int)p.x;
        if (pi < 0 || pi >= a.n_poses) { status = -2; break; }
        if (!(e.w > 0.0f)) { status = -3; break; }
        est_transform(e, a.poses + 7 * pi);
        break;
      }
      case 8: {
        const int k = (int)p.x, pi = (int)p.y;
        if (k < 0 || k >= 4 || k == cur || pi < 0 || pi >= a.n_poses) { status = -2; break; }
        if (!(slots[k].w > 0.0f)) { status = -3; break; }
        Estimator o = slots[k];
        est_transform(o, a.poses + 7 * pi);
        est_merge(e, o);
        break;
      }
      default: status = -2;
    }
  }
  *a.result = status ? status : nq;
}

// ---- batched fit over CSR index sets -------------------------------------------------------
struct FitArgs {
  const float* xyz;      // [n_points][3]
  const int* offsets;    // [n_regions + 1]
  const int* indices;    // [offsets[n_regions]]
  const float* weights;  // nullable: per index entry (weighted AddPoint)
  const float* hints;    // nullable: [n_regions][3] SetNormalOrientation before the points
  int n_regions;
  float* query;          // [n_regions][10]
  float* state;          // nullable: [n_regions][10]
};

constexpr int kFitWarps = 8;

__global__ void __launch_bounds__(32 * kFitWarps) fit_planes_csr_kernel(FitArgs a) {
  __shared__ float s_stage[kFitWarps][32 * 11];
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* stage = s_stage[warp];
  for (int r = blockIdx.x * kFitWarps + warp; r < a.n_regions; r += gridDim.x * kFitWarps) {
    const int b = a.offsets[r], e = a.offsets[r + 1];
    float accj = 0.0f;  // lane j < 10: accumulator j
    for (int i0 = b; i0 < e; i0 += 32) {
      const int i = i0 + lane;
      if (i < e) {
        const V3 p = ld3(a.xyz, a.indices[i]);
        float* st = stage + lane * 11;
        if (a.weights) {
          const float w = a.weights[i];
          st[0] = p.x * p.x * w; st[1] = p.x * p.y * w; st[2] = p.x * p.z * w;
          st[3] = p.y * p.y * w; st[4] = p.y * p.z * w; st[5] = p.z * p.z * w;
          st[6] = p.x * w; st[7] = p.y * w; st[8] = p.z * w; st[9] = w;
        } else {
          st[0] = p.x * p.x; st[1] = p.x * p.y; st[2] = p.x * p.z;
          st[3] = p.y * p.y; st[4] = p.y * p.z; st[5] = p.z * p.z;
          st[6] = p.x; st[7] = p.y; st[8] = p.z; st[9] = 1.0f;
        }
      }
      __syncwarp();
      if (lane < 10) {
        const int m = min(32, e - i0);
        for (int k = 0; k < m; ++k) accj += stage[k * 11 + lane];
      }
      __syncwarp();
    }
    // every lane gets the ten sums; the solve runs redundantly, lane 0 writes
    Estimator est;
    est_clear(est);
#pragma unroll
    for (int j = 0; j < 6; ++j) est.acc[j] = __shfl_sync(FULL, accj, j);
    est.csum = V3{__shfl_sync(FULL, accj, 6), __shfl_sync(FULL, accj, 7), __shfl_sync(FULL, accj, 8)};
    est.w = __shfl_sync(FULL, accj, 9);
    if (a.hints) est.plane.n = ld3(a.hints, r);  // SetNormalOrientation on a cleared estimator: plane invalid -> hint
    est.up_to_date = !(e > b);
    float q[10];
    est_query(est, q);
    if (lane == 0) {
      for (int j = 0; j < 10; ++j) a.query[10 * (size_t)r + j] = q[j];
      if (a.state) {
        for (int j = 0; j < 6; ++j) a.state[10 * (size_t)r + j] = est.acc[j];
        a.state[10 * (size_t)r + 6] = est.csum.x; a.state[10 * (size_t)r + 7] = est.csum.y;
        a.state[10 * (size_t)r + 8] = est.csum.z; a.state[10 * (size_t)r + 9] = est.w;
      }
    }
  }
}

// ---- Transform / Merge of many accumulator states ------------------------------------------
// states[i] <- Transform(states[i], pose i or the shared pose 0); err |= 1 where w <= 0 (CHECK).
__global__ void __launch_bounds__(128) estimator_transform_kernel(float* states, int n, const double* poses,
                                                                  int pose_stride, int* err) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Estimator e;
  est_clear(e);
  float* s = states + 10 * (size_t)i;
  for (int j = 0; j < 6; ++j) e.acc[j] = s[j];
  e.csum = V3{s[6], s[7], s[8]};
  e.w = s[9];
  e.up_to_date = false;
  if (!(e.w > 0.0f)) { atomicOr(err, 1); return; }
  est_transform(e, poses + (size_t)pose_stride * i);
  for (int j = 0; j < 6; ++j) s[j] = e.acc[j];
  s[6] = e.csum.x; s[7] = e.csum.y; s[8] = e.csum.z; s[9] = e.w;
}

// dst[i] <- Merge(dst[i], src[i] [transformed by pose i first]); elementwise sums (.cc:128-140)
__global__ void __launch_bounds__(128) estimator_merge_kernel(float* dst, const float* src, int n, const double* poses,
                                                              int pose_stride, int* err) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Estimator o;
  est_clear(o);
  const float* s = src + 10 * (size_t)i;
  for (int j = 0; j < 6; ++j) o.acc[j] = s[j];
  o.csum = V3{s[6], s[7], s[8]};
  o.w = s[9];
  o.up_to_date = false;
  if (poses) {
    if (!(o.w > 0.0f)) { atomicOr(err, 1); return; }
    est_transform(o, poses + (size_t)pose_stride * i);
  }
  float* d = dst + 10 * (size_t)i;
  for (int j = 0; j < 6; ++j) d[j] += o.acc[j];
  d[6] += o.csum.x; d[7] += o.csum.y; d[8] += o.csum.z; d[9] += o.w;
}

}  // namespace pcseg
