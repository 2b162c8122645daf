// pcseg_math.cuh -- device arithmetic shared by all kernels.
//
// Everything here is fp32, round-to-nearest, compiled with -fmad=false so that
// no multiply-add is contracted (the reference binary is baseline x86-64,
// .bazelrc:1).  Reduction orders follow Eigen 3.4.0's fixed-size unroller:
// sum/dot/squaredNorm/trace of 3 elements = e0 + (e1 + e2).
#pragma once
#include <cfloat>
#include <cmath>
#include <cuda_runtime.h>
#include "cr_trig.cuh"

namespace pcseg {

struct V3 {
  float x, y, z;
};

__device__ __forceinline__ V3 ld3(const float* __restrict__ p, int idx) {
  const float* q = p + 3 * (size_t)idx;
  return V3{__ldg(q), __ldg(q + 1), __ldg(q + 2)};
}
__device__ __forceinline__ V3 vsub(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 vadd(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 vscale(float s, V3 a) { return V3{s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ V3 vdiv(V3 a, float s) { return V3{a.x / s, a.y / s, a.z / s}; }
// IEEE division / square root as CALLS: the inline expansions (~12 / ~10 instructions each) make
// the eigen-solve ~5 KB larger, and the planar growth is bound by instruction fetch.
// Same operations, same results.
// COMPACT = true is used by the planar growth (measured 35.7 -> 34.6 ms), false by the
// issue-bound normals kernel (where the calls cost 5 %).
__device__ __noinline__ float div_rn_call(float a, float b) { return a / b; }
__device__ __noinline__ float sqrt_rn_call(float a) { return sqrtf(a); }
template <bool COMPACT>
__device__ __forceinline__ float div_rn(float a, float b) { return COMPACT ? div_rn_call(a, b) : a / b; }
template <bool COMPACT>
__device__ __forceinline__ float sqrt_rn(float a) { return COMPACT ? sqrt_rn_call(a) : sqrtf(a); }
template <bool COMPACT>
__device__ __forceinline__ V3 vdiv_c(V3 a, float s) {
  return V3{div_rn<COMPACT>(a.x, s), div_rn<COMPACT>(a.y, s), div_rn<COMPACT>(a.z, s)};
}
__device__ __forceinline__ V3 vneg(V3 a) { return V3{-a.x, -a.y, -a.z}; }
__device__ __forceinline__ float vdot(V3 a, V3 b) { return a.x * b.x + (a.y * b.y + a.z * b.z); }
__device__ __forceinline__ float vsqn(V3 a) { return a.x * a.x + (a.y * a.y + a.z * a.z); }
__device__ __forceinline__ V3 vcross(V3 a, V3 b) {
  return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
__device__ __forceinline__ bool finite1(float v) { return fabsf(v) <= FLT_MAX; }  // false for NaN/Inf
__device__ __forceinline__ bool vfinite(V3 a) { return finite1(a.x) && finite1(a.y) && finite1(a.z); }
__device__ __forceinline__ bool vhasnan(V3 a) { return (a.x != a.x) || (a.y != a.y) || (a.z != a.z); }

// libm float trig of the reference (Eigen computeDirect: atan2f, cosf, sinf;
// ps/planar_region.h:399 acosf) evaluated in double and rounded once.
__device__ __forceinline__ float cr_atan2f(float y, float x) { return (float)atan2((double)y, (double)x); }
__device__ __forceinline__ float cr_cosf(float x) { return (float)cos((double)x); }
__device__ __forceinline__ float cr_sinf(float x) { return (float)sin((double)x); }
__device__ __forceinline__ void cr_sincosf(float x, float* s, float* c) {
  double sd, cd;
  sincos((double)x, &sd, &cd);
  *s = (float)sd;
  *c = (float)cd;
}
__device__ __forceinline__ float cr_acosf(float x) { return (float)acos((double)x); }

// Hyperplane(n, e): offset = -n.dot(e); absDistance = |n.dot(p) + offset|
struct Plane {
  V3 n;
  float d;
};
__device__ __forceinline__ Plane make_plane(V3 n, V3 e) { return Plane{n, -vdot(n, e)}; }
__device__ __forceinline__ float plane_abs_distance(const Plane& pl, V3 p) { return fabsf(vdot(pl.n, p) + pl.d); }

// Quaternion * vector, Eigen QuaternionBase::_transformVector
__device__ __forceinline__ V3 qrotate(float qx, float qy, float qz, float qw, V3 v) {
  V3 qv{qx, qy, qz};
  V3 uv = vcross(qv, v);
  uv = vadd(uv, uv);
  return vadd(vadd(v, vscale(qw, uv)), vcross(qv, uv));
}

// ---------------------------------------------------------------------------
// Eigen 3.4.0 SelfAdjointEigenSolver<Matrix3f>::computeDirect (restated).
// Symmetric input given by its lower triangle; outputs eigenvalues (ascending)
// and eigenvector column 0 only (the reference reads eigenvalues()[0], [1] and
// eigenvectors().col(0): ps/plane_estimator.cc:204-207).
// ---------------------------------------------------------------------------
struct Sym3 {
  float m00, m10, m20, m11, m21, m22;
};

__device__ __forceinline__ V3 sym_col(const Sym3& s, int c) {
  if (c == 0) return V3{s.m00, s.m10, s.m20};
  if (c == 1) return V3{s.m10, s.m11, s.m21};
  return V3{s.m20, s.m21, s.m22};
}

template <bool COMPACT>
__device__ __forceinline__ void extract_kernel(const Sym3& t, V3* res, V3* representative) {
  int i0 = 0;
  float best = fabsf(t.m00);
  float v1 = fabsf(t.m11);
  if (v1 > best) { best = v1; i0 = 1; }
  float v2 = fabsf(t.m22);
  if (v2 > best) { best = v2; i0 = 2; }
  V3 r = sym_col(t, i0);
  V3 a = sym_col(t, (i0 + 1) % 3);
  V3 b = sym_col(t, (i0 + 2) % 3);
  V3 c0 = vcross(r, a);
  V3 c1 = vcross(r, b);
  float n0 = vsqn(c0);
  float n1 = vsqn(c1);
  *representative = r;
  if (n0 > n1) *res = vdiv_c<COMPACT>(c0, sqrt_rn<COMPACT>(n0));
  else *res = vdiv_c<COMPACT>(c1, sqrt_rn<COMPACT>(n1));
}

template <bool COMPACT>
__device__ __noinline__ void eig3_direct(const Sym3& mat, float evals[3], V3* evec0) {
  float shift = div_rn<COMPACT>(mat.m00 + (mat.m11 + mat.m22), 3.0f);
  Sym3 s = mat;
  s.m00 -= shift;
  s.m11 -= shift;
  s.m22 -= shift;
  float scale = fmaxf(fmaxf(fmaxf(fabsf(s.m00), fabsf(s.m10)), fmaxf(fabsf(s.m20), fabsf(s.m11))),
                      fmaxf(fabsf(s.m21), fabsf(s.m22)));
  if (scale > 0.0f) {
    s.m00 = div_rn<COMPACT>(s.m00, scale); s.m10 = div_rn<COMPACT>(s.m10, scale); s.m20 = div_rn<COMPACT>(s.m20, scale);
    s.m11 = div_rn<COMPACT>(s.m11, scale); s.m21 = div_rn<COMPACT>(s.m21, scale); s.m22 = div_rn<COMPACT>(s.m22, scale);
  }
  // computeRoots
  const float s_inv3 = 1.0f / 3.0f;
  const float s_sqrt3 = sqrtf(3.0f);
  float c0 = s.m00 * s.m11 * s.m22 + 2.0f * s.m10 * s.m20 * s.m21 - s.m00 * s.m21 * s.m21 -
             s.m11 * s.m20 * s.m20 - s.m22 * s.m10 * s.m10;
  float c1 = s.m00 * s.m11 - s.m10 * s.m10 + s.m00 * s.m22 - s.m20 * s.m20 + s.m11 * s.m22 -
             s.m21 * s.m21;
  float c2 = s.m00 + s.m11 + s.m22;
  float c2_over_3 = c2 * s_inv3;
  float a_over_3 = (c2 * c2_over_3 - c1) * s_inv3;
  a_over_3 = (a_over_3 < 0.0f) ? 0.0f : a_over_3;  // numext::maxi(a, 0)
  float half_b = 0.5f * (c0 + c2_over_3 * (2.0f * c2_over_3 * c2_over_3 - c1));
  float q = a_over_3 * a_over_3 * a_over_3 - half_b * half_b;
  q = (q < 0.0f) ? 0.0f : q;
  float rho = sqrt_rn<COMPACT>(a_over_3);
  float theta = cr_atan2f_pos(sqrt_rn<COMPACT>(q), half_b) * s_inv3;  // sqrt(q) >= +0: atan2f in [0, pi]
  float cos_theta, sin_theta;
  cr_sincosf_small(theta, &sin_theta, &cos_theta);
  evals[0] = c2_over_3 - rho * (cos_theta + s_sqrt3 * sin_theta);
  evals[1] = c2_over_3 - rho * (cos_theta - s_sqrt3 * sin_theta);
  evals[2] = c2_over_3 + 2.0f * rho * cos_theta;

  V3 e0;
  if ((evals[2] - evals[0]) <= FLT_EPSILON) {
    e0 = V3{1.0f, 0.0f, 0.0f};
  } else {
    float d0 = evals[2] - evals[1];
    float d1 = evals[1] - evals[0];
    int k = 0, l = 2;
    if (d0 > d1) { k = 2; l = 0; d0 = d1; }
    Sym3 t = s;
    float ek = (k == 0) ? evals[0] : evals[2];  // selects, not indexing: evals stays in registers
    t.m00 -= ek; t.m11 -= ek; t.m22 -= ek;
    V3 vk, vl;
    extract_kernel<COMPACT>(t, &vk, &vl);
    if (k == 0) {
      e0 = vk;
    } else {
      if (d0 <= 2.0f * FLT_EPSILON * d1) {
        float dd = vdot(vk, vl);
        vl = V3{vl.x - dd * vl.x, vl.y - dd * vl.y, vl.z - dd * vl.z};
        float z = vsqn(vl);
        if (z > 0.0f) vl = vdiv_c<COMPACT>(vl, sqrt_rn<COMPACT>(z));
      } else {
        t = s;
        float el = (l == 0) ? evals[0] : evals[2];
        t.m00 -= el; t.m11 -= el; t.m22 -= el;
        V3 dummy;
        extract_kernel<COMPACT>(t, &vl, &dummy);
      }
      e0 = vl;
    }
  }
  *evec0 = e0;
  evals[0] = evals[0] * scale + shift;
  evals[1] = evals[1] * scale + shift;
  evals[2] = evals[2] * scale + shift;
}

// ---------------------------------------------------------------------------
// PlaneEstimator state + ComputePlane (ps/plane_estimator.cc:46-65,184-229).
// ---------------------------------------------------------------------------
struct Estimator {
  float acc[6];
  V3 csum;
  float w;
  Plane plane;     // plane_ (its normal is the orientation hint)
  V3 centroid;
  float curvature;
  bool valid;
  bool up_to_date;
};

__device__ __forceinline__ void est_clear(Estimator& e) {
  for (int i = 0; i < 6; ++i) e.acc[i] = 0.0f;
  e.csum = V3{0.0f, 0.0f, 0.0f};
  e.w = 0.0f;
  e.plane = Plane{V3{1.0f, 0.0f, 0.0f}, 0.0f};
  e.centroid = V3{0.0f, 0.0f, 0.0f};
  e.curvature = 0.0f;
  e.valid = false;
  e.up_to_date = true;
}

__device__ __forceinline__ void est_add(Estimator& e, V3 p) {
  e.acc[0] += p.x * p.x;
  e.acc[1] += p.x * p.y;
  e.acc[2] += p.x * p.z;
  e.acc[3] += p.y * p.y;
  e.acc[4] += p.y * p.z;
  e.acc[5] += p.z * p.z;
  e.csum = vadd(e.csum, p);
  e.w += 1.0f;
  e.up_to_date = false;
}

// ComputePlane split in two: est_raw is everything that depends only on the
// sums (centroid, covariance, eigen-solve); est_apply is the part that also
// depends on the PREVIOUS plane normal (eigenvector sign, invalid-plane
// fallback).  The planar growth computes est_raw for several speculative
// re-fits concurrently and applies them in order (grow_planar_pipe.cuh).
struct RawFit {
  V3 centroid;
  V3 ev;          // eigenvector of the smallest eigenvalue
  float smallest; // eigenvalues()[0]
  float trace;
  bool valid;     // eigenvalues()[1] > FLT_MIN
};

// Caller guarantees w > 0 (the reference CHECK-aborts otherwise).
template <bool COMPACT = false>
__device__ __forceinline__ RawFit est_raw(const float acc[6], V3 csum, float w) {
  RawFit r;
  r.centroid = vdiv_c<COMPACT>(csum, w);
  float a0 = div_rn<COMPACT>(acc[0], w), a1 = div_rn<COMPACT>(acc[1], w), a2 = div_rn<COMPACT>(acc[2], w);
  float a3 = div_rn<COMPACT>(acc[3], w), a4 = div_rn<COMPACT>(acc[4], w), a5 = div_rn<COMPACT>(acc[5], w);
  Sym3 cov;
  cov.m00 = a0 - r.centroid.x * r.centroid.x;
  cov.m10 = a1 - r.centroid.x * r.centroid.y;
  cov.m20 = a2 - r.centroid.x * r.centroid.z;
  cov.m11 = a3 - r.centroid.y * r.centroid.y;
  cov.m21 = a4 - r.centroid.y * r.centroid.z;
  cov.m22 = a5 - r.centroid.z * r.centroid.z;
  float evals[3];
  eig3_direct<COMPACT>(cov, evals, &r.ev);
  r.valid = evals[1] > FLT_MIN;
  r.smallest = evals[0];
  r.trace = cov.m00 + (cov.m11 + cov.m22);
  return r;
}

__device__ __forceinline__ void est_apply(Estimator& e, const RawFit& r) {
  e.up_to_date = true;
  e.centroid = r.centroid;
  if (r.valid) {
    if (vdot(e.plane.n, r.ev) > 0.0f) e.plane = make_plane(r.ev, e.centroid);
    else e.plane = make_plane(vneg(r.ev), e.centroid);
    if (r.trace > r.smallest && r.smallest > FLT_MIN) e.curvature = fabsf(r.smallest / r.trace);
    else e.curvature = 0.0f;
    e.valid = true;
  } else {
    e.curvature = 0.0f;
    e.plane = make_plane(e.plane.n, e.centroid);
    e.valid = false;
  }
}

template <bool COMPACT = false>
__device__ __forceinline__ void est_compute(Estimator& e) {
  const RawFit r = est_raw<COMPACT>(e.acc, e.csum, e.w);
  est_apply(e, r);
}

// SetNormalOrientation, ps/plane_estimator.cc:113-126
__device__ __forceinline__ void est_set_orientation(Estimator& e, V3 hint) {
  if (!e.up_to_date) est_compute(e);
  if (e.valid) {
    if (vdot(e.plane.n, hint) < 0.0f) {
      e.plane.n = V3{e.plane.n.x * -1.0f, e.plane.n.y * -1.0f, e.plane.n.z * -1.0f};
      e.plane.d *= -1.0f;
    }
  } else {
    e.plane.n = hint;
  }
}

}  // namespace pcseg
