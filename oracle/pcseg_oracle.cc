// pcseg_oracle.cc -- CPU ORACLE (test infrastructure, never shipped, never
// linked into libpcseg_b200.so).  Plain C++17 restatement of the reference's
// geometric segmentation hot path; each function cites the reference
// file:line it follows ("ps/" = pointcloud_segmentation/ in the reference).
//
// Build: g++ -std=c++17 -O2 -ffp-contract=off -fno-fast-math (no FMA
// contraction: the reference is built for baseline x86-64, .bazelrc:1).
//
// Third-party arithmetic that is NOT in the reference tree and is restated
// here from its published algorithm (bit-level parity with the real
// dependency is unpinned):
//   * Eigen 3.4.0 (WORKSPACE:46-52): 3-vector reductions e0+(e1+e2),
//     Hyperplane, cross, Quaternion*vector,
//     SelfAdjointEigenSolver<Matrix3f>::computeDirect.
//   * x_edr_eigenmath 1.0.0 (WORKSPACE:66-74): PoseFromPlane, Square, Degrees.
//   * x_edr_mobility_collision 1.0.0 (WORKSPACE:86-93): ConvexHull (restated
//     as Andrew's monotone chain, CCW, collinear points dropped).
//   * libm float trig (atan2f/cosf/sinf/acosf): by default evaluated in double
//     and rounded once to float (== a correctly-rounded libm such as glibc >=
//     2.41 CORE-MATH up to double-rounding); oracle_set_native_libm(1) switches
//     to this box's native float libm.
#include "pcseg_oracle.h"

#include <algorithm>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <map>
#include <queue>
#include <set>
#include <utility>
#include <unordered_map>
#include <vector>

namespace {

constexpr int kUnlabeled = -1;                       // ps/segmentation.h:36
constexpr int kAlreadyInUnexaminedPointsQueue = -3;  // ps/segmentation.h:41
constexpr int kAlreadyExamedPoint = -4;              // ps/segmentation.h:43

bool g_native_libm = false;

inline float o_atan2f(float y, float x) {
  return g_native_libm ? atan2f(y, x)
                       : static_cast<float>(atan2(static_cast<double>(y),
                                                  static_cast<double>(x)));
}
inline float o_cosf(float x) {
  return g_native_libm ? cosf(x)
                       : static_cast<float>(cos(static_cast<double>(x)));
}
inline float o_sinf(float x) {
  return g_native_libm ? sinf(x)
                       : static_cast<float>(sin(static_cast<double>(x)));
}
inline float o_acosf(float x) {
  return g_native_libm ? acosf(x)
                       : static_cast<float>(acos(static_cast<double>(x)));
}

// ---------------------------------------------------------------------------
// Eigen-3.4.0-ordered 3-vector arithmetic.
// ---------------------------------------------------------------------------
struct V3 {
  float x, y, z;
  float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
inline V3 operator-(const V3& a, const V3& b) {
  return {a.x - b.x, a.y - b.y, a.z - b.z};
}
inline V3 operator+(const V3& a, const V3& b) {
  return {a.x + b.x, a.y + b.y, a.z + b.z};
}
inline V3 operator*(float s, const V3& a) { return {s * a.x, s * a.y, s * a.z}; }
inline V3 operator/(const V3& a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline V3 neg(const V3& a) { return {-a.x, -a.y, -a.z}; }
// redux unroller for 3 elements: e0 + (e1 + e2)
inline float dot(const V3& a, const V3& b) {
  return a.x * b.x + (a.y * b.y + a.z * b.z);
}
inline float sqnorm(const V3& a) { return a.x * a.x + (a.y * a.y + a.z * a.z); }
inline V3 cross(const V3& a, const V3& b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
inline bool all_finite(const V3& a) {
  return std::isfinite(a.x) && std::isfinite(a.y) && std::isfinite(a.z);
}
inline bool has_nan(const V3& a) {
  return std::isnan(a.x) || std::isnan(a.y) || std::isnan(a.z);
}
inline V3 normalized(const V3& a) {
  float z = sqnorm(a);
  if (z > 0.0f) return a / sqrtf(z);
  return a;
}

// Eigen::Hyperplane<float,3>: coeffs = (normal, offset)
struct Plane {
  V3 n;
  float d;
  Plane() : n{1.0f, 0.0f, 0.0f}, d(0.0f) {}
  Plane(const V3& normal, const V3& e) : n(normal), d(-dot(normal, e)) {}
  float signedDistance(const V3& p) const { return dot(n, p) + d; }
  float absDistance(const V3& p) const { return fabsf(signedDistance(p)); }
};

struct Quat {
  float x, y, z, w;
};
// Eigen QuaternionBase::_transformVector
inline V3 rotate(const Quat& q, const V3& v) {
  V3 qv{q.x, q.y, q.z};
  V3 uv = cross(qv, v);
  uv = uv + uv;
  return (v + q.w * uv) + cross(qv, uv);
}

struct Cloud3 {  // Cloud<Vector3f, ColMajor> view, ps/cloud.h:74-178
  const float* data;
  int rows, cols;
  V3 at(int index) const {
    return {data[3 * index], data[3 * index + 1], data[3 * index + 2]};
  }
  V3 at(int row, int col) const { return at(col * rows + row); }
  int size() const { return rows * cols; }
  int lin(int row, int col) const { return col * rows + row; }
};

// ---------------------------------------------------------------------------
// Eigen 3.4.0 SelfAdjointEigenSolver<Matrix3f>::computeDirect restated
// (Eigen/src/Eigenvalues/SelfAdjointEigenSolver.h,
//  direct_selfadjoint_eigenvalues<SolverType,3,false>).  m is column-major.
// ---------------------------------------------------------------------------
struct M3 {
  float m[9];
  float& operator()(int r, int c) { return m[c * 3 + r]; }
  float operator()(int r, int c) const { return m[c * 3 + r]; }
  V3 col(int c) const { return {m[c * 3], m[c * 3 + 1], m[c * 3 + 2]}; }
  void set_col(int c, const V3& v) {
    m[c * 3] = v.x;
    m[c * 3 + 1] = v.y;
    m[c * 3 + 2] = v.z;
  }
};

void compute_roots(const M3& m, float roots[3]) {
  const float s_inv3 = 1.0f / 3.0f;
  const float s_sqrt3 = sqrtf(3.0f);
  float c0 = m(0, 0) * m(1, 1) * m(2, 2) + 2.0f * m(1, 0) * m(2, 0) * m(2, 1) -
             m(0, 0) * m(2, 1) * m(2, 1) - m(1, 1) * m(2, 0) * m(2, 0) -
             m(2, 2) * m(1, 0) * m(1, 0);
  float c1 = m(0, 0) * m(1, 1) - m(1, 0) * m(1, 0) + m(0, 0) * m(2, 2) -
             m(2, 0) * m(2, 0) + m(1, 1) * m(2, 2) - m(2, 1) * m(2, 1);
  float c2 = m(0, 0) + m(1, 1) + m(2, 2);
  float c2_over_3 = c2 * s_inv3;
  float a_over_3 = (c2 * c2_over_3 - c1) * s_inv3;
  a_over_3 = std::max(a_over_3, 0.0f);
  float half_b = 0.5f * (c0 + c2_over_3 * (2.0f * c2_over_3 * c2_over_3 - c1));
  float q = a_over_3 * a_over_3 * a_over_3 - half_b * half_b;
  q = std::max(q, 0.0f);
  float rho = sqrtf(a_over_3);
  float theta = o_atan2f(sqrtf(q), half_b) * s_inv3;
  float cos_theta = o_cosf(theta);
  float sin_theta = o_sinf(theta);
  roots[0] = c2_over_3 - rho * (cos_theta + s_sqrt3 * sin_theta);
  roots[1] = c2_over_3 - rho * (cos_theta - s_sqrt3 * sin_theta);
  roots[2] = c2_over_3 + 2.0f * rho * cos_theta;
}

void extract_kernel(const M3& mat, V3* res, V3* representative) {
  int i0 = 0;
  float best = fabsf(mat(0, 0));
  for (int i = 1; i < 3; ++i) {
    float v = fabsf(mat(i, i));
    if (v > best) {
      best = v;
      i0 = i;
    }
  }
  *representative = mat.col(i0);
  V3 c0 = cross(*representative, mat.col((i0 + 1) % 3));
  V3 c1 = cross(*representative, mat.col((i0 + 2) % 3));
  float n0 = sqnorm(c0);
  float n1 = sqnorm(c1);
  if (n0 > n1)
    *res = c0 / sqrtf(n0);
  else
    *res = c1 / sqrtf(n1);
}

void eig3_direct(const M3& mat, float evals[3], M3* evecs) {
  // trace = e0 + (e1 + e2)
  float shift = (mat(0, 0) + (mat(1, 1) + mat(2, 2))) / 3.0f;
  M3 s;
  for (int c = 0; c < 3; ++c)
    for (int r = 0; r < 3; ++r) s(r, c) = (r >= c) ? mat(r, c) : mat(c, r);
  s(0, 0) -= shift;
  s(1, 1) -= shift;
  s(2, 2) -= shift;
  float scale = fabsf(s.m[0]);
  for (int i = 1; i < 9; ++i) scale = std::max(scale, fabsf(s.m[i]));
  if (scale > 0.0f)
    for (int i = 0; i < 9; ++i) s.m[i] /= scale;
  compute_roots(s, evals);
  if ((evals[2] - evals[0]) <= FLT_EPSILON) {
    for (int i = 0; i < 9; ++i) evecs->m[i] = 0.0f;
    (*evecs)(0, 0) = (*evecs)(1, 1) = (*evecs)(2, 2) = 1.0f;
  } else {
    M3 tmp = s;
    float d0 = evals[2] - evals[1];
    float d1 = evals[1] - evals[0];
    int k = 0, l = 2;
    if (d0 > d1) {
      std::swap(k, l);
      d0 = d1;
    }
    V3 vk, vl;
    tmp(0, 0) -= evals[k];
    tmp(1, 1) -= evals[k];
    tmp(2, 2) -= evals[k];
    extract_kernel(tmp, &vk, &vl);
    if (d0 <= 2.0f * FLT_EPSILON * d1) {
      float dd = dot(vk, vl);
      vl = V3{vl.x - dd * vl.x, vl.y - dd * vl.y, vl.z - dd * vl.z};
      float z = sqnorm(vl);
      if (z > 0.0f) vl = vl / sqrtf(z);
    } else {
      tmp = s;
      tmp(0, 0) -= evals[l];
      tmp(1, 1) -= evals[l];
      tmp(2, 2) -= evals[l];
      V3 dummy;
      extract_kernel(tmp, &vl, &dummy);
    }
    evecs->set_col(k, vk);
    evecs->set_col(l, vl);
    evecs->set_col(1, normalized(cross(evecs->col(2), evecs->col(0))));
  }
  for (int i = 0; i < 3; ++i) {
    evals[i] *= scale;
    evals[i] += shift;
  }
}

// ---------------------------------------------------------------------------
// PlaneEstimator, ps/plane_estimator.h:37-120, ps/plane_estimator.cc
// ---------------------------------------------------------------------------
class PlaneEstimator {
 public:
  PlaneEstimator() { Clear(); }
  void Clear() {  // .cc:46-53
    cumulative_weights_ = 0.0f;
    plane_up_to_date_ = true;
    plane_valid_ = false;
    cumulative_centroid_ = {0, 0, 0};
    for (float& a : accumulator_) a = 0.0f;
    plane_.n = {1.0f, 0.0f, 0.0f};
    plane_.d = 0.0f;
  }
  void AddPoint(const V3& p) {  // .cc:55-65
    accumulator_[0] += p.x * p.x;
    accumulator_[1] += p.x * p.y;
    accumulator_[2] += p.x * p.z;
    accumulator_[3] += p.y * p.y;
    accumulator_[4] += p.y * p.z;
    accumulator_[5] += p.z * p.z;
    cumulative_centroid_ = cumulative_centroid_ + p;
    cumulative_weights_ += 1.0f;
    plane_up_to_date_ = false;
  }
  void AddPoint(const V3& p, float w) {  // .cc:67-77
    accumulator_[0] += p.x * p.x * w;
    accumulator_[1] += p.x * p.y * w;
    accumulator_[2] += p.x * p.z * w;
    accumulator_[3] += p.y * p.y * w;
    accumulator_[4] += p.y * p.z * w;
    accumulator_[5] += p.z * p.z * w;
    cumulative_centroid_ = cumulative_centroid_ + V3{p.x * w, p.y * w, p.z * w};
    cumulative_weights_ += w;
    plane_up_to_date_ = false;
  }
  const Plane& GetPlane() const {  // .cc:85-90
    if (!plane_up_to_date_) ComputePlane();
    return plane_;
  }
  bool PlaneValid() const {  // .cc:92-97
    if (!plane_up_to_date_) ComputePlane();
    return plane_valid_;
  }
  const V3& Centroid() const {  // .cc:99-104
    if (!plane_up_to_date_) ComputePlane();
    return centroid_;
  }
  float Curvature() const {  // .cc:106-111
    if (!plane_up_to_date_) ComputePlane();
    return curvature_;
  }
  void SetNormalOrientation(const V3& hint) {  // .cc:113-126
    if (PlaneValid()) {
      if (dot(plane_.n, hint) < 0.0f) {
        plane_.n = {plane_.n.x * -1.0f, plane_.n.y * -1.0f, plane_.n.z * -1.0f};
        plane_.d *= -1.0f;
      }
    } else {
      plane_.n = hint;
    }
  }
  void Merge(const PlaneEstimator& o) {  // .cc:128-133
    for (int i = 0; i < 6; ++i) accumulator_[i] += o.accumulator_[i];
    cumulative_centroid_ = cumulative_centroid_ + o.cumulative_centroid_;
    cumulative_weights_ += o.cumulative_weights_;
    plane_up_to_date_ = false;
  }
  // Transform(const Pose3d& new_pose_current), .cc:142-182.  pose = quaternion xyzw + translation in
  // double: rotationMatrix() is evaluated in double and cast to float (:159-160), the pose itself is
  // cast to float before it is applied to the centroid (:164).
  void Transform(const double q[4], const double t[3]) {
    if (!(cumulative_weights_ > 0.0f)) abort();  // CHECK_GT(cumulative_weights_, 0.0f)
    if (!plane_up_to_date_) centroid_ = cumulative_centroid_ / cumulative_weights_;
    float accu[6];
    for (int i = 0; i < 6; ++i) accu[i] = accumulator_[i] / cumulative_weights_;
    M3 cov;
    cov.m[0] = accu[0] - centroid_.x * centroid_.x;
    cov.m[1] = accu[1] - centroid_.x * centroid_.y;
    cov.m[2] = accu[2] - centroid_.x * centroid_.z;
    cov.m[4] = accu[3] - centroid_.y * centroid_.y;
    cov.m[5] = accu[4] - centroid_.y * centroid_.z;
    cov.m[8] = accu[5] - centroid_.z * centroid_.z;
    cov.m[3] = cov.m[1];
    cov.m[6] = cov.m[2];
    cov.m[7] = cov.m[5];
    // Quaterniond::toRotationMatrix (Geometry/Quaternion.h), then .cast<float>()
    const double tx = 2.0 * q[0], ty = 2.0 * q[1], tz = 2.0 * q[2];
    const double twx = tx * q[3], twy = ty * q[3], twz = tz * q[3];
    const double txx = tx * q[0], txy = ty * q[0], txz = tz * q[0];
    const double tyy = ty * q[1], tyz = tz * q[1], tzz = tz * q[2];
    M3 rot;
    rot(0, 0) = (float)(1.0 - (tyy + tzz)); rot(0, 1) = (float)(txy - twz); rot(0, 2) = (float)(txz + twy);
    rot(1, 0) = (float)(txy + twz); rot(1, 1) = (float)(1.0 - (txx + tzz)); rot(1, 2) = (float)(tyz - twx);
    rot(2, 0) = (float)(txz - twy); rot(2, 1) = (float)(tyz + twx); rot(2, 2) = (float)(1.0 - (txx + tyy));
    // rotation * covariance * rotation.transpose(): lazy coefficient products, each a 3-term
    // reduction e0 + (e1 + e2); the left product is evaluated into a temporary first
    M3 rc, out;
    for (int c = 0; c < 3; ++c)
      for (int r = 0; r < 3; ++r)
        rc(r, c) = rot(r, 0) * cov(0, c) + (rot(r, 1) * cov(1, c) + rot(r, 2) * cov(2, c));
    for (int c = 0; c < 3; ++c)
      for (int r = 0; r < 3; ++r)
        out(r, c) = rc(r, 0) * rot(c, 0) + (rc(r, 1) * rot(c, 1) + rc(r, 2) * rot(c, 2));
    const Quat qf{(float)q[0], (float)q[1], (float)q[2], (float)q[3]};
    const V3 tf{(float)t[0], (float)t[1], (float)t[2]};
    const V3 nc = rotate(qf, centroid_) + tf;
    accumulator_[0] = out.m[0] + nc.x * nc.x;
    accumulator_[1] = out.m[1] + nc.x * nc.y;
    accumulator_[2] = out.m[2] + nc.x * nc.z;
    accumulator_[3] = out.m[4] + nc.y * nc.y;
    accumulator_[4] = out.m[5] + nc.y * nc.z;
    accumulator_[5] = out.m[8] + nc.z * nc.z;
    for (float& a : accumulator_) a *= cumulative_weights_;
    cumulative_centroid_ = V3{nc.x * cumulative_weights_, nc.y * cumulative_weights_, nc.z * cumulative_weights_};
    plane_up_to_date_ = false;
  }
  void MergeWithPose(const PlaneEstimator& o, const double q[4], const double t[3]) {  // .cc:135-140
    PlaneEstimator other(o);
    other.Transform(q, t);
    Merge(other);
  }
  float weight() const { return cumulative_weights_; }
  void state(float out[10]) const {
    for (int i = 0; i < 6; ++i) out[i] = accumulator_[i];
    out[6] = cumulative_centroid_.x;
    out[7] = cumulative_centroid_.y;
    out[8] = cumulative_centroid_.z;
    out[9] = cumulative_weights_;
  }

 private:
  void ComputePlane() const {  // .cc:184-229
    if (!(cumulative_weights_ > 0.0f)) {
      // CHECK_GT(cumulative_weights_, 0.0f) -> process abort in the reference.
      abort();
    }
    plane_up_to_date_ = true;
    centroid_ = cumulative_centroid_ / cumulative_weights_;
    float accu[6];
    for (int i = 0; i < 6; ++i) accu[i] = accumulator_[i] / cumulative_weights_;
    M3 cov;
    cov.m[0] = accu[0] - centroid_.x * centroid_.x;
    cov.m[1] = accu[1] - centroid_.x * centroid_.y;
    cov.m[2] = accu[2] - centroid_.x * centroid_.z;
    cov.m[4] = accu[3] - centroid_.y * centroid_.y;
    cov.m[5] = accu[4] - centroid_.y * centroid_.z;
    cov.m[8] = accu[5] - centroid_.z * centroid_.z;
    cov.m[3] = cov.m[1];
    cov.m[6] = cov.m[2];
    cov.m[7] = cov.m[5];
    float evals[3];
    M3 evecs;
    eig3_direct(cov, evals, &evecs);
    if (evals[1] > FLT_MIN) {
      float smallest = evals[0];
      V3 ev = evecs.col(0);
      if (dot(plane_.n, ev) > 0.0f)
        plane_ = Plane(ev, centroid_);
      else
        plane_ = Plane(neg(ev), centroid_);
      float trace = cov.m[0] + (cov.m[4] + cov.m[8]);
      if (trace > smallest && smallest > FLT_MIN)
        curvature_ = fabsf(smallest / trace);
      else
        curvature_ = 0.0f;
      plane_valid_ = true;
    } else {
      curvature_ = 0.0f;
      plane_ = Plane(plane_.n, centroid_);
      plane_valid_ = false;
    }
  }
  mutable Plane plane_;
  mutable V3 centroid_{0, 0, 0};
  mutable float curvature_ = 0.0f;
  mutable bool plane_up_to_date_;
  mutable bool plane_valid_;
  V3 cumulative_centroid_;
  float accumulator_[6];
  float cumulative_weights_;
};

// ---------------------------------------------------------------------------
// ps/algorithms.h:106-257
// ---------------------------------------------------------------------------
int FindNormalSupportNeighbors(const Cloud3& pc, int col, int row, float min_d,
                               float max_d, bool diag, std::vector<int>* out) {
  const float kMin2 = min_d * min_d;
  const float kMax2 = max_d * max_d;
  out->clear();
  const V3 center = pc.at(row, col);
  if (!all_finite(center)) return 0;
  int min_row = std::max(row - 1, 0);
  int max_row = std::min(row + 1, pc.rows - 1);
  int min_col = std::max(col - 1, 0);
  int max_col = std::min(col + 1, pc.cols - 1);
  out->push_back(pc.lin(row, col));
  // upper neighbor
  int index = pc.lin(row - 1, col);
  for (int r = row - 1; r >= 0; --r, index -= 1) {
    const V3 p = pc.at(index);
    if (!all_finite(p)) continue;
    const float d2 = sqnorm(center - p);
    if (d2 >= kMin2 && d2 <= kMax2) {
      out->push_back(index);
      min_row = r;
      break;
    }
  }
  // left neighbor
  index = pc.lin(row, col - 1);
  for (int c = col - 1; c >= 0; --c, index -= pc.rows) {
    const V3 p = pc.at(index);
    if (!all_finite(p)) continue;
    const float d2 = sqnorm(center - p);
    if (d2 >= kMin2 && d2 <= kMax2) {
      out->push_back(index);
      min_col = c;
      break;
    }
  }
  // lower neighbor
  index = pc.lin(row + 1, col);
  for (int r = row + 1; r < pc.rows; ++r, index += 1) {
    const V3 p = pc.at(index);
    if (!all_finite(p)) continue;
    const float d2 = sqnorm(center - p);
    if (d2 >= kMin2 && d2 <= kMax2) {
      out->push_back(index);
      max_row = r;
      break;
    }
  }
  // right neighbor
  index = pc.lin(row, col + 1);
  for (int c = col + 1; c < pc.cols; ++c, index += pc.rows) {
    const V3 p = pc.at(index);
    if (!all_finite(p)) continue;
    const float d2 = sqnorm(center - p);
    if (d2 >= kMin2 && d2 <= kMax2) {
      out->push_back(index);
      max_col = c;
      break;
    }
  }
  if (!diag) return static_cast<int>(out->size());
  auto try_diag = [&](int r, int c) {
    int idx = pc.lin(r, c);
    const V3 p = pc.at(idx);
    const float d2 = sqnorm(center - p);
    if (all_finite(p) && d2 >= kMin2 && d2 <= kMax2) out->push_back(idx);
  };
  if (min_col != col) {
    if (min_row != row) try_diag(min_row, min_col);
    if (max_row != row) try_diag(max_row, min_col);
  }
  if (max_col != col) {
    if (min_row != row) try_diag(min_row, max_col);
    if (max_row != row) try_diag(max_row, max_col);
  }
  return static_cast<int>(out->size());
}

// ---------------------------------------------------------------------------
// ps/algorithms.h:264-292
// ---------------------------------------------------------------------------
V3 CumulativePolygonNormal(const Cloud3& pc, const int* indices, int size) {
  V3 normal{0.0f, 0.0f, 0.0f};
  for (int i = 0; i < size; ++i) {
    int a = indices ? indices[i] : i;
    int b = indices ? indices[(i + 1) % size] : (i + 1) % size;
    normal = normal + cross(pc.at(a), pc.at(b));
  }
  return normal;
}
float PolygonArea(const Cloud3& pc, const int* indices, int size) {
  return 0.5f * sqrtf(sqnorm(CumulativePolygonNormal(pc, indices, size)));
}

// ---------------------------------------------------------------------------
// eigenmath::PoseFromPlane (x_edr_eigenmath 1.0.0, not in the tree; restated:
// z axis = plane normal via Quaternion::FromTwoVectors(UnitZ, normal),
// origin = -offset * normal) and collision::ConvexHull (x_edr_mobility_collision
// 1.0.0, not in the tree; restated as Andrew's monotone chain in double).
// ps/algorithms.h:526-549 with patch P1 (output buffer sized to the hull).
// ---------------------------------------------------------------------------
Quat FromTwoVectorsUnitZ(const V3& b) {
  V3 v0{0.0f, 0.0f, 1.0f};
  V3 v1 = normalized(b);
  float c = dot(v1, v0);
  if (c < -1.0f + FLT_EPSILON) {
    return Quat{1.0f, 0.0f, 0.0f, 0.0f};  // rotation by pi about x
  }
  V3 axis = cross(v0, v1);
  float s = sqrtf((1.0f + c) * 2.0f);
  float invs = 1.0f / s;
  return Quat{axis.x * invs, axis.y * invs, axis.z * invs, s * 0.5f};
}

struct P2d {
  double x, y;
};
inline double cross2(const P2d& o, const P2d& a, const P2d& b) {
  return (a.x - o.x) * (b.y - o.y) - (a.y - o.y) * (b.x - o.x);
}
std::vector<P2d> MonotoneChain(std::vector<P2d> pts) {
  std::sort(pts.begin(), pts.end(), [](const P2d& a, const P2d& b) {
    return a.x < b.x || (a.x == b.x && a.y < b.y);
  });
  pts.erase(std::unique(pts.begin(), pts.end(),
                        [](const P2d& a, const P2d& b) {
                          return a.x == b.x && a.y == b.y;
                        }),
            pts.end());
  int n = static_cast<int>(pts.size());
  if (n < 3) return pts;
  std::vector<P2d> h(2 * n);
  int k = 0;
  for (int i = 0; i < n; ++i) {
    while (k >= 2 && cross2(h[k - 2], h[k - 1], pts[i]) <= 0) --k;
    h[k++] = pts[i];
  }
  for (int i = n - 2, t = k + 1; i >= 0; --i) {
    while (k >= t && cross2(h[k - 2], h[k - 1], pts[i]) <= 0) --k;
    h[k++] = pts[i];
  }
  h.resize(k - 1);
  return h;
}

void PlanarConvexHull(const Cloud3& pc, const std::vector<int>& indices,
                      const Plane& plane, std::vector<V3>* hull) {
  Quat q = FromTwoVectorsUnitZ(plane.n);
  V3 t = (-plane.d) * plane.n;
  Quat qi{-q.x, -q.y, -q.z, q.w};
  V3 ti = neg(rotate(qi, t));
  std::vector<P2d> plane_points(indices.size());
  for (size_t i = 0; i < indices.size(); ++i) {
    V3 pp = rotate(qi, pc.at(indices[i])) + ti;
    plane_points[i] = {pp.x, pp.y};
  }
  std::vector<P2d> ch = MonotoneChain(plane_points);
  hull->resize(ch.size());  // patch P1
  for (size_t i = 0; i < ch.size(); ++i) {
    V3 pp{static_cast<float>(ch[i].x), static_cast<float>(ch[i].y), 0.0f};
    (*hull)[i] = rotate(q, pp) + t;
  }
}

// ---------------------------------------------------------------------------
// Result container shared by planar / cluster / mean-shift segmentations.
// ---------------------------------------------------------------------------
}  // namespace

struct oracle_regions {
  std::vector<int> inlier_offsets{0};
  std::vector<int> inlier_indices;
  std::vector<int> label_ids;
  std::vector<int> seed_indices;
  std::vector<float> planes;
  std::vector<float> centroids;
  std::vector<float> areas;
  std::vector<float> estimators;
  std::vector<int> boundary_offsets{0};
  std::vector<int> boundary_indices;
  std::vector<int> disc_offsets{0};
  std::vector<int> disc_indices;
  std::vector<int> hull_offsets{0};
  std::vector<float> hull_points;
  std::vector<float> seed_points;
  std::vector<float> bboxes;
  int count() const { return static_cast<int>(label_ids.size()); }
};

namespace {

// ---------------------------------------------------------------------------
// PlanarRegion, ps/planar_region.h:87-465, ps/planar_region.cc
// ---------------------------------------------------------------------------
struct Neighbor {
  int delta_x, delta_y, delta_index;
};
std::vector<Neighbor> CreateNeighborhood(bool use8, int rows) {  // .cc:26-45
  if (use8)
    return {{-1, 0, -rows}, {-1, -1, -rows - 1}, {0, -1, -1},
            {1, -1, rows - 1}, {1, 0, rows},     {1, 1, rows + 1},
            {0, 1, 1},         {-1, 1, -rows + 1}};
  return {{-1, 0, -rows}, {0, -1, -1}, {1, 0, rows}, {0, 1, 1}};
}

struct PlanarInput {
  Quat robot_q;
  Cloud3 points;
  Cloud3 normals;
};

class PlanarRegion {
 public:
  explicit PlanarRegion(const oracle_planar_config& cfg) : config_(cfg) {}

  void Init(const PlanarInput& in, int seed, std::queue<int>* q) {  // .h:144-159
    const V3 seed_normal = in.normals.at(seed);
    const V3 seed_point = in.points.at(seed);
    seed_point_index_ = seed;
    plane_ = Plane(seed_normal, seed_point);
    plane_estimator_.Clear();
    plane_estimator_.SetNormalOrientation(seed_normal);
    centroid_ = seed_point;
    area_ = 0.0f;
    inlier_indices_.clear();
    projected_boundary_points_.clear();
    q->push(seed);
  }

  bool CheckInlierAndUpdate(const PlanarInput& in, int cand, int* labels,
                            int current_label_id) {  // .h:163-184
    const V3 point = in.points.at(cand);
    if (plane_.absDistance(point) < config_.max_plane_distance) {
      inlier_indices_.push_back(cand);
      plane_estimator_.AddPoint(point);
      if (inlier_indices_.size() %
              static_cast<size_t>(config_.plane_model_reestimation_period) ==
          0) {
        plane_ = plane_estimator_.GetPlane();
      }
      labels[cand] = current_label_id;
      label_id_ = current_label_id;
      return true;
    }
    labels[cand] = kUnlabeled;
    return false;
  }

  void AddNeighbors(const PlanarInput& in, int point_index, int* labels,
                    std::queue<int>* q) {  // .h:229-255
    const int rows = in.points.rows, cols = in.points.cols;
    int row = point_index % rows, col = point_index / rows;
    for (int dc = -1; dc <= 1; ++dc) {
      for (int dr = -1; dr <= 1; ++dr) {
        if (std::abs(dc + dr) != 1) continue;
        const int cc = col + dc, rc = row + dr;
        if (cc < 0 || cc >= cols || rc < 0 || rc >= rows ||
            labels[cc * rows + rc] != kUnlabeled)
          continue;
        labels[cc * rows + rc] = kAlreadyInUnexaminedPointsQueue;
        q->push(cc * rows + rc);
      }
    }
  }

  bool CheckAndFinalize(const PlanarInput& in, const int* labels) {  // .h:189-224
    if (inlier_indices_.size() <
        static_cast<size_t>(config_.min_region_inliers))
      return false;
    plane_ = plane_estimator_.GetPlane();
    centroid_ = plane_estimator_.Centroid();
    for (int i = static_cast<int>(inlier_indices_.size()) - 1; i >= 0; --i) {
      if (FindLabeledRegionBoundary(inlier_indices_[i], in, labels)) break;
    }
    // detail::CheckMinRowsAndCols, .cc:91-106
    {
      const int rows = in.points.rows, cols = in.points.cols;
      int min_x = cols, max_x = 0, min_y = rows, max_y = 0;
      for (int index : boundary_indices_) {
        int x = index / rows, y = index % rows;
        min_x = std::min(x, min_x);
        max_x = std::max(x, max_x);
        min_y = std::min(y, min_y);
        max_y = std::max(y, max_y);
      }
      if (!(((max_x - min_x) > 3) && ((max_y - min_y) > 3))) return false;
    }
    if (boundary_indices_.empty()) return false;
    PlanarConvexHull(in.points, boundary_indices_, plane_,
                     &projected_boundary_points_);
    if (projected_boundary_points_.size() < 3) return false;
    area_ = PolygonArea(in.points, boundary_indices_.data(),
                        static_cast<int>(boundary_indices_.size()));
    return area_ >= config_.min_region_area;
  }

  bool FindLabeledRegionBoundary(int start_idx, const PlanarInput& in,
                                 const int* labels) {  // .h:295-425
    const int rows = in.points.rows, cols = in.points.cols;
    const std::vector<Neighbor> directions = CreateNeighborhood(true, rows);
    boundary_indices_.clear();
    discontinuous_boundary_indices_.clear();
    int curr_idx = start_idx;
    int curr_x = start_idx / rows;
    int curr_y = start_idx % rows;
    const int region_label = labels[start_idx];
    // detail::FindInitialPredecessorDirection, .cc:67-82
    int direction_idx = -1;
    for (int i = 0; i < 8; ++i) {
      const int x = curr_x + directions[i].delta_x;
      const int y = curr_y + directions[i].delta_y;
      const int index = curr_idx + directions[i].delta_index;
      if (x >= 0 && x < cols && y >= 0 && y < rows &&
          labels[index] != region_label) {
        direction_idx = i;
        break;
      }
    }
    if (direction_idx < 0) return false;
    const float min_range_sq =
        config_.discontinuity_min_range * config_.discontinuity_min_range;
    const float max_range_sq =
        config_.discontinuity_max_range * config_.discontinuity_max_range;
    bool found_start_idx = false;
    while (true) {
      // detail::FindNextBoundaryPointDirection, .cc:47-65
      int new_direction_index = -1;
      for (int delta_idx = 1; delta_idx <= 8; ++delta_idx) {
        const int nd = (direction_idx + delta_idx) % 8;
        const int x = curr_x + directions[nd].delta_x;
        const int y = curr_y + directions[nd].delta_y;
        const int index = curr_idx + directions[nd].delta_index;
        if (x >= 0 && x < cols && y >= 0 && y < rows &&
            labels[index] == region_label) {
          new_direction_index = nd;
          break;
        }
      }
      if (new_direction_index < 0) abort();  // CHECK_GE
      direction_idx = (new_direction_index + 4) % 8;
      curr_idx += directions[new_direction_index].delta_index;
      curr_x += directions[new_direction_index].delta_x;
      curr_y += directions[new_direction_index].delta_y;
      if (found_start_idx) {
        if (curr_idx == boundary_indices_.front()) break;
        found_start_idx = false;
      }
      boundary_indices_.push_back(curr_idx);
      const V3 pc = in.points.at(curr_idx);
      if (sqnorm(pc) > min_range_sq && sqnorm(pc) < max_range_sq) {
        const std::vector<Neighbor> dd = CreateNeighborhood(false, rows);
        auto in_bounds = [&](int idx, const Neighbor& n) {  // .cc:84-89
          int x = (idx / rows) + n.delta_x;
          int y = (idx % rows) + n.delta_y;
          return x >= 0 && x < cols && y >= 0 && y < rows;
        };
        int cnt = 0;
        for (const Neighbor& n : dd) cnt += in_bounds(curr_idx, n) ? 1 : 0;
        if (cnt == 4) {
          for (const Neighbor& n : dd) {
            const V3 pn = in.points.at(curr_idx + n.delta_index);
            if (std::isnan(pc.z) || std::isnan(pn.z)) continue;
            if (!in_bounds(curr_idx, n) ||
                region_label != labels[curr_idx + n.delta_index])
              continue;
            const V3 pt_delta = pc - pn;
            const V3 pt_delta_robot = rotate(in.robot_q, pt_delta);
            const float cos_angle = dot(in.normals.at(curr_idx),
                                        in.normals.at(curr_idx + n.delta_index));
            // eigenmath::Degrees (not in tree): radians * 180 / pi, restated in
            // double then rounded to float.
            const float deg = static_cast<float>(
                static_cast<double>(o_acosf(cos_angle)) * (180.0 / M_PI));
            if (fabsf(deg) < config_.discontinuity_normal_angle_diff &&
                fabsf(pt_delta_robot.z) < config_.discontinuity_z_diff)
              continue;
            if (fabsf(pt_delta_robot.z) / sqrtf(sqnorm(pt_delta_robot)) <
                config_.discontinuity_z_ratio)
              continue;
            discontinuous_boundary_indices_.insert(curr_idx);
          }
        }
      }
      if (curr_idx == start_idx) found_start_idx = true;
    }
    return true;
  }

  void Export(oracle_regions* out) const {
    out->inlier_indices.insert(out->inlier_indices.end(), inlier_indices_.begin(),
                               inlier_indices_.end());
    out->inlier_offsets.push_back(static_cast<int>(out->inlier_indices.size()));
    out->label_ids.push_back(label_id_);
    out->seed_indices.push_back(seed_point_index_);
    out->planes.insert(out->planes.end(),
                       {plane_.n.x, plane_.n.y, plane_.n.z, plane_.d});
    out->centroids.insert(out->centroids.end(),
                          {centroid_.x, centroid_.y, centroid_.z});
    out->areas.push_back(area_);
    float st[10];
    plane_estimator_.state(st);
    out->estimators.insert(out->estimators.end(), st, st + 10);
    out->boundary_indices.insert(out->boundary_indices.end(),
                                 boundary_indices_.begin(),
                                 boundary_indices_.end());
    out->boundary_offsets.push_back(
        static_cast<int>(out->boundary_indices.size()));
    out->disc_indices.insert(out->disc_indices.end(),
                             discontinuous_boundary_indices_.begin(),
                             discontinuous_boundary_indices_.end());
    out->disc_offsets.push_back(static_cast<int>(out->disc_indices.size()));
    for (const V3& h : projected_boundary_points_)
      out->hull_points.insert(out->hull_points.end(), {h.x, h.y, h.z});
    out->hull_offsets.push_back(static_cast<int>(out->hull_points.size() / 3));
  }

  const std::vector<int>& inlier_indices() const { return inlier_indices_; }

 private:
  int label_id_ = 0;
  oracle_planar_config config_;
  PlaneEstimator plane_estimator_;
  V3 centroid_{0, 0, 0};
  Plane plane_;
  int seed_point_index_ = 0;
  float area_ = 0.0f;
  std::vector<int> inlier_indices_;
  std::vector<int> boundary_indices_;
  std::set<int> discontinuous_boundary_indices_;  // flat_hash_set: compare as set
  std::vector<V3> projected_boundary_points_;
};

// ---------------------------------------------------------------------------
// ClusterRegion, ps/cluster_region.h:47-176
// ---------------------------------------------------------------------------
struct Pair {
  int source_index, candidate_index;
};

class ClusterRegion {
 public:
  explicit ClusterRegion(const oracle_cluster_config& cfg) : config_(cfg) {}
  void Init(const Cloud3&, int seed, std::queue<Pair>* q) {  // :73-81
    inlier_indices_.clear();
    q->push(Pair{seed, seed});
  }
  bool CheckInlierAndUpdate(const Cloud3& in, const Pair pr, int* labels,
                            int current_label_id) {  // :85-107
    const V3 cand = in.at(pr.candidate_index);
    const V3 src = in.at(pr.source_index);
    if (sqnorm(cand - src) < config_.squared_distance_threshold) {
      inlier_indices_.push_back(pr.candidate_index);
      labels[pr.candidate_index] = current_label_id;
      label_id_ = current_label_id;
      return true;
    }
    labels[pr.candidate_index] = kUnlabeled;
    return false;
  }
  bool CheckAndFinalize() const {  // :111-116
    return !(inlier_indices_.size() <
             static_cast<size_t>(config_.min_region_inliers));
  }
  void AddNeighbors(const Cloud3& points, const Pair& pr, int* labels,
                    std::queue<Pair>* q) {  // :122-150
    const int point_index = pr.candidate_index;
    const int rows = points.rows, cols = points.cols;
    int row = point_index % rows, col = point_index / rows;
    const int hw = config_.half_search_window;
    for (int dc = -hw; dc <= hw; ++dc) {
      for (int dr = -hw; dr <= hw; ++dr) {
        if (dc == 0 && dr == 0) continue;
        const int cc = col + dc, rc = row + dr;
        if (cc < 0 || cc >= cols || rc < 0 || rc >= rows ||
            labels[cc * rows + rc] != kUnlabeled)
          continue;
        labels[cc * rows + rc] = kAlreadyInUnexaminedPointsQueue;
        q->push(Pair{point_index, cc * rows + rc});
      }
    }
  }
  void Export(oracle_regions* out) const {
    out->inlier_indices.insert(out->inlier_indices.end(), inlier_indices_.begin(),
                               inlier_indices_.end());
    out->inlier_offsets.push_back(static_cast<int>(out->inlier_indices.size()));
    out->label_ids.push_back(label_id_);
    if (has_seed_)
      out->seed_points.insert(out->seed_points.end(), {seed_.x, seed_.y, seed_.z});
  }
  const std::vector<int>& inlier_indices() const { return inlier_indices_; }
  std::vector<int>* mutable_inlier_indices() { return &inlier_indices_; }
  int label_id() const { return label_id_; }
  int* mutable_label_id() { return &label_id_; }
  const oracle_cluster_config& config() const { return config_; }
  void SetSeed(const V3& s) {
    seed_ = s;
    has_seed_ = true;
  }
  const V3& Seed() const { return seed_; }

 private:
  int label_id_ = 0;
  oracle_cluster_config config_;
  std::vector<int> inlier_indices_;
  V3 seed_{0, 0, 0};
  bool has_seed_ = false;
};

// ---------------------------------------------------------------------------
// SegmentRegions driver, ps/segmentation.h:239-292
// ---------------------------------------------------------------------------
// diagnostics: growth attempts of the last SegmentRegions calls by inlier-count bucket
// (0, 1, 2-7, 8-31, 32-255, 256-4095, >= 4096) and queue pops per bucket
static long long g_attempt_hist[16];
template <typename RegionT, typename InputT, typename NeighborT, typename CfgT>
void SegmentRegions(const InputT& input, const int* seeds, int n_seeds,
                    const CfgT& cfg, oracle_regions* regions, int* labels,
                    int n_labels, int initial_id_offset) {
  std::queue<NeighborT> unexamined;
  RegionT region(cfg);
  int num_seed_points = n_seeds;
  while (num_seed_points > 0) {
    const int seed = seeds[num_seed_points - 1];
    --num_seed_points;
    if (labels[seed] != kUnlabeled) continue;
    region.Init(input, seed, &unexamined);
    long long pops = 0;
    while (!unexamined.empty()) {
      const NeighborT cand = unexamined.front();
      unexamined.pop();
      ++pops;
      if (region.CheckInlierAndUpdate(input, cand, labels,
                                      regions->count() + initial_id_offset)) {
        region.AddNeighbors(input, cand, labels, &unexamined);
      }
    }
    {
      const size_t ni = region.inlier_indices().size();
      const int b = ni == 0 ? 0 : ni == 1 ? 1 : ni < 8 ? 2 : ni < 32 ? 3 : ni < 256 ? 4 : ni < 4096 ? 5 : 6;
      ++g_attempt_hist[b];
      g_attempt_hist[8 + b] += pops;
    }
    bool ok;
    if constexpr (std::is_same<RegionT, PlanarRegion>::value)
      ok = region.CheckAndFinalize(input, labels);
    else
      ok = region.CheckAndFinalize();
    if (ok) {
      region.Export(regions);
    } else {
      for (const int i : region.inlier_indices()) labels[i] = kAlreadyExamedPoint;
    }
  }
  for (int i = 0; i < n_labels; ++i)
    if (labels[i] == kAlreadyExamedPoint) labels[i] = kUnlabeled;
}

// ---------------------------------------------------------------------------
// Mean shift, ps/mean_shift_segmentation.h, .cc
// ---------------------------------------------------------------------------
struct PointIndex {  // .h:54-77, .cc:22-34 (Vector2f: x = row, y = col)
  float row, col;
  int RowIndex() const { return static_cast<int>(std::round(row)); }
  int ColIndex() const { return static_cast<int>(std::round(col)); }
};
struct SeedPointAndIndex {
  V3 seed_point;
  PointIndex index;
};

}  // namespace

// ===========================================================================
// C interface
// ===========================================================================
extern "C" {

void oracle_set_native_libm(int on) { g_native_libm = on != 0; }

int oracle_normals_organized(const float* xyz, int rows, int cols,
                             const double sensor_t[3],
                             const oracle_normals_params* params, int row_begin,
                             int row_end, int col_begin, int col_end,
                             float* normals, float* curvature) {
  // ps/algorithms.h:329-366
  const float kInvalid = std::numeric_limits<float>::quiet_NaN();
  Cloud3 points{xyz, rows, cols};
  PlaneEstimator est;
  std::vector<int> nbrs;
  const V3 sensor{static_cast<float>(sensor_t[0]), static_cast<float>(sensor_t[1]),
                  static_cast<float>(sensor_t[2])};
  for (int col = col_begin; col < col_end; ++col) {
    for (int row = row_begin; row < row_end; ++row) {
      const V3 point = points.at(row, col);
      float* normal = normals + 3 * (col * rows + row);
      float* curv = curvature ? curvature + (col * rows + row) : nullptr;
      if (!all_finite(point) ||
          FindNormalSupportNeighbors(points, col, row,
                                     params->min_neighbor_distance,
                                     params->max_neighbor_distance,
                                     params->include_diagonal_neighbors != 0,
                                     &nbrs) < params->min_num_support_neighbors) {
        normal[0] = normal[1] = normal[2] = kInvalid;
        if (curv) *curv = kInvalid;
        continue;
      }
      est.SetNormalOrientation(sensor - point);
      for (int idx : nbrs) est.AddPoint(points.at(idx));
      if (!est.PlaneValid()) {
        normal[0] = normal[1] = normal[2] = kInvalid;
        if (curv) *curv = kInvalid;
        continue;  // quirk Q3: no Clear()
      }
      const Plane& pl = est.GetPlane();
      normal[0] = pl.n.x;
      normal[1] = pl.n.y;
      normal[2] = pl.n.z;
      if (curv) *curv = est.Curvature();
      est.Clear();
    }
  }
  return 0;
}

int oracle_find_normal_support_neighbors(const float* xyz, int rows, int cols,
                                         int col, int row, float min_d,
                                         float max_d, int diag, int* out) {
  Cloud3 points{xyz, rows, cols};
  std::vector<int> v;
  int n = FindNormalSupportNeighbors(points, col, row, min_d, max_d, diag != 0, &v);
  for (size_t i = 0; i < v.size(); ++i) out[i] = v[i];
  return n;
}

int oracle_plane_estimator_replay(const int* ops, const float* args, int n_ops,
                                  float* out, int out_cap) {
  return oracle_plane_estimator_replay2(ops, args, n_ops, nullptr, 0, out, out_cap);
}

int oracle_plane_estimator_replay2(const int* ops, const float* args, int n_ops,
                                   const double* poses, int n_poses,
                                   float* out, int out_cap) {
  PlaneEstimator slots[4];
  int cur = 0;
  int nq = 0;
  for (int i = 0; i < n_ops; ++i) {
    PlaneEstimator& est = slots[cur];
    V3 p{args[4 * i], args[4 * i + 1], args[4 * i + 2]};
    float w = args[4 * i + 3];
    switch (ops[i]) {
      case 0: est.AddPoint(p); break;
      case 1: est.AddPoint(p, w); break;
      case 2: est.SetNormalOrientation(p); break;
      case 3: {
        if ((nq + 1) * 10 > out_cap) return -1;
        float* o = out + 10 * nq++;
        if (!(est.weight() > 0.0f)) {
          // PlaneValid() right after Clear(): up to date & invalid, .cc:46-53
          o[0] = est.PlaneValid() ? 1.0f : 0.0f;
          o[1] = 0.0f;
          o[2] = o[3] = o[4] = 0.0f;
          const Plane& pl = est.GetPlane();
          o[5] = pl.n.x; o[6] = pl.n.y; o[7] = pl.n.z; o[8] = pl.d;
          o[9] = est.weight();
          break;
        }
        o[0] = est.PlaneValid() ? 1.0f : 0.0f;
        o[1] = est.Curvature();
        const V3& c = est.Centroid();
        o[2] = c.x; o[3] = c.y; o[4] = c.z;
        const Plane& pl = est.GetPlane();
        o[5] = pl.n.x; o[6] = pl.n.y; o[7] = pl.n.z; o[8] = pl.d;
        o[9] = est.weight();
        break;
      }
      case 4: est.Clear(); break;
      case 5: {  // select the active estimator
        const int k = (int)p.x;
        if (k < 0 || k >= 4) return -2;
        cur = k;
        break;
      }
      case 6: {  // Merge(other)
        const int k = (int)p.x;
        if (k < 0 || k >= 4 || k == cur) return -2;
        est.Merge(slots[k]);
        break;
      }
      case 7: {  // Transform(pose)
        const int pi = (int)p.x;
        if (pi < 0 || pi >= n_poses) return -2;
        if (!(est.weight() > 0.0f)) return -3;  // CHECK_GT
        est.Transform(poses + 7 * pi, poses + 7 * pi + 4);
        break;
      }
      case 8: {  // Merge(other, this_pose_other)
        const int k = (int)p.x, pi = (int)p.y;
        if (k < 0 || k >= 4 || k == cur || pi < 0 || pi >= n_poses) return -2;
        if (!(slots[k].weight() > 0.0f)) return -3;
        est.MergeWithPose(slots[k], poses + 7 * pi, poses + 7 * pi + 4);
        break;
      }
      default: return -2;
    }
  }
  return nq;
}

void oracle_eig3(const float* a, float* evals, float* evecs) {
  M3 m;
  m(0, 0) = a[0]; m(1, 0) = a[1]; m(2, 0) = a[2];
  m(1, 1) = a[3]; m(2, 1) = a[4]; m(2, 2) = a[5];
  m(0, 1) = a[1]; m(0, 2) = a[2]; m(1, 2) = a[4];
  M3 ev;
  for (float& v : ev.m) v = 0.0f;
  eig3_direct(m, evals, &ev);
  for (int i = 0; i < 9; ++i) evecs[i] = ev.m[i];
}

float oracle_polygon_area(const float* xyz, int n_points, const int* indices,
                          int n_indices, float* cumulative_normal) {
  Cloud3 pc{xyz, 1, n_points};
  int size = indices ? n_indices : n_points;
  V3 n = CumulativePolygonNormal(pc, indices, size);
  if (cumulative_normal) {
    cumulative_normal[0] = n.x;
    cumulative_normal[1] = n.y;
    cumulative_normal[2] = n.z;
  }
  return 0.5f * sqrtf(sqnorm(n));
}

void oracle_prepare_unlabeled_labels(int* labels, int n) {
  for (int i = 0; i < n; ++i) labels[i] = kUnlabeled;
}

void oracle_debug_attempt_hist(long long out[16], int reset) {
  for (int i = 0; i < 16; ++i) out[i] = g_attempt_hist[i];
  if (reset) for (int i = 0; i < 16; ++i) g_attempt_hist[i] = 0;
}

int oracle_segment_planar(const float* xyz, const float* normals, int rows,
                          int cols, const float robot_q_xyzw[4],
                          const int* seeds, int n_seeds,
                          const oracle_planar_config* cfg, int id_offset,
                          int* labels, oracle_regions** out) {
  if (rows <= 1) return 1;  // CHECK_GT(points.Rows(), 1), ps/planar_region.h:264
  if (cfg->plane_model_reestimation_period <= 0 || cfg->min_region_inliers < 0)
    return 2;  // quirk Q8: UB / size_t wrap in the reference
  PlanarInput in{Quat{robot_q_xyzw[0], robot_q_xyzw[1], robot_q_xyzw[2],
                      robot_q_xyzw[3]},
                 Cloud3{xyz, rows, cols}, Cloud3{normals, rows, cols}};
  auto* res = new oracle_regions();
  SegmentRegions<PlanarRegion, PlanarInput, int, oracle_planar_config>(
      in, seeds, n_seeds, *cfg, res, labels, rows * cols, id_offset);
  *out = res;
  return 0;
}

int oracle_segment_cluster(const float* xyz, int rows, int cols,
                           const int* seeds, int n_seeds,
                           const oracle_cluster_config* cfg, int id_offset,
                           int* labels, oracle_regions** out) {
  if (rows <= 1) return 1;  // CHECK_GT(points.Rows(), 1), ps/cluster_region.h:154
  if (cfg->min_region_inliers < 0 || cfg->half_search_window < 0) return 2;
  Cloud3 in{xyz, rows, cols};
  auto* res = new oracle_regions();
  SegmentRegions<ClusterRegion, Cloud3, Pair, oracle_cluster_config>(
      in, seeds, n_seeds, *cfg, res, labels, rows * cols, id_offset);
  *out = res;
  return 0;
}

// Fixed-radius Euclidean clustering of an UNORGANIZED cloud (extension; semantics defined in
// x-edr-pointcloud-segmentation_b200/csrc/voxel_cluster.cuh).  Plain
// restatement: voxel buckets in a std::unordered_map, sequential union-find, regions numbered by
// smallest member index, inlier lists ascending, centroid summed in the documented lane order.
int oracle_cluster_unorganized(const float* xyz, int n, float radius,
                               int min_region_inliers, int id_offset,
                               int* labels, oracle_regions** out) {
  if (n <= 0 || !(radius > 0.0f) || min_region_inliers < 1) return 2;
  const float r2 = radius * radius;
  auto coord = [&](float v) {
    float q = std::floor(v / radius);
    q = std::fmin(std::fmax(q, -1048576.0f), 1048575.0f);
    return static_cast<long long>(q);
  };
  auto key = [](long long x, long long y, long long z) {
    return ((x + 1048576) << 42) | ((y + 1048576) << 21) | (z + 1048576);
  };
  std::vector<char> part(n, 0);
  std::unordered_map<long long, std::vector<int>> grid;
  for (int i = 0; i < n; ++i) {
    const float* p = xyz + 3 * static_cast<size_t>(i);
    if (labels[i] != kUnlabeled || !std::isfinite(p[0]) || !std::isfinite(p[1]) || !std::isfinite(p[2])) continue;
    part[i] = 1;
    grid[key(coord(p[0]), coord(p[1]), coord(p[2]))].push_back(i);
  }
  std::vector<int> parent(n);
  for (int i = 0; i < n; ++i) parent[i] = i;
  auto find = [&](int x) {
    while (parent[x] != x) { parent[x] = parent[parent[x]]; x = parent[x]; }
    return x;
  };
  for (int i = 0; i < n; ++i) {
    if (!part[i]) continue;
    const float* p = xyz + 3 * static_cast<size_t>(i);
    const long long vx = coord(p[0]), vy = coord(p[1]), vz = coord(p[2]);
    for (long long dz = -1; dz <= 1; ++dz)
      for (long long dy = -1; dy <= 1; ++dy)
        for (long long dx = -1; dx <= 1; ++dx) {
          const long long nx = vx + dx, ny = vy + dy, nz = vz + dz;
          if (nx < -1048576 || nx > 1048575 || ny < -1048576 || ny > 1048575 || nz < -1048576 || nz > 1048575) continue;
          auto it = grid.find(key(nx, ny, nz));
          if (it == grid.end()) continue;
          for (int j : it->second) {
            if (j >= i) continue;
            const float* q = xyz + 3 * static_cast<size_t>(j);
            const float d0 = q[0] - p[0], d1 = q[1] - p[1], d2 = q[2] - p[2];
            if (d0 * d0 + (d1 * d1 + d2 * d2) < r2) {   // ps/cluster_region.h:94 order
              int a = find(i), b = find(j);
              if (a != b) { if (a < b) std::swap(a, b); parent[a] = b; }   // root = smallest member
            }
          }
        }
  }
  std::vector<int> count(n, 0), root(n, -1);
  for (int i = 0; i < n; ++i)
    if (part[i]) { root[i] = find(i); ++count[root[i]]; }
  auto* res = new oracle_regions();
  std::vector<int> ordinal(n, -1);
  int R = 0;
  for (int i = 0; i < n; ++i)
    if (part[i] && root[i] == i && count[i] >= min_region_inliers) ordinal[i] = R++;
  std::vector<std::vector<int>> lists(R);
  for (int i = 0; i < n; ++i)
    if (part[i] && ordinal[root[i]] >= 0) {
      labels[i] = id_offset + ordinal[root[i]];
      lists[ordinal[root[i]]].push_back(i);
    }
  for (int r = 0; r < R; ++r) {
    const std::vector<int>& l = lists[r];
    res->inlier_indices.insert(res->inlier_indices.end(), l.begin(), l.end());
    res->inlier_offsets.push_back(static_cast<int>(res->inlier_indices.size()));
    res->label_ids.push_back(id_offset + r);
    constexpr int kLanes = 256;   // vox_stats_kernel: thread t of 256 adds the elements t, t + 256, ...
    double lane[kLanes][3] = {};
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (size_t k = 0; k < l.size(); ++k) {   // lane k % 256 adds its elements in ascending order
      const float* p = xyz + 3 * static_cast<size_t>(l[k]);
      for (int c = 0; c < 3; ++c) {
        lane[k % kLanes][c] += static_cast<double>(p[c]);
        mn[c] = std::fmin(mn[c], p[c]);
        mx[c] = std::fmax(mx[c], p[c]);
      }
    }
    double tot[3] = {0.0, 0.0, 0.0};
    for (int k = 0; k < kLanes; ++k)
      for (int c = 0; c < 3; ++c) tot[c] += lane[k][c];
    for (int c = 0; c < 3; ++c) res->centroids.push_back(static_cast<float>(tot[c] / static_cast<double>(l.size())));
    for (int c = 0; c < 3; ++c) res->bboxes.push_back(mn[c]);
    for (int c = 0; c < 3; ++c) res->bboxes.push_back(mx[c]);
  }
  *out = res;
  return 0;
}

int oracle_mean_shift(const float* xyz, int rows, int cols,
                      const oracle_cluster_config* cfg, int iterations,
                      int id_offset, int sort_mode, int* labels,
                      oracle_regions** out, float* modes_out, int modes_cap,
                      int* n_modes) {
  // ps/mean_shift_segmentation.h:207-330
  constexpr double kSquareDistanceThreshold = 1;            // :33
  constexpr int kHalfSearchWindow = 5;                      // :39
  constexpr double kIntensityRatioThreshold = 0.5;          // :43
  constexpr double kSquaredCentroidDistanceThreshold = 1.0; // :47
  constexpr double kSquaredNeighorDistanceThreshold = 0.2 * 0.2;  // :51
  if (cfg->min_region_inliers < 0 || cfg->half_search_window < 0) return 2;
  Cloud3 pc{xyz, rows, cols};
  const int n = rows * cols;
  auto* regions = new oracle_regions();
  std::vector<double> intensities(n, 1.0);
  std::vector<SeedPointAndIndex> seeds;
  std::vector<int> seed_cloud_index;
  for (int index = 0; index < n; ++index) {
    if (labels[index] == kUnlabeled && !has_nan(pc.at(index))) {
      int row = index % rows, col = index / rows;
      seeds.push_back({pc.at(index),
                       PointIndex{static_cast<float>(row), static_cast<float>(col)}});
      seed_cloud_index.push_back(index);
    }
  }
  std::vector<bool> valid(seeds.size(), true);
  for (int iteration = 0; iteration < iterations; ++iteration) {
    for (size_t i = 0; i < seeds.size(); ++i) {
      if (!valid[i]) continue;
      // GetNeighborsShift, :97-131 fused with GetWeightedShift, :134-146
      // (kernel weights are all 1.0; 1.0 * v is exact, quirk Q9).
      const SeedPointAndIndex& c = seeds[i];
      const int row = c.index.RowIndex();
      const int col = c.index.ColIndex();
      int count = 0;
      V3 shift_distance{0.0f, 0.0f, 0.0f};
      float shift_row = 0.0f, shift_col = 0.0f;
      for (int dc = -kHalfSearchWindow; dc <= kHalfSearchWindow; ++dc) {
        for (int dr = -kHalfSearchWindow; dr <= kHalfSearchWindow; ++dr) {
          const int cc = col + dc, rc = row + dr;
          if (cc < 0 || cc >= cols || rc < 0 || rc >= rows ||
              labels[cc * rows + rc] != kUnlabeled || has_nan(pc.at(rc, cc)))
            continue;
          const V3 d = pc.at(rc, cc) - c.seed_point;
          const float ir = static_cast<float>(rc) - c.index.row;
          const float ic = static_cast<float>(cc) - c.index.col;
          if (static_cast<double>(sqnorm(d)) > kSquareDistanceThreshold) continue;
          shift_distance = shift_distance + d;
          shift_row += ir;
          shift_col += ic;
          ++count;
        }
      }
      const double total_weight = static_cast<double>(count);
      if (total_weight < kIntensityRatioThreshold * kHalfSearchWindow *
                             kHalfSearchWindow * 4) {
        valid[i] = false;
        continue;
      }
      const float tw = static_cast<float>(total_weight);
      seeds[i].seed_point = seeds[i].seed_point + shift_distance / tw;
      seeds[i].index.row += shift_row / tw;
      seeds[i].index.col += shift_col / tw;
      intensities[i] = total_weight;
    }
  }
  if (n_modes) *n_modes = static_cast<int>(seeds.size());
  if (modes_out) {
    for (size_t i = 0; i < seeds.size() && static_cast<int>(i) < modes_cap; ++i) {
      float* o = modes_out + 8 * i;
      o[0] = seeds[i].seed_point.x;
      o[1] = seeds[i].seed_point.y;
      o[2] = seeds[i].seed_point.z;
      o[3] = seeds[i].index.row;
      o[4] = seeds[i].index.col;
      o[5] = static_cast<float>(intensities[i]);
      o[6] = valid[i] ? 1.0f : 0.0f;
      o[7] = static_cast<float>(seed_cloud_index[i]);
    }
  }
  // :262-280
  std::vector<std::pair<SeedPointAndIndex, double>> sorted;
  sorted.reserve(seeds.size());
  for (size_t i = 0; i < seeds.size(); ++i)
    if (valid[i]) sorted.push_back(std::make_pair(seeds[i], intensities[i]));
  auto cmp = [](const std::pair<SeedPointAndIndex, double>& l,
                const std::pair<SeedPointAndIndex, double>& r) {
    return l.second < r.second;
  };
  if (sort_mode == 1)
    std::stable_sort(sorted.begin(), sorted.end(), cmp);
  else
    std::sort(sorted.begin(), sorted.end(), cmp);
  // :282-329
  std::vector<bool> valid_sorted(sorted.size(), true);
  for (size_t i = 0; i < sorted.size(); ++i) {
    if (!valid_sorted[i]) continue;
    ClusterRegion region(*cfg);
    *region.mutable_label_id() = regions->count() + id_offset;
    const SeedPointAndIndex& ci = sorted[i].first;
    region.SetSeed(ci.seed_point);
    std::queue<int> unexamined;
    // point_cloud.LinearizeIndex(RowIndex, ColIndex): unchecked, :296-297
    unexamined.emplace(ci.index.ColIndex() * rows + ci.index.RowIndex());
    bool first = true;
    while (!unexamined.empty()) {
      const int center = unexamined.front();
      unexamined.pop();
      // AddNeighborsAndGrowReigon, :152-199
      const int hw = region.config().half_search_window;
      int row = center % rows, col = center / rows;
      for (int dc = -hw; dc <= hw; ++dc) {
        for (int dr = -hw; dr <= hw; ++dr) {
          if (!first && dc == 0 && dr == 0) continue;
          const int cc = col + dc, rc = row + dr;
          if (cc < 0 || cc >= cols || rc < 0 || rc >= rows ||
              labels[cc * rows + rc] != kUnlabeled || has_nan(pc.at(rc, cc)))
            continue;
          const V3 cand = pc.at(rc, cc);
          if (static_cast<double>(sqnorm(cand - region.Seed())) >
              kSquaredCentroidDistanceThreshold) {
            if (first || static_cast<double>(sqnorm(cand - pc.at(center))) >
                             kSquaredNeighorDistanceThreshold)
              continue;
          }
          const int inlier = cc * rows + rc;
          region.mutable_inlier_indices()->push_back(inlier);
          labels[inlier] = region.label_id();
          unexamined.push(inlier);
        }
      }
      first = false;
    }
    if (region.CheckAndFinalize()) {
      for (size_t j = i + 1; j < sorted.size(); ++j) {
        if (valid_sorted[j] &&
            static_cast<double>(sqnorm(sorted[j].first.seed_point - ci.seed_point)) <
                kSquaredCentroidDistanceThreshold)
          valid_sorted[j] = false;
      }
      region.Export(regions);
    } else {
      for (const int k : region.inlier_indices()) labels[k] = kUnlabeled;
    }
  }
  *out = regions;
  return 0;
}

int oracle_seeds_plane_support(const float* xyz, const float* normals, int rows,
                               int cols, int neighborhood_size,
                               float max_plane_distance, int min_support,
                               int* seeds_out, int cap) {
  // ps/segmentation.h:190-230 + :54-73; AtUnsafe(col,row) passes (col,row) as
  // (row,col): element read is data[row_arg... ] = data[second*rows + first].
  if (rows <= 1) return -1;
  if (rows > cols) return -2;  // transposed reads leave the buffer (UB in ref)
  Cloud3 points{xyz, rows, cols}, nrm{normals, rows, cols};
  const int half = neighborhood_size / 2;
  std::multimap<int, int> ranked;
  if (rows < neighborhood_size || cols < neighborhood_size) return 0;
  for (int col = 0; col < cols; ++col) {
    for (int row = 0; row < rows; ++row) {
      // AtUnsafe(row=col, col=row) -> index = row * Rows + col
      const V3 point = points.at(row * rows + col);
      const V3 normal = nrm.at(row * rows + col);
      if (!all_finite(normal) || !all_finite(point)) continue;
      Plane plane(normal, point);
      int num = 0;
      const int min_row = std::max(0, row - half);
      const int max_row = std::min(rows - 1, row + half);
      const int min_col = std::max(0, col - half);
      const int max_col = std::min(cols - 1, col + half);
      for (int sc = min_col; sc <= max_col; ++sc)
        for (int sr = min_row; sr <= max_row; ++sr)
          if (plane.absDistance(points.at(sr * rows + sc)) < max_plane_distance)
            ++num;
      if (num >= min_support) ranked.emplace(num, col * rows + row);
    }
  }
  int k = 0;
  for (const auto& kv : ranked) {
    if (k < cap) seeds_out[k] = kv.second;
    ++k;
  }
  return k;
}

int oracle_seeds_average_normals(const float* xyz, const float* normals,
                                 int rows, int cols, int neighborhood_size,
                                 int min_num_valid_normals,
                                 float min_avg_normal_length, int* seeds_out,
                                 int cap) {
  // ps/segmentation.h:136-184 + :78-120 (Q1: AtUnsafe(col - n, row + i - k)
  // reads data[(row+i-k) * Rows + (col-n)]).
  (void)xyz;
  if (rows <= 1) return -1;
  if (rows > cols) return -2;  // transposed reads leave the buffer (UB in ref)
  Cloud3 nrm{normals, rows, cols};
  const int half = neighborhood_size / 2;
  const float kSq = min_avg_normal_length * min_avg_normal_length;
  int k = 0;
  if (rows < neighborhood_size || cols < neighborhood_size) return 0;
  std::vector<V3> avg_row(neighborhood_size);
  std::vector<int> nvalid_row(neighborhood_size);
  for (int row = half; row < rows - half; ++row) {
    for (int i = 0; i < neighborhood_size; ++i) {
      avg_row[i] = {0, 0, 0};
      nvalid_row[i] = 0;
    }
    for (int col = 0; col < cols; ++col) {
      V3 avg{0, 0, 0};
      int num_valid = 0;
      for (int i = 0; i < neighborhood_size; ++i) {
        if (col >= neighborhood_size) {
          const V3 leaving = nrm.at((row + i - half) * rows + (col - neighborhood_size));
          if (all_finite(leaving)) {
            avg_row[i] = avg_row[i] - leaving;
            --nvalid_row[i];
          }
        }
        const V3 entering = nrm.at((row + i - half) * rows + col);
        if (all_finite(entering)) {
          avg_row[i] = avg_row[i] + entering;
          ++nvalid_row[i];
        }
        const int new_num = num_valid + nvalid_row[i];
        // (int * Vector3f + int * Vector3f) / int, ints converted to float
        const float a = static_cast<float>(num_valid);
        const float b = static_cast<float>(nvalid_row[i]);
        avg = (a * avg + b * avg_row[i]) / static_cast<float>(new_num);
        num_valid = new_num;
      }
      const float sq = sqnorm(avg);
      if (col >= (neighborhood_size - 1) && num_valid >= min_num_valid_normals &&
          sq >= kSq) {
        if (k < cap) seeds_out[k] = (col * rows + row) - half;
        ++k;
      }
    }
  }
  return k;
}

// The sort of ps/mean_shift_segmentation.h:276-280 alone: order_out[k] = position (in the input
// list) of the element that std::sort (mode 0; this box's libstdc++, quirk Q7) or std::stable_sort
// (mode 1, patch P2) puts at place k when the list is compared by intensity only.
void oracle_sort_order(const double* intensities, int n, int sort_mode, int* order_out) {
  std::vector<std::pair<int, double>> v(n);
  for (int i = 0; i < n; ++i) v[i] = std::make_pair(i, intensities[i]);
  auto cmp = [](const std::pair<int, double>& l, const std::pair<int, double>& r) { return l.second < r.second; };
  if (sort_mode == 1) std::stable_sort(v.begin(), v.end(), cmp);
  else std::sort(v.begin(), v.end(), cmp);
  for (int i = 0; i < n; ++i) order_out[i] = v[i].first;
}

// FindSeedPointsFromLastPlanarRegions, ps/planar_region.h:478-519.  prev_* describe the previous
// regions (inlier_indices().size(), centroid(), plane().normal()); the multimap orders them by
// inlier count, insertion order within a count.  Returns the number of seeds appended.
int oracle_seeds_last_planar_regions(const float* xyz, const float* normals, int n_points, int n_prev,
                                     const int* prev_inlier_counts, const float* prev_centroids,
                                     const float* prev_normals, const float q_xyzw[4], const float t[3],
                                     float max_distance, float max_normal_difference_angle, int* seeds_out) {
  const float kMaxSq = max_distance * max_distance;
  const float kCos = o_cosf(max_normal_difference_angle);
  const Quat q{q_xyzw[0], q_xyzw[1], q_xyzw[2], q_xyzw[3]};
  const V3 tr{t[0], t[1], t[2]};
  struct SeedPoint { V3 point, normal; };
  std::multimap<int, SeedPoint> past;
  for (int i = 0; i < n_prev; ++i) {
    const V3 c{prev_centroids[3 * i], prev_centroids[3 * i + 1], prev_centroids[3 * i + 2]};
    const V3 n{prev_normals[3 * i], prev_normals[3 * i + 1], prev_normals[3 * i + 2]};
    past.emplace(prev_inlier_counts[i], SeedPoint{rotate(q, c) + tr, rotate(q, n)});
  }
  int k = 0;
  for (const auto& kv : past) {
    const SeedPoint& sp = kv.second;
    float min_sq = kMaxSq;
    int min_idx = -1;
    for (int j = 0; j < n_points; ++j) {
      const V3 p{xyz[3 * j], xyz[3 * j + 1], xyz[3 * j + 2]};
      const V3 nr{normals[3 * j], normals[3 * j + 1], normals[3 * j + 2]};
      const float sq = sqnorm(sp.point - p);
      if (sq < min_sq && dot(sp.normal, nr) > kCos) {
        min_sq = sq;
        min_idx = j;
      }
    }
    if (min_idx >= 0) seeds_out[k++] = min_idx;
  }
  return k;
}

// MultichannelCloud::TransformInPlace, ps/multichannel_cloud.cc:53-68.  Pose3d::cast<float>() first
// (:55); points: Pose3f * p = q * p + t (eigenmath::Pose3, x_edr_eigenmath 1.0.0, restated:
// Eigen quaternion-vector product, then + translation); normals: rotationMatrix() * n with Eigen's
// Quaternion::toRotationMatrix and the fixed-size 3-term reduction e0 + (e1 + e2).
void oracle_transform_in_place(float* points, float* normals, int n, const double q_xyzw[4], const double t[3]) {
  const Quat q{(float)q_xyzw[0], (float)q_xyzw[1], (float)q_xyzw[2], (float)q_xyzw[3]};
  const V3 tr{(float)t[0], (float)t[1], (float)t[2]};
  const float tx = 2.0f * q.x, ty = 2.0f * q.y, tz = 2.0f * q.z;
  const float twx = tx * q.w, twy = ty * q.w, twz = tz * q.w;
  const float txx = tx * q.x, txy = ty * q.x, txz = tz * q.x;
  const float tyy = ty * q.y, tyz = tz * q.y, tzz = tz * q.z;
  const float r00 = 1.0f - (tyy + tzz), r01 = txy - twz, r02 = txz + twy;
  const float r10 = txy + twz, r11 = 1.0f - (txx + tzz), r12 = tyz - twx;
  const float r20 = txz - twy, r21 = tyz + twx, r22 = 1.0f - (txx + tyy);
  for (int i = 0; i < n; ++i) {
    if (points) {
      float* p = points + 3 * (size_t)i;
      const V3 w = rotate(q, V3{p[0], p[1], p[2]}) + tr;
      p[0] = w.x; p[1] = w.y; p[2] = w.z;
    }
    if (normals) {
      float* p = normals + 3 * (size_t)i;
      const V3 v{p[0], p[1], p[2]};
      p[0] = r00 * v.x + (r01 * v.y + r02 * v.z);
      p[1] = r10 * v.x + (r11 * v.y + r12 * v.z);
      p[2] = r20 * v.x + (r21 * v.y + r22 * v.z);
    }
  }
}

int oracle_regions_count(const oracle_regions* r) { return r->count(); }
const int* oracle_regions_inlier_offsets(const oracle_regions* r) { return r->inlier_offsets.data(); }
const int* oracle_regions_inlier_indices(const oracle_regions* r) { return r->inlier_indices.data(); }
const int* oracle_regions_label_ids(const oracle_regions* r) { return r->label_ids.data(); }
const int* oracle_regions_seed_indices(const oracle_regions* r) { return r->seed_indices.data(); }
const float* oracle_regions_planes(const oracle_regions* r) { return r->planes.data(); }
const float* oracle_regions_centroids(const oracle_regions* r) { return r->centroids.data(); }
const float* oracle_regions_areas(const oracle_regions* r) { return r->areas.data(); }
const float* oracle_regions_estimators(const oracle_regions* r) { return r->estimators.data(); }
const int* oracle_regions_boundary_offsets(const oracle_regions* r) { return r->boundary_offsets.data(); }
const int* oracle_regions_boundary_indices(const oracle_regions* r) { return r->boundary_indices.data(); }
const int* oracle_regions_discontinuous_offsets(const oracle_regions* r) { return r->disc_offsets.data(); }
const int* oracle_regions_discontinuous_indices(const oracle_regions* r) { return r->disc_indices.data(); }
const int* oracle_regions_hull_offsets(const oracle_regions* r) { return r->hull_offsets.data(); }
const float* oracle_regions_hull_points(const oracle_regions* r) { return r->hull_points.data(); }
const float* oracle_regions_bboxes(const oracle_regions* r) { return r->bboxes.data(); }
const float* oracle_regions_seed_points(const oracle_regions* r) { return r->seed_points.data(); }
void oracle_regions_free(oracle_regions* r) { delete r; }

}  // extern "C"
