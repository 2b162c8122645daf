// TEST INFRASTRUCTURE ONLY.  Stand-in for eigenmath/pose3.h (x_edr_eigenmath 1.0.0):
// rigid transform = unit quaternion + translation; p' = q * p + t.
#ifndef ORACLE_SHIMS_EIGENMATH_POSE3_H_
#define ORACLE_SHIMS_EIGENMATH_POSE3_H_
#include "eigenmath/types.h"
namespace eigenmath {
template <typename T>
class Pose3 {
 public:
  Pose3() : t_(Vector3<T>::Zero()), q_(Quaternion<T>::Identity()) {}
  Pose3(const Quaternion<T>& q, const Vector3<T>& t) : t_(t), q_(q) {}
  static Pose3 Identity() { return Pose3(); }
  const Vector3<T>& translation() const { return t_; }
  Vector3<T>& translation() { return t_; }
  const Quaternion<T>& quaternion() const { return q_; }
  void setQuaternion(const Quaternion<T>& q) { q_ = q; }
  Matrix3<T> rotationMatrix() const { return q_.toRotationMatrix(); }
  template <typename U>
  Pose3<U> cast() const { return Pose3<U>(q_.template cast<U>(), t_.template cast<U>()); }
  Vector3<T> operator*(const Vector3<T>& p) const { return q_ * p + t_; }
  Pose3 operator*(const Pose3& o) const { return Pose3(q_ * o.q_, q_ * o.t_ + t_); }
  // The rotation of a pose is a UNIT quaternion (eigenmath keeps it in an SO3), so its inverse is the
  // conjugate -- no division by the squared norm.  x_edr_eigenmath is not vendored; the CPU
  // restatement (pcseg_oracle.cc: PlanarConvexHull) and the CUDA path (make_plane_frame) state the
  // same choice, and tests/test_gpu_configs.py::test_c3 (33 regions, hull vertex counts) keeps the
  // three in agreement.
  Pose3 inverse() const {
    const Quaternion<T> qi = q_.conjugate();
    return Pose3(qi, -(qi * t_));
  }
 private:
  Vector3<T> t_;
  Quaternion<T> q_;
};
using Pose3f = Pose3<float>;
using Pose3d = Pose3<double>;
}  // namespace eigenmath
#endif
