// TEST INFRASTRUCTURE ONLY.  Stand-in for eigenmath/plane_conversions.h (x_edr_eigenmath 1.0.0,
// not vendored).  Restated: the plane frame has its z axis along the plane normal
// (Quaternion::FromTwoVectors(UnitZ, normal)) and its origin at the point of the plane closest
// to the world origin (-offset * normal).  Used only by PlanarConvexHull (ps/algorithms.h:530).
#ifndef ORACLE_SHIMS_EIGENMATH_PLANE_CONVERSIONS_H_
#define ORACLE_SHIMS_EIGENMATH_PLANE_CONVERSIONS_H_
#include "eigenmath/pose3.h"
namespace eigenmath {
template <typename T>
Pose3<T> PoseFromPlane(const Eigen::Hyperplane<T, 3>& plane) {
  const Vector3<T> n = plane.normal();
  const Quaternion<T> q = Quaternion<T>::FromTwoVectors(Vector3<T>(T(0), T(0), T(1)), n);
  return Pose3<T>(q, (-plane.offset()) * n);
}
}  // namespace eigenmath
#endif
