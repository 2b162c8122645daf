// TEST INFRASTRUCTURE ONLY.  Stand-in for x_edr_eigenmath 1.0.0 eigenmath/types.h
// (WORKSPACE:66-74 of the reference; not vendored): typedefs over the Eigen shim.
#ifndef ORACLE_SHIMS_EIGENMATH_TYPES_H_
#define ORACLE_SHIMS_EIGENMATH_TYPES_H_
#include "eigenmath/shim_eigen.h"
namespace eigenmath {
template <typename T> using Vector2 = Eigen::Matrix<T, 2, 1>;
template <typename T> using Vector3 = Eigen::Matrix<T, 3, 1>;
template <typename T> using Matrix3 = Eigen::Matrix<T, 3, 3>;
template <typename T> using Quaternion = Eigen::Quaternion<T>;
using Vector2f = Eigen::Vector2f;
using Vector2d = Eigen::Vector2d;
using Vector3f = Eigen::Vector3f;
using Vector3d = Eigen::Vector3d;
using Vector6f = Eigen::Matrix<float, 6, 1>;
using Matrix3f = Eigen::Matrix3f;
using Matrix3d = Eigen::Matrix3d;
using VectorXd = Eigen::VectorXd;
using Quaternionf = Eigen::Quaternionf;
using Quaterniond = Eigen::Quaterniond;
using Plane3f = Eigen::Hyperplane<float, 3>;
using Plane3d = Eigen::Hyperplane<double, 3>;
}  // namespace eigenmath
#endif
