// shim_eigen.h -- TEST INFRASTRUCTURE ONLY (oracle/_ref build).
//
// Minimal stand-in for the part of Eigen 3.4.0 (WORKSPACE:46-52 of the
// reference; the dependency is not vendored and not in this image) that the
// reference's hot-path sources touch, so that those sources can be compiled IN
// PLACE from /root/reference (see oracle/Makefile).  Eager evaluation, fixed
// sizes, column-major storage.  Arithmetic follows Eigen 3.4.0's published
// behaviour for non-vectorised fixed-size types:
//   * 3-element reductions (sum, dot, squaredNorm, trace) use the fixed-size
//     unroller's split [0,1) + [1,3)  =>  e0 + (e1 + e2)  (Core/Redux.h);
//   * v / s divides every coefficient, v * s multiplies every coefficient;
//   * Hyperplane(n, e): offset = -n.dot(e), no normalisation (Geometry/Hyperplane.h);
//   * Quaternion * vector: uv = q.vec x v; uv += uv; v + w*uv + q.vec x uv;
//   * SelfAdjointEigenSolver<Matrix3f>::computeDirect: closed-form 3x3 path
//     (Eigenvalues/SelfAdjointEigenSolver.h, direct_selfadjoint_eigenvalues<.,3,false>).
// Scalars of another arithmetic type are converted to the matrix scalar first
// (Eigen's promote_scalar_arg rule).  Bit-level equality with a real Eigen
// build is NOT pinned.
#ifndef ORACLE_SHIMS_EIGENMATH_SHIM_EIGEN_H_
#define ORACLE_SHIMS_EIGENMATH_SHIM_EIGEN_H_

#include <cfloat>
#include <cmath>
#include <cstddef>
#include <limits>
#include <memory>
#include <type_traits>
#include <vector>

namespace shim {
// 0 = float trig evaluated in double and rounded once (default; matches the
// oracle port and the CUDA path), 1 = this box's libm float functions.
inline int& native_libm() {
  static int v = 0;
  return v;
}
inline float atan2_f(float y, float x) {
  return native_libm() ? ::atan2f(y, x) : static_cast<float>(::atan2(static_cast<double>(y), static_cast<double>(x)));
}
inline float cos_f(float x) { return native_libm() ? ::cosf(x) : static_cast<float>(::cos(static_cast<double>(x))); }
inline float sin_f(float x) { return native_libm() ? ::sinf(x) : static_cast<float>(::sin(static_cast<double>(x))); }
inline double atan2_f(double y, double x) { return ::atan2(y, x); }
inline double cos_f(double x) { return ::cos(x); }
inline double sin_f(double x) { return ::sin(x); }
}  // namespace shim

namespace Eigen {

using Index = std::ptrdiff_t;
constexpr int Dynamic = -1;
enum ComputationInfo { Success = 0, NumericalIssue = 1, NoConvergence = 2, InvalidInput = 3 };
enum { ComputeFullU = 0x04, ComputeThinU = 0x08, ComputeFullV = 0x10, ComputeThinV = 0x20 };
struct NoQRPreconditioner {};
template <typename M, typename P = void>
class JacobiSVD;  // only named by ICP templates the hot path never instantiates

template <typename T>
using aligned_allocator = std::allocator<T>;

template <typename Derived>
struct MatrixBase {};

template <typename T, int R, int C>
class Matrix : public MatrixBase<Matrix<T, R, C>> {
 public:
  using Scalar = T;
  static constexpr int N = R * C;
  T a[N];

  Matrix() {}
  template <typename A, typename B,
            typename = std::enable_if_t<std::is_arithmetic<A>::value && std::is_arithmetic<B>::value && N == 2>>
  Matrix(A x, B y) {
    a[0] = static_cast<T>(x);
    a[1] = static_cast<T>(y);
  }
  template <typename A, typename B, typename D,
            typename = std::enable_if_t<std::is_arithmetic<A>::value && N == 3>>
  Matrix(A x, B y, D z) {
    a[0] = static_cast<T>(x);
    a[1] = static_cast<T>(y);
    a[2] = static_cast<T>(z);
  }
  template <typename A, typename B, typename D, typename E,
            typename = std::enable_if_t<std::is_arithmetic<A>::value && N == 4>>
  Matrix(A x, B y, D z, E w) {
    a[0] = static_cast<T>(x);
    a[1] = static_cast<T>(y);
    a[2] = static_cast<T>(z);
    a[3] = static_cast<T>(w);
  }

  static Matrix Constant(T v) {
    Matrix m;
    for (int i = 0; i < N; ++i) m.a[i] = v;
    return m;
  }
  static Matrix Zero() { return Constant(T(0)); }
  static Matrix Ones() { return Constant(T(1)); }
  static Matrix Identity() {
    Matrix m = Zero();
    for (int i = 0; i < (R < C ? R : C); ++i) m.a[i * R + i] = T(1);
    return m;
  }
  void setZero() { *this = Zero(); }
  void setIdentity() { *this = Identity(); }

  T& x() { return a[0]; }
  T& y() { return a[1]; }
  T& z() { return a[2]; }
  T& w() { return a[3]; }
  const T& x() const { return a[0]; }
  const T& y() const { return a[1]; }
  const T& z() const { return a[2]; }
  const T& w() const { return a[3]; }
  T& operator[](Index i) { return a[i]; }
  const T& operator[](Index i) const { return a[i]; }
  T& operator()(Index i) { return a[i]; }
  const T& operator()(Index i) const { return a[i]; }
  T& operator()(Index r, Index c) { return a[c * R + r]; }
  const T& operator()(Index r, Index c) const { return a[c * R + r]; }
  T& coeffRef(Index i) { return a[i]; }
  const T& coeffRef(Index i) const { return a[i]; }
  const T& coeff(Index i) const { return a[i]; }
  T& coeffRef(Index r, Index c) { return a[c * R + r]; }
  const T& coeff(Index r, Index c) const { return a[c * R + r]; }
  Index size() const { return N; }
  Index rows() const { return R; }
  Index cols() const { return C; }

  Matrix<T, R, 1> col(Index c) const {
    Matrix<T, R, 1> v;
    for (int r = 0; r < R; ++r) v.a[r] = a[c * R + r];
    return v;
  }
  void setCol(Index c, const Matrix<T, R, 1>& v) {
    for (int r = 0; r < R; ++r) a[c * R + r] = v.a[r];
  }
  Matrix<T, C, R> transpose() const {
    Matrix<T, C, R> t;
    for (int r = 0; r < R; ++r)
      for (int c = 0; c < C; ++c) t.a[r * C + c] = a[c * R + r];
    return t;
  }
  template <typename U>
  Matrix<U, R, C> cast() const {
    Matrix<U, R, C> m;
    for (int i = 0; i < N; ++i) m.a[i] = static_cast<U>(a[i]);
    return m;
  }

  Matrix operator-() const {
    Matrix m;
    for (int i = 0; i < N; ++i) m.a[i] = -a[i];
    return m;
  }
  Matrix& operator+=(const Matrix& o) {
    for (int i = 0; i < N; ++i) a[i] += o.a[i];
    return *this;
  }
  Matrix& operator-=(const Matrix& o) {
    for (int i = 0; i < N; ++i) a[i] -= o.a[i];
    return *this;
  }
  template <typename S, typename = std::enable_if_t<std::is_arithmetic<S>::value>>
  Matrix& operator*=(S s) {
    const T t = static_cast<T>(s);
    for (int i = 0; i < N; ++i) a[i] *= t;
    return *this;
  }
  template <typename S, typename = std::enable_if_t<std::is_arithmetic<S>::value>>
  Matrix& operator/=(S s) {
    const T t = static_cast<T>(s);
    for (int i = 0; i < N; ++i) a[i] /= t;
    return *this;
  }
  friend Matrix operator+(const Matrix& l, const Matrix& r) {
    Matrix m;
    for (int i = 0; i < N; ++i) m.a[i] = l.a[i] + r.a[i];
    return m;
  }
  friend Matrix operator-(const Matrix& l, const Matrix& r) {
    Matrix m;
    for (int i = 0; i < N; ++i) m.a[i] = l.a[i] - r.a[i];
    return m;
  }
  template <typename S, typename = std::enable_if_t<std::is_arithmetic<S>::value>>
  friend Matrix operator*(const Matrix& l, S s) {
    const T t = static_cast<T>(s);
    Matrix m;
    for (int i = 0; i < N; ++i) m.a[i] = l.a[i] * t;
    return m;
  }
  template <typename S, typename = std::enable_if_t<std::is_arithmetic<S>::value>>
  friend Matrix operator*(S s, const Matrix& r) {
    const T t = static_cast<T>(s);
    Matrix m;
    for (int i = 0; i < N; ++i) m.a[i] = t * r.a[i];
    return m;
  }
  template <typename S, typename = std::enable_if_t<std::is_arithmetic<S>::value>>
  friend Matrix operator/(const Matrix& l, S s) {
    const T t = static_cast<T>(s);
    Matrix m;
    for (int i = 0; i < N; ++i) m.a[i] = l.a[i] / t;
    return m;
  }

  // Redux.h fixed-size unroller: redux(start, len) = redux(start, len/2) op redux(start+len/2, len-len/2)
  static T ReduxSum(const T* v, int start, int len) {
    if (len == 1) return v[start];
    const int half = len / 2;
    return ReduxSum(v, start, half) + ReduxSum(v, start + half, len - half);
  }
  T sum() const { return ReduxSum(a, 0, N); }
  T dot(const Matrix& o) const {
    T p[N];
    for (int i = 0; i < N; ++i) p[i] = a[i] * o.a[i];
    return ReduxSum(p, 0, N);
  }
  T squaredNorm() const { return dot(*this); }
  T norm() const { return std::sqrt(squaredNorm()); }
  T trace() const {
    T d[R];
    for (int i = 0; i < R; ++i) d[i] = a[i * R + i];
    return ReduxSum(d, 0, R);
  }
  Matrix normalized() const {  // Core/Dot.h: n / sqrt(z) if z > 0
    const T z = squaredNorm();
    if (z > T(0)) return *this / std::sqrt(z);
    return *this;
  }
  void normalize() {
    const T z = squaredNorm();
    if (z > T(0)) *this /= std::sqrt(z);
  }
  Matrix cross(const Matrix& b) const {
    static_assert(N == 3, "cross is for 3-vectors");
    return Matrix(a[1] * b.a[2] - a[2] * b.a[1], a[2] * b.a[0] - a[0] * b.a[2], a[0] * b.a[1] - a[1] * b.a[0]);
  }
  bool allFinite() const {
    for (int i = 0; i < N; ++i)
      if (!std::isfinite(a[i])) return false;
    return true;
  }
  bool hasNaN() const {
    for (int i = 0; i < N; ++i)
      if (std::isnan(a[i])) return true;
    return false;
  }
  T determinant() const {
    static_assert(R == 3 && C == 3, "determinant only for 3x3");
    return a[0] * (a[4] * a[8] - a[7] * a[5]) - a[3] * (a[1] * a[8] - a[7] * a[2]) + a[6] * (a[1] * a[5] - a[4] * a[2]);
  }
};

// Lazy coefficient product of Eigen for small fixed sizes: (row . col) reduced by the unroller.
template <typename T, int R, int K, int C>
Matrix<T, R, C> operator*(const Matrix<T, R, K>& l, const Matrix<T, K, C>& r) {
  Matrix<T, R, C> m;
  for (int c = 0; c < C; ++c)
    for (int i = 0; i < R; ++i) {
      T p[K];
      for (int k = 0; k < K; ++k) p[k] = l(i, k) * r(k, c);
      m(i, c) = Matrix<T, K, 1>::ReduxSum(p, 0, K);
    }
  return m;
}

using Vector2f = Matrix<float, 2, 1>;
using Vector2d = Matrix<double, 2, 1>;
using Vector3f = Matrix<float, 3, 1>;
using Vector3d = Matrix<double, 3, 1>;
using Matrix3f = Matrix<float, 3, 3>;
using Matrix3d = Matrix<double, 3, 3>;

// Dynamic double vector: only Ones(n), size(), sum(), operator[] are used
// (ps/mean_shift_segmentation.h:89-93,136-146,242-244).
class VectorXd {
 public:
  VectorXd() {}
  static VectorXd Ones(Index n) {
    VectorXd v;
    v.d_.assign(static_cast<size_t>(n), 1.0);
    return v;
  }
  Index size() const { return static_cast<Index>(d_.size()); }
  double operator[](Index i) const { return d_[static_cast<size_t>(i)]; }
  double sum() const {  // all entries are 1.0: every summation order is exact
    double s = 0.0;
    for (double v : d_) s += v;
    return s;
  }

 private:
  std::vector<double> d_;
};

// Geometry/Hyperplane.h, AmbientDim = 3: coeffs = (normal, offset).
template <typename T, int Dim>
class Hyperplane {
  static_assert(Dim == 3, "only 3-D planes");

 public:
  using VectorType = Matrix<T, 3, 1>;
  using Coefficients = Matrix<T, 4, 1>;
  Hyperplane() {}
  Hyperplane(const VectorType& n, const VectorType& e) {
    normal() = n;
    offset() = -n.dot(e);
  }
  Hyperplane(const VectorType& n, const T& d) {
    normal() = n;
    offset() = d;
  }
  VectorType& normal() { return *reinterpret_cast<VectorType*>(c_.a); }
  const VectorType& normal() const { return *reinterpret_cast<const VectorType*>(c_.a); }
  T& offset() { return c_.a[3]; }
  const T& offset() const { return c_.a[3]; }
  Coefficients& coeffs() { return c_; }
  const Coefficients& coeffs() const { return c_; }
  T signedDistance(const VectorType& p) const { return normal().dot(p) + offset(); }
  T absDistance(const VectorType& p) const { return std::abs(signedDistance(p)); }
  VectorType projection(const VectorType& p) const { return p - signedDistance(p) * normal(); }

 private:
  Coefficients c_;
};

// Geometry/Quaternion.h (coefficients stored x, y, z, w).
template <typename T>
class Quaternion {
 public:
  using Vec3 = Matrix<T, 3, 1>;
  using Mat3 = Matrix<T, 3, 3>;
  Quaternion() {}
  Quaternion(T w, T x, T y, T z) : x_(x), y_(y), z_(z), w_(w) {}
  explicit Quaternion(const Mat3& m) {  // Shepperd's method as in Eigen's quaternionbase_assign_impl
    T t = m.trace();
    if (t > T(0)) {
      t = std::sqrt(t + T(1));
      w_ = T(0.5) * t;
      t = T(0.5) / t;
      x_ = (m(2, 1) - m(1, 2)) * t;
      y_ = (m(0, 2) - m(2, 0)) * t;
      z_ = (m(1, 0) - m(0, 1)) * t;
    } else {
      int i = 0;
      if (m(1, 1) > m(0, 0)) i = 1;
      if (m(2, 2) > m(i, i)) i = 2;
      const int j = (i + 1) % 3, k = (j + 1) % 3;
      t = std::sqrt(m(i, i) - m(j, j) - m(k, k) + T(1));
      T q[3];
      q[i] = T(0.5) * t;
      t = T(0.5) / t;
      w_ = (m(k, j) - m(j, k)) * t;
      q[j] = (m(j, i) + m(i, j)) * t;
      q[k] = (m(k, i) + m(i, k)) * t;
      x_ = q[0];
      y_ = q[1];
      z_ = q[2];
    }
  }
  static Quaternion Identity() { return Quaternion(T(1), T(0), T(0), T(0)); }
  T x() const { return x_; }
  T y() const { return y_; }
  T z() const { return z_; }
  T w() const { return w_; }
  Vec3 vec() const { return Vec3(x_, y_, z_); }
  Quaternion conjugate() const { return Quaternion(w_, -x_, -y_, -z_); }
  T squaredNorm() const { return (x_ * x_ + y_ * y_) + (z_ * z_ + w_ * w_); }
  Quaternion normalized() const {
    const T n = std::sqrt(squaredNorm());
    return Quaternion(w_ / n, x_ / n, y_ / n, z_ / n);
  }
  Quaternion inverse() const {
    const T n2 = squaredNorm();
    if (n2 > T(0)) return Quaternion(w_ / n2, -x_ / n2, -y_ / n2, -z_ / n2);
    return Quaternion(T(0), T(0), T(0), T(0));
  }
  // QuaternionBase::_transformVector
  Vec3 operator*(const Vec3& v) const {
    Vec3 uv = vec().cross(v);
    uv += uv;
    return v + w_ * uv + vec().cross(uv);
  }
  Quaternion operator*(const Quaternion& b) const {
    const Quaternion& a = *this;
    return Quaternion(a.w_ * b.w_ - a.x_ * b.x_ - a.y_ * b.y_ - a.z_ * b.z_,
                      a.w_ * b.x_ + a.x_ * b.w_ + a.y_ * b.z_ - a.z_ * b.y_,
                      a.w_ * b.y_ + a.y_ * b.w_ + a.z_ * b.x_ - a.x_ * b.z_,
                      a.w_ * b.z_ + a.z_ * b.w_ + a.x_ * b.y_ - a.y_ * b.x_);
  }
  Mat3 toRotationMatrix() const {
    Mat3 r;
    const T tx = T(2) * x_, ty = T(2) * y_, tz = T(2) * z_;
    const T twx = tx * w_, twy = ty * w_, twz = tz * w_;
    const T txx = tx * x_, txy = ty * x_, txz = tz * x_;
    const T tyy = ty * y_, tyz = tz * y_, tzz = tz * z_;
    r(0, 0) = T(1) - (tyy + tzz);
    r(0, 1) = txy - twz;
    r(0, 2) = txz + twy;
    r(1, 0) = txy + twz;
    r(1, 1) = T(1) - (txx + tzz);
    r(1, 2) = tyz - twx;
    r(2, 0) = txz - twy;
    r(2, 1) = tyz + twx;
    r(2, 2) = T(1) - (txx + tyy);
    return r;
  }
  template <typename U>
  Quaternion<U> cast() const {
    return Quaternion<U>(static_cast<U>(w_), static_cast<U>(x_), static_cast<U>(y_), static_cast<U>(z_));
  }
  // Quaternion::FromTwoVectors(a, b) (Geometry/Quaternion.h setFromTwoVectors), non-degenerate branch
  // plus the c ~ -1 fallback reduced to a fixed perpendicular axis.
  static Quaternion FromTwoVectors(const Vec3& a, const Vec3& b) {
    const Vec3 v0 = a.normalized();
    const Vec3 v1 = b.normalized();
    const T c = v1.dot(v0);
    if (c < T(-1) + std::numeric_limits<T>::epsilon()) return Quaternion(T(0), T(1), T(0), T(0));
    const Vec3 axis = v0.cross(v1);
    const T s = std::sqrt((T(1) + c) * T(2));
    const T invs = T(1) / s;
    return Quaternion(s * T(0.5), axis.x() * invs, axis.y() * invs, axis.z() * invs);
  }

 private:
  T x_, y_, z_, w_;
};
using Quaternionf = Quaternion<float>;
using Quaterniond = Quaternion<double>;

// Eigenvalues/SelfAdjointEigenSolver.h: only the closed-form 3x3 path
// (computeDirect -> direct_selfadjoint_eigenvalues<Solver, 3, false>::run) is provided.
template <typename MatrixType>
class SelfAdjointEigenSolver {
 public:
  using Scalar = typename MatrixType::Scalar;
  using VectorType = Matrix<Scalar, 3, 1>;
  SelfAdjointEigenSolver() : info_(InvalidInput) {}
  ComputationInfo info() const { return info_; }
  const VectorType& eigenvalues() const { return eivals_; }
  const MatrixType& eigenvectors() const { return eivecs_; }

  SelfAdjointEigenSolver& computeDirect(const MatrixType& mat) {
    // shift to the mean eigenvalue, map the coefficients to [-1, 1]
    const Scalar shift = mat.trace() / Scalar(3);
    MatrixType scaled;  // mat.selfadjointView<Lower>()
    for (int c = 0; c < 3; ++c)
      for (int r = 0; r < 3; ++r) scaled(r, c) = (r >= c) ? mat(r, c) : mat(c, r);
    for (int i = 0; i < 3; ++i) scaled(i, i) -= shift;
    Scalar scale = Scalar(0);  // cwiseAbs().maxCoeff()
    for (int i = 0; i < 9; ++i) {
      const Scalar v = std::abs(scaled.a[i]);
      if (v > scale) scale = v;
    }
    if (scale > Scalar(0)) scaled /= scale;
    ComputeRoots(scaled, &eivals_);
    if ((eivals_(2) - eivals_(0)) <= std::numeric_limits<Scalar>::epsilon()) {
      eivecs_.setIdentity();
    } else {
      MatrixType tmp = scaled;
      Scalar d0 = eivals_(2) - eivals_(1);
      Scalar d1 = eivals_(1) - eivals_(0);
      int k = 0, l = 2;
      if (d0 > d1) {
        k = 2;
        l = 0;
        d0 = d1;
      }
      VectorType vk, vl;
      for (int i = 0; i < 3; ++i) tmp(i, i) -= eivals_(k);
      ExtractKernel(tmp, &vk, &vl);  // vl = representative column
      if (d0 <= 2 * std::numeric_limits<Scalar>::epsilon() * d1) {
        vl -= vk.dot(vl) * vl;  // sic: as written in Eigen 3.4.0
        vl.normalize();
      } else {
        tmp = scaled;
        for (int i = 0; i < 3; ++i) tmp(i, i) -= eivals_(l);
        VectorType dummy;
        ExtractKernel(tmp, &vl, &dummy);
      }
      eivecs_.setCol(k, vk);
      eivecs_.setCol(l, vl);
      eivecs_.setCol(1, eivecs_.col(2).cross(eivecs_.col(0)).normalized());
    }
    eivals_ *= scale;
    for (int i = 0; i < 3; ++i) eivals_(i) += shift;
    info_ = Success;
    return *this;
  }

 private:
  static void ComputeRoots(const MatrixType& m, VectorType* roots) {
    const Scalar s_inv3 = Scalar(1) / Scalar(3);
    const Scalar s_sqrt3 = std::sqrt(Scalar(3));
    const Scalar c0 = m(0, 0) * m(1, 1) * m(2, 2) + Scalar(2) * m(1, 0) * m(2, 0) * m(2, 1) -
                      m(0, 0) * m(2, 1) * m(2, 1) - m(1, 1) * m(2, 0) * m(2, 0) - m(2, 2) * m(1, 0) * m(1, 0);
    const Scalar c1 = m(0, 0) * m(1, 1) - m(1, 0) * m(1, 0) + m(0, 0) * m(2, 2) - m(2, 0) * m(2, 0) +
                      m(1, 1) * m(2, 2) - m(2, 1) * m(2, 1);
    const Scalar c2 = m(0, 0) + m(1, 1) + m(2, 2);
    const Scalar c2_over_3 = c2 * s_inv3;
    Scalar a_over_3 = (c2 * c2_over_3 - c1) * s_inv3;
    a_over_3 = a_over_3 > Scalar(0) ? a_over_3 : Scalar(0);  // numext::maxi(a_over_3, 0)
    const Scalar half_b = Scalar(0.5) * (c0 + c2_over_3 * (Scalar(2) * c2_over_3 * c2_over_3 - c1));
    Scalar q = a_over_3 * a_over_3 * a_over_3 - half_b * half_b;
    q = q > Scalar(0) ? q : Scalar(0);
    const Scalar rho = std::sqrt(a_over_3);
    const Scalar theta = shim::atan2_f(std::sqrt(q), half_b) * s_inv3;
    const Scalar cos_theta = shim::cos_f(theta);
    const Scalar sin_theta = shim::sin_f(theta);
    (*roots)(0) = c2_over_3 - rho * (cos_theta + s_sqrt3 * sin_theta);
    (*roots)(1) = c2_over_3 - rho * (cos_theta - s_sqrt3 * sin_theta);
    (*roots)(2) = c2_over_3 + Scalar(2) * rho * cos_theta;
  }
  static void ExtractKernel(const MatrixType& mat, VectorType* res, VectorType* representative) {
    int i0 = 0;  // mat.diagonal().cwiseAbs().maxCoeff(&i0): first maximum
    Scalar best = std::abs(mat(0, 0));
    for (int i = 1; i < 3; ++i) {
      const Scalar v = std::abs(mat(i, i));
      if (v > best) {
        best = v;
        i0 = i;
      }
    }
    *representative = mat.col(i0);
    const VectorType c0 = representative->cross(mat.col((i0 + 1) % 3));
    const VectorType c1 = representative->cross(mat.col((i0 + 2) % 3));
    const Scalar n0 = c0.squaredNorm();
    const Scalar n1 = c1.squaredNorm();
    if (n0 > n1) *res = c0 / std::sqrt(n0);
    else *res = c1 / std::sqrt(n1);
  }
  VectorType eivals_;
  MatrixType eivecs_;
  ComputationInfo info_;
};

}  // namespace Eigen
#endif  // ORACLE_SHIMS_EIGENMATH_SHIM_EIGEN_H_
