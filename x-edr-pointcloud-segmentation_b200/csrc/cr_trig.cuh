// cr_trig.cuh -- float trig of the eigen-solve, evaluated in double and
// rounded once, on the restricted domains the hot path needs.
//
// Eigen's computeDirect (called from ps/plane_estimator.cc:202-203) evaluates
//   theta = atan2f(sqrt(q), half_b) / 3;  cosf(theta);  sinf(theta)
// with sqrt(q) >= 0, hence atan2f in [0, pi] and theta in [0, pi/3].  The
// policy for libm results is "evaluate in double to ~1e-16 relative, round to
// float once" (DESIGN.md); CUDA's generic double atan2/sincos do that but sit
// on the serial chain of the planar growth (one re-fit per 30 inliers), so
// these are branch-light restatements for exactly those domains:
//   * atan2: octant folding, then atan(mn/mx) = k*pi/16 + atan(t) with
//     t = (mn - c_k mx) / (mx + c_k mn), c_k = tan(k pi/16), |t| <= tan(pi/32),
//     one double division, 9-term odd polynomial;
//   * sin / cos on [0, 1.06]: direct polynomials, no range reduction.
// Coefficients: tools/gen_trig_coeffs.py (Chebyshev fits, mpmath 60 digits).
// Arguments outside the fast domain take CUDA's generic double functions.
#pragma once
#include <cmath>
#include <cuda_runtime.h>

namespace pcseg {

__device__ __forceinline__ double fast_atan2_pos(double y, double x) {
  // y >= 0, not both zero, both finite.
  const double ax = fabs(x);
  const bool swap = y > ax;
  const double mn = swap ? ax : y;
  const double mx = swap ? y : ax;
  // k = round(atan(mn/mx) / (pi/16)) through 4 threshold compares
  int k = 0;
  k += (mn > mx * 0x1.936bb8c5b2da2p-4) ? 1 : 0;
  k += (mn > mx * 0x1.36a08355c63dcp-2) ? 1 : 0;
  k += (mn > mx * 0x1.11ab7190834ecp-1) ? 1 : 0;
  k += (mn > mx * 0x1.a43002ae42850p-1) ? 1 : 0;
  double c = 0.0, base = 0.0;
  if (k == 1) { c = 0x1.975f5e0553158p-3; base = 0x1.921fb54442d18p-3; }
  if (k == 2) { c = 0x1.a827999fcef32p-2; base = 0x1.921fb54442d18p-2; }
  if (k == 3) { c = 0x1.561b82ab7f990p-1; base = 0x1.2d97c7f3321d2p-1; }
  if (k == 4) { c = 1.0; base = 0x1.921fb54442d18p-1; }
  const double num = fma(-c, mx, mn);
  const double den = fma(c, mn, mx);
  const double t = num / den;
  const double u = t * t;
  double p = -0x1.9ddcab1d76a95p-5;
  p = fma(p, u, 0x1.e18ec24c24a0bp-5);
  p = fma(p, u, -0x1.1110a4e9d9875p-4);
  p = fma(p, u, 0x1.3b13b096c7c1ep-4);
  p = fma(p, u, -0x1.745d17453eb8cp-4);
  p = fma(p, u, 0x1.c71c71c71c282p-4);
  p = fma(p, u, -0x1.2492492492492p-3);
  p = fma(p, u, 0x1.999999999999ap-3);
  p = fma(p, u, -0x1.5555555555555p-2);
  double r = base + fma(t * u, p, t);  // atan(mn/mx) in [0, pi/4]
  if (swap) r = 0x1.921fb54442d18p+0 - r;        // pi/2 - r
  if (x < 0.0) r = 0x1.921fb54442d18p+1 - r;     // pi - r
  return r;
}

// atan2f(y, x) for y >= +0 (the only case computeDirect produces).
__device__ __forceinline__ float cr_atan2f_pos(float y, float x) {
  const float m = fmaxf(y, fabsf(x));
  if (!(m >= 1.0e-30f && m <= 1.0e30f) || !(y >= 0.0f))
    return (float)atan2((double)y, (double)x);  // zeros, NaN, extreme scales, negative y
  return (float)fast_atan2_pos((double)y, (double)x);
}

// sinf / cosf for 0 <= x <= 1.06.
__device__ __forceinline__ void cr_sincosf_small(float xf, float* s, float* c) {
  if (!(xf >= 0.0f && xf <= 1.06f)) {
    double sd, cd;
    sincos((double)xf, &sd, &cd);
    *s = (float)sd;
    *c = (float)cd;
    return;
  }
  const double x = (double)xf;
  const double u = x * x;
  double ps = 0x1.6da3b556355a8p-66;
  ps = fma(ps, u, -0x1.2f44d6ce2f789p-57);
  ps = fma(ps, u, 0x1.952c7095f0e19p-49);
  ps = fma(ps, u, -0x1.ae7f3e6e1a0d8p-41);
  ps = fma(ps, u, 0x1.6124613a84477p-33);
  ps = fma(ps, u, -0x1.ae64567f544cbp-26);
  ps = fma(ps, u, 0x1.71de3a556c734p-19);
  ps = fma(ps, u, -0x1.a01a01a01a01ap-13);
  ps = fma(ps, u, 0x1.1111111111111p-7);
  ps = fma(ps, u, -0x1.5555555555555p-3);
  double pc = -0x1.0a2a7baa1e83cp-70;
  pc = fma(pc, u, 0x1.e53c3c88c84eap-62);
  pc = fma(pc, u, -0x1.682781f22e5c2p-53);
  pc = fma(pc, u, 0x1.ae7f3e6fcf4bap-45);
  pc = fma(pc, u, -0x1.93974a8c0618cp-37);
  pc = fma(pc, u, 0x1.1eed8eff8d890p-29);
  pc = fma(pc, u, -0x1.27e4fb7789f5cp-22);
  pc = fma(pc, u, 0x1.a01a01a01a01ap-16);
  pc = fma(pc, u, -0x1.6c16c16c16c17p-10);
  pc = fma(pc, u, 0x1.5555555555555p-5);
  *s = (float)fma(x * u, ps, x);
  *c = (float)fma(u * u, pc, fma(-0.5, u, 1.0));
}

}  // namespace pcseg
