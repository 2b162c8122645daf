#!/usr/bin/env python
"""Generates the double-precision polynomial coefficients of csrc/cr_trig.cuh
(near-minimax Chebyshev fits computed with mpmath at 60 digits).

  atan(t)  = t + t^3 * A(t^2)            |t| <= tan(pi/32) * 1.02
  sin(x)   = x + x^3 * S(x^2)            0 <= x <= 1.06
  cos(x)   = 1 - x^2/2 + x^4 * C(x^2)    0 <= x <= 1.06
  acos: handled through atan2 in the kernel (one call per boundary cell).
"""
import mpmath as mp

mp.mp.dps = 60


def cheb_fit(f, a, b, n):
    """Coefficients (monomial basis in u, ascending) of the degree n-1 Chebyshev
    interpolant of f(u) on [a, b]."""
    nodes = [mp.cos(mp.pi * (2 * k + 1) / (2 * n)) for k in range(n)]
    us = [(a + b) / 2 + (b - a) / 2 * x for x in nodes]
    fs = [f(u) for u in us]
    # solve Vandermonde in high precision (n is small)
    V = mp.matrix(n, n)
    for i in range(n):
        for j in range(n):
            V[i, j] = us[i] ** j
    c = mp.lu_solve(V, mp.matrix(fs))
    return [c[i] for i in range(n)]


def maxerr(f, poly, a, b, samples=4001):
    worst = mp.mpf(0)
    for i in range(samples):
        u = a + (b - a) * i / (samples - 1)
        p = mp.mpf(0)
        for c in reversed(poly):
            p = p * u + mp.mpf(float(c))
        worst = max(worst, abs(p - f(u)))
    return worst


def emit(name, coeffs):
    print("// %s" % name)
    for i, c in enumerate(coeffs):
        print("  %s,  // u^%d" % (float(c).hex(), i))


if __name__ == "__main__":
    tmax = mp.tan(mp.pi / 32) * mp.mpf("1.02")

    def fa(u):  # (atan(t) - t) / t^3, u = t^2
        if u == 0:
            return -mp.mpf(1) / 3
        t = mp.sqrt(u)
        return (mp.atan(t) - t) / t ** 3
    A = cheb_fit(fa, mp.mpf(0), tmax ** 2, 9)
    print("atan err (abs, in units of t^3):", mp.nstr(maxerr(fa, A, mp.mpf(0), tmax ** 2), 5))
    emit("ATAN", A)

    xmax = mp.mpf("1.06")

    def fs(u):
        if u == 0:
            return -mp.mpf(1) / 6
        x = mp.sqrt(u)
        return (mp.sin(x) - x) / x ** 3
    S = cheb_fit(fs, mp.mpf(0), xmax ** 2, 10)
    print("sin err:", mp.nstr(maxerr(fs, S, mp.mpf(0), xmax ** 2), 5))
    emit("SIN", S)

    def fc(u):
        if u == 0:
            return mp.mpf(1) / 24
        x = mp.sqrt(u)
        return (mp.cos(x) - 1 + u / 2) / u ** 2
    Cc = cheb_fit(fc, mp.mpf(0), xmax ** 2, 10)
    print("cos err:", mp.nstr(maxerr(fc, Cc, mp.mpf(0), xmax ** 2), 5))
    emit("COS", Cc)
    print("// tan(k*pi/16), k*pi/16 for k = 0..4; thresholds tan((2j+1)pi/32)")
    for k in range(5):
        print("  %s, %s," % (float(mp.tan(k * mp.pi / 16)).hex(), float(k * mp.pi / 16).hex()))
    for j in range(4):
        print("  %s," % float(mp.tan((2 * j + 1) * mp.pi / 32)).hex())
    print("pi", float(mp.pi).hex(), "pi/2", float(mp.pi / 2).hex())
