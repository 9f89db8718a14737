#!/usr/bin/env python
"""Generates the constants of the fp64 model math in torchimize_b200/csrc/models.cuh (mpmath, 200-bit arithmetic):

  kExp2Tab    T[j] = 2^(j/32) as hi + lo pairs, j = 0..31, and the Cody-Waite split of ln2/32 (hi with 24 trailing zero bits)
  kErfcxPoly  erfcx(x) = exp(x^2) erfc(x) on x in [0, 6] as a degree-20 polynomial in s = 1 - 3 x / (x + 3)  (s in [-1, 1]):
              Chebyshev interpolant converted to the monomial basis; prints the approximation error

tests/test_model_math_cpu.py checks the header's values and the accuracy of the same fma sequence compiled as C.
"""
import math

import mpmath as mp

mp.mp.prec = 200


def exp_table(nt=32):
    rows = []
    for j in range(nt):
        v = mp.mpf(2) ** (mp.mpf(j) / nt)
        hi = float(v)
        rows.append((hi, float(v - mp.mpf(hi))))
    step = mp.log(2) / nt
    m, e = math.frexp(float(step))
    mi = int(m * (1 << 53)) & ~((1 << 24) - 1)
    hi = math.ldexp(mi, e - 53)
    return rows, hi, float(step - mp.mpf(hi)), float(nt / mp.log(2))


def erfcx_poly(deg=20, c=3, xmax=6):
    c, xmax = mp.mpf(c), mp.mpf(xmax)
    rlo, rhi = 1 / (xmax + c), 1 / c
    mid, half = (rlo + rhi) / 2, (rhi - rlo) / 2                # s = (1/(x + c) - mid) / half = 9 / (x + 3) - 2 = 1 - 3 x / (x + 3)
    erfcx = lambda x: mp.exp(x * x) * mp.erfc(x)
    f = lambda s: erfcx(1 / (mid + half * s) - c)
    n = deg + 1
    nodes = [mp.cos(mp.pi * (k + mp.mpf(1) / 2) / n) for k in range(n)]
    fs = [f(t) for t in nodes]
    cheb = [2 / mp.mpf(n) * sum(fs[k] * mp.cos(mp.pi * j * (k + mp.mpf(1) / 2) / n) for k in range(n)) for j in range(n)]
    cheb[0] /= 2
    T = [[mp.mpf(1)], [mp.mpf(0), mp.mpf(1)]]
    for k in range(2, n):
        a = [mp.mpf(0)] + [2 * v for v in T[k - 1]]
        b = T[k - 2] + [mp.mpf(0)] * (len(a) - len(T[k - 2]))
        T.append([x - y for x, y in zip(a, b)])
    mono = [mp.mpf(0)] * n
    for j in range(n):
        for i, v in enumerate(T[j]):
            mono[i] += cheb[j] * v
    worst = mp.mpf(0)
    for i in range(2000):
        x = xmax * mp.mpf(i + 0.5) / 2000
        s = (1 / (x + c) - mid) / half
        acc = mp.mpf(0)
        for v in reversed(mono):
            acc = acc * s + v
        worst = max(worst, abs(acc / erfcx(x) - 1))
    return [float(v) for v in mono], float(worst)


if __name__ == '__main__':
    rows, hi, lo, inv = exp_table()
    print('// kExp2Tab[32]: {hi, lo} of 2^(j/32)')
    for j in range(0, 32, 2):
        print('    ' + ', '.join('{%r, %r}' % r for r in rows[j:j + 2]) + ',')
    print(f'// 32/ln2 = {inv!r}; ln2/32 = {hi!r} + {lo!r}')
    mono, err = erfcx_poly()
    print(f'// kErfcxPoly[21] (max relative approximation error {err:.2e}):')
    for j in range(0, 21, 4):
        print('    ' + ', '.join(repr(v) for v in mono[j:j + 4]) + ',')
