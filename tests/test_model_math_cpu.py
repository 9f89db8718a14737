"""fp64 model math of csrc/models.cuh (exp_nonpos, one_plus_erf_core) checked on the CPU: the constants are read out of the
CUDA header, the SAME fma sequence is compiled as plain C (gcc, -ffp-contract=off, libm fma = one rounding like DFMA) and
compared with long-double libm over the whole argument range.  Pins the tables: a wrong entry or coefficient fails here,
without a GPU.  (The GPU side of the same check: tests/test_gpu_parity.py::test_fp64_exp_nonpos_accuracy and the goldens.)"""
import math
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HDR = os.path.join(ROOT, 'torchimize_b200', 'csrc', 'models.cuh')

C_MODEL = r'''
#include <math.h>
#include <stdio.h>
#include <stdint.h>
#include <string.h>
static const double TAB[32][2] = {%(tab)s};
static const double EP[5] = {%(ep)s};
static const double GP[21] = {%(gp)s};
static double exp_nonpos(double x) {
  const double magic = 6755399441055744.0;
  double t = fma(x, %(inv)s, magic);
  int64_t bits; memcpy(&bits, &t, 8); const int n = (int)(uint32_t)bits;
  t -= magic;
  double r = fma(t, %(lnhi)s, x); r = fma(t, %(lnlo)s, r);
  double q = fma(r, EP[0], EP[1]);
  for (int i = 2; i < 5; ++i) q = fma(q, r, EP[i]);
  const double p = fma(q, r * r, r);
  const double m = fma(TAB[n & 31][0], p, TAB[n & 31][1]) + TAB[n & 31][0];
  const double e = ldexp(m, n >> 5);
  return x >= -707.0 ? e : (x < -707.0 ? 0.0 : x);
}
static double rcp(double w) {                      /* MUFU.RCP64H (>= 20 good bits) + two Newton steps */
  double r = (double)(float)(1.0 / w) * (1.0 + 4e-7);
  double e = fma(-w, r, 1.0); r = fma(r, e, r);
  e = fma(-w, r, 1.0); r = fma(r, e, r);
  return r;
}
static double one_plus_erf(double x) {
  if (fabs(x) >= 6.0) return x > 0.0 ? 2.0 : 0.0;
  const double ax = fabs(x), E = exp_nonpos(-(x * x));
  const double r = rcp(ax + 3.0);
  const double s = fma(-3.0, ax * r, 1.0);
  double g = GP[20];
  for (int i = 19; i >= 0; --i) g = fma(g, s, GP[i]);
  const double c = E * g;
  return x < 0.0 ? c : 2.0 - c;
}
int main(void) {
  double we = 0, wp = 0, wn = 0;
  const long N = 2000000;
  for (long i = 0; i < N; ++i) {
    const double x = -706.9 * ((double)i + 0.37) / N;
    const long double ref = expl((long double)x);
    const double d = (double)ref, ulp = nextafter(d, INFINITY) - d;
    const double err = (double)(fabsl((long double)exp_nonpos(x) - ref) / ulp);
    if (err > we) we = err;
  }
  for (long i = 0; i < N; ++i) {
    const double x = -6.5 + 13.0 * ((double)i + 0.31) / N;
    const long double ref = x < 0 ? erfcl(-(long double)x) : 1.0L + erfl((long double)x);
    const double got = one_plus_erf(x);
    const long double err = fabsl((long double)got - ref);
    if (x >= 0) { const double d = (double)ref, ulp = nextafter(d, INFINITY) - d; if ((double)(err / ulp) > wp) wp = (double)(err / ulp); }
    else if ((double)err > wn) wn = (double)err;
  }
  printf("%%.4f %%.4f %%.4e %%.17g %%.17g %%.17g\n", we, wp, wn, exp_nonpos(-800.0), exp_nonpos(0.0), one_plus_erf(NAN));
  return 0;
}
'''


def _array(src, name):
    m = re.search(name + r'\[\d+\]\s*=\s*\{(.*?)\};', src, re.S)
    assert m, name
    return [float(v) for v in re.findall(r'-?\d+\.?\d*(?:[eE][-+]?\d+)?', m.group(1))]


def test_tables_match_their_definition():
    src = open(HDR).read()
    tab = _array(src, 'kExp2Tab')
    assert len(tab) == 64
    for j in range(32):
        hi, lo = tab[2 * j], tab[2 * j + 1]
        assert hi == 2.0 ** (j / 32) or abs(hi - 2.0 ** (j / 32)) <= math.ulp(hi), j      # hi: 2^(j/32) rounded
        assert abs(lo) <= math.ulp(hi) / 2 * 1.0000001, j                                  # lo: the rounding remainder
    assert _array(src, 'kExpPoly') == [1 / 720, 1 / 120, 1 / 24, 1 / 6, 0.5]
    assert len(_array(src, 'kErfcxPoly')) == 21


@pytest.mark.skipif(shutil.which('gcc') is None, reason='needs gcc')
def test_c_model_of_the_fp64_model_math_is_accurate(tmp_path):
    src = open(HDR).read()
    tab = _array(src, 'kExp2Tab')
    body = src[src.index('__device__ __forceinline__ double exp_nonpos(double x)'):]
    inv = re.search(r'fma\(x, ([0-9.eE+-]+), magic\)', body).group(1)
    lnhi = re.search(r'double r = fma\(t, (-[0-9.eE+-]+), x\)', body).group(1)
    lnlo = re.search(r'r = fma\(t, (-[0-9.eE+-]+), r\)', body).group(1)
    c = C_MODEL % dict(tab=', '.join('{%r, %r}' % (tab[2 * j], tab[2 * j + 1]) for j in range(32)),
                       ep=', '.join(repr(v) for v in _array(src, 'kExpPoly')),
                       gp=', '.join(repr(v) for v in _array(src, 'kErfcxPoly')), inv=inv, lnhi=lnhi, lnlo=lnlo)
    cfile, exe = tmp_path / 'mm.c', tmp_path / 'mm'
    cfile.write_text(c)
    subprocess.run(['gcc', '-O2', '-ffp-contract=off', '-o', str(exe), str(cfile), '-lm'], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()
    exp_ulp, erf_pos_ulp, erf_neg_abs = float(out[0]), float(out[1]), float(out[2])
    assert exp_ulp <= 0.6, f'exp_nonpos: {exp_ulp} ulp'                       # 0.55 measured
    assert erf_pos_ulp <= 2.0, f'1 + erf(x), x >= 0: {erf_pos_ulp} ulp'       # 1.6 measured (CUDA erf: 2 ulp documented)
    assert erf_neg_abs <= 3e-16, f'1 + erf(x), x < 0: {erf_neg_abs} absolute'  # 2.6e-16 measured
    assert float(out[3]) == 0.0 and float(out[4]) == 1.0 and math.isnan(float(out[5]))
