"""CPU: the C-ABI library loads, exports every symbol include/nlls_b200.h declares, and rejects
bad arguments before touching the device (no compute calls here)."""
import ctypes
import os
import re

import pytest
import torch

from conftest import ROOT
from torchimize_b200 import _lib


def _declared():
    src = open(os.path.join(ROOT, 'include', 'nlls_b200.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(nlls_[a-z0-9_]+)\s*\(', src)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = _declared()
    assert len(names) >= 14
    for n in names:
        assert hasattr(lib, n), f'{n} declared in nlls_b200.h but not exported'
    # the ctypes table binds exactly the declared functions (+ last_error)
    assert set(_lib.SIGNATURES) | {'nlls_last_error'} == set(names)


def test_abi_version_and_ctrl_size():
    lib = _lib.load()
    assert lib.nlls_abi_version() == 2
    hdr = open(os.path.join(ROOT, 'include', 'nlls_b200.h')).read()
    assert '#define NLLS_B200_ABI_VERSION 2' in hdr

    class Ctrl(ctypes.Structure):
        _fields_ = [('done', ctypes.c_int), ('iters', ctypes.c_int), ('singular', ctypes.c_int),
                    ('error', ctypes.c_int), ('red', ctypes.c_double * 4)]
    assert ctypes.sizeof(Ctrl) == _lib.CTRL_BYTES


def test_argument_validation_needs_no_device():
    lib = _lib.load()
    rc = lib.nlls_normal_eq(7, None, None, None, None, 1, 1, 1, 1, 1, None, None, None, None, None)
    assert rc == -1 and b'dtype' in lib.nlls_last_error()
    rc = lib.nlls_normal_eq(_lib.F64, None, None, None, None, 1, 4, 1, 8, 33, None, None, None, None, None)
    assert rc == -2
    rc = lib.nlls_normal_eq(_lib.F64, None, None, None, None, 1, 4, 1, 8, 4, None, None, None, None, None)
    assert rc == -3
    rc = lib.nlls_model_normal_eq(9, _lib.F64, 8, 4, 8, 8, 1, 0, 8, 1, 1, 1, 4, None, None, 8, 8, 8, None, None, None)
    assert rc == -5
    assert lib.nlls_iter_end(None, 0, 0.0, None, None) == -3


def test_product_refuses_cpu_tensors():
    """No CPU fallback: the drop-ins raise instead of silently computing elsewhere."""
    from torchimize_b200.functions import lsq_lma_parallel, lsq_gna_parallel, gradient_descent_parallel
    p = torch.zeros(3, 4, dtype=torch.float64)
    for fn in (lsq_lma_parallel, lsq_gna_parallel, gradient_descent_parallel):
        with pytest.raises(_lib.NllsError):
            fn(p, lambda q: q[:, None, :], lambda q: q)
    with pytest.raises(AssertionError):
        lsq_lma_parallel(torch.zeros(4), lambda q: q)          # reference's only validation (lma:74)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, 'torchimize_b200')
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                txt = open(os.path.join(dp, f)).read()
                assert 'oracle' not in txt.replace('# oracle', ''), f'{f} mentions the oracle'
