"""nlls_model_normal_eq_static (include/nlls_b200.h): the fused normal-equations kernel with its scratch — the two [B,N]
model-value ping-pong buffers — in a static __device__ array of the library, for callers that cannot allocate device memory.
No counterpart in the reference; the check is against the existing entry point
nlls_model_normal_eq with caller-owned buffers, through the C ABI.

Collected last on purpose (file name): a failure here hides nothing after it.

Tolerance: the two entry points run the SAME kernel on the same inputs, and the kernel's summation order is fixed, so the outputs
are compared bit for bit (the tolerance the repo's own same-kernel tests use, e.g. test_gpu_retire's run-to-run check)."""
import os
import sys

import pytest
import torch

from helpers import DT

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
from torchimize_b200 import _lib                                   # noqa: E402

pytestmark = pytest.mark.gpu
DEV = 'cuda'
LIMIT = 1048576                                                    # NLLS_STATIC_SCRATCH_ELEMS


def _batch(B, C, N, seed, dt):
    from bench import make_batch
    p0, t, y = make_batch(B, C, N, seed, torch.device(DEV), dt)
    return p0.contiguous(), t.contiguous(), y.contiguous(), torch.linspace(1.0, 2.0, C, dtype=dt, device=DEV)


def _two_iterations(entry, p0, t, y, w, own_buffers):
    """iterations 0 and 1 of one batch through `entry`: the second call reads the model values the first one stored"""
    B, C, N = y.shape
    dt = y.dtype
    ctrl = _lib.new_ctrl(DEV)
    m = torch.full((2, B, N), float('nan'), dtype=dt, device=DEV) if own_buffers else None
    outs = []
    for it, p in enumerate((p0, (p0 * 1.01).contiguous())):
        g = torch.full((B, 4), float('nan'), dtype=dt, device=DEV)
        H = torch.full((B, 4, 4), float('nan'), dtype=dt, device=DEV)
        st = torch.full((B, 4), float('nan'), dtype=torch.float64, device=DEV)
        ctrl.view(torch.int32)[1] = it
        head = (_lib.MODEL_EMG, _lib.dtype_code(p), _lib.ptr(p), 4, _lib.ptr(t), _lib.ptr(y), C * N, N, _lib.ptr(w), C, B, C, N)
        bufs = (_lib.ptr(m[0]), _lib.ptr(m[1])) if own_buffers else ()
        _lib.call(entry, *head, *bufs, _lib.ptr(g), _lib.ptr(H), _lib.ptr(st), _lib.ptr(ctrl), _lib.stream_ptr())
        torch.cuda.synchronize()
        outs.append((g, H, st))
    return outs


def _same(a, b):
    return bool(((a == b) | (torch.isnan(a) & torch.isnan(b))).all())


@pytest.mark.parametrize('dn', ['f64', 'f32'])
def test_static_scratch_entry_point_matches_the_caller_buffer_entry_point(dn):
    dt = DT[dn]
    # the third reuses the first shape with other data: whatever the array still holds must not leak in; the last one sits exactly
    # on the limit (B * N == NLLS_STATIC_SCRATCH_ELEMS)
    shapes = [(257, 3, 1024, 11), (64, 5, 1000, 12), (257, 3, 1024, 13), (LIMIT // 1024, 2, 1024, 14)]
    for B, C, N, seed in shapes:
        p0, t, y, w = _batch(B, C, N, seed, dt)
        want = _two_iterations('nlls_model_normal_eq', p0, t, y, w, own_buffers=True)
        got = _two_iterations('nlls_model_normal_eq_static', p0, t, y, w, own_buffers=False)
        again = _two_iterations('nlls_model_normal_eq_static', p0, t, y, w, own_buffers=False)     # two calls in a row, twice
        for it in range(2):
            for name, a, b, c in zip(('g', 'H', 'stats'), want[it], got[it], again[it]):
                assert not bool(torch.isnan(a).any()), f'{name}: the existing entry point left NaN fill in place'
                assert _same(a, b), f'{dn} B={B} C={C} N={N} iteration {it}: {name} differs, max |d| {(a - b).abs().max():.3e}'
                assert _same(a, c), f'{dn} B={B} C={C} N={N} iteration {it} (repeat): {name} differs'
        # iteration 1's fcon sum really came from the stored model values of iteration 0
        assert bool((want[1][2][:, 1] > 0).all()) and _same(want[1][2][:, 1], got[1][2][:, 1])


def test_static_scratch_entry_point_allocates_nothing():
    dt = torch.float64
    p0, t, y, w = _batch(300, 3, 1024, 21, dt)
    _two_iterations('nlls_model_normal_eq_static', p0, t, y, w, own_buffers=False)        # modules loaded, address looked up
    B, C, N = y.shape
    ctrls = [_lib.new_ctrl(DEV) for _ in range(4)]           # every torch-side tensor exists before the measured calls
    for it, c in enumerate(ctrls):
        c.view(torch.int32)[1] = it
    g = torch.empty(B, 4, dtype=dt, device=DEV)
    H = torch.empty(B, 4, 4, dtype=dt, device=DEV)
    st = torch.empty(B, 4, dtype=torch.float64, device=DEV)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    held0 = torch.cuda.memory_reserved()
    for ctrl in ctrls:
        _lib.call('nlls_model_normal_eq_static', _lib.MODEL_EMG, _lib.dtype_code(p0), _lib.ptr(p0), 4, _lib.ptr(t), _lib.ptr(y),
                  C * N, N, _lib.ptr(w), C, B, C, N, _lib.ptr(g), _lib.ptr(H), _lib.ptr(st), _lib.ptr(ctrl), _lib.stream_ptr())
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_reserved() == held0
    assert free1 >= free0, f'device memory in use grew by {free0 - free1} bytes across calls of the no-allocation entry point'


def test_static_scratch_entry_point_rejects_sizes_beyond_the_limit():
    dt = torch.float64
    p0, t, y, w = _batch(8, 2, 1024, 31, dt)
    B, C, N = y.shape
    g = torch.empty(B, 4, dtype=dt, device=DEV)
    H = torch.empty(B, 4, 4, dtype=dt, device=DEV)
    st = torch.empty(B, 4, dtype=torch.float64, device=DEV)
    ctrl = _lib.new_ctrl(DEV)

    def call(B_claimed):
        _lib.call('nlls_model_normal_eq_static', _lib.MODEL_EMG, _lib.dtype_code(p0), _lib.ptr(p0), 4, _lib.ptr(t), _lib.ptr(y),
                  C * N, N, _lib.ptr(w), C, B_claimed, C, N, _lib.ptr(g), _lib.ptr(H), _lib.ptr(st), _lib.ptr(ctrl),
                  _lib.stream_ptr())

    with pytest.raises(_lib.NllsError, match='NLLS_STATIC_SCRATCH_ELEMS'):       # rejected on the host: nothing is launched,
        call(LIMIT // N + 1)                                                     # the (small) tensors are never touched
    with pytest.raises(_lib.NllsError, match='NLLS_STATIC_SCRATCH_ELEMS'):
        call(2 ** 40)
    call(B)                                                                      # at a legal size the same tensors go through
    torch.cuda.synchronize()
    assert bool(torch.isfinite(g).all())
