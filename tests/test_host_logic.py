"""Host-side logic that needs no GPU: the forward-mode numeric-Jacobian fallback (reference: functions/jacobian.py:36-57,
here P JVP passes) against the oracle's reverse-mode formulation, incl. the fallback for callables forward-mode AD
cannot trace; callable binding (lma:62-72)."""
import torch

from helpers import load
from oracle import nlls_oracle as O, models as M
from torchimize_b200.functions.jacobian import batch_jacobian_approx_t, jacobian_approx_t
from torchimize_b200.functions.parallel._engine import bind_callables


def _emg():
    gold = load('synthetic')
    p, t, y = (torch.as_tensor(gold[k]) for k in ('p0', 't', 'y'))
    return p, (lambda q: M.emg_cost(q, t, y))


def test_forward_mode_batch_jacobian_matches_reverse_mode_and_analytic():
    p, fun = _emg()
    J = batch_jacobian_approx_t(p, fun)
    J_rev = O.autograd_batch_jacobian(p, fun)
    gold = load('synthetic')
    J_an = M.emg_jac(p, torch.as_tensor(gold['t']), torch.as_tensor(gold['y']))
    assert J.shape == J_rev.shape == J_an.shape
    scale = J_an.abs().max()
    assert (J - J_rev).abs().max() <= 1e-12 * scale
    assert (J - J_an).abs().max() <= 1e-12 * scale


def test_fallback_to_reverse_mode_when_forward_mode_cannot_trace():
    class Square(torch.autograd.Function):                          # no jvp rule: forward mode raises
        @staticmethod
        def forward(ctx, x):
            ctx.save_for_backward(x)
            return x * x

        @staticmethod
        def backward(ctx, g):
            (x,) = ctx.saved_tensors
            return 2 * x * g

    p = torch.tensor([[1., 2.], [3., 4.], [5., 6.]], dtype=torch.float64)
    fun = lambda q: (Square.apply(q)[:, None, :] * torch.ones(1, 2, 1, dtype=q.dtype)).contiguous()   # [B, C=2, N=2]
    J = batch_jacobian_approx_t(p, fun)
    want = torch.zeros(3, 2, 2, 2, dtype=torch.float64)
    for b in range(3):
        for n in range(2):
            want[b, :, n, n] = 2 * p[b, n]
    assert torch.equal(J, want)


def test_single_problem_jacobian_and_binding():
    p = torch.tensor([1., 2., 3.], dtype=torch.float64)
    J = jacobian_approx_t(p, lambda q: torch.stack([q[0] * q[1], q[2] ** 2]))
    assert torch.allclose(J, torch.tensor([[2., 1., 0.], [0., 0., 6.]], dtype=torch.float64))
    calls = []
    f = lambda q, a, b: calls.append(('f', a, b)) or q
    j = lambda q, a, b: calls.append(('j', a, b)) or q
    fun, jac = bind_callables(f, j, (1, 2))
    fun(p), jac(p)
    assert calls == [('f', 1, 2), ('j', 1, 2)]
    fun0, _ = bind_callables(lambda q: q + 1, None, ())              # args empty: function is used as is (lma:62-64)
    assert torch.equal(fun0(p), p + 1)


def test_track_struct_matches_the_header_and_is_inserted_before_the_stream_argument():
    """nlls_track (include/nlls_b200.h): 3 pointers + 2 ints + a 64-bit count = 40 bytes; every TRACKED entry point has the
    struct pointer as its last argument before `stream`."""
    import ctypes
    import os
    import re
    from conftest import ROOT
    from torchimize_b200 import _lib
    assert ctypes.sizeof(_lib.Track) == 40
    assert [f[0] for f in _lib.Track._fields_] == ['pstat', 'active', 'count', 'retire', 'reserved', 'B']
    hdr = open(os.path.join(ROOT, 'include', 'nlls_b200.h')).read()
    decl = {m.group(1): m.group(2) for m in re.finditer(r'\nint (nlls_\w+)\(([^;]*?)\);', hdr, re.S)}
    with_track = {n for n, a in decl.items() if 'const nlls_track* track' in a}
    assert with_track == set(_lib.TRACKED)
    for n in with_track:
        args = [a.strip() for a in re.sub(r'/\*.*?\*/', '', decl[n], flags=re.S).split(',')]
        assert args[-1].startswith('void* stream') and 'nlls_track' in args[-2], n
        assert _lib.SIGNATURES[n][-2] is _lib._tp


def test_bound_model_callables_are_recognised_and_validated_without_a_gpu():
    import pytest
    from torchimize_b200 import _lib
    from torchimize_b200.models import EMGModel, bound_model_of, fused_model_of
    emg = EMGModel()
    x, y = torch.zeros(8), torch.zeros(3, 8)
    bm = emg.bind(x, y)
    assert bound_model_of(bm.cost, bm.jac, 'gn') == (emg, (x, y))
    assert bound_model_of(bm.cost, None, 'gd') == (emg, (x, y))            # jac_function=None: the model's analytic Jacobian
    assert bound_model_of(bm.cost, emg.bind(x, y).jac) is None                # callables of two different bindings
    assert bound_model_of(emg.cost, emg.jac) is None and fused_model_of(bm.cost, bm.jac) is None
    assert bound_model_of(lambda p: p, None) is None
    with pytest.raises(_lib.NllsError, match='CUDA'):                         # no CPU path, with or without binding
        bm.cost(torch.zeros(3, 4))


def test_solver_mode_is_validated_before_anything_touches_the_device():
    import pytest
    from torchimize_b200 import _lib
    from torchimize_b200.functions import lsq_gna_parallel
    with pytest.raises(_lib.NllsError, match='CUDA'):
        lsq_gna_parallel(torch.zeros(2, 4), lambda p: p, mode='retire')
