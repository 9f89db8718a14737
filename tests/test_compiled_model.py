"""Compiled user models (torchimize_b200.models.compile_model): a CUDA device functor is
compiled into a plugin library exporting the nlls_model_* entry points; the drop-in solvers then run the fused
kernels for it.  CPU part: the plugin builds (nvcc cross-compiles) and exports the entry points; GPU part: parity
with the oracle on the same seeded inputs and bit-identity with the built-in EMG kernels."""
import ctypes
import os

import pytest
import torch

from conftest import ROOT
from torchimize_b200 import _lib

EXAMPLE = os.path.join(ROOT, 'torchimize_b200', 'user_models', 'gauss_offset.cuh')
EMG_CLONE = 'namespace nlls { template <typename T> struct UserModel : EmgModel<T> {}; }\n'
DEV = 'cuda:0'


def test_plugin_builds_and_exports_the_model_entry_points():
    from torchimize_b200.models import compile_model_file
    m = compile_model_file(EXAMPLE)
    assert m.n_params == (4,) and m.model_id == _lib.MODEL_USER
    so = _lib.build_plugin(open(EXAMPLE).read(), 'gauss_offset')
    assert os.path.exists(so) and os.path.dirname(os.path.dirname(so)) == _lib.JIT_DIR       # in-tree, cached
    lib = ctypes.CDLL(so)
    for name in _lib.PLUGIN_ENTRY_POINTS + ('nlls_abi_version', 'nlls_last_error', 'nlls_user_model_params'):
        assert hasattr(lib, name), name
    assert lib.nlls_abi_version() == _lib.load().nlls_abi_version()


def test_broken_functor_raises_with_the_compiler_message():
    from torchimize_b200.models import compile_model
    with pytest.raises(_lib.NllsError, match='error'):
        compile_model('namespace nlls { template <typename T> struct UserModel { static constexpr int P = 2; }; }', 'broken')


def _gauss_batch(B, C, N, seed, dt):
    g = torch.Generator().manual_seed(seed)
    x = torch.linspace(-4, 6, N, dtype=torch.float64)
    gt = torch.tensor([3., 1., .8, .5], dtype=torch.float64) * (1 + .1 * (torch.rand(B, 4, generator=g, dtype=torch.float64) - .5))
    p0 = gt * (1 + .2 * (torch.rand(B, 4, generator=g, dtype=torch.float64) - .5))
    clean = gt[:, 0, None] * torch.exp(-(x - gt[:, 1, None]) ** 2 / (2 * gt[:, 2, None] ** 2)) + gt[:, 3, None]
    y = clean[:, None, :] + .01 * torch.randn(B, C, N, generator=g, dtype=torch.float64)
    return p0.to(dt), x.to(dt), y.to(dt), gt


@pytest.mark.gpu
@pytest.mark.parametrize('solver', ['lma', 'gna', 'gda'])
@pytest.mark.parametrize('shape', [(24, 3, 512), (40, 1, 1000), (5, 2, 130)])
def test_compiled_gauss_offset_matches_the_oracle_fp64(solver, shape):
    from oracle import nlls_oracle as O, models as M
    from torchimize_b200.functions import lsq_lma_parallel, lsq_gna_parallel, gradient_descent_parallel
    from torchimize_b200.models import compile_model_file
    from helpers import compare_plist
    B, C, N = shape
    model = compile_model_file(EXAMPLE)
    p0, x, y, gt = _gauss_batch(B, C, N, 11, torch.float64)
    w = torch.linspace(1., 1.5, C, dtype=torch.float64)
    kw = dict(ftol=1e-8, ptol=1e-8, gtol=1e-8, max_iter=40)
    ours = dict(lma=lsq_lma_parallel, gna=lsq_gna_parallel, gda=gradient_descent_parallel)[solver]
    ref = getattr(O, solver)(p0.clone(), M.gauss_offset_cost, M.gauss_offset_jac, args=(x, y), wvec=w, **kw)
    for fused in (True, False):                                   # fused kernels of the plugin / tensor path through its eval kernel
        p = p0.clone().to(DEV)
        got = ours(p, model.cost, model.jac, args=(x.to(DEV), y.to(DEV)), wvec=w.to(DEV), fused=fused, **kw)
        compare_plist(got, torch.stack(ref.p_list), 'f64', f'compiled gauss_offset {solver} fused={fused}')
    if solver != 'gda':
        assert torch.allclose(p.cpu(), gt, atol=5e-2)


@pytest.mark.gpu
@pytest.mark.parametrize('dn', ['f64', 'f32'])
def test_compiled_emg_clone_is_bit_identical_to_the_builtin_kernels(dn):
    from torchimize_b200.functions import lsq_lma_parallel, lsq_gna_parallel, gradient_descent_parallel
    from torchimize_b200.models import EMGModel, compile_model
    from bench import make_batch
    dt = torch.float64 if dn == 'f64' else torch.float32
    clone, builtin = compile_model(EMG_CLONE, 'emg_clone'), EMGModel()
    p0, t, y = make_batch(64, 3, 1024, 7, 'cpu', dt)
    w = torch.ones(3, dtype=dt, device=DEV)
    for solver, kw in ((lsq_lma_parallel, dict(meth='marq')), (lsq_gna_parallel, {}), (gradient_descent_parallel, {})):
        outs = []
        for m in (clone, builtin):
            p = p0.clone().to(DEV)
            outs.append(torch.stack(solver(p, m.cost, m.jac, args=(t.to(DEV), y.to(DEV)), wvec=w, max_iter=25, **kw)))
        assert outs[0].shape == outs[1].shape
        assert bool(((outs[0] == outs[1]) | (torch.isnan(outs[0]) & torch.isnan(outs[1]))).all())


@pytest.mark.gpu
def test_compiled_wide_model_runs_the_tensor_core_fused_kernel_like_the_builtin():
    """A compiled functor with P = 8 (two EMG peaks) takes k_model_normal_eq_mma in the plugin: bit-identical to MultiEMGModel(2)."""
    from torchimize_b200.functions import lsq_lma_parallel, lsq_gna_parallel, gradient_descent_parallel
    from torchimize_b200.models import MultiEMGModel, compile_model
    from test_gpu_multi_emg import multi_emg_batch
    clone = compile_model('namespace nlls { template <typename T> struct UserModel : MultiEmgModel<T, 2> {}; }\n', 'emg2_clone')
    assert clone.n_params == (8,) and clone.fused_solvers == ('gn', 'lm', 'gd')
    builtin = MultiEMGModel(2)
    p0, t, y, _ = multi_emg_batch(16, 2, 1024, 4, torch.float64, C=2)
    w = torch.ones(2, dtype=torch.float64, device=DEV)
    for solver, kw in ((lsq_lma_parallel, dict(meth='lev')), (lsq_gna_parallel, {}), (gradient_descent_parallel, {})):
        outs = []
        for m in (clone, builtin):
            p = p0.clone().to(DEV)
            outs.append(torch.stack(solver(p, m.cost, m.jac, args=(t.to(DEV), y.to(DEV)), wvec=w, max_iter=20, **kw)))
        assert outs[0].shape == outs[1].shape and bool((outs[0] == outs[1]).all())


@pytest.mark.gpu
def test_compiled_model_rejects_a_wrong_parameter_count():
    from torchimize_b200.functions import lsq_gna_parallel
    from torchimize_b200.models import compile_model_file
    model = compile_model_file(EXAMPLE)
    p0, x, y, _ = _gauss_batch(4, 1, 64, 3, torch.float64)
    with pytest.raises(_lib.NllsError):
        lsq_gna_parallel(p0[:, :3].contiguous().to(DEV), model.cost, model.jac, args=(x.to(DEV), y.to(DEV)))
