"""Built-in fit models with fused CUDA residual+Jacobian kernels (K4).

A model object supplies the two callables the reference's solvers expect —
``function(p, x, y) -> [B,C,N]`` and ``jac_function(p, x, y) -> [B,C,N,P]``
(README.rst:97-98) — backed by ``nlls_model_eval``.  When BOTH callables handed to a solver are
the bound ``cost`` / ``jac`` methods of the same model object, the solver skips materialising
J altogether and runs the fused normal-equation kernels instead (``nlls_model_normal_eq*``).

    emg = EMGModel(n_costs=2)
    coeffs = lsq_lma_parallel(p, emg.cost, emg.jac, args=(t, y), wvec=torch.ones(2, ...))

Models (definitions follow the reference's tests):
  EMGModel            p=[alpha, mu, sigma, eta]   tests/emg.py:19-49, tests/unit_test_parallel.py:81-118
  SkewGaussModel      p=[norm, mean, sigm, skew(, offset)]   tests/unit_test_skewed_gaussian.py:43-51
Cost is the squared error (y - m)^2 used as residual, as in tests/unit_test_parallel.py:40,55-56.
"""
import torch

from . import _lib


class FusedModel:
    """Base: data layout handling shared by all built-in models."""
    model_id = -1
    n_params = ()
    lib = None            # plugin library of a compiled user model; None = libnlls_b200.so
    fused_solvers = ('gn', 'lm', 'lm2', 'gd')   # solvers with fused kernels for this model; the others take the tensor path

    def __init__(self, n_costs=None):
        # n_costs: number of cost channels when y is [B,N] (one data row shared by all costs, as in
        # the reference's multi_cost_batch); ignored when y is [B,C,N].
        self.n_costs = n_costs

    # -- layout -----------------------------------------------------------------------------
    def layout(self, p, x, y):
        """-> (x, y, y_bstride, y_cstride, C, N) with x, y contiguous in p's dtype on p's device."""
        _lib.require_cuda(p, x, y)
        if p.dim() != 2 or p.shape[1] not in self.n_params:
            raise _lib.NllsError(f'{type(self).__name__}: p must be [B,P] with P in {self.n_params}, got {tuple(p.shape)}')
        x = x.to(dtype=p.dtype).contiguous()
        y = y.to(dtype=p.dtype).contiguous()
        N = x.shape[0]
        if y.dim() == 2:
            C = 1 if self.n_costs is None else int(self.n_costs)
            ybs, ycs = N, 0
        elif y.dim() == 3:
            C = y.shape[1]
            ybs, ycs = C * N, N
        else:
            raise _lib.NllsError(f'y must be [B,N] or [B,C,N], got {tuple(y.shape)}')
        if y.shape[0] != p.shape[0] or y.shape[-1] != N or x.dim() != 1:
            raise _lib.NllsError(f'shape mismatch: p {tuple(p.shape)}, x {tuple(x.shape)}, y {tuple(y.shape)}')
        return x, y, ybs, ycs, C, N

    # -- the reference-style callables ------------------------------------------------------
    def _eval(self, p, x, y, want_jac):
        x, y, ybs, ycs, C, N = self.layout(p, x, y)
        pc = p.contiguous()
        B, P = pc.shape
        f = torch.empty((B, C, N), dtype=p.dtype, device=p.device)
        J = torch.empty((B, C, N, P), dtype=p.dtype, device=p.device) if want_jac else None
        _lib.call('nlls_model_eval', self.model_id, _lib.dtype_code(p), _lib.ptr(pc), P, _lib.ptr(x), _lib.ptr(y),
                  ybs, ycs, B, C, N, _lib.ptr(f), _lib.ptr(J), None, _lib.stream_ptr(), lib=self.lib)
        return f, J

    def cost(self, p, x, y):
        """function(p, x, y) -> residuals [B,C,N] = (y - m)^2."""
        return self._eval(p, x, y, False)[0]

    def jac(self, p, x, y):
        """jac_function(p, x, y) -> Jacobian [B,C,N,P] of ``cost``."""
        return self._eval(p, x, y, True)[1]

    def cost_and_jac(self, p, x, y):
        return self._eval(p, x, y, True)

    def bind(self, x, y):
        """The model with its data bound: ``.cost(p)`` / ``.jac(p)`` — the one-argument callables that
        lsq_gna_parallel_plain / gradient_descent_parallel_plain / newton_step_parallel take (gna:94-101, gda:90-97,
        newton_parallel.py:21-26).  The solvers recognise them and run the fused kernels (no Jacobian tensor)."""
        return BoundModel(self, x, y)


class BoundModel:
    """FusedModel + data (see FusedModel.bind)."""

    def __init__(self, model, x, y):
        self.model, self.args = model, (x, y)

    def cost(self, p):
        return self.model.cost(p, *self.args)

    def jac(self, p):
        return self.model.jac(p, *self.args)


class EMGModel(FusedModel):
    model_id = _lib.MODEL_EMG
    n_params = (4,)


class SkewGaussModel(FusedModel):
    model_id = _lib.MODEL_SKEWGAUSS
    n_params = (4, 5)


class MultiEMGModel(FusedModel):
    """Sum of ``n_peaks`` (2..4) EMG peaks, p = [alpha_1, mu_1, sigma_1, eta_1, ...] (P = 4 n_peaks; SURVEY.md §8d, the
    large-P workload of cfg-4).  Gauss-Newton and Levenberg-Marquardt run the tensor-core fused kernel
    (k_model_normal_eq_mma: the model is evaluated once per sample and the per-warp 32 x 16 tile of k rows is reduced
    with fp64 DMMA; J is never materialised), so does gradient_descent_parallel (cost-summed rows); the two-pass LM
    schedule and ``fused=False`` take the tensor path (``nlls_model_eval`` + K1, also on the fp64 tensor cores for P > 8)."""
    model_id = _lib.MODEL_MULTI_EMG
    fused_solvers = ('gn', 'lm', 'gd')

    def __init__(self, n_peaks=4, n_costs=None):
        super().__init__(n_costs)
        if n_peaks not in (2, 3, 4):
            raise _lib.NllsError('MultiEMGModel: n_peaks must be 2, 3 or 4')
        self.n_params = (4 * n_peaks,)


class CompiledModel(FusedModel):
    """A user-supplied CUDA device functor compiled into a plugin library."""
    model_id = _lib.MODEL_USER

    def __init__(self, lib, n_params, name, n_costs=None):
        super().__init__(n_costs)
        self.lib, self.n_params, self.name = lib, (int(n_params),), name
        if int(n_params) > 6:                  # tensor-core fused kernel: no two-pass LM schedule (like MultiEMGModel)
            self.fused_solvers = ('gn', 'lm', 'gd')

    def __repr__(self):
        return f'CompiledModel({self.name!r}, P={self.n_params[0]})'


def compile_model(source, name='user', n_costs=None):
    """Compile a model written as a CUDA device functor and return it as a fused model object.

    ``source`` is CUDA C++ defining ``template <typename T> struct UserModel`` in namespace ``nlls`` with
    ``static constexpr int P`` (<= 16 parameters; above 6 the fp64 tensor-core fused kernel runs GN / LM and GD takes the tensor path), ``load(const T* p)``, ``T value(T x) const`` and
    ``void eval(T x, T& m, T (&k)[P]) const`` with ``k = -dm/dp`` (see torchimize_b200/user_models/gauss_offset.cuh;
    every helper of csrc/models.cuh is in scope).  The cost is the squared error used as residual, as for the
    built-in models.  nvcc compiles csrc/fused_models.cu for that functor (sm_100a only, ~1 min once, cached in
    torchimize_b200/_jit/); the returned object's ``cost`` / ``jac`` go to any of the drop-in solvers, which then
    run the fused kernels of the plugin: no Jacobian tensor, no Python callable inside the iteration.
    """
    so = _lib.build_plugin(source, name)
    lib = _lib.load_plugin(so)
    return CompiledModel(lib, lib.nlls_user_model_params(), name, n_costs)


def compile_model_file(path, name=None, n_costs=None):
    import os
    with open(path) as f:
        return compile_model(f.read(), name or os.path.splitext(os.path.basename(path))[0], n_costs)


def model_owner_of(function, jac_function):
    """The FusedModel whose bound ``cost`` / ``jac`` the two callables are (fused kernels or not), else None: lets the
    tensor path evaluate residuals and Jacobian with ONE nlls_model_eval launch instead of two."""
    owner = getattr(function, '__self__', None)
    if not isinstance(owner, FusedModel) or getattr(function, '__func__', None) is not FusedModel.cost:
        return None
    if getattr(jac_function, '__self__', None) is owner and getattr(jac_function, '__func__', None) is FusedModel.jac:
        return owner
    return None


def fused_model_of(function, jac_function, solver=None):
    """The FusedModel both callables belong to (and that has fused kernels for `solver`), else None (=> tensor path)."""
    owner = getattr(function, '__self__', None)
    if not isinstance(owner, FusedModel) or getattr(function, '__func__', None) is not FusedModel.cost:
        return None
    if solver is not None and solver not in owner.fused_solvers:
        return None
    if jac_function is None:
        return owner          # analytic Jacobian of the built-in model stands in for autograd's
    if getattr(jac_function, '__self__', None) is owner and getattr(jac_function, '__func__', None) is FusedModel.jac:
        return owner
    return None


def bound_model_of(function, jac_function, solver=None):
    """(model, args) when both callables are the ``cost`` / ``jac`` of one BoundModel whose model has fused kernels for
    `solver`, else None."""
    owner = getattr(function, '__self__', None)
    if not isinstance(owner, BoundModel) or getattr(function, '__func__', None) is not BoundModel.cost:
        return None
    if solver is not None and solver not in owner.model.fused_solvers:
        return None
    if jac_function is not None and not (getattr(jac_function, '__self__', None) is owner and
                                         getattr(jac_function, '__func__', None) is BoundModel.jac):
        return None
    return owner.model, owner.args
