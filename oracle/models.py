"""TEST INFRASTRUCTURE ONLY — CPU (torch, any float dtype) restatement of the model
callables the reference's tests feed into the batched solvers.

Vectorised over the batch dimension so the oracle can run at B >> 6; the reference's
own test versions loop over B in Python (tests/unit_test_parallel.py:67-79, 95-103).
``tests/golden/make_golden.py`` asserts these are bit-identical to those loop versions.

EMG ("exponentially modified / skewed Gaussian"), parameters p = [alpha, mu, sigma, eta]:
    m(t) = alpha * exp(-(t-mu)^2 / (2 sigma^2)) * (1 + erf(eta (t-mu) / (sigma sqrt 2)))
      reference: tests/emg.py:20-21, tests/unit_test_parallel.py:81-93
    cost  f = (y - m)^2                                tests/unit_test_parallel.py:40,55-56
    jac   J = 2 (y - m) * [-G A, alpha dmu, alpha dsigma, alpha deta]
                                                       tests/unit_test_parallel.py:105-118
      with the partials of tests/emg.py:23-49 (all carry the sign of d(-m)/dp).

Skewed Gaussian, p = [norm, mean, sigm, skew (, offset)]:
    m(x) = norm/(sigm sqrt(2 pi)) * exp(-(x-mean)^2/(2 sigm^2)) * (1+erf(skew (x-mean)/(sigm sqrt 2)))
      reference: tests/unit_test_skewed_gaussian.py:43-51  (the reference differentiates it
      with autograd; the analytic Jacobian below is checked against autograd in tests/).
    The optional 5th parameter is a constant baseline offset (SURVEY.md §8d, cfg-3 reading).
"""
import math
import torch

_SQRT2 = 2 ** .5
_SQRT_2_OVER_PI = (2 / math.pi) ** .5


def _cols(p):
    # p [B,P] -> P tensors [B,1] that broadcast against t [N]
    return [p[:, i, None] for i in range(p.shape[1])]


# ----------------------------------------------------------------------------- EMG
def emg_terms(p, t):
    """gauss G, asymmetry A, erf-derivative kernel E; each [B,N]. tests/emg.py:20,21,26"""
    alpha, mu, sigma, eta = _cols(p)
    d = t - mu
    G = torch.exp(-.5 * d ** 2 / sigma ** 2)
    A = 1 + torch.erf(eta * d / (sigma * _SQRT2))
    E = torch.exp(-.5 * (d ** 2 / sigma ** 2) * eta ** 2)
    return G, A, E


def emg_model(p, t):
    """[B,N] model values; tests/unit_test_parallel.py:81-93."""
    alpha = p[:, 0, None]
    G, A, _ = emg_terms(p, t)
    return alpha * G * A


def _as_bcn(y, like):
    return y if y.dim() == 3 else y[:, None, :]


def emg_cost(p, t, y, n_costs=None):
    """Squared-error 'residuals' [B,C,N].  y is [B,N] (shared by all costs, as in the
    reference's multi_cost_batch) or [B,C,N]; n_costs repeats a [B,N] y over C."""
    r = _as_bcn(y, p) - emg_model(p, t)[:, None, :]
    f = r ** 2
    if n_costs is not None and f.shape[1] == 1:
        f = f.expand(-1, n_costs, -1).contiguous()
    return f


def emg_jac(p, t, y, n_costs=None):
    """Analytic Jacobian of emg_cost, [B,C,N,4]; tests/unit_test_parallel.py:105-118."""
    alpha, mu, sigma, eta = _cols(p)
    d = t - mu
    G, A, E = emg_terms(p, t)
    # d(1+erf)/d{mu,sigma,eta}: tests/emg.py:27-29
    dA_mu = E * -1 * _SQRT_2_OVER_PI * eta / sigma
    dA_sigma = E * -1 * _SQRT_2_OVER_PI * eta * d / sigma ** 2
    dA_eta = E * _SQRT_2_OVER_PI * d / sigma
    # dG/d{mu,sigma}: tests/emg.py:23-24
    dG_mu = G * d / sigma ** 2
    dG_sigma = G * d ** 2 / sigma ** 3
    # negated product rule: tests/emg.py:31-49
    k_alpha = -1 * G * A
    k_mu = -1 * (G * dA_mu + dG_mu * A)
    k_sigma = -1 * (G * dA_sigma + dG_sigma * A)
    k_eta = -1 * G * dA_eta
    dm = torch.stack([k_alpha, alpha * k_mu, alpha * k_sigma, alpha * k_eta], dim=-1)  # [B,N,4]
    c = 2 * (_as_bcn(y, p) - (alpha * G * A)[:, None, :])                               # [B,C,N]
    J = c[..., None] * dm[:, None, :, :]
    if n_costs is not None and J.shape[1] == 1:
        J = J.expand(-1, n_costs, -1, -1).contiguous()
    return J


# ----------------------------------------------------------------- skewed Gaussian
def skewgauss_model(p, x):
    """[B,N]; tests/unit_test_skewed_gaussian.py:43-51 (+ optional offset p[:,4])."""
    c = _cols(p)
    norm, mean, sigm, skew = c[:4]
    amp = norm / (sigm * (2 * math.pi) ** .5)
    spread = torch.exp((-(x - mean) ** 2.0) / (2 * sigm ** 2.0))
    skw = 1 + torch.erf((skew * (x - mean)) / (sigm * _SQRT2))
    m = amp * spread * skw
    if len(c) == 5:
        m = m + c[4]
    return m


def skewgauss_cost(p, x, y):
    """[B,1,N] squared errors; cost_fun of tests/unit_test_skewed_gaussian.py:35."""
    return ((y - skewgauss_model(p, x)) ** 2)[:, None, :]


def skewgauss_jac(p, x, y):
    """Analytic [B,1,N,P] Jacobian of skewgauss_cost (derived; reference uses autograd)."""
    c = _cols(p)
    norm, mean, sigm, skew = c[:4]
    d = x - mean
    amp = norm / (sigm * (2 * math.pi) ** .5)
    S = torch.exp(-(d ** 2) / (2 * sigm ** 2))
    K = 1 + torch.erf(skew * d / (sigm * _SQRT2))
    E = torch.exp(-(skew ** 2) * d ** 2 / (2 * sigm ** 2))
    dm = [S * K / (sigm * (2 * math.pi) ** .5),
          amp * S * (d / sigm ** 2 * K - _SQRT_2_OVER_PI * skew / sigm * E),
          amp * S * (K * (d ** 2 / sigm ** 3 - 1 / sigm) - _SQRT_2_OVER_PI * skew * d / sigm ** 2 * E),
          amp * S * _SQRT_2_OVER_PI * E * d / sigm]
    m = amp * S * K
    if len(c) == 5:
        m = m + c[4]
        dm.append(torch.ones_like(m))
    dm = torch.stack(dm, dim=-1)
    r = y - m
    return (-2 * r[..., None] * dm)[:, None, :, :]


# ------------------------------------------------------- Gaussian on a constant baseline
# Not a reference model: the CPU statement of torchimize_b200/user_models/gauss_offset.cuh, the example functor
# of the compiled-model path, written with the reference's cost convention
# (squared error used as residual, tests/unit_test_parallel.py:40,105-118).  p = [a, mu, sigma, c].
def gauss_offset_model(p, x):
    a, mu, sigma, c = _cols(p)
    return a * torch.exp(-(x - mu) ** 2 / (2 * sigma ** 2)) + c


def gauss_offset_cost(p, x, y, n_costs=None):
    f = (_as_bcn(y, p) - gauss_offset_model(p, x)[:, None, :]) ** 2
    if n_costs is not None and f.shape[1] == 1:
        f = f.expand(-1, n_costs, -1).contiguous()
    return f


def gauss_offset_jac(p, x, y, n_costs=None):
    a, mu, sigma, c = _cols(p)
    d = x - mu
    G = torch.exp(-d ** 2 / (2 * sigma ** 2))
    dm = torch.stack([G, a * G * d / sigma ** 2, a * G * d ** 2 / sigma ** 3, torch.ones_like(G)], dim=-1)   # [B,N,4]
    r = _as_bcn(y, p) - (a * G + c)[:, None, :]
    J = -2 * r[..., None] * dm[:, None, :, :]
    if n_costs is not None and J.shape[1] == 1:
        J = J.expand(-1, n_costs, -1, -1).contiguous()
    return J


# ----------------------------------------------------------------------- sum of K EMG peaks
# SURVEY.md §8d large-P workload (cfg-4: 4 peaks = 16 parameters): the reference's EMG model and partials
# (tests/emg.py:19-49) applied per peak; model values summed in peak order, Jacobian blocks concatenated.
def multi_emg_model(p, t):
    K = p.shape[1] // 4
    m = emg_model(p[:, :4], t)
    for j in range(1, K):
        m = m + emg_model(p[:, 4 * j:4 * j + 4], t)
    return m


def multi_emg_cost(p, t, y, n_costs=None):
    f = (_as_bcn(y, p) - multi_emg_model(p, t)[:, None, :]) ** 2
    if n_costs is not None and f.shape[1] == 1:
        f = f.expand(-1, n_costs, -1).contiguous()
    return f


def multi_emg_jac(p, t, y, n_costs=None):
    K = p.shape[1] // 4
    yb = _as_bcn(y, p)
    c = 2 * (yb - multi_emg_model(p, t)[:, None, :])                                 # [B,C,N]
    blocks = []
    for j in range(K):
        pj = p[:, 4 * j:4 * j + 4]
        # emg_jac(pj, t, y_j) = 2 (y_j - m_j) * dm_j: with y_j := m_j + 1/2 the factor is exactly 1 -> dm_j [B,1,N,4]
        alpha, mu, sigma, eta = _cols(pj)
        d = t - mu
        G, A, E = emg_terms(pj, t)
        dA_mu = E * -1 * _SQRT_2_OVER_PI * eta / sigma
        dA_sigma = E * -1 * _SQRT_2_OVER_PI * eta * d / sigma ** 2
        dA_eta = E * _SQRT_2_OVER_PI * d / sigma
        dG_mu = G * d / sigma ** 2
        dG_sigma = G * d ** 2 / sigma ** 3
        blocks.append(torch.stack([-1 * G * A, alpha * (-1 * (G * dA_mu + dG_mu * A)), alpha * (-1 * (G * dA_sigma + dG_sigma * A)),
                                   alpha * (-1 * G * dA_eta)], dim=-1))
    dm = torch.cat(blocks, dim=-1)                                                   # [B,N,4K]
    J = c[..., None] * dm[:, None, :, :]
    if n_costs is not None and J.shape[1] == 1:
        J = J.expand(-1, n_costs, -1, -1).contiguous()
    return J


# ------------------------------------------------------------- single-problem EMG callables
# tests/unit_test_analytical_jacobian.py:35,50-73: p [4], t [N], y [N] -> f [N] = (y - emg(p, t))^2, J [N,4].
def emg_single_cost(p, t, y):
    return emg_cost(p[None], t, y[None])[0, 0]


def emg_single_jac(p, t, y):
    return emg_jac(p[None], t, y[None])[0, 0]
