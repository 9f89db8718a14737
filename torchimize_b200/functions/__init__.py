"""Drop-in surface of the batched path: same names as torchimize/functions/__init__.py:4-8.
(The single-problem solvers of torchimize.functions.single are out of scope — SURVEY.md §8f.)"""
from .parallel.gda_fun_parallel import gradient_descent_parallel, gradient_descent_parallel_plain
from .parallel.gna_fun_parallel import lsq_gna_parallel, lsq_gna_parallel_plain
from .parallel.lma_fun_parallel import lsq_lma_parallel
from .parallel.newton_parallel import newton_step_parallel
from .fun_dims import test_fun_dims_parallel
from .jacobian import batch_jacobian_approx_t, jacobian_approx_t
