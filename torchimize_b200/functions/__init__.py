"""Drop-in surface: same names as torchimize/functions/__init__.py:1-8 (single-problem solvers: SURVEY.md §8f rank 3)."""
from .single.gda_fun_single import gradient_descent
from .single.gna_fun_single import lsq_gna
from .single.lma_fun_single import lsq_lma
from .parallel.gda_fun_parallel import gradient_descent_parallel, gradient_descent_parallel_plain
from .parallel.gna_fun_parallel import lsq_gna_parallel, lsq_gna_parallel_plain
from .parallel.lma_fun_parallel import lsq_lma_parallel
from .parallel.newton_parallel import newton_step_parallel
from .fun_dims import test_fun_dims_parallel
from .jacobian import batch_jacobian_approx_t, jacobian_approx_t
