"""torchimize_b200 — B200 (sm_100a) batched nonlinear-least-squares core behind the call
signatures of hahnec/torchimize's parallel solvers (torchimize/functions/__init__.py:4-8).

The compute path is hand-written CUDA reached through the C ABI in include/nlls_b200.h
(libnlls_b200.so, loaded with ctypes).  There is no CPU fallback: importing works anywhere,
calling a solver without the library or without a CUDA tensor raises.
"""
__version__ = '0.1.0'
