#!/usr/bin/env python
"""Builds, in-tree (torchimize_b200/_jit/), the plugin libraries the GPU tests compile (tests/test_compiled_model.py),
so that a GPU box spends its minutes on kernels instead of on nvcc.  Run after every change of csrc/ (the cache key
hashes the kernel sources).  Several models compile in parallel."""
import os
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from torchimize_b200 import _lib                                  # noqa: E402

MODELS = {
    'gauss_offset': open(os.path.join(ROOT, 'torchimize_b200', 'user_models', 'gauss_offset.cuh')).read(),
    'emg_clone': 'namespace nlls { template <typename T> struct UserModel : EmgModel<T> {}; }\n',
    'emg2_clone': 'namespace nlls { template <typename T> struct UserModel : MultiEmgModel<T, 2> {}; }\n',
}

def build_all(prune=True):
    """compile the three plugins in parallel (cached by source hash); prune: drop builds of older kernel sources"""
    import shutil
    keep, built = set(), []
    with ThreadPoolExecutor(len(MODELS)) as ex:
        for so in ex.map(lambda kv: _lib.build_plugin(kv[1], kv[0]), MODELS.items()):
            built.append(so)
            keep.add(os.path.basename(os.path.dirname(so)))
    if prune:
        for d in os.listdir(_lib.JIT_DIR):                           # builds of older kernel sources: dead weight in the snapshot
            if d not in keep:
                shutil.rmtree(os.path.join(_lib.JIT_DIR, d), ignore_errors=True)
    return built


if __name__ == '__main__':
    for so in build_all():
        print('built', os.path.relpath(so, ROOT))
