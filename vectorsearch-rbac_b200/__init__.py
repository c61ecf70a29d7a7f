"""vectorsearch-rbac_b200 — B200-native RBAC-filtered kNN engine (one hot path of rjzhb/VectorSearch-RBAC).

Layout:
  csrc/        CUDA kernels (sm_100a) and the C-ABI implementation -> librbacknn.so
  _lib.py      ctypes binding of include/rbacknn.h (fails loudly when the .so is missing)
  engine.py    thin Python handle over the C-ABI (Index, Scan)
  harness.py   mirror of the reference's Python search / ground-truth functions for this path
  synth.py     seeded synthetic corpora, RBAC generators and query sets
  sharded.py   one-process-per-GPU document-striped search with an all-gather merge

The engine names (Index, Scan, METRIC_L2, ...) resolve on first use and load librbacknn.so then (ImportError
if it is missing: there is no fallback).  `import vectorsearch_rbac_b200.synth` alone — the input generators, which
`bench.py --impl reference` needs — does not load the library.
"""
import importlib as _importlib

_ENGINE_NAMES = ("Index", "MultiIndex", "Exchange", "Scan", "IvfflatPages", "RkError", "METRIC_L2", "METRIC_IP", "FILTER_PARTITION", "FILTER_BITMASK", "FILTER_NONE",
                 "ENGINE_AUTO", "ENGINE_SCAN", "ENGINE_TENSOR", "LAYOUT_AUTO", "LAYOUT_ROLE", "LAYOUT_COMBINATION",
                 "LAYOUT_DYNAMIC")
_SUBMODULES = ("_lib", "engine", "harness", "synth", "sharded", "dynamic_layout", "build")


def __getattr__(name):
    if name in _ENGINE_NAMES:
        return getattr(_importlib.import_module(__name__ + ".engine"), name)
    if name in _SUBMODULES:
        return _importlib.import_module(__name__ + "." + name)
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")


def __dir__():
    return sorted(list(globals()) + list(_ENGINE_NAMES) + list(_SUBMODULES))
