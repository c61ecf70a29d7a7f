"""vectorsearch-rbac_b200 — B200-native RBAC-filtered kNN engine (one hot path of rjzhb/VectorSearch-RBAC).

Layout:
  csrc/        CUDA kernels (sm_100a) and the C-ABI implementation -> librbacknn.so
  _lib.py      ctypes binding of include/rbacknn.h (fails loudly when the .so is missing)
  engine.py    thin Python handle over the C-ABI (Index, Scan)
  harness.py   mirror of the reference's Python search / ground-truth functions for this path
  synth.py     seeded synthetic corpora, RBAC generators and query sets
  sharded.py   one-process-per-GPU document-striped search with an all-gather merge
"""
from . import _lib  # noqa: F401
from .engine import Index, Scan, RkError, METRIC_L2, METRIC_IP, FILTER_PARTITION, FILTER_BITMASK, FILTER_NONE  # noqa: F401
from .engine import ENGINE_AUTO, ENGINE_SCAN, ENGINE_TENSOR  # noqa: F401
from .engine import LAYOUT_AUTO, LAYOUT_ROLE, LAYOUT_COMBINATION, LAYOUT_DYNAMIC  # noqa: F401
