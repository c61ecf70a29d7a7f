"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/rbacknn.h declares, and fails loudly (no CPU fallback) when there is no GPU."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_functions():
    src = open(os.path.join(ROOT, "include", "rbacknn.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rk_[a-z_0-9]+)\s*\(", src)))


def test_header_symbols_all_exported_and_bound():
    from vectorsearch_rbac_b200 import _lib
    names = _declared_functions()
    assert len(names) >= 25
    lib = C.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/rbacknn.h but not exported by librbacknn.so"
    assert set(names) == set(_lib.SIGNATURES), set(names) ^ set(_lib.SIGNATURES)
    assert lib.rk_abi_version() == _lib.RK_ABI_VERSION == 3


def test_struct_sizes_match_header():
    from vectorsearch_rbac_b200 import _lib
    assert C.sizeof(_lib.RkConfig) == 16 * 4
    assert C.sizeof(_lib.RkScanOpts) == 8 * 4
    assert C.sizeof(_lib.RkStats) == 8 * 8 + 4 * 4 + 8 + 9 * 8


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import vectorsearch_rbac_b200 as vr
    with pytest.raises(vr.RkError) as ei:
        vr.Index(128)
    assert ei.value.code == -5 and "no CPU path" in str(ei.value)


def test_argument_validation_needs_no_gpu():
    from vectorsearch_rbac_b200 import _lib
    cfg = _lib.RkConfig()
    cfg.abi_version, cfg.dim, cfg.metric, cfg.device, cfg.k_max = _lib.RK_ABI_VERSION, 0, 0, 0, 10
    assert not _lib.lib.rk_open(C.byref(cfg))
    assert "dim 0" in _lib.last_error()
    cfg.dim, cfg.k_max = 128, 5000
    assert not _lib.lib.rk_open(C.byref(cfg))
    assert "k_max" in _lib.last_error()
    cfg.abi_version = 99
    assert not _lib.lib.rk_open(C.byref(cfg))
    assert "ABI" in _lib.last_error()
    assert _lib.lib.rk_finalize(None) == -1


def test_product_package_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "vectorsearch-rbac_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dp, fn)).read()
                assert "oracle" not in text.lower(), f"{fn} mentions the oracle"
