"""ctypes binding of include/rbacknn.h.

There is no Python or CPU fallback: if librbacknn.so is missing this module raises at import, and
without a CUDA device rk_open() fails with RK_ERR_NO_DEVICE.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# RBACKNN_LIB: another build of the same library (kernel experiments); the default is the in-tree build
LIB_PATH = os.environ.get("RBACKNN_LIB") or os.path.join(_HERE, "librbacknn.so")

RK_ABI_VERSION = 3


class RkConfig(C.Structure):
    _fields_ = [("abi_version", C.c_int32), ("dim", C.c_int32), ("metric", C.c_int32), ("device", C.c_int32),
                ("k_max", C.c_int32), ("reserved", C.c_int32 * 11)]


class RkScanOpts(C.Structure):
    _fields_ = [("k", C.c_int32), ("probes", C.c_int32), ("max_probes", C.c_int32), ("filter", C.c_int32),
                ("null_query", C.c_int32), ("layout", C.c_int32), ("keep_duplicates", C.c_int32),
                ("engine", C.c_int32)]


class RkStats(C.Structure):
    _fields_ = [("rows", C.c_uint64), ("docs", C.c_uint64), ("partitions", C.c_uint64), ("device_bytes", C.c_uint64),
                ("launches", C.c_uint64), ("segments", C.c_uint64), ("rows_scanned", C.c_uint64),
                ("algo_bytes", C.c_uint64), ("scan_ms", C.c_float), ("select_ms", C.c_float), ("probe_ms", C.c_float),
                ("reserved_f", C.c_float), ("total_launches", C.c_uint64), ("tensor_redo", C.c_uint64),
                ("tensor_exact", C.c_uint64), ("tensor_plan", C.c_uint64), ("tensor_flop", C.c_uint64), ("tensor_plan_bytes", C.c_uint64),
                ("segment_kinds", C.c_uint64), ("single_fused", C.c_uint64), ("ivf_assign_tc_rows", C.c_uint64),
                ("ivf_assign_fallback_rows", C.c_uint64)]


# every symbol include/rbacknn.h declares: name -> (restype, argtypes)
_P = C.c_void_p
SIGNATURES = {
    "rk_open": (_P, [C.POINTER(RkConfig)]),
    "rk_close": (None, [_P]),
    "rk_last_error": (C.c_char_p, []),
    "rk_abi_version": (C.c_int, []),
    "rk_set_option": (C.c_int, [_P, C.c_char_p, C.c_int64]),
    "rk_load_rows": (C.c_int, [_P, _P, _P, _P, C.c_size_t]),
    "rk_load_rows_device": (C.c_int, [_P, _P, _P, _P, C.c_size_t]),
    "rk_reserve_rows": (C.c_int, [_P, C.c_size_t]),
    "rk_set_doc_roles": (C.c_int, [_P, _P, _P, C.c_size_t]),
    "rk_set_user_roles": (C.c_int, [_P, _P, _P, C.c_size_t]),
    "rk_build_role_partitions": (C.c_int, [_P]),
    "rk_build_combination_partitions": (C.c_int, [_P]),
    "rk_define_partition": (C.c_int, [_P, C.c_uint32, _P, C.c_size_t]),
    "rk_map_roleset_to_partitions": (C.c_int, [_P, _P, C.c_size_t, _P, C.c_size_t]),
    "rk_finalize": (C.c_int, [_P]),
    "rk_build_ivf": (C.c_int, [_P, C.c_int, C.c_uint64]),
    "rk_build_ivf_from_centers": (C.c_int, [_P, C.c_int, _P]),
    "rk_build_ivf_from_lists": (C.c_int, [_P, C.c_int, _P, _P]),
    "rk_ivf_info": (C.c_int, [_P, C.POINTER(C.c_int), _P, _P]),
    "rk_ivfpages_open": (C.c_int, [_P, C.c_size_t, C.POINTER(_P)]),
    "rk_ivfpages_info": (C.c_int, [_P, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "rk_ivfpages_read": (C.c_int, [_P, _P, _P, _P, _P, _P]),
    "rk_ivfpages_close": (None, [_P]),
    "rk_beginscan": (_P, [_P]),
    "rk_rescan": (C.c_int, [_P, _P, C.c_int32, C.c_uint32, C.POINTER(RkScanOpts)]),
    "rk_gettuple": (C.c_int, [_P, C.POINTER(C.c_uint64), C.POINTER(C.c_uint32), C.POINTER(C.c_double)]),
    "rk_gettuple_batch": (C.c_int, [_P, C.c_int, _P, _P, _P]),
    "rk_endscan": (None, [_P]),
    "rk_search_batch": (C.c_int, [_P, C.c_size_t, _P, _P, C.POINTER(RkScanOpts), _P, _P, _P, _P]),
    "rk_search_batch_device": (C.c_int, [_P, C.c_size_t, _P, _P, C.POINTER(RkScanOpts), _P, _P, _P, _P]),
    "rk_search_batch_shard": (C.c_int, [_P, C.c_size_t, _P, _P, C.POINTER(RkScanOpts), _P, _P, _P]),
    "rk_ivf_probe_device": (C.c_int, [_P, C.c_size_t, _P, C.POINTER(RkScanOpts), _P]),
    "rk_search_batch_shard_probed": (C.c_int, [_P, C.c_size_t, _P, _P, C.POINTER(RkScanOpts), _P, _P, _P, _P]),
    "rk_merge_shards": (C.c_int, [_P, C.c_size_t, C.c_int, C.c_int, C.c_size_t, _P, _P, _P, _P, _P, _P, _P]),
    "rk_exchange_create": (C.c_int, [_P, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(_P)]),
    "rk_exchange_destroy": (C.c_int, [_P]),
    "rk_exchange_get_handle": (C.c_int, [_P, _P]),
    "rk_exchange_base": (_P, [_P]),
    "rk_exchange_block_bytes": (C.c_size_t, [_P]),
    "rk_exchange_connect_ipc": (C.c_int, [_P, C.c_int, _P]),
    "rk_exchange_connect_ptr": (C.c_int, [_P, C.c_int, _P, C.c_int]),
    "rk_search_batch_shard_push": (C.c_int, [_P, _P, C.c_size_t, _P, _P, C.POINTER(RkScanOpts), _P]),
    "rk_merge_shards_pull": (C.c_int, [_P, _P, C.c_size_t, C.c_int, _P, _P, _P, _P, _P]),
    "rk_exchange_status": (C.c_int, [_P, _P]),
    "rk_multi_open": (_P, [C.POINTER(RkConfig), _P, C.c_int32, C.c_size_t]),
    "rk_multi_close": (None, [_P]),
    "rk_multi_n_gpus": (C.c_int, [_P]),
    "rk_multi_shard": (_P, [_P, C.c_int]),
    "rk_multi_load_rows": (C.c_int, [_P, _P, _P, _P, C.c_size_t]),
    "rk_multi_set_doc_roles": (C.c_int, [_P, _P, _P, C.c_size_t]),
    "rk_multi_set_user_roles": (C.c_int, [_P, _P, _P, C.c_size_t]),
    "rk_multi_build_role_partitions": (C.c_int, [_P]),
    "rk_multi_build_combination_partitions": (C.c_int, [_P]),
    "rk_multi_define_partition": (C.c_int, [_P, C.c_uint32, _P, C.c_size_t]),
    "rk_multi_map_roleset_to_partitions": (C.c_int, [_P, _P, C.c_size_t, _P, C.c_size_t]),
    "rk_multi_finalize": (C.c_int, [_P]),
    "rk_multi_build_ivf": (C.c_int, [_P, C.c_int, C.c_uint64]),
    "rk_multi_search_batch": (C.c_int, [_P, C.c_size_t, _P, _P, C.POINTER(RkScanOpts), _P, _P, _P, _P]),
    "rk_search_rows_noalloc": (C.c_int, [_P, C.c_size_t, _P, C.c_int32, _P, _P, _P, _P]),
    "rk_set_stream": (C.c_int, [_P, _P]),
    "rk_synchronize": (C.c_int, [_P]),
    "rk_get_stats": (C.c_int, [_P, C.POINTER(RkStats)]),
    "rk_enable_timing": (C.c_int, [_P, C.c_int]),
    "rk_get_scan_times": (C.c_int, [_P, _P, C.c_int]),
}


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python __graft_entry__.py` (nvcc, sm_100a). "
            "This engine has no CPU or pure-Python fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)   # AttributeError here == header / library mismatch
        fn.restype = res
        fn.argtypes = args
    if lib.rk_abi_version() != RK_ABI_VERSION:
        raise ImportError(f"librbacknn ABI {lib.rk_abi_version()} != binding {RK_ABI_VERSION}")
    return lib


lib = _load()


def last_error() -> str:
    return (lib.rk_last_error() or b"").decode()
