"""engine.py — Python handle over the C-ABI (include/rbacknn.h).  Host logic only: every distance,
mask test and top-k runs in librbacknn.so on the GPU."""
from __future__ import annotations

import ctypes as C
from typing import Iterable, Optional, Sequence, Tuple

import numpy as np

from . import _lib
from ._lib import RkConfig, RkScanOpts, RkStats, lib

METRIC_L2, METRIC_IP = 0, 1
FILTER_PARTITION, FILTER_BITMASK, FILTER_NONE = 0, 1, 2
LAYOUT_AUTO, LAYOUT_ROLE, LAYOUT_COMBINATION, LAYOUT_DYNAMIC = 0, 1, 2, 3
ENGINE_AUTO, ENGINE_SCAN, ENGINE_TENSOR = 0, 1, 2


class RkError(RuntimeError):
    """A negative rk_status from the library; `.code` holds it (RK_ERR_DIM = -2, ...)."""

    def __init__(self, code: int, msg: str):
        super().__init__(f"[rk {code}] {msg}")
        self.code = code


def _check(rc: int) -> int:
    if rc < 0:
        raise RkError(rc, _lib.last_error())
    return rc


def _u32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.uint32)


def make_opts(k: int, probes: int = 0, max_probes: int = 0, filter: int = FILTER_PARTITION, null_query: bool = False,
              layout: int = LAYOUT_AUTO, keep_duplicates: bool = False, engine: int = ENGINE_AUTO) -> RkScanOpts:
    o = RkScanOpts()
    o.engine = int(engine)
    o.k, o.probes, o.max_probes, o.filter = int(k), int(probes), int(max_probes), int(filter)
    o.null_query, o.layout, o.keep_duplicates = int(bool(null_query)), int(layout), int(bool(keep_duplicates))
    return o


class Index:
    """One GPU-resident index: rows + permissions + partition layouts (+ optional IVFFlat lists)."""

    def __init__(self, dim: int, metric: int = METRIC_L2, device: int = 0, k_max: int = 128):
        cfg = RkConfig()
        cfg.abi_version, cfg.dim, cfg.metric, cfg.device, cfg.k_max = _lib.RK_ABI_VERSION, dim, metric, device, k_max
        self._h = lib.rk_open(C.byref(cfg))
        if not self._h:
            raise RkError(-5, _lib.last_error())
        self.dim, self.metric, self.device, self.k_max = dim, metric, device, k_max

    # -- lifecycle
    def close(self) -> None:
        if getattr(self, "_h", None):
            lib.rk_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- loading
    def reserve_rows(self, n: int) -> None:
        _check(lib.rk_reserve_rows(self._h, n))

    def load_rows(self, x: np.ndarray, tid: np.ndarray, doc: np.ndarray) -> None:
        x = np.ascontiguousarray(x, dtype=np.float32)
        if x.ndim != 2 or x.shape[1] != self.dim:
            raise RkError(-2, f"different vector dimensions {self.dim} and {x.shape[-1] if x.ndim else 0}")
        tid = np.ascontiguousarray(tid, dtype=np.uint64)
        doc = _u32(doc)
        assert len(tid) == len(doc) == x.shape[0]
        _check(lib.rk_load_rows(self._h, x.ctypes.data, tid.ctypes.data, doc.ctypes.data, x.shape[0]))

    def load_rows_device(self, d_x_ptr: int, n: int, tid: np.ndarray, doc: np.ndarray) -> None:
        """Rows already resident in HBM: d_x_ptr is a raw device address of n*dim fp32."""
        tid = np.ascontiguousarray(tid, dtype=np.uint64)
        doc = _u32(doc)
        assert len(tid) == len(doc) == n
        _check(lib.rk_load_rows_device(self._h, d_x_ptr, tid.ctypes.data, doc.ctypes.data, n))

    def set_doc_roles(self, doc: Iterable[int], role: Iterable[int]) -> None:
        d, r = _u32(doc), _u32(role)
        _check(lib.rk_set_doc_roles(self._h, d.ctypes.data, r.ctypes.data, len(d)))

    def set_user_roles(self, user: Iterable[int], role: Iterable[int]) -> None:
        u, r = _u32(user), _u32(role)
        _check(lib.rk_set_user_roles(self._h, u.ctypes.data, r.ctypes.data, len(u)))

    def build_role_partitions(self) -> None:
        _check(lib.rk_build_role_partitions(self._h))

    def build_combination_partitions(self) -> None:
        _check(lib.rk_build_combination_partitions(self._h))

    def define_partition(self, part_id: int, docs: Iterable[int]) -> None:
        d = _u32(list(docs))
        _check(lib.rk_define_partition(self._h, part_id, d.ctypes.data, len(d)))

    def map_roleset_to_partitions(self, roles: Sequence[int], parts: Sequence[int]) -> None:
        r, p = _u32(list(roles)), _u32(list(parts))
        _check(lib.rk_map_roleset_to_partitions(self._h, r.ctypes.data, len(r), p.ctypes.data, len(p)))

    def finalize(self) -> None:
        _check(lib.rk_finalize(self._h))

    # -- IVFFlat
    def build_ivf(self, lists: int, seed: int = 42) -> None:
        _check(lib.rk_build_ivf(self._h, lists, seed))

    def build_ivf_from_centers(self, centers: np.ndarray) -> None:
        c = np.ascontiguousarray(centers, dtype=np.float32)
        assert c.ndim == 2 and c.shape[1] == self.dim
        _check(lib.rk_build_ivf_from_centers(self._h, c.shape[0], c.ctypes.data))

    def build_ivf_from_lists(self, centers: np.ndarray, row_list: np.ndarray) -> None:
        """Centres and the row -> list assignment (by load order) of an index that already exists, e.g. IvfflatPages.read()."""
        c = np.ascontiguousarray(centers, dtype=np.float32)
        rl = np.ascontiguousarray(row_list, dtype=np.uint32)
        assert c.ndim == 2 and c.shape[1] == self.dim
        _check(lib.rk_build_ivf_from_lists(self._h, c.shape[0], c.ctypes.data, rl.ctypes.data))

    def ivf_info(self, nrows: int) -> Tuple[np.ndarray, np.ndarray]:
        n = C.c_int(0)
        _check(lib.rk_ivf_info(self._h, C.byref(n), None, None))
        centers = np.empty((n.value, self.dim), dtype=np.float32)
        row_list = np.empty(nrows, dtype=np.uint32)
        _check(lib.rk_ivf_info(self._h, C.byref(n), centers.ctypes.data, row_list.ctypes.data))
        return centers, row_list

    def ivf_centers(self) -> np.ndarray:
        """The list centres [lists, dim] of the built IVFFlat index."""
        n = C.c_int(0)
        _check(lib.rk_ivf_info(self._h, C.byref(n), None, None))
        centers = np.empty((n.value, self.dim), dtype=np.float32)
        _check(lib.rk_ivf_info(self._h, C.byref(n), centers.ctypes.data, None))
        return centers

    # -- search
    def search_batch(self, Q: np.ndarray, user_ids: Optional[Sequence[int]], k: int, probes: int = 0,
                     filter: int = FILTER_PARTITION, layout: int = LAYOUT_AUTO, keep_duplicates: bool = False,
                     null_query: bool = False, max_probes: int = 0, engine: int = ENGINE_AUTO):
        """nq independent filtered top-k scans (host buffers, copies inside).  Returns
        (tid[nq,k] uint64, doc[nq,k] uint32, dist[nq,k] float64, count[nq] int32)."""
        Q = np.ascontiguousarray(Q, dtype=np.float32)
        if Q.ndim == 1:
            Q = Q[None, :]
        if Q.shape[1] != self.dim:
            raise RkError(-2, f"different vector dimensions {self.dim} and {Q.shape[1]}")
        nq = Q.shape[0]
        users = None if user_ids is None else _u32(user_ids)
        o = make_opts(k, probes, max_probes, filter, null_query, layout, keep_duplicates, engine)
        tid = np.empty((nq, k), dtype=np.uint64)
        doc = np.empty((nq, k), dtype=np.uint32)
        dist = np.empty((nq, k), dtype=np.float64)
        cnt = np.zeros(nq, dtype=np.int32)
        _check(lib.rk_search_batch(self._h, nq, Q.ctypes.data, None if users is None else users.ctypes.data,
                                   C.byref(o), tid.ctypes.data, doc.ctypes.data, dist.ctypes.data, cnt.ctypes.data))
        return tid, doc, dist, cnt

    def search_batch_device(self, dQ_ptr: int, nq: int, user_ids: Sequence[int], opts: RkScanOpts, d_tid: int,
                            d_doc: int, d_dist: int, d_count: int) -> None:
        """Device-pointer variant (asynchronous on the index stream); pointers are raw addresses,
        e.g. torch.Tensor.data_ptr()."""
        users = _u32(user_ids)
        _check(lib.rk_search_batch_device(self._h, nq, dQ_ptr, users.ctypes.data, C.byref(opts), d_tid, d_doc, d_dist,
                                          d_count))

    def search_batch_shard(self, dQ_ptr: int, nq: int, user_ids: Sequence[int], opts: RkScanOpts, d_key: int,
                           d_tid: int, d_doc: int) -> None:
        users = _u32(user_ids)
        _check(lib.rk_search_batch_shard(self._h, nq, dQ_ptr, users.ctypes.data, C.byref(opts), d_key, d_tid, d_doc))

    def ivf_lists(self) -> int:
        n = C.c_int(0)
        _check(lib.rk_ivf_info(self._h, C.byref(n), None, None))
        return int(n.value)

    def ivf_probe_device(self, dQ_ptr: int, nq: int, opts: RkScanOpts, d_lists: int) -> None:
        """List choice alone: min(opts.probes, lists) list numbers per query, nearest first, into d_lists (device u32)."""
        _check(lib.rk_ivf_probe_device(self._h, nq, dQ_ptr, C.byref(opts), d_lists))

    def search_batch_shard_probed(self, dQ_ptr: int, nq: int, user_ids: Sequence[int], opts: RkScanOpts, d_lists: int,
                                  d_key: int, d_tid: int, d_doc: int) -> None:
        users = _u32(user_ids)
        _check(lib.rk_search_batch_shard_probed(self._h, nq, dQ_ptr, users.ctypes.data, C.byref(opts), d_lists, d_key,
                                                d_tid, d_doc))

    def merge_shards(self, nq: int, k: int, n_shards: int, d_key: int, d_tid: int, d_doc: int, d_out_tid: int,
                     d_out_doc: int, d_out_dist: int, d_out_count: int, shard_stride_bytes: int = 0) -> None:
        _check(lib.rk_merge_shards(self._h, nq, k, n_shards, shard_stride_bytes, d_key, d_tid, d_doc, d_out_tid,
                                   d_out_doc, d_out_dist, d_out_count))

    def begin_scan(self) -> "Scan":
        return Scan(self)

    # -- plumbing
    def set_option(self, name: str, value: int) -> None:
        """Tuning / test switch of this index (rk_set_option)."""
        _check(lib.rk_set_option(self._h, name.encode(), int(value)))

    def set_stream(self, cuda_stream: Optional[int]) -> None:
        _check(lib.rk_set_stream(self._h, cuda_stream))

    def synchronize(self) -> None:
        _check(lib.rk_synchronize(self._h))

    def enable_timing(self, on: bool = True) -> None:
        _check(lib.rk_enable_timing(self._h, int(on)))

    def scan_times_ms(self, max_n: int = 64) -> np.ndarray:
        """CUDA-event durations of the most recent scan-kernel launches (timing must be enabled)."""
        out = np.zeros(max_n, dtype=np.float32)
        n = _check(lib.rk_get_scan_times(self._h, out.ctypes.data, max_n))
        return out[:n]

    def stats(self) -> dict:
        s = RkStats()
        _check(lib.rk_get_stats(self._h, C.byref(s)))
        return {f: getattr(s, f) for f, _ in RkStats._fields_ if not f.startswith("reserved")}


class Exchange:
    """Ring of receive slots in this rank's HBM that the other ranks' GPUs write their shard results into over NVLink
    (rk_exchange_* in include/rbacknn.h): push = local search + stores into every peer's slot + arrival flags; pull = wait
    for the peers' flags inside the merge kernel + G-way merge + acknowledgement.  No collective call on the data path."""

    def __init__(self, index: Index, nq_max: int, k_max: int, world: int, rank: int, depth: int = 4):
        self.index, self.world, self.rank, self.depth = index, world, rank, depth
        h = C.c_void_p()
        _check(lib.rk_exchange_create(index._h, nq_max, k_max, world, rank, depth, C.byref(h)))
        self._h = h

    def close(self) -> None:
        if getattr(self, "_h", None):
            lib.rk_exchange_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def handle(self) -> bytes:
        """CUDA IPC handle of the ring (64 bytes) for the other processes."""
        buf = C.create_string_buffer(64)
        _check(lib.rk_exchange_get_handle(self._h, buf))
        return buf.raw

    def base(self) -> int:
        return int(lib.rk_exchange_base(self._h) or 0)

    def connect_ipc(self, peer: int, handle: bytes) -> None:
        _check(lib.rk_exchange_connect_ipc(self._h, peer, C.create_string_buffer(handle, 64)))

    def connect_ptr(self, peer: int, base: int, device: int = -1) -> None:
        _check(lib.rk_exchange_connect_ptr(self._h, peer, base, device))

    def push(self, dQ_ptr: int, nq: int, user_ids: Sequence[int], opts: RkScanOpts, d_lists: int = 0) -> None:
        users = _u32(user_ids)
        _check(lib.rk_search_batch_shard_push(self.index._h, self._h, nq, dQ_ptr, users.ctypes.data, C.byref(opts), d_lists or None))

    def pull(self, nq: int, k: int, d_out_tid: int, d_out_doc: int, d_out_dist: int, d_out_count: int, stream: int = 0) -> None:
        _check(lib.rk_merge_shards_pull(self.index._h, self._h, nq, k, stream or None, d_out_tid, d_out_doc, d_out_dist, d_out_count))

    def status(self, stream: int = 0) -> None:
        """Synchronises the stream; raises if a wait for a peer timed out."""
        _check(lib.rk_exchange_status(self._h, stream or None))


class MultiIndex:
    """Several GPUs of one box behind one handle, one process (rk_multi_* in include/rbacknn.h): document-striped shards, one
    host thread per device, the peer exchange between them.  Mirrors Index for loading, layouts and search_batch."""

    def __init__(self, dim: int, devices: Sequence[int], metric: int = METRIC_L2, k_max: int = 128, nq_max: int = 1024):
        cfg = RkConfig()
        cfg.abi_version, cfg.dim, cfg.metric, cfg.device, cfg.k_max = _lib.RK_ABI_VERSION, dim, metric, int(devices[0]), k_max
        dv = np.ascontiguousarray(devices, dtype=np.int32)
        self._h = lib.rk_multi_open(C.byref(cfg), dv.ctypes.data, len(dv), nq_max)
        if not self._h:
            raise RkError(-5, _lib.last_error())
        self.dim, self.metric, self.devices, self.k_max = dim, metric, list(map(int, devices)), k_max

    def close(self) -> None:
        if getattr(self, "_h", None):
            lib.rk_multi_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def load_rows(self, x: np.ndarray, tid: np.ndarray, doc: np.ndarray) -> None:
        x = np.ascontiguousarray(x, dtype=np.float32)
        if x.ndim != 2 or x.shape[1] != self.dim:
            raise RkError(-2, f"expected {self.dim} dimensions, not {x.shape[-1]}")
        tid = np.ascontiguousarray(tid, dtype=np.uint64)
        doc = _u32(doc)
        _check(lib.rk_multi_load_rows(self._h, x.ctypes.data, tid.ctypes.data, doc.ctypes.data, x.shape[0]))

    def set_doc_roles(self, doc: Iterable[int], role: Iterable[int]) -> None:
        d, r = _u32(doc), _u32(role)
        _check(lib.rk_multi_set_doc_roles(self._h, d.ctypes.data, r.ctypes.data, len(d)))

    def set_user_roles(self, user: Iterable[int], role: Iterable[int]) -> None:
        u, r = _u32(user), _u32(role)
        _check(lib.rk_multi_set_user_roles(self._h, u.ctypes.data, r.ctypes.data, len(u)))

    def build_role_partitions(self) -> None:
        _check(lib.rk_multi_build_role_partitions(self._h))

    def build_combination_partitions(self) -> None:
        _check(lib.rk_multi_build_combination_partitions(self._h))

    def define_partition(self, part_id: int, docs: Iterable[int]) -> None:
        d = _u32(docs)
        _check(lib.rk_multi_define_partition(self._h, part_id, d.ctypes.data, len(d)))

    def map_roleset_to_partitions(self, roles: Sequence[int], parts: Sequence[int]) -> None:
        r, p = _u32(roles), _u32(parts)
        _check(lib.rk_multi_map_roleset_to_partitions(self._h, r.ctypes.data, len(r), p.ctypes.data, len(p)))

    def finalize(self) -> None:
        _check(lib.rk_multi_finalize(self._h))

    def build_ivf(self, lists: int, seed: int = 42) -> None:
        _check(lib.rk_multi_build_ivf(self._h, lists, seed))

    def search_batch(self, Q: np.ndarray, user_ids: Optional[Sequence[int]], k: int, probes: int = 0, filter: int = FILTER_PARTITION,
                     layout: int = LAYOUT_AUTO, engine: int = ENGINE_AUTO, keep_duplicates: bool = False):
        Q = np.ascontiguousarray(Q, dtype=np.float32).reshape(-1, self.dim)
        nq = Q.shape[0]
        users = None if user_ids is None else _u32(user_ids)
        o = make_opts(k, probes=probes, filter=filter, layout=layout, engine=engine, keep_duplicates=keep_duplicates)
        tid = np.empty((nq, k), dtype=np.uint64)
        doc = np.empty((nq, k), dtype=np.uint32)
        dist = np.empty((nq, k), dtype=np.float64)
        cnt = np.zeros(nq, dtype=np.int32)
        _check(lib.rk_multi_search_batch(self._h, nq, Q.ctypes.data, None if users is None else users.ctypes.data, C.byref(o),
                                         tid.ctypes.data, doc.ctypes.data, dist.ctypes.data, cnt.ctypes.data))
        return tid, doc, dist, cnt


class Scan:
    """begin / rescan / gettuple / endscan — the pgvector index-scan contract
    (pgvector/src/ivfflat.h:335-338, ivfscan.c:241-406)."""

    def __init__(self, index: Index):
        self._ix = index
        self._h = lib.rk_beginscan(index._h)
        if not self._h:
            raise RkError(-4, _lib.last_error())

    def rescan(self, q: Optional[np.ndarray], user_id: int, k: int, probes: int = 0, max_probes: int = 0,
               filter: int = FILTER_PARTITION, layout: int = LAYOUT_AUTO) -> None:
        o = make_opts(k, probes, max_probes, filter, q is None, layout)
        if q is None:
            _check(lib.rk_rescan(self._h, None, 0, user_id, C.byref(o)))
        else:
            q = np.ascontiguousarray(q, dtype=np.float32).reshape(-1)
            _check(lib.rk_rescan(self._h, q.ctypes.data, q.shape[0], user_id, C.byref(o)))

    def gettuple(self):
        """Returns (tid, doc, distance) or None when exhausted."""
        t, d, x = C.c_uint64(0), C.c_uint32(0), C.c_double(0)
        rc = _check(lib.rk_gettuple(self._h, C.byref(t), C.byref(d), C.byref(x)))
        return (t.value, d.value, x.value) if rc == 1 else None

    def gettuple_batch(self, max_n: int):
        """Up to max_n tuples in one call: (tid uint64[n], doc uint32[n], distance float64[n]); n == 0 when exhausted."""
        tid = np.empty(max_n, dtype=np.uint64)
        doc = np.empty(max_n, dtype=np.uint32)
        dist = np.empty(max_n, dtype=np.float64)
        n = _check(lib.rk_gettuple_batch(self._h, max_n, tid.ctypes.data, doc.ctypes.data, dist.ctypes.data))
        return tid[:n], doc[:n], dist[:n]

    def __iter__(self):
        while True:
            r = self.gettuple()
            if r is None:
                return
            yield r

    def end(self) -> None:
        if self._h:
            lib.rk_endscan(self._h)
            self._h = None

    def __del__(self):
        try:
            self.end()
        except Exception:
            pass


class IvfflatPages:
    """The main fork of a pgvector ivfflat index relation, parsed (rk_ivfpages_*; pgvector/src/ivfscan.c:45-101, 128-168).
    Host code: needs no device."""

    def __init__(self, rel: bytes):
        self._buf = np.frombuffer(rel, dtype=np.uint8)          # keeps the bytes alive: the handle points into them
        h = C.c_void_p()
        _check(lib.rk_ivfpages_open(self._buf.ctypes.data, self._buf.size, C.byref(h)))
        self._h = h
        d, l, n, u = C.c_int32(), C.c_int32(), C.c_uint64(), C.c_uint64()
        _check(lib.rk_ivfpages_info(self._h, C.byref(d), C.byref(l), C.byref(n), C.byref(u)))
        self.dim, self.lists, self.tuples, self.unused_items = d.value, l.value, n.value, u.value

    def read(self):
        """-> centers [lists, dim], list_start [lists + 1], row_list [tuples], x [tuples, dim], tid [tuples]"""
        centers = np.empty((self.lists, self.dim), dtype=np.float32)
        list_start = np.empty(self.lists + 1, dtype=np.uint32)
        row_list = np.empty(self.tuples, dtype=np.uint32)
        x = np.empty((self.tuples, self.dim), dtype=np.float32)
        tid = np.empty(self.tuples, dtype=np.uint64)
        _check(lib.rk_ivfpages_read(self._h, centers.ctypes.data, list_start.ctypes.data, row_list.ctypes.data, x.ctypes.data, tid.ctypes.data))
        return centers, list_start, row_list, x, tid

    def close(self) -> None:
        if self._h:
            lib.rk_ivfpages_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
