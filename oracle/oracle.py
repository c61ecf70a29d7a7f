"""oracle.py — numpy/ctypes front end of the CPU checker (oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs, never by the product package.  Parity status: PINNED (see oracle.c header).

Layout logic restated from the reference (paths relative to its checkout):
  role partitions ......... controller/baseline/prefilter/initialize_partitions.py:302-311
  combination partitions .. initialize_partitions.py:549-555, 682-693
  user -> partitions ...... prefilter_role.py:120-124; prefilter_combination_role.py:122-133;
                            controller/dynamic_partition/search.py:44-60
  RLS predicate ........... controller/baseline/pg_row_security/row_level_security.py:55-65
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
METRIC_L2, METRIC_IP = 0, 1

_DISTFN = C.CFUNCTYPE(C.c_float, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float))


def build(force: bool = False) -> None:
    """Compile the C restatement (and oracle/_ref when /root/reference is present)."""
    if force or not os.path.exists(os.path.join(_HERE, "liboracle.so")) or \
            os.path.getmtime(os.path.join(_HERE, "liboracle.so")) < os.path.getmtime(os.path.join(_HERE, "oracle.c")) or \
            not os.path.exists(os.path.join(_HERE, "libsynthgen.so")):
        subprocess.run(["make", "-C", _HERE, "all"], check=True, capture_output=True)


def _fptr(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_float))


class _Lib:
    def __init__(self, path: str):
        self.lib = C.CDLL(path)
        L = self.lib
        L.orc_l2sq.restype = C.c_float
        L.orc_l2sq.argtypes = [C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        L.orc_ip.restype = C.c_float
        L.orc_ip.argtypes = [C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        L.orc_operator_value.restype = C.c_double
        L.orc_operator_value.argtypes = [C.c_int, C.c_float]
        L.orc_order_value.restype = C.c_double
        L.orc_order_value.argtypes = [C.c_int, C.c_float]
        L.orc_check_dims.argtypes = [C.c_int, C.c_int]
        L.orc_check_finite.argtypes = [C.c_void_p, C.c_size_t]
        L.orc_search_batch.restype = C.c_int
        L.orc_search_batch.argtypes = [
            C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p,
            C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
            C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_ivf_assign.argtypes = [C.c_void_p, C.c_int, C.c_size_t, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orc_ivf_order_lists.restype = C.c_int
        L.orc_ivf_order_lists.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                          C.c_void_p, C.c_void_p]
        L.orc_ivf_scan_batch.restype = C.c_long
        L.orc_ivf_scan_batch.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_long,
                                         C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_recall.restype = C.c_double
        L.orc_recall.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]


_libs: Dict[str, _Lib] = {}


def lib(variant: str = "strict") -> _Lib:
    """variant: 'strict' (liboracle.so) or 'pgvflags' (pgvector Makefile flags)."""
    if variant not in _libs:
        build()
        name = {"strict": "liboracle.so", "pgvflags": "liboracle_pgvflags.so"}[variant]
        _libs[variant] = _Lib(os.path.join(_HERE, name))
    return _libs[variant]


class RefLoops:
    """The reference's own distance loops from oracle/_ref (None if that build is absent)."""

    def __init__(self, path: str):
        self.lib = C.CDLL(path)
        for n in ("pgvref_l2sq", "pgvref_ip"):
            f = getattr(self.lib, n)
            f.restype = C.c_float
            f.argtypes = [C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        for n in ("pgvref_l2_distance", "pgvref_neg_ip"):
            f = getattr(self.lib, n)
            f.restype = C.c_double
            f.argtypes = [C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float)]

    def fn_ptrs(self) -> Tuple[int, int]:
        return (C.cast(self.lib.pgvref_l2sq, C.c_void_p).value, C.cast(self.lib.pgvref_ip, C.c_void_p).value)


def ref_loops(o0: bool = False) -> Optional[RefLoops]:
    p = os.path.join(_HERE, "_ref", "libpgvref_O0.so" if o0 else "libpgvref.so")
    return RefLoops(p) if os.path.exists(p) else None


# ----------------------------------------------------------------------------------------------
# scalar helpers (golden-vector surface)

def l2_distance(a: Sequence[float], b: Sequence[float]) -> float:
    """SQL l2_distance / `<->` (vector.c:568-578). Raises ValueError on dim mismatch (vector.c:61-67)."""
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    if lib().lib.orc_check_dims(len(a), len(b)) != 0:
        raise ValueError(f"different vector dimensions {len(a)} and {len(b)}")
    s = lib().lib.orc_l2sq(len(a), _fptr(a), _fptr(b))
    return lib().lib.orc_operator_value(METRIC_L2, s)


def inner_product(a: Sequence[float], b: Sequence[float]) -> float:
    """SQL inner_product (vector.c:611-621)."""
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    if lib().lib.orc_check_dims(len(a), len(b)) != 0:
        raise ValueError(f"different vector dimensions {len(a)} and {len(b)}")
    return float(lib().lib.orc_ip(len(a), _fptr(a), _fptr(b)))


def negative_inner_product(a, b) -> float:
    """SQL `<#>` (vector.c:626-636)."""
    return -inner_product(a, b)


# ----------------------------------------------------------------------------------------------
# layouts

class Rbac:
    """PermissionAssignment + UserRoles as plain arrays."""

    def __init__(self, doc_role_pairs: np.ndarray, user_role_pairs: np.ndarray):
        self.doc_role = np.asarray(doc_role_pairs, dtype=np.int64).reshape(-1, 2)   # (doc, role)
        self.user_role = np.asarray(user_role_pairs, dtype=np.int64).reshape(-1, 2)  # (user, role)

    def roles_of_user(self, user: int) -> List[int]:
        """SELECT role_id FROM userroles WHERE user_id = %s (prefilter_role.py:120), sorted."""
        return sorted(set(self.user_role[self.user_role[:, 0] == user, 1].tolist()))

    def docs_of_role(self, role: int) -> np.ndarray:
        return np.unique(self.doc_role[self.doc_role[:, 1] == role, 0])


def role_partition_rows(row_doc: np.ndarray, rbac: Rbac) -> Dict[int, np.ndarray]:
    """documentblocks_role_<r>: SELECT DISTINCT rows JOIN PermissionAssignment WHERE role_id = r
    (initialize_partitions.py:302-311). Returns role -> ascending row indices."""
    out = {}
    for r in np.unique(rbac.doc_role[:, 1]):
        docs = rbac.docs_of_role(int(r))
        out[int(r)] = np.nonzero(np.isin(row_doc, docs))[0].astype(np.uint32)
    return out


def combination_partition_rows(row_doc: np.ndarray, rbac: Rbac) -> Dict[Tuple[int, ...], np.ndarray]:
    """One table per distinct array_agg(role_id ORDER BY role_id) (initialize_partitions.py:549-555),
    filled with the DISTINCT union WHERE role_id = ANY(roles) (:682-693)."""
    combos = {tuple(rbac.roles_of_user(int(u))) for u in np.unique(rbac.user_role[:, 0])}
    out = {}
    for combo in sorted(combos):
        docs = np.unique(rbac.doc_role[np.isin(rbac.doc_role[:, 1], combo), 0])
        out[combo] = np.nonzero(np.isin(row_doc, docs))[0].astype(np.uint32)
    return out


def allowed_rows(row_doc: np.ndarray, rbac: Rbac, user: int) -> np.ndarray:
    """RLS predicate (row_level_security.py:55-65): EXISTS pa JOIN ur WHERE pa.document_id = row.document_id
    AND ur.user_id = user."""
    roles = rbac.roles_of_user(user)
    docs = np.unique(rbac.doc_role[np.isin(rbac.doc_role[:, 1], roles), 0])
    return np.nonzero(np.isin(row_doc, docs))[0].astype(np.uint32)


def dynamic_partition_plan(row_doc: np.ndarray, rbac: Rbac, partitions: Dict[int, Iterable[int]],
                           roleset_parts: Dict[Tuple[int, ...], Sequence[int]], users: Sequence[int]):
    """The per-query scans of dynamic_partition_search (controller/dynamic_partition/search.py:44-111):
    user -> sorted role set (:44-52) -> partition ids from CombRolePartitions WHERE comb_role = roles (:54-60; a role set
    without a row gets NO partition) -> one `ORDER BY distance LIMIT k` per partition table (:93-104), on which PostgreSQL
    applies the RLS policy EXISTS(user_accessible_documents ...) unless the partition is pure
    (load_result_to_database.py:591-622) — for a pure partition the predicate holds for every row, so applying it always
    gives the same rows.  Returns (row lists, per-query list indices) for search_batch(dedup=True), whose merge is
    merge_results (:347-364).  The layout need NOT cover a user's visible set: rows outside the user's partitions are
    simply never scanned, exactly as in the reference."""
    row_doc = np.asarray(row_doc)
    part_rows = {int(p): np.nonzero(np.isin(row_doc, np.asarray(list(d), dtype=np.int64)))[0].astype(np.uint32)
                 for p, d in partitions.items()}
    plist: List[np.ndarray] = []
    index_of: Dict[Tuple[Tuple[int, ...], int], int] = {}
    qparts: List[List[int]] = []
    allowed_cache: Dict[Tuple[int, ...], np.ndarray] = {}
    for u in users:
        roles = tuple(rbac.roles_of_user(int(u)))
        mine: List[int] = []
        for pid in (roleset_parts.get(roles, ()) if roles else ()):
            key = (roles, int(pid))
            if key not in index_of:
                if roles not in allowed_cache:
                    allowed_cache[roles] = allowed_rows(row_doc, rbac, int(u))
                rows = np.intersect1d(part_rows[int(pid)], allowed_cache[roles]).astype(np.uint32)
                index_of[key] = len(plist)
                plist.append(rows)
            if len(plist[index_of[key]]):
                mine.append(index_of[key])
        qparts.append(mine)
    return plist, qparts


# ----------------------------------------------------------------------------------------------
# batched search

def search_batch(X: np.ndarray, tid: np.ndarray, partitions: Sequence[np.ndarray],
                 query_parts: Sequence[Sequence[int]], Q: np.ndarray, metric: int, k: int, dedup: bool = True,
                 nthreads: int = 1, variant: str = "strict", ref: Optional[RefLoops] = None):
    """For each query i: ORDER BY distance LIMIT k over every partition in query_parts[i], then the
    harness merge (prefilter_role.py:174-189).  Returns (tid[nq,k], row[nq,k], dist[nq,k] float64,
    sum[nq,k] float32, count[nq])."""
    X = np.ascontiguousarray(X, dtype=np.float32)
    Q = np.ascontiguousarray(Q, dtype=np.float32).reshape(-1, X.shape[1])
    tid = np.ascontiguousarray(tid, dtype=np.uint64)
    nq = Q.shape[0]
    part_off = np.zeros(len(partitions) + 1, dtype=np.uint64)
    for i, p in enumerate(partitions):
        part_off[i + 1] = part_off[i] + len(p)
    part_rows = (np.concatenate([np.asarray(p, dtype=np.uint32) for p in partitions])
                 if len(partitions) and part_off[-1] > 0 else np.zeros(1, dtype=np.uint32))
    part_rows = np.ascontiguousarray(part_rows, dtype=np.uint32)
    qoff = np.zeros(nq + 1, dtype=np.uint64)
    for i, ps in enumerate(query_parts):
        qoff[i + 1] = qoff[i] + len(ps)
    qpart = np.ascontiguousarray(
        np.concatenate([np.asarray(list(ps), dtype=np.uint32) for ps in query_parts]) if qoff[-1] > 0
        else np.zeros(1, dtype=np.uint32), dtype=np.uint32)
    out_tid = np.empty((nq, k), dtype=np.uint64)
    out_row = np.empty((nq, k), dtype=np.uint32)
    out_dist = np.empty((nq, k), dtype=np.float64)
    out_sum = np.empty((nq, k), dtype=np.float32)
    out_count = np.zeros(nq, dtype=np.int32)
    f_l2, f_ip = (ref.fn_ptrs() if ref is not None else (None, None))
    rc = lib(variant).lib.orc_search_batch(
        X.ctypes.data, X.shape[1], tid.ctypes.data, X.shape[0], part_off.ctypes.data, part_rows.ctypes.data,
        nq, Q.ctypes.data, qoff.ctypes.data, qpart.ctypes.data, metric, k, int(dedup), f_l2, f_ip, nthreads,
        out_tid.ctypes.data, out_row.ctypes.data, out_dist.ctypes.data, out_sum.ctypes.data,
        out_count.ctypes.data)
    if rc != 0:
        raise RuntimeError(f"orc_search_batch failed: {rc}")
    return out_tid, out_row, out_dist, out_sum, out_count


# ----------------------------------------------------------------------------------------------
# IVFFlat

def ivf_assign(X: np.ndarray, centers: np.ndarray, metric: int) -> np.ndarray:
    X = np.ascontiguousarray(X, dtype=np.float32)
    centers = np.ascontiguousarray(centers, dtype=np.float32)
    out = np.empty(X.shape[0], dtype=np.uint32)
    lib().lib.orc_ivf_assign(X.ctypes.data, X.shape[1], X.shape[0], centers.ctypes.data, centers.shape[0], metric,
                             out.ctypes.data)
    return out


def ivf_lists_csr(row_list: np.ndarray, lists: int):
    order = np.argsort(row_list, kind="stable").astype(np.uint32)
    counts = np.bincount(row_list, minlength=lists)
    off = np.zeros(lists + 1, dtype=np.uint64)
    off[1:] = np.cumsum(counts)
    return off, np.ascontiguousarray(order)


def ivf_order_lists(centers: np.ndarray, q: Optional[np.ndarray], metric: int, max_probes: int):
    """GetScanLists (ivfscan.c:36-107). q=None is the NULL query."""
    centers = np.ascontiguousarray(centers, dtype=np.float32)
    lists, dim = centers.shape
    null = q is None
    qq = np.zeros(dim, dtype=np.float32) if null else np.ascontiguousarray(q, dtype=np.float32)
    n = min(max_probes, lists)
    out = np.empty(n, dtype=np.int32)
    outd = np.empty(n, dtype=np.float64)
    got = lib().lib.orc_ivf_order_lists(centers.ctypes.data, lists, dim, qq.ctypes.data, metric, max_probes,
                                        int(null), out.ctypes.data, outd.ctypes.data)
    return out[:got], outd[:got]


def ivf_scan(X, tid, centers, list_off, list_rows, q, metric, probes, max_probes=None, allowed=None, limit=None):
    """ivfflatgettuple stream (ivfscan.c:339-389): yields batches; each batch is the sorted tuples of the
    next `probes` lists; iterative scan continues up to max_probes.  Returns list of (tid, row, dist)
    arrays, one per batch; with `limit` stops once that many tuples were produced (executor LIMIT)."""
    X = np.ascontiguousarray(X, dtype=np.float32)
    tid = np.ascontiguousarray(tid, dtype=np.uint64)
    lists = centers.shape[0]
    probes = min(probes, lists)                       # ivfscan.c:262-263
    maxp = probes if max_probes is None else min(max(max_probes, probes), lists)  # :257-266
    order, _ = ivf_order_lists(centers, q, metric, maxp)
    null = q is None
    qq = np.zeros(X.shape[1], dtype=np.float32) if null else np.ascontiguousarray(q, dtype=np.float32)
    amask = None if allowed is None else np.ascontiguousarray(allowed, dtype=np.uint8)
    batches = []
    produced = 0
    idx = 0
    while idx < maxp:
        pl = np.ascontiguousarray(order[idx: idx + probes], dtype=np.int32)
        idx += len(pl)
        cap = int(sum(int(list_off[l + 1] - list_off[l]) for l in pl)) + 1
        ot = np.empty(cap, dtype=np.uint64)
        orow = np.empty(cap, dtype=np.uint32)
        od = np.empty(cap, dtype=np.float64)
        n = lib().lib.orc_ivf_scan_batch(X.ctypes.data, X.shape[1], tid.ctypes.data, list_off.ctypes.data,
                                         list_rows.ctypes.data, pl.ctypes.data, len(pl),
                                         None if amask is None else amask.ctypes.data, qq.ctypes.data, metric,
                                         int(null), cap, ot.ctypes.data, orow.ctypes.data, od.ctypes.data)
        batches.append((ot[:n].copy(), orow[:n].copy(), od[:n].copy()))
        produced += n
        if limit is not None and produced >= limit:
            break
    return batches


def gen_sift(rows: np.ndarray, dim: int = 128, seed: int = 1234, n_clusters: int = 4096, nthreads: int = 1,
             out: Optional[np.ndarray] = None) -> np.ndarray:
    """SIFT-shaped synthetic rows on the CPU (synthgen.c), bit-identical to the package's torch generator
    (synth.sift_like); used by `bench.py --impl reference` so that arm needs neither the GPU nor librbacknn.so."""
    build()
    if "synthgen" not in _libs:
        L = C.CDLL(os.path.join(_HERE, "libsynthgen.so"))
        L.orc_gen_sift.restype = C.c_int
        L.orc_gen_sift.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_int]
        _libs["synthgen"] = L
    rows = np.ascontiguousarray(rows, dtype=np.int64)
    if out is None:
        out = np.empty((len(rows), dim), dtype=np.float32)
    assert out.flags.c_contiguous and out.shape == (len(rows), dim) and out.dtype == np.float32
    if _libs["synthgen"].orc_gen_sift(rows.ctypes.data, len(rows), dim, seed, n_clusters, out.ctypes.data, nthreads) != 0:
        raise MemoryError("orc_gen_sift")
    return out


def recall(truth: Iterable[int], pred: Iterable[int]) -> float:
    """compute_recall (common_function.py:1154-1160) over tuple ids."""
    t = np.ascontiguousarray(list(truth), dtype=np.uint64)
    p = np.ascontiguousarray(list(pred), dtype=np.uint64)
    if len(t) == 0:
        return 0.0
    return float(lib().lib.orc_recall(t.ctypes.data, len(t), p.ctypes.data if len(p) else None, len(p)))
