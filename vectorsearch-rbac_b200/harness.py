"""harness.py — the reference harness's Python entry points for this path, served by the engine.

The unchanged harness calls `search_func(user_id=, query_vector=, topk=, statistics_type=)` and expects
`(rows, seconds)` with rows = `(block_id, document_id, block_content, distance)` tuples in ascending distance
(basic_benchmark/common_function.py:1366-1385; controller/baseline/prefilter/prefilter_role.py:22-26, 162-195;
controller/baseline/prefilter/prefilter_combination_role.py:22-40; controller/dynamic_partition/search.py:31-111),
computes recall over `(document_id, block_id)` sets (common_function.py:1154-1160, 1394-1395) and caches ground
truth as a JSON list of `{'query': {...}, 'ground_truth': rows}` (common_function.py:831-849).

`Searcher` gives those functions over an engine.Index: same names, argument meaning and return shapes, so a
maintainer swaps `from controller.baseline.prefilter.prefilter_role import search_documents_role_partition` for
`searcher.search_documents_role_partition`.  `block_content` is not stored on the device: the
slot is filled from an optional host-side lookup (`content_of`), else b"".

There is no CPU fallback here: every call goes through librbacknn.so.
"""
from __future__ import annotations

import json
import os
import time
from typing import Callable, Iterable, List, Optional, Sequence, Tuple

import numpy as np

from . import engine as _engine


def parse_vector(query_vector) -> np.ndarray:
    """The harness passes vectors as pgvector text '[x1,x2,...]' (prefilter_role.py:162-172 interpolates the string
    into SQL) or as sequences; both become fp32.  NaN / Inf are rejected like `vector_in` does
    (pgvector/src/vector.c:102-113)."""
    if isinstance(query_vector, str):
        v = np.asarray(json.loads(query_vector), dtype=np.float32)
    else:
        v = np.asarray(query_vector, dtype=np.float32)
    if v.ndim != 1:
        raise ValueError("query vector must be one-dimensional")
    if not np.all(np.isfinite(v)):
        raise ValueError("NaN / infinite value not allowed in vector")
    return v


def compute_recall(true_results: Iterable, predicted_results: Iterable) -> float:
    """|GT ∩ pred| / |GT| over sets (common_function.py:1154-1160)."""
    true_set, predicted_set = set(true_results), set(predicted_results)
    return len(true_set & predicted_set) / len(true_set)


def result_keys(rows: Sequence[Tuple]) -> set:
    """The harness's identity of a result row: (document_id, block_id) (common_function.py:1394-1395)."""
    return set((r[1], r[0]) for r in rows)


class Searcher:
    """The reference's per-strategy search functions over one finalized engine.Index."""

    def __init__(self, index, content_of: Optional[Callable[[int, int], bytes]] = None):
        self.index = index
        self.content_of = content_of

    # -- one query ------------------------------------------------------------------------------
    def _search(self, user_id: int, query_vector, topk: int, flt: int, layout: int, probes: int = 0):
        q = parse_vector(query_vector)[None, :]
        t0 = time.time()
        tid, doc, dist, cnt = self.index.search_batch(q, [int(user_id)], int(topk), probes=probes, filter=flt, layout=layout)
        seconds = time.time() - t0
        n = int(cnt[0])
        rows = [(int(tid[0, i]), int(doc[0, i]), self.content_of(int(tid[0, i]), int(doc[0, i])) if self.content_of else b"",
                 float(dist[0, i])) for i in range(n)]
        return rows, seconds

    def search_documents_role_partition(self, user_id, query_vector, topk=5, statistics_type="sql"):
        """prefilter_role.py:22-26 — one exact scan per role partition of the user, merged, deduplicated, first k."""
        return self._search(user_id, query_vector, topk, _engine.FILTER_PARTITION, _engine.LAYOUT_ROLE)

    def search_documents_combination_partition(self, user_id, query_vector, topk=5, statistics_type="sql"):
        """prefilter_combination_role.py:22-40 — the partition of the user's exact role set."""
        return self._search(user_id, query_vector, topk, _engine.FILTER_PARTITION, _engine.LAYOUT_COMBINATION)

    def search_documents_rls(self, user_id, query_vector, topk=5, statistics_type="sql"):
        """controller/baseline/pg_row_security/row_level_security.py:55-65 — the whole table under the row-level policy."""
        return self._search(user_id, query_vector, topk, _engine.FILTER_BITMASK, 0)

    def dynamic_partition_search(self, user_id, query_vector, topk=5, statistics_type="sql"):
        """controller/dynamic_partition/search.py:31-111 — role set -> partitions, row filter unless pure, merge."""
        return self._search(user_id, query_vector, topk, _engine.FILTER_PARTITION, _engine.LAYOUT_DYNAMIC)

    # -- batches --------------------------------------------------------------------------------
    def ground_truth_func_batch(self, queries: Sequence[dict], batch: int = 1024) -> List[List[Tuple]]:
        """Exact filtered ground truth for a list of `{'user_id', 'query_vector', 'topk'}` dicts — the job of
        `ground_truth_func_batch` (common_function.py:917-1139: per-role FAISS IndexFlatL2, union, dedup keeping the
        smallest distance, sort, first k), here one batched pass of the tensor-core engine with the exact re-score."""
        out: List[List[Tuple]] = [None] * len(queries)
        by_k = {}
        for i, q in enumerate(queries):
            by_k.setdefault(int(q.get("topk", 5)), []).append(i)
        for k, idx in by_k.items():
            for a in range(0, len(idx), batch):
                part = idx[a:a + batch]
                Q = np.stack([parse_vector(queries[i]["query_vector"]) for i in part])
                users = [int(queries[i]["user_id"]) for i in part]
                tid, doc, dist, cnt = self.index.search_batch(Q, users, k, filter=_engine.FILTER_BITMASK)
                for j, i in enumerate(part):
                    n = int(cnt[j])
                    out[i] = [(int(tid[j, t]), int(doc[j, t]), b"", float(dist[j, t])) for t in range(n)]
        return out


def save_ground_truth_cache(queries: Sequence[dict], results: Sequence[Sequence[Tuple]], cache_file: str) -> None:
    """The harness's cache format (common_function.py:831-849): a JSON list of
    {'query': {'user_id', 'query_vector', 'topk'}, 'ground_truth': [[block_id, document_id, content, distance], ...]};
    bytes content is written as text so the file stays JSON."""
    data = []
    for q, rows in zip(queries, results):
        qv = q["query_vector"]
        data.append({
            "query": {"user_id": int(q["user_id"]), "query_vector": qv if isinstance(qv, str) else [float(x) for x in qv],
                      "topk": int(q.get("topk", 5))},
            "ground_truth": [[int(r[0]), int(r[1]), r[2].decode("utf-8", "replace") if isinstance(r[2], (bytes, bytearray)) else r[2],
                              float(r[3])] for r in rows]})
    with open(cache_file, "w") as f:
        json.dump(data, f)


def load_ground_truth_cache(cache_file: str) -> Tuple[List[dict], List[List[Tuple]]]:
    with open(cache_file) as f:
        data = json.load(f)
    return [d["query"] for d in data], [[tuple(r) for r in d["ground_truth"]] for d in data]


# ---- "pointer" ground-truth cache (basic_benchmark/compute_ground_truth.py:35-59, read by the C++ benches:
# acorn_benchmark/src/benchmark_utils.cpp:26-76) ------------------------------------------------------------------

def to_pointer_results(results: Iterable[Iterable]) -> List[List[List[int]]]:
    """Per query the [block_id, document_id] pairs of its result rows, in result order; rows that are empty, shorter than
    two fields or whose ids are not integers are skipped (compute_ground_truth.py:35-53)."""
    out: List[List[List[int]]] = []
    for rows in results:
        formatted: List[List[int]] = []
        for entry in rows or []:
            if not entry or len(entry) < 2:
                continue
            try:
                formatted.append([int(entry[0]), int(entry[1])])
            except (TypeError, ValueError):
                continue
        out.append(formatted)
    return out


def save_pointer_cache(results: Iterable[Iterable], cache_file: str) -> None:
    """JSON list (one entry per query) of [block_id, document_id] lists, indent 2 (compute_ground_truth.py:56-59) — the file
    the C++ benches load as their ground truth."""
    d = os.path.dirname(cache_file)
    if d:
        os.makedirs(d, exist_ok=True)
    with open(cache_file, "w", encoding="utf-8") as f:
        json.dump(to_pointer_results(results), f, indent=2)


def load_pointer_cache(cache_file: str) -> Optional[List[List[List[int]]]]:
    """None when the file is missing, is not JSON or does not hold a list (compute_ground_truth.py:62-73)."""
    if not os.path.exists(cache_file):
        return None
    try:
        with open(cache_file, encoding="utf-8") as f:
            data = json.load(f)
    except json.JSONDecodeError:
        return None
    return data if isinstance(data, list) else None


def pointer_cache_as_flat_pairs(pointer: Sequence[Sequence[Sequence[int]]], queries: Sequence[dict]) -> Optional[List[Tuple[int, int]]]:
    """What load_groundtruth_from_cache builds from the file (benchmark_utils.cpp:26-76): the first `topk` pairs of every
    query, flattened, as (document_id, block_id); None when the cache does not match the query list (wrong length, or a query
    with fewer than topk pairs) — the C++ reader then recomputes the ground truth."""
    if not queries or len(pointer) != len(queries):
        return None
    flat: List[Tuple[int, int]] = []
    for q, entry in zip(queries, pointer):
        topk = max(int(q.get("topk", 5)), 0)
        if not isinstance(entry, (list, tuple)) or len(entry) < topk:
            return None
        for pair in entry[:topk]:
            if not isinstance(pair, (list, tuple)) or len(pair) < 2:
                return None
            flat.append((int(pair[1]), int(pair[0])))
    return flat
