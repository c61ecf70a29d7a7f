"""dynamic_layout.py — a dynamic-partition layout for the C5 workload (BASELINE.json configs[4]).

Restates the greedy splitter of the reference's AnonySys optimiser, single-role stage
(controller/dynamic_partition/hnsw/AnonySys_dynamic_partition.py:425-672, `split_comb_roles` with
combination_mode=False):

  * start with ONE partition holding every document, every role "homed" there (:428-438);
  * while the stored documents stay within alpha x the corpus (:441, :475): take the largest partition that
    still serves more than one role (:443-454), try to move each of its roles r into a NEW partition holding
    exactly r's documents (:523-529), shrink the old partition to the documents its remaining roles need
    (:547-562), score the move by  (relative query-time change) / (relative storage growth)  with the
    reference's sign rules (:590-614) and apply the best one that makes queries faster (:622-665);
  * a user's role set is served by the home partitions of its roles (`CombRolePartitions`,
    controller/dynamic_partition/load_result_to_database.py:293-299); rows of a partition the user may not
    see are dropped by the per-row predicate unless the partition is pure for that role set (:591-604).

What is NOT the reference's: the query-time model.  The optimiser prices a partition with an HNSW latency
model whose fitted constants (k, beta, a, b in `parameter_hnsw.json`, :795-832) are not shipped with the
repository.  Here a partition costs what it costs this engine — one pass over its rows — so
query time of a role set = sum of the sizes of its partitions, weighted by how many users hold the role set
(`comb_role_weights`, uniform users).

`split_combinations` restates the second stage (combination_mode=True, :597-636 with `update_comb_role_tracker_stage2`,
:312-423): whole role COMBINATIONS of the largest partition move into new partitions and every combination that shares a role
with the moved one re-chooses, among the partitions it used so far plus the new one, the cheapest subset that still covers
its documents.  `import_reference_layout` takes a layout the reference's own optimiser produced (`partition_assignment`,
`comb_role_tracker`, `document_to_index` as load_result_to_database.py:191-338 consumes them).
"""
from __future__ import annotations

from itertools import combinations as _subsets
from typing import Dict, Iterable, List, Mapping, Optional, Sequence, Set, Tuple

import numpy as np


def split_roles(doc_roles: np.ndarray, user_roles: np.ndarray, alpha: float = 2.0, max_splits: int = 10_000
                ) -> Tuple[Dict[int, np.ndarray], Dict[Tuple[int, ...], List[int]], dict]:
    """doc_roles [(document, role)], user_roles [(user, role)] -> (partitions, roleset_partitions, info).

    partitions          {partition id: sorted document ids}
    roleset_partitions  {sorted role tuple of some user: sorted partition ids that serve it}
    info                storage factor, number of splits, query cost before / after (rows scanned per user, mean)
    """
    doc_roles = np.asarray(doc_roles, dtype=np.int64).reshape(-1, 2)
    user_roles = np.asarray(user_roles, dtype=np.int64).reshape(-1, 2)
    docs = np.unique(doc_roles[:, 0])
    roles = np.unique(np.concatenate([doc_roles[:, 1], user_roles[:, 1]]))
    d_ix = {int(d): i for i, d in enumerate(docs)}
    r_ix = {int(r): i for i, r in enumerate(roles)}
    D, R = len(docs), len(roles)
    role_docs = np.zeros((R, D), dtype=bool)
    role_docs[[r_ix[int(r)] for r in doc_roles[:, 1]], [d_ix[int(d)] for d in doc_roles[:, 0]]] = True

    # distinct role sets of the users and how many users hold each (the reference's comb_role_weights)
    by_user: Dict[int, List[int]] = {}
    for u, r in user_roles:
        by_user.setdefault(int(u), []).append(int(r))
    combo_w: Dict[Tuple[int, ...], int] = {}
    for rs in by_user.values():
        key = tuple(sorted(set(rs)))
        combo_w[key] = combo_w.get(key, 0) + 1
    combos = sorted(combo_w)
    M = np.zeros((len(combos), R), dtype=bool)
    for c, key in enumerate(combos):
        M[c, [r_ix[r] for r in key]] = True
    w = np.asarray([combo_w[key] for key in combos], dtype=np.float64)

    part_docs: List[np.ndarray] = [role_docs.any(axis=0)]       # partition 0: every document some role can see
    home = np.zeros(R, dtype=np.int64)                          # role -> its partition
    n_docs0 = int(part_docs[0].sum())

    def sizes():
        return np.asarray([int(p.sum()) for p in part_docs], dtype=np.float64)

    def query_cost(home_, size_):
        # per role set: sum of the sizes of the distinct partitions its roles are homed in
        cost = np.zeros(len(combos))
        for pid in np.unique(home_):
            cost += size_[pid] * (M[:, home_ == pid].any(axis=1))
        return cost

    cost0 = float((query_cost(home, sizes()) * w).sum() / w.sum())
    splits = 0
    while sizes().sum() <= alpha * n_docs0 and splits < max_splits:
        sz = sizes()
        # the largest partition that still serves more than one role
        target = -1
        for pid in np.argsort(-sz, kind="stable"):
            if int((home == pid).sum()) > 1:
                target = int(pid)
                break
        if target < 0:
            break
        in_p = np.flatnonzero(home == target)
        cover = role_docs[in_p].sum(axis=0)                      # how many of the partition's roles need each document
        cnt_p = M[:, in_p].sum(axis=1)                           # how many of a role set's roles live in the partition
        before = float((query_cost(home, sz) * w).sum())
        storage = float(sz.sum())
        best = None
        for r in in_p:
            t_size = float(role_docs[r].sum())
            rem_size = float(((cover - role_docs[r]) > 0).sum())
            has_r = M[:, r]
            delta = np.where(has_r, -sz[target] + t_size + np.where(cnt_p > 1, rem_size, 0.0),
                             np.where(cnt_p > 0, -sz[target] + rem_size, 0.0))
            q_delta = float((delta * w).sum()) / before          # relative query-time change (:579-582)
            growth = (t_size + rem_size - sz[target]) / storage  # relative storage growth (:572-573)
            flag = -100.0 if growth < 0 else 1.0                 # (:585-588)
            if q_delta < 0:
                score = flag * q_delta / (growth + 1e-10)
                if best is None or score < best[0]:
                    best = (score, int(r))
        if best is None:
            break
        r = best[1]
        part_docs.append(role_docs[r].copy())
        home[r] = len(part_docs) - 1
        part_docs[target] = (cover - role_docs[r]) > 0
        splits += 1

    partitions = {pid: docs[np.flatnonzero(p)] for pid, p in enumerate(part_docs) if p.any()}
    roleset_partitions = {key: sorted({int(home[r_ix[r]]) for r in key if int(home[r_ix[r]]) in partitions}) for key in combos}
    sz = sizes()
    info = {"alpha": alpha, "splits": splits, "partitions": len(partitions), "storage_factor": float(sz.sum()) / max(n_docs0, 1),
            "mean_docs_scanned_before": cost0, "mean_docs_scanned_after": float((query_cost(home, sz) * w).sum() / w.sum())}
    return partitions, roleset_partitions, info


def apply_layout(index, partitions: Dict[int, Sequence[int]], roleset_partitions: Dict[Tuple[int, ...], Sequence[int]]) -> None:
    """Registers the layout with an engine.Index (rk_define_partition / rk_map_roleset_to_partitions); call before
    finalize()."""
    for pid, d in partitions.items():
        index.define_partition(int(pid), np.asarray(d, dtype=np.uint32))
    for roles, parts in roleset_partitions.items():
        index.map_roleset_to_partitions(list(roles), list(parts))


def import_reference_layout(partition_assignment: Mapping[int, Iterable[int]], comb_role_tracker: Mapping[Tuple[int, ...], object],
                            document_to_index: Optional[Mapping[int, int]] = None
                            ) -> Tuple[Dict[int, np.ndarray], Dict[Tuple[int, ...], List[int]]]:
    """A layout as the reference's optimiser hands it to its loader (initialize_partitions_and_role_mappings,
    controller/dynamic_partition/load_result_to_database.py:286-338) -> (partitions, roleset_partitions) for `apply_layout`.

    partition_assignment  {partition id: document INDICES}; `document_to_index` {document id: index} maps them back — an index
                          without a document is dropped, as create_and_populate_partition_table does (:226-228).  Without a
                          map the entries are taken as document ids.
    comb_role_tracker     {role combination: {partition id: roles served there}} (or any iterable of partition ids): the
                          loader stores sorted(combination) -> every partition id it iterates over
                          (insert_comb_role_partition_mapping, :253-284).
    """
    index_to_document = None if document_to_index is None else {int(v): int(k) for k, v in document_to_index.items()}
    partitions: Dict[int, np.ndarray] = {}
    for pid, idxs in partition_assignment.items():
        if index_to_document is None:
            docs = {int(i) for i in idxs}
        else:
            docs = {index_to_document[int(i)] for i in idxs if int(i) in index_to_document}
        partitions[int(pid)] = np.asarray(sorted(docs), dtype=np.int64)
    roleset_partitions: Dict[Tuple[int, ...], List[int]] = {}
    for comb, parts in comb_role_tracker.items():
        key = tuple(sorted(int(r) for r in (comb if isinstance(comb, (tuple, list, set, frozenset)) else (comb,))))
        have = roleset_partitions.setdefault(key, [])
        for pid in parts:                       # iterating a dict yields its partition ids, exactly what the loader inserts
            if int(pid) not in have:
                have.append(int(pid))
    for key in roleset_partitions:
        roleset_partitions[key].sort()
    return partitions, roleset_partitions


def split_combinations(doc_roles: np.ndarray, user_roles: np.ndarray, partitions: Mapping[int, Sequence[int]],
                       roleset_partitions: Mapping[Tuple[int, ...], Sequence[int]], alpha: float = 2.0, max_splits: int = 10_000,
                       max_subset: int = 4) -> Tuple[Dict[int, np.ndarray], Dict[Tuple[int, ...], List[int]], dict]:
    """Second stage of the optimiser (combination_mode=True), applied to the layout the first stage left.

    Loop while the stored documents stay within alpha x the corpus (:441): in the largest partition that serves more than one
    role combination ENTIRELY (:443-454) try, for every such combination c, a new partition holding exactly c's documents
    (:523-529); every combination sharing a role with c then picks, among its previous partitions and the new one, the
    covering subset with the smallest query cost (update_comb_role_tracker_stage2, :340-391 — here cost = documents scanned,
    the sum of the subset's sizes, in place of the HNSW latency model; subsets of up to `max_subset` partitions); the old
    partition keeps only what the roles still served there need (:547-562).  A move is scored by (relative query-time change
    of the combinations) / (relative storage growth) with the sign rule for shrinking storage (:585-598), must make queries
    faster, and the best one is applied (:622-665).
    """
    doc_roles = np.asarray(doc_roles, dtype=np.int64).reshape(-1, 2)
    user_roles = np.asarray(user_roles, dtype=np.int64).reshape(-1, 2)
    role_docs: Dict[int, Set[int]] = {}
    for d, r in doc_roles:
        role_docs.setdefault(int(r), set()).add(int(d))
    by_user: Dict[int, Set[int]] = {}
    for u, r in user_roles:
        by_user.setdefault(int(u), set()).add(int(r))
    weight: Dict[Tuple[int, ...], int] = {}
    for rs in by_user.values():
        key = tuple(sorted(rs))
        weight[key] = weight.get(key, 0) + 1
    combos = sorted(weight)
    comb_docs = {c: set().union(*[role_docs.get(r, set()) for r in c]) for c in combos}
    part: Dict[int, Set[int]] = {int(p): {int(x) for x in d} for p, d in partitions.items()}
    n_docs0 = len(set().union(*part.values())) if part else 0
    # tracker: combination -> {partition: roles it gets there}; a role is served by the smallest of the combination's
    # partitions that holds all its documents, else by all of them (:393-421)
    def assign_roles(c: Tuple[int, ...], pids: Sequence[int]) -> Dict[int, Set[int]]:
        tr: Dict[int, Set[int]] = {int(p): set() for p in pids}
        for r in c:
            full = [p for p in pids if role_docs.get(r, set()) <= part[p]]
            if full:
                tr[min(full, key=lambda p: (len(part[p]), p))].add(r)
            else:
                for p in pids:
                    tr[p].add(r)
        return tr
    tracker = {c: assign_roles(c, [int(p) for p in roleset_partitions.get(c, ()) if int(p) in part]) for c in combos}

    def cost_of(tr: Mapping[Tuple[int, ...], Mapping[int, Set[int]]], sizes: Mapping[int, int], which: Iterable[Tuple[int, ...]]) -> float:
        return float(sum(weight[c] * sum(sizes[p] for p in tr[c]) for c in which))

    def best_cover(c: Tuple[int, ...], cand: Sequence[int], pt: Mapping[int, Set[int]]) -> Optional[Tuple[int, ...]]:
        need, best, best_cost = comb_docs[c], None, None
        for n in range(1, min(len(cand), max_subset) + 1):
            for sub in _subsets(sorted(cand), n):
                if not need <= set().union(*[pt[p] for p in sub]):
                    continue
                cst = sum(len(pt[p]) for p in sub)
                if best_cost is None or cst < best_cost:
                    best, best_cost = sub, cst
        return best

    cost0 = cost_of(tracker, {p: len(d) for p, d in part.items()}, combos) / max(1, sum(weight.values()))
    splits = 0
    while sum(len(d) for d in part.values()) <= alpha * n_docs0 and splits < max_splits:
        target, whole = None, []
        for pid in sorted(part, key=lambda p: (-len(part[p]), p)):
            whole = [c for c in combos if pid in tracker[c] and set(tracker[c][pid]) == set(c)]      # served entirely there
            if len(whole) > 1:
                target = pid
                break
        if target is None:
            break
        new_pid = max(part) + 1
        involved = [c for c in combos if target in tracker[c]]
        sizes = {p: len(d) for p, d in part.items()}
        before = cost_of(tracker, sizes, involved)
        storage = float(sum(sizes.values()))
        best = None
        for c in whole:
            pt = dict(part)
            pt[new_pid] = set(comb_docs[c])
            tr = dict(tracker)
            for a in combos:
                if not set(a) & set(c):
                    continue
                if set(tr[a]) == {new_pid}:
                    continue
                sub = best_cover(a, list(set(tr[a]) | {new_pid}), pt)
                if sub is not None:
                    tr[a] = {p: set() for p in sub}
                    for r in a:
                        full = [p for p in sub if role_docs.get(r, set()) <= pt[p]]
                        if full:
                            tr[a][min(full, key=lambda p: (len(pt[p]), p))].add(r)
                        else:
                            for p in sub:
                                tr[a][p].add(r)
            still = set()
            for a in combos:
                if target in tr[a]:
                    still |= tr[a][target]
            pt[target] = part[target] & set().union(*[role_docs.get(r, set()) for r in still]) if still else set()
            sz = {p: len(d) for p, d in pt.items()}
            for a in combos:                     # a combination never keeps a partition that went empty
                if target in tr[a] and not pt[target]:
                    tr[a] = {p: v for p, v in tr[a].items() if p != target}
            after = cost_of(tr, sz, involved)
            q_delta = (after - before) / before if before > 0 else 0.0
            growth = (sum(sz.values()) - storage) / storage if storage > 0 else 0.0
            flag = -100.0 if growth < 0 else 1.0
            if q_delta < 0:
                score = flag * q_delta / (growth + 1e-10)
                if best is None or score < best[0]:
                    best = (score, c, pt, tr)
        if best is None:
            break
        _, _, part, tracker = best
        part = {p: d for p, d in part.items() if d}
        splits += 1

    out_parts = {p: np.asarray(sorted(d), dtype=np.int64) for p, d in part.items() if d}
    out_map = {c: sorted(p for p in tracker[c] if p in out_parts) for c in combos}
    sizes = {p: len(d) for p, d in part.items()}
    info = {"alpha": alpha, "combination_splits": splits, "partitions": len(out_parts),
            "storage_factor": sum(sizes.values()) / max(n_docs0, 1), "mean_docs_scanned_before": cost0,
            "mean_docs_scanned_after": cost_of(tracker, sizes, combos) / max(1, sum(weight.values()))}
    return out_parts, out_map, info
