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
(`comb_role_weights`, uniform users).  The stage that splits whole role COMBINATIONS (combination_mode=True,
:615-620) is not restated.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

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
