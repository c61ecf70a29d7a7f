"""Shared helpers for the parity tests: small seeded datasets loaded into both the engine and the
oracle, and the tie-aware top-k comparison the north star defines (ids bit-exact except for ties
within 1e-5 relative fp32 distance)."""
from __future__ import annotations

import numpy as np
import torch

from oracle import oracle as orc

RTOL = 1e-5   # BASELINE.json north_star: "ties within a stated fp32 distance tolerance of 1e-5 relative"


def make_dataset(kind="sift", n=6000, dim=128, rows_per_doc=20, num_users=40, num_roles=12, rbac="tree", seed=3,
                 m_roles=3, tree=(3, 2, 3)):
    import vectorsearch_rbac_b200.synth as synth
    X = synth.corpus(kind, torch.arange(n), dim, seed=seed).numpy()
    tid, doc = synth.row_ids(n, rows_per_doc)
    doc_ids = np.unique(doc)
    if rbac == "tree":
        ur, dr = synth.tree_rbac(num_users, num_roles, doc_ids.tolist(), h=tree[0], b0=tree[1], b1=tree[2], seed=seed)
    else:
        ur, dr = synth.random_rbac(num_users, num_roles, doc_ids.tolist(), m_roles=m_roles,
                                   m_perms=max(2, 3 * len(doc_ids) // num_roles), seed=seed)
    return dict(X=X, tid=tid, doc=doc, user_roles=ur, doc_roles=dr, dim=dim)


def load_index(ds, metric=0, k_max=128, role_parts=True, combo_parts=True, device=0, perm=None):
    import vectorsearch_rbac_b200 as vr
    ix = vr.Index(ds["dim"], metric=metric, device=device, k_max=k_max)
    X, tid, doc = ds["X"], ds["tid"], ds["doc"]
    if perm is not None:
        X, tid, doc = X[perm], tid[perm], doc[perm]
    half = len(X) // 2
    ix.load_rows(X[:half], tid[:half], doc[:half])      # chunked load
    ix.load_rows(X[half:], tid[half:], doc[half:])
    ix.set_doc_roles(ds["doc_roles"][:, 0], ds["doc_roles"][:, 1])
    ix.set_user_roles(ds["user_roles"][:, 0], ds["user_roles"][:, 1])
    if role_parts:
        ix.build_role_partitions()
    if combo_parts:
        ix.build_combination_partitions()
    return ix


def oracle_role_search(ds, Q, users, metric, k, dedup=True, mode="role"):
    """Oracle result for the role-partition / combination / bitmask semantics."""
    rb = orc.Rbac(ds["doc_roles"], ds["user_roles"])
    if mode == "role":
        parts = orc.role_partition_rows(ds["doc"], rb)
        keys = sorted(parts)
        idx = {r: i for i, r in enumerate(keys)}
        plist = [parts[r] for r in keys]
        qparts = [[idx[r] for r in rb.roles_of_user(int(u)) if r in idx and len(parts[r])] for u in users]
    elif mode == "combination":
        parts = orc.combination_partition_rows(ds["doc"], rb)
        keys = sorted(parts)
        idx = {c: i for i, c in enumerate(keys)}
        plist = [parts[c] for c in keys]
        qparts = []
        for u in users:
            c = tuple(rb.roles_of_user(int(u)))
            qparts.append([idx[c]] if c in idx and len(parts[c]) else [])
    elif mode == "bitmask":
        uniq = sorted(set(int(u) for u in users))
        idx = {u: i for i, u in enumerate(uniq)}
        plist = [orc.allowed_rows(ds["doc"], rb, u) for u in uniq]
        qparts = [[idx[int(u)]] if len(plist[idx[int(u)]]) else [] for u in users]
    elif mode == "none":
        plist = [np.arange(len(ds["X"]), dtype=np.uint32)]
        qparts = [[0] for _ in users]
    else:
        raise ValueError(mode)
    return orc.search_batch(ds["X"], ds["tid"], plist, qparts, Q, metric, k, dedup=dedup)


def assert_topk_matches(got, exp, exact, ext=None, rtol=RTOL):
    """got = (tid, doc, dist, count) from the engine; exp = (tid, row, dist, sum, count) from the oracle.
    exact=True (integer-valued data): ids and float8 distances bit-identical.
    exact=False: distances within rtol; ids identical except where the oracle's distances are within
    rtol of each other (ties)."""
    g_tid, _, g_dist, g_cnt = got
    e_tid, _, e_dist, _, e_cnt = exp
    np.testing.assert_array_equal(g_cnt, e_cnt)
    if exact:
        np.testing.assert_array_equal(g_tid, e_tid)
        np.testing.assert_array_equal(g_dist.view(np.uint64), e_dist.view(np.uint64))
        return
    for q in range(len(g_cnt)):
        c = int(g_cnt[q])
        gd, ed = g_dist[q, :c], e_dist[q, :c]
        np.testing.assert_allclose(gd, ed, rtol=rtol, atol=1e-30)
        assert np.all(np.diff(gd) >= -abs(gd[:-1]) * rtol) if c > 1 else True
        for i in range(c):
            if g_tid[q, i] == e_tid[q, i]:
                continue
            # allowed only if the oracle has a near-tie covering this position
            tol = rtol * max(abs(ed[i]), 1e-30) * 4
            near = np.abs(e_dist[q, :c] - ed[i]) <= tol
            cand = set(e_tid[q, :c][near].tolist())
            boundary = i >= c - 1 or abs(ed[c - 1] - ed[i]) <= tol
            assert g_tid[q, i] in cand or boundary, (q, i, g_tid[q, i], e_tid[q, i], gd[i], ed[i])
