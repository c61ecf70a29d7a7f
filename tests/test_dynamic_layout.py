"""Dynamic-partition layout (C5): the greedy role splitter restated from the reference's AnonySys optimiser
(controller/dynamic_partition/hnsw/AnonySys_dynamic_partition.py:425-672) and its replay on the engine
(controller/dynamic_partition/search.py:31-111, 347-364: user -> role set -> partitions -> filtered scans ->
merge / dedup)."""
import numpy as np
import pytest

from oracle import oracle as orc
from tests.util import assert_topk_matches, load_index, make_dataset, oracle_role_search


def _docs_of_roles(dr):
    rd = {}
    for d, r in dr:
        rd.setdefault(int(r), set()).add(int(d))
    return rd


@pytest.mark.parametrize("alpha", [1.0, 1.5, 2.0, 3.0])
def test_split_roles_invariants(alpha):
    import vectorsearch_rbac_b200.synth as synth
    from vectorsearch_rbac_b200.dynamic_layout import split_roles
    docs = list(range(1, 601))
    ur, dr = synth.random_rbac(80, 16, docs, m_roles=3, m_perms=9 * len(docs) // 16, seed=5)
    parts, rp, info = split_roles(dr, ur, alpha=alpha)
    rd = _docs_of_roles(dr)
    n0 = len(set(int(d) for d in dr[:, 0]))
    stored = sum(len(p) for p in parts.values())
    assert abs(stored / n0 - info["storage_factor"]) < 1e-9
    # the loop checks the budget BEFORE a split (AnonySys_dynamic_partition.py:441), so only the last split may overshoot
    assert stored <= alpha * n0 + max(len(v) for v in rd.values()) + 1
    assert info["mean_docs_scanned_after"] <= info["mean_docs_scanned_before"]
    users = {}
    for u, r in ur:
        users.setdefault(int(u), set()).add(int(r))
    for roles in users.values():
        key = tuple(sorted(roles))
        need = set().union(*[rd.get(r, set()) for r in key])
        have = set().union(*[set(parts[p].tolist()) for p in rp[key]]) if rp[key] else set()
        assert need <= have                       # a user's partitions hold every document the user may see
    if alpha == 1.0:
        assert info["splits"] <= 1


def test_split_roles_disjoint_roles_split_completely():
    """Roles with pairwise disjoint documents: every split removes rows from the big partition at zero storage
    growth, so the splitter peels roles off until one role per partition is left."""
    from vectorsearch_rbac_b200.dynamic_layout import split_roles
    dr = np.asarray([(d, 1 + (d - 1) // 10) for d in range(1, 61)])          # 6 roles x 10 documents
    ur = np.asarray([(u, 1 + (u - 1) % 6) for u in range(1, 13)])
    parts, rp, info = split_roles(dr, ur, alpha=2.0)
    assert info["partitions"] == 6 and abs(info["storage_factor"] - 1.0) < 1e-9
    assert info["mean_docs_scanned_after"] == 10.0
    assert all(len(v) == 1 for v in rp.values())


@pytest.mark.gpu
@pytest.mark.parametrize("engine", ["scan", "tensor"])
def test_dynamic_layout_search_matches_oracle(engine):
    import vectorsearch_rbac_b200 as vr
    from vectorsearch_rbac_b200.dynamic_layout import apply_layout, split_roles
    ds = make_dataset("sift", n=16000, dim=128, rows_per_doc=20, rbac="random", num_users=60, num_roles=14)
    parts, rp, info = split_roles(ds["doc_roles"], ds["user_roles"], alpha=2.0)
    assert info["partitions"] > 1
    ix = load_index(ds, role_parts=False, combo_parts=False, k_max=128)
    apply_layout(ix, parts, rp)
    ix.finalize()
    rng = np.random.default_rng(4)
    Q = ds["X"][rng.integers(0, len(ds["X"]), 96)].copy()
    users = np.unique(ds["user_roles"][:, 0])[rng.integers(0, 60, 96)].astype(np.uint32)
    users[5] = 777777                                  # unknown user: empty result (prefilter_role.py:120-124)
    eng = vr.ENGINE_SCAN if engine == "scan" else vr.ENGINE_TENSOR
    got = ix.search_batch(Q, users, 50, filter=vr.FILTER_PARTITION, layout=vr.LAYOUT_DYNAMIC, engine=eng)
    # partitions overlap and impure ones are filtered per row: the answer is the user's visible rows, deduplicated
    exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, 50, mode="bitmask")
    assert_topk_matches(got, exp, exact=True)
    assert got[3][5] == 0
    ix.close()
