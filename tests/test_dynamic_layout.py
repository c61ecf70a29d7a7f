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


def test_import_reference_layout_follows_the_loader():
    """load_result_to_database.py:226-228 (document indices -> ids, unknown indices dropped) and :253-284, :322-325
    (sorted(combination) -> every partition id the tracker entry iterates over)."""
    from vectorsearch_rbac_b200.dynamic_layout import import_reference_layout
    document_to_index = {101: 0, 102: 1, 103: 2, 104: 3}
    partition_assignment = {0: {0, 1, 2, 3, 9}, 1: {2, 3}, 2: set()}
    comb_role_tracker = {(2, 1): {0: {1}, 1: {2}}, (3,): {1: {3}}, frozenset({4}): [0]}
    parts, rp = import_reference_layout(partition_assignment, comb_role_tracker, document_to_index)
    assert {p: d.tolist() for p, d in parts.items()} == {0: [101, 102, 103, 104], 1: [103, 104], 2: []}
    assert rp == {(1, 2): [0, 1], (3,): [1], (4,): [0]}
    parts2, _ = import_reference_layout({5: [7, 3]}, {}, None)             # no index map: entries are document ids
    assert parts2[5].tolist() == [3, 7]


def test_split_combinations_second_stage():
    """combination_mode=True (AnonySys_dynamic_partition.py:597-636, :312-423): every combination still covers its documents,
    queries do not get slower, the budget is checked before a split."""
    import vectorsearch_rbac_b200.synth as synth
    from vectorsearch_rbac_b200.dynamic_layout import split_combinations, split_roles
    docs = list(range(1, 401))
    ur, dr = synth.random_rbac(60, 12, docs, m_roles=3, m_perms=9 * len(docs) // 12, seed=3)
    parts, rp, info1 = split_roles(dr, ur, alpha=1.6)
    p2, m2, info2 = split_combinations(dr, ur, parts, rp, alpha=2.0)
    rd = _docs_of_roles(dr)
    for comb, ps in m2.items():
        need = set().union(*[rd.get(r, set()) for r in comb])
        have = set().union(*[set(p2[p].tolist()) for p in ps]) if ps else set()
        assert need <= have
    assert info2["combination_splits"] >= 1 and info2["partitions"] == len(p2)
    assert info2["mean_docs_scanned_after"] < info2["mean_docs_scanned_before"]
    n0 = len(set(int(d) for d in dr[:, 0]))
    biggest = max(len(set().union(*[rd.get(r, set()) for r in c])) for c in m2)
    assert sum(len(p) for p in p2.values()) <= 2.0 * n0 + biggest
    # nothing to do when every partition serves at most one combination entirely
    dr1 = np.asarray([(d, 1 + (d - 1) // 10) for d in range(1, 31)])
    ur1 = np.asarray([(1, 1), (2, 2), (3, 3)])
    pa, ra, _ = split_roles(dr1, ur1, alpha=2.0)
    pb, rb, ib = split_combinations(dr1, ur1, pa, ra, alpha=2.0)
    assert ib["combination_splits"] == 0 and {k: v.tolist() for k, v in pb.items()} == {k: v.tolist() for k, v in pa.items()}


@pytest.mark.gpu
def test_second_stage_and_imported_layouts_search_like_the_reference():
    """A layout from the combination stage, and the same layout round-tripped through the reference's export form
    (partition_assignment over document indices + comb_role_tracker), replayed on the engine against the oracle's restatement
    of dynamic_partition_search (search.py:44-111, 347-364)."""
    import vectorsearch_rbac_b200 as vr
    from vectorsearch_rbac_b200.dynamic_layout import apply_layout, import_reference_layout, split_combinations, split_roles
    ds = make_dataset("sift", n=12000, dim=64, rows_per_doc=20, rbac="random", num_users=50, num_roles=10)
    parts, rp, _ = split_roles(ds["doc_roles"], ds["user_roles"], alpha=1.5)
    p2, m2, info = split_combinations(ds["doc_roles"], ds["user_roles"], parts, rp, alpha=2.2)
    assert info["combination_splits"] >= 1
    doc_ids = np.unique(ds["doc"])
    document_to_index = {int(d): i for i, d in enumerate(doc_ids)}
    exported = {p: {document_to_index[int(d)] for d in docs} for p, docs in p2.items()}
    tracker = {c: {p: set(c) for p in ps} for c, ps in m2.items()}
    p3, m3 = import_reference_layout(exported, tracker, document_to_index)
    assert {p: d.tolist() for p, d in p3.items()} == {p: d.tolist() for p, d in p2.items()} and m3 == m2
    ix = load_index(ds, role_parts=False, combo_parts=False, k_max=128)
    apply_layout(ix, p3, m3)
    ix.finalize()
    rng = np.random.default_rng(9)
    Q = ds["X"][rng.integers(0, len(ds["X"]), 64)].copy()
    users = np.unique(ds["user_roles"][:, 0])[rng.integers(0, 50, 64)].astype(np.uint32)
    got = ix.search_batch(Q, users, 40, filter=vr.FILTER_PARTITION, layout=vr.LAYOUT_DYNAMIC, engine=vr.ENGINE_SCAN)
    rb = orc.Rbac(ds["doc_roles"], ds["user_roles"])
    plist, qparts = orc.dynamic_partition_plan(ds["doc"], rb, {p: d.tolist() for p, d in p3.items()}, m3, users)
    exp = orc.search_batch(ds["X"], ds["tid"], plist, qparts, Q, orc.METRIC_L2, 40, dedup=True)
    assert_topk_matches(got, exp, exact=True)
    ix.close()
