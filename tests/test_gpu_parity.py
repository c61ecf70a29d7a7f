"""GPU parity: the CUDA path (through the C-ABI) against the oracle on the same seeded inputs."""
import numpy as np
import pytest

from oracle import oracle as orc
from tests.util import assert_topk_matches, load_index, make_dataset, oracle_role_search

pytestmark = pytest.mark.gpu


def _queries(ds, nq, seed=11, users=None):
    rng = np.random.default_rng(seed)
    rows = rng.integers(0, len(ds["X"]), nq)
    all_users = np.unique(ds["user_roles"][:, 0]) if users is None else np.asarray(users)
    return ds["X"][rows].copy(), all_users[rng.integers(0, len(all_users), nq)].astype(np.uint32)


@pytest.mark.parametrize("k", [1, 10, 100])
def test_role_partition_exact_sift(k):
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=6000, dim=128, rbac="tree")
    ix = load_index(ds, metric=vr.METRIC_L2)
    ix.finalize()
    Q, users = _queries(ds, 64)
    got = ix.search_batch(Q, users, k, filter=vr.FILTER_PARTITION, layout=vr.LAYOUT_ROLE)
    exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, k, mode="role")
    assert_topk_matches(got, exp, exact=True)
    # every query vector is a corpus row the user may or may not see; distances are exact integers' roots
    assert np.all(got[3] > 0)


@pytest.mark.parametrize("mode,flt,layout", [("role", 0, 1), ("combination", 0, 2), ("bitmask", 1, 0), ("none", 2, 0)])
def test_multi_role_users_all_filters(mode, flt, layout):
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=5000, dim=64, rows_per_doc=10, rbac="random", num_users=30, num_roles=10)
    ix = load_index(ds, metric=vr.METRIC_L2)
    ix.finalize()
    Q, users = _queries(ds, 48)
    got = ix.search_batch(Q, users, 20, filter=flt, layout=layout)
    exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, 20, mode=mode)
    assert_topk_matches(got, exp, exact=True)


def test_partition_and_bitmask_modes_agree():
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=8000, dim=128, rbac="random", num_users=25, num_roles=9)
    ix = load_index(ds)
    ix.finalize()
    Q, users = _queries(ds, 40)
    a = ix.search_batch(Q, users, 50, filter=vr.FILTER_PARTITION, layout=vr.LAYOUT_ROLE)
    b = ix.search_batch(Q, users, 50, filter=vr.FILTER_PARTITION, layout=vr.LAYOUT_COMBINATION)
    c = ix.search_batch(Q, users, 50, filter=vr.FILTER_BITMASK)
    for x in (b, c):
        np.testing.assert_array_equal(a[0], x[0])
        np.testing.assert_array_equal(a[2], x[2])
        np.testing.assert_array_equal(a[3], x[3])


@pytest.mark.parametrize("dim", [3, 96, 300, 768])
def test_inner_product_float_data(dim):
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("cohere", n=3000, dim=dim, rows_per_doc=7, rbac="random", num_users=20, num_roles=8)
    ix = load_index(ds, metric=vr.METRIC_IP)
    ix.finalize()
    Q, users = _queries(ds, 32)
    got = ix.search_batch(Q, users, 25, filter=vr.FILTER_BITMASK)
    exp = oracle_role_search(ds, Q, users, orc.METRIC_IP, 25, mode="bitmask")
    assert_topk_matches(got, exp, exact=False)


def test_l2_float_data_and_unsorted_load_order():
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("cohere", n=4000, dim=128, rows_per_doc=9, rbac="tree")
    perm = np.random.default_rng(1).permutation(len(ds["X"]))      # documents arrive interleaved
    ix = load_index(ds, metric=vr.METRIC_L2, perm=perm)
    ix.finalize()
    Q, users = _queries(ds, 32)
    got = ix.search_batch(Q, users, 30, filter=vr.FILTER_PARTITION, layout=vr.LAYOUT_ROLE)
    exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, 30, mode="role")
    assert_topk_matches(got, exp, exact=False)


def test_edge_cases():
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=900, dim=16, rows_per_doc=3, rbac="tree", num_users=10, num_roles=5)
    ix = load_index(ds, k_max=1024)
    ix.finalize()
    Q, users = _queries(ds, 8)
    # unknown user -> no roles -> empty result (prefilter_role.py:120-124)
    t, d, x, c = ix.search_batch(Q[:2], np.array([99999, users[0]], dtype=np.uint32), 10, layout=vr.LAYOUT_ROLE)
    assert c[0] == 0 and np.all(np.isinf(x[0])) and np.all(t[0] == np.iinfo(np.uint64).max) and c[1] > 0
    # k larger than the visible set: count < k, padded
    got = ix.search_batch(Q, users, 1024, filter=vr.FILTER_BITMASK)
    exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, 1024, mode="bitmask")
    assert_topk_matches(got, exp, exact=True)
    assert np.all(got[3] < 1024)
    # dimension mismatch and non-finite input are rejected like pgvector does
    with pytest.raises(vr.RkError, match="different vector dimensions 16 and 15"):
        ix.search_batch(np.zeros((1, 15), np.float32), users[:1], 5)
    bad = Q[:1].copy(); bad[0, 3] = np.nan
    with pytest.raises(vr.RkError) as ei:
        ix.search_batch(bad, users[:1], 5)
    assert ei.value.code == -3
    with pytest.raises(vr.RkError):
        ix.search_batch(Q[:1], users[:1], 2000)      # k > k_max
    # ground-truth merge keeps duplicates (common_function.py:713-714)


def test_keep_duplicates_ground_truth_variant():
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=3000, dim=32, rows_per_doc=10, rbac="random", num_users=20, num_roles=6, m_roles=3)
    ix = load_index(ds)
    ix.finalize()
    Q, users = _queries(ds, 24)
    got = ix.search_batch(Q, users, 15, layout=vr.LAYOUT_ROLE, keep_duplicates=True)
    exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, 15, dedup=False, mode="role")
    assert_topk_matches(got, exp, exact=True)


def test_scan_iterator_contract():
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=2000, dim=128, rbac="tree")
    ix = load_index(ds)
    ix.finalize()
    Q, users = _queries(ds, 3)
    sc = ix.begin_scan()
    for i in range(3):                        # one scan handle, rescanned per query
        sc.rescan(Q[i], int(users[i]), k=12, filter=vr.FILTER_PARTITION, layout=vr.LAYOUT_ROLE)
        rows = list(sc)
        exp = oracle_role_search(ds, Q[i:i + 1], users[i:i + 1], orc.METRIC_L2, 12, mode="role")
        assert [r[0] for r in rows] == exp[0][0][:exp[4][0]].tolist()
        assert [r[2] for r in rows] == exp[2][0][:exp[4][0]].tolist()
        assert sc.gettuple() is None
    with pytest.raises(vr.RkError, match="different vector dimensions 128 and 5"):
        sc.rescan(np.zeros(5, np.float32), 1, k=3)
    sc.end()


def test_large_scan_many_segments_and_compactions():
    """One query over 200k rows: many segments per CTA, several buffer compactions, 148-way merge."""
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=200_000, dim=32, rows_per_doc=50, rbac="tree", num_users=4, num_roles=3)
    ix = load_index(ds, role_parts=False, combo_parts=False)
    ix.finalize()
    Q, users = _queries(ds, 4)
    got = ix.search_batch(Q, users, 100, filter=vr.FILTER_NONE)
    exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, 100, mode="none")
    assert_topk_matches(got, exp, exact=True)
    # descending-quality data: rows sorted so that every later row beats the running threshold
    order = np.argsort(-((ds["X"] - Q[0]) ** 2).sum(1), kind="stable")
    ds2 = dict(ds, X=ds["X"][order].copy())
    ix2 = load_index(ds2, role_parts=False, combo_parts=False)
    ix2.finalize()
    got = ix2.search_batch(Q[:1], users[:1], 100, filter=vr.FILTER_NONE)
    exp = oracle_role_search(ds2, Q[:1], users[:1], orc.METRIC_L2, 100, mode="none")
    assert_topk_matches(got, exp, exact=True)


@pytest.mark.parametrize("world", [2, 3])
def test_shard_exchange_form_and_merge(world):
    """G shards (document d -> shard d mod G) searched separately on one GPU, their packed exchange
    buffers concatenated exactly as an all-gather would lay them out, then rk_merge_shards with
    shard_stride_bytes: must equal the oracle over the whole corpus."""
    import torch
    import vectorsearch_rbac_b200 as vr
    from vectorsearch_rbac_b200 import sharded
    ds = make_dataset("sift", n=9000, dim=64, rows_per_doc=15, rbac="random", num_users=20, num_roles=8)
    Q, users = _queries(ds, 24)
    k, nq = 40, 24
    dev = torch.device("cuda", 0)
    dQ = torch.from_numpy(Q).to(dev)
    per = sharded.packed_bytes(nq, k)
    recv = torch.empty(per * world, dtype=torch.uint8, device=dev)
    shards = []
    for r in range(world):
        mine = sharded.local_row_mask(ds["doc"], r, world)
        local = dict(ds, X=ds["X"][mine], tid=ds["tid"][mine], doc=ds["doc"][mine])
        ix = load_index(local, role_parts=False, combo_parts=False)
        ix.finalize()
        shards.append(ix)
        s = sharded.ShardedSearcher(ix, r, world, device=dev)
        opts = sharded.make_opts(k, filter=vr.FILTER_BITMASK)
        s._cuda_local_search(dQ, users, opts, recv[r * per:(r + 1) * per])
        ix.synchronize()
    s = sharded.ShardedSearcher(shards[0], 0, world, device=dev)
    b = s._get_bufs(nq, k)
    s._cuda_merge(recv, nq, k, (b["tid"], b["doc"], b["dist"], b["cnt"]))
    shards[0].synchronize()
    got = (b["tid"].cpu().numpy().view(np.uint64), b["doc"].cpu().numpy(), b["dist"].cpu().numpy(), b["cnt"].cpu().numpy())
    exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, k, mode="bitmask")
    assert_topk_matches(got, exp, exact=True)
    # document ids ride along
    row_of_tid = {int(t): i for i, t in enumerate(ds["tid"])}
    for q in range(nq):
        for i in range(int(got[3][q])):
            assert got[1][q, i] == ds["doc"][row_of_tid[int(got[0][q, i])]]
    for ix in shards:
        ix.close()


@pytest.mark.parametrize("rbac", ["tree", "random"])
def test_document_order_by_class_does_not_change_results(rbac):
    """rk_finalize orders the documents (= the rows) by (role mask, id) — option class_order, default 1 — so that what a user may
    see is a few runs of rows; with it the planner cuts a masked scan inside those runs (no per-extent mask test).  Results are
    those of the by-id order and of the oracle, bit for bit, on every engine / filter; tree RBAC gives a few runs per user
    (the run-based plan: fewer segments than extents walked), random RBAC thousands (the per-extent mask path)."""
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=60000, dim=64, rows_per_doc=10, rbac=rbac, num_users=60, num_roles=20)
    rng = np.random.default_rng(9)
    users = np.unique(ds["user_roles"][:, 0])
    Q = ds["X"][rng.integers(0, len(ds["X"]), 48)].copy()
    qu = users[rng.integers(0, len(users), len(Q))].astype(np.uint32)
    out = {}
    for co in (0, 2):
        ix = load_index(ds, k_max=64)
        ix.set_option("class_order", co)
        ix.finalize()
        res = []
        for flt, lay, eng in ((vr.FILTER_BITMASK, 0, vr.ENGINE_SCAN), (vr.FILTER_BITMASK, 0, vr.ENGINE_TENSOR),
                              (vr.FILTER_PARTITION, vr.LAYOUT_ROLE, vr.ENGINE_SCAN), (vr.FILTER_PARTITION, vr.LAYOUT_COMBINATION, vr.ENGINE_SCAN),
                              (vr.FILTER_NONE, 0, vr.ENGINE_SCAN)):
            res.append(ix.search_batch(Q, qu, 40, filter=flt, layout=lay, engine=eng))
        res.append(ix.search_batch(Q[:1], qu[:1], 40, filter=vr.FILTER_BITMASK, engine=vr.ENGINE_SCAN))      # fused single-query launch
        out[co] = res
        if co == 2:
            exp = oracle_role_search(ds, Q, qu, orc.METRIC_L2, 40, mode="bitmask")
            assert_topk_matches(res[0], exp, exact=True)
        ix.close()
    for a, b in zip(out[0], out[2]):
        for x, y in zip(a, b):
            np.testing.assert_array_equal(x, y)
