"""GPU parity of the batched tensor-core engine (tcgen05, RK_ENGINE_TENSOR): it must return exactly
what the scan engine and the oracle return — integer-valued data bit for bit, float data within the
north-star tolerance — including the queries it hands back to the exact path."""
import numpy as np
import pytest

from oracle import oracle as orc
from tests.util import assert_topk_matches, load_index, make_dataset, oracle_role_search

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["one-cta", "cta-pair"])
def scoring_kernel(request, monkeypatch):
    """Every case runs on both scoring kernels: rk_bscan_kernel (cta_group::1, 128-query tiles) and
    rk_bscan2_kernel (cta_group::2 CTA pairs, 256-query tiles)."""
    monkeypatch.setenv("RK_TENSOR_PAIR", "1" if request.param == "cta-pair" else "0")
    return request.param


def _queries(ds, nq, seed=11):
    rng = np.random.default_rng(seed)
    rows = rng.integers(0, len(ds["X"]), nq)
    users = np.unique(ds["user_roles"][:, 0])
    return ds["X"][rows].copy(), users[rng.integers(0, len(users), nq)].astype(np.uint32)


@pytest.mark.parametrize("nq,k", [(5, 10), (128, 100), (300, 30), (40, 384)])
def test_tensor_engine_sift_bitmask_exact(nq, k):
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=20000, dim=128, rows_per_doc=20, rbac="random", num_users=40, num_roles=12)
    ix = load_index(ds, role_parts=False, combo_parts=False, k_max=384)
    ix.finalize()
    Q, users = _queries(ds, nq)
    got = ix.search_batch(Q, users, k, filter=vr.FILTER_BITMASK, engine=vr.ENGINE_TENSOR)
    st = ix.stats()
    exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, k, mode="bitmask")
    assert_topk_matches(got, exp, exact=True)
    ref = ix.search_batch(Q, users, k, filter=vr.FILTER_BITMASK, engine=vr.ENGINE_SCAN)
    for a, b in zip(got, ref):
        np.testing.assert_array_equal(a, b)
    assert st["tensor_exact"] == 1
    assert st["tensor_redo"] <= nq // 4          # the tensor path itself answered (ties can send a few back)
    ix.close()


@pytest.mark.parametrize("metric,dim", [(0, 96), (1, 768), (1, 100)])
def test_tensor_engine_float_data(metric, dim):
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("cohere", n=12000, dim=dim, rows_per_doc=8, rbac="random", num_users=30, num_roles=10)
    ix = load_index(ds, metric=metric, role_parts=False, combo_parts=False, k_max=256)
    ix.finalize()
    Q, users = _queries(ds, 150)
    k = 25
    got = ix.search_batch(Q, users, k, filter=vr.FILTER_BITMASK, engine=vr.ENGINE_TENSOR)
    ref = ix.search_batch(Q, users, k, filter=vr.FILTER_BITMASK, engine=vr.ENGINE_SCAN)
    for a, b in zip(got, ref):                   # exact re-scoring uses the scan's summation order: identical bits
        np.testing.assert_array_equal(a, b)
    exp = oracle_role_search(ds, Q, users, metric, k, mode="bitmask")
    assert_topk_matches(got, exp, exact=False)
    assert ix.stats()["tensor_exact"] == 0
    ix.close()


@pytest.mark.parametrize("metric,dim,kind", [(1, 768, "cohere"), (0, 512, "cohere"), (0, 256, "sift"), (1, 300, "cohere2")])
def test_bf16_and_tf32_nomination_give_the_scan_engines_answer(metric, dim, kind):
    """The CTA-pair kernel nominates in bf16 (tcgen05.mma kind::f16 on bf16 shadow copies) or TF32; the exact fp32 re-score and
    the sufficiency proof make the ANSWER the scan engine's either way (pgvector/src/vector.c:549-563, 596-606 arithmetic), only
    the number of refused queries may differ.  Integers up to 255 (SIFT shape) are exact in bf16: no error margin at all."""
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset(kind, n=14000, dim=dim, rows_per_doc=8, rbac="random", num_users=30, num_roles=10)
    ix = load_index(ds, metric=metric, role_parts=False, combo_parts=False, k_max=256)
    ix.finalize()
    ix.set_option("tensor_pair", 1)              # bf16 nomination exists on the CTA-pair kernel only
    Q, users = _queries(ds, 120)
    k = 40
    ref = ix.search_batch(Q, users, k, filter=vr.FILTER_BITMASK, engine=vr.ENGINE_SCAN)
    exp = oracle_role_search(ds, Q, users, metric, k, mode="bitmask")
    assert_topk_matches(ref, exp, exact=kind == "sift")
    for bf16 in (1, 0):
        ix.set_option("tensor_bf16", bf16)
        got = ix.search_batch(Q, users, k, filter=vr.FILTER_BITMASK, engine=vr.ENGINE_TENSOR)
        st = ix.stats()
        assert bool(st["tensor_plan"] & 32) and bool(st["tensor_plan"] & 64) == bool(bf16)      # pair kernel; bf16 as asked
        for a, b in zip(got, ref):
            np.testing.assert_array_equal(a, b)
        if kind == "sift":
            assert st["tensor_exact"] == 1
    ix.close()


def test_tensor_engine_unfiltered_and_small_visible_sets():
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=9000, dim=64, rows_per_doc=30, rbac="tree", num_users=12, num_roles=6)
    ix = load_index(ds, k_max=256)
    ix.finalize()
    Q, users = _queries(ds, 40)
    users[0] = 4242                                    # unknown user: sees nothing
    for flt, mode in ((vr.FILTER_NONE, "none"), (vr.FILTER_BITMASK, "bitmask"), (vr.FILTER_PARTITION, "role")):
        got = ix.search_batch(Q, users, 200, filter=flt, layout=vr.LAYOUT_ROLE, engine=vr.ENGINE_TENSOR)
        ref = ix.search_batch(Q, users, 200, filter=flt, layout=vr.LAYOUT_ROLE, engine=vr.ENGINE_SCAN)
        for a, b in zip(got, ref):
            np.testing.assert_array_equal(a, b)
    assert got[3][0] == 0
    with pytest.raises(vr.RkError):
        ix.search_batch(Q, users, 400, engine=vr.ENGINE_TENSOR)        # k > 384
    ix.close()


@pytest.mark.parametrize("plan", ["dense", "grouped", "atoms"])
def test_tensor_engine_plans_agree(plan, monkeypatch, scoring_kernel):
    """Dense / per-user-mask / per-document-class (atoms) plans walk different tiles with different query
    groupings; all must return the oracle's answer (a query appears in several atoms' units and its
    lists are merged per query)."""
    import vectorsearch_rbac_b200 as vr
    monkeypatch.setenv("RK_TENSOR_PLAN", plan)
    ds = make_dataset("sift", n=30000, dim=64, rows_per_doc=25, rbac="tree", num_users=60, num_roles=20)
    ix = load_index(ds, k_max=256)
    ix.finalize()
    Q, users = _queries(ds, 200, seed=5)
    users[3] = 999999                                  # unknown user
    k = 50
    got = ix.search_batch(Q, users, k, filter=vr.FILTER_BITMASK, engine=vr.ENGINE_TENSOR)
    exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, k, mode="bitmask")
    assert_topk_matches(got, exp, exact=True)
    assert got[3][3] == 0
    assert ix.stats()["tensor_plan"] & 15 == {"dense": 1, "grouped": 2, "atoms": 3}[plan]
    assert bool(ix.stats()["tensor_plan"] & 32) == (scoring_kernel == "cta-pair")
    ix.close()


@pytest.mark.parametrize("kind,metric,flt", [("sift", 0, "none"), ("cohere", 1, "bitmask")])
def test_tensor_engine_sampling_pass(kind, metric, flt, monkeypatch):
    """The sampling pass seeds every unit's threshold with the kp-th best value of every S-th tile;
    the answer must not change (thresholds are upper bounds), with and without it."""
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset(kind, n=60000, dim=32, rows_per_doc=50, rbac="random", num_users=20, num_roles=6)
    ix = load_index(ds, metric=metric, role_parts=False, combo_parts=False, k_max=256)
    ix.finalize()
    Q, users = _queries(ds, 64, seed=9)
    k = 10
    f = vr.FILTER_NONE if flt == "none" else vr.FILTER_BITMASK
    ix.set_option("tensor_plan", 1)                   # dense
    ix.set_option("tensor_sample", 0)
    base = ix.search_batch(Q, users, k, filter=f, engine=vr.ENGINE_TENSOR)
    units0 = ix.stats()["segments"]
    ix.set_option("tensor_sample", 1)
    ix.set_option("tensor_sdiv", 4)                   # sample every 4th.. tile already on this small corpus
    ix.set_option("tensor_smin", 64)
    got = ix.search_batch(Q, users, k, filter=f, engine=vr.ENGINE_TENSOR)
    assert ix.stats()["tensor_plan"] & 31 == 1 + 16    # dense plan, sampling pass ran
    assert ix.stats()["segments"] > units0
    for a, b in zip(got, base):
        np.testing.assert_array_equal(a, b)
    exp = oracle_role_search(ds, Q, users, metric, k, mode=flt)
    assert_topk_matches(got, exp, exact=(kind == "sift"))
    ix.close()


def test_engine_auto_picks_tensor_for_batches():
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=8000, dim=128, rows_per_doc=20, rbac="tree", num_users=30, num_roles=10)
    ix = load_index(ds, k_max=128)
    ix.finalize()
    Q, users = _queries(ds, 64, seed=2)
    got = ix.search_batch(Q, users, 20, filter=vr.FILTER_PARTITION, layout=vr.LAYOUT_ROLE)          # ENGINE_AUTO
    assert ix.stats()["tensor_plan"] != 0                                         # went through the tensor engine
    ref = ix.search_batch(Q, users, 20, filter=vr.FILTER_PARTITION, layout=vr.LAYOUT_ROLE, engine=vr.ENGINE_SCAN)
    for a, b in zip(got, ref):
        np.testing.assert_array_equal(a, b)
    small = ix.search_batch(Q[:8], users[:8], 20, filter=vr.FILTER_BITMASK)      # below the batch threshold: scan
    assert ix.stats()["tensor_plan"] == 0
    for a, b in zip(small, [x[:8] for x in ref]):
        np.testing.assert_array_equal(a, b)
    ix.close()


def test_tensor_engine_long_units_many_compactions(monkeypatch):
    """Long units with a full tile of queries: every query's candidate buffer fills and is compacted many
    times while the other warps keep appending (regression: a count read before the other warps' appends
    of the tile had landed let a buffer overflow into its neighbour's best candidates)."""
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("cohere", n=160000, dim=32, rows_per_doc=40, rbac="random", num_users=24, num_roles=6)
    ix = load_index(ds, metric=1, role_parts=False, combo_parts=False, k_max=128)
    ix.finalize()
    Q, users = _queries(ds, 128, seed=21)
    ix.set_option("tensor_plan", 1)
    ix.set_option("tensor_sample", 0)
    ix.set_option("tensor_tpu", 100000)                  # one unit per query tile: the whole corpus in one pass
    for flt in (vr.FILTER_NONE, vr.FILTER_BITMASK):
        got = ix.search_batch(Q, users, 100, filter=flt, engine=vr.ENGINE_TENSOR)
        ref = ix.search_batch(Q, users, 100, filter=flt, engine=vr.ENGINE_SCAN)
        for a, b in zip(got, ref):
            np.testing.assert_array_equal(a, b)
    ix.close()


@pytest.mark.parametrize("dim,lay", [(128, "bitmask"), (256, "role")])
def test_short_nomination_lists_on_exact_data_and_compaction_marks(dim, lay):
    """Integer rows + integer queries: from the second batch on the nomination list is k + 28 entries (option tensor_kp_exact);
    a batch of NON-integer queries that follows is still answered exactly (the proof refuses what the short list cannot cover);
    the compaction-round mark (tensor_round_mark) and tensor_kp_exact = -1 change nothing in the results.  One-CTA kernel (128
    floats) and CTA-pair kernel (256)."""
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=40000, dim=dim, rows_per_doc=20, rbac="tree", num_users=40, num_roles=12)
    ix = load_index(ds, combo_parts=False, k_max=128)
    ix.finalize()
    flt, layout, mode = (vr.FILTER_BITMASK, 0, "bitmask") if lay == "bitmask" else (vr.FILTER_PARTITION, vr.LAYOUT_ROLE, "role")
    k = 50
    for rep, (mark, kpx) in enumerate([(512, 28), (512, 28), (768, 28), (200, 4), (512, -1)]):
        ix.set_option("tensor_round_mark", mark)
        ix.set_option("tensor_kp_exact", kpx)
        Q, users = _queries(ds, 96, seed=20 + rep)
        got = ix.search_batch(Q, users, k, filter=flt, layout=layout, engine=vr.ENGINE_TENSOR)
        st = ix.stats()
        exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, k, mode=mode)
        assert_topk_matches(got, exp, exact=True)
        assert st["tensor_exact"] == 1 and st["tensor_redo"] <= 24
    # non-integer queries right after an exact batch: the short list meets a non-zero error bound
    ix.set_option("tensor_kp_exact", 28)
    Q, users = _queries(ds, 96, seed=31)
    Qf = (Q + np.random.default_rng(1).uniform(-0.4, 0.4, Q.shape)).astype(np.float32)
    got = ix.search_batch(Qf, users, k, filter=flt, layout=layout, engine=vr.ENGINE_TENSOR)
    assert ix.stats()["tensor_exact"] == 0
    ref = ix.search_batch(Qf, users, k, filter=flt, layout=layout, engine=vr.ENGINE_SCAN)
    for a, b in zip(got, ref):
        np.testing.assert_array_equal(a, b)
    exp = oracle_role_search(ds, Qf, users, orc.METRIC_L2, k, mode=mode)
    assert_topk_matches(got, exp, exact=False)
    # and the batch after it is back on the full list, then short again
    for _ in range(2):
        Q, users = _queries(ds, 96, seed=33)
        got = ix.search_batch(Q, users, k, filter=flt, layout=layout, engine=vr.ENGINE_TENSOR)
        assert_topk_matches(got, oracle_role_search(ds, Q, users, orc.METRIC_L2, k, mode=mode), exact=True)
    ix.close()
