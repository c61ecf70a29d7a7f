"""GPU parity of the batched tensor-core engine (tcgen05, RK_ENGINE_TENSOR): it must return exactly
what the scan engine and the oracle return — integer-valued data bit for bit, float data within the
north-star tolerance — including the queries it hands back to the exact path."""
import numpy as np
import pytest

from oracle import oracle as orc
from tests.util import assert_topk_matches, load_index, make_dataset, oracle_role_search

pytestmark = pytest.mark.gpu


def _queries(ds, nq, seed=11):
    rng = np.random.default_rng(seed)
    rows = rng.integers(0, len(ds["X"]), nq)
    users = np.unique(ds["user_roles"][:, 0])
    return ds["X"][rows].copy(), users[rng.integers(0, len(users), nq)].astype(np.uint32)


@pytest.mark.parametrize("nq,k", [(5, 10), (128, 100), (300, 30)])
def test_tensor_engine_sift_bitmask_exact(nq, k):
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=20000, dim=128, rows_per_doc=20, rbac="random", num_users=40, num_roles=12)
    ix = load_index(ds, role_parts=False, combo_parts=False, k_max=256)
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
        ix.search_batch(Q, users, 300, engine=vr.ENGINE_TENSOR)        # k > 256
    ix.close()
