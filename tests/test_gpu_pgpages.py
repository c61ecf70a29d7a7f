"""An ivfflat index FILE served from HBM: pages (written by tests/pgpages.py as pgvector lays them out) ->
rk_ivfpages_read -> rk_load_rows + rk_build_ivf_from_lists -> list-probe search, against the oracle's restatement of
ivfscan.c over the SAME stored lists.  Some tuples sit in a list that is not their nearest centre's (as after a centre drifted
or an operator re-clustered): ivfscan.c:128-168 scans the lists as stored, and so must the engine."""
import numpy as np
import pytest

from oracle import oracle as orc
from tests import pgpages as pg
from tests.util import make_dataset

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("metric,dim", [(0, 64), (1, 24)])
def test_index_file_to_hbm_matches_oracle_over_the_stored_lists(metric, dim):
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=5000, dim=dim, rows_per_doc=10, rbac="random", num_users=20, num_roles=8)
    n, lists = len(ds["X"]), 20
    rng = np.random.default_rng(5)
    heap_tid = ((np.arange(n, dtype=np.uint64) // 120) << np.uint64(16)) | (np.arange(n, dtype=np.uint64) % 120 + 1)
    centers = ds["X"][rng.choice(n, lists, replace=False)].copy()
    stored = orc.ivf_assign(ds["X"], centers, metric).astype(np.int64)
    moved = rng.choice(n, n // 20, replace=False)
    stored[moved] = (stored[moved] + 1 + rng.integers(0, lists - 1, len(moved))) % lists
    tuples = [[(ds["X"][i], int(heap_tid[i] >> 16), int(heap_tid[i] & 0xFFFF)) for i in np.flatnonzero(stored == l)] for l in range(lists)]
    rel = pg.write_index(centers, tuples)

    pages = vr.IvfflatPages(rel)
    c, list_start, row_list, x, tid = pages.read()
    pages.close()
    assert pages.tuples == n
    row_of_tid = {int(t): i for i, t in enumerate(heap_tid)}
    src = np.array([row_of_tid[int(t)] for t in tid])                 # the heap row behind every index tuple
    ix = vr.Index(dim, metric=metric, device=0, k_max=128)
    ix.load_rows(x, tid, ds["doc"][src])
    ix.set_doc_roles(ds["doc_roles"][:, 0], ds["doc_roles"][:, 1])
    ix.set_user_roles(ds["user_roles"][:, 0], ds["user_roles"][:, 1])
    ix.finalize()
    ix.build_ivf_from_lists(c, row_list)
    c2, rl2 = ix.ivf_info(n)
    np.testing.assert_array_equal(rl2, row_list)
    np.testing.assert_array_equal(c2, centers)

    off, rows = orc.ivf_lists_csr(stored.astype(np.uint32), lists)
    users = np.unique(ds["user_roles"][:, 0])
    Q = ds["X"][rng.integers(0, n, 10)].copy()
    qu = users[rng.integers(0, len(users), len(Q))].astype(np.uint32)
    rb = orc.Rbac(ds["doc_roles"], ds["user_roles"])
    k = 40
    for probes in (1, 4, lists):
        for flt in (vr.FILTER_NONE, vr.FILTER_BITMASK):
            got = ix.search_batch(Q, qu, k, probes=probes, filter=flt)
            for qi in range(len(Q)):
                allowed = None
                if flt == vr.FILTER_BITMASK:
                    allowed = np.zeros(n, dtype=np.uint8)
                    allowed[orc.allowed_rows(ds["doc"], rb, int(qu[qi]))] = 1
                (t, r, d), = orc.ivf_scan(ds["X"], heap_tid, centers, off, rows, Q[qi], metric, probes=probes, allowed=allowed)
                m = min(k, len(t))
                assert got[3][qi] == m
                np.testing.assert_array_equal(got[0][qi, :m], t[:m])
                np.testing.assert_array_equal(got[2][qi, :m].view(np.uint64), d[:m].view(np.uint64))
    # a row_list that names a list the index does not have is refused
    ix2 = vr.Index(dim, metric=metric, device=0, k_max=16)
    ix2.load_rows(x[:100], tid[:100], ds["doc"][src][:100])
    ix2.finalize()
    with pytest.raises(vr.RkError):
        ix2.build_ivf_from_lists(c, np.full(100, lists, dtype=np.uint32))
    ix2.close()
    ix.close()
