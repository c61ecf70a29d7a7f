"""GPU parity of the IVFFlat path (list choice, list scan with the fused permission mask, iterative
refill) against the oracle's restatement of pgvector/src/ivfscan.c, plus the recall properties the
reference's own TAP tests assert (pgvector/test/t/003_ivfflat_vector_build_recall.pl:104-116,
005_ivfflat_query_recall.pl:32-41)."""
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


def _allowed_mask(ds, user):
    rb = orc.Rbac(ds["doc_roles"], ds["user_roles"])
    m = np.zeros(len(ds["X"]), dtype=np.uint8)
    m[orc.allowed_rows(ds["doc"], rb, int(user))] = 1
    return m


def _oracle_ivf_topk(ds, centers, off, rows, q, metric, probes, k, allowed=None):
    (t, r, d), = orc.ivf_scan(ds["X"], ds["tid"], centers, off, rows, q, metric, probes=probes, allowed=allowed)
    return t[:k], d[:k]


@pytest.mark.parametrize("metric", [0, 1])
def test_ivf_given_centers_assignment_and_probe_parity(metric):
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=6000, dim=64, rows_per_doc=10, rbac="random", num_users=20, num_roles=8)
    lists = 24
    centers = ds["X"][np.random.default_rng(2).choice(len(ds["X"]), lists, replace=False)].copy()
    ix = load_index(ds, metric=metric, role_parts=False, combo_parts=False)
    ix.finalize()
    ix.build_ivf_from_centers(centers)
    c2, row_list = ix.ivf_info(len(ds["X"]))
    np.testing.assert_array_equal(c2, centers)
    exp_list = orc.ivf_assign(ds["X"], centers, metric)
    np.testing.assert_array_equal(row_list, exp_list)             # integer data: first minimum wins, bit-exact
    off, rows = orc.ivf_lists_csr(exp_list, lists)
    Q, users = _queries(ds, 12)
    k = 30
    for probes in (1, 5, lists):
        got_n = ix.search_batch(Q, users, k, probes=probes, filter=vr.FILTER_NONE)
        got_m = ix.search_batch(Q, users, k, probes=probes, filter=vr.FILTER_BITMASK)
        for qi in range(len(Q)):
            t, d = _oracle_ivf_topk(ds, centers, off, rows, Q[qi], metric, probes, k)
            n = len(t)
            assert got_n[3][qi] == n
            np.testing.assert_array_equal(got_n[0][qi, :n], t)
            np.testing.assert_array_equal(got_n[2][qi, :n].view(np.uint64), d.view(np.uint64))
            t, d = _oracle_ivf_topk(ds, centers, off, rows, Q[qi], metric, probes, k, allowed=_allowed_mask(ds, users[qi]))
            n = len(t)
            assert got_m[3][qi] == n
            np.testing.assert_array_equal(got_m[0][qi, :n], t)
            np.testing.assert_array_equal(got_m[2][qi, :n].view(np.uint64), d.view(np.uint64))
    # probes == lists scans every tuple: identical to the exact filtered scan
    exact = ix.search_batch(Q, users, k, probes=0, filter=vr.FILTER_BITMASK)
    full = ix.search_batch(Q, users, k, probes=lists, filter=vr.FILTER_BITMASK)
    for a, b in zip(exact, full):
        np.testing.assert_array_equal(a, b)
    ix.close()


def test_ivf_iterative_scan_and_null_query():
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=1000, dim=16, rows_per_doc=5, rbac="tree", num_users=10, num_roles=5)   # < k rows in total
    lists = 6
    centers = ds["X"][np.random.default_rng(3).choice(len(ds["X"]), lists, replace=False)].copy()
    ix = load_index(ds, role_parts=False, combo_parts=False, k_max=1024)
    ix.finalize()
    ix.build_ivf_from_centers(centers)
    off, rows = orc.ivf_lists_csr(orc.ivf_assign(ds["X"], centers, 0), lists)
    Q, users = _queries(ds, 2)
    k = 1024                                                       # >= any list: a batch is never truncated
    sc = ix.begin_scan()
    # probes=2, max_probes=5: batches of 2, 2, 1 lists, each sorted, streamed back to back (ivfscan.c:375-381)
    sc.rescan(Q[0], int(users[0]), k=k, probes=2, max_probes=5, filter=vr.FILTER_NONE)
    got = [(r[0], r[2]) for r in sc]
    exp = []
    for t, r, d in orc.ivf_scan(ds["X"], ds["tid"], centers, off, rows, Q[0], 0, probes=2, max_probes=5):
        exp += list(zip(t.tolist(), d.tolist()))
    assert got == exp
    # without max_probes only the first batch comes back
    sc.rescan(Q[1], int(users[1]), k=k, probes=3, filter=vr.FILTER_NONE)
    got = [r[0] for r in sc]
    (t, r, d), = orc.ivf_scan(ds["X"], ds["tid"], centers, off, rows, Q[1], 0, probes=3)
    assert got == t.tolist()
    # ORDER BY v <-> NULL: every distance is 0, lists in ascending number (ivfscan.c:181-200)
    sc.rescan(None, int(users[0]), k=k, probes=2, filter=vr.FILTER_NONE)
    res = list(sc)
    (t, r, d), = orc.ivf_scan(ds["X"], ds["tid"], centers, off, rows, None, 0, probes=2)
    assert sorted(x[0] for x in res) == sorted(t.tolist()) and all(x[2] == 0.0 for x in res)
    sc.end()
    ix.close()


@pytest.mark.parametrize("metric", [0, 1])
def test_ivf_kmeans_build_recall_properties(metric):
    """003_ivfflat_vector_build_recall.pl: 3-d uniform data, limit 20, 100 lists: recall >= 0.71 at probes=1,
    >= 0.95 at probes=10, == 1.00 at probes=lists; 005: a row queried with itself comes back first."""
    import vectorsearch_rbac_b200 as vr
    rng = np.random.default_rng(0)
    n, lists, k = 30000, 100, 20
    X = rng.random((n, 3), dtype=np.float32)
    ds = dict(X=X, tid=np.arange(1, n + 1, dtype=np.uint64), doc=(np.arange(n) // 10 + 1).astype(np.uint32),
              doc_roles=np.array([[d, 1] for d in range(1, n // 10 + 1)]), user_roles=np.array([[1, 1]]), dim=3)
    ix = load_index(ds, metric=metric, role_parts=False, combo_parts=False)
    ix.finalize()
    ix.build_ivf(lists, seed=42)
    centers, row_list = ix.ivf_info(n)
    assert centers.shape == (lists, 3) and np.isfinite(centers).all()
    # every row sits in the list of its nearest centre (ivfbuild.c:141-198), up to fp32 near-ties
    exp = orc.ivf_assign(X, centers, metric)
    assert (row_list != exp).mean() < 1e-3
    Q = rng.random((100, 3), dtype=np.float32)
    users = np.ones(100, dtype=np.uint32)
    truth = ix.search_batch(Q, users, k, probes=0, filter=vr.FILTER_NONE)
    rec = {}
    for probes in (1, 10, lists):
        got = ix.search_batch(Q, users, k, probes=probes, filter=vr.FILTER_NONE)
        rec[probes] = np.mean([orc.recall(truth[0][i], got[0][i][:got[3][i]]) for i in range(100)])
    assert rec[lists] == 1.0
    if metric == 0:
        assert rec[1] >= 0.71 and rec[10] >= 0.95, rec
        self_q = ix.search_batch(X[:50], np.ones(50, np.uint32), 1, probes=1, filter=vr.FILTER_NONE)
        assert np.array_equal(self_q[0][:, 0], ds["tid"][:50])
    else:
        assert rec[10] >= rec[1]
    ix.close()


def test_ivf_float_data_bitmask_and_errors():
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("cohere", n=5000, dim=96, rows_per_doc=8, rbac="random", num_users=16, num_roles=6)
    ix = load_index(ds, metric=vr.METRIC_L2, role_parts=False, combo_parts=False)
    ix.finalize()
    with pytest.raises(vr.RkError):
        ix.search_batch(ds["X"][:1], np.array([1], np.uint32), 5, probes=2)       # no index yet
    ix.build_ivf(32, seed=7)
    with pytest.raises(vr.RkError):
        ix.build_ivf(32, seed=7)                                                    # already built
    centers, row_list = ix.ivf_info(len(ds["X"]))
    off, rows = orc.ivf_lists_csr(row_list, 32)
    Q, users = _queries(ds, 10)
    k = 20
    got = ix.search_batch(Q, users, k, probes=6, filter=vr.FILTER_BITMASK)
    for qi in range(len(Q)):
        t, d = _oracle_ivf_topk(ds, centers, off, rows, Q[qi], 0, 6, k, allowed=_allowed_mask(ds, users[qi]))
        n = int(got[3][qi])
        # list choice may differ only where two centre distances tie within fp32 rounding; distances must agree
        if n == len(t):
            np.testing.assert_allclose(got[2][qi, :n], d, rtol=1e-5)
    ix.close()


@pytest.mark.parametrize("metric,dim", [(0, 100), (1, 100), (0, 772), (0, 1540)])
def test_ivf_probe_batched_path_matches_single_query_path(metric, dim, monkeypatch):
    """The batched probe (keys kernel + radix-select ordering) must choose the lists the one-block-per-query kernel
    (full sort of (key, list)) chooses: same arithmetic per (query, centre), ties to the lower list number
    (ivfscan.c:36-107 pushes lists in order and keeps the earlier one on equal distance).  Repeated centres make ties
    that straddle the probe cut; dims that are not a multiple of 128 exercise the column tail."""
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=4000, dim=dim, rows_per_doc=8, rbac="random", num_users=16, num_roles=6)
    rng = np.random.default_rng(5)
    base = ds["X"][rng.choice(len(ds["X"]), 75, replace=False)]
    centers = np.concatenate([base, base, base[:37], base]).copy()       # 262 lists, every centre two to four times
    lists = len(centers)
    ix = load_index(ds, metric=metric, role_parts=False, combo_parts=False)
    ix.finalize()
    ix.build_ivf_from_centers(centers)
    exp_list = orc.ivf_assign(ds["X"], centers, metric)
    off, rows = orc.ivf_lists_csr(exp_list, lists)
    Q, users = _queries(ds, 37)
    k = 25
    for probes in (1, 3, 6, 40, lists):
        ix.set_option("ivf_probe_single", 0)
        got_b = ix.search_batch(Q, users, k, probes=probes, filter=vr.FILTER_BITMASK)
        ix.set_option("ivf_probe_single", 1)
        got_s = ix.search_batch(Q, users, k, probes=probes, filter=vr.FILTER_BITMASK)
        for a, b in zip(got_b, got_s):
            np.testing.assert_array_equal(a, b)
        if dim == 100:
            for qi in range(0, len(Q), 6):
                t, d = _oracle_ivf_topk(ds, centers, off, rows, Q[qi], metric, probes, k, allowed=_allowed_mask(ds, users[qi]))
                n = len(t)
                assert got_b[3][qi] == n
                np.testing.assert_array_equal(got_b[0][qi, :n], t)
    # NULL query in a batch: every centre is at distance 0, so the choice is lists 0 .. probes-1 on both paths
    # (ivfscan.c:196-200, 181-185) and every tuple of those lists ties at 0: ascending tuple id
    ix.set_option("ivf_probe_single", 0)
    got_b = ix.search_batch(Q[:9], users[:9], k, probes=3, filter=vr.FILTER_NONE, null_query=True)
    ix.set_option("ivf_probe_single", 1)
    got_s = ix.search_batch(Q[:9], users[:9], k, probes=3, filter=vr.FILTER_NONE, null_query=True)
    for a, b in zip(got_b, got_s):
        np.testing.assert_array_equal(a, b)
    (t, r, d), = orc.ivf_scan(ds["X"], ds["tid"], centers, off, rows, None, metric, probes=3)
    n = min(k, len(t))
    assert got_b[3][0] == n
    np.testing.assert_array_equal(got_b[0][0, :n], t[:n])
    ix.close()


@pytest.mark.parametrize("world", [2, 3])
def test_ivf_sharded_lists_match_single_index(world):
    """centres replicated, every list's rows striped with their documents (d mod G).  G shards built
    from the same centres, searched with probes > 0, exchanged and merged must return exactly what one index holding
    the whole lists returns (every rank chooses the same lists; a row lives on one rank)."""
    import torch
    import vectorsearch_rbac_b200 as vr
    from vectorsearch_rbac_b200 import sharded
    ds = make_dataset("sift", n=9000, dim=64, rows_per_doc=15, rbac="random", num_users=20, num_roles=8)
    lists = 40
    centers = ds["X"][np.random.default_rng(3).choice(len(ds["X"]), lists, replace=False)].copy()
    full = load_index(ds, role_parts=False, combo_parts=False)
    full.finalize()
    full.build_ivf_from_centers(centers)
    np.testing.assert_array_equal(full.ivf_centers(), centers)
    Q, users = _queries(ds, 24)
    k, nq = 40, 24
    dev = torch.device("cuda", 0)
    dQ = torch.from_numpy(Q).to(dev)
    per = sharded.packed_bytes(nq, k)
    shards = []
    for r in range(world):
        mine = sharded.local_row_mask(ds["doc"], r, world)
        local = dict(ds, X=ds["X"][mine], tid=ds["tid"][mine], doc=ds["doc"][mine])
        ix = load_index(local, role_parts=False, combo_parts=False)
        ix.finalize()
        ix.build_ivf_from_centers(centers)
        shards.append(ix)
    for probes in (1, 4, lists):
        for flt in (vr.FILTER_BITMASK, vr.FILTER_NONE):
            recv = torch.empty(per * world, dtype=torch.uint8, device=dev)
            opts = sharded.make_opts(k, probes=probes, filter=flt)
            # probes == 4: the list choice comes from outside (each rank chooses for nq / G queries and the choices are
            # all-gathered — here: shard r chooses for its slice, the slices are concatenated)
            ext = None
            if probes == 4:
                P = min(probes, lists)
                ext = torch.zeros((nq, P), dtype=torch.int32, device=dev)
                cnt = (nq + world - 1) // world
                for r, ix in enumerate(shards):
                    lo, hi = min(nq, r * cnt), min(nq, (r + 1) * cnt)
                    if hi > lo:
                        ix.ivf_probe_device(dQ[lo:hi].data_ptr(), hi - lo, opts, ext[lo:hi].data_ptr())
                        ix.synchronize()
            for r, ix in enumerate(shards):
                s = sharded.ShardedSearcher(ix, r, world, device=dev)
                if ext is None:
                    s._cuda_local_search(dQ, users, opts, recv[r * per:(r + 1) * per])
                else:
                    key, doc, tid = sharded.split_packed(recv[r * per:(r + 1) * per], nq, k)
                    ix.search_batch_shard_probed(dQ.data_ptr(), nq, users, opts, ext.data_ptr(), key.data_ptr(), tid.data_ptr(),
                                                 doc.data_ptr())
                ix.synchronize()
            s = sharded.ShardedSearcher(shards[0], 0, world, device=dev)
            b = s._get_bufs(nq, k)
            s._cuda_merge(recv, nq, k, (b["tid"], b["doc"], b["dist"], b["cnt"]))
            shards[0].synchronize()
            got = (b["tid"].cpu().numpy().view(np.uint64), b["doc"].cpu().numpy(), b["dist"].cpu().numpy(), b["cnt"].cpu().numpy())
            exp = full.search_batch(Q, users, k, probes=probes, filter=flt)
            np.testing.assert_array_equal(got[3], exp[3])
            for q in range(nq):
                n = int(exp[3][q])
                np.testing.assert_array_equal(got[0][q, :n], exp[0][q, :n])
                np.testing.assert_array_equal(got[1][q, :n], exp[1][q, :n])
                np.testing.assert_array_equal(got[2][q, :n].view(np.uint64), exp[2][q, :n].view(np.uint64))
    for ix in shards + [full]:
        ix.close()


@pytest.mark.parametrize("kind,dim,metric,lists", [("sift", 64, 0, 300), ("sift", 128, 1, 256), ("cohere2", 96, 0, 512), ("cohere", 100, 1, 260),
                                                     ("cohere2", 768, 0, 1024)])
def test_tensor_core_assignment_equals_the_fp32_kernel_and_the_oracle(kind, dim, metric, lists):
    """The build's nearest-centre assignment (pgvector/src/ivfbuild.c:141-198, ivfkmeans.c:258-494) on the tensor cores
    (rk_assign_tc.cu: TF32 nomination, exact re-score, proof, refused rows through the fp32 kernel) must give, row for row,
    what the fp32 kernel gives; on integer-valued data both equal the oracle's orc_ivf_assign (first minimum wins)."""
    import vectorsearch_rbac_b200 as vr
    n = 20000 if dim <= 128 else 9000
    ds = make_dataset(kind, n=n, dim=dim, rows_per_doc=10, rbac="random", num_users=20, num_roles=8)
    rng = np.random.default_rng(4)
    centers = ds["X"][rng.choice(n, lists, replace=False)].copy()
    if kind != "sift":
        centers += rng.normal(0, 0.01, centers.shape).astype(np.float32)      # centres are means, not rows
    out = {}
    for tc in (1, 0):
        ix = load_index(ds, metric=metric, role_parts=False, combo_parts=False)
        ix.finalize()
        ix.set_option("ivf_assign_tc", tc)
        ix.build_ivf_from_centers(centers)
        st = ix.stats()
        assert st["ivf_assign_tc_rows"] == (n if tc else 0)
        if tc:
            # the proof must carry most rows (duplicated rows tie exactly with their centre's twin: those are refused)
            assert st["ivf_assign_fallback_rows"] <= n // 2, st["ivf_assign_fallback_rows"]
        out[tc] = ix.ivf_info(n)[1]
        ix.close()
    np.testing.assert_array_equal(out[1], out[0])
    if kind == "sift":
        np.testing.assert_array_equal(out[1], orc.ivf_assign(ds["X"], centers, metric))


def test_ivf_build_with_tensor_core_kmeans_reaches_the_same_recall():
    """rk_build_ivf end to end (k-means iterations + final assignment) with the tensor-core assignment on and off.  The
    assignment is bit-identical; the centre update sums rows with floating-point atomics, whose order differs from run to run,
    so centres agree to rounding and only rows on a knife's edge between two centres may change list."""
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("cohere2", n=30000, dim=64, rows_per_doc=10, rbac="random", num_users=20, num_roles=8)
    res = {}
    for tc in (1, 0):
        ix = load_index(ds, metric=0, role_parts=False, combo_parts=False)
        ix.finalize()
        ix.set_option("ivf_assign_tc", tc)
        ix.build_ivf(256, seed=7)
        res[tc] = ix.ivf_info(len(ds["X"]))
        if tc:
            assert ix.stats()["ivf_assign_tc_rows"] > len(ds["X"])              # the k-means iterations used it as well
        ix.close()
    np.testing.assert_allclose(res[1][0], res[0][0], rtol=1e-3, atol=1e-5)
    assert np.mean(res[1][1] == res[0][1]) > 0.995


def _splitmix(h):
    M = (1 << 64) - 1
    h &= M
    h ^= h >> 30; h = (h * 0xBF58476D1CE4E5B9) & M
    h ^= h >> 27; h = (h * 0x94D049BB133111EB) & M
    h ^= h >> 31
    return h


def test_kmeanspp_seeding_follows_initcenters():
    """pgvector/src/ivfkmeans.c:21-89 restated in numpy with the engine's documented random numbers (first centre
    splitmix(seed) % n, then u_i = splitmix(seed, i) / 2^53): weight = min(weight, l2_distance^2) as a float, choice = sum * u,
    next centre = first sample whose running total reaches the choice.  Integer-valued data: every weight and every running
    total is an exact integer, so the GPU's block-wise double sums pick exactly the rows the sequential loop picks."""
    import vectorsearch_rbac_b200 as vr
    n, dim, lists, seed = 3000, 32, 40, 12345
    ds = make_dataset("sift", n=n, dim=dim, rows_per_doc=10, rbac="random", num_users=20, num_roles=8)
    ix = load_index(ds, metric=0, role_parts=False, combo_parts=False)
    ix.set_option("class_order", 0)                  # index order = load order (by document id), the order the restatement walks
    ix.finalize()
    ix.set_option("ivf_sample_identity", 1)          # the sample = the table in index order (n <= 10000: everything is sampled)
    ix.set_option("ivf_kmeans_iters", 0)             # the seeds become the centres
    ix.build_ivf(lists, seed=seed)
    got = ix.ivf_centers()
    ix.close()
    M = (1 << 64) - 1
    S = ds["X"].astype(np.float64)
    chosen = [(_splitmix((seed * 0x9E3779B97F4A7C15 + 0x632BE59BD9B4E019) & M) >> 32) % n]
    weight = np.full(n, np.finfo(np.float32).max, dtype=np.float32)
    for i in range(lists - 1):
        d = np.sqrt(((S - S[chosen[i]]) ** 2).sum(axis=1))
        d2 = d * d
        upd = d2 < weight.astype(np.float64)
        weight[upd] = d2[upd].astype(np.float32)
        cum = np.cumsum(weight.astype(np.float64))
        u = (_splitmix((seed + 0x9E3779B97F4A7C15 * (i + 1)) & M) >> 11) / 9007199254740992.0
        choice = cum[-1] * u
        j = int(np.searchsorted(cum, choice, side="left"))
        chosen.append(min(j, n - 1))
    np.testing.assert_array_equal(got, ds["X"][chosen])
    assert len(set(chosen)) == lists                  # a chosen row's weight is 0: it is not drawn again


@pytest.mark.parametrize("metric", [0, 1])
def test_kmeanspp_build_is_reproducible_and_not_worse_than_first_rows(metric):
    """Same seed, same centres (no floating-point atomics in the seeding); and the k-means objective after the build is not worse
    than from the first sampled rows on clustered data."""
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("cohere2", n=20000, dim=48, rows_per_doc=10, rbac="random", num_users=20, num_roles=8)
    X = ds["X"] / np.linalg.norm(ds["X"], axis=1, keepdims=True) if metric else ds["X"]

    def build(pp, iters):
        ix = load_index(ds, metric=metric, role_parts=False, combo_parts=False)
        ix.finalize()
        ix.set_option("ivf_kmeanspp", pp)
        ix.set_option("ivf_kmeans_iters", iters)
        ix.build_ivf(128, seed=3)
        c, rl = ix.ivf_info(len(X))
        ix.close()
        return c, rl

    a, _ = build(1, 0)
    b, _ = build(1, 0)
    np.testing.assert_array_equal(a, b)
    assert len(np.unique(a, axis=0)) == 128
    obj = {}
    for pp in (1, 0):
        c, rl = build(pp, 25)
        cc = c / np.maximum(np.linalg.norm(c, axis=1, keepdims=True), 1e-30) if metric else c
        obj[pp] = float(((X - cc[rl]) ** 2).sum())
    assert obj[1] <= obj[0] * 1.02, obj
