"""GPU parity pins for the paths round 1 left unpinned ("What's weak" #1): role masks of more than one
64-bit word through every engine, the IVFFlat per-row-mask fallback, the reference's golden vectors THROUGH the
C-ABI, the dynamic-partition search against a restatement of the reference's own search loop with a layout that does
not cover the user's visible set, maximum dimensions, the streaming half of ivfflatgettuple, and the boundary's
threading / validation rules.  Everything is compared with oracle/."""
import math
import threading

import numpy as np
import pytest

from oracle import oracle as orc
from tests.util import assert_topk_matches, load_index, make_dataset, oracle_role_search

pytestmark = pytest.mark.gpu

SEG_EXTENTS, SEG_ROWS, SEG_ROWMASK, SEG_IVFQ = 1 << 0, 1 << 2, 1 << 3, 1 << 4


def _queries(ds, nq, seed=11):
    rng = np.random.default_rng(seed)
    rows = rng.integers(0, len(ds["X"]), nq)
    users = np.unique(ds["user_roles"][:, 0])
    return ds["X"][rows].copy(), users[rng.integers(0, len(users), nq)].astype(np.uint32)


def _wide_dataset(roles, rbac, n=24000, dim=64, users=None, kind="sift"):
    """100 roles -> 2 mask words (what the BASELINE configs run), 300 roles -> 5.  Tree RBAC gives one role per user,
    so users whose only role sits in an upper mask word exist; random RBAC gives up to three roles anywhere."""
    tree = (4, 3, 4) if roles <= 120 else (5, 3, 4)
    return make_dataset(kind, n=n, dim=dim, rows_per_doc=10, rbac=rbac, num_users=users or roles * 2, num_roles=roles, tree=tree)


def _upper_word_users(ds, word):
    roles = np.unique(np.concatenate([ds["doc_roles"][:, 1], ds["user_roles"][:, 1]]))
    dense = {int(r): i for i, r in enumerate(roles)}
    return [int(u) for u, r in ds["user_roles"] if dense[int(r)] >= 64 * word]


# ---- (a) role masks of 2 and 5 words -------------------------------------------------------------------------------

@pytest.mark.parametrize("roles,rbac", [(100, "tree"), (100, "random"), (300, "tree"), (300, "random")])
def test_wide_role_masks_scan_engine(roles, rbac):
    """rk_scan.cu:148-153, 229-234 (mask loops over W words) and the extent tables of every layout."""
    import vectorsearch_rbac_b200 as vr
    ds = _wide_dataset(roles, rbac)
    ix = load_index(ds)
    ix.finalize()
    Q, users = _queries(ds, 96)
    up = _upper_word_users(ds, 1 if roles == 100 else 4)
    assert up, "dataset must hold users whose role sits in the last mask word"
    users[:8] = np.asarray(up[:8] * 8, dtype=np.uint32)[:8]
    for mode, flt, layout in (("role", 0, 1), ("combination", 0, 2), ("bitmask", 1, 0)):
        got = ix.search_batch(Q, users, 30, filter=flt, layout=layout, engine=vr.ENGINE_SCAN)
        exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, 30, mode=mode)
        assert_topk_matches(got, exp, exact=True)
    assert np.all(got[3][:8] > 0)
    ix.close()


@pytest.mark.parametrize("plan", [1, 2, 3])
@pytest.mark.parametrize("pair", [0, 1])
@pytest.mark.parametrize("roles,rbac", [(100, "tree"), (300, "tree"), (300, "random")])
def test_wide_role_masks_tensor_engine(roles, rbac, pair, plan):
    """Both tcgen05 kernels, all three plans: the epilogue's first-two-words vote, its W > 2 per-row pass
    (rk_bscan_dev.cuh / rk_bscan.cu "more than 128 roles") and the 4-stage shared-memory layout of the pair kernel."""
    import vectorsearch_rbac_b200 as vr
    ds = _wide_dataset(roles, rbac, n=20000)
    ix = load_index(ds, role_parts=False, combo_parts=False, k_max=128)
    ix.finalize()
    ix.set_option("tensor_pair", pair)
    ix.set_option("tensor_plan", plan)
    Q, users = _queries(ds, 160, seed=3)
    up = _upper_word_users(ds, 1 if roles == 100 else 4)
    users[:8] = np.asarray((up * 8)[:8], dtype=np.uint32)
    got = ix.search_batch(Q, users, 40, filter=vr.FILTER_BITMASK, engine=vr.ENGINE_TENSOR)
    st = ix.stats()
    exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, 40, mode="bitmask")
    assert_topk_matches(got, exp, exact=True)
    if plan != 3 or rbac == "tree":       # random RBAC has too many document classes for the atoms plan: planner falls back
        assert st["tensor_plan"] & 15 == plan
    assert bool(st["tensor_plan"] & 32) == bool(pair)
    assert st["tensor_redo"] <= 40
    ix.close()


@pytest.mark.parametrize("roles,rbac,kind", [(100, "tree", SEG_IVFQ), (100, "random", SEG_ROWMASK), (300, "tree", SEG_ROWMASK)])
def test_wide_role_masks_ivf_list_scan_paths(roles, rbac, kind):
    """<= 128 document classes: class-ordered lists (word c1 of the visible-class set covers classes 64..127);
    more: the per-row mask test RK_SEG_ROWMASK (rk_ivf.cu, rk_scan.cu:282-325).  The stats say which path ran."""
    import vectorsearch_rbac_b200 as vr
    ds = _wide_dataset(roles, rbac, n=16000, dim=32)
    lists = 20
    centers = ds["X"][np.random.default_rng(2).choice(len(ds["X"]), lists, replace=False)].copy()
    ix = load_index(ds, role_parts=False, combo_parts=False)
    ix.finalize()
    ix.build_ivf_from_centers(centers)
    off, rows = orc.ivf_lists_csr(orc.ivf_assign(ds["X"], centers, 0), lists)
    Q, users = _queries(ds, 24)
    up = _upper_word_users(ds, 1 if roles == 100 else 4)
    users[:6] = np.asarray((up * 6)[:6], dtype=np.uint32)
    rb = orc.Rbac(ds["doc_roles"], ds["user_roles"])
    k = 25
    for probes in (1, 4, lists):
        got = ix.search_batch(Q, users, k, probes=probes, filter=vr.FILTER_BITMASK)
        assert ix.stats()["segment_kinds"] == kind
        for qi in range(len(Q)):
            allowed = np.zeros(len(ds["X"]), dtype=np.uint8)
            allowed[orc.allowed_rows(ds["doc"], rb, int(users[qi]))] = 1
            (t, r, d), = orc.ivf_scan(ds["X"], ds["tid"], centers, off, rows, Q[qi], 0, probes=probes, allowed=allowed)
            n = min(k, len(t))
            assert got[3][qi] == n
            np.testing.assert_array_equal(got[0][qi, :n], t[:n])
            np.testing.assert_array_equal(got[2][qi, :n].view(np.uint64), d[:n].view(np.uint64))
    ix.close()


def test_ivf_many_classes_concurrent_scans():
    """the visible-class table is shared by every rk_scan handle.  More than 64 classes, two host threads,
    each with its own scan handle, many rescans: every answer equals the oracle's."""
    import vectorsearch_rbac_b200 as vr
    ds = _wide_dataset(100, "tree", n=12000, dim=32)
    lists = 12
    centers = ds["X"][np.random.default_rng(4).choice(len(ds["X"]), lists, replace=False)].copy()
    ix = load_index(ds, role_parts=False, combo_parts=False)
    ix.finalize()
    ix.build_ivf_from_centers(centers)
    off, rows = orc.ivf_lists_csr(orc.ivf_assign(ds["X"], centers, 0), lists)
    Q, users = _queries(ds, 40, seed=8)
    rb = orc.Rbac(ds["doc_roles"], ds["user_roles"])
    exp = []
    for qi in range(len(Q)):
        allowed = np.zeros(len(ds["X"]), dtype=np.uint8)
        allowed[orc.allowed_rows(ds["doc"], rb, int(users[qi]))] = 1
        (t, r, d), = orc.ivf_scan(ds["X"], ds["tid"], centers, off, rows, Q[qi], 0, probes=3, allowed=allowed)
        exp.append(t[:15].tolist())
    errors = []

    def worker(lo, hi):
        try:
            sc = ix.begin_scan()
            for qi in list(range(lo, hi)) * 3:
                sc.rescan(Q[qi], int(users[qi]), k=15, probes=3, filter=vr.FILTER_BITMASK)
                got = [sc.gettuple() for _ in range(len(exp[qi]))]
                if [g[0] for g in got] != exp[qi]:
                    errors.append(qi)
            sc.end()
        except Exception as e:      # noqa: BLE001
            errors.append(repr(e))

    th = [threading.Thread(target=worker, args=(0, 20)), threading.Thread(target=worker, args=(20, 40))]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errors, errors
    ix.close()


# ---- (c) the reference's golden vectors through rk_* ----------------------------------------------------------------

def _tiny_index(rows, metric, k_max=8):
    import vectorsearch_rbac_b200 as vr
    X = np.asarray(rows, dtype=np.float32)
    n = len(X)
    ix = vr.Index(X.shape[1], metric=metric, k_max=k_max)
    ix.load_rows(X, np.arange(1, n + 1, dtype=np.uint64), np.ones(n, dtype=np.uint32))
    ix.set_doc_roles([1], [1])
    ix.set_user_roles([1], [1])
    ix.build_role_partitions()
    ix.finalize()
    return ix


def test_vector_type_goldens_through_the_c_abi():
    """pgvector/test/expected/vector_type.out:373-429, each value produced by rk_search_batch on a one-row table."""
    import vectorsearch_rbac_b200 as vr
    u = np.array([1], dtype=np.uint32)

    def dist(row, q, metric, engine=vr.ENGINE_SCAN, flt=vr.FILTER_BITMASK):
        ix = _tiny_index([row], metric)
        t, d, x, c = ix.search_batch(np.asarray([q], np.float32), u, 1, filter=flt, layout=vr.LAYOUT_ROLE, engine=engine)
        ix.close()
        assert c[0] == 1 and t[0, 0] == 1
        return x[0, 0]

    for flt in (vr.FILTER_BITMASK, vr.FILTER_PARTITION, vr.FILTER_NONE):
        assert dist([3, 4], [0, 0], 0, flt=flt) == 5.0                              # :373-377, :399-403 (`<->`)
        assert dist([0, 1], [0, 0], 0, flt=flt) == 1.0                              # :379-383
        assert dist([1, 1, 1, 1, 1, 1, 1, 4, 5], [1] * 9, 0, flt=flt) == 5.0        # :393-397
        assert dist([3, 4], [1, 2], 1, flt=flt) == -11.0                            # :405-409, :425-429 (`<#>` = -ip)
        assert dist(list(range(1, 10)), [1] * 9, 1, flt=flt) == -45.0               # :419-423
        # fp32 overflow in the accumulator -> Infinity (:387-391) / inner product Infinity, so `<#>` = -Infinity (:413-417);
        # +Inf keys sit right below the engine's sentinel key and must still come back as a real tuple
        assert dist([-3e38], [3e38], 0, flt=flt) == math.inf
        assert dist([3e38], [3e38], 1, flt=flt) == -math.inf
    # the same through the tensor engine (TF32 value inf - inf = NaN nominates nothing: the scan must take over)
    assert dist([-3e38], [3e38], 0, engine=vr.ENGINE_TENSOR) == math.inf
    assert dist([3e38], [3e38], 1, engine=vr.ENGINE_TENSOR) == -math.inf
    assert dist([3, 4], [0, 0], 0, engine=vr.ENGINE_TENSOR) == 5.0
    # dimension mismatch: "different vector dimensions 2 and 1" (:385-386, :411-412)
    ix = _tiny_index([[1, 2]], 0)
    with pytest.raises(vr.RkError, match="different vector dimensions 2 and 1") as ei:
        ix.search_batch(np.zeros((1, 1), np.float32), u, 1)
    assert ei.value.code == -2
    sc = ix.begin_scan()
    with pytest.raises(vr.RkError, match="different vector dimensions 2 and 1"):
        sc.rescan(np.zeros(1, np.float32), 1, k=1)
    sc.end()
    ix.close()


@pytest.mark.parametrize("metric,order", [(0, [[1, 2, 3], [1, 2, 4], [1, 1, 1], [0, 0, 0]]),
                                           (1, [[1, 2, 4], [1, 2, 3], [1, 1, 1], [0, 0, 0]])])
def test_ivfflat_vector_goldens_lists1_through_the_c_abi(metric, order):
    """ivfflat_vector.out:3-14 (L2) / :39-50 (inner product): lists = 1, ORDER BY val <-> / <#> '[3,3,3]'; :16-20 NULL
    query returns every row.  Through rk_build_ivf_from_centers + rk_rescan / rk_gettuple."""
    import vectorsearch_rbac_b200 as vr
    rows = [[0, 0, 0], [1, 2, 3], [1, 1, 1], [1, 2, 4]]
    ix = _tiny_index(rows, metric)
    ix.build_ivf_from_centers(np.asarray(rows[:3], np.float32).mean(axis=0, keepdims=True))
    sc = ix.begin_scan()
    sc.rescan(np.array([3, 3, 3], np.float32), 1, k=8, probes=1, filter=vr.FILTER_NONE)
    got = [rows[int(t) - 1] for t, _, _ in sc]
    assert got == order
    sc.rescan(None, 1, k=8, probes=1, filter=vr.FILTER_NONE)
    res = list(sc)
    assert len(res) == 4 and all(r[2] == 0.0 for r in res)
    # the same with a batch limit smaller than the list: the iterator streams the rest (ivfscan.c:375-388)
    sc.rescan(np.array([3, 3, 3], np.float32), 1, k=1, probes=1, filter=vr.FILTER_BITMASK)
    assert [rows[int(t) - 1] for t, _, _ in sc] == order
    sc.end()
    ix.close()


def test_ivfflat_vector_goldens_iterative_through_the_c_abi():
    """ivfflat_vector.out:86-111: lists = 3 (one row each), iterative_scan = relaxed_order: all three rows by default,
    one with max_probes = 1, two with max_probes = 2."""
    import vectorsearch_rbac_b200 as vr
    rows = [[0, 0, 0], [1, 2, 3], [1, 1, 1]]
    ix = _tiny_index(rows, 0)
    ix.build_ivf_from_centers(np.asarray(rows, np.float32))
    sc = ix.begin_scan()
    q = np.array([3, 3, 3], np.float32)
    for maxp, exp in ((32768, [[1, 2, 3], [1, 1, 1], [0, 0, 0]]), (1, [[1, 2, 3]]), (2, [[1, 2, 3], [1, 1, 1]])):
        sc.rescan(q, 1, k=4, probes=1, max_probes=maxp, filter=vr.FILTER_NONE)
        assert [rows[int(t) - 1] for t, _, _ in sc] == exp
    sc.end()
    ix.close()


# ---- (d) dynamic partitions against the reference's own search loop ------------------------------------------------

def test_dynamic_layout_that_does_not_cover_the_visible_set():
    """controller/dynamic_partition/search.py:44-111, 347-364 restated in oracle.dynamic_partition_plan: a layout exported
    by the reference need not hold every document a role set may see (and a role set may have no CombRolePartitions row at
    all); rows outside the user's partitions are never scanned, impure partitions are filtered per row, overlapping
    partitions are merged with dedup."""
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=12000, dim=64, rows_per_doc=12, rbac="random", num_users=40, num_roles=10)
    rb = orc.Rbac(ds["doc_roles"], ds["user_roles"])
    docs = np.unique(ds["doc"])
    rng = np.random.default_rng(6)
    parts = {p: sorted(rng.choice(docs, size=len(docs) // 3, replace=False).tolist()) for p in range(1, 7)}   # overlapping
    rolesets = sorted({tuple(rb.roles_of_user(int(u))) for u in np.unique(ds["user_roles"][:, 0])})
    rp = {}
    for i, rs in enumerate(rolesets):
        if i % 7 == 3:
            continue                                     # no CombRolePartitions row for this role set
        rp[rs] = sorted(rng.choice(np.arange(1, 7), size=1 + i % 3, replace=False).tolist())
    ix = load_index(ds, role_parts=False, combo_parts=False, k_max=128)
    for p, d in parts.items():
        ix.define_partition(p, d)
    for rs, ps in rp.items():
        ix.map_roleset_to_partitions(rs, ps)
    ix.finalize()
    Q, users = _queries(ds, 80, seed=4)
    users[2] = 424242
    plist, qparts = orc.dynamic_partition_plan(ds["doc"], rb, parts, rp, users)
    exp = orc.search_batch(ds["X"], ds["tid"], plist, qparts, Q, orc.METRIC_L2, 40, dedup=True)
    got = ix.search_batch(Q, users, 40, filter=vr.FILTER_PARTITION, layout=vr.LAYOUT_DYNAMIC, engine=vr.ENGINE_SCAN)
    assert_topk_matches(got, exp, exact=True)
    # not the bitmask answer: the layout leaves visible rows out
    full = oracle_role_search(ds, Q, users, orc.METRIC_L2, 40, mode="bitmask")
    assert any(not np.array_equal(full[0][i], exp[0][i]) for i in range(len(Q)))
    assert got[3][2] == 0 and (got[3] == 0).sum() > 1            # unknown user, and role sets without partitions
    # the default engine must give the same (it keeps the dynamic layout on the scan)
    auto = ix.search_batch(Q, users, 40, filter=vr.FILTER_PARTITION, layout=vr.LAYOUT_DYNAMIC)
    for a, b in zip(auto, got):
        np.testing.assert_array_equal(a, b)
    ix.close()


# ---- (e) maximum dimensions ------------------------------------------------------------------------------------------

@pytest.mark.parametrize("dim,metric", [(2000, 0), (2048, 0), (2048, 1), (1999, 1)])
def test_maximum_dimensions(dim, metric):
    """IVFFLAT_MAX_DIM = 2000 (ivfflat.h:24) and the engine's RK_MAX_DIM = 2048: scan, tensor and (dim <= 2000) IVF."""
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=3000, dim=dim, rows_per_doc=10, rbac="tree", num_users=12, num_roles=6)
    ds["X"] = np.minimum(ds["X"], 63.0)                 # keeps every fp32 partial sum exact at 2048 dimensions
    ix = load_index(ds, metric=metric, role_parts=False, combo_parts=False)
    ix.finalize()
    Q, users = _queries(ds, 40)
    exp = oracle_role_search(ds, Q, users, metric, 20, mode="bitmask")
    for eng in (vr.ENGINE_SCAN, vr.ENGINE_TENSOR):
        got = ix.search_batch(Q, users, 20, filter=vr.FILTER_BITMASK, engine=eng)
        assert_topk_matches(got, exp, exact=True)
    if dim <= 2000:
        lists = 8
        centers = ds["X"][:lists].copy()
        ix.build_ivf_from_centers(centers)
        full = ix.search_batch(Q, users, 20, probes=lists, filter=vr.FILTER_BITMASK)
        assert_topk_matches(full, exp, exact=True)
    else:
        with pytest.raises(vr.RkError, match="2000 dimensions"):
            ix.build_ivf_from_centers(ds["X"][:4].copy())
    ix.close()
    with pytest.raises(vr.RkError):
        vr.Index(2049)


# ---- boundary: streaming iterator, threads, validation ---------------------------------------------------------------

@pytest.mark.parametrize("flt", ["none", "bitmask"])
def test_iterator_streams_whole_probe_batches(flt):
    """ivfflatgettuple (ivfscan.c:375-388) hands back EVERY tuple of the probed lists, in order, and only then refills
    from the next `probes` lists.  k = 10 per GPU pass over lists of thousands of tuples, against orc.ivf_scan."""
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=20000, dim=32, rows_per_doc=10, rbac="tree", num_users=10, num_roles=5)
    lists = 6
    centers = ds["X"][np.random.default_rng(3).choice(len(ds["X"]), lists, replace=False)].copy()
    ix = load_index(ds, role_parts=False, combo_parts=False, k_max=256)
    ix.finalize()
    ix.build_ivf_from_centers(centers)
    off, rows = orc.ivf_lists_csr(orc.ivf_assign(ds["X"], centers, 0), lists)
    Q, users = _queries(ds, 2)
    rb = orc.Rbac(ds["doc_roles"], ds["user_roles"])
    allowed = None
    if flt == "bitmask":
        allowed = np.zeros(len(ds["X"]), dtype=np.uint8)
        allowed[orc.allowed_rows(ds["doc"], rb, int(users[0]))] = 1
    f = vr.FILTER_NONE if flt == "none" else vr.FILTER_BITMASK
    exp = []
    for t, r, d in orc.ivf_scan(ds["X"], ds["tid"], centers, off, rows, Q[0], 0, probes=2, max_probes=5, allowed=allowed):
        exp += list(zip(t.tolist(), d.tolist()))
    assert len(exp) > 2000
    sc = ix.begin_scan()
    sc.rescan(Q[0], int(users[0]), k=200, probes=2, max_probes=5, filter=f)
    got = [(r[0], r[2]) for r in sc]
    assert got == exp
    # a consumer that stops early (LIMIT): the first 35 tuples with k = 10
    sc.rescan(Q[0], int(users[0]), k=10, probes=2, max_probes=5, filter=f)
    got = [sc.gettuple() for _ in range(35)]
    assert [(g[0], g[2]) for g in got] == exp[:35]
    # exact scans stand for ORDER BY ... LIMIT k: they end after k tuples
    sc.rescan(Q[1], int(users[1]), k=10, filter=f)
    assert len(list(sc)) == 10
    sc.end()
    ix.close()


def test_batch_entry_points_from_two_threads():
    """rk_search_batch* share the index's workspace: concurrent callers serialise on a mutex instead of racing."""
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=20000, dim=64, rows_per_doc=10, rbac="random", num_users=30, num_roles=9)
    ix = load_index(ds, role_parts=False, combo_parts=False, k_max=128)
    ix.finalize()
    Q, users = _queries(ds, 200, seed=13)
    exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, 25, mode="bitmask")
    out = {}

    def worker(name, lo, hi, eng):
        try:
            for _ in range(6):
                out[name] = ix.search_batch(Q[lo:hi], users[lo:hi], 25, filter=vr.FILTER_BITMASK, engine=eng)
        except Exception as e:      # noqa: BLE001
            out[name] = e

    th = [threading.Thread(target=worker, args=("a", 0, 100, vr.ENGINE_TENSOR)),
          threading.Thread(target=worker, args=("b", 100, 200, vr.ENGINE_SCAN))]
    [t.start() for t in th]
    [t.join() for t in th]
    for name, lo, hi in (("a", 0, 100), ("b", 100, 200)):
        assert not isinstance(out[name], Exception), out[name]
        assert_topk_matches(out[name], tuple(x[lo:hi] for x in exp), exact=True)
    ix.close()


def test_device_resident_queries_are_validated():
    """vector_in rejects NaN / Inf (vector.c:102-113); for device-resident queries the check is a kernel and the error
    surfaces at the next rk_synchronize."""
    import torch
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=3000, dim=32, rows_per_doc=10, rbac="tree", num_users=10, num_roles=5)
    ix = load_index(ds, role_parts=False, combo_parts=False)
    ix.finalize()
    Q, users = _queries(ds, 8)
    dev = torch.device("cuda", 0)
    k = 5
    o = vr.engine.make_opts(k, filter=vr.FILTER_BITMASK, engine=vr.ENGINE_SCAN)
    tid = torch.empty((8, k), dtype=torch.int64, device=dev); doc = torch.empty((8, k), dtype=torch.int32, device=dev)
    dist = torch.empty((8, k), dtype=torch.float64, device=dev); cnt = torch.empty(8, dtype=torch.int32, device=dev)
    dQ = torch.from_numpy(Q).to(dev)
    ix.search_batch_device(dQ.data_ptr(), 8, users, o, tid.data_ptr(), doc.data_ptr(), dist.data_ptr(), cnt.data_ptr())
    ix.synchronize()
    exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, k, mode="bitmask")
    np.testing.assert_array_equal(tid.cpu().numpy().view(np.uint64), exp[0])
    dQ[3, 7] = float("nan")
    ix.search_batch_device(dQ.data_ptr(), 8, users, o, tid.data_ptr(), doc.data_ptr(), dist.data_ptr(), cnt.data_ptr())
    with pytest.raises(vr.RkError) as ei:
        ix.synchronize()
    assert ei.value.code == -3
    dQ[3, 7] = 1.0
    ix.search_batch_device(dQ.data_ptr(), 8, users, o, tid.data_ptr(), doc.data_ptr(), dist.data_ptr(), cnt.data_ptr())
    ix.synchronize()                                  # the flag was cleared
    with pytest.raises(vr.RkError, match="unknown option"):
        ix.set_option("no_such_switch", 1)
    ix.close()


# ---- the one-launch single-query path ---------------------------------------------------------------------------------

@pytest.mark.parametrize("kind,dim,metric", [("sift", 128, 0), ("cohere", 96, 1), ("sift", 772, 0)])
def test_single_query_fused_launch(kind, dim, metric):
    """A pgvector scan is one query per rescan (ivfscan.c:339-389): nq = 1 runs as ONE launch (scan + merge by the last
    CTA), through rk_search_batch and through rk_rescan / rk_gettuple, for every filter whose plan cannot repeat a tuple."""
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset(kind, n=60000, dim=dim, rows_per_doc=20, rbac="tree", num_users=40, num_roles=12)
    ix = load_index(ds, metric=metric, k_max=1024)
    ix.finalize()
    Q, users = _queries(ds, 12, seed=17)
    users[1] = 31337                                           # unknown user: empty answer
    exact = kind == "sift"
    sc = ix.begin_scan()
    for mode, flt, layout in (("role", 0, 1), ("combination", 0, 2), ("bitmask", 1, 0), ("none", 2, 0)):
        for k in (1, 100, 1000):
            exp = oracle_role_search(ds, Q, users, metric, k, mode=mode)
            for qi in range(len(Q)):
                got = ix.search_batch(Q[qi:qi + 1], users[qi:qi + 1], k, filter=flt, layout=layout)
                st = ix.stats()
                assert st["single_fused"] == 1 and st["launches"] == 1, st
                assert_topk_matches(got, tuple(x[qi:qi + 1] for x in exp), exact=exact)
            if k == 100:
                for qi in range(0, len(Q), 3):
                    sc.rescan(Q[qi], int(users[qi]), k=k, filter=flt, layout=layout)
                    rows = list(sc)
                    n = int(exp[4][qi])
                    assert len(rows) == n
                    if exact:
                        assert [r[0] for r in rows] == exp[0][qi][:n].tolist()
                        assert [r[2] for r in rows] == exp[2][qi][:n].tolist()
    sc.end()
    # the two-launch path gives the same bits
    a = ix.search_batch(Q[:1], users[:1], 100, filter=vr.FILTER_BITMASK)
    ix.set_option("single_fused", 0)
    b = ix.search_batch(Q[:1], users[:1], 100, filter=vr.FILTER_BITMASK)
    assert ix.stats()["single_fused"] == 0 and ix.stats()["launches"] == 2
    for x, y in zip(a, b):
        np.testing.assert_array_equal(x, y)
    ix.close()


def test_single_query_fused_ties_and_fallbacks():
    """Ties at the k-th key: the answer takes the smallest tuple ids among them; more ties than the fused merge holds
    (> 2048) make the call fall back to the two-launch path; a user with several role partitions needs the dedup merge."""
    import vectorsearch_rbac_b200 as vr
    ds = make_dataset("sift", n=40000, dim=16, rows_per_doc=10, rbac="random", num_users=30, num_roles=8)
    ds["X"][5000:9000] = ds["X"][5000]                       # 4000 identical rows
    ds["X"][20000:20300] = ds["X"][20000]                    # 300 identical rows
    ix = load_index(ds, k_max=512)
    ix.finalize()
    users = np.unique(ds["user_roles"][:, 0]).astype(np.uint32)
    for row, k in ((5000, 50), (5000, 100), (20000, 100), (20000, 400), (123, 100)):
        Q = ds["X"][row:row + 1].copy()
        got = ix.search_batch(Q, users[:1], k, filter=vr.FILTER_NONE)
        exp = oracle_role_search(ds, Q, users[:1], orc.METRIC_L2, k, mode="none")
        assert_topk_matches(got, exp, exact=True)
    assert ix.stats()["single_fused"] == 1
    rb = orc.Rbac(ds["doc_roles"], ds["user_roles"])
    multi = next(int(u) for u in users if len(rb.roles_of_user(int(u))) > 1)
    one = next(int(u) for u in users if len(rb.roles_of_user(int(u))) == 1)
    Q = ds["X"][777:778].copy()
    for u, fused in ((multi, 0), (one, 1)):
        us = np.array([u], dtype=np.uint32)
        got = ix.search_batch(Q, us, 60, filter=vr.FILTER_PARTITION, layout=vr.LAYOUT_ROLE)
        assert ix.stats()["single_fused"] == fused
        assert_topk_matches(got, oracle_role_search(ds, Q, us, orc.METRIC_L2, 60, mode="role"), exact=True)
        got = ix.search_batch(Q, us, 60, filter=vr.FILTER_PARTITION, layout=vr.LAYOUT_ROLE, keep_duplicates=True)
        assert ix.stats()["single_fused"] == 1                # the ground-truth merge keeps duplicates: nothing to dedup
        assert_topk_matches(got, oracle_role_search(ds, Q, us, orc.METRIC_L2, 60, dedup=False, mode="role"), exact=True)
    ix.close()
