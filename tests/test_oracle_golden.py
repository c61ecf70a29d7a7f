"""Pins oracle/ against every known-answer vector the reference holds for the hot path.

Sources (relative to the reference checkout):
  pgvector/test/expected/vector_type.out:373-429      distance functions / operators
  pgvector/test/expected/ivfflat_vector.out:3-20,39-50,86-111   IVFFlat order, NULL query, probes
  pgvector/test/t/003_ivfflat_vector_build_recall.pl:104-116     recall == 1.0 at probes == lists
  pgvector/test/t/005_ivfflat_query_recall.pl:32-41              self query is its own top-1
and, when oracle/_ref was built from the reference's vector.c, the reference's own loops.
"""
import math

import numpy as np
import pytest

from oracle import oracle as orc


def test_l2_distance_goldens():
    assert orc.l2_distance([0, 0], [3, 4]) == 5.0                       # vector_type.out:373-377
    assert orc.l2_distance([0, 0], [0, 1]) == 1.0                       # :379-383
    with pytest.raises(ValueError, match="different vector dimensions 2 and 1"):
        orc.l2_distance([1, 2], [3])                                    # :385-386
    with np.errstate(over="ignore"):
        assert math.isinf(orc.l2_distance([3e38], [-3e38]))             # :387-391 fp32 overflow
    assert orc.l2_distance([1] * 9, [1, 1, 1, 1, 1, 1, 1, 4, 5]) == 5.0  # :393-397


def test_inner_product_goldens():
    assert orc.inner_product([1, 2], [3, 4]) == 11.0                    # :405-409
    with pytest.raises(ValueError, match="different vector dimensions 2 and 1"):
        orc.inner_product([1, 2], [3])                                  # :411-412
    assert math.isinf(orc.inner_product([3e38], [3e38]))                # :413-417
    assert orc.inner_product([1] * 9, list(range(1, 10))) == 45.0       # :419-423
    assert orc.negative_inner_product([1, 2], [3, 4]) == -11.0          # :425-429


@pytest.mark.parametrize("o0", [False, True])
def test_port_matches_reference_loops(o0):
    ref = orc.ref_loops(o0=o0)
    if ref is None:
        pytest.skip("oracle/_ref not built (reference checkout absent)")
    import ctypes as C
    rng = np.random.default_rng(5)
    fp = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))
    # goldens through the reference's own code
    a = np.array([1] * 9, dtype=np.float32)
    b = np.array([1, 1, 1, 1, 1, 1, 1, 4, 5], dtype=np.float32)
    assert ref.lib.pgvref_l2_distance(9, fp(a), fp(b)) == 5.0
    c = np.arange(1, 10, dtype=np.float32)
    assert ref.lib.pgvref_neg_ip(9, fp(a), fp(c)) == -45.0
    for dim in (1, 3, 16, 127, 128, 300, 768):
        # integer-valued (SIFT-shaped): every summation order is exact -> bit equality for both builds
        x = rng.integers(0, 256, dim).astype(np.float32)
        y = rng.integers(0, 256, dim).astype(np.float32)
        assert ref.lib.pgvref_l2sq(dim, fp(x), fp(y)) == orc.lib().lib.orc_l2sq(dim, fp(x), fp(y))
        assert ref.lib.pgvref_ip(dim, fp(x), fp(y)) == orc.lib().lib.orc_ip(dim, fp(x), fp(y))
        # real-valued: the -O0 build is the strict order bit for bit; the vectorised build agrees to 1e-5
        x = rng.standard_normal(dim).astype(np.float32)
        y = rng.standard_normal(dim).astype(np.float32)
        s_ref = ref.lib.pgvref_l2sq(dim, fp(x), fp(y))
        s_port = orc.lib().lib.orc_l2sq(dim, fp(x), fp(y))
        if o0:
            assert s_ref == s_port
            assert ref.lib.pgvref_ip(dim, fp(x), fp(y)) == orc.lib().lib.orc_ip(dim, fp(x), fp(y))
        else:
            assert abs(s_ref - s_port) <= 1e-5 * abs(s_port)


def _ivf_table():
    # ivfflat_vector.out:3-9: three rows, index built with lists=1, then [1,2,4] inserted
    X = np.array([[0, 0, 0], [1, 2, 3], [1, 1, 1], [1, 2, 4]], dtype=np.float32)
    tid = np.arange(4, dtype=np.uint64)
    return X, tid


def test_ivfflat_order_l2_and_ip_lists1():
    X, tid = _ivf_table()
    centers = X[:3].mean(axis=0, keepdims=True).astype(np.float32)      # one list holds everything
    off, rows = orc.ivf_lists_csr(orc.ivf_assign(X, centers, orc.METRIC_L2), 1)
    q = np.array([3, 3, 3], dtype=np.float32)
    (t, r, d), = orc.ivf_scan(X, tid, centers, off, rows, q, orc.METRIC_L2, probes=1)
    assert [X[i].tolist() for i in r] == [[1, 2, 3], [1, 2, 4], [1, 1, 1], [0, 0, 0]]   # :9-14
    (t, r, d), = orc.ivf_scan(X, tid, centers, off, rows, q, orc.METRIC_IP, probes=1)
    assert [X[i].tolist() for i in r] == [[1, 2, 4], [1, 2, 3], [1, 1, 1], [0, 0, 0]]   # :45-50
    # NULL query returns every indexed row (:16-20, :52-56)
    (t, r, d), = orc.ivf_scan(X, tid, centers, off, rows, None, orc.METRIC_L2, probes=1)
    assert len(t) == 4 and np.all(d == 0.0)


def test_ivfflat_iterative_probes():
    # ivfflat_vector.out:86-111: lists=3, one row per list, relaxed_order iterative scan
    X = np.array([[0, 0, 0], [1, 2, 3], [1, 1, 1]], dtype=np.float32)
    tid = np.arange(3, dtype=np.uint64)
    centers = X.copy()
    off, rows = orc.ivf_lists_csr(orc.ivf_assign(X, centers, orc.METRIC_L2), 3)
    q = np.array([3, 3, 3], dtype=np.float32)
    got = lambda maxp: [X[i].tolist() for b in orc.ivf_scan(X, tid, centers, off, rows, q, orc.METRIC_L2, 1, maxp)
                        for i in b[1]]
    assert got(32768) == [[1, 2, 3], [1, 1, 1], [0, 0, 0]]      # default max_probes: all lists (:89-96)
    assert got(1) == [[1, 2, 3]]                                # max_probes = 1 (:97-102)
    assert got(2) == [[1, 2, 3], [1, 1, 1]]                     # max_probes = 2 (:103-109)


def test_ivfflat_recall_full_probes_is_exact():
    # 003_ivfflat_vector_build_recall.pl: 3-d uniform data, k (limit) 20, 100 lists; probes == lists -> recall 1.0
    rng = np.random.default_rng(42)
    n, lists, k = 5000, 100, 20
    X = rng.random((n, 3), dtype=np.float32)
    tid = np.arange(n, dtype=np.uint64)
    centers = X[rng.choice(n, lists, replace=False)].copy()
    for metric in (orc.METRIC_L2, orc.METRIC_IP):
        off, rows = orc.ivf_lists_csr(orc.ivf_assign(X, centers, metric), lists)
        Q = rng.random((20, 3), dtype=np.float32)
        et, _, ed, _, _ = orc.search_batch(X, tid, [np.arange(n, dtype=np.uint32)], [[0]] * len(Q), Q, metric, k)
        for i, q in enumerate(Q):
            (t, r, d), = orc.ivf_scan(X, tid, centers, off, rows, q, metric, probes=lists)
            assert orc.recall(et[i], t[:k]) == 1.0
            np.testing.assert_array_equal(d[:k], ed[i])
            lo = orc.ivf_scan(X, tid, centers, off, rows, q, metric, probes=1)[0][0]
            assert orc.recall(et[i], lo[:k]) <= 1.0
    # 005_ivfflat_query_recall.pl:32-41 — a row queried with its own vector comes back first (L2)
    off, rows = orc.ivf_lists_csr(orc.ivf_assign(X, centers, orc.METRIC_L2), lists)
    for i in (0, 17, 4999):
        t = orc.ivf_scan(X, tid, centers, off, rows, X[i], orc.METRIC_L2, probes=1)[0][0]
        assert t[0] == i


def test_merge_dedup_semantics():
    # prefilter_role.py:174-189: two role partitions sharing a document -> duplicates removed, k kept
    rng = np.random.default_rng(0)
    X = rng.integers(0, 256, (60, 8)).astype(np.float32)
    tid = (np.arange(60) + 1).astype(np.uint64)
    pa = np.arange(0, 40, dtype=np.uint32)
    pb = np.arange(20, 60, dtype=np.uint32)
    q = X[25:26]
    t, r, d, s, c = orc.search_batch(X, tid, [pa, pb], [[0, 1]], q, orc.METRIC_L2, 10, dedup=True)
    assert c[0] == 10 and len(set(t[0].tolist())) == 10 and t[0][0] == 26 and d[0][0] == 0.0
    exact = np.sqrt(((X - q) ** 2).sum(1).astype(np.float32).astype(np.float64))
    order = np.lexsort((tid, exact))[:10]
    np.testing.assert_array_equal(t[0], tid[order])
    # ground-truth variant keeps duplicates (common_function.py:713-714)
    t2, *_ = orc.search_batch(X, tid, [pa, pb], [[0, 1]], q, orc.METRIC_L2, 10, dedup=False)
    assert t2[0][0] == 26 and t2[0][1] == 26
    # unknown user: no partitions -> empty result (prefilter_role.py:120-124)
    t3, _, d3, _, c3 = orc.search_batch(X, tid, [pa, pb], [[]], q, orc.METRIC_L2, 10)
    assert c3[0] == 0 and np.all(np.isinf(d3))
