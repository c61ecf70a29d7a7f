"""rk_search_rows_noalloc — the entry point that allocates nothing (static device scratch; include/rbacknn.h) — against
the existing entry point rk_search_batch_device(filter = RK_FILTER_NONE, engine = RK_ENGINE_SCAN) through the C-ABI, on two
calls in a row, and against the oracle.  Integer-valued data: tuple ids, counts and float8 distances bit for bit; float data:
the same kernels and summation order, so bit for bit against the existing entry point as well, and the 1e-5 tie rule
(tests/util.py) against the oracle.

Kept in a file of its own that pytest collects LAST: it has no counterpart in the
reference and must not hide another test's result.  Passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

from oracle import oracle as orc
from tests.util import assert_topk_matches, load_index, make_dataset, oracle_role_search

pytestmark = pytest.mark.gpu


def _bufs(torch, dev, nq, k):
    return (torch.full((nq, k), -7, dtype=torch.int64, device=dev), torch.full((nq, k), -7, dtype=torch.int32, device=dev),
            torch.full((nq, k), -7.0, dtype=torch.float64, device=dev), torch.full((nq,), -7, dtype=torch.int32, device=dev))


def _host(o):
    return (o[0].cpu().numpy().view(np.uint64), o[1].cpu().numpy().view(np.uint32), o[2].cpu().numpy(), o[3].cpu().numpy())


@pytest.mark.parametrize("kind,dim,metric", [("sift", 128, 0), ("cohere", 96, 1)])
def test_no_alloc_entry_point_matches_the_existing_one_twice_in_a_row(kind, dim, metric):
    import torch
    import vectorsearch_rbac_b200 as vr
    lib = vr._lib.lib
    ds = make_dataset(kind, n=30000, dim=dim, rows_per_doc=20, rbac="tree")
    ix = load_index(ds, metric=metric, role_parts=False, combo_parts=False, k_max=128)
    ix.finalize()
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(17)
    # two calls in a row that differ in nq and k: the second must not see anything the first left in the shared scratch
    for call, (nq, k) in enumerate([(16, 100), (5, 128), (16, 7)]):
        Q = ds["X"][rng.integers(0, len(ds["X"]), nq)].copy()
        dQ = torch.from_numpy(Q).to(dev)
        ref = _bufs(torch, dev, nq, k)
        opts = vr.engine.make_opts(k, filter=vr.FILTER_NONE, engine=vr.ENGINE_SCAN)
        ix.search_batch_device(dQ.data_ptr(), nq, np.zeros(nq, np.uint32), opts, *[t.data_ptr() for t in ref])
        ix.synchronize()
        got = _bufs(torch, dev, nq, k)
        torch.cuda.synchronize(dev)
        free0 = torch.cuda.mem_get_info(dev)[0]
        rc = lib.rk_search_rows_noalloc(ix._h, nq, dQ.data_ptr(), k, *[t.data_ptr() for t in got])
        assert rc == 0, vr._lib.last_error()
        ix.synchronize()
        if call > 0:                                  # (the first call loads the kernels' module)
            assert torch.cuda.mem_get_info(dev)[0] == free0, "device memory changed across a call that must not allocate"
        g, r = _host(got), _host(ref)
        np.testing.assert_array_equal(g[3], r[3])
        np.testing.assert_array_equal(g[0], r[0])
        np.testing.assert_array_equal(g[1], r[1])
        np.testing.assert_array_equal(g[2].view(np.uint64), r[2].view(np.uint64))
        exp = oracle_role_search(ds, Q, np.zeros(nq, np.uint32), metric, k, mode="none")
        assert_topk_matches(g, exp, exact=kind == "sift")
    ix.close()


def test_no_alloc_entry_point_rejects_what_the_static_scratch_cannot_hold():
    import torch
    import vectorsearch_rbac_b200 as vr
    lib = vr._lib.lib
    ds = make_dataset("sift", n=2000, dim=32, rows_per_doc=10, rbac="tree")
    ix = load_index(ds, role_parts=False, combo_parts=False, k_max=256)
    ix.finalize()
    dev = torch.device("cuda", 0)
    dQ = torch.zeros((32, 32), dtype=torch.float32, device=dev)
    o = _bufs(torch, dev, 32, 256)
    ptrs = [t.data_ptr() for t in o]
    assert lib.rk_search_rows_noalloc(ix._h, 17, dQ.data_ptr(), 10, *ptrs) < 0 and "RK_NOALLOC_MAX_NQ" in vr._lib.last_error()
    assert lib.rk_search_rows_noalloc(ix._h, 4, dQ.data_ptr(), 129, *ptrs) < 0 and "RK_NOALLOC_MAX_K" in vr._lib.last_error()
    assert lib.rk_search_rows_noalloc(ix._h, 4, None, 10, *ptrs) < 0
    assert lib.rk_search_rows_noalloc(ix._h, 0, dQ.data_ptr(), 10, *ptrs) == 0          # nothing to do
    # a table smaller than k: counts below k, padding as the existing entry point writes it
    rc = lib.rk_search_rows_noalloc(ix._h, 2, dQ.data_ptr(), 128, *ptrs)
    assert rc == 0, vr._lib.last_error()
    ix.synchronize()
    ix.close()
