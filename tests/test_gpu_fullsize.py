"""BASELINE.json's FULL sizes on the GPU (C3: 10 M x 128 L2 bitmask scan; C1: 1 M x 128 role partitions; C2: 1 M x 768 inner product,
combination roles), checked through properties that do not need a CPU pass over the whole corpus:
  * the two scoring engines (CUDA-core scan, tensor-core nomination + exact re-score) return the same bits,
  * results are ordered by (distance, tuple id) and every returned document is one the user may see,
  * a query that IS a visible corpus row finds that row at distance 0 (L2),
  * the filtered k-th distance is never below the unfiltered one, rank by rank,
  * one query through the fused single-query launch equals its row of the batch,
  * and, for a few queries, the oracle over the user's visible rows only (regenerated from the seeded generator — a few hundred
    thousand rows instead of ten million) gives the same tuple ids and float8 bits.
The small-size parity tests compare every row with the oracle; bench.py's cpu_baseline leg does it at full size for the whole step."""
import numpy as np
import pytest
import torch

from oracle import oracle as orc

pytestmark = pytest.mark.gpu

K = 100
SHAPES = {
    "c3": dict(rows=10_000_000, dim=128, kind="sift", rbac="tree", metric=0, filter="bitmask"),
    "c1": dict(rows=1_000_000, dim=128, kind="sift", rbac="tree", metric=0, filter="role"),
    "c2": dict(rows=1_000_000, dim=768, kind="cohere", rbac="random", metric=1, filter="combination"),
}


def _build(wl):
    import vectorsearch_rbac_b200 as vr
    import vectorsearch_rbac_b200.synth as synth
    dev = torch.device("cuda", 0)
    rows, dim = wl["rows"], wl["dim"]
    tid, doc = synth.row_ids(rows, 100)
    doc_ids = np.arange(1, (rows + 99) // 100 + 1)
    if wl["rbac"] == "tree":
        ur, dr = synth.tree_rbac(1000, 100, doc_ids.tolist(), h=4, b0=3, b1=4, seed=42)
    else:
        ur, dr = synth.random_rbac(1000, 100, doc_ids.tolist(), 3, 9 * len(doc_ids) // 100, seed=42)
    ix = vr.Index(dim, metric=wl["metric"], device=0, k_max=128)
    ix.reserve_rows(rows)
    for s in range(0, rows, 1 << 20):
        idx = np.arange(s, min(rows, s + (1 << 20)))
        xb = synth.corpus(wl["kind"], torch.as_tensor(idx, device=dev), dim, seed=1234)
        ix.load_rows_device(xb.data_ptr(), len(idx), tid[idx], doc[idx])
        del xb
    ix.set_doc_roles(dr[:, 0], dr[:, 1])
    ix.set_user_roles(ur[:, 0], ur[:, 1])
    if wl["filter"] == "role":
        ix.build_role_partitions()
    if wl["filter"] == "combination":
        ix.build_combination_partitions()
    ix.finalize()
    return ix, tid, doc, ur, dr, dev


@pytest.mark.parametrize("shape", ["c3", "c1", "c2"])
def test_full_size_properties(shape):
    import vectorsearch_rbac_b200 as vr
    import vectorsearch_rbac_b200.synth as synth
    wl = SHAPES[shape]
    ix, tid, doc, ur, dr, dev = _build(wl)
    nq = 64
    qrows, qusers = synth.queries(wl["rows"], np.arange(1, 1001), nq, seed=7)
    rb = orc.Rbac(dr, ur)
    for qi in range(0, nq, 8):                                       # every eighth query is a row its user may see (property 4)
        vis = orc.allowed_rows(doc, rb, int(qusers[qi]))
        if len(vis):
            qrows[qi] = vis[len(vis) // 3]
    Q = synth.corpus(wl["kind"], torch.as_tensor(qrows, device=dev), wl["dim"], seed=1234).cpu().numpy()
    flt, lay = {"bitmask": (vr.FILTER_BITMASK, 0), "role": (vr.FILTER_PARTITION, vr.LAYOUT_ROLE),
                "combination": (vr.FILTER_PARTITION, vr.LAYOUT_COMBINATION)}[wl["filter"]]
    scan = ix.search_batch(Q, qusers, K, filter=flt, layout=lay, engine=vr.ENGINE_SCAN)
    tens = ix.search_batch(Q, qusers, K, filter=flt, layout=lay, engine=vr.ENGINE_TENSOR)
    assert ix.stats()["tensor_plan"] != 0                          # the tensor engine really took the batch
    # (1) engine independence, bit for bit
    np.testing.assert_array_equal(scan[3], tens[3])
    for qi in range(nq):
        n = scan[3][qi]
        np.testing.assert_array_equal(scan[0][qi, :n], tens[0][qi, :n])
        np.testing.assert_array_equal(scan[2][qi, :n].view(np.uint64), tens[2][qi, :n].view(np.uint64))
    # (2) order and (3) permission
    roles_of_doc = {}
    for d_, r_ in dr:
        roles_of_doc.setdefault(int(d_), set()).add(int(r_))
    for qi in range(nq):
        n = scan[3][qi]
        t, dd, ds = scan[0][qi, :n], scan[1][qi, :n], scan[2][qi, :n]
        assert np.all(np.diff(ds) >= 0)
        same = np.flatnonzero(np.diff(ds) == 0)
        assert np.all(t[same + 1] > t[same])                          # ties: ascending tuple id (no row is returned twice)
        mine = set(rb.roles_of_user(int(qusers[qi])))
        assert all(roles_of_doc.get(int(x), set()) & mine for x in dd)
    # (4) a query that is a visible corpus row finds it at distance 0 (L2: the smallest possible value)
    if wl["metric"] == 0:
        hits = 0
        for qi in range(nq):
            mine = set(rb.roles_of_user(int(qusers[qi])))
            if roles_of_doc.get(int(doc[qrows[qi]]), set()) & mine:
                zero = scan[0][qi, :scan[3][qi]][scan[2][qi, :scan[3][qi]] == 0.0]
                assert int(tid[qrows[qi]]) in set(int(x) for x in zero)
                hits += 1
        assert hits > 0
    # (5) filtering can only push the k-th neighbour away, rank by rank (L2 and <#> both order ascending)
    free = ix.search_batch(Q, qusers, K, filter=vr.FILTER_NONE, engine=vr.ENGINE_SCAN)
    for qi in range(nq):
        n = scan[3][qi]
        assert free[3][qi] == K
        assert np.all(free[2][qi, :n] <= scan[2][qi, :n])
    # (6) the fused single-query launch returns the batch's row
    for qi in (0, 17, 40):
        one = ix.search_batch(Q[qi:qi + 1], qusers[qi:qi + 1], K, filter=flt, layout=lay, engine=vr.ENGINE_SCAN)
        n = scan[3][qi]
        assert one[3][0] == n
        np.testing.assert_array_equal(one[0][0, :n], scan[0][qi, :n])
        np.testing.assert_array_equal(one[2][0, :n].view(np.uint64), scan[2][qi, :n].view(np.uint64))
    # (7) the oracle over the visible rows only
    exact = wl["kind"] == "sift"                                     # integer-valued rows: float8 bits must agree
    for qi in (3, 29):
        vis = orc.allowed_rows(doc, rb, int(qusers[qi])).astype(np.int64)
        if len(vis) == 0:
            assert scan[3][qi] == 0
            continue
        Xv = synth.corpus(wl["kind"], torch.as_tensor(vis, device=dev), wl["dim"], seed=1234).cpu().numpy()
        et, er, ed, es, ec = orc.search_batch(Xv, tid[vis], [np.arange(len(vis), dtype=np.uint32)], [[0]], Q[qi:qi + 1], wl["metric"], K)
        n = int(ec[0])
        assert scan[3][qi] == n
        if exact:
            np.testing.assert_array_equal(scan[0][qi, :n], et[0, :n])
            np.testing.assert_array_equal(scan[2][qi, :n].view(np.uint64), ed[0, :n].view(np.uint64))
        else:
            np.testing.assert_allclose(scan[2][qi, :n], ed[0, :n], rtol=1e-5, atol=1e-6)      # north-star tolerance: 1e-5 relative
            far = np.abs(ed[0, 1:n] - ed[0, :n - 1]) > 1e-5 * np.maximum(1.0, np.abs(ed[0, 1:n]))
            clear = np.concatenate([[True], far]) & np.concatenate([far, [True]])              # ranks no tie could swap
            np.testing.assert_array_equal(scan[0][qi, :n][clear], et[0, :n][clear])
    ix.close()


def test_full_size_ivfflat_properties():
    """C4 at full size (10 M x 768, nlist 4096, tree RBAC fused into the list scan): every returned tuple sits in a list the query
    probed (ivfscan.c:128-168) and is visible to the user; results are ordered; recall@100 against the exact filtered result
    does not fall as nprobe grows; probing every list IS the exact result; the iterator (k = 10 per refill) streams the same
    tuples a k = 120 batch returns (ivfscan.c:375-388)."""
    import vectorsearch_rbac_b200 as vr
    import vectorsearch_rbac_b200.synth as synth
    wl = dict(rows=10_000_000, dim=768, kind="cohere2", rbac="tree", metric=0, filter="bitmask")
    ix, tid, doc, ur, dr, dev = _build(wl)
    lists = 4096
    ix.build_ivf(lists, seed=42)
    _, row_list = ix.ivf_info(wl["rows"])
    nq = 64
    qrows, qusers = synth.queries(wl["rows"], np.arange(1, 1001), nq, seed=7)
    Qd = synth.corpus(wl["kind"], torch.as_tensor(qrows, device=dev), wl["dim"], seed=1234).contiguous()
    Q = Qd.cpu().numpy()
    rb = orc.Rbac(dr, ur)
    roles_of_doc = {}
    for d_, r_ in dr:
        roles_of_doc.setdefault(int(d_), set()).add(int(r_))
    exact = ix.search_batch(Q, qusers, K, probes=0, filter=vr.FILTER_BITMASK, engine=vr.ENGINE_SCAN)
    assert np.array_equal(tid, np.arange(len(tid), dtype=tid.dtype) + tid[0])       # synth.row_ids: tuple id = row number + const
    prev = -1.0
    for probes in (1, 4, 16, 64):
        d_lists = torch.empty((nq, probes), dtype=torch.int32, device=dev)
        ix.ivf_probe_device(Qd.data_ptr(), nq, vr.engine.make_opts(K, probes=probes, filter=vr.FILTER_BITMASK), d_lists.data_ptr())
        ix.synchronize()
        chosen = d_lists.cpu().numpy()
        got = ix.search_batch(Q, qusers, K, probes=probes, filter=vr.FILTER_BITMASK)
        rec = []
        for qi in range(nq):
            n = got[3][qi]
            t, dd, ds = got[0][qi, :n], got[1][qi, :n], got[2][qi, :n]
            assert np.all(np.diff(ds) >= 0)
            assert np.all(np.isin(row_list[(t - tid[0]).astype(np.int64)], chosen[qi]))
            mine = set(rb.roles_of_user(int(qusers[qi])))
            assert all(roles_of_doc.get(int(x), set()) & mine for x in dd)
            ne = exact[3][qi]
            if ne:
                rec.append(len(set(t.tolist()) & set(exact[0][qi, :ne].tolist())) / ne)
        r = float(np.mean(rec))
        assert r >= prev - 1e-9
        prev = r
    assert prev > 0.99                                                # nprobe 64 of 4096 on the two-level corpus: 0.9996 measured
    full = ix.search_batch(Q[:8], qusers[:8], K, probes=lists, filter=vr.FILTER_BITMASK)
    for a, b in zip(full, [e[:8] for e in exact]):
        np.testing.assert_array_equal(a, b)
    # the iterator streams a whole probe batch in order, ten tuples per refill
    batch = ix.search_batch(Q[:2], qusers[:2], 120, probes=16, filter=vr.FILTER_BITMASK)
    sc = vr.Scan(ix)
    for qi in range(2):
        sc.rescan(Q[qi], int(qusers[qi]), 10, probes=16, filter=vr.FILTER_BITMASK)
        n = batch[3][qi]
        pulled = [sc.gettuple() for _ in range(n)]
        np.testing.assert_array_equal(np.array([p[0] for p in pulled], dtype=np.uint64), batch[0][qi, :n])
        np.testing.assert_array_equal(np.array([p[2] for p in pulled]).view(np.uint64), batch[2][qi, :n].view(np.uint64))
    sc.end()
    ix.close()
