#!/usr/bin/env python
"""check_engines.py — tensor engine vs scan engine (and the oracle on a few queries) on one bench workload.
Usage: python tools/check_engines.py --workload c2 [--nq 200] [--rows N]"""
import argparse, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench, vectorsearch_rbac_b200 as vr

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="c2")
ap.add_argument("--nq", type=int, default=200)
ap.add_argument("--rows", type=int, default=0)
ap.add_argument("--k", type=int, default=100)
ap.add_argument("--cpu", type=int, default=8, help="queries also checked against the oracle")
a = ap.parse_args()
wl = dict(bench.WORKLOADS[a.workload])
if a.rows:
    wl["rows"] = a.rows
dev = torch.device("cuda", 0)
tid, doc, ur, dr = bench.build_rbac(wl)
metric = vr.METRIC_L2 if wl["metric"] == "l2" else vr.METRIC_IP
ix = vr.Index(wl["dim"], metric=metric, device=0, k_max=max(128, a.k))
X = np.empty((wl["rows"], wl["dim"]), dtype=np.float32)
for s in range(0, wl["rows"], 1 << 18):
    idx = np.arange(s, min(wl["rows"], s + (1 << 18)))
    xb = bench.gen_rows(wl, idx, dev)
    ix.load_rows_device(xb.data_ptr(), len(idx), tid[idx], doc[idx])
    X[idx] = xb.cpu().numpy()
ix.set_doc_roles(dr[:, 0], dr[:, 1]); ix.set_user_roles(ur[:, 0], ur[:, 1])
ix.finalize()
Q, users = bench.make_queries(wl, a.nq, dev)
Qn = Q.cpu().numpy()
res = {}
for name, eng in (("scan", vr.ENGINE_SCAN), ("tensor", vr.ENGINE_TENSOR)):
    res[name] = ix.search_batch(Qn, users, a.k, filter=vr.FILTER_BITMASK, engine=eng)
    st = ix.stats()
    print(name, "plan", st["tensor_plan"], "redo", st["tensor_redo"], "units", st["segments"])
s_, t_ = res["scan"], res["tensor"]
bad = [q for q in range(a.nq) if not (np.array_equal(s_[0][q], t_[0][q]) and np.array_equal(s_[2][q].view(np.uint64), t_[2][q].view(np.uint64)))]
print("queries where tensor != scan:", len(bad), bad[:10])
for q in bad[:3]:
    d = np.flatnonzero(s_[0][q] != t_[0][q])
    print(" q", q, "first diff pos", d[:5], "scan", s_[2][q][d[:3]], "tensor", t_[2][q][d[:3]], "counts", s_[3][q], t_[3][q])
for q in bad[:5]:
    miss = sorted(set(s_[0][q].tolist()) - set(t_[0][q].tolist()))
    print(" q", q, "user", users[q], "rows the tensor engine missed (tid, tile, row in tile):", [(m, (m - 1) // 128, (m - 1) % 128) for m in miss])
if a.cpu:
    from tests.util import assert_topk_matches
    parts, qparts = bench.cpu_partitions(doc, ur, dr, users[:a.cpu])
    (c_tid, c_dist, c_cnt), kind = bench.cpu_search(X, tid, parts, qparts, Qn[:a.cpu], a.k, 8, wl["metric"])
    for name in ("scan", "tensor"):
        g = res[name]
        try:
            assert_topk_matches((g[0][:a.cpu], None, g[2][:a.cpu], g[3][:a.cpu]), (c_tid, None, c_dist, None, c_cnt), exact=wl["kind"] == "sift")
            print(name, "vs oracle (%s): ok" % kind)
        except AssertionError as e:
            print(name, "vs oracle (%s): MISMATCH" % kind, str(e)[:300])
            print("  gpu", g[2][0][:5], g[0][0][:5], "cpu", c_dist[0][:5], c_tid[0][:5])
ix.close()
