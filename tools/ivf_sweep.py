#!/usr/bin/env python
"""ivf_sweep.py — config C4: IVFFlat role-filtered search, nprobe sweep, recall@k vs the exact filtered
ground truth (the engine's own exact scan, which the parity tests pin to the oracle).
Usage: python tools/ivf_sweep.py [--rows 10000000 --dim 768 --lists 4096 --nq 1000 --k 100]"""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench, vectorsearch_rbac_b200 as vr

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=10_000_000)
ap.add_argument("--dim", type=int, default=768)
ap.add_argument("--kind", default="cohere2")
ap.add_argument("--lists", type=int, default=4096)
ap.add_argument("--nq", type=int, default=1000)
ap.add_argument("--k", type=int, default=100)
ap.add_argument("--metric", default="l2")
ap.add_argument("--probes", default="1,2,4,8,16,32,64,128,256")
ap.add_argument("--reps", type=int, default=3)
a = ap.parse_args()
wl = (a.rows, a.dim, a.kind, 100, 1000, 100, "c4")
dev = torch.device("cuda", 0)
metric = vr.METRIC_L2 if a.metric == "l2" else vr.METRIC_IP
tid, doc, ur, dr = bench.build_rbac(wl)
ix = vr.Index(a.dim, metric=metric, device=0, k_max=max(128, a.k))
ix.reserve_rows(a.rows)
t0 = time.time()
for s in range(0, a.rows, 1 << 19):
    idx = np.arange(s, min(a.rows, s + (1 << 19)))
    xb = bench.gen_rows(wl, idx, dev)
    ix.load_rows_device(xb.data_ptr(), len(idx), tid[idx], doc[idx])
    del xb
ix.set_doc_roles(dr[:, 0], dr[:, 1]); ix.set_user_roles(ur[:, 0], ur[:, 1])
ix.finalize()
t_load = time.time() - t0
t0 = time.time()
ix.build_ivf(a.lists, seed=42)
torch.cuda.synchronize()
t_build = time.time() - t0
ix.enable_timing(True)
Q, users = bench.make_queries(wl, a.nq, dev)
Q = Q.contiguous()
o = [torch.empty((a.nq, a.k), dtype=t, device=dev) for t in (torch.int64, torch.int32, torch.float64)] + [torch.empty(a.nq, dtype=torch.int32, device=dev)]

def run(probes):
    opts = vr.engine.make_opts(a.k, probes=probes, filter=vr.FILTER_BITMASK)
    ts = []
    for _ in range(a.reps + 1):
        torch.cuda.synchronize(); t = time.perf_counter()
        ix.search_batch_device(Q.data_ptr(), a.nq, users, opts, *[x.data_ptr() for x in o])
        ix.synchronize(); ts.append(time.perf_counter() - t)
    st = ix.stats()
    return min(ts[1:]), o[0].cpu().numpy().copy(), o[3].cpu().numpy().copy(), st

t_exact, gt, gt_cnt, st = run(0)
print(json.dumps(dict(rows=a.rows, dim=a.dim, lists=a.lists, nq=a.nq, k=a.k, load_s=round(t_load, 1), build_s=round(t_build, 1),
                      exact_qps=a.nq / t_exact, exact_scan_ms=st["scan_ms"], exact_gbs=st["algo_bytes"] / st["scan_ms"] / 1e6)), flush=True)
for p in [int(x) for x in a.probes.split(",")]:
    if p > a.lists:
        continue
    t, got, cnt, st = run(p)
    rec = np.mean([len(set(gt[i, :gt_cnt[i]].tolist()) & set(got[i, :cnt[i]].tolist())) / max(1, gt_cnt[i]) for i in range(a.nq)])
    print(json.dumps(dict(nprobe=p, qps=a.nq / t, recall=float(rec), scan_ms=st["scan_ms"], probe_ms=st["probe_ms"], select_ms=st["select_ms"],
                          segments=st["segments"], est_gbs=st["algo_bytes"] / max(st["scan_ms"] + st["probe_ms"], 1e-9) / 1e6)), flush=True)
ix.close()
