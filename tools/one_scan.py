#!/usr/bin/env python
"""one_scan.py — a few launches of one scan configuration (for ncu captures).
Usage: python tools/one_scan.py --rows 2000000 --nq 16 --filter none|bitmask|partition [--reps 3]"""
import argparse, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench, vectorsearch_rbac_b200 as vr

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=2_000_000)
ap.add_argument("--dim", type=int, default=128)
ap.add_argument("--kind", default="sift")
ap.add_argument("--nq", type=int, default=16)
ap.add_argument("--k", type=int, default=100)
ap.add_argument("--filter", default="none")
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--engine", default="scan")
ap.add_argument("--metric", default="l2")
ap.add_argument("--rbac", default="tree")
a = ap.parse_args()
wl = (a.rows, a.dim, a.kind, 100, 1000, 100, "one")
dev = torch.device("cuda", 0)
tid, doc, ur, dr = bench.build_rbac(wl)
if a.rbac == "random":
    import vectorsearch_rbac_b200.synth as synth
    ndoc = (a.rows + 99) // 100
    ur, dr = synth.random_rbac(1000, 100, list(range(1, ndoc + 1)), 3, 9 * ndoc // 100, seed=42)
metric = vr.METRIC_L2 if a.metric == "l2" else vr.METRIC_IP
ix = vr.Index(a.dim, metric=metric, device=0, k_max=max(128, a.k))
for s in range(0, a.rows, 1 << 20):
    idx = np.arange(s, min(a.rows, s + (1 << 20)))
    xb = bench.gen_rows(wl, idx, dev)
    ix.load_rows_device(xb.data_ptr(), len(idx), tid[idx], doc[idx])
ix.set_doc_roles(dr[:, 0], dr[:, 1]); ix.set_user_roles(ur[:, 0], ur[:, 1])
ix.build_role_partitions()
if a.filter == "combination":
    ix.build_combination_partitions()
ix.finalize(); ix.enable_timing(True)
Q, users = bench.make_queries(wl, a.nq, dev)
Q = Q.contiguous()
flt = {"none": vr.FILTER_NONE, "bitmask": vr.FILTER_BITMASK, "partition": vr.FILTER_PARTITION, "combination": vr.FILTER_PARTITION}[a.filter]
eng = {"scan": vr.ENGINE_SCAN, "tensor": vr.ENGINE_TENSOR}[a.engine]
opts = vr.engine.make_opts(a.k, filter=flt, layout=(vr.LAYOUT_COMBINATION if a.filter == "combination" else vr.LAYOUT_ROLE) if flt == vr.FILTER_PARTITION else 0, engine=eng)
o = [torch.empty((a.nq, a.k), dtype=t, device=dev) for t in (torch.int64, torch.int32, torch.float64)] + [torch.empty(a.nq, dtype=torch.int32, device=dev)]
import time
for _ in range(a.reps):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    ix.search_batch_device(Q.data_ptr(), a.nq, users, opts, *[t.data_ptr() for t in o])
    ix.synchronize()
    wall = (time.perf_counter() - t0) * 1e3
    st = ix.stats()
    flop = 2.0 * a.nq * a.rows * a.dim
    print("wall_ms %.3f qps %.0f redo %d exact %d units %d dense_TFLOPs %.1f" % (wall, a.nq / wall * 1e3, st["tensor_redo"], st["tensor_exact"], st["segments"], flop / (st["scan_ms"] * 1e-3) / 1e12 if st["scan_ms"] > 0 else 0))
    print("scan_ms %.4f select_ms %.4f GB/s %.1f segs %d" % (st["scan_ms"], st["select_ms"], st["algo_bytes"] / st["scan_ms"] / 1e6, st["segments"]))
ix.close()
