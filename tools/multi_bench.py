#!/usr/bin/env python
"""multi_bench.py — the C3 step through rk_multi_search_batch: ONE process, G GPUs behind one C-ABI handle (host buffers in and
out, host threads + peer exchange inside librbacknn.so).  Usage: python tools/multi_bench.py --gpus 8 [--rows N] [--steps K]"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
import vectorsearch_rbac_b200 as vr  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--gpus", type=int, default=torch.cuda.device_count())
ap.add_argument("--workload", default="c3")
ap.add_argument("--rows", type=int, default=0)
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--nq", type=int, default=1000)
ap.add_argument("--k", type=int, default=100)
ap.add_argument("--engine", default="scan", choices=["scan", "auto"])
a = ap.parse_args()
wl = dict(bench.WORKLOADS[a.workload])
if a.rows:
    wl["rows"] = a.rows
dev = torch.device("cuda", 0)
tid, doc, ur, dr = bench.build_rbac(wl)
mi = vr.MultiIndex(wl["dim"], list(range(a.gpus)), metric=vr.METRIC_L2 if wl["metric"] == "l2" else vr.METRIC_IP, k_max=max(128, a.k), nq_max=a.nq)
t0 = time.perf_counter()
for s in range(0, wl["rows"], 1 << 20):
    idx = np.arange(s, min(wl["rows"], s + (1 << 20)))
    mi.load_rows(bench.gen_rows(wl, idx, dev).cpu().numpy(), tid[idx], doc[idx])
mi.set_doc_roles(dr[:, 0], dr[:, 1])
mi.set_user_roles(ur[:, 0], ur[:, 1])
mi.finalize()
load_s = time.perf_counter() - t0
Q, users = bench.make_queries(wl, a.nq, dev)
Qn = Q.cpu().numpy()
eng = vr.ENGINE_SCAN if a.engine == "scan" else vr.ENGINE_AUTO
for _ in range(3):
    out = mi.search_batch(Qn, users, a.k, filter=vr.FILTER_BITMASK, engine=eng)
t0 = time.perf_counter()
for _ in range(a.steps):
    out = mi.search_batch(Qn, users, a.k, filter=vr.FILTER_BITMASK, engine=eng)
dt = time.perf_counter() - t0
bench.emit(({"what": "rk_multi_search_batch (one process, host buffers, synchronous)", "workload": a.workload, "rows": wl["rows"], "n_gpus": a.gpus,
                  "engine": a.engine, "queries_per_s": a.nq * a.steps / dt, "ms_per_step": dt / a.steps * 1e3, "load_s": load_s,
                  "checksum_tid": int(out[0].astype(np.uint64).sum() % (1 << 61)), "count_sum": int(out[3].sum())}))
mi.close()
