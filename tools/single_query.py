#!/usr/bin/env python
"""single_query.py — the single-query legs of bench.py on their own (C3 or C1 shape): device / e2e / rescan microseconds of
one query per role depth, L2 flushed.  Usage: python tools/single_query.py [--workload c3] [--rows N] [--filter bitmask|partition]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
import vectorsearch_rbac_b200 as vr  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="c3")
ap.add_argument("--rows", type=int, default=0)
ap.add_argument("--filter", default="bitmask")
ap.add_argument("--k", type=int, default=100)
ap.add_argument("--reps", type=int, default=20)
ap.add_argument("--fused", type=int, default=1)
ap.add_argument("--zc", type=int, default=1)
ap.add_argument("--debug", type=int, default=0)
ap.add_argument("--segs", type=int, default=0)
ap.add_argument("--seed", type=int, default=0)
ap.add_argument("--seedmin", type=int, default=2048)
a = ap.parse_args()
wl = dict(bench.WORKLOADS[a.workload])
if a.rows:
    wl["rows"] = a.rows
dev = torch.device("cuda", 0)
tid, doc, ur, dr = bench.build_rbac(wl)
ix = vr.Index(wl["dim"], metric=vr.METRIC_L2 if wl["metric"] == "l2" else vr.METRIC_IP, device=0, k_max=max(128, a.k))
ix.reserve_rows(wl["rows"])
for s in range(0, wl["rows"], 1 << 20):
    idx = np.arange(s, min(wl["rows"], s + (1 << 20)))
    xb = bench.gen_rows(wl, idx, dev)
    ix.load_rows_device(xb.data_ptr(), len(idx), tid[idx], doc[idx])
ix.set_doc_roles(dr[:, 0], dr[:, 1])
ix.set_user_roles(ur[:, 0], ur[:, 1])
if a.filter == "partition":
    ix.build_role_partitions()
ix.finalize()
ix.set_option("single_fused", a.fused)
ix.set_option("single_zero_copy", a.zc)
ix.set_option("single_segs", a.segs)
ix.set_option("single_seed", a.seed)
ix.set_option("single_seed_min", a.seedmin)
if a.debug:
    # phase timeline of a few fused launches per case (stderr), L2 flushed
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    Q, _ = bench.make_queries(wl, 4, dev)
    Qn = Q.cpu().numpy()
    for sz, us_ in [(0, [0])] + bench.depth_users(ur, dr):
        f = vr.FILTER_NONE if sz == 0 else (vr.FILTER_BITMASK if a.filter == "bitmask" else vr.FILTER_PARTITION)
        lay = vr.LAYOUT_ROLE if f == vr.FILTER_PARTITION else 0
        for mode in ("rotating", "flushed"):
            for i in range(10):
                if mode == "flushed":
                    flush.zero_(); torch.cuda.synchronize()
                ix.set_option("single_debug", a.debug if i >= 8 else 0)
                if i >= 8:
                    print(f"--- docs visible {sz} {mode} rep {i}", file=sys.stderr)
                ix.search_batch(Qn[i % 4:i % 4 + 1], np.array([us_[i % len(us_)]], np.uint32), a.k, filter=f, layout=lay, engine=vr.ENGINE_SCAN)
    ix.set_option("single_debug", 0)
    del flush
flt = vr.FILTER_BITMASK if a.filter == "bitmask" else vr.FILTER_PARTITION
lay = 0 if a.filter == "bitmask" else vr.LAYOUT_ROLE
res = bench.single_query_legs(ix, wl, ur, dr, dev, a.k, flt, lay, reps=a.reps)
for leg in res["legs"]:
    print(json.dumps(leg))
    print("%-52s roof %6.1f | rotating dev %6.1f e2e %6.1f rescan %6.1f | flushed dev %6.1f e2e %6.1f rescan %6.1f" % (
        leg["case"][:52], leg["roofline_us"], leg["rotating"]["device_us"], leg["rotating"]["e2e_us"], leg["rotating"]["rescan_gettuple_us"],
        leg["flushed"]["device_us"], leg["flushed"]["e2e_us"], leg["flushed"]["rescan_gettuple_us"]), file=sys.stderr)
ix.close()
