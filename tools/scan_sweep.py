#!/usr/bin/env python
"""scan_sweep.py — single-query (and small-batch) filtered exact scans at several selectivities:
scan-kernel CUDA-event time vs algorithmic bytes -> GB/s and fraction of the measured
HBM peak.  Usage: python tools/scan_sweep.py [--rows N] [--dim D] [--nq 1,8,64]"""
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--dim", type=int, default=128)
    ap.add_argument("--kind", default="sift")
    ap.add_argument("--nq", default="1,16")
    ap.add_argument("--k", type=int, default=100)
    ap.add_argument("--reps", type=int, default=10)
    a = ap.parse_args()
    wl = (a.rows, a.dim, a.kind, 100, 1000, 100, "sweep")
    dev = torch.device("cuda", 0)
    tid, doc, ur, dr = bench.build_rbac(wl)
    ix = vr.Index(a.dim, metric=vr.METRIC_L2, device=0, k_max=max(128, a.k))
    ix.reserve_rows(a.rows)
    for s in range(0, a.rows, 1 << 20):
        idx = np.arange(s, min(a.rows, s + (1 << 20)))
        xb = bench.gen_rows(wl, idx, dev)
        ix.load_rows_device(xb.data_ptr(), len(idx), tid[idx], doc[idx])
    ix.set_doc_roles(dr[:, 0], dr[:, 1])
    ix.set_user_roles(ur[:, 0], ur[:, 1])
    ix.build_role_partitions()
    ix.finalize()
    ix.enable_timing(True)
    peak, _ = bench.peaks()
    # one user per tree depth: visible rows differ by role depth
    nroles_docs = {}
    for d, r in dr:
        nroles_docs[r] = nroles_docs.get(r, 0) + 1
    by_size = sorted(set(nroles_docs.values()))
    users = []
    for sz in by_size:
        r = next(r for r, c in nroles_docs.items() if c == sz)
        users.append(int(ur[ur[:, 1] == r][0, 0]))
    out = []
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for nq in [int(x) for x in a.nq.split(",")]:
        Qall, _ = bench.make_queries(wl, nq, dev)
        Qall = Qall.contiguous()
        o_tid = torch.empty((nq, a.k), dtype=torch.int64, device=dev)
        o_doc = torch.empty((nq, a.k), dtype=torch.int32, device=dev)
        o_dist = torch.empty((nq, a.k), dtype=torch.float64, device=dev)
        o_cnt = torch.empty(nq, dtype=torch.int32, device=dev)
        cases = [("none", vr.FILTER_NONE, 0, users[0])]
        for u in users:
            cases.append(("bitmask", vr.FILTER_BITMASK, 0, u))
            cases.append(("partition", vr.FILTER_PARTITION, vr.LAYOUT_ROLE, u))
        for name, flt, layout, u in cases:
            opts = vr.engine.make_opts(a.k, filter=flt, layout=layout)
            us = np.full(nq, u, dtype=np.uint32)
            ts = []
            for rep in range(a.reps + 2):
                flush.zero_()       # evict the row store from L2 between repetitions
                ix.search_batch_device(Qall.data_ptr(), nq, us, opts, o_tid.data_ptr(), o_doc.data_ptr(), o_dist.data_ptr(), o_cnt.data_ptr())
                ix.synchronize()
                st = ix.stats()
                if rep >= 2:
                    ts.append((st["scan_ms"], st["select_ms"]))
            ms = float(np.median([t[0] for t in ts]))
            sel = float(np.median([t[1] for t in ts]))
            gbs = st["algo_bytes"] / (ms * 1e-3) / 1e9
            rec = dict(nq=nq, filter=name, user=u, rows_scanned=st["rows_scanned"], segments=st["segments"], algo_mb=st["algo_bytes"] / 1e6,
                       scan_ms=ms, select_ms=sel, gbs=gbs, frac=gbs / peak)
            out.append(rec)
            print(json.dumps(rec), flush=True)
    ix.close()


if __name__ == "__main__":
    main()
