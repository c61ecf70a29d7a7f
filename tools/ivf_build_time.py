#!/usr/bin/env python
"""ivf_build_time.py — C4 IVFFlat build (k-means + final assignment) with the tensor-core assignment on / off: seconds, rows the
proof refused, and whether both give the same lists for the same centres.  Usage: python tools/ivf_build_time.py [--rows N]"""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench, vectorsearch_rbac_b200 as vr

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=10_000_000)
ap.add_argument("--lists", type=int, default=4096)
a = ap.parse_args()
wl = dict(bench.WORKLOADS["c4"]); wl["rows"] = a.rows
dev = torch.device("cuda", 0)
tid, doc, ur, dr = bench.build_rbac(wl)


def make_index(tc):
    ix = vr.Index(wl["dim"], metric=vr.METRIC_L2, device=0, k_max=128)
    ix.reserve_rows(a.rows)
    for s in range(0, a.rows, 1 << 18):
        idx = np.arange(s, min(a.rows, s + (1 << 18)))
        xb = bench.gen_rows(wl, idx, dev); torch.cuda.synchronize()
        ix.load_rows_device(xb.data_ptr(), len(idx), tid[idx], doc[idx]); del xb
    ix.set_doc_roles(dr[:, 0], dr[:, 1]); ix.set_user_roles(ur[:, 0], ur[:, 1]); ix.finalize()
    ix.set_option("ivf_assign_tc", tc)
    return ix


out = {}
for tc in (1, 0):
    ix = make_index(tc)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    ix.build_ivf(a.lists, seed=42)
    torch.cuda.synchronize(); t_build = time.perf_counter() - t0
    s1 = ix.stats()
    out[tc] = dict(build_s=t_build, tc_rows=s1["ivf_assign_tc_rows"], refused_rows=s1["ivf_assign_fallback_rows"])
    if tc:
        c_tc, lists_tc = ix.ivf_info(a.rows)
    ix.close(); del ix; torch.cuda.empty_cache()
# the final assignment alone, both ways, from the SAME centres: seconds and equality of the lists
res = {}
for tc in (1, 0):
    ix = make_index(tc)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    ix.build_ivf_from_centers(c_tc)
    torch.cuda.synchronize(); out[tc]["assign_and_pack_s"] = time.perf_counter() - t0
    out[tc]["assign_refused_rows"] = ix.stats()["ivf_assign_fallback_rows"]
    res[tc] = ix.ivf_info(a.rows)[1]
    ix.close(); del ix; torch.cuda.empty_cache()
bench.emit({"rows": a.rows, "lists": a.lists, "tensor_core": out[1], "fp32_kernel": out[0],
            "same_lists_for_the_same_centres": bool(np.array_equal(res[1], res[0]) and np.array_equal(res[1], lists_tc))})
