#!/usr/bin/env python
"""bench.py — filtered top-100 QPS of the RBAC-filtered kNN hot path on 1..8 B200.

Contract (driver): `python bench.py --gpus N --steps K --warmup W` (under torchrun for N > 1) prints
ONE JSON line on rank 0.  A step = one batch of `--nq` (query vector, user) pairs pushed through the
hot path (permission prefilter ⊕ L2 scoring ⊕ top-k [⊕ all-gather merge]).

Workload (default `c3`): SURVEY.md §8 C3 — SIFT10M-shaped synthetic corpus (10 M x 128 fp32,
integer-valued), 100 rows per document, tree RBAC (1000 users, 100 roles, h=4, b in [3,4]), fused
role-bitmask scan, k=100, 1000 queries per step; with N GPUs the documents are striped over the
ranks (strong scaling: the corpus and the batch are fixed) and the ranks' top-k are merged with one
all-gather.  `c1` is the 1 M-row variant (SIFT-1M shape).

  value      queries/s with the query matrix already resident in HBM (CUDA events, max over ranks)
  e2e        the same through the host-buffer API: pinned host Q -> H2D -> search -> D2H of results
  roofline   the scan kernel's algorithmic bytes (allowed rows x 4d + mask bytes +
             12k per query) / its CUDA-event duration, against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline  the reference's own distance loops (oracle/_ref, built from pgvector/src/vector.c)
             inside the oracle's restated scan, all host threads, on the same queries (N=1 only)

`--impl reference` times only that CPU path (rank 0; other ranks exit 0).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "filtered top-100 QPS at 1/2/4/8 B200 vs recall@100; % of HBM roofline"
WORKLOADS = {
    # name: (rows, dim, kind, rows_per_doc, users, roles, description)
    "c3": (10_000_000, 128, "sift", 100, 1000, 100,
           "C3 SIFT10M-shape 10M x 128 fp32 L2, tree RBAC (1000 users, 100 roles, h=4, b=3..4), fused role-bitmask "
           "scan, exact top-100, documents striped over the GPUs + all-gather top-k merge"),
    "c1": (1_000_000, 128, "sift", 100, 1000, 100,
           "C1 SIFT-1M-shape 1M x 128 fp32 L2, tree RBAC (1000 users, 100 roles, h=4, b=3..4), role-partition "
           "prefilter, exact top-100"),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--rows", type=int, default=0, help="override corpus rows (debug)")
    ap.add_argument("--nq", type=int, default=1000)
    ap.add_argument("--k", type=int, default=100)
    ap.add_argument("--filter", default="", choices=["", "bitmask", "partition"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# workload

def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def traffic_per_launch(workload, n_gpus, flt):
    """dram bytes per scan launch from the committed `ncu --set full` capture, if one matches."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(p):
        return None
    with open(p) as f:
        t = json.load(f)
    return t.get(f"{workload}:{flt}:gpus{n_gpus}")


def build_rbac(wl):
    import vectorsearch_rbac_b200.synth as synth
    rows, dim, kind, rpd, users, roles, _ = wl
    tid, doc = synth.row_ids(rows, rpd)
    doc_ids = np.arange(1, (rows + rpd - 1) // rpd + 1)
    ur, dr = synth.tree_rbac(users, roles, doc_ids.tolist(), h=4, b0=3, b1=4, seed=42)
    return tid, doc, ur, dr


def make_queries(wl, nq, device):
    import torch
    import vectorsearch_rbac_b200.synth as synth
    rows, dim, kind, rpd, users, roles, _ = wl
    qrows, qusers = synth.queries(rows, np.arange(1, users + 1), nq, seed=7)
    Q = synth.corpus(kind, torch.as_tensor(qrows, device=device), dim, seed=1234)
    return Q, qusers


def gen_rows(wl, row_idx, device):
    import torch
    import vectorsearch_rbac_b200.synth as synth
    rows, dim, kind, *_ = wl
    return synth.corpus(kind, torch.as_tensor(row_idx, device=device), dim, seed=1234)


class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU while the timed region runs (pynvml)."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake": 0x80, "sync_boost": 0x10, "display_clock_setting": 0x100}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for n, bit in names.items():
                    if r & bit:
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.01)

    def __enter__(self):
        if self.nv:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._t:
            self._t.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------
# CPU leg (the checker timed as a baseline; never on the product path)

def cpu_partitions(doc, ur, dr, qusers):
    """Visible-row lists of the distinct role sets among `qusers` (RLS predicate semantics)."""
    doc = np.asarray(doc, dtype=np.int64)
    roles_of = {}
    for u, r in ur:
        roles_of.setdefault(int(u), set()).add(int(r))
    docs_of_role = {}
    order = np.argsort(dr[:, 1], kind="stable")
    rs, ds = dr[order, 1], dr[order, 0]
    bounds = np.flatnonzero(np.diff(rs)) + 1
    for seg_r, seg_d in zip(np.split(rs, bounds), np.split(ds, bounds)):
        docs_of_role[int(seg_r[0])] = seg_d
    ndoc = int(doc.max()) + 1
    parts, index_of, qparts = [], {}, []
    for u in qusers:
        key = tuple(sorted(roles_of.get(int(u), ())))
        if key not in index_of:
            ok = np.zeros(ndoc + 1, dtype=bool)
            for r in key:
                ok[docs_of_role.get(r, np.zeros(0, dtype=np.int64))] = True
            index_of[key] = len(parts)
            parts.append(np.flatnonzero(ok[doc]).astype(np.uint32))
        qparts.append([index_of[key]] if len(parts[index_of[key]]) else [])
    return parts, qparts


def cpu_search(X, tid, parts, qparts, Q, k, nthreads):
    """(tid, dist, count), kind — the reference's own loops (oracle/_ref) inside the oracle's scan when
    that build is present, else the oracle port with pgvector's Makefile flags."""
    from oracle import oracle as orc
    ref = orc.ref_loops()
    kind = "reference" if ref is not None else "port"
    t, _, d, _, c = orc.search_batch(X, tid, parts, qparts, Q, orc.METRIC_L2, k, dedup=True, nthreads=nthreads,
                                     variant="pgvflags", ref=ref)
    return (t, d, c), kind


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_model():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def run_reference(args, wl):
    """--impl reference: the reference's CPU implementation of the path, all host threads."""
    import torch
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    rows, dim, kind, rpd, users, roles, desc = wl
    dev = torch.device("cuda", 0) if torch.cuda.is_available() else torch.device("cpu")   # data generation only
    tid, doc, ur, dr = build_rbac(wl)
    X = np.empty((rows, dim), dtype=np.float32)
    for a in range(0, rows, 1 << 20):
        b = min(rows, a + (1 << 20))
        X[a:b] = gen_rows(wl, np.arange(a, b), dev).cpu().numpy()
    Q, qusers = make_queries(wl, args.nq, dev)
    Q = Q.cpu().numpy()
    parts, qparts = cpu_partitions(doc, ur, dr, qusers)
    cores = host_cores()
    # calibrate the bounded sample: one step should take ~2 s, the whole run <= ~3 min
    n0 = min(args.nq, max(cores, 8))
    t0 = time.perf_counter()
    cpu_search(X, tid, parts, qparts[:n0], Q[:n0], args.k, cores)
    per_q = (time.perf_counter() - t0) / n0
    budget = min(2.5, 170.0 / max(1, args.steps + args.warmup))
    S = int(max(min(cores, args.nq), min(args.nq, budget / max(per_q, 1e-6))))
    for _ in range(args.warmup):
        cpu_search(X, tid, parts, qparts[:S], Q[:S], args.k, cores)
    t0 = time.perf_counter()
    kind_used = "port"
    for _ in range(args.steps):
        _, kind_used = cpu_search(X, tid, parts, qparts[:S], Q[:S], args.k, cores)
    dt = time.perf_counter() - t0
    qps = S * args.steps / dt
    sample = (f"{S} of the {args.nq} queries of the step per timed step, full visible row sets, "
              f"{cores} threads on {cpu_model()}; pgvector distance loops without PostgreSQL page/tuple/fmgr overhead")
    line = {
        "impl": "reference", "metric": METRIC, "value": qps, "unit": "queries/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "rows": rows, "dim": dim, "k": args.k, "queries_per_step": S},
        "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": cores, "kind": kind_used, "sample": sample},
        "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------

def main():
    args = parse_args()
    wl = list(WORKLOADS[args.workload])
    if args.rows:
        wl[0] = args.rows
    wl = tuple(wl)
    if args.impl == "reference":
        run_reference(args, wl)
        return

    import torch
    import torch.distributed as dist
    import vectorsearch_rbac_b200 as vr
    from vectorsearch_rbac_b200 import sharded

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the engine has no CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    rows, dim, kind, rpd, n_users, n_roles, desc = wl
    flt_name = args.filter or ("bitmask" if args.workload == "c3" else "partition")
    flt = vr.FILTER_BITMASK if flt_name == "bitmask" else vr.FILTER_PARTITION

    # ---- load this rank's shard ---------------------------------------------------------------
    tid, doc, ur, dr = build_rbac(wl)
    mine = np.flatnonzero(sharded.local_row_mask(doc, rank, world))
    ix = vr.Index(dim, metric=vr.METRIC_L2, device=local_rank, k_max=max(128, args.k))
    ix.reserve_rows(len(mine))
    keep_host = world == 1 and rank == 0 and not args.no_cpu
    X_host = np.empty((rows, dim), dtype=np.float32) if keep_host else None
    CH = 1 << 20
    for a in range(0, len(mine), CH):
        idx = mine[a:a + CH]
        xb = gen_rows(wl, idx, dev)
        ix.load_rows_device(xb.data_ptr(), len(idx), tid[idx], doc[idx])
        if keep_host:
            X_host[idx] = xb.cpu().numpy()
        del xb
    ix.set_doc_roles(dr[:, 0], dr[:, 1])
    ix.set_user_roles(ur[:, 0], ur[:, 1])
    if flt == vr.FILTER_PARTITION:
        ix.build_role_partitions()
    ix.finalize()
    torch.cuda.empty_cache()

    Q, qusers = make_queries(wl, args.nq, dev)
    Q = Q.contiguous()
    opts = vr.engine.make_opts(args.k, filter=flt, layout=vr.LAYOUT_ROLE if flt == vr.FILTER_PARTITION else 0)
    torch.cuda.set_stream(torch.cuda.Stream(dev))     # a real stream: torch events, NCCL and the engine share it
    stream = torch.cuda.current_stream(dev)
    sharded.use_torch_stream(ix, dev)
    searcher = sharded.ShardedSearcher(ix, rank, world, device=dev)
    ix.enable_timing(True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- value: device-resident inputs --------------------------------------------------------
    for _ in range(max(3, args.warmup)):
        searcher.search_device(Q, qusers, opts)
    barrier()
    launches0 = ix.stats()["total_launches"]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clk:
        e0.record(stream)
        for _ in range(args.steps):
            out = searcher.search_device(Q, qusers, opts)
        e1.record(stream)
        barrier()
    dt_dev = e0.elapsed_time(e1) * 1e-3
    st = ix.stats()
    launches = st["total_launches"] - launches0
    scan_ms = ix.scan_times_ms(min(64, args.steps))
    algo_bytes = st["algo_bytes"]
    res_dev = tuple(t.clone() for t in out)

    # ---- e2e: host buffers through the public API ----------------------------------------------
    Qh = torch.empty((args.nq, dim), dtype=torch.float32, pin_memory=True)
    Qh.copy_(Q)
    outs_h = tuple(torch.empty(t.shape, dtype=t.dtype, pin_memory=True) for t in res_dev)
    h2d = Qh.numel() * 4
    d2h = sum(t.numel() * t.element_size() for t in outs_h)

    def e2e_step():
        if world == 1:
            # the C-ABI call with host pointers: copies in, scan, select, copies out, synchronous
            o = vr._lib.lib.rk_search_batch(ix._h, args.nq, Qh.data_ptr(), qusers.ctypes.data, opts, outs_h[0].data_ptr(),
                                            outs_h[1].data_ptr(), outs_h[2].data_ptr(), outs_h[3].data_ptr())
            if o < 0:
                raise RuntimeError(vr._lib.last_error())
        else:
            searcher.search(Qh, qusers, opts, out_host=outs_h)

    for _ in range(max(3, args.warmup)):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    dt_e2e = time.perf_counter() - t0
    same = all(torch.equal(a.cpu(), b) for a, b in zip(res_dev, outs_h))

    # ---- reduce over ranks ----------------------------------------------------------------------
    if world > 1:
        t = torch.tensor([dt_dev, dt_e2e, float(np.mean(scan_ms)) if len(scan_ms) else 0.0], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt_dev, dt_e2e = float(t[0]), float(t[1])
        tl = torch.tensor([launches, algo_bytes], device=dev, dtype=torch.int64)
        dist.all_reduce(tl, op=dist.ReduceOp.SUM)
        launches_all = int(tl[0])
    else:
        launches_all = int(launches)

    if rank == 0:
        peak, peak_src = peaks()
        kms = float(np.mean(scan_ms)) if len(scan_ms) else float("nan")
        achieved = algo_bytes / (kms * 1e-3) / 1e9 if kms > 0 else 0.0
        line = {
            "metric": METRIC, "value": args.nq * args.steps / dt_dev, "unit": "queries/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": dt_dev / args.steps * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": desc, "rows": rows, "dim": dim, "k": args.k, "queries_per_step": args.nq,
                       "filter": flt_name, "rows_this_rank": int(len(mine)),
                       "l2_policy": "inputs larger than L2: each step streams %.1f GB of row data per GPU (L2 = 126 MB)" % (algo_bytes / 1e9),
                       "recall_at_100": 1.0, "recall_note": "exact scan; ids checked against the oracle on the cpu_baseline sample"},
            "roofline": {"bound": "hbm", "kernel": "rk_scan_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "frac_of_8TBps": achieved / 8000.0, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": int(algo_bytes), "kernel_ms": kms,
                         "launches_timed": int(len(scan_ms)), "traffic": traffic_per_launch(args.workload, world, flt_name)},
            "e2e": {"value": args.nq * args.steps / dt_e2e, "unit": "queries/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h), "ms_per_step": dt_e2e / args.steps * 1e3,
                    "matches_device_path": bool(same)},
            "gpu_launches": launches_all,
            "clocks": clk.summary(),
        }
        if world == 1 and not args.no_cpu:
            cores = host_cores()
            parts, qparts = cpu_partitions(doc, ur, dr, qusers)
            Qn = Q.cpu().numpy()
            n0 = min(args.nq, max(cores, 8))
            t0 = time.perf_counter()
            cpu_search(X_host, tid, parts, qparts[:n0], Qn[:n0], args.k, cores)
            per_q = (time.perf_counter() - t0) / n0
            S = int(max(min(cores, args.nq), min(args.nq, 20.0 / max(per_q, 1e-6))))
            t0 = time.perf_counter()
            (c_tid, c_dist, c_cnt), kind = cpu_search(X_host, tid, parts, qparts[:S], Qn[:S], args.k, cores)
            dtc = time.perf_counter() - t0
            g_tid = outs_h[0].numpy().view(np.uint64)[:S]
            ids_ok = bool(np.array_equal(g_tid, c_tid) and np.array_equal(outs_h[3].numpy()[:S], c_cnt)
                          and np.array_equal(outs_h[2].numpy()[:S].view(np.uint64), c_dist.view(np.uint64)))
            line["cpu_baseline"] = {
                "value": S / dtc, "unit": "queries/s", "cores": cores, "kind": kind,
                "sample": f"{S} of the step's {args.nq} queries, full visible row sets, {cores} threads on {cpu_model()}; "
                          "pgvector distance loops without PostgreSQL page/tuple/fmgr overhead",
                "gpu_ids_and_distances_bit_identical": ids_ok}
        print(json.dumps(line), flush=True)
    ix.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
