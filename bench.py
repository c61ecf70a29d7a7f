#!/usr/bin/env python
"""bench.py — filtered top-100 QPS of the RBAC-filtered kNN hot path on 1..8 B200.

Contract (driver): `python bench.py --gpus N --steps K --warmup W` (under torchrun for N > 1) prints
ONE JSON line on rank 0.  A step = one batch of `--nq` (query vector, user) pairs pushed through the
hot path (permission prefilter ⊕ L2 scoring ⊕ top-k [⊕ all-gather merge]).

Workloads (default `c3`):
  c3  SIFT10M-shaped synthetic corpus (10 M x 128 fp32, integer-valued), 100 rows per document, tree RBAC
      (1000 users, 100 roles, h=4, b in [3,4]), fused role-bitmask scan — one bandwidth-bound pass per
      query, as the config specifies — k=100, 1000 queries per step; with N GPUs the documents are striped
      over the ranks (strong scaling: corpus and batch are fixed) and the ranks' top-k are merged with one
      all-gather.  The line also carries a "batched" object: the same step through RK_ENGINE_AUTO, which
      sends a batch this size to the tcgen05 engine (identical results).
  c1  the 1 M-row variant (SIFT-1M shape), role-partition prefilter.
  c2  Cohere-shaped 1 M x 768 fp32 inner product, random RBAC, combination-role filter, batched
      tensor-core engine (the headline engine of this workload).

  value      queries/s with the query matrix already resident in HBM (CUDA events, max over ranks)
  e2e        the same through the host-buffer API: pinned host Q -> H2D -> search -> D2H of results
  roofline   the dominant kernel against the roofline that bounds it: scan kernel — algorithmic bytes
             (allowed rows x 4d + mask bytes + 12k per query) / CUDA-event duration against
             MEASURED_PEAKS.json hbm_gbs; tensor kernel — F = 2 nq_g R_g d against a cuBLAS TF32 peak
             measured live, or the plan's bytes against hbm_gbs when its intensity is below the ridge
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
    # SURVEY.md §8 configs.  engine = what the headline timed region runs; c1 / c3 also report the batched
    # (RK_ENGINE_AUTO -> tensor) numbers of the same step in a "batched" object.
    "c3": dict(rows=10_000_000, dim=128, kind="sift", rpd=100, users=1000, roles=100, rbac="tree", metric="l2",
               filter="bitmask", engine="scan",
               desc="C3 SIFT10M-shape 10M x 128 fp32 L2, tree RBAC (1000 users, 100 roles, h=4, b=3..4), fused role-bitmask "
                    "scan (one pass per query), exact top-100, documents striped over the GPUs + all-gather top-k merge"),
    "c1": dict(rows=1_000_000, dim=128, kind="sift", rpd=100, users=1000, roles=100, rbac="tree", metric="l2",
               filter="partition", engine="scan",
               desc="C1 SIFT-1M-shape 1M x 128 fp32 L2, tree RBAC (1000 users, 100 roles, h=4, b=3..4), role-partition "
                    "prefilter, exact top-100"),
    "c2": dict(rows=1_000_000, dim=768, kind="cohere", rpd=100, users=1000, roles=100, rbac="random", metric="ip",
               filter="combination", engine="tensor",
               desc="C2 Cohere-Wikipedia shape 1M x 768 fp32 inner product, random RBAC (1000 users, 100 roles, <=3 roles per "
                    "user, 450..900 documents per role), combination-role (user-level) filter, batched tensor-core exact top-100"),
    "c5": dict(rows=1_000_000, dim=128, kind="sift", rpd=100, users=1000, roles=100, rbac="random", metric="l2",
               filter="dynamic", engine="scan",
               desc="C5 dynamic-partition workload replayed as GPU filtered scans: SIFT-1M-shape 1M x 128 fp32 L2, random RBAC "
                    "(1000 users, 100 roles, <=3 roles per user), partitions from the AnonySys greedy role splitter restated with "
                    "a linear scan-cost model at storage alpha = 2.0 (the optimiser's fitted HNSW constants are not shipped), "
                    "per-row permission predicate inside impure partitions, exact top-100 (recall 1.0 >= the 0.95 target)"),
    "c4": dict(rows=10_000_000, dim=768, kind="cohere2", rpd=100, users=1000, roles=100, rbac="tree", metric="l2",
               filter="bitmask", engine="scan", ivf_lists=4096,
               desc="C4 IVFFlat role-filtered search: synthetic Cohere shape 10M x 768 fp32 L2 (two-level mixture: 64 topics x 1024 "
                    "sub-topics, so that recall depends on nprobe), tree RBAC mask fused into the list scan, nlist = 4096 (GPU k-means "
                    "build), one nprobe point per run (--nprobe; sweep: tools/ivf_sweep.py), top-100, recall@100 against the exact "
                    "filtered result (the engine's, and the oracle's on a sample)"),
}


# The contract is ONE JSON line on stdout.  Libraries print there too (NCCL's "NCCL version ..." banner at the first
# collective), so everything else written to fd 1 is sent to stderr and only emit() reaches the real stdout.
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(line):
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


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
    ap.add_argument("--filter", default="", choices=["", "bitmask", "partition", "combination", "dynamic"])
    ap.add_argument("--engine", default="", choices=["", "scan", "tensor", "auto"], help="override the workload's headline engine")
    ap.add_argument("--nprobe", type=int, default=16, help="workload c4: lists probed per query")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--pipeline", action="store_true", help="(kept for compatibility: pipelining is the default with the peer exchange)")
    ap.add_argument("--no-pipeline", action="store_true", help="N > 1: run scan, exchange and merge of a step back to back on one stream")
    ap.add_argument("--peer-overlap", default="deferred", choices=["deferred", "side"],
                    help="pipelined peer exchange: deferred = merge of step i-1 after the push of step i on one stream; side = merge on a side stream")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="N > 1: peer = stores into the other GPUs' receive rings over NVLink + flags (rk_exchange.cu); nccl = all_gather_into_tensor")
    ap.add_argument("--no-batched", action="store_true", help="skip the batched (RK_ENGINE_AUTO) leg of a scan-engine workload")
    ap.add_argument("--emu-world", type=int, default=0, help="debug, one GPU: load only the stripe rank --emu-rank of this many ranks would hold")
    ap.add_argument("--emu-rank", type=int, default=0)
    ap.add_argument("--no-replicas", action="store_true", help="N > 1: skip the query-split replica leg")
    ap.add_argument("--no-single", action="store_true", help="skip the single-query legs (N = 1, tree RBAC workloads)")
    ap.add_argument("--single-reps", type=int, default=16)
    ap.add_argument("--no-legs", action="store_true", help="default c3 line at N = 1: skip the short legs of c1 / c2 / c5 / c4")
    ap.add_argument("--legs-budget-s", type=int, default=200, help="no further short leg is started after this many seconds")
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


def as_workload(wl):
    """Accepts a WORKLOADS entry or the legacy tuple (rows, dim, kind, rows_per_doc, users, roles, desc) the tools pass."""
    if isinstance(wl, dict):
        return wl
    rows, dim, kind, rpd, users, roles, desc = wl
    return dict(rows=rows, dim=dim, kind=kind, rpd=rpd, users=users, roles=roles, rbac="tree", metric="l2", filter="bitmask",
                engine="scan", desc=desc)


def build_rbac(wl):
    import vectorsearch_rbac_b200.synth as synth
    wl = as_workload(wl)
    tid, doc = synth.row_ids(wl["rows"], wl["rpd"])
    ndoc = (wl["rows"] + wl["rpd"] - 1) // wl["rpd"]
    doc_ids = np.arange(1, ndoc + 1)
    if wl["rbac"] == "random":
        # RandomRBACDataGenerator(m_roles=3, m_perms=9*docs/roles)
        ur, dr = synth.random_rbac(wl["users"], wl["roles"], doc_ids.tolist(), 3, 9 * ndoc // wl["roles"], seed=42)
    else:
        ur, dr = synth.tree_rbac(wl["users"], wl["roles"], doc_ids.tolist(), h=4, b0=3, b1=4, seed=42)
    return tid, doc, ur, dr


def make_queries(wl, nq, device):
    import torch
    import vectorsearch_rbac_b200.synth as synth
    wl = as_workload(wl)
    qrows, qusers = synth.queries(wl["rows"], np.arange(1, wl["users"] + 1), nq, seed=7)
    Q = synth.corpus(wl["kind"], torch.as_tensor(qrows, device=device), wl["dim"], seed=1234)
    return Q, qusers


def gen_rows(wl, row_idx, device):
    import torch
    import vectorsearch_rbac_b200.synth as synth
    wl = as_workload(wl)
    return synth.corpus(wl["kind"], torch.as_tensor(row_idx, device=device), wl["dim"], seed=1234)


def tf32_peak(device):
    """cuBLAS TF32 GEMM peak measured the way MEASURED_PEAKS.json measures bf16: 8192^3, best of 10, CUDA events
    (the driver's file has no TF32 figure; SURVEY.md §8d asks for one measured the same way)."""
    import torch
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        a = torch.randn(8192, 8192, device=device)
        b = torch.randn(8192, 8192, device=device)
        for _ in range(3):
            a @ b
        best = float("inf")
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            a @ b
            e1.record()
            torch.cuda.synchronize(device)
            best = min(best, e0.elapsed_time(e1))
        return 2.0 * 8192 ** 3 / (best * 1e-3) / 1e12
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev


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

    def _sample(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake": 0x80, "sync_boost": 0x10, "display_clock_setting": 0x100}
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

    def _run(self):
        # The first sample is taken by __enter__ (right after the warm-up steps, before the first timed launch); the thread
        # then samples every 10 ms.  An NVML query takes the driver about a millisecond and competes with the launching
        # thread, which a timed region of a few 1-ms steps cannot absorb (8 ranks, C4: 1.36 ms per step with a query
        # inside the region against 0.9 ms without), so regions shorter than 10 ms carry the pre-region sample only.
        while not self._stop.wait(0.01):
            self._sample()

    def __enter__(self):
        if self.nv:
            self._sample()
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
# single-query legs: what one pgvector-shaped scan costs (one query per rescan, ivfscan.c:339-389)

def depth_users(ur, dr, per_depth=4):
    """Per distinct visible-set size (tree RBAC: role depth t = 1..4 sees t x rows / sets), smallest first: up to `per_depth`
    users whose roles share as few documents as possible (roles under different top-level ancestors share none), so that
    cycling through them re-reads nothing from L2.  Returns [(documents visible, [users])]."""
    docs_of_role = {}
    for d, r in dr:
        docs_of_role.setdefault(int(r), set()).add(int(d))
    user_of_role = {}
    for u, r in ur:
        user_of_role.setdefault(int(r), int(u))
    by_size = {}
    for r, ds in docs_of_role.items():
        if r in user_of_role:
            by_size.setdefault(len(ds), []).append(r)
    out = []
    for sz in sorted(by_size):
        picked, seen = [], set()
        for r in sorted(by_size[sz]):
            if not (docs_of_role[r] & seen):
                picked.append(r)
                seen |= docs_of_role[r]
            if len(picked) == per_depth:
                break
        out.append((sz, [user_of_role[r] for r in picked]))
    return out


def single_query_legs(ix, wl, ur, dr, dev, k, flt, layout, reps=24, hbm_peak=None):
    """Device and end-to-end time of ONE query through the host-buffer C-ABI: an unfiltered full-corpus query and filtered
    queries per role depth.  Two cache states per case:
      rotating  consecutive calls come from different users whose visible sets are disjoint and together exceed the 126 MB
                L2 (the full corpus exceeds it by itself): row data comes from HBM, the kernel's code stays cached — the
                steady state of a server answering one query after another;
      flushed   a 512 MB write before every call evicts everything, code included (cold start).
    device_us = CUDA events around the call's kernels (one fused launch per query); e2e_us = wall clock of
    rk_search_batch(nq = 1) with host buffers, transfers inside; rescan_us = rk_rescan + rk_gettuple_batch on a scan handle.
    frac = (algorithmic bytes / measured HBM peak) / time."""
    import torch
    import vectorsearch_rbac_b200 as vr
    wl = as_workload(wl)
    peak = hbm_peak or peaks()[0]
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    Q, _ = make_queries(wl, 8, dev)
    Qh = torch.empty(Q.shape, dtype=torch.float32, pin_memory=True)
    Qh.copy_(Q)
    torch.cuda.synchronize(dev)
    o_tid = torch.empty((1, k), dtype=torch.int64, pin_memory=True)
    o_doc = torch.empty((1, k), dtype=torch.int32, pin_memory=True)
    o_dist = torch.empty((1, k), dtype=torch.float64, pin_memory=True)
    o_cnt = torch.empty(1, dtype=torch.int32, pin_memory=True)
    cases = [("unfiltered full corpus", vr.FILTER_NONE, 0, [0])]
    for t, (sz, users) in enumerate(depth_users(ur, dr), 1):
        cases.append((f"filtered, role depth {t} ({sz} documents visible)", flt, layout, users))
    sc = ix.begin_scan()
    legs = []
    lib = vr._lib.lib
    qn = Qh.numpy()
    for name, f, lay, users in cases:
        opts = vr.engine.make_opts(k, filter=f, layout=lay, engine=vr.ENGINE_SCAN)
        us = [np.array([u], dtype=np.uint32) for u in users]

        def call(i):
            rc = lib.rk_search_batch(ix._h, 1, Qh[i % 8].data_ptr(), us[i % len(us)].ctypes.data, opts, o_tid.data_ptr(),
                                     o_doc.data_ptr(), o_dist.data_ptr(), o_cnt.data_ptr())
            if rc < 0:
                raise RuntimeError(vr._lib.last_error())

        leg = {"case": name, "users_rotated": len(users)}
        for mode in ("rotating", "flushed"):
            def pre():
                if mode == "flushed":
                    flush.zero_()
                    torch.cuda.synchronize(dev)
            for i in range(2 * len(us) + 2):
                call(i)
            dev_us, e2e_us, rs_us = [], [], []
            ix.enable_timing(True)
            for i in range(reps):
                pre()
                call(i)
                st = ix.stats()
                dev_us.append((st["scan_ms"] + (st["select_ms"] if not st["single_fused"] else 0.0)) * 1e3)
            ix.enable_timing(False)
            st = ix.stats()
            for i in range(reps):
                pre()
                t0 = time.perf_counter()
                call(i)
                e2e_us.append((time.perf_counter() - t0) * 1e6)
            for i in range(reps):
                pre()
                t0 = time.perf_counter()
                sc.rescan(qn[i % 8], users[i % len(users)], k=k, filter=f, layout=lay)
                got = sc.gettuple_batch(k)
                rs_us.append((time.perf_counter() - t0) * 1e6)
            roof_us = st["algo_bytes"] / (peak * 1e9) * 1e6
            d, e, r = float(np.median(dev_us)), float(np.median(e2e_us)), float(np.median(rs_us))
            leg.update({"rows_scanned": int(st["rows_scanned"]), "algorithmic_bytes": int(st["algo_bytes"]), "roofline_us": roof_us,
                        "launches_per_query": int(st["launches"]), "fused_single_launch": bool(st["single_fused"]),
                        "segments": int(st["segments"]), "count": int(o_cnt[0]), "rescan_count": int(len(got[0]))})
            leg[mode] = {"device_us": d, "frac_device": roof_us / d if d > 0 else None, "e2e_us": e,
                         "frac_e2e": roof_us / e if e > 0 else None, "rescan_gettuple_us": r}
        legs.append(leg)
    sc.end()
    del flush
    return {"reps": reps, "k": k, "hbm_peak_gbs": peak, "legs": legs,
            "cache_states": "rotating = consecutive calls from users with disjoint visible sets (together > L2), code cached; "
                            "flushed = 512 MB write before every call (data and code cold)"}


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


def cpu_search(X, tid, parts, qparts, Q, k, nthreads, metric="l2"):
    """(tid, dist, count), kind — the reference's own loops (oracle/_ref) inside the oracle's scan when
    that build is present, else the oracle port with pgvector's Makefile flags."""
    from oracle import oracle as orc
    ref = orc.ref_loops()
    kind = "reference" if ref is not None else "port"
    m = orc.METRIC_L2 if metric == "l2" else orc.METRIC_IP
    t, _, d, _, c = orc.search_batch(X, tid, parts, qparts, Q, m, k, dedup=True, nthreads=nthreads, variant="pgvflags", ref=ref)
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


def make_config(wl, args, world):
    """The workload-defining keys of the JSON line — identical in the b200 arm and the reference arm (what differs
    between the arms, engine and per-rank sizes, goes under "run")."""
    wl = as_workload(wl)
    gb = wl["rows"] * wl["dim"] * 4 / max(1, world) / 1e9
    cfg = {"workload": wl["desc"], "rows": wl["rows"], "dim": wl["dim"], "k": args.k, "queries_per_step": args.nq,
           "filter": args.filter or wl["filter"], "metric_op": wl["metric"], "rbac": wl["rbac"], "users": wl["users"], "roles": wl["roles"],
           "l2_policy": "no flush between steps: the row store (%.2f GB per GPU at N = %d) is %.0fx the 126 MB L2.  Queries of one step that "
                        "walk the same rows share them through L2 (scheduled that way), so DRAM traffic is below the algorithmic bytes: "
                        "roofline.frac is algorithmic bytes / time, roofline.dram_frac is the measured DRAM bytes / time; the "
                        "single_query legs (one query, L2 flushed or rotated out) are the no-reuse HBM numbers" % (gb, world, gb * 1e3 / 126.0)}
    if wl.get("ivf_lists"):
        cfg.update({"lists": int(wl["ivf_lists"]), "nprobe": args.nprobe})
    return cfg


def run_reference(args, wl):
    """--impl reference: the reference's CPU implementation of the path (its own distance loops from oracle/_ref inside
    the oracle's restated scan + top-k + merge), all host threads, same config / metric / unit as the b200 arm.  Uses the
    CPU only: the corpus comes from oracle/synthgen.c (bit-identical to the package's generator), and neither
    librbacknn.so nor CUDA is loaded."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if wl.get("ivf_lists"):
        # the CPU leg of c4 scans the lists of the index the GPU arm builds (k-means centres are part of the workload);
        # it is timed inside that arm's cpu_baseline
        emit({"impl": "reference", "unavailable": "workload c4: the IVFFlat CPU leg runs inside the b200 arm (cpu_baseline) on the index it builds"})
        return
    from oracle import oracle as orc
    import vectorsearch_rbac_b200.synth as synth          # input generators only; does not load librbacknn.so
    rows, dim = wl["rows"], wl["dim"]
    cores = host_cores()
    tid, doc, ur, dr = build_rbac(wl)
    qrows, qusers = synth.queries(rows, np.arange(1, wl["users"] + 1), args.nq, seed=7)
    if wl["kind"] == "sift":
        X = orc.gen_sift(np.arange(rows, dtype=np.int64), dim, 1234, nthreads=cores)
        Q = orc.gen_sift(qrows, dim, 1234, nthreads=1)
        gen = "oracle/synthgen.c on the host (bit-identical to the b200 arm's corpus)"
    else:
        # the Cohere shape is floating point (log / cos): generated with torch on the host; last-bit differences from the
        # GPU-generated corpus of the b200 arm do not matter to a timing
        import torch
        X = np.empty((rows, dim), dtype=np.float32)
        for a in range(0, rows, 1 << 18):
            b = min(rows, a + (1 << 18))
            X[a:b] = synth.corpus(wl["kind"], torch.arange(a, b), dim, seed=1234).numpy()
        Q = synth.corpus(wl["kind"], torch.as_tensor(qrows), dim, seed=1234).numpy()
        gen = "torch on the host"
    parts, qparts = cpu_partitions(doc, ur, dr, qusers)
    # every step times ALL queries of the b200 arm's step unless that cannot end within ~3 minutes; then a bounded prefix
    n0 = min(args.nq, max(cores, 8))
    t0 = time.perf_counter()
    cpu_search(X, tid, parts, qparts[:n0], Q[:n0], args.k, cores, wl["metric"])
    per_q = (time.perf_counter() - t0) / n0
    total = max(1, args.steps + args.warmup)
    S = args.nq if per_q * args.nq * total <= 170.0 else int(max(min(cores, args.nq), min(args.nq, 170.0 / total / max(per_q, 1e-6))))
    for _ in range(args.warmup):
        cpu_search(X, tid, parts, qparts[:S], Q[:S], args.k, cores, wl["metric"])
    t0 = time.perf_counter()
    kind_used = "port"
    for _ in range(args.steps):
        _, kind_used = cpu_search(X, tid, parts, qparts[:S], Q[:S], args.k, cores, wl["metric"])
    dt = time.perf_counter() - t0
    qps = S * args.steps / dt
    sample = (f"{S} of the {args.nq} queries of the step per timed step, full visible row sets, "
              f"{cores} threads on {cpu_model()}; pgvector distance loops without PostgreSQL page/tuple/fmgr overhead; corpus from {gen}")
    line = {
        "impl": "reference", "metric": METRIC, "value": qps, "unit": "queries/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": make_config(wl, args, args.gpus),
        "run": {"queries_timed_per_step": S, "threads": cores},
        "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": cores, "kind": kind_used, "sample": sample},
        "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def cpu_single_thread(X, tid, parts, qparts, Q, k, metric, budget_s=4.0):
    """The harness's own protocol (basic_benchmark/common_function.py:1366-1382, 1414): per query two warm-up calls and one
    timed call on ONE thread, QPS = 1 / mean latency; with the reference's distance loops at pgvector's Makefile flags
    (-O2 -ftree-vectorize ...) and as the reference ships them (compile_pgvector.sh:7: -g -O0)."""
    from oracle import oracle as orc
    m = orc.METRIC_L2 if metric == "l2" else orc.METRIC_IP
    out = {"protocol": "per query 2 warm-up calls + 1 timed call, one thread, QPS = 1 / mean latency (common_function.py:1366-1382, 1414)"}
    for name, o0 in (("O2_pgvector_makefile_flags", False), ("O0_as_shipped_compile_pgvector_sh", True)):
        ref = orc.ref_loops(o0=o0)
        if ref is None:
            out[name] = None
            continue
        lat, t_start = [], time.perf_counter()
        for i in range(len(Q)):
            if not qparts[i]:
                continue
            p = [parts[qparts[i][0]]]
            for rep in range(3):
                t0 = time.perf_counter()
                orc.search_batch(X, tid, p, [[0]], Q[i:i + 1], m, k, dedup=True, nthreads=1, variant="pgvflags", ref=ref)
                dt = time.perf_counter() - t0
            lat.append(dt)
            if time.perf_counter() - t_start > budget_s and len(lat) >= 2:
                break
        out[name] = {"value": 1.0 / float(np.mean(lat)), "unit": "queries/s", "queries": len(lat), "mean_latency_ms": float(np.mean(lat)) * 1e3}
    return out


# ------------------------------------------------------------------------------------------------

def run_b200(args, wl_key, env, steps, cpu_budget_s, leg=False):
    """One workload on this rank's GPU; returns the JSON line (rank 0) or None.  leg=True: a short run for the "workloads"
    object of the default line (no single-query legs, no single-thread CPU protocol)."""
    import torch
    import torch.distributed as dist
    import vectorsearch_rbac_b200 as vr
    from vectorsearch_rbac_b200 import sharded

    wl = dict(WORKLOADS[wl_key])
    if args.rows and not leg:
        wl["rows"] = args.rows
    rank, world, local_rank, dev = env["rank"], env["world"], env["local_rank"], env["dev"]
    rows, dim = wl["rows"], wl["dim"]
    flt_name = (args.filter if not leg else "") or wl["filter"]
    metric = vr.METRIC_L2 if wl["metric"] == "l2" else vr.METRIC_IP
    flt = {"bitmask": vr.FILTER_BITMASK, "partition": vr.FILTER_PARTITION, "combination": vr.FILTER_PARTITION,
           "dynamic": vr.FILTER_PARTITION}[flt_name]
    layout = {"bitmask": 0, "partition": vr.LAYOUT_ROLE, "combination": vr.LAYOUT_COMBINATION, "dynamic": vr.LAYOUT_DYNAMIC}[flt_name]
    engine_name = (args.engine if not leg else "") or wl["engine"]
    engine = {"scan": vr.ENGINE_SCAN, "tensor": vr.ENGINE_TENSOR, "auto": vr.ENGINE_AUTO}[engine_name]
    no_cpu = args.no_cpu

    # ---- load this rank's shard ---------------------------------------------------------------
    tid, doc, ur, dr = build_rbac(wl)
    mine = np.flatnonzero(sharded.local_row_mask(doc, rank, world))
    if args.emu_world > 1 and world == 1:
        mine = np.flatnonzero(sharded.local_row_mask(doc, args.emu_rank, args.emu_world))
        no_cpu = True
    ix = vr.Index(dim, metric=metric, device=local_rank, k_max=max(128, args.k))
    ix.reserve_rows(len(mine))
    ivf_lists = int(wl.get("ivf_lists", 0))
    keep_host = world == 1 and rank == 0 and not no_cpu and not ivf_lists
    X_host = np.empty((rows, dim), dtype=np.float32) if keep_host else None
    CH = 1 << 20 if dim <= 256 else 1 << 18
    for a0 in range(0, len(mine), CH):
        idx = mine[a0:a0 + CH]
        xb = gen_rows(wl, idx, dev)
        torch.cuda.current_stream(dev).synchronize()      # the engine loads on its own stream: the rows must be complete
        ix.load_rows_device(xb.data_ptr(), len(idx), tid[idx], doc[idx])
        if keep_host:
            X_host[idx] = xb.cpu().numpy()
        del xb
    ix.set_doc_roles(dr[:, 0], dr[:, 1])
    ix.set_user_roles(ur[:, 0], ur[:, 1])
    if flt_name == "partition":
        ix.build_role_partitions()
    if flt_name == "combination":
        ix.build_combination_partitions()
    layout_info = None
    if flt_name == "dynamic":
        from vectorsearch_rbac_b200 import dynamic_layout
        parts_d, roleset_parts, layout_info = dynamic_layout.split_roles(dr, ur, alpha=2.0)
        dynamic_layout.apply_layout(ix, parts_d, roleset_parts)
    ix.finalize()
    ivf_build_s = None
    if ivf_lists:
        t_b = time.perf_counter()
        # N > 1: rank 0 trains the centres on its stripe, all ranks assign their rows to them (sharded.build_ivf_sharded)
        sharded.build_ivf_sharded(ix, rank, world, ivf_lists, seed=42, device=dev)
        torch.cuda.synchronize()
        ivf_build_s = time.perf_counter() - t_b
    torch.cuda.empty_cache()

    Q, qusers = make_queries(wl, args.nq, dev)
    Q = Q.contiguous()
    stream = torch.cuda.current_stream(dev)
    sharded.use_torch_stream(ix, dev)
    searcher = sharded.ShardedSearcher(ix, rank, world, device=dev)
    peer_ok = world > 1 and args.exchange == "peer" and searcher.enable_peer_exchange(args.nq, max(args.k, 1))
    searcher.peer_overlap = args.peer_overlap
    # with the peer exchange the merge of step i runs on a side stream under the scan of step i + 1 (no collective's CTAs
    # compete with the persistent scan kernel); with NCCL the serial order measured faster (round 1) and stays the default
    pipelined = world > 1 and not args.no_pipeline and (peer_ok or args.pipeline)
    ix.enable_timing(True)
    Qh = torch.empty((args.nq, dim), dtype=torch.float32, pin_memory=True)
    Qh.copy_(Q)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def measure(eng):
        """One engine over the step: device-resident `value` leg, then the host-buffer `e2e` leg."""
        opts = vr.engine.make_opts(args.k, probes=args.nprobe if ivf_lists else 0, filter=flt, layout=layout, engine=eng)
        # N > 1: scan, all-gather and merge of a step run back to back on one stream (stages_ms reports each).  With
        # --pipeline the exchange of step i moves to a side stream under the scan of step i + 1.
        step_fn = searcher.search_device_pipelined if pipelined else searcher.search_device
        if ivf_lists and world == 1:
            def step_fn(Qd, users, o):       # the index-probe path of the C-ABI (one GPU)
                bf = searcher._get_bufs(Qd.shape[0], o.k)
                ix.search_batch_device(Qd.data_ptr(), Qd.shape[0], users, o, bf["tid"].data_ptr(), bf["doc"].data_ptr(),
                                       bf["dist"].data_ptr(), bf["cnt"].data_ptr())
                return bf["tid"], bf["doc"], bf["dist"], bf["cnt"]
        for _ in range(max(3, args.warmup)):
            step_fn(Q, qusers, opts)
        searcher.flush()
        barrier()
        launches0 = ix.stats()["total_launches"]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        searcher.stage_events = [] if world > 1 else None
        with ClockSampler(local_rank) as clk:
            e0.record(stream)
            t_host = 0.0
            for _ in range(steps):
                th0 = time.perf_counter()
                out = step_fn(Q, qusers, opts)
                t_host += time.perf_counter() - th0
            searcher.flush()
            e1.record(stream)
            barrier()
        r = {"dt_dev": e0.elapsed_time(e1) * 1e-3, "clk": clk.summary(), "stages": searcher.stage_ms(), "pipelined": pipelined,
             "host_issue_ms": t_host / steps * 1e3}
        searcher.stage_events = None
        if ivf_lists:
            ix.synchronize()                 # collects the last launch's probe / list-scan event times
        st = ix.stats()
        r["stats"] = st
        r["launches"] = st["total_launches"] - launches0
        r["scan_ms"] = ix.scan_times_ms(min(64, steps))
        res_dev = tuple(t.clone() for t in out)
        r["res_dev"] = res_dev
        outs_h = tuple(torch.empty(t.shape, dtype=t.dtype, pin_memory=True) for t in res_dev)
        r["h2d"] = Qh.numel() * 4
        r["d2h"] = sum(t.numel() * t.element_size() for t in outs_h)

        def e2e_step():
            if world == 1:
                # the C-ABI call with host pointers: copies in, search, copies out, synchronous
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
        for _ in range(steps):
            e2e_step()
        barrier()
        r["dt_e2e"] = time.perf_counter() - t0
        r["same"] = all(torch.equal(x.cpu(), y) for x, y in zip(res_dev, outs_h))
        r["outs_h"] = outs_h
        if world > 1:
            t = torch.tensor([r["dt_dev"], r["dt_e2e"]], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            r["dt_dev"], r["dt_e2e"] = float(t[0]), float(t[1])
            tl = torch.tensor([r["launches"]], device=dev, dtype=torch.int64)
            dist.all_reduce(tl, op=dist.ReduceOp.SUM)
            r["launches"] = int(tl[0])
        return r

    def with_traffic(rf, key):
        """roofline.traffic = dram bytes of one launch of this kernel from the committed `ncu --set full` capture
        (profiles/traffic.json); dram_frac = those bytes over the live kernel time against the same peak; l2_share = the part
        of the algorithmic bytes that did not come from DRAM."""
        tr = traffic_per_launch(wl_key, world, key) if (not args.rows or leg) and args.emu_world <= 1 else None
        rf["traffic"] = tr
        if tr and rf.get("kernel_ms") and rf.get("unit") == "GB/s":
            algo = rf.get("algorithmic_bytes_per_launch") or rf.get("bytes_per_launch")
            rf["dram_GBps"] = tr / (rf["kernel_ms"] * 1e-3) / 1e9
            rf["dram_frac"] = rf["dram_GBps"] / rf["peak"]
            if algo:
                rf["l2_share"] = max(0.0, 1.0 - tr / float(algo))
        if rf.get("bound") == "hbm" and rf.get("frac", 0) > 1.0:
            rf["note"] = ("algorithmic bytes / time is above the HBM peak because part of the algorithmic bytes is served from L2 "
                          "(queries of one document class stream the same contiguous rows at the same time): dram_frac is the "
                          "DRAM-based fraction, l2_share the part that did not come from DRAM")
        return rf

    def roofline_of(r, eng_name):
        """The dominant kernel of the step against the roofline that bounds it."""
        st = r["stats"]
        kms = float(np.mean(r["scan_ms"])) if len(r["scan_ms"]) else float("nan")
        if ivf_lists:
            # IVFFlat probe: SURVEY.md §8(d) B = lists x 4d + rows of the probed lists the user may see x 4d + 12k per query;
            # time = the centre-scoring kernel + the list-scan kernel (CUDA events of both, last launch)
            peak, peak_src = peaks()
            tms = float(st["scan_ms"] + st["probe_ms"])
            ach = st["algo_bytes"] / (tms * 1e-3) / 1e9 if tms > 0 else 0.0
            return {"bound": "hbm", "kernel": "rk_ivf_probe_kernel + rk_scan_kernel (list segments)", "achieved": ach, "peak": peak,
                    "unit": "GB/s", "frac": ach / peak, "frac_of_8TBps": ach / 8000.0, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": int(st["algo_bytes"]), "kernel_ms": tms, "probe_ms": float(st["probe_ms"]),
                    "list_scan_ms": float(st["scan_ms"]), "launches_timed": 1, "traffic": None}
        if st["tensor_plan"] == 0:
            # CUDA-core scan, one pass per query: HBM.  Bytes per SURVEY.md §8(d): allowed rows x 4d + mask bytes + 12k, per query
            peak, peak_src = peaks()
            ach = st["algo_bytes"] / (kms * 1e-3) / 1e9 if kms > 0 else 0.0
            return with_traffic({"bound": "hbm", "kernel": "rk_scan_kernel", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                                 "frac_of_8TBps": ach / 8000.0, "peak_source": peak_src, "algorithmic_bytes_per_launch": int(st["algo_bytes"]),
                                 "kernel_ms": kms, "launches_timed": int(len(r["scan_ms"]))}, flt_name + ":" + eng_name)
        # tensor engine.  SURVEY.md §8(d), batched exact: F = 2 nq_g R_g d over the groups the plan forms
        # (dense plan: all queries x all rows), bytes 4d (R_g + nq_g); the launch sits in whichever regime its
        # arithmetic intensity F / bytes puts it (ridge = TF32 peak / HBM peak).
        plan = {1: "dense", 2: "per user mask", 3: "per document class (atoms)"}[st["tensor_plan"] & 15]
        hbm, hbm_src = peaks()
        tf32 = env.setdefault("tf32_peak", None) or tf32_peak(dev)
        env["tf32_peak"] = tf32
        flop = float(st["tensor_flop"])
        byts = float(st["algo_bytes"])
        intensity = flop / max(byts, 1.0)
        ridge = tf32 * 1e12 / (hbm * 1e9)
        common = {"kernel": "rk_bscan2_kernel (CTA pairs, tcgen05 cta_group::2)" if st["tensor_plan"] & 32 else "rk_bscan_kernel",
                  "tensor_redo": int(st["tensor_redo"]), "plan": plan + (" + sampling pass" if st["tensor_plan"] & 16 else ""),
                  "kernel_ms": kms, "launches_timed": int(len(r["scan_ms"])), "flop_per_launch": flop, "bytes_per_launch": byts,
                  "plan_bytes_per_launch": float(st["tensor_plan_bytes"]),
                  "intensity_flop_per_byte": intensity, "ridge_flop_per_byte": ridge}
        if intensity >= ridge and st["tensor_plan"] & 64:
            # bf16 nomination: against the driver-measured sustained bf16 peak (MEASURED_PEAKS.json), the TF32 figure beside it
            ach = flop / (kms * 1e-3) / 1e12 if kms > 0 else 0.0
            pk = peaks_all().get("bf16_tflops_sustained") or 1338.0
            return with_traffic({"bound": "tensor", "achieved": ach, "peak": pk, "unit": "TFLOP/s", "frac": ach / pk,
                                 "peak_source": "measured (MEASURED_PEAKS.json bf16_tflops_sustained)" if peaks_all().get("bf16_tflops_sustained") else "fallback 1338",
                                 "operands": "bf16 shadow copies of rows and queries, fp32 accumulation in TMEM; exact fp32 re-score + sufficiency proof after",
                                 "frac_of_tf32_peak_measured_live": ach / tf32, "tf32_peak_measured": tf32, **common}, flt_name + ":" + eng_name)
        if intensity >= ridge:
            ach = flop / (kms * 1e-3) / 1e12 if kms > 0 else 0.0
            bf16 = peaks_all().get("bf16_tflops_sustained")
            return with_traffic({"bound": "tensor", "achieved": ach, "peak": tf32, "unit": "TFLOP/s", "frac": ach / tf32,
                                 "peak_source": "measured live: cuBLAS TF32 (torch.matmul, allow_tf32) 8192^3, best of 10 — MEASURED_PEAKS.json has no TF32 figure",
                                 "frac_of_nominal_1100": ach / 1100.0,
                                 "frac_of_measured_bf16_sustained": (ach / bf16) if bf16 else None, **common}, flt_name + ":" + eng_name)
        ach = byts / (kms * 1e-3) / 1e9 if kms > 0 else 0.0
        return with_traffic({"bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s", "frac": ach / hbm, "peak_source": hbm_src,
                             "tf32_tflops_achieved": flop / (kms * 1e-3) / 1e12 if kms > 0 else 0.0, "tf32_peak_measured": tf32, **common},
                            flt_name + ":" + eng_name)

    main_r = measure(engine)
    batched_r = None
    if engine == vr.ENGINE_SCAN and not args.no_batched and not ivf_lists:
        batched_r = measure(vr.ENGINE_AUTO)

    ex_exact = None
    if ivf_lists:
        # the engine's own exact filtered result (probes = 0) of the step's queries, for recall@k: a sharded search at
        # N > 1, so every rank takes part
        opts0 = vr.engine.make_opts(args.k, filter=flt, layout=layout, engine=vr.ENGINE_AUTO)
        ex_exact = [t.cpu().numpy() for t in searcher.search_device(Q, qusers, opts0)]
        barrier()

    # ---- replicas (N > 1; SURVEY.md §8e "replicas only ... separate line"): every GPU holds the WHOLE corpus (it fits: C3 is
    # 5.12 GB) and answers its own 1/N slice of the step's queries; no exchange at all.  Reported beside the sharded headline,
    # never instead of it: sharding is what a corpus larger than one GPU needs, replication is the better layout when it fits.
    replicas = None
    if world > 1 and not leg and not args.no_replicas and not ivf_lists and rows * dim * 4 <= 40e9:
        ixr = vr.Index(dim, metric=metric, device=local_rank, k_max=max(128, args.k))
        ixr.reserve_rows(rows)
        for a0 in range(0, rows, CH):
            idx = np.arange(a0, min(rows, a0 + CH))
            xb = gen_rows(wl, idx, dev)
            torch.cuda.current_stream(dev).synchronize()
            ixr.load_rows_device(xb.data_ptr(), len(idx), tid[idx], doc[idx])
            del xb
        ixr.set_doc_roles(dr[:, 0], dr[:, 1])
        ixr.set_user_roles(ur[:, 0], ur[:, 1])
        if flt_name == "partition":
            ixr.build_role_partitions()
        if flt_name == "combination":
            ixr.build_combination_partitions()
        ixr.finalize()
        sharded.use_torch_stream(ixr, dev)
        per = (args.nq + world - 1) // world
        lo, hi = min(args.nq, rank * per), min(args.nq, (rank + 1) * per)
        nloc = hi - lo
        rb = [torch.empty((max(nloc, 1), args.k), dtype=t, device=dev) for t in (torch.int64, torch.int32, torch.float64)] + \
             [torch.empty(max(nloc, 1), dtype=torch.int32, device=dev)]
        Qloc = Q[lo:hi].contiguous()
        replicas = {}
        for name_, eng_ in (("scan", vr.ENGINE_SCAN), ("auto", vr.ENGINE_AUTO)):
            o_ = vr.engine.make_opts(args.k, filter=flt, layout=layout, engine=eng_)

            def rstep():
                if nloc:
                    ixr.search_batch_device(Qloc.data_ptr(), nloc, qusers[lo:hi], o_, *[t.data_ptr() for t in rb])
            for _ in range(max(3, args.warmup)):
                rstep()
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(steps):
                rstep()
            e1.record(stream)
            barrier()
            dt_ = torch.tensor([e0.elapsed_time(e1) * 1e-3], device=dev, dtype=torch.float64)
            dist.all_reduce(dt_, op=dist.ReduceOp.MAX)
            # the slice's answers against the sharded path's answers for the same queries
            same_ = torch.tensor([int(nloc == 0 or all(torch.equal(a[:nloc], b[lo:hi]) for a, b in zip(rb, main_r["res_dev"])))], device=dev)
            dist.all_reduce(same_, op=dist.ReduceOp.MIN)
            replicas[name_] = {"value": args.nq * steps / float(dt_[0]), "unit": "queries/s", "ms_per_step": float(dt_[0]) / steps * 1e3,
                               "identical_to_sharded_results": bool(int(same_[0]))}
        replicas["layout"] = ("every GPU holds all %d rows (%.2f GB) and answers %d of the step's %d queries; no exchange; value = all queries of the "
                              "step / slowest rank's device time" % (rows, rows * dim * 4 / 1e9, per, args.nq))
        ixr.close()
        del ixr, rb
        torch.cuda.empty_cache()

    single = None
    if rank == 0 and world == 1 and not leg and not args.no_single and not ivf_lists and wl["rbac"] == "tree":
        ix.enable_timing(False)
        single = single_query_legs(ix, wl, ur, dr, dev, args.k, flt, layout, reps=args.single_reps)

    line = None
    if rank == 0:
        st = main_r["stats"]
        line = {
            "metric": METRIC, "value": args.nq * steps / main_r["dt_dev"], "unit": "queries/s", "n_gpus": world,
            "steps": steps, "warmup": max(3, args.warmup), "ms_per_step": main_r["dt_dev"] / steps * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32" if st["tensor_plan"] == 0 else ("bf16" if st["tensor_plan"] & 64 else "tf32") + " nomination + f32 exact re-score",
            "data": "synthetic",
            "config": make_config(wl, args, world),
            "run": {"engine": engine_name, "rows_this_rank": int(len(mine)), "segments": int(st["segments"]),
                    "host_issue_ms_per_step": main_r["host_issue_ms"],
                    "recall_at_100": 1.0, "recall_note": "exact search; ids checked against the oracle on the cpu_baseline sample"},
            "roofline": roofline_of(main_r, engine_name),
            "e2e": {"value": args.nq * steps / main_r["dt_e2e"], "unit": "queries/s", "h2d_bytes_per_step": int(main_r["h2d"]),
                    "d2h_bytes_per_step": int(main_r["d2h"]), "ms_per_step": main_r["dt_e2e"] / steps * 1e3,
                    "matches_device_path": bool(main_r["same"])},
            "gpu_launches": int(main_r["launches"]),
            "clocks": main_r["clk"],
        }
        if layout_info is not None:
            line["run"]["dynamic_layout"] = layout_info
        if main_r.get("stages"):
            # rank 0's mean device time per stage of a step (the all-gather also absorbs waiting for the slowest rank)
            line["stages_ms"] = main_r["stages"]
        if world > 1:
            line["run"]["exchange"] = searcher.exchange_kind + (
                ("; one stream, merge of step i-1 after the push of step i (no wait for the slowest rank, nothing competes with the scan for SMs)"
                 if peer_ok and searcher.peer_overlap == "deferred" else "; merge of step i on a side stream under the scan of step i + 1")
                if main_r["pipelined"] else "; exchange + merge after every scan, same stream")
            if peer_ok:
                searcher._ex.status()                   # raises if any wait for a peer timed out
        if single is not None:
            line["single_query"] = single
        if replicas is not None:
            line["replicas"] = replicas
        if batched_r is not None:
            # the same step through RK_ENGINE_AUTO (batches of >= 32 exact scans go to the tcgen05 engine): identical
            # results, every row a group of queries shares read once
            line["batched"] = {
                "engine": "auto -> tensor" if batched_r["stats"]["tensor_plan"] else "auto -> scan",
                "value": args.nq * steps / batched_r["dt_dev"], "unit": "queries/s", "ms_per_step": batched_r["dt_dev"] / steps * 1e3,
                "e2e": {"value": args.nq * steps / batched_r["dt_e2e"], "unit": "queries/s", "ms_per_step": batched_r["dt_e2e"] / steps * 1e3},
                "roofline": roofline_of(batched_r, "auto"), "gpu_launches": int(batched_r["launches"]),
                "identical_to_headline_results": bool(all(torch.equal(x, y) for x, y in zip(main_r["outs_h"], batched_r["outs_h"]))),
                "tensor_redo": int(batched_r["stats"]["tensor_redo"]), "clocks": batched_r["clk"]}
        if ivf_lists:
            ivf_roofline(line, args, wl, ix, vr, torch, dev, mine, doc, ur, dr, qusers, Q, main_r, ex_exact, ivf_lists, ivf_build_s, world, flt)
            line["roofline"] = with_traffic(line["roofline"], flt_name + ":" + engine_name) if args.nprobe == 16 else line["roofline"]
        if ivf_lists and (no_cpu or world != 1):
            line["run"].pop("_exact_tid", None)
            line["run"].pop("_exact_cnt", None)
        if ivf_lists and not no_cpu and world == 1:
            ivf_cpu_leg(line, args, wl, ix, dev, tid, doc, ur, dr, qusers, Q, main_r, ivf_lists, rows, dim, CH)
        if world == 1 and not no_cpu and not ivf_lists:
            cores = host_cores()
            outs_h = main_r["outs_h"]
            parts, qparts = cpu_partitions(doc, ur, dr, qusers)
            Qn = Q.cpu().numpy()
            n0 = min(args.nq, max(cores, 8))
            t0 = time.perf_counter()
            cpu_search(X_host, tid, parts, qparts[:n0], Qn[:n0], args.k, cores, wl["metric"])
            per_q = (time.perf_counter() - t0) / n0
            S = int(max(min(cores, args.nq), min(args.nq, cpu_budget_s / max(per_q, 1e-6))))
            t0 = time.perf_counter()
            (c_tid, c_dist, c_cnt), kind = cpu_search(X_host, tid, parts, qparts[:S], Qn[:S], args.k, cores, wl["metric"])
            dtc = time.perf_counter() - t0
            g_tid = outs_h[0].numpy().view(np.uint64)[:S]
            g_dist = outs_h[2].numpy()[:S]
            g_cnt = outs_h[3].numpy()[:S]
            if wl["kind"] == "sift":
                # integer-valued data: ids, counts and float8 distances bit for bit
                dist_note = "bit-identical"
                ids_ok = bool(np.array_equal(g_tid, c_tid) and np.array_equal(g_cnt, c_cnt))
                dist_ok = bool(np.array_equal(g_dist.view(np.uint64), c_dist.view(np.uint64)))
            else:
                # float data: the parity tests' rule (tests/util.py) — distances within 1e-5 relative, ids identical
                # except inside oracle ties of that width
                from tests.util import assert_topk_matches
                dist_note = "within 1e-5 relative; ids identical outside oracle ties of that width"
                try:
                    assert_topk_matches((g_tid, None, g_dist, g_cnt), (c_tid, None, c_dist, None, c_cnt), exact=False)
                    ids_ok = dist_ok = True
                except AssertionError as e:
                    ids_ok = dist_ok = False
                    dist_note += " — MISMATCH: " + str(e)[:200]
            line["cpu_baseline"] = {
                "value": S / dtc, "unit": "queries/s", "cores": cores, "kind": kind,
                "sample": f"{S} of the step's {args.nq} queries, full visible row sets, {cores} threads on {cpu_model()}; "
                          "pgvector distance loops without PostgreSQL page/tuple/fmgr overhead",
                "gpu_ids_match": ids_ok, "gpu_distances": dist_note, "gpu_distances_match": dist_ok}
            if not leg:
                line["cpu_baseline"]["single_thread"] = cpu_single_thread(X_host, tid, parts, qparts, Qn, args.k, wl["metric"])
    if searcher._ex is not None:
        barrier()                      # no peer may still be writing into this rank's ring
        searcher._ex.close()
    ix.close()
    del searcher, Q, Qh, X_host
    torch.cuda.empty_cache()
    return line


def peaks_all():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f)
    return {}


def ivf_roofline(line, args, wl, ix, vr, torch, dev, mine, doc, ur, dr, qusers, Q, main_r, ex, ivf_lists, ivf_build_s, world, flt):
    """C4: recall@k against the engine's own exact filtered result (probes = 0, pinned to the oracle by the parity tests) and
    the list scan's algorithmic bytes per SURVEY.md §8(d): per query lists x 4d (centres) + VISIBLE rows of the probed lists
    x 4d + mask bytes + 12k.  The engine's own figure counts every row of the probed lists, so the visible rows are
    recounted here from the index's list assignment (numpy bookkeeping on the engine's own list choice; the oracle is
    only the CPU leg's)."""
    dim = wl["dim"]
    got_t, got_c = main_r["outs_h"][0].numpy(), main_r["outs_h"][3].numpy()
    rec = float(np.mean([len(set(ex[0][i, :ex[3][i]].tolist()) & set(got_t[i, :got_c[i]].tolist())) / max(1, int(ex[3][i]))
                         for i in range(args.nq)]))
    _, row_list_ = ix.ivf_info(len(mine))            # N > 1: this rank's stripe of every list (rank 0's GPU is reported)
    lrows_ = np.argsort(row_list_, kind="stable")
    off_ = np.searchsorted(row_list_[lrows_], np.arange(ivf_lists + 1))
    P_ = min(args.nprobe, ivf_lists)
    pl_ = torch.empty((args.nq, P_), dtype=torch.int32, device=dev)
    ix.ivf_probe_device(Q.data_ptr(), args.nq, vr.engine.make_opts(args.k, probes=args.nprobe, filter=flt), pl_.data_ptr())
    ix.synchronize()
    pl_ = pl_.cpu().numpy()
    doc_np = np.asarray(doc)[mine]
    roles_of_, docs_of_ = {}, {}
    for u_, r_ in ur:
        roles_of_.setdefault(int(u_), set()).add(int(r_))
    for d_, r_ in dr:
        docs_of_.setdefault(int(r_), []).append(int(d_))
    vis_cache = {}
    probed_rows = visible_rows = 0
    for i in range(args.nq):
        u_ = int(qusers[i])
        if u_ not in vis_cache:
            v_ = np.zeros(int(np.asarray(doc).max()) + 2, dtype=bool)
            for r_ in roles_of_.get(u_, ()):
                v_[np.asarray(docs_of_.get(r_, []), dtype=np.int64)] = True
            vis_cache[u_] = v_
        for l_ in pl_[i]:
            rr = lrows_[int(off_[l_]):int(off_[l_ + 1])]
            probed_rows += len(rr)
            visible_rows += int(vis_cache[u_][doc_np[rr]].sum())
    rb_ = dim * 4
    b_centres = args.nq * ivf_lists * rb_
    # mask bytes M: lists ordered by document class (<= 128 distinct role masks) read one offset per class and probed
    # list; otherwise row -> document -> role mask, 20 B per probed row
    doc_sets = {}
    for d_, r_ in dr:
        doc_sets.setdefault(int(d_), set()).add(int(r_))
    n_cls = len(set(frozenset(v) for v in doc_sets.values()))
    m_bytes = args.nq * args.nprobe * (n_cls + 1) * 4 if n_cls <= 128 else probed_rows * 20
    b_lists = visible_rows * rb_ + m_bytes + args.nq * 12 * args.k
    rf = line["roofline"]
    # dominant kernel: the list scan (HBM: visible rows of the probed lists).  The centre scoring shares the
    # centre matrix across the batch (a 12.6 MB matrix kept in L2, fp32 FMA work), so it is reported beside
    # it with its own rate and not added into the HBM figure
    lms, pms = rf["list_scan_ms"] * 1e-3, rf["probe_ms"] * 1e-3
    fl_centres = 2.0 * args.nq * ivf_lists * dim * (2 if wl["metric"] == "l2" else 1)
    rf.update({"kernel": "rk_scan_kernel (class-ordered list segments)", "kernel_ms": rf["list_scan_ms"],
               "algorithmic_bytes_per_launch": int(b_lists), "achieved": b_lists / lms / 1e9 if lms > 0 else 0.0,
               "frac": b_lists / lms / 1e9 / rf["peak"] if lms > 0 else 0.0,
               "frac_of_8TBps": b_lists / lms / 1e9 / 8000.0 if lms > 0 else 0.0,
               "probed_rows": int(probed_rows), "visible_rows": int(visible_rows), "document_classes": n_cls,
               "list_bytes": int(b_lists),
               "probe": {"kernels": "rk_ivf_keys_kernel + rk_ivf_order_kernel", "ms": rf["probe_ms"],
                         "fp32_TFLOPs": fl_centres / pms / 1e12 if pms > 0 else None,
                         "centre_bytes_once": int(ivf_lists * rb_), "centre_bytes_per_query_definition": int(b_centres),
                         "survey_8d_GBps_both_kernels": (b_centres + b_lists) / (lms + pms) / 1e9 if lms + pms > 0 else None},
               "note": "SURVEY.md 8(d) counts the centres once per query (lists x 4d); the batched probe reads them once per "
                       "16 queries from L2, so that figure would exceed the HBM peak and is reported only under probe; the "
                       "list scan reads only the rows the user may see (one extent per visible document class and probed list); "
                       "a read-only stream can exceed the copy-measured peak, and queries probing the same lists share rows in L2 "
                       "(traffic = dram bytes of the launch)"})
    if world > 1:
        rf["probe"]["ms"] = None
        rf["probe"]["fp32_TFLOPs"] = None
        rf["probe"]["note"] = ("N > 1: the list choice is split by query across the ranks (nq / N queries each) and "
                               "all-gathered; its time is part of stages_ms.local_search")
        line["run"]["ivf_sharding"] = ("centres trained on rank 0's stripe and broadcast; every list's rows striped with "
                                       "their documents (d mod N); list choice split by query + all-gather")
    line["run"]["_exact_tid"], line["run"]["_exact_cnt"] = ex[0], ex[3]
    line["run"].update({"recall_at_100": rec, "ivf_build_s": ivf_build_s,
                        "recall_note": "against the engine's exact filtered top-k of the same queries (probes = 0)"})


def ivf_cpu_leg(line, args, wl, ix, dev, tid, doc, ur, dr, qusers, Q, main_r, ivf_lists, rows, dim, CH):
    """C4's CPU leg: the oracle's restatement of ivfscan.c with the reference's distance loops, same centres and lists, on a
    sample of the queries; only the rows of the lists the sample probes are brought to the host.  Also the full-size parity
    check of the step, and recall@k of the sample against the ORACLE's exact filtered top-k."""
    from oracle import oracle as orc
    from tests.util import assert_topk_matches
    S = min(args.nq, 32)
    m = orc.METRIC_L2 if wl["metric"] == "l2" else orc.METRIC_IP
    centers, row_list = ix.ivf_info(rows)
    Qn = Q.cpu().numpy()[:S]
    need = set()
    for q in Qn:
        need.update(int(x) for x in orc.ivf_order_lists(centers, q, m, args.nprobe)[0])
    need_rows = np.flatnonzero(np.isin(row_list, np.fromiter(need, dtype=np.uint32)))
    X_sub = np.empty((len(need_rows), dim), dtype=np.float32)
    for a0 in range(0, len(need_rows), CH):
        X_sub[a0:a0 + CH] = gen_rows(wl, need_rows[a0:a0 + CH], dev).cpu().numpy()
    off, lrows = orc.ivf_lists_csr(row_list[need_rows], ivf_lists)
    tid_sub, doc_sub = tid[need_rows], np.asarray(doc)[need_rows]
    roles_of, docs_of = {}, {}
    for u, r in ur:
        roles_of.setdefault(int(u), set()).add(int(r))
    for d_, r in dr:
        docs_of.setdefault(int(r), []).append(int(d_))
    c_tid = np.full((S, args.k), np.uint64(0xFFFFFFFFFFFFFFFF), dtype=np.uint64)
    c_dist = np.full((S, args.k), np.inf)
    c_cnt = np.zeros(S, dtype=np.int32)
    t0 = time.perf_counter()
    for i in range(S):
        vis = np.zeros(int(doc_sub.max()) + 2, dtype=np.uint8)
        for r in roles_of.get(int(qusers[i]), ()):
            dd = np.asarray(docs_of.get(r, []), dtype=np.int64)
            vis[dd[dd < len(vis)]] = 1
        (t_, _, d_), = orc.ivf_scan(X_sub, tid_sub, centers, off, lrows, Qn[i], m, probes=args.nprobe, allowed=vis[doc_sub])
        n = min(args.k, len(t_))
        c_tid[i, :n], c_dist[i, :n], c_cnt[i] = t_[:n], d_[:n], n
    dtc = time.perf_counter() - t0
    g_tid = main_r["outs_h"][0].numpy().view(np.uint64)[:S]
    g_dist, g_cnt = main_r["outs_h"][2].numpy()[:S], main_r["outs_h"][3].numpy()[:S]
    note = "within 1e-5 relative; ids identical outside oracle ties of that width"
    try:
        assert_topk_matches((g_tid, None, g_dist, g_cnt), (c_tid, None, c_dist, None, c_cnt), exact=False)
        ok = True
    except AssertionError as e:
        ok = False
        note += " — MISMATCH: " + str(e)[:200]
    # recall@k against the ORACLE's exact filtered top-k (pgvector's exact scan restated on the CPU over every row the user may
    # see) on a few of the queries: the engine's own exact result, which the line's recall is taken against, is thereby checked
    # at full size as well
    S2 = min(S, 4)
    rec_o, same_exact = [], True
    ex_t, ex_c = line["run"].get("_exact_tid"), line["run"].get("_exact_cnt")
    doc_all = np.asarray(doc)
    for i in range(S2):
        vis = np.zeros(int(doc_all.max()) + 2, dtype=bool)
        for r in roles_of.get(int(qusers[i]), ()):
            vis[np.asarray(docs_of.get(r, []), dtype=np.int64)] = True
        vrows = np.flatnonzero(vis[doc_all])
        Xv = np.empty((len(vrows), dim), dtype=np.float32)
        for a0 in range(0, len(vrows), CH):
            Xv[a0:a0 + CH] = gen_rows(wl, vrows[a0:a0 + CH], dev).cpu().numpy()
        t_, _, _, _, c_ = orc.search_batch(Xv, tid[vrows], [np.arange(len(vrows), dtype=np.uint32)], [[0]] if len(vrows) else [[]], Qn[i:i + 1],
                                           m, args.k, dedup=False, nthreads=host_cores(), variant="pgvflags", ref=orc.ref_loops())
        truth = set(t_[0, :int(c_[0])].tolist())
        rec_o.append(len(truth & set(g_tid[i, :int(g_cnt[i])].tolist())) / max(1, len(truth)))
        if ex_t is not None:
            same_exact = same_exact and set(ex_t[i, :int(ex_c[i])].view(np.uint64).tolist()) == truth
        del Xv
    line["run"]["recall_at_100_vs_oracle_exact"] = {"value": float(np.mean(rec_o)) if rec_o else None, "queries": S2,
                                                    "engine_exact_ids_equal_oracle_exact": bool(same_exact) if ex_t is not None else None}
    line["run"].pop("_exact_tid", None)
    line["run"].pop("_exact_cnt", None)
    line["cpu_baseline"] = {
        "value": S / dtc, "unit": "queries/s", "cores": 1, "kind": "port",
        "sample": f"{S} of the step's {args.nq} queries, nprobe {args.nprobe}, the oracle's restatement of ivfscan.c (centre scoring, "
                  f"full sort of the probed lists' visible tuples) on one thread of {cpu_model()}, the GPU build's centres and lists; "
                  "no PostgreSQL page/tuple/fmgr overhead",
        "gpu_ids_match": ok, "gpu_distances": note, "gpu_distances_match": ok}


def leg_summary(line):
    """What a short leg contributes to the default line's "workloads" object."""
    rf = line["roofline"]
    out = {"workload": line["config"]["workload"], "value": line["value"], "unit": line["unit"], "ms_per_step": line["ms_per_step"],
           "steps": line["steps"], "e2e": line["e2e"]["value"], "engine": line["run"]["engine"], "dtype": line["dtype"],
           "roofline": {k: rf.get(k) for k in ("bound", "kernel", "achieved", "peak", "unit", "frac", "kernel_ms", "dram_frac", "l2_share", "traffic",
                                               "plan", "frac_of_measured_bf16_sustained") if rf.get(k) is not None},
           "gpu_launches": line["gpu_launches"], "recall_at_100": line["run"].get("recall_at_100"),
           "recall_at_100_vs_oracle_exact": line["run"].get("recall_at_100_vs_oracle_exact")}
    if "cpu_baseline" in line:
        cb = line["cpu_baseline"]
        out["cpu_baseline"] = {k: cb.get(k) for k in ("value", "cores", "kind", "gpu_ids_match", "gpu_distances_match")}
    if "batched" in line:
        b = line["batched"]
        out["batched"] = {"value": b["value"], "e2e": b["e2e"]["value"], "frac": b["roofline"]["frac"], "bound": b["roofline"]["bound"],
                          "identical_to_headline_results": b["identical_to_headline_results"]}
    if line["config"].get("nprobe"):
        out["nprobe"] = line["config"]["nprobe"]
    return out


def main():
    args = parse_args()
    wl = dict(WORKLOADS[args.workload])
    if args.rows:
        wl["rows"] = args.rows
    if args.impl == "reference":
        run_reference(args, wl)
        return

    import torch
    import torch.distributed as dist

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
    torch.cuda.set_stream(torch.cuda.Stream(dev, priority=-1))     # a real stream: torch events, NCCL and the engine share it
    env = {"rank": rank, "world": world, "local_rank": local_rank, "dev": dev}
    t_start = time.perf_counter()
    line = run_b200(args, args.workload, env, args.steps, cpu_budget_s=20.0)

    # the default line (C3, one GPU) also carries short legs of the other single-GPU configs of BASELINE.json
    if rank == 0 and world == 1 and args.workload == "c3" and not args.rows and not args.no_legs:
        legs = {}
        for key in ("c1", "c2", "c5", "c4"):
            if time.perf_counter() - t_start > args.legs_budget_s:
                legs[key] = {"skipped": "time budget of the default run (%d s) used up" % args.legs_budget_s}
                continue
            try:
                legs[key] = leg_summary(run_b200(args, key, env, min(args.steps, 5), cpu_budget_s=3.0, leg=True))
            except Exception as e:            # a failing leg must not take the headline line with it
                legs[key] = {"error": repr(e)[:300]}
        line["workloads"] = legs
    if rank == 0:
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
