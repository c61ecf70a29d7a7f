"""sharded.py — one process per GPU: a document-striped corpus searched by every rank, the ranks'
local top-k exchanged with ONE all-gather over NVLink and merged by a G-way select kernel.

The reference has no multi-GPU path; the closest thing it touches is FAISS's
host-thread shards with a CPU merge (logical_partition_benchmark/benchmark/faiss/faiss/
IndexShards.cpp:197-256).  Here:

  partitioning   document d lives on rank  d mod G  (a document's rows and its role mask stay
                 together, services/read_dataset_function.py:338-339 keeps them contiguous), so
                 every role / combination / dynamic partition is split ~evenly over all ranks.
  local search   Index.search_batch_shard -> per query k (fp32 key, doc, tid) triples, ascending,
                 written into ONE packed device buffer  [key nq*k f32 | doc nq*k u32 | tid nq*k u64].
  exchange       torch.distributed.all_gather_into_tensor of that buffer (16*nq*k bytes per rank).
  merge          Index.merge_shards with shard_stride_bytes = 16*nq*k on every rank.

A physical row lives on exactly one rank, so the merge needs no dedup; ties are broken by tuple id
exactly as on one GPU, which makes the result independent of G.

The exchange itself has two forms: NCCL (`all_gather_into_tensor`, the plumbing default on CPU / gloo) and the peer
exchange (`enable_peer_exchange`: every rank's select output is stored straight into the other GPUs' receive rings over
NVLink and the merge kernel waits on arrival flags — rk_exchange.cu), which takes the collective, its launch and its
implicit barrier off the data path so that the merge of step i runs under the scan of step i + 1.

`local_search` / `merge` are injectable so the exchange logic is testable with the gloo backend on
CPU (tests/test_sharded_gloo.py); the defaults are the CUDA paths and there is no CPU fallback.
"""
from __future__ import annotations

from typing import Callable, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.distributed as dist

from . import engine as _engine


def owner_of_doc(doc_id: np.ndarray, world: int) -> np.ndarray:
    """Rank that stores a document: d mod G."""
    return (np.asarray(doc_id, dtype=np.int64) % world).astype(np.int32)


def local_row_mask(row_doc: np.ndarray, rank: int, world: int) -> np.ndarray:
    return owner_of_doc(row_doc, world) == rank


def packed_bytes(nq: int, k: int) -> int:
    return 16 * nq * k


def split_packed(buf: torch.Tensor, nq: int, k: int) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """Views (key f32[nq,k], doc i32[nq,k], tid i64[nq,k]) into one rank's packed exchange buffer."""
    n = nq * k
    key = buf[: 4 * n].view(torch.float32).view(nq, k)
    doc = buf[4 * n: 8 * n].view(torch.int32).view(nq, k)
    tid = buf[8 * n: 16 * n].view(torch.int64).view(nq, k)
    return key, doc, tid


def build_ivf_sharded(index, rank: int, world: int, lists: int, seed: int = 42, group=None,
                      device: Optional[torch.device] = None) -> np.ndarray:
    """IVFFlat over a document-striped corpus ("same centroids replicated on all GPUs, each list's rows
    striped").  Rank 0 trains the centres on its own stripe — documents are striped d mod G, so the stripe is a
    uniform 1/G sample of the table pgvector would sample from (ivfbuild.c:404-441) — the centres are broadcast, and
    every other rank assigns its rows to them (`build_ivf_from_centers`, ivfbuild.c:141-198).  All ranks then score
    the same centres with the same arithmetic, choose the same lists for a query, scan their stripe of those lists,
    and the usual all-gather + merge returns what one GPU holding the whole lists would return.
    Returns the centres.  `device` = where the broadcast buffer lives (cuda for NCCL, cpu for gloo)."""
    if world == 1:
        index.build_ivf(lists, seed=seed)
        return index.ivf_centers()
    dev = device if device is not None else torch.device("cuda", index.device)
    if rank == 0:
        index.build_ivf(lists, seed=seed)
        c = torch.from_numpy(index.ivf_centers()).to(dev)
    else:
        c = torch.empty((lists, index.dim), dtype=torch.float32, device=dev)
    dist.broadcast(c, src=0, group=group)
    centers = c.cpu().numpy()
    if rank != 0:
        index.build_ivf_from_centers(centers)
    return centers


class ShardedSearcher:
    """Search over this rank's shard + all-gather merge.  `index` is this rank's engine.Index (already
    loaded with the rows of the documents it owns, and finalized)."""

    def __init__(self, index, rank: int, world: int, group=None, device: Optional[torch.device] = None,
                 local_search: Optional[Callable] = None, merge: Optional[Callable] = None, shard_probe: bool = True):
        self.index, self.rank, self.world, self.group = index, rank, world, group
        # IVFFlat (opts.probes > 0) on G > 1 ranks: the centres are replicated, so the list choice of the batch is split by
        # query — rank r scores the centres for queries [r * ceil(nq / G), ...) only — and all-gathered (4 * probes bytes
        # per query) before every rank scans its stripe of the chosen lists.  False: every rank chooses for all queries.
        self.shard_probe = shard_probe
        self._ivf_lists = None
        self.device = device if device is not None else torch.device("cuda", index.device)
        self._local_search = local_search or self._cuda_local_search
        self._merge = merge or self._cuda_merge
        self._bufs = {}
        # optional per-stage CUDA events of every call (local search | all-gather | merge), for bench.py's "stages_ms"
        self.stage_events = None
        self._ex = None                  # engine.Exchange once enable_peer_exchange succeeded on every rank
        self.peer_overlap = "deferred"   # pipelined peer exchange: "deferred" = merge of step i-1 after the push of step i, one
                                         # stream; "side" = merge of step i on a low-priority side stream under the scan of step i+1
        self._deferred = None
        self.exchange_kind = "nccl all_gather_into_tensor" if world > 1 else "none (one GPU)"

    def stage_ms(self):
        """Mean device time of the three stages over the calls recorded since stage_events was last reset."""
        ev = self.stage_events or []
        if not ev:
            return None
        torch.cuda.synchronize(self.device)
        n = float(len(ev))
        return {"local_search": sum(a.elapsed_time(b) for a, b, _, _ in ev) / n,
                "all_gather": sum(b.elapsed_time(c) for _, b, c, _ in ev) / n,
                "merge": sum(c.elapsed_time(d) for _, _, c, d in ev) / n}

    # -- default (CUDA) stages ---------------------------------------------------------------
    def _cuda_probe_exchange(self, dQ: torch.Tensor, opts) -> torch.Tensor:
        """Device tensor [>= nq, P] int32 (P = min(probes, lists)): row q = the lists of query q, nearest first."""
        nq = dQ.shape[0]
        if self._ivf_lists is None:
            self._ivf_lists = self.index.ivf_lists()
        P = min(int(opts.probes), self._ivf_lists)
        cnt = (nq + self.world - 1) // self.world
        keyb = ("probe", nq, P)
        if keyb not in self._bufs:
            self._bufs[keyb] = (torch.zeros((cnt, P), dtype=torch.int32, device=self.device),
                                torch.empty((cnt * self.world, P), dtype=torch.int32, device=self.device))
        send, recv = self._bufs[keyb]
        lo = min(nq, self.rank * cnt)
        hi = min(nq, lo + cnt)
        if hi > lo:
            self.index.ivf_probe_device(dQ[lo:hi].data_ptr(), hi - lo, opts, send.data_ptr())
        dist.all_gather_into_tensor(recv, send, group=self.group)
        return recv

    def _cuda_local_search(self, dQ: torch.Tensor, users: np.ndarray, opts, send: torch.Tensor) -> None:
        nq = dQ.shape[0]
        key, doc, tid = split_packed(send, nq, opts.k)
        if int(opts.probes) > 0 and self.world > 1 and self.shard_probe and dist.is_initialized():
            lists = self._cuda_probe_exchange(dQ, opts)
            self.index.search_batch_shard_probed(dQ.data_ptr(), nq, users, opts, lists.data_ptr(), key.data_ptr(),
                                                 tid.data_ptr(), doc.data_ptr())
            return
        self.index.search_batch_shard(dQ.data_ptr(), nq, users, opts, key.data_ptr(), tid.data_ptr(), doc.data_ptr())

    def _cuda_merge(self, recv: torch.Tensor, nq: int, k: int, out) -> None:
        o_tid, o_doc, o_dist, o_cnt = out
        n = nq * k
        base = recv.data_ptr()
        self.index.merge_shards(nq, k, self.world, base, base + 8 * n, base + 4 * n, o_tid.data_ptr(), o_doc.data_ptr(),
                                o_dist.data_ptr(), o_cnt.data_ptr(), shard_stride_bytes=packed_bytes(nq, k))

    # -- peer exchange -----------------------------------------------------------------------
    def enable_peer_exchange(self, nq_max: int, k_max: int, depth: int = 4, make_exchange: Optional[Callable] = None) -> bool:
        """Sets up the NVLink peer exchange (one process per GPU: the rings' CUDA IPC handles travel through
        all_gather_object).  All ranks must call it.  Returns False — and leaves the NCCL all-gather in place — unless every
        rank could map every other rank's ring."""
        if self.world == 1:
            return False
        make_exchange = make_exchange or (lambda: _engine.Exchange(self.index, nq_max, k_max, self.world, self.rank, depth))
        ex, ok = None, 1
        try:
            ex = make_exchange()
            mine = ex.handle()
        except Exception:
            ok, mine = 0, b""
        handles = [None] * self.world
        dist.all_gather_object(handles, mine, group=self.group)
        if ok:
            try:
                for g, h in enumerate(handles):
                    if g != self.rank:
                        if not h:
                            raise RuntimeError("peer without a ring")
                        ex.connect_ipc(g, h)
            except Exception:
                ok = 0
        oks = [None] * self.world
        dist.all_gather_object(oks, ok, group=self.group)
        if not all(oks):
            if ex is not None:
                ex.close()
            return False
        self._ex = ex
        self.exchange_kind = "peer stores over NVLink into every rank's receive ring + arrival flags (rk_exchange.cu), depth %d" % depth
        return True

    def _peer_push(self, dQ: torch.Tensor, users: np.ndarray, opts) -> int:
        """Local search into this rank's block of the next ring slot + push to the peers.  Returns the ring position.  The
        slot's previous use (depth steps ago) may still be read by a merge on the side stream: order after it."""
        if not hasattr(self, "_ring_done"):
            self._ring_done, self._ring_n = [None] * self._ex.depth, 0
        r = self._ring_n % self._ex.depth
        self._ring_n += 1
        if self._ring_done[r] is not None:
            torch.cuda.current_stream(self.device).wait_event(self._ring_done[r])
            self._ring_done[r] = None
        lists = 0
        if int(opts.probes) > 0 and self.shard_probe:
            lists = self._cuda_probe_exchange(dQ, opts).data_ptr()
        self._ex.push(dQ.data_ptr(), dQ.shape[0], users, opts, lists)
        return r

    # -- buffers -----------------------------------------------------------------------------
    def _get_bufs(self, nq: int, k: int, which: int = 0):
        keyb = (nq, k, which)
        if keyb not in self._bufs:
            dev = self.device
            nb = packed_bytes(nq, k)
            self._bufs[keyb] = dict(
                send=torch.empty(nb, dtype=torch.uint8, device=dev),
                recv=torch.empty(nb * self.world, dtype=torch.uint8, device=dev),
                tid=torch.empty((nq, k), dtype=torch.int64, device=dev),
                doc=torch.empty((nq, k), dtype=torch.int32, device=dev),
                dist=torch.empty((nq, k), dtype=torch.float64, device=dev),
                cnt=torch.empty(nq, dtype=torch.int32, device=dev))
        return self._bufs[keyb]

    # -- public ------------------------------------------------------------------------------
    def search_device(self, dQ: torch.Tensor, users: Sequence[int], opts):
        """dQ [nq, dim] fp32 resident on this rank's device (identical on all ranks).  Returns device
        tensors (tid i64 [nq,k] — the uint64 tuple ids bit-cast, doc i32, dist f64, count i32); every
        rank holds the full merged result.  Asynchronous on the current stream."""
        nq, k = dQ.shape[0], opts.k
        users = np.ascontiguousarray(users, dtype=np.uint32)
        b = self._get_bufs(nq, k)
        rec = self.stage_events is not None and self.device.type == "cuda"
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)] if rec else None
        if rec: ev[0].record()
        if self._ex is not None:
            # peer exchange, same stream: the merge kernel itself waits for the peers' arrival flags
            self._issue_deferred_pull()           # (pulls happen in push order)
            self._peer_push(dQ, users, opts)
            if rec: ev[1].record(); ev[2].record()
            self._ex.pull(nq, k, b["tid"].data_ptr(), b["doc"].data_ptr(), b["dist"].data_ptr(), b["cnt"].data_ptr())
            if rec:
                ev[3].record()
                self.stage_events.append(tuple(ev))
            return b["tid"], b["doc"], b["dist"], b["cnt"]
        self._local_search(dQ, users, opts, b["send"])
        if rec: ev[1].record()
        if self.world > 1:
            dist.all_gather_into_tensor(b["recv"], b["send"], group=self.group)
            recv = b["recv"]
        else:
            recv = b["send"]
        if rec: ev[2].record()
        self._merge(recv, nq, k, (b["tid"], b["doc"], b["dist"], b["cnt"]))
        if rec:
            ev[3].record()
            self.stage_events.append(tuple(ev))
        return b["tid"], b["doc"], b["dist"], b["cnt"]

    def search_device_pipelined(self, dQ: torch.Tensor, users: Sequence[int], opts):
        """Same result as search_device, but the exchange (all-gather + G-way merge) of this batch runs on a side
        stream while the current stream is free to start the NEXT batch's scan: with short scans (8 GPUs: 4 ms) the
        exchange and the wait for the slowest rank would otherwise sit between two scans.  Two buffer sets alternate.
        The returned tensors are complete once `flush()` (or the next-but-one call) has ordered the side stream
        before the current one."""
        if self.world == 1 or self.device.type != "cuda":
            return self.search_device(dQ, users, opts)
        nq, k = dQ.shape[0], opts.k
        users = np.ascontiguousarray(users, dtype=np.uint32)
        if not hasattr(self, "_side"):
            # lower priority than the stream the scans run on: the merge's blocks fill SMs as the persistent scan's CTAs retire
            self._side = torch.cuda.Stream(device=self.device, priority=0)
            self._pipe_n = 0
            self._pipe_done = [None, None]
        s = self._pipe_n & 1
        self._pipe_n += 1
        b = self._get_bufs(nq, k, s)
        main = torch.cuda.current_stream(self.device)
        if self._ex is not None and self.peer_overlap == "deferred":
            # One stream, the merge one step behind: [scan i | select i | push i | merge i-1].  By the time merge i-1 runs
            # every peer's push of step i-1 has long arrived (no wait on the slowest rank), and nothing competes with the
            # persistent scan kernel for SMs — a merge on a side stream delays the start of some of the next scan's CTAs,
            # which costs more than it hides (8 GPUs: scan 4.04 ms against 3.88 alone).
            rec = self.stage_events is not None
            ev4 = [torch.cuda.Event(enable_timing=True) for _ in range(4)] if rec else None
            if rec: ev4[0].record(main)
            self._peer_push(dQ, users, opts)
            if rec: ev4[1].record(main); ev4[2].record(main)
            self._issue_deferred_pull()
            if rec:
                ev4[3].record(main)
                self.stage_events.append(tuple(ev4))
            self._deferred = (nq, k, b)
            return b["tid"], b["doc"], b["dist"], b["cnt"]
        if self._ex is not None:
            # the outputs of set s were last written by the merge issued two calls ago: same side stream, so already ordered
            rec = self.stage_events is not None
            ev4 = [torch.cuda.Event(enable_timing=True) for _ in range(4)] if rec else None
            if rec: ev4[0].record(main)
            ring = self._peer_push(dQ, users, opts)
            ev = torch.cuda.Event()
            ev.record(main)
            if rec: ev4[1].record(main)
            self._side.wait_event(ev)
            with torch.cuda.stream(self._side):
                if rec: ev4[2].record(self._side)
                self._ex.pull(nq, k, b["tid"].data_ptr(), b["doc"].data_ptr(), b["dist"].data_ptr(), b["cnt"].data_ptr(),
                              stream=self._side.cuda_stream)
                if rec: ev4[3].record(self._side)
                done = torch.cuda.Event()
                done.record(self._side)
            if rec:
                self.stage_events.append(tuple(ev4))
            self._pipe_done[s] = done
            self._ring_done[ring] = done
            return b["tid"], b["doc"], b["dist"], b["cnt"]
        if self._pipe_done[s] is not None:
            main.wait_event(self._pipe_done[s])          # the exchange that last read this set's send buffer
        self._local_search(dQ, users, opts, b["send"])
        ev = torch.cuda.Event()
        ev.record(main)
        self._side.wait_event(ev)
        with torch.cuda.stream(self._side):
            dist.all_gather_into_tensor(b["recv"], b["send"], group=self.group)
            self.index.set_stream(self._side.cuda_stream)
            try:
                self._merge(b["recv"], nq, k, (b["tid"], b["doc"], b["dist"], b["cnt"]))
            finally:
                self.index.set_stream(main.cuda_stream if main.cuda_stream else 1)
            done = torch.cuda.Event()
            done.record(self._side)
        self._pipe_done[s] = done
        return b["tid"], b["doc"], b["dist"], b["cnt"]

    def _issue_deferred_pull(self) -> None:
        d = getattr(self, "_deferred", None)
        if d is not None:
            nq, k, b = d
            self._ex.pull(nq, k, b["tid"].data_ptr(), b["doc"].data_ptr(), b["dist"].data_ptr(), b["cnt"].data_ptr())
            self._deferred = None

    def flush(self) -> None:
        """Orders every pipelined exchange issued so far before the current stream."""
        self._issue_deferred_pull()
        for d in getattr(self, "_pipe_done", []):
            if d is not None:
                torch.cuda.current_stream(self.device).wait_event(d)

    def search(self, Q_host: torch.Tensor, users: Sequence[int], opts, out_host=None):
        """End-to-end call with HOST buffers: Q_host is a (pinned) CPU tensor [nq, dim]; results come
        back as CPU tensors (pinned when `out_host` supplies them).  Synchronous."""
        dQ = Q_host.to(self.device, non_blocking=True)
        tid, doc, dst, cnt = self.search_device(dQ, users, opts)
        if out_host is None:
            out_host = tuple(torch.empty(t.shape, dtype=t.dtype, pin_memory=self.device.type == "cuda") for t in (tid, doc, dst, cnt))
        for h, d in zip(out_host, (tid, doc, dst, cnt)):
            h.copy_(d, non_blocking=True)
        if self.device.type == "cuda":
            torch.cuda.current_stream(self.device).synchronize()
        return out_host


def use_torch_stream(index, device: Optional[torch.device] = None) -> None:
    """Run the index's kernels on torch's current CUDA stream so torch events and NCCL collectives
    order with them."""
    h = torch.cuda.current_stream(device).cuda_stream
    index.set_stream(h if h else 1)     # 0 is the legacy default stream: name it explicitly (cudaStreamLegacy)


def make_opts(*a, **kw):
    return _engine.make_opts(*a, **kw)
