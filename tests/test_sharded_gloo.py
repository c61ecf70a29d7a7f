"""World-size-2 gloo test of the sharded search plumbing (vectorsearch_rbac_b200/sharded.py):
document striping, the packed exchange buffer, the all-gather and the [G][nq][k] merge layout.
The GPU stages are replaced by checker stand-ins (the oracle scans each rank's shard; a numpy
G-way merge) — this exercises the host-side logic only; the CUDA stages are covered by -m gpu."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import oracle as orc
from tests.util import make_dataset, oracle_role_search


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from vectorsearch_rbac_b200 import sharded
        ds = make_dataset("sift", n=3000, dim=32, rows_per_doc=10, rbac="random", num_users=12, num_roles=6)
        k, nq = 15, 10
        rng = np.random.default_rng(5)
        Q = ds["X"][rng.integers(0, len(ds["X"]), nq)].copy()
        users = np.unique(ds["user_roles"][:, 0])[rng.integers(0, 12, nq)].astype(np.uint32)
        mine = sharded.local_row_mask(ds["doc"], rank, world)
        assert mine.sum() > 0 and set(np.unique(ds["doc"][mine]) % world) == {rank}
        local = dict(ds, X=ds["X"][mine], tid=ds["tid"][mine], doc=ds["doc"][mine])

        def local_search(dQ, us, opts, send):
            t, _, _, s, c = oracle_role_search(local, dQ.numpy(), us, orc.METRIC_L2, opts.k, mode="bitmask")
            key, doc, tid = sharded.split_packed(send, dQ.shape[0], opts.k)
            key.copy_(torch.from_numpy(s))                       # fp32 ordering value, +Inf padding
            tid.copy_(torch.from_numpy(t.view(np.int64)))        # UINT64_MAX padding
            doc.zero_()

        def merge(recv, nq_, k_, out):
            o_tid, o_doc, o_dist, o_cnt = out
            per = sharded.packed_bytes(nq_, k_)
            keys, tids = [], []
            for g in range(world):
                kk, _, tt = sharded.split_packed(recv[g * per:(g + 1) * per], nq_, k_)
                keys.append(kk.numpy()); tids.append(tt.numpy().view(np.uint64))
            keys = np.concatenate(keys, axis=1); tids = np.concatenate(tids, axis=1)
            for q in range(nq_):
                order = np.lexsort((tids[q], keys[q]))[:k_]
                kq, tq = keys[q][order], tids[q][order]
                valid = np.isfinite(kq)
                o_cnt[q] = int(valid.sum())
                o_tid[q] = torch.from_numpy(tq.view(np.int64))
                o_dist[q] = torch.from_numpy(np.sqrt(kq.astype(np.float64)))

        sh = sharded.ShardedSearcher(None, rank, world, device=torch.device("cpu"), local_search=local_search, merge=merge)
        opts = sharded.make_opts(k, filter=1)
        tid, doc, dst, cnt = sh.search(torch.from_numpy(Q), users, opts)
        exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, k, mode="bitmask")
        np.testing.assert_array_equal(cnt.numpy(), exp[4])
        for q in range(nq):
            c = int(exp[4][q])
            np.testing.assert_array_equal(tid.numpy().view(np.uint64)[q, :c], exp[0][q, :c])
            np.testing.assert_array_equal(dst.numpy()[q, :c], exp[2][q, :c])

        # IVFFlat over the stripes: rank 0 trains, everybody assigns to rank 0's centres (stand-in index objects:
        # k-means and assignment are CUDA stages covered by -m gpu; this checks who calls what with which centres)
        class FakeIndex:
            dim = 32
            device = 0
            def __init__(self): self.trained = False; self.given = None
            def build_ivf(self, lists, seed=42):
                self.trained = True
                self._c = np.random.default_rng(seed + 1000 * rank).random((lists, self.dim), dtype=np.float32)
            def ivf_centers(self): return self._c
            def build_ivf_from_centers(self, c): self.given = np.array(c, copy=True)
        fi = FakeIndex()
        centres = sharded.build_ivf_sharded(fi, rank, world, lists=7, seed=42, device=torch.device("cpu"))
        expect = np.random.default_rng(42).random((7, 32), dtype=np.float32)       # what rank 0 "trained"
        np.testing.assert_array_equal(centres, expect)
        assert fi.trained == (rank == 0)
        if rank != 0:
            np.testing.assert_array_equal(fi.given, expect)
        # peer-exchange handshake (rk_exchange.cu's rings are CUDA; this checks who learns whose handle, and that one rank
        # failing to map a ring sends EVERY rank back to the all-gather)
        class FakeExchange:
            depth = 4
            def __init__(self, fail_connect=False): self.seen, self.closed, self.fail_connect = {}, False, fail_connect
            def handle(self): return bytes([rank]) * 64
            def connect_ipc(self, g, h):
                if self.fail_connect:
                    raise RuntimeError("cannot map")
                self.seen[g] = h
            def close(self): self.closed = True
        sh2 = sharded.ShardedSearcher(None, rank, world, device=torch.device("cpu"), local_search=local_search, merge=merge)
        fx = FakeExchange()
        assert sh2.enable_peer_exchange(10, 15, make_exchange=lambda: fx) is True
        assert sh2._ex is fx and fx.seen == {g: bytes([g]) * 64 for g in range(world) if g != rank}
        assert "peer stores" in sh2.exchange_kind
        sh3 = sharded.ShardedSearcher(None, rank, world, device=torch.device("cpu"), local_search=local_search, merge=merge)
        fx3 = FakeExchange(fail_connect=(rank == 1))
        assert sh3.enable_peer_exchange(10, 15, make_exchange=lambda: fx3) is False
        assert sh3._ex is None and fx3.closed and "nccl" in sh3.exchange_kind
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_sharded_allgather_merge_world2(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))


def test_owner_of_doc_stripes_evenly():
    from vectorsearch_rbac_b200 import sharded
    docs = np.arange(1, 1001)
    for g in (1, 2, 4, 8):
        own = sharded.owner_of_doc(docs, g)
        counts = np.bincount(own, minlength=g)
        assert counts.min() >= 1000 // g and counts.max() <= 1000 // g + 1
