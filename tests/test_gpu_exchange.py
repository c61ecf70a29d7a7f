"""Peer exchange (rk_exchange.cu): G ranks emulated on one GPU — every rank owns a shard index, a stream and a receive
ring; the rings are connected by pointer (what cudaIpcOpenMemHandle yields between processes).  Every rank's merged
result must equal the oracle over the whole corpus on every step, across ring wrap-arounds (the acknowledgement path), with
the merge on a second stream, for both engines."""
import numpy as np
import pytest

from oracle import oracle as orc
from tests.util import assert_topk_matches, load_index, make_dataset, oracle_role_search

pytestmark = pytest.mark.gpu


def _setup(world, nq, k, depth, n=9000):
    import torch
    import vectorsearch_rbac_b200 as vr
    from vectorsearch_rbac_b200 import sharded
    ds = make_dataset("sift", n=n, dim=64, rows_per_doc=15, rbac="random", num_users=20, num_roles=8)
    dev = torch.device("cuda", 0)
    ranks = []
    for r in range(world):
        mine = sharded.local_row_mask(ds["doc"], r, world)
        local = dict(ds, X=ds["X"][mine], tid=ds["tid"][mine], doc=ds["doc"][mine])
        ix = load_index(local, role_parts=False, combo_parts=False)
        ix.finalize()
        st, side = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        ix.set_stream(st.cuda_stream)
        ex = vr.engine.Exchange(ix, nq, k, world, r, depth)
        outs = [dict(tid=torch.empty((nq, k), dtype=torch.int64, device=dev), doc=torch.empty((nq, k), dtype=torch.int32, device=dev),
                     dist=torch.empty((nq, k), dtype=torch.float64, device=dev), cnt=torch.empty(nq, dtype=torch.int32, device=dev))
                for _ in range(2)]
        ranks.append(dict(ix=ix, ex=ex, st=st, side=side, outs=outs))
    for a in ranks:
        for g, b in enumerate(ranks):
            a["ex"].connect_ptr(g, b["ex"].base(), 0)
    return ds, dev, ranks


@pytest.mark.parametrize("world,engine", [(2, "scan"), (3, "scan"), (3, "auto")])
def test_peer_exchange_matches_oracle_over_ring_wraps(world, engine):
    import torch
    import vectorsearch_rbac_b200 as vr
    nq, k, depth = 40, 30, 2
    ds, dev, ranks = _setup(world, nq, k, depth)
    rng = np.random.default_rng(11)
    all_users = np.unique(ds["user_roles"][:, 0])
    eng = vr.ENGINE_SCAN if engine == "scan" else vr.ENGINE_AUTO
    for step in range(2 * depth + 1):
        nq_s = nq if step % 2 == 0 else nq - 7                      # the slot layout follows the call's nq
        Q = ds["X"][rng.integers(0, len(ds["X"]), nq_s)].copy()
        users = all_users[rng.integers(0, len(all_users), nq_s)].astype(np.uint32)
        dQ = torch.from_numpy(Q).to(dev)
        torch.cuda.synchronize()
        opts = vr.engine.make_opts(k, filter=vr.FILTER_BITMASK, engine=eng)
        for a in ranks:                                              # every rank: local search + push to the peers
            a["ex"].push(dQ.data_ptr(), nq_s, users, opts)
        # (one GPU stands in for all ranks here: a merge spinning on a flag could hold the SMs the peer's scan still needs,
        # which cannot happen when every rank has its own GPU — so the pushes are drained before the merges are issued)
        torch.cuda.synchronize()
        for a in ranks:                                              # every rank: wait for the peers' flags + merge, side stream
            o = a["outs"][step & 1]
            ev = torch.cuda.Event()
            ev.record(a["st"])
            a["side"].wait_event(ev)
            a["ex"].pull(nq_s, k, o["tid"].data_ptr(), o["doc"].data_ptr(), o["dist"].data_ptr(), o["cnt"].data_ptr(),
                         stream=a["side"].cuda_stream)
        exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, k, mode="bitmask")
        for a in ranks:
            a["ex"].status(a["side"].cuda_stream)
            o = a["outs"][step & 1]
            got = (o["tid"][:nq_s].cpu().numpy().view(np.uint64), o["doc"][:nq_s].cpu().numpy(), o["dist"][:nq_s].cpu().numpy(),
                   o["cnt"][:nq_s].cpu().numpy())
            assert_topk_matches(got, exp, exact=True)
    for a in ranks:
        a["ex"].close()
        a["ix"].close()


def test_peer_exchange_argument_checks():
    import vectorsearch_rbac_b200 as vr
    ds, dev, ranks = _setup(2, 8, 10, 2, n=1500)
    a = ranks[0]
    with pytest.raises(vr.RkError):                                  # nothing pushed yet
        o = a["outs"][0]
        a["ex"].pull(8, 10, o["tid"].data_ptr(), o["doc"].data_ptr(), o["dist"].data_ptr(), o["cnt"].data_ptr())
    with pytest.raises(vr.RkError):                                  # beyond the ring's capacity
        a["ex"].push(0, 9, np.zeros(9, np.uint32), vr.engine.make_opts(10, filter=vr.FILTER_BITMASK))
    with pytest.raises(vr.RkError):
        vr.engine.Exchange(a["ix"], 8, 10, 2, 5, 2)                  # rank outside the world
    for r in ranks:
        r["ex"].close()
        r["ix"].close()


def _two_gpu_worker(rank, world, port, out_dir):
    import os
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import vectorsearch_rbac_b200 as vr
        from vectorsearch_rbac_b200 import sharded
        torch.cuda.set_device(rank)
        dev = torch.device("cuda", rank)
        ds = make_dataset("sift", n=20000, dim=64, rows_per_doc=15, rbac="random", num_users=20, num_roles=8)
        mine = sharded.local_row_mask(ds["doc"], rank, world)
        local = dict(ds, X=ds["X"][mine], tid=ds["tid"][mine], doc=ds["doc"][mine])
        ix = load_index(local, role_parts=False, combo_parts=False, device=rank)
        ix.finalize()
        torch.cuda.set_stream(torch.cuda.Stream(dev, priority=-1))
        sharded.use_torch_stream(ix, dev)
        sh = sharded.ShardedSearcher(ix, rank, world, device=dev)
        nq, k = 64, 50
        assert sh.enable_peer_exchange(nq, k, depth=3), "CUDA IPC mapping of the peers' rings failed"
        rng = np.random.default_rng(3)
        all_users = np.unique(ds["user_roles"][:, 0])
        opts = sharded.make_opts(k, filter=vr.FILTER_BITMASK, engine=vr.ENGINE_SCAN)
        steps = []
        for step in range(8):                                        # pipelined: the merge one step behind / on a side stream
            sh.flush()
            sh.peer_overlap = "deferred" if step < 4 else "side"
            Q = ds["X"][rng.integers(0, len(ds["X"]), nq)].copy()
            users = all_users[rng.integers(0, len(all_users), nq)].astype(np.uint32)
            dQ = torch.from_numpy(Q).to(dev)
            out = sh.search_device_pipelined(dQ, users, opts) if step % 4 != 3 else sh.search_device(dQ, users, opts)
            sh.flush()
            torch.cuda.synchronize(dev)
            steps.append((Q, users, tuple(t.cpu().numpy() for t in out)))
        sh._ex.status()
        for Q, users, (tid, doc, dst, cnt) in steps:
            exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, k, mode="bitmask")
            assert_topk_matches((tid.view(np.uint64), doc, dst, cnt), exp, exact=True)
        # several steps in flight without a host sync in between (the ring and the acknowledgements do the flow control)
        Q = ds["X"][rng.integers(0, len(ds["X"]), nq)].copy()
        users = all_users[rng.integers(0, len(all_users), nq)].astype(np.uint32)
        dQ = torch.from_numpy(Q).to(dev)
        sh.flush()
        sh.peer_overlap = "deferred"
        for _ in range(12):
            out = sh.search_device_pipelined(dQ, users, opts)
        sh.flush()
        torch.cuda.synchronize(dev)
        sh._ex.status()
        exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, k, mode="bitmask")
        tid, doc, dst, cnt = (t.cpu().numpy() for t in out)
        assert_topk_matches((tid.view(np.uint64), doc, dst, cnt), exp, exact=True)
        dist.barrier()
        sh._ex.close()
        ix.close()
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_peer_exchange_two_processes_two_gpus(tmp_path):
    """The real thing: one process per GPU, rings mapped through CUDA IPC, merges pipelined under the next scan."""
    import socket
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mp.spawn(_two_gpu_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(2))


@pytest.mark.parametrize("n_shards", [2, 3])
def test_multi_index_one_process_matches_oracle(n_shards):
    """rk_multi_*: several shards behind one handle in one process (here all on cuda:0; real boxes list distinct devices).  Role,
    combination, bitmask and dynamic filters on both engines against the oracle over the whole corpus."""
    import torch
    import vectorsearch_rbac_b200 as vr
    from vectorsearch_rbac_b200.dynamic_layout import apply_layout, split_roles
    ds = make_dataset("sift", n=9000, dim=64, rows_per_doc=15, rbac="random", num_users=20, num_roles=8)
    devs = [0] * n_shards if torch.cuda.device_count() < n_shards else list(range(n_shards))
    mi = vr.MultiIndex(64, devs, metric=vr.METRIC_L2, k_max=64, nq_max=64)
    half = len(ds["X"]) // 2
    mi.load_rows(ds["X"][:half], ds["tid"][:half], ds["doc"][:half])
    mi.load_rows(ds["X"][half:], ds["tid"][half:], ds["doc"][half:])
    mi.set_doc_roles(ds["doc_roles"][:, 0], ds["doc_roles"][:, 1])
    mi.set_user_roles(ds["user_roles"][:, 0], ds["user_roles"][:, 1])
    mi.build_role_partitions()
    mi.build_combination_partitions()
    parts, rp, _ = split_roles(ds["doc_roles"], ds["user_roles"], alpha=2.0)
    apply_layout(mi, parts, rp)
    mi.finalize()
    rng = np.random.default_rng(21)
    all_users = np.unique(ds["user_roles"][:, 0])
    for step in range(6):                                            # more steps than the ring is deep
        nq = 48 if step % 2 == 0 else 33
        Q = ds["X"][rng.integers(0, len(ds["X"]), nq)].copy()
        users = all_users[rng.integers(0, len(all_users), nq)].astype(np.uint32)
        flt, layout, mode = [(vr.FILTER_PARTITION, vr.LAYOUT_ROLE, "role"), (vr.FILTER_PARTITION, vr.LAYOUT_COMBINATION, "combination"),
                             (vr.FILTER_BITMASK, 0, "bitmask"), (vr.FILTER_PARTITION, vr.LAYOUT_DYNAMIC, "bitmask")][step % 4]
        eng = vr.ENGINE_SCAN if step < 4 else vr.ENGINE_AUTO
        got = mi.search_batch(Q, users, 40, filter=flt, layout=layout, engine=eng)
        exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, 40, mode=mode)
        assert_topk_matches(got, exp, exact=True)
    with pytest.raises(vr.RkError):
        mi.search_batch(np.zeros((65, 64), np.float32), np.ones(65, np.uint32), 40, filter=vr.FILTER_BITMASK)     # beyond nq_max
    mi.close()
