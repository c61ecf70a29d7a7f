import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench, vectorsearch_rbac_b200 as vr
wl = dict(bench.WORKLOADS["c2"])
dev = torch.device("cuda", 0)
tid, doc, ur, dr = bench.build_rbac(wl)
ix = vr.Index(wl["dim"], metric=vr.METRIC_IP, device=0, k_max=128)
for s in range(0, wl["rows"], 1 << 18):
    idx = np.arange(s, min(wl["rows"], s + (1 << 18)))
    xb = bench.gen_rows(wl, idx, dev)
    ix.load_rows_device(xb.data_ptr(), len(idx), tid[idx], doc[idx])
ix.set_doc_roles(dr[:, 0], dr[:, 1]); ix.set_user_roles(ur[:, 0], ur[:, 1])
ix.finalize()
Q, users = bench.make_queries(wl, 1000, dev)
Qn = Q.cpu().numpy()
ref = ix.search_batch(Qn, users, 100, filter=vr.FILTER_BITMASK, engine=vr.ENGINE_SCAN)
def run(sel, label, **env):
    for k_, v_ in env.items(): os.environ[k_] = v_
    got = ix.search_batch(Qn[sel], users[sel], 100, filter=vr.FILTER_BITMASK, engine=vr.ENGINE_TENSOR)
    for k_ in env: os.environ.pop(k_)
    qi = np.arange(1000)[sel]
    bad = [i for i in range(len(got[0])) if not np.array_equal(got[0][i], ref[0][qi[i]])]
    print(label, "plan", ix.stats()["tensor_plan"], "units", ix.stats()["segments"], "redo", ix.stats()["tensor_redo"], "nbad", len(bad), [int(qi[i]) for i in bad[:6]])
    for i in bad[:2]:
        g, r = got[0][i], ref[0][qi[i]]
        miss = sorted(set(r.tolist()) - set(g.tolist())); extra = sorted(set(g.tolist()) - set(r.tolist()))
        print("   q", int(qi[i]), "dups in result", len(g) - len(set(g.tolist())), "missed (tid,tile,row)", [(m, (m - 1) // 128, (m - 1) % 128) for m in miss[:6]], "extra", extra[:4],
              "ranks of missed", [int(np.flatnonzero(r == m)[0]) for m in miss[:6]])
w = slice(768, 896)
run(w, "base", RK_TENSOR_PLAN="dense")
for tpu in ("6", "20", "100", "400", "2000"):
    run(w, "tpu " + tpu, RK_TENSOR_PLAN="dense", RK_TENSOR_TPU=tpu)
run(slice(0, 1000), "full dense", RK_TENSOR_PLAN="dense")
run(slice(0, 1000), "full grouped", RK_TENSOR_PLAN="grouped")
run(slice(0, 1000), "full auto")
ix.close()
