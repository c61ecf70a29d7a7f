"""The boundary is plain C: a C99 program compiled by gcc against include/rbacknn.h and linked with librbacknn.so
(no C++ or torch types anywhere in the header).  Without a GPU it must fail loudly; with one it runs a scan through
begin / rescan / gettuple / endscan (pgvector/src/ivfflat.h:335-338) whose rows are checked against the oracle."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "vectorsearch-rbac_b200")
SRC = os.path.join(ROOT, "tests", "c_client", "client.c")


def _build(tmp_path):
    exe = str(tmp_path / "rk_client")
    cmd = ["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), SRC, "-o", exe,
           "-L", PKG, "-lrbacknn", f"-Wl,-rpath,{PKG}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_c99_client_compiles_and_fails_loudly_without_a_device(tmp_path):
    import torch
    exe = _build(tmp_path)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: the scan mode is tested instead")
    r = subprocess.run([exe, "nodevice"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "no CPU path" in r.stdout


@pytest.mark.gpu
def test_c99_client_scan_matches_oracle(tmp_path):
    from oracle import oracle as orc
    exe = _build(tmp_path)
    r = subprocess.run([exe, "scan"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = r.stdout.strip().splitlines()
    X = np.array([[0, 0, 0, 0], [1, 0, 0, 0], [0, 2, 0, 0], [0, 0, 3, 0], [4, 0, 0, 0], [0, 0, 0, 5]], dtype=np.float32)
    tid = np.arange(101, 107, dtype=np.uint64)
    # user 7 -> role 1 -> documents 1, 2 -> rows 0..3
    et, _, ed, _, ec = orc.search_batch(X, tid, [np.arange(4, dtype=np.uint32)], [[0]], np.zeros((1, 4), np.float32), orc.METRIC_L2, 8)
    rows = [ln.split() for ln in lines if ln[0].isdigit()]
    assert [int(a[0]) for a in rows] == et[0][:ec[0]].tolist() == [101, 102, 103, 104]
    assert [float(a[2]) for a in rows] == ed[0][:ec[0]].tolist()
    assert [int(a[1]) for a in rows] == [1, 1, 2, 2]
    assert "err different vector dimensions 4 and 3" in lines
    assert "batch 4 101 102 103 104" in lines
