"""Host logic: the seeded generators follow the reference's procedures."""
import os

import numpy as np
import torch

from vectorsearch_rbac_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sift_like_is_integer_valued_and_position_addressable():
    a = synth.corpus("sift", torch.arange(0, 500), 128, seed=5)
    b = synth.corpus("sift", torch.arange(250, 500), 128, seed=5)
    assert torch.equal(a[250:], b)                       # any row range regenerates identically
    assert a.min() >= 0 and a.max() <= 255 and torch.equal(a, a.round())
    # exactness bound of SURVEY.md §7B: 128 * 255^2 < 2^24
    assert 128 * 255 ** 2 < 2 ** 24


def test_cohere_like_shape():
    c = synth.corpus("cohere", torch.arange(2000), 768, seed=5)
    n = c.norm(dim=1)
    assert 0.8 < n.mean() < 1.3 and torch.isfinite(c).all()


def test_row_ids_follow_sift_ingest():
    tid, doc = synth.row_ids(250, 100)
    assert tid[0] == 1 and tid[-1] == 250 and doc[0] == 1 and doc[99] == 1 and doc[100] == 2 and doc[-1] == 3


def test_tree_rbac_structure():
    docs = list(range(1, 10001))
    ur, dr = synth.tree_rbac(1000, 100, docs, h=4, b0=3, b1=4, seed=42)
    # every user has exactly one role, users split evenly (tree_based_rbac_data_generator.py:164-187)
    assert len(ur) == 1000 and len(np.unique(ur[:, 0])) == 1000
    per_role = np.bincount(ur[:, 1])[1:]
    assert per_role.max() - per_role[per_role > 0].min() <= 1
    # every document is assigned; a role at depth t sees t disjoint sets of ~100 documents (:88-142)
    assert set(dr[:, 0].tolist()) == set(docs)
    sizes = np.bincount(dr[:, 1])[1:]
    assert set((sizes // 100).tolist()) <= {1, 2, 3, 4, 5} and sizes.min() >= 100
    assert (ur == synth.tree_rbac(1000, 100, docs, seed=42)[0]).all()       # seeded
    # a child's documents include its parent's
    by_role = {r: set(dr[dr[:, 1] == r, 0].tolist()) for r in np.unique(dr[:, 1])}
    big = max(by_role, key=lambda r: len(by_role[r]))
    assert any(by_role[r] < by_role[big] for r in by_role if r != big)


def test_random_rbac_structure():
    docs = list(range(1, 501))
    ur, dr = synth.random_rbac(200, 20, docs, m_roles=3, m_perms=60, seed=1)
    n_roles = np.bincount(ur[:, 0])[1:]
    assert n_roles.min() >= 1 and n_roles.max() <= 3
    assert set(dr[:, 0].tolist()) == set(docs)            # all documents assigned at least once
    per_role = np.bincount(dr[:, 1])[1:]
    assert per_role.max() <= 60 and per_role.min() >= 30


def test_queries_seeded():
    r, u = synth.queries(1000, [5, 6, 7], 50, seed=7)
    r2, u2 = synth.queries(1000, [5, 6, 7], 50, seed=7)
    assert (r == r2).all() and (u == u2).all() and r.max() < 1000 and set(u.tolist()) <= {5, 6, 7}


def test_cpu_generator_of_the_reference_arm_is_bit_identical():
    """oracle/synthgen.c (what `bench.py --impl reference` fills its host matrix with, so that arm needs neither the GPU
    nor librbacknn.so) against the package's torch generator, bit for bit."""
    import torch
    from oracle import oracle as orc
    rows = np.concatenate([np.arange(0, 3000), np.random.default_rng(5).integers(0, 10_000_000, 3000)]).astype(np.int64)
    for dim in (128, 96):
        a = synth.sift_like(torch.as_tensor(rows), dim, 1234).numpy()
        b = orc.gen_sift(rows, dim, 1234, nthreads=3)
        assert a.dtype == b.dtype and np.array_equal(a, b)


def test_importing_the_generators_does_not_load_the_engine_library():
    """`bench.py --impl reference` imports vectorsearch_rbac_b200.synth only: the package must not pull librbacknn.so in."""
    import subprocess
    import sys
    code = ("import sys; sys.path.insert(0, %r); import vectorsearch_rbac_b200.synth; "
            "assert 'librbacknn' not in open('/proc/self/maps').read(); "
            "assert 'vectorsearch_rbac_b200._lib' not in sys.modules") % ROOT
    subprocess.run([sys.executable, "-c", code], check=True)


def test_two_level_cohere_shape():
    """The C4 corpus: finite, norms of order one, any row range regenerates identically, rows of one sub-topic closer to each
    other than rows of different topics."""
    rows = torch.arange(0, 4000)
    c = synth.corpus("cohere2", rows, 96, seed=5)
    assert torch.isfinite(c).all() and 0.5 < c.norm(dim=1).mean() < 1.5
    assert torch.equal(synth.corpus("cohere2", torch.arange(1000, 1500), 96, seed=5), c[1000:1500])
    h = synth._lsr(synth._mix64(rows ^ synth._s64(5 * 7919 + 21)), 11)
    macro = (h % 64).numpy()
    d = torch.cdist(c[:400], c[:400]).numpy()
    same = macro[:400, None] == macro[None, :400]
    off = ~np.eye(400, dtype=bool)
    assert d[same & off].mean() < d[~same].mean()
