"""synth.py — seeded synthetic inputs of the shapes the reference benchmarks use.

The reference seeds nothing (random.randint / random.shuffle / np.random.choice without a seed), so
the distributions are restated here with explicit seeds:
  corpus ids ....... block_id = i + 1, document_id = i // rows_per_doc + 1
                     (services/read_dataset_function.py:338-339)
  tree RBAC ........ services/rbac_generator/tree_based_rbac_data_generator.py:21-217
  random RBAC ...... services/rbac_generator/random_rbac_data_generator.py:26-82
  queries .......... services/read_dataset_function.py:737-749 (uniform corpus row + uniform user)

Vectors come from a counter-based hash, so any row range can be generated independently (each GPU
rank generates only its own documents) and identically on CPU and GPU for the integer-valued
SIFT shape.
"""
from __future__ import annotations

import random
from typing import List, Sequence, Tuple

import numpy as np
import torch

_M64 = (1 << 64) - 1


def _s64(c: int) -> int:
    c &= _M64
    return c - (1 << 64) if c >= (1 << 63) else c


_C1, _C2, _C3 = _s64(0x9E3779B97F4A7C15), _s64(0xBF58476D1CE4E5B9), _s64(0x94D049BB133111EB)


def _lsr(x: torch.Tensor, s: int) -> torch.Tensor:
    return (x >> s) & ((1 << (64 - s)) - 1)


def _mix64(x: torch.Tensor) -> torch.Tensor:
    """splitmix64 finaliser on int64 tensors (wrap-around arithmetic)."""
    x = x * _C1
    x = (x ^ _lsr(x, 30)) * _C2
    x = (x ^ _lsr(x, 27)) * _C3
    return x ^ _lsr(x, 31)


def _grid(rows: torch.Tensor, dim: int, salt: int) -> torch.Tensor:
    cols = torch.arange(dim, dtype=torch.int64, device=rows.device)
    return _mix64((rows[:, None] * dim + cols[None, :]) ^ _s64(salt))


def sift_like(rows: torch.Tensor, dim: int = 128, seed: int = 1234, n_clusters: int = 4096) -> torch.Tensor:
    """SIFT-shaped vectors: non-negative integers 0..255 stored as fp32, skewed towards small
    values, clustered.  Every partial sum of squared differences is < 2**24, so fp32 distances are
    exact in any summation order (SURVEY.md §7 hard part B)."""
    rows = rows.to(torch.int64)
    cl = _lsr(_mix64(rows ^ _s64(seed * 7919 + 1)), 11) % n_clusters
    cbits = _grid(cl, dim, seed * 31 + 2)
    base = ((cbits & 255) * (_lsr(cbits, 8) & 255)) >> 8              # 0..254, mean ~63
    nbits = _grid(rows, dim, seed * 131 + 3)
    noise = (nbits & 31) + (_lsr(nbits, 5) & 31) - 31                # triangular, -31..31
    return (base + noise).clamp_(0, 255).to(torch.float32)


def _normal(bits: torch.Tensor) -> torch.Tensor:
    u1 = ((_lsr(bits, 40) & 0xFFFFFF).to(torch.float32) + 1.0) / 16777217.0
    u2 = (_lsr(bits, 16) & 0xFFFFFF).to(torch.float32) / 16777216.0
    return torch.sqrt(-2.0 * torch.log(u1)) * torch.cos(6.283185307179586 * u2)


def cohere_like(rows: torch.Tensor, dim: int = 768, seed: int = 1234, n_clusters: int = 1024,
                sigma: float = 0.3) -> torch.Tensor:
    """Cohere-Wikipedia-shaped vectors: un-normalised fp32, mixture of Gaussian clusters with
    centres ~ N(0, I/dim) and within-cluster sigma/sqrt(dim); norms ~ 1."""
    rows = rows.to(torch.int64)
    cl = _lsr(_mix64(rows ^ _s64(seed * 7919 + 11)), 11) % n_clusters
    centre = _normal(_grid(cl, dim, seed * 31 + 12))
    noise = _normal(_grid(rows, dim, seed * 131 + 13))
    return ((centre + sigma * noise) * (1.0 / dim ** 0.5)).to(torch.float32)


def cohere_two_level(rows: torch.Tensor, dim: int = 768, seed: int = 1234, n_macro: int = 64, n_micro: int = 1024,
                     sigma_macro: float = 0.5, sigma_micro: float = 0.5, topic_scale: float = 0.5) -> torch.Tensor:
    """A harder Cohere-shaped corpus for the IVFFlat workload: topics (macro-clusters, centres ~ topic_scale N(0, I/dim)) made
    of sub-topics (micro-clusters at sigma_macro / sqrt(dim) from the topic centre) made of rows (sigma_micro / sqrt(dim) from
    the sub-topic centre).  A row's nearest neighbours then spread over many sub-topics of its topic, and k-means with lists
    cut through them — recall depends on nprobe the way it does on real embeddings, which a mixture of tight, mutually
    equidistant Gaussian blobs (cohere_like: recall 0.9999 at nprobe 16 of 4096) does not show.  The defaults (64 topics x
    1024 sub-topics, all three scales 0.5) were chosen on the C4 shape (10 M rows, 4096 lists, tree RBAC): recall@100
    0.16 / 0.52 / 0.89 / 0.9995 / 1.0 at nprobe 1 / 4 / 16 / 64 / 256."""
    import os
    if os.environ.get("RK_SYNTH"):          # tuning aid: "topic_scale,sigma_macro,sigma_micro,n_macro,n_micro"
        t = os.environ["RK_SYNTH"].split(",")
        topic_scale, sigma_macro, sigma_micro, n_macro, n_micro = float(t[0]), float(t[1]), float(t[2]), int(t[3]), int(t[4])
    rows = rows.to(torch.int64)
    h = _lsr(_mix64(rows ^ _s64(seed * 7919 + 21)), 11)
    macro = h % n_macro
    micro = macro * n_micro + (_lsr(h, 24) % n_micro)
    c1 = _normal(_grid(macro, dim, seed * 31 + 22))
    c2 = _normal(_grid(micro, dim, seed * 37 + 23))
    noise = _normal(_grid(rows, dim, seed * 131 + 24))
    return ((topic_scale * c1 + sigma_macro * c2 + sigma_micro * noise) * (1.0 / dim ** 0.5)).to(torch.float32)


def corpus(kind: str, rows: torch.Tensor, dim: int, seed: int = 1234, chunk: int = 1 << 18) -> torch.Tensor:
    fn = {"sift": sift_like, "cohere": cohere_like, "cohere2": cohere_two_level}[kind]
    out = torch.empty((rows.numel(), dim), dtype=torch.float32, device=rows.device)
    for a in range(0, rows.numel(), chunk):
        out[a:a + chunk] = fn(rows[a:a + chunk], dim, seed)
    return out


def row_ids(n_rows: int, rows_per_doc: int = 100) -> Tuple[np.ndarray, np.ndarray]:
    """(block_id, document_id) of rows 0..n-1 as the SIFT ingest assigns them."""
    i = np.arange(n_rows, dtype=np.int64)
    return (i + 1).astype(np.uint64), (i // rows_per_doc + 1).astype(np.uint32)


# ----------------------------------------------------------------------------------------------
# RBAC generators -> (user_roles[U,2] = (user, role), doc_roles[P,2] = (document, role))

def tree_rbac(num_users: int, num_roles: int, doc_ids: Sequence[int], h: int = 4, b0: int = 3, b1: int = 4,
              seed: int = 42) -> Tuple[np.ndarray, np.ndarray]:
    """Tree-based RBAC: roles form a DFS-built tree of height <= h with b0..b1 children per node;
    the documents are shuffled and split evenly over the tree's nodes; a role sees its own set
    plus those of all its ancestors; users are split evenly over the roles in DFS order, one role
    each."""
    rng = random.Random(seed)
    pending = list(range(1, num_roles + 1))
    children = {0: []}
    order: List[int] = []            # DFS pre-order of attached roles
    parent = {}

    def grow(node: int, level: int) -> None:
        if level >= h or not pending:
            return
        for _ in range(min(rng.randint(b0, b1), len(pending))):
            if not pending:
                break
            c = pending.pop(0)
            children[node].append(c)
            children[c] = []
            parent[c] = node
            order.append(c)
            grow(c, level + 1)

    grow(0, 0)
    docs = list(doc_ids)
    rng.shuffle(docs)
    n_sets = len(order)
    size = len(docs) // n_sets
    own = {}
    for i, r in enumerate(order):
        own[r] = docs[i * size:] if i == n_sets - 1 else docs[i * size:(i + 1) * size]
    doc_roles = []
    for r in order:
        a = r
        while a != 0:                 # own set + every ancestor's
            doc_roles.extend((d, r) for d in own[a])
            a = parent[a]
    users = np.arange(1, num_users + 1)
    user_roles = []
    for r, chunk in zip(order, np.array_split(users, len(order))):
        user_roles.extend((int(u), r) for u in chunk)
    return np.asarray(user_roles, dtype=np.int64).reshape(-1, 2), np.asarray(doc_roles, dtype=np.int64).reshape(-1, 2)


def random_rbac(num_users: int, num_roles: int, doc_ids: Sequence[int], m_roles: int, m_perms: int,
                seed: int = 42) -> Tuple[np.ndarray, np.ndarray]:
    """Random RBAC: each user gets 1..m_roles distinct roles; each role gets m_perms/2..m_perms
    distinct documents (role permission sets are pairwise different); documents left unassigned
    go to random roles that still have room."""
    rng = random.Random(seed)
    docs = list(doc_ids)
    roles = list(range(1, num_roles + 1))
    user_roles = []
    for u in range(1, num_users + 1):
        for r in rng.sample(roles, rng.randint(1, m_roles)):
            user_roles.append((u, r))
    perms = {}
    seen = set()
    left = set(docs)
    for r in roles:
        while True:
            pick = tuple(sorted(rng.sample(docs, rng.randint(m_perms // 2, m_perms))))
            if pick not in seen:
                seen.add(pick)
                perms[r] = set(pick)
                left.difference_update(pick)
                break
    for d in sorted(left):
        while True:
            r = rng.choice(roles)
            if len(perms[r]) < m_perms:
                perms[r].add(d)
                break
    doc_roles = [(d, r) for r in roles for d in sorted(perms[r])]
    return np.asarray(user_roles, dtype=np.int64).reshape(-1, 2), np.asarray(doc_roles, dtype=np.int64).reshape(-1, 2)


def queries(n_rows: int, user_ids: Sequence[int], nq: int, seed: int = 7) -> Tuple[np.ndarray, np.ndarray]:
    """nq (corpus row, user) pairs drawn uniformly with replacement (zipf_param = 0 path)."""
    rng = np.random.default_rng(seed)
    rows = rng.integers(0, n_rows, size=nq)
    users = np.asarray(user_ids)[rng.integers(0, len(user_ids), size=nq)]
    return rows.astype(np.int64), users.astype(np.uint32)
