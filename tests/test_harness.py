"""The reference harness's entry points (search_func signature, recall, ground-truth cache format) over the engine:
controller/baseline/prefilter/prefilter_role.py:22-26, 174-189; basic_benchmark/common_function.py:831-849, 1154-1160,
1366-1395."""
import json

import numpy as np
import pytest

from oracle import oracle as orc
from tests.util import load_index, make_dataset, oracle_role_search


def test_parse_vector_and_recall_and_cache(tmp_path):
    from vectorsearch_rbac_b200 import harness
    np.testing.assert_array_equal(harness.parse_vector("[1, 2.5, -3]"), np.asarray([1, 2.5, -3], dtype=np.float32))
    with pytest.raises(ValueError):
        harness.parse_vector("[1, NaN]")                      # vector.c:102-113
    with pytest.raises(ValueError):
        harness.parse_vector([[1, 2]])
    gt = [(7, 1, b"", 0.5), (9, 1, b"", 0.7), (3, 2, b"", 0.9), (4, 2, b"", 1.0)]
    pred = [(7, 1, b"", 0.5), (3, 2, b"", 0.9), (11, 5, b"", 0.95)]
    assert harness.compute_recall(harness.result_keys(gt), harness.result_keys(pred)) == 0.5
    queries = [{"user_id": 3, "query_vector": "[1,2]", "topk": 4}, {"user_id": 4, "query_vector": [0.5, 1.5]}]
    f = tmp_path / "ground_truth_cache.json"
    harness.save_ground_truth_cache(queries, [gt, pred], str(f))
    raw = json.load(open(f))
    assert raw[0]["query"] == {"user_id": 3, "query_vector": "[1,2]", "topk": 4} and raw[1]["query"]["topk"] == 5
    assert raw[0]["ground_truth"][0] == [7, 1, "", 0.5]
    q2, r2 = harness.load_ground_truth_cache(str(f))
    assert [int(r[0]) for r in r2[1]] == [7, 3, 11] and q2[1]["user_id"] == 4


def test_pointer_cache_format(tmp_path):
    """compute_ground_truth.py:35-73 (writer / reader) and the C++ reader's view of the file (benchmark_utils.cpp:26-76)."""
    from vectorsearch_rbac_b200 import harness
    res = [[(7, 1, b"x", 0.5), (9, 1, b"", 0.7), (), ("bad", 2, b"", 0.1), (3,)], None, [(4, 2, b"", 1.0)]]
    assert harness.to_pointer_results(res) == [[[7, 1], [9, 1]], [], [[4, 2]]]
    f = tmp_path / "sub" / "ground_truth_cache.json"
    harness.save_pointer_cache(res, str(f))
    text = open(f).read()
    assert text.startswith("[\n  [\n    [\n      7,")                       # json.dump(..., indent=2)
    assert harness.load_pointer_cache(str(f)) == [[[7, 1], [9, 1]], [], [[4, 2]]]
    assert harness.load_pointer_cache(str(tmp_path / "missing.json")) is None
    (tmp_path / "broken.json").write_text("{not json")
    assert harness.load_pointer_cache(str(tmp_path / "broken.json")) is None
    (tmp_path / "dict.json").write_text("{}")
    assert harness.load_pointer_cache(str(tmp_path / "dict.json")) is None
    ptr = harness.load_pointer_cache(str(f))
    qs = [{"user_id": 1, "topk": 2}, {"user_id": 2, "topk": 0}, {"user_id": 3, "topk": 1}]
    assert harness.pointer_cache_as_flat_pairs(ptr, qs) == [(1, 7), (1, 9), (2, 4)]          # (document_id, block_id)
    assert harness.pointer_cache_as_flat_pairs(ptr, qs[:2]) is None                           # length mismatch
    assert harness.pointer_cache_as_flat_pairs(ptr, [qs[0], {"topk": 1}, qs[2]]) is None      # a query short of topk pairs


@pytest.mark.gpu
def test_search_funcs_match_oracle():
    import vectorsearch_rbac_b200 as vr
    from vectorsearch_rbac_b200 import harness
    ds = make_dataset("sift", n=6000, dim=32, rows_per_doc=20, rbac="random", num_users=30, num_roles=8)
    ix = load_index(ds)
    ix.finalize()
    s = harness.Searcher(ix)
    rng = np.random.default_rng(1)
    rows = rng.integers(0, len(ds["X"]), 6)
    users = np.unique(ds["user_roles"][:, 0])[rng.integers(0, 30, 6)].astype(np.uint32)
    Q = ds["X"][rows].copy()
    exp = oracle_role_search(ds, Q, users, orc.METRIC_L2, 7, mode="role")
    exp_c = oracle_role_search(ds, Q, users, orc.METRIC_L2, 7, mode="combination")
    queries = []
    for i in range(6):
        text = "[" + ",".join(repr(float(x)) for x in Q[i]) + "]"       # the harness passes pgvector text
        got, secs = s.search_documents_role_partition(user_id=int(users[i]), query_vector=text, topk=7, statistics_type="sql")
        assert secs > 0 and len(got) == int(exp[4][i])
        assert [r[0] for r in got] == exp[0][i, :len(got)].tolist()
        np.testing.assert_array_equal(np.asarray([r[3] for r in got]), exp[2][i, :len(got)])
        got_c, _ = s.search_documents_combination_partition(user_id=int(users[i]), query_vector=Q[i], topk=7)
        assert [r[0] for r in got_c] == exp_c[0][i, :len(got_c)].tolist()
        got_r, _ = s.search_documents_rls(user_id=int(users[i]), query_vector=Q[i], topk=7)
        assert harness.compute_recall(harness.result_keys(got_c), harness.result_keys(got_r)) == 1.0
        queries.append({"user_id": int(users[i]), "query_vector": text, "topk": 7})
    gt = s.ground_truth_func_batch(queries)
    for i in range(6):
        assert [r[0] for r in gt[i]] == exp_c[0][i, :len(gt[i])].tolist()
    ix.close()
