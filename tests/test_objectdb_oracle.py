"""CPU tests of the object-DB builder / evaluator restatement (oracle/objectdb_oracle.py) and of the product's host logic
(index plan, Evaluator's numpy interface) against the fixtures generated from the unmodified reference
(tests/golden/db_builder.npz)."""
import os

import numpy as np
import pytest

from oracle import objectdb_oracle as orc
from tests import helpers

G = np.load(os.path.join(helpers.GOLDEN, "db_builder.npz"))
CASES = {"pad": (14, 4), "trunc": (9, 4), "wide": (12, 16)}


def test_oracle_mean_descriptors_vs_reference():
    inp = orc.db_builder_inputs()
    assert np.array_equal(orc.mean_descriptors(inp["desc"], inp["idxs"]), G["avg"])
    assert np.array_equal(orc.mean_descriptors(inp["scores"], inp["idxs"]), G["avg_scores"])


@pytest.mark.parametrize("tag", list(CASES))
def test_oracle_leaf_builder_vs_reference(tag):
    n_target, num_leaf = CASES[tag]
    inp = orc.db_builder_inputs()
    np.random.seed(5)
    d, s = orc.pad_features3d_random(G["avg"].T, G["avg_scores"], n_target)
    dl, sl = orc.build_features3d_leaves(inp["desc"].T, inp["scores"], inp["idxs"], n_target, num_leaf)
    assert np.array_equal(d, G[f"{tag}_avg"]) and np.array_equal(s, G[f"{tag}_avg_scores"])
    assert np.array_equal(dl, G[f"{tag}_leaves"]) and np.array_equal(sl, G[f"{tag}_leaf_scores"])


@pytest.mark.parametrize("tag", list(CASES))
def test_product_leaf_plan_reproduces_the_reference_choice(tag):
    """onepose_b200.objectdb.leaf_plan consumes np.random exactly as data_utils.py:183-192 does."""
    from onepose_b200.objectdb import leaf_plan
    n_target, num_leaf = CASES[tag]
    inp = orc.db_builder_inputs()
    np.random.seed(5)
    plan = leaf_plan(inp["idxs"], n_target, num_leaf)
    assert plan.shape == (n_target * num_leaf,) and plan.dtype == np.int64
    src = np.concatenate([inp["desc"], np.ones((1, 256), np.float32)])          # row -1 = dustbin
    assert np.array_equal(src[plan].T, G[f"{tag}_leaves"])
    # every real leaf of a point appears at most once, leaves stay inside their own point
    upper = np.cumsum(inp["idxs"]); lower = upper - inp["idxs"]
    for n in range(min(n_target, len(inp["idxs"]))):
        row = plan[n * num_leaf:(n + 1) * num_leaf]; real = row[row >= 0]
        assert len(set(real.tolist())) == len(real) and np.all((real >= lower[n]) & (real < upper[n]))
        assert len(real) == min(num_leaf, inp["idxs"][n])


def test_pose_errors_and_evaluator_vs_reference():
    from onepose_b200.evaluator import Evaluator, query_pose_error
    inp = orc.db_builder_inputs()
    errs = np.asarray([orc.query_pose_error(p, g) for p, g in zip(inp["poses_pred"], inp["poses_gt"])])
    np.testing.assert_allclose(errs, G["pose_errs"], rtol=0, atol=0)
    mine = np.asarray([query_pose_error(p, g) for p, g in zip(inp["poses_pred"], inp["poses_gt"])])
    np.testing.assert_allclose(mine, G["pose_errs"], rtol=0, atol=0)
    s = orc.evaluator_summary(inp["poses_pred"], inp["poses_gt"])
    assert [s["cmd1"], s["cmd3"], s["cmd5"]] == G["cmd"].tolist()
    ev = Evaluator()
    for p, g in zip(inp["poses_pred"], inp["poses_gt"]):
        ev.evaluate(np.vstack([p, [0, 0, 0, 1]]), g)                            # 4x4 prediction, 3x4 ground truth
    ev.evaluate(None, inp["poses_gt"][0])                                       # a failed frame counts as a miss
    r = ev.summarize(verbose=False)
    n = len(inp["poses_pred"])
    assert [r["cmd1"], r["cmd3"], r["cmd5"]] == [float(c) * n / (n + 1) for c in G["cmd"]]
    assert 0 < G["cmd"][0] < G["cmd"][2] <= 1                                   # the fixture separates the three thresholds
