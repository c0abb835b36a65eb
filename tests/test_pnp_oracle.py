"""CPU tests: the C restatement of cv2's EPnP-RANSAC against cv2's own outputs (golden + live)."""
import os

import numpy as np
import pytest

from onepose_b200 import synth
from oracle import pnp_oracle as po
from tests import helpers

mg = helpers.load_make_golden()


def test_cv_rng_stream_known_answer():
    # cv::RNG((uint64)-1) multiply-with-carry: first draws are fixed numbers
    s = po.cv_sample_stream(1000, 3)
    assert s.shape == (3, 5) and s.dtype == np.int32
    state = 0xFFFFFFFFFFFFFFFF
    exp = []
    for _ in range(5):
        state = ((state & 0xFFFFFFFF) * 4164903690 + (state >> 32)) & 0xFFFFFFFFFFFFFFFF
        exp.append((state & 0xFFFFFFFF) % 1000)
    assert len(set(exp)) == 5 and list(s[0]) == exp
    for row in po.cv_sample_stream(6, 50):
        assert len(set(row.tolist())) == 5 and row.max() < 6


@pytest.mark.parametrize("case", list(mg.PNP_CASES))
def test_ransac_restatement_vs_golden(case):
    seed, n, ratio, scale = mg.PNP_CASES[case]
    g = np.load(os.path.join(helpers.GOLDEN, case + ".npz"))
    pr = synth.make_pnp_problem(seed, n, ratio)
    pose, pose_h, inl = po.ransac_pnp_restated(pr["K"], pr["pts2d"], pr["pts3d"], scale=scale)
    assert np.array_equal(np.asarray(inl).ravel(), g["inliers"])        # bit-exact inlier set
    r_err, t_err = po.pose_error(pose, g["pose"])
    assert r_err < 1e-4 and t_err < 1e-6                                  # deg, cm
    # reference tolerance of north_star vs ground truth: the problem is well posed
    r_gt, t_gt = po.pose_error(pose, pr["pose_gt"])
    assert r_gt < 0.5 and t_gt < 0.5


@pytest.mark.parametrize("case", list(mg.PNP_CASES))
def test_minimal_sample_epnp_vs_golden(case):
    seed, n, ratio, scale = mg.PNP_CASES[case]
    g = np.load(os.path.join(helpers.GOLDEN, case + ".npz"))
    pr = synth.make_pnp_problem(seed, n, ratio)
    X32 = (pr["pts3d"] * scale).astype(np.float32); uv32 = pr["pts2d"].astype(np.float32)
    # 5-point EPnP has a 2-dim exact null space of M^T M whose basis is decided by rounding noise
    # (one-sided Jacobi), so a minority of minimal samples land on a slightly different Gauss-Newton
    # end point than cv2; the bulk must agree to rounding and none may be far off.
    diffs = []
    for s, rt in zip(g["samples"], g["sample_rt"]):
        rv, tv = po.hypothesis(X32, uv32, pr["K"], s)
        diffs.append(max(np.abs(rv - rt[:3]).max(), np.abs(tv - rt[3:]).max() / np.abs(rt[3:]).max()))
    diffs = np.array(diffs)
    assert np.median(diffs) < 1e-9 and (diffs < 1e-6).mean() >= 0.6 and diffs.max() < 2e-2, diffs


def test_sequential_select_replays_ransac():
    pr = synth.make_pnp_problem(21, 1500, 0.4, scale=1000.0)
    r = po.ransac_epnp(pr["pts3d"], pr["pts2d"], pr["K"], return_samples=True)
    X32 = pr["pts3d"].astype(np.float32); uv32 = pr["pts2d"].astype(np.float32)
    stream = po.cv_sample_stream(len(X32), 4096)
    assert np.array_equal(stream[:r["iters"]], r["samples"])
    counts = []
    for s in stream[:r["iters"] + 64]:
        rv, tv = po.hypothesis(X32, uv32, pr["K"], s)
        counts.append(int((po.reproj_err(X32, uv32, pr["K"], rv, tv) <= np.float32(25)).sum()))
    best, iters = po.sequential_select(len(X32), counts)
    assert iters == r["iters"]
    rv, tv = po.hypothesis(X32, uv32, pr["K"], stream[best])
    mask = po.reproj_err(X32, uv32, pr["K"], rv, tv) <= np.float32(25)
    assert np.array_equal(np.nonzero(mask)[0], r["inliers"])


def test_restatement_vs_live_cv2():
    cv2 = pytest.importorskip("cv2")
    for seed, (n, ratio) in enumerate([(400, 0.7), (900, 0.45), (50, 0.9)]):
        pr = synth.make_pnp_problem(100 + seed, n, ratio)
        a = po.ransac_pnp_cv2(pr["K"], pr["pts2d"], pr["pts3d"], scale=1000)
        b = po.ransac_pnp_restated(pr["K"], pr["pts2d"], pr["pts3d"], scale=1000)
        assert np.array_equal(np.asarray(a[2]).ravel(), np.asarray(b[2]).ravel())
        r_err, t_err = po.pose_error(b[0], a[0])
        assert r_err < 1e-4 and t_err < 1e-6   # acos near 1 resolves ~1e-6 deg


def test_wrapper_edge_cases_vs_golden():
    g = np.load(os.path.join(helpers.GOLDEN, "pnp_edge.npz"))
    for n in (0, 3):     # cv2.error -> identity pose, [] (eval_utils.py:40-42)
        assert np.array_equal(g[f"pose_{n}"], np.eye(4)[:3]) and len(g[f"inl_{n}"]) == 0
    for n in (4, 5):     # direct solve, all points returned as inliers
        assert list(g[f"inl_{n}"]) == list(range(n))
    for n in (6, 50):    # all outliers: cv2 reports failure, the wrapper returns []
        assert len(g[f"inl_out{n}"]) == 0
    assert g["p3p_poses"].shape == (8, 3, 4) and len(g["inl_lowinl"]) == 80


def test_edge_golden_vs_live_cv2_and_restatement():
    """The committed edge fixture against the image's cv2 (the library the reference calls) and the C restatement."""
    pytest.importorskip("cv2")
    g = np.load(os.path.join(helpers.GOLDEN, "pnp_edge.npz"))
    mg = helpers.load_make_golden()
    for seed in range(8):
        pr = synth.make_pnp_problem(900 + seed, 4, 1.0, px_noise=0.5 if seed % 2 else 0.0)
        pose, _, inl = po.ransac_pnp_cv2(pr["K"], pr["pts2d"], pr["pts3d"], scale=1000)
        assert list(np.asarray(inl).ravel()) == [0, 1, 2, 3]
        r_err, t_err = po.pose_error(pose, g["p3p_poses"][seed])
        assert r_err < 1e-4 and t_err < 1e-6
    for n in (6, 50):
        p2, p3 = mg.all_outlier_problem(n)
        assert len(po.ransac_pnp_cv2(synth.INTRINSIC, p2, p3, scale=1000)[2]) == 0
        assert len(po.ransac_pnp_restated(synth.INTRINSIC, p2, p3, scale=1000)[2]) == 0
    pr = synth.make_pnp_problem(4, 400, 0.2)
    ref = po.ransac_epnp(pr["pts3d"] * 1000.0, pr["pts2d"], pr["K"])
    assert ref["iters"] == 10000 and np.array_equal(ref["inliers"], g["inl_lowinl"])
