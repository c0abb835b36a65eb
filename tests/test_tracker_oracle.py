"""CPU: the tracker-geometry oracle (oracle/tracker_oracle.py) against the outputs of the unmodified reference functions
(tests/golden/tracker.npz) and its LM solver's own sanity (the solver has no reference counterpart here: DeepLM is absent)."""
import os

import cv2
import numpy as np
import pytest

from oracle import tracker_oracle as tro

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "tracker.npz"))


def tri_inputs():
    d = tro.tracker_inputs(0)
    Ks, Ts = [], []
    for c in (1, 2):
        Ks.append(np.array([[d["cams_gt"][c, 6], 0, d["cams_gt"][c, 7]], [0, d["cams_gt"][c, 6], d["cams_gt"][c, 8]], [0, 0, 1]]))
        Tcw = np.eye(4); Tcw[:3, :3] = cv2.Rodrigues(d["cams_gt"][c, :3])[0]; Tcw[:3, 3] = d["cams_gt"][c, 3:6]
        Ts.append(np.linalg.inv(Tcw))
    return d, Ks, Ts


@pytest.mark.parametrize("seed", [0, 1])
def test_rotate_and_residual(seed):
    d = tro.tracker_inputs(seed)
    pts, cams = d["points0"][d["pt_idx"]], d["cams0"][d["cam_idx"], :6]
    np.testing.assert_allclose(tro.angle_axis_rotate_point(cams[:, :3], pts), G[f"rot_{seed}"], rtol=0, atol=1e-15)
    np.testing.assert_allclose(tro.snavely_residual(pts, cams, d["features"]), G[f"res_{seed}"], rtol=0, atol=1e-10)
    np.testing.assert_array_equal(G[f"res_{seed}"], G[f"res3d_{seed}"])


def test_project_and_triangulate():
    d, Ks, Ts = tri_inputs()
    K = np.array([[600.0, 0, 256], [0, 600, 256], [0, 0, 1]])
    RT = np.concatenate([cv2.Rodrigues(d["cams_gt"][1, :3])[0], d["cams_gt"][1, 3:6, None]], 1)
    xy, depth = tro.project(d["points_gt"], K, RT)
    np.testing.assert_allclose(xy, G["proj_xy"], rtol=0, atol=1e-10)
    np.testing.assert_allclose(depth, G["proj_depth"], rtol=0, atol=1e-14)
    P1, P2 = Ks[0] @ np.linalg.inv(Ts[0])[:3], Ks[1] @ np.linalg.inv(Ts[1])[:3]
    X = tro.triangulate(P1, P2, G["tri_x1"], G["tri_x2"])
    np.testing.assert_allclose(X, G["tri_X"], rtol=0, atol=1e-9)
    assert np.abs(X - d["points_gt"]).max() < 1e-2             # 0.3 px noise: millimetres


def test_lm_reduces_cost_and_recovers_structure():
    d = tro.tracker_inputs(2, noise_px=0.0, perturb=0.005)
    pts, cams, log = tro.ba_lm(d["points0"], d["cams0"], d["features"], d["pt_idx"], d["cam_idx"], num_iterations=8, num_success=8)
    costs = [c for c, _, _ in log]
    assert all(ok for _, _, ok in log[:3]) and log[-1][1] < 1e-6 * costs[0]
    r = tro.snavely_residual(pts[d["pt_idx"]], cams[d["cam_idx"]], d["features"])
    assert np.abs(r).max() < 1e-3                               # noise-free: the reprojection error vanishes (gauge aside)


def test_jacobians_match_finite_differences():
    d = tro.tracker_inputs(3)
    pts, cams = d["points0"][d["pt_idx"]], d["cams0"][d["cam_idx"], :6]
    r, Jc, Jp = tro.jacobians(pts, cams, d["features"])
    h = 1e-6
    for j in range(6):
        cp, cm = cams.copy(), cams.copy(); cp[:, j] += h; cm[:, j] -= h
        fd = (tro.snavely_residual(pts, cp, d["features"]) - tro.snavely_residual(pts, cm, d["features"])) / (2 * h)
        np.testing.assert_allclose(Jc[:, :, j], fd, rtol=1e-5, atol=1e-4)
    for j in range(3):
        pp, pm = pts.copy(), pts.copy(); pp[:, j] += h; pm[:, j] -= h
        fd = (tro.snavely_residual(pp, cams, d["features"]) - tro.snavely_residual(pm, cams, d["features"])) / (2 * h)
        np.testing.assert_allclose(Jp[:, :, j], fd, rtol=1e-5, atol=1e-4)
