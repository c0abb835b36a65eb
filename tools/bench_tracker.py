"""Row f4 measurement: tracker geometry kernels with the CPU restatement timed beside them.

    python tools/bench_tracker.py

* triangulation and Snavely residual at 1 M correspondences / observations (HBM-streaming float64 kernels): time, algorithmic GB/s;
  CPU: cv2.triangulatePoints / the numpy oracle on a bounded sample;
* one windowed bundle adjustment at the tracker's size (20 cameras, 2000 points, 8 observations each, 5 LM iterations): time on the
  device (no host round trip inside the solve); CPU: the oracle's dense LM on a 10x smaller window."""
import json
import os
import sys
import time

import cv2
import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from onepose_b200 import tracker  # noqa: E402
from oracle import tracker_oracle as tro  # noqa: E402


def timed(fn, iters=10):
    fn(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def main():
    res = {}
    rs = np.random.RandomState(0)
    n = 1 << 20
    d = tro.tracker_inputs(0, n_cams=20, n_pts=2000, obs_per_pt=8)
    # residual, gathered in the kernel
    pt_idx = torch.tensor(rs.randint(0, 2000, n), device="cuda"); cam_idx = torch.tensor(rs.randint(0, 20, n), device="cuda")
    feats = torch.tensor(np.c_[rs.uniform(0, 512, (n, 2)), np.full(n, 600.0), np.full(n, 256.0), np.full(n, 256.0)], device="cuda")
    pts, cams = torch.tensor(d["points_gt"], device="cuda"), torch.tensor(d["cams_gt"][:, :6], device="cuda").contiguous()
    ms = timed(lambda: tracker.SnavelyReprojectionErrorV2(pts, cams, feats, pt_idx, cam_idx))
    res["snavely_residual_1M"] = {"ms": round(ms, 4), "alg_GB_s": round(n * (40 + 16 + 16) / ms / 1e6, 1)}
    k = 1 << 16
    t0 = time.perf_counter()
    tro.snavely_residual(d["points_gt"][pt_idx[:k].cpu().numpy()], d["cams_gt"][cam_idx[:k].cpu().numpy(), :6], feats[:k].cpu().numpy())
    res["snavely_residual_1M"]["cpu_port_ms_scaled"] = round((time.perf_counter() - t0) * 1e3 * n / k, 1)
    # triangulation
    K = np.array([[600.0, 0, 256], [0, 600, 256], [0, 0, 1]])
    T1 = np.eye(4); T2 = np.eye(4); T2[:3, :3] = cv2.Rodrigues(np.array([0.0, 0.3, 0.0]))[0]; T2[:3, 3] = [0.2, 0, 0.02]
    X = rs.uniform(-0.1, 0.1, (n, 3)) + [0, 0, 0.6]
    P = [K @ np.linalg.inv(T)[:3] for T in (T1, T2)]
    xs = []
    for Pi in P:
        h = np.c_[X, np.ones(n)] @ Pi.T
        xs.append(h[:, :2] / h[:, 2:])
    x1, x2 = torch.tensor(xs[0], device="cuda"), torch.tensor(xs[1], device="cuda")
    ms = timed(lambda: tracker.apply_triangulation(K, K, T1, T2, x1, x2))
    res["triangulate_1M"] = {"ms": round(ms, 4), "alg_GB_s": round(n * (32 + 24) / ms / 1e6, 1)}
    t0 = time.perf_counter()
    cv2.triangulatePoints(P[0], P[1], xs[0][:k].T, xs[1][:k].T)
    res["triangulate_1M"]["cpu_cv2_ms_scaled"] = round((time.perf_counter() - t0) * 1e3 * n / k, 1)
    # bundle adjustment
    def ba():
        p = torch.tensor(d["points0"], device="cuda"); c = torch.tensor(d["cams0"][:, :6], device="cuda").contiguous()
        tracker.ba_solve(p, c, d["features"], d["pt_idx"], d["cam_idx"])
    ms = timed(ba, 5)
    res["ba_20cams_2000pts_16000obs_5it"] = {"ms_incl_host_setup": round(ms, 3)}
    small = tro.tracker_inputs(0, n_cams=8, n_pts=200, obs_per_pt=8)
    t0 = time.perf_counter()
    tro.ba_lm(small["points0"], small["cams0"], small["features"], small["pt_idx"], small["cam_idx"])
    res["ba_20cams_2000pts_16000obs_5it"]["cpu_port_dense_lm_8cams_200pts_ms"] = round((time.perf_counter() - t0) * 1e3, 1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
