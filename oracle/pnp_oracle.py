"""ctypes front-end of ``oracle/pnp_oracle.c`` plus the reference-shaped wrappers.

TEST INFRASTRUCTURE ONLY (see the header of ``pnp_oracle.c``).  ``ransac_pnp_restated`` follows
``src/utils/eval_utils.py:18-42`` with the C restatement in place of ``cv2.solvePnPRansac``;
``ransac_pnp_cv2`` is the same wrapper over the image's own cv2 (the third-party library the
reference calls), used as the live pin and as the timed CPU baseline.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import build_oracle

_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build_oracle.build())
        _lib.oracle_epnp.restype = C.c_double
        _lib.oracle_ransac_epnp.restype = C.c_int
        _lib.oracle_sequential_select.restype = C.c_int
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def cv_sample_stream(npoints, count, model_points=5):
    """First ``count`` minimal samples cv2's RANSAC draws for ``npoints`` correspondences."""
    out = np.empty((count, model_points), dtype=np.int32)
    lib().oracle_cv_sample_stream(int(npoints), int(count), int(model_points), _p(out, C.c_int32))
    return out


def epnp(pws, us, fu=1.0, fv=1.0, uc=0.0, vc=0.0):
    pws = np.ascontiguousarray(pws, dtype=np.float64)
    us = np.ascontiguousarray(us, dtype=np.float64)
    R = np.empty(9); t = np.empty(3)
    err = lib().oracle_epnp(len(pws), _p(pws, C.c_double), _p(us, C.c_double), C.c_double(fu), C.c_double(fv),
                            C.c_double(uc), C.c_double(vc), _p(R, C.c_double), _p(t, C.c_double))
    return R.reshape(3, 3), t, err


def hypothesis(X32, uv32, K, idx):
    X32 = np.ascontiguousarray(X32, dtype=np.float32); uv32 = np.ascontiguousarray(uv32, dtype=np.float32)
    K = np.ascontiguousarray(K, dtype=np.float64); idx = np.ascontiguousarray(idx, dtype=np.int32)
    rvec = np.empty(3); tvec = np.empty(3)
    lib().oracle_hypothesis(_p(X32, C.c_float), _p(uv32, C.c_float), _p(K, C.c_double), _p(idx, C.c_int32),
                            len(idx), _p(rvec, C.c_double), _p(tvec, C.c_double))
    return rvec, tvec


def reproj_err(X32, uv32, K, rvec, tvec):
    X32 = np.ascontiguousarray(X32, dtype=np.float32); uv32 = np.ascontiguousarray(uv32, dtype=np.float32)
    K = np.ascontiguousarray(K, dtype=np.float64)
    rvec = np.ascontiguousarray(rvec, dtype=np.float64).ravel(); tvec = np.ascontiguousarray(tvec, dtype=np.float64).ravel()
    err = np.empty(len(X32), dtype=np.float32)
    lib().oracle_reproj_err(len(X32), _p(X32, C.c_float), _p(uv32, C.c_float), _p(K, C.c_double),
                            _p(rvec, C.c_double), _p(tvec, C.c_double), _p(err, C.c_float))
    return err


def sequential_select(npoints, counts, confidence=0.99, model_points=5, max_iters=10000):
    counts = np.ascontiguousarray(counts, dtype=np.int32)
    it = C.c_int(0)
    best = lib().oracle_sequential_select(int(npoints), _p(counts, C.c_int32), len(counts), C.c_double(confidence),
                                          model_points, max_iters, C.byref(it))
    return best, it.value


def rodrigues_to_mat(rvec):
    r = np.ascontiguousarray(rvec, dtype=np.float64).ravel(); R = np.empty(9)
    lib().oracle_rodrigues_to_mat(_p(r, C.c_double), _p(R, C.c_double))
    return R.reshape(3, 3)


def ransac_epnp(pts3d, pts2d, K, thr=5.0, confidence=0.99, max_iters=10000, return_samples=False):
    """Restated ``cv2.solvePnPRansac(..., flags=SOLVEPNP_EPNP)`` for n > 5 correspondences."""
    X = np.ascontiguousarray(pts3d, dtype=np.float64); uv = np.ascontiguousarray(pts2d, dtype=np.float64)
    K = np.ascontiguousarray(K, dtype=np.float64)
    n = len(X)
    rvec = np.zeros(3); tvec = np.zeros(3)
    inl = np.empty(max(n, 1), dtype=np.int32)
    n_inl = C.c_int(0); iters = C.c_int(0)
    samples = np.empty((max_iters, 5), dtype=np.int32) if return_samples else None
    ok = lib().oracle_ransac_epnp(n, _p(X, C.c_double), _p(uv, C.c_double), _p(K, C.c_double), C.c_double(thr),
                                  C.c_double(confidence), max_iters, _p(rvec, C.c_double), _p(tvec, C.c_double),
                                  _p(inl, C.c_int32), C.byref(n_inl), C.byref(iters),
                                  _p(samples, C.c_int32) if return_samples else None)
    out = dict(ok=bool(ok), rvec=rvec, tvec=tvec, inliers=inl[:n_inl.value].copy(), iters=iters.value)
    if return_samples:
        out["samples"] = samples[:iters.value].copy()
    return out


def ransac_pnp_restated(K, pts_2d, pts_3d, scale=1):
    """eval_utils.py:18-42 with the restated RANSAC (only n > 5 is restated; other sizes -> cv2 wrapper)."""
    pts_2d = np.ascontiguousarray(np.asarray(pts_2d).astype(np.float64))
    pts_3d = np.ascontiguousarray(np.asarray(pts_3d).astype(np.float64)) * scale
    K = np.asarray(K).astype(np.float64)
    if len(pts_2d) <= 5:
        return ransac_pnp_cv2(K, pts_2d, pts_3d / scale, scale)
    r = ransac_epnp(pts_3d, pts_2d, K)
    R = rodrigues_to_mat(r["rvec"])
    pose = np.concatenate([R, (r["tvec"] / scale)[:, None]], axis=-1)
    pose_homo = np.concatenate([pose, np.array([[0, 0, 0, 1.0]])], axis=0)
    inliers = r["inliers"][:, None] if r["ok"] else []
    return pose, pose_homo, inliers


def ransac_pnp_cv2(K, pts_2d, pts_3d, scale=1):
    """eval_utils.py:18-42 verbatim in behaviour, on the image's cv2 (identity pose + [] on cv2.error)."""
    import cv2
    dist = np.zeros((8, 1), dtype=np.float64)
    p2 = np.ascontiguousarray(np.asarray(pts_2d).astype(np.float64))
    p3 = np.ascontiguousarray(np.asarray(pts_3d).astype(np.float64)) * scale
    Kd = np.asarray(K).astype(np.float64)
    try:
        _, rvec, tvec, inl = cv2.solvePnPRansac(p3, p2, Kd, dist, reprojectionError=5,
                                                iterationsCount=10000, flags=cv2.SOLVEPNP_EPNP)
        R = cv2.Rodrigues(rvec)[0]
        pose = np.concatenate([R, tvec / scale], axis=-1)
        return pose, np.concatenate([pose, [[0, 0, 0, 1.0]]], axis=0), ([] if inl is None else inl)
    except cv2.error:
        return np.eye(4)[:3], np.eye(4), []


def pose_error(pose_pred, pose_gt):
    """Rotation (deg) / translation (cm) error, eval_utils.py:45-63."""
    pose_pred = np.asarray(pose_pred)[:3]; pose_gt = np.asarray(pose_gt)[:3]
    t_err = np.linalg.norm(pose_pred[:, 3] - pose_gt[:, 3]) * 100
    tr = np.trace(pose_pred[:, :3] @ pose_gt[:, :3].T)
    tr = min(tr, 3.0)
    return np.rad2deg(np.arccos(np.clip((tr - 1.0) / 2.0, -1, 1))), t_err
