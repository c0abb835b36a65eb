"""Tracker-side geometry  (SURVEY.md 8(f) row f4).

Mirrors the geometric core of the reference's (work-in-progress) ``BATracker`` (src/tracker/ba_tracker.py) and its helpers
(src/tracker/tracking_utils.py); the image side (LK optical flow, video writer, key-frame book-keeping) is out of scope.

* ``project``                      -- tracking_utils.py:74-89
* ``SnavelyReprojectionErrorV2``   -- tracking_utils.py:143-169 (+ ``AngleAxisRotatePoint`` :91-140)
* ``apply_triangulation``          -- ba_tracker.py:267-273 (cv2.triangulatePoints)
* ``flow_pose``                    -- the geometric half of ``flow_track`` (ba_tracker.py:311-320): PnP on the flow-tracked key-frame
                                      correspondences (the second ``ransac_PnP`` call site) + mean reprojection distance
* ``apply_ba``                     -- ba_tracker.py:358-441, same arguments and return pair.  The reference calls DeepLM's ``Solve``
                                      (``from DeepLM.TorchLM.solver import Solve``, third-party, not vendored, no pinned version);
                                      here ``op_ba_lm``: Levenberg-Marquardt + Schur complement on the device (csrc/tracker.cu).
* ``get_cam_param`` / ``get_cam_params_back`` -- ba_tracker.py:443-466 (BAL camera vector <-> K, pose)

float64 CUDA tensors; no CPU fallback.
"""
import numpy as np
import torch

from ._lib import call, lib, ptr, stream_ptr


def _dev(x, device=None):
    t = torch.as_tensor(np.asarray(x) if not isinstance(x, torch.Tensor) else x)
    dev = torch.device(device) if device is not None else (t.device if t.is_cuda else torch.device("cuda"))
    if dev.type != "cuda":
        raise RuntimeError("onepose_b200.tracker: CUDA only (no CPU fallback)")
    return t.to(device=dev, dtype=torch.float64).contiguous()


def project(xyz, K, RT, need_depth=False, ref2d=None):
    """tracking_utils.py:74-89.  xyz [N,3], K [3,3], RT [3,4] -> xy [N,2] (, depth [N]).  With ``ref2d`` [N,2] the mean distance
    |xy - ref2d| is returned as a third / second element (flow_track's kpt_dist), computed in the same kernel."""
    xyz = _dev(xyz); K = _dev(K, xyz.device); RT = _dev(RT, xyz.device)[:3, :4].contiguous()
    n = xyz.shape[0]
    xy = torch.empty(n, 2, dtype=torch.float64, device=xyz.device)
    depth = torch.empty(n, dtype=torch.float64, device=xyz.device) if need_depth else None
    ref = _dev(ref2d, xyz.device) if ref2d is not None else None
    dist = torch.zeros(1, dtype=torch.float64, device=xyz.device) if ref is not None else None
    with torch.cuda.device(xyz.device):
        call("op_project_points", ptr(xyz), ptr(K), ptr(RT), n, ptr(ref), ptr(xy), ptr(depth), ptr(dist), stream_ptr())
    out = (xy, depth) if need_depth else (xy,)
    if ref is not None:
        out = out + (dist[0] / max(n, 1),)
    return out if len(out) > 1 else out[0]


def SnavelyReprojectionErrorV2(points_ob, cameras_ob, features, ptIdx=None, camIdx=None):
    """tracking_utils.py:143-169: residual [n,2].  As in the reference ``points_ob`` / ``cameras_ob`` are per observation ([n,3] / [n,6]
    or [n,1,.]); with ``ptIdx`` / ``camIdx`` they are the variable tables and the gather happens in the kernel."""
    points_ob = _dev(points_ob); cameras_ob = _dev(cameras_ob, points_ob.device); features = _dev(features, points_ob.device)
    if points_ob.dim() == 3:
        points_ob, cameras_ob = points_ob[:, 0, :].contiguous(), cameras_ob[:, 0, :].contiguous()
    cameras_ob = cameras_ob[:, :6].contiguous()
    n = features.shape[0]
    pi = torch.as_tensor(ptIdx).to(points_ob.device, torch.int64).contiguous() if ptIdx is not None else None
    ci = torch.as_tensor(camIdx).to(points_ob.device, torch.int64).contiguous() if camIdx is not None else None
    out = torch.empty(n, 2, dtype=torch.float64, device=points_ob.device)
    with torch.cuda.device(points_ob.device):
        call("op_snavely_residual", ptr(points_ob), ptr(cameras_ob), ptr(features), ptr(pi), ptr(ci), n, ptr(out), stream_ptr())
    return out


def apply_triangulation(K1, K2, Tcw1, Tcw2, kpt2d_1, kpt2d_2):
    """ba_tracker.py:267-273: poses are camera-to-world 4x4 (inverted for the projection matrices, as the reference does)."""
    K1 = _dev(K1); dev = K1.device
    P1 = (K1 @ torch.linalg.inv(_dev(Tcw1, dev))[:3, :]).contiguous()
    P2 = (_dev(K2, dev) @ torch.linalg.inv(_dev(Tcw2, dev))[:3, :]).contiguous()
    x1, x2 = _dev(kpt2d_1, dev), _dev(kpt2d_2, dev)
    X = torch.empty(x1.shape[0], 3, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        call("op_triangulate_points", ptr(P1), ptr(P2), ptr(x1), ptr(x2), x1.shape[0], ptr(X), stream_ptr())
    return X


def flow_pose(K_crop, mkpts2d_query, kpt3ds_kf, scale=1):
    """ba_tracker.py:311-320 without the images: pose from the flow-tracked key-frame correspondences + mean reprojection distance.
    -> (pose_init [3,4], pose_init_homo [4,4], inliers, kpt_dist)."""
    from .pnp import ransac_PnP
    pose, pose_homo, inliers = ransac_PnP(K_crop, mkpts2d_query, kpt3ds_kf, scale=scale)
    _, kpt_dist = project(kpt3ds_kf, K_crop, pose, ref2d=mkpts2d_query)
    return pose, pose_homo, inliers, float(kpt_dist)


def get_cam_param(K, pose):
    """ba_tracker.py:459-466: (K, pose) -> BAL vector [rvec(3), t(3), f, k1, k2]."""
    import cv2
    K, pose = np.asarray(K), np.asarray(pose)
    return np.concatenate([cv2.Rodrigues(pose[:3, :3])[0].flatten(), pose[:3, 3], [K[0, 0], K[0, 2], K[1, 2]]])


def get_cam_params_back(cam_params):
    """ba_tracker.py:443-457."""
    import cv2
    cam_params = np.asarray(cam_params)
    f, k1, k2 = cam_params[6:9]
    pose = np.eye(4)
    pose[:3, :3] = cv2.Rodrigues(cam_params[:3])[0]
    pose[:3, 3] = cam_params[3:6]
    return np.array([[f, 0, k1], [0, f, k2], [0, 0, 1]]), pose


def ba_solve(points, cam_pose, features, ptIdx, camIdx, numIterations=5, numSuccessIterations=5, lambda0=1e-4, return_state=False):
    """The ``Solve(variables=[points, cam_pose], constants=[features], indices=[ptIdx, camIdx], fn=SnavelyReprojectionErrorV2, ...)``
    call of ba_tracker.py:401-407: ``points`` [P,3] and ``cam_pose`` [C,6] float64 CUDA tensors are optimised IN PLACE."""
    if not (points.is_cuda and cam_pose.is_cuda and points.dtype == torch.float64 and cam_pose.dtype == torch.float64):
        raise RuntimeError("ba_solve: float64 CUDA tensors only")
    dev = points.device
    P, O = points.shape[0], features.shape[0]
    pt = np.asarray(ptIdx.cpu() if isinstance(ptIdx, torch.Tensor) else ptIdx).astype(np.int64)
    cm_all = np.asarray(camIdx.cpu() if isinstance(camIdx, torch.Tensor) else camIdx).astype(np.int64)
    used, cm = np.unique(cm_all, return_inverse=True)                         # only the cameras of the window are variables
    d_used = torch.as_tensor(used).to(dev)
    cam_all, cam_pose = cam_pose, cam_pose[d_used].contiguous()
    C = cam_pose.shape[0]
    order = np.argsort(pt, kind="stable")                                     # host plumbing: observations grouped by point
    off = np.zeros(P + 1, np.int32); np.add.at(off, pt + 1, 1); off = np.cumsum(off).astype(np.int32)
    i32 = lambda a: torch.as_tensor(np.ascontiguousarray(a, dtype=np.int32)).to(dev)
    d_cm, d_pt, d_off, d_obs = i32(cm), i32(pt), i32(off), i32(order)
    feats = _dev(features, dev)
    nbytes = lib.op_ba_workspace_bytes(C, P, O)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    state = torch.empty(8, dtype=torch.float64, device=dev)
    assert points.is_contiguous() and cam_pose.is_contiguous()
    with torch.cuda.device(dev):
        call("op_ba_lm", ptr(cam_pose), ptr(points), ptr(feats), ptr(d_cm), ptr(d_pt), ptr(d_off), ptr(d_obs), C, P, O, int(numIterations),
             int(numSuccessIterations), float(lambda0), ptr(ws), nbytes, ptr(state), stream_ptr())
    cam_all[d_used] = cam_pose
    return state if return_state else None


def apply_ba(kpt2ds, kpt2d3d_ids, kpt2d_fids, kpt3d_list, cams, win_size=20, verbose=False):
    """ba_tracker.py:358-441: same arguments (+ the tracker's ``win_size`` attribute) and return pair (points_opt [P,3], cam_opt [F,9])
    as numpy arrays.  Cameras are BAL vectors [rvec, t, f, k1, k2]; only observations of mapped points inside the window enter."""
    kpt2ds, kpt2d3d_ids, kpt2d_fids = np.asarray(kpt2ds), np.asarray(kpt2d3d_ids), np.asarray(kpt2d_fids)
    kpt3d_list, cams = np.asarray(kpt3d_list, dtype=np.float64), np.asarray(cams, dtype=np.float64)
    points = torch.tensor(kpt3d_list, device="cuda", dtype=torch.float64)
    cam_pose = torch.tensor(cams[:, :6], device="cuda", dtype=torch.float64)
    valid2d_idx = np.where(kpt2d3d_ids != -1)[0]
    if np.max(kpt2d_fids) > win_size:
        valid_idx_fid = np.where(kpt2d_fids > np.max(kpt2d_fids) - win_size)
        valid2d_idx = np.intersect1d(valid_idx_fid, valid2d_idx)
    ks = cams[np.array(kpt2d_fids[valid2d_idx], dtype=int), 6:]
    features = np.concatenate([kpt2ds[valid2d_idx], ks], axis=1)
    ba_solve(points, cam_pose, features, kpt2d3d_ids[valid2d_idx], kpt2d_fids[valid2d_idx], numIterations=5, numSuccessIterations=5)
    cam_opt = np.concatenate([cam_pose.cpu().numpy(), cams[:, 6:]], axis=1)
    return points.cpu().numpy(), cam_opt
