"""Drop-in ``ransac_PnP`` and its batched form on the sm_100a PnP-RANSAC kernels.

Reference: ``src/utils/eval_utils.py:18-42`` --
``ransac_PnP(K, pts_2d, pts_3d, scale=1) -> (pose [3,4], pose_homo [4,4], inliers)``, a wrapper over
``cv2.solvePnPRansac(reprojectionError=5, iterationsCount=10000, flags=SOLVEPNP_EPNP)``.  Here the
hypothesis solves, inlier counting, winner selection and the refit run in ``op_pnp_ransac``; the host
only generates the minimal-sample index stream, which is cv2's own (``cv::RNG(2^64-1)``
multiply-with-carry, five distinct indices per iteration), so results can be compared to cv2 sample
for sample.  Failure behaviour follows the wrapper: fewer than 4 points (cv2.error) gives the
identity pose and ``[]`` and nothing ever raises.  No CPU fallback exists: a CUDA device is required.
"""
from __future__ import annotations

import functools

import numpy as np
import torch

from ._lib import call, lib, ptr, stream_ptr

REPROJ_ERROR = 5.0          # eval_utils.py:28
MAX_ITERS = 10000           # eval_utils.py:29
CONFIDENCE = 0.99           # cv2 default
# Upper bound of hypotheses per frame = cv2's iterationsCount: the sequential rule may run that far when the best inlier ratio stays
# below ~26 % (adaptive niters > 4096).  They are evaluated in adaptive rounds [0,32), [32,256), [256,1024), [1024,4096), [4096,8192),
# [8192,10000): a frame whose replayed bound is reached costs nothing in later rounds, so only hard frames pay for the tail.
DEFAULT_HYPOTHESES = MAX_ITERS

_M32 = 0xFFFFFFFF


@functools.lru_cache(maxsize=256)
def cv_sample_stream(npoints: int, count: int) -> np.ndarray:
    """First ``count`` 5-point samples that cv2's RANSAC draws for ``npoints`` correspondences.

    cv::RNG((uint64)-1): state = (state & 0xffffffff) * 4164903690 + (state >> 32); the draw is the low
    32 bits modulo ``npoints``; duplicates inside a sample are redrawn in order (getSubset).
    """
    out = np.zeros((count, 5), dtype=np.int32)
    if npoints < 5:
        return out
    state = 0xFFFFFFFFFFFFFFFF
    for it in range(count):
        idx = []
        while len(idx) < 5:
            state = ((state & _M32) * 4164903690 + (state >> 32)) & 0xFFFFFFFFFFFFFFFF
            v = (state & _M32) % npoints
            if v not in idx:
                idx.append(v)
        out[it] = idx
    return out


LM_ITERS = 20               # cv2's solvePnPRefineLM default iteration cap


def ransac_pnp_batched(K, pts2d, pts3d, counts, samples=None, sample_set=None, scale=1.0,
                       hypotheses=DEFAULT_HYPOTHESES, return_debug=False, refine=None):
    """Batched PnP-RANSAC on device.

    K [F,3,3] float64, pts2d [F,S,2] / pts3d [F,S,3] float64 (pts3d already multiplied by ``scale``),
    counts [F] int32 valid rows per frame, all CUDA tensors.  ``samples`` [n_sets,H,5] int32 with
    ``sample_set`` [F] selecting a set per frame (host-supplied samples); when ``None`` cv2's own stream is
    generated on the device for each frame's count.
    Returns dict(pose [F,3,4] float64 with t/scale, inlier_mask [F,S] bool, n_inliers, iters_run).
    """
    if not pts2d.is_cuda:
        raise RuntimeError("onepose_b200.pnp runs on CUDA only (no CPU fallback)")
    dev = pts2d.device
    F, S = pts2d.shape[0], pts2d.shape[1]
    pts2d = pts2d.to(torch.float64).contiguous()
    pts3d = pts3d.to(torch.float64).contiguous()
    K = K.to(device=dev, dtype=torch.float64).contiguous()
    counts = counts.to(device=dev, dtype=torch.int32).contiguous()
    if samples is None:          # cv2's own stream, generated lazily on the device inside op_pnp_ransac
        H = int(hypotheses)
    else:
        samples = samples.to(device=dev, dtype=torch.int32).contiguous()
        sample_set = sample_set.to(device=dev, dtype=torch.int32).contiguous()
        H = samples.shape[1]
    pose = torch.empty(F, 3, 4, dtype=torch.float64, device=dev)
    mask = torch.empty(F, S, dtype=torch.uint8, device=dev)
    n_inl = torch.empty(F, dtype=torch.int32, device=dev)
    iters = torch.empty(F, dtype=torch.int32, device=dev)
    hyp_counts = torch.empty(F, H, dtype=torch.int32, device=dev)
    hyp_pose = torch.empty(F, H, 6, dtype=torch.float64, device=dev)      # only evaluated hypotheses are ever read back
    nbytes = lib.op_pnp_workspace_bytes(F, S, H)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        call("op_pnp_ransac", ptr(pts2d), ptr(pts3d), ptr(counts), F, S, ptr(K), ptr(samples), ptr(sample_set), H,
                                REPROJ_ERROR, CONFIDENCE, MAX_ITERS, float(scale), ptr(pose), ptr(mask), ptr(n_inl),
                                ptr(iters), ptr(hyp_counts), ptr(hyp_pose), ptr(ws), nbytes, stream_ptr())
    out = dict(pose=pose, inlier_mask=mask.view(torch.bool), n_inliers=n_inl, iters_run=iters)      # the kernel writes 0 / 1 bytes
    if refine not in (None, "lm"):
        raise ValueError(f"refine={refine!r}: None (cv2's result, the parity mode) or 'lm'")
    if refine == "lm":
        # opt-in polish (north_star item 4): Levenberg-Marquardt on the winner's inliers, float64, analytic Jacobian (csrc/pnp_lm.cuh)
        rms = torch.empty(F, 2, dtype=torch.float64, device=dev)
        with torch.cuda.device(dev):
            call("op_pnp_refine_lm", ptr(pts2d), ptr(pts3d), ptr(counts), F, S, ptr(K), ptr(mask), ptr(n_inl), float(scale), LM_ITERS, ptr(pose),
                 ptr(rms), stream_ptr())
        out["reproj_rms"] = rms
    if return_debug:
        out.update(hyp_counts=hyp_counts, hyp_pose=hyp_pose)
    return out


def ransac_PnP(K, pts_2d, pts_3d, scale=1, device=None, refine=None):
    """Same signature / return triple as the reference ``ransac_PnP`` (eval_utils.py:18-42).

    numpy or torch inputs; numpy float64 outputs: ``pose`` [3,4], ``pose_homo`` [4,4] and ``inliers`` (int32
    [n,1] index array like cv2's, or ``[]`` on failure).
    """
    dev = torch.device(device) if device is not None else (
        pts_2d.device if isinstance(pts_2d, torch.Tensor) and pts_2d.is_cuda else torch.device("cuda"))
    as_t = lambda a: (a.detach() if isinstance(a, torch.Tensor) else torch.as_tensor(np.asarray(a))).to(
        device=dev, dtype=torch.float64)
    p2, p3, Kt = as_t(pts_2d).reshape(-1, 2), as_t(pts_3d).reshape(-1, 3) * float(scale), as_t(K).reshape(3, 3)
    n = p2.shape[0]
    if n < 4 or p3.shape[0] != n:          # cv2.error branch of the reference
        print("CV ERROR")
        return np.eye(4)[:3], np.eye(4), []
    S = max(n, 1)
    r = ransac_pnp_batched(Kt[None], p2[None], p3[None], torch.tensor([n], dtype=torch.int32, device=dev),
                           scale=float(scale), hypotheses=_hypotheses_for(n), refine=refine)
    pose = r["pose"][0].cpu().numpy()
    pose_homo = np.concatenate([pose, np.array([[0, 0, 0, 1.0]])], axis=0)
    n_in = int(r["n_inliers"][0])
    inliers = [] if n_in == 0 else torch.nonzero(r["inlier_mask"][0, :S]).to(torch.int32).cpu().numpy()
    return pose, pose_homo, inliers


def _hypotheses_for(n):
    # the sequential rule never needs more than MAX_ITERS samples; one launch evaluates a fixed block
    return DEFAULT_HYPOTHESES if n > 5 else 1


def gather_matches_device(matches0, scores0, kpts2d, kpts3d, scale=1000.0):
    """Device-side valid-match gather of the caller (inference.py:147-152): matches0 [B,N2] int64, scores0 [B,N2], kpts2d [B,N2,2],
    kpts3d [B,N3,3] (batch stride may be 0).  Returns dict(mkpts2d [B,N2,2] f64, mkpts3d_scaled [B,N2,3] f64 = kpts3d * scale, mconf,
    match_query_idx, n_matches), valid rows first in ascending query order, zeros behind them."""
    dev = matches0.device
    B, N2 = matches0.shape
    N3 = kpts3d.shape[1]
    k2 = kpts2d.float().contiguous()
    shared3 = kpts3d.shape[0] == 1 or kpts3d.stride(0) == 0
    k3 = kpts3d[:1].float().contiguous() if shared3 else kpts3d.float().contiguous()
    out2d = torch.empty(B, N2, 2, dtype=torch.float64, device=dev)
    out3d = torch.empty(B, N2, 3, dtype=torch.float64, device=dev)
    conf = torch.empty(B, N2, dtype=torch.float32, device=dev)
    qidx = torch.empty(B, N2, dtype=torch.int32, device=dev)
    count = torch.empty(B, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        call("op_gather_matches", ptr(matches0), ptr(scores0.contiguous()), ptr(k2), ptr(k3), 0 if shared3 else N3 * 3,
                                    B, N2, N3, float(scale), ptr(out2d), ptr(out3d), ptr(conf), ptr(qidx), ptr(count),
                                    stream_ptr())
    return dict(mkpts2d=out2d, mkpts3d_scaled=out3d, mconf=conf, match_query_idx=qidx, n_matches=count)


def pose_from_gathered(g, K, scale=1000.0, hypotheses=DEFAULT_HYPOTHESES, refine=None):
    """PnP-RANSAC on gathered matches (inference.py:153-155).  K [3,3] or [B,3,3]; returns the pose keys of ``pred``.
    refine="lm": optional Levenberg-Marquardt polish of the pose on the winner's inliers (off by default: cv2's last step is EPnP)."""
    dev = g["mkpts2d"].device
    B = g["mkpts2d"].shape[0]
    Kt = torch.as_tensor(np.asarray(K) if not isinstance(K, torch.Tensor) else K).to(device=dev, dtype=torch.float64)
    Kt = Kt.reshape(-1, 3, 3)
    if Kt.shape[0] == 1:
        Kt = Kt.expand(B, 3, 3)
    r = ransac_pnp_batched(Kt, g["mkpts2d"], g["mkpts3d_scaled"], g["n_matches"], scale=scale, hypotheses=hypotheses, refine=refine)
    return dict(mkpts2d=g["mkpts2d"], mkpts3d=g["mkpts3d_scaled"] / scale, mconf=g["mconf"], n_matches=g["n_matches"],
                match_query_idx=g["match_query_idx"], pose=r["pose"], inliers=r["inlier_mask"], n_inliers=r["n_inliers"],
                ransac_iters=r["iters_run"])


def pose_from_matches(matches0, scores0, kpts2d, kpts3d, K, scale=1000.0, hypotheses=DEFAULT_HYPOTHESES, refine=None):
    """Device-side version of the caller's gather + PnP (inference.py:147-155).

    matches0 [B,N2] int64, scores0 [B,N2], kpts2d [B,N2,2], kpts3d [B,N3,3] (batch stride may be 0),
    K [3,3] or [B,3,3].  Returns the extra ``pred`` keys documented in ``GATsSuperGlue.forward``.
    """
    return pose_from_gathered(gather_matches_device(matches0, scores0, kpts2d, kpts3d, scale), K, scale, hypotheses, refine)
