"""Seeded synthetic inputs for the GATsSPG + PnP hot path.

Everything is drawn from ``numpy.random.default_rng`` (PCG64) so the same seed gives the same
bits on the build container and on the GPU box.  The layouts are the public ones the reference
consumes (``GATs_SuperGlue.py:209-218``, produced by ``inference.py:80-94``):

* ``descriptors2d_query`` [B, 256, N2], ``descriptors3d_db`` [B, 256, N3],
  ``descriptors2d_db`` [B, 256, N3*L] with the L leaves of point n in columns n*L .. n*L+L-1,
* ``keypoints2d`` [B, N2, 2] pixels, ``keypoints3d`` [B, N3, 3] metres.

The object model follows the on-disk DB conventions of the reference: the 3D descriptor is the
un-normalised mean of the real leaves (``src/sfm/postprocess/feature_process.py:297-305``) and
missing leaves are the all-ones dustbin descriptor (``src/utils/data_utils.py:175-190``).
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np
import torch

# model hparams the hot path reads (configs/experiment/train_GATsSPG.yaml:44-60)
DEFAULT_HPARAMS = {
    "descriptor_dim": 256,
    "keypoints_encoder": [32, 64, 128],
    "sinkhorn_iterations": 100,
    "match_threshold": 0.2,
    "match_type": "softmax",
    "scale_factor": 0.07,
    "include_self": True,
    "with_linear_transform": False,
    "additional": False,
}

GNN_LAYER_KINDS = ("GATs", "self", "cross") * 4

# named shapes of BASELINE.json's configs
CONFIGS = {
    "cpu_ref": dict(n2=1024, n3=4096, leaves=8, frames=1),
    "headline": dict(n2=2048, n3=8192, leaves=8, frames=64),
    "large": dict(n2=4096, n3=32768, leaves=16, frames=1),
    "pnp_stress": dict(ncorr=10000, inlier_ratio=0.3, hypotheses=4096, frames=64),
    "sweep": dict(n2=2048, n3=8192, leaves=8, objects=8, frames=512),
}

INTRINSIC = np.array([[600.0, 0, 256.0], [0, 600.0, 256.0], [0, 0, 1.0]])


def weight_shapes(hparams=DEFAULT_HPARAMS):
    """Name -> shape of every parameter of the reference ``GATsSuperGlue`` state_dict."""
    D = hparams["descriptor_dim"]
    shapes = {}
    for tag, inp in (("kenc_2d", 3), ("kenc_3d", 4)):
        chans = [inp] + list(hparams["keypoints_encoder"]) + [D]
        for i in range(1, len(chans)):
            idx = (i - 1) * 3                     # Conv, InstanceNorm, ReLU triples of MLP()
            shapes[f"{tag}.encoder.{idx}.weight"] = (chans[i], chans[i - 1], 1)
            shapes[f"{tag}.encoder.{idx}.bias"] = (chans[i],)
    for i, kind in enumerate(GNN_LAYER_KINDS):
        p = f"gnn.layers.{i}."
        if kind == "GATs":
            shapes[p + "W"] = (D, D)
            shapes[p + "a"] = (2 * D, 1)
        else:
            shapes[p + "attn.merge.weight"] = (D, D, 1)
            shapes[p + "attn.merge.bias"] = (D,)
            for j in range(3):
                shapes[p + f"attn.proj.{j}.weight"] = (D, D, 1)
                shapes[p + f"attn.proj.{j}.bias"] = (D,)
            shapes[p + "mlp.0.weight"] = (2 * D, 2 * D, 1)
            shapes[p + "mlp.0.bias"] = (2 * D,)
            shapes[p + "mlp.3.weight"] = (D, 2 * D, 1)
            shapes[p + "mlp.3.bias"] = (D,)
    shapes["final_proj.weight"] = (D, D, 1)
    shapes["final_proj.bias"] = (D,)
    shapes["bin_score"] = ()
    return shapes


def make_weights(seed=0, damped=False, hparams=DEFAULT_HPARAMS):
    """Random-init weights with the reference's init distributions.

    Conv1d: U(+-1/sqrt(fan_in)) (torch default), GATs W/a: xavier normal with gain 1.414
    (GATs.py:51-54), ``mlp.3.bias = 0`` (GATs_SuperGlue.py:128), ``bin_score = 1`` (:200).
    ``damped=True`` is the weight set of SURVEY.md section 8(c): ``mlp.3.weight *= 0.01`` and
    ``final_proj.bias = 0`` so planted correspondences survive the 0.2 threshold and PnP gets a
    non-empty input.
    """
    rng = np.random.default_rng(seed)
    out = {}
    for name, shp in weight_shapes(hparams).items():
        if name == "bin_score":
            arr = np.array(1.0)
        elif name.endswith(".W") or name.endswith(".a"):
            std = 1.414 * math.sqrt(2.0 / (shp[0] + shp[1]))
            arr = rng.standard_normal(shp) * std
        else:
            fan_in = shp[1] if len(shp) == 3 else None
            if fan_in is None:                      # bias: fan_in of its conv = matching weight
                wshape = weight_shapes(hparams)[name[:-4] + "weight"]
                fan_in = wshape[1]
            bound = 1.0 / math.sqrt(fan_in)
            arr = rng.uniform(-bound, bound, shp)
            if name.endswith("mlp.3.bias"):
                arr = np.zeros(shp)
        if damped:
            if name.endswith("mlp.3.weight"):
                arr = arr * 0.01
            if name == "final_proj.bias":
                arr = np.zeros(shp)
        out[name] = torch.from_numpy(np.asarray(arr, dtype=np.float32).copy())
    return out


def _unit(x):
    return x / np.linalg.norm(x, axis=-1, keepdims=True)


@dataclass
class ObjectModel:
    """One object's SfM database in the public layout (inference.py:113-130)."""
    object_id: int
    base: np.ndarray             # [N3, 256] unit "true" descriptors (generator state, not an input)
    keypoints3d: np.ndarray      # [N3, 3] metres
    descriptors3d_db: np.ndarray  # [256, N3]
    descriptors2d_db: np.ndarray  # [256, N3*L]
    num_leaf: int


def make_object(object_id, n3, num_leaf, sigma=0.3, p_pad=0.3, dim=256):
    rng = np.random.default_rng(1000 + object_id)
    base = _unit(rng.standard_normal((n3, dim)))
    leaves = _unit(base[:, None, :] + sigma * rng.standard_normal((n3, num_leaf, dim)) / 16.0)
    real = np.ones((n3, num_leaf), dtype=bool)
    real[:, 1:] = rng.random((n3, num_leaf - 1)) >= p_pad      # leaf 0 is always real
    avg = (leaves * real[..., None]).sum(1) / real.sum(1, keepdims=True)   # mean of real leaves
    leaves = np.where(real[..., None], leaves, 1.0)            # dustbin = all ones
    kp3 = rng.uniform(-0.1, 0.1, (n3, 3))
    return ObjectModel(
        object_id=object_id, base=base.astype(np.float32),
        keypoints3d=kp3.astype(np.float32),
        descriptors3d_db=np.ascontiguousarray(avg.T.astype(np.float32)),
        descriptors2d_db=np.ascontiguousarray(
            leaves.reshape(n3 * num_leaf, dim).T.astype(np.float32)),
        num_leaf=num_leaf)


def rodrigues(rvec):
    """Axis-angle -> rotation matrix (float64)."""
    rvec = np.asarray(rvec, dtype=np.float64)
    th = np.linalg.norm(rvec)
    if th < 1e-12:
        return np.eye(3)
    k = rvec / th
    Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + math.sin(th) * Kx + (1 - math.cos(th)) * (Kx @ Kx)


@dataclass
class QueryFrame:
    keypoints2d: np.ndarray         # [N2, 2]
    descriptors2d_query: np.ndarray  # [256, N2]
    pose_gt: np.ndarray             # [3, 4] (metres)
    planted: np.ndarray             # [N2] index of the matching 3D point or -1
    K: np.ndarray                   # [3, 3]


def make_frame(obj: ObjectModel, frame_id, n2, sigma=0.3, px_noise=0.5, dim=256):
    rng = np.random.default_rng(obj.object_id * 1_000_000 + frame_id + 7)
    n3 = obj.base.shape[0]
    rvec = rng.standard_normal(3) * 0.5
    R = rodrigues(rvec)
    t = np.array([0.02, -0.01, 0.5])
    n_plant = min(n2 // 2, n3)
    perm = rng.permutation(n3)[:n_plant]
    desc = np.empty((n2, dim))
    desc[:n_plant] = _unit(obj.base[perm] + sigma * rng.standard_normal((n_plant, dim)) / 16.0)
    desc[n_plant:] = _unit(rng.standard_normal((n2 - n_plant, dim)))
    cam = obj.keypoints3d[perm].astype(np.float64) @ R.T + t
    uv = (cam[:, :2] / cam[:, 2:3]) * np.array([INTRINSIC[0, 0], INTRINSIC[1, 1]]) \
        + np.array([INTRINSIC[0, 2], INTRINSIC[1, 2]])
    kp2 = np.empty((n2, 2))
    kp2[:n_plant] = uv + px_noise * rng.standard_normal((n_plant, 2))
    kp2[n_plant:] = rng.uniform(0, 512, (n2 - n_plant, 2))
    planted = np.full(n2, -1, dtype=np.int64)
    planted[:n_plant] = perm
    return QueryFrame(
        keypoints2d=kp2.astype(np.float32),
        descriptors2d_query=np.ascontiguousarray(desc.T.astype(np.float32)),
        pose_gt=np.concatenate([R, t[:, None]], axis=1), planted=planted, K=INTRINSIC.copy())


def pack_batch(obj: ObjectModel, frames, device="cpu", share_db=True):
    """The reference's ``pack_data`` (inference.py:80-94) for a batch of frames of one object.

    With ``share_db`` the static DB tensors are ``expand``-ed (stride-0 batch axis) instead of
    copied per frame, which is how a batched caller hands one object to many frames.
    """
    B = len(frames)
    q = torch.from_numpy(np.stack([f.descriptors2d_query for f in frames])).to(device)
    k2 = torch.from_numpy(np.stack([f.keypoints2d for f in frames])).to(device)
    d3 = torch.from_numpy(obj.descriptors3d_db)[None].to(device)
    lv = torch.from_numpy(obj.descriptors2d_db)[None].to(device)
    k3 = torch.from_numpy(obj.keypoints3d)[None].to(device)
    if share_db:
        d3, lv, k3 = (x.expand(B, *x.shape[1:]) for x in (d3, lv, k3))
    else:
        d3, lv, k3 = (x.repeat(B, *([1] * (x.dim() - 1))) for x in (d3, lv, k3))
    return {"keypoints2d": k2, "keypoints3d": k3, "descriptors2d_query": q,
            "descriptors3d_db": d3, "descriptors2d_db": lv}


def make_pnp_problem(seed, ncorr=10000, inlier_ratio=0.3, px_noise=0.5, scale=1.0):
    """Config 4: synthetic 2D-3D correspondences with a known pose and outlier fraction."""
    rng = np.random.default_rng(50_000 + seed)
    R = rodrigues(rng.standard_normal(3) * 0.5)
    t = np.array([0.02, -0.01, 0.5])
    X = rng.uniform(-0.1, 0.1, (ncorr, 3))
    cam = X @ R.T + t
    uv = cam[:, :2] / cam[:, 2:3] * 600.0 + 256.0 + px_noise * rng.standard_normal((ncorr, 2))
    n_in = int(round(ncorr * inlier_ratio))
    is_inlier = np.zeros(ncorr, dtype=bool)
    is_inlier[rng.permutation(ncorr)[:n_in]] = True
    uv = np.where(is_inlier[:, None], uv, rng.uniform(0, 512, (ncorr, 2)))
    return dict(K=INTRINSIC.copy(), pts2d=uv, pts3d=X * scale,
                pose_gt=np.concatenate([R, t[:, None]], 1), is_inlier=is_inlier)
