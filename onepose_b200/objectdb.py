"""Object DB: on-disk format, leaf-tensor builder and device-resident handle  (SURVEY.md 8(f) row f2).

Mirrors the reference's DB side of ``inference_core`` (inference.py:113-130) and the helpers it calls:

* ``pad_features3d_random``      -- data_utils.py:143-160 (same name, arguments, return pair)
* ``build_features3d_leaves``    -- data_utils.py:163-205 (same name, arguments, return pair; consumes ``np.random`` exactly as
                                    the reference does, so a seeded run gives the same leaves)
* ``mean_descriptors`` / ``mean_scores`` / ``save_3d_anno`` -- feature_process.py:191-195,297-319
* files ``anno_3d_average.npz`` / ``anno_3d_collect.npz`` {keypoints3d [N3,3], descriptors3d [256,n], scores3d [n,1]} and
  ``idxs.npy`` (leaf count per 3D point), feature_process.py:355-363

The index plan (which collected descriptor becomes leaf j of point n) is host logic, as in the reference; the tensors are
built on the GPU by ``op_gather_rows`` / ``op_segment_mean`` directly in the matcher's point-major layout, and
``ObjectDB.register(model)`` hands that layout to ``GATsSuperGlue`` so that no per-frame re-upload or re-transposition of the
72 MiB DB happens (the reference does both in ``pack_data``, inference.py:80-94).  CUDA only; no CPU fallback.
"""
import numpy as np
import torch

from ._lib import call, ptr, stream_ptr

_D = 256


# ----------------------------------------------------------------------------------------------- host-side index plans
def leaf_plan(idxs, n_target_shape, num_leaf, rng=np.random):
    """Affiliation plan of build_features3d_leaves (data_utils.py:180-197,199-203): for each of the n_target_shape points
    num_leaf indices into the collected descriptors; -1 = the all-ones dustbin.  Draws ``rng.permutation`` once per point,
    in point order, on the same arrays as the reference, so ``np.random.seed(s)`` reproduces the reference's choice."""
    idxs = np.asarray(idxs).astype(np.int64).ravel()
    orig_num = idxs.shape[0]
    dustbin_id = int(idxs.sum())                      # the reference appends the dustbin column after the last descriptor
    upper = np.cumsum(idxs, axis=0)
    lower = np.insert(upper[:-1], 0, 0) if orig_num else upper
    plan = []
    for start, end in zip(lower, upper):
        if num_leaf > end - start:
            ids = np.arange(start, end).tolist()
            ids += [dustbin_id] * int(num_leaf - (end - start))
            plan.append(rng.permutation(np.array(ids)))
        else:
            plan.append(rng.permutation(np.arange(start, end))[:num_leaf])
    plan = np.concatenate(plan, axis=0).astype(np.int64) if plan else np.zeros(0, np.int64)
    assert plan.shape[0] == orig_num * num_leaf
    plan[plan == dustbin_id] = -1
    n_pad = n_target_shape - orig_num
    if n_pad < 0:
        plan = plan[:num_leaf * n_target_shape]
    else:
        plan = np.concatenate([plan, np.full(n_pad * num_leaf, -1, np.int64)])
    return plan


def _as_cuda_f32(x, device):
    t = torch.as_tensor(np.asarray(x) if not isinstance(x, torch.Tensor) else x)
    return t.to(device=device, dtype=torch.float32).contiguous()


def _device(device):
    dev = torch.device(device if device is not None else "cuda")
    if dev.type != "cuda":
        raise RuntimeError("onepose_b200.objectdb builds the DB on a CUDA device only (no CPU fallback)")
    return dev


def gather_rows(src_pm, plan, device=None):
    """src_pm [n,256] fp32 CUDA, plan int64 [m] (-1 = ones) -> [m,256] (op_gather_rows)."""
    dev = src_pm.device
    idx = torch.as_tensor(plan, dtype=torch.int64).to(dev)
    out = torch.empty(idx.numel(), _D, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        call("op_gather_rows", ptr(src_pm), ptr(idx), src_pm.shape[0], idx.numel(), ptr(out), stream_ptr())
    return out


# ----------------------------------------------------------------------------------------------- reference-named helpers
def pad_features3d_random(descriptors, scores, n_target_shape, device=None):
    """data_utils.py:143-160: pad (all-ones descriptors, zero scores) or truncate to n_target_shape points.
    descriptors [256,n], scores [n,1] -> ([256,n_target], [n_target,1]) CUDA fp32."""
    dev = _device(device if device is not None else (descriptors.device if isinstance(descriptors, torch.Tensor) and descriptors.is_cuda else None))
    d = _as_cuda_f32(descriptors, dev)
    s = _as_cuda_f32(scores, dev)
    n = d.shape[1]
    plan = np.arange(n_target_shape, dtype=np.int64)
    plan[plan >= n] = -1
    out = gather_rows(d.t().contiguous(), plan).t()                         # channel-major view of the point-major tensor
    s_out = torch.zeros(n_target_shape, 1, dtype=torch.float32, device=dev)
    s_out[:min(n, n_target_shape)] = s[:n_target_shape]
    return out, s_out


def build_features3d_leaves(descriptors, scores, idxs, n_target_shape, num_leaf, device=None, rng=np.random, return_point_major=False):
    """data_utils.py:163-205.  descriptors [256, sum(idxs)], scores [sum(idxs),1], idxs [n] leaf counts ->
    (descriptors [256, n_target*num_leaf], scores [n_target*num_leaf, 1]); leaf j of point n is column n*num_leaf+j.
    The descriptor result is a channel-major VIEW of the point-major tensor the matcher wants; with
    ``return_point_major`` the [n_target*num_leaf, 256] tensor itself is returned as a third element."""
    dev = _device(device if device is not None else (descriptors.device if isinstance(descriptors, torch.Tensor) and descriptors.is_cuda else None))
    plan = leaf_plan(idxs, n_target_shape, num_leaf, rng)
    d = _as_cuda_f32(descriptors, dev)
    s = _as_cuda_f32(scores, dev).reshape(-1, 1)
    pm = gather_rows(d.t().contiguous(), plan)
    pidx = torch.as_tensor(plan).to(dev)
    s_out = torch.where(pidx[:, None] >= 0, s[pidx.clamp_min(0)], torch.zeros((), device=dev))
    if return_point_major:
        return pm.t(), s_out, pm
    return pm.t(), s_out


def _segment_mean(rows, idxs, device):
    dev = _device(device)
    x = _as_cuda_f32(rows, dev)
    x = x.reshape(x.shape[0], -1)
    offs = torch.as_tensor(np.concatenate([[0], np.cumsum(np.asarray(idxs).astype(np.int64))])).to(dev)
    out = torch.empty(offs.numel() - 1, x.shape[1], dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        call("op_segment_mean", ptr(x), ptr(offs), offs.numel() - 1, x.shape[1], ptr(out), stream_ptr())
    return out


def mean_descriptors(descriptors, idxs, device=None):
    """feature_process.py:297-305: [sum(idxs),256] leaf descriptors -> [n,256] per-point means (not re-normalised)."""
    return _segment_mean(descriptors, idxs, device)


def mean_scores(scores, idxs, device=None):
    """feature_process.py:308-319."""
    return _segment_mean(scores, idxs, device)


def save_3d_anno(xyzs, descriptors, scores, out_path):
    """feature_process.py:191-195: descriptors [n,256] are stored transposed ([256,n])."""
    to_np = lambda a: a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
    np.savez(out_path, keypoints3d=to_np(xyzs), descriptors3d=to_np(descriptors).transpose(1, 0), scores3d=to_np(scores))


# ----------------------------------------------------------------------------------------------- device-resident handle
class ObjectDB:
    """One object's matching DB resident on a GPU: what inference_core builds at inference.py:113-130."""

    def __init__(self, keypoints3d, avg_descriptors3d, leaves_pm, num_leaf, avg_scores=None, leaf_scores=None):
        self.keypoints3d = keypoints3d                  # [N3,3] fp32
        self.avg_pm = avg_descriptors3d                 # [N3,256] point-major
        self.leaves_pm = leaves_pm                      # [N3*L,256] point-major, leaf j of point n at row n*L+j
        self.num_leaf = int(num_leaf)
        self.avg_scores, self.leaf_scores = avg_scores, leaf_scores
        # the reference-format (channel-major) tensors are views: no second copy of the 64 MiB leaf tensor
        self.descriptors3d_db = self.avg_pm.t()
        self.descriptors2d_db = self.leaves_pm.t()

    @property
    def num_3d(self):
        return self.keypoints3d.shape[0]

    @classmethod
    def from_arrays(cls, keypoints3d, avg_descriptors3d, avg_scores3d, clt_descriptors3d, clt_scores3d, idxs, num_leaf,
                    device=None, rng=np.random):
        """Arrays in the on-disk layout (descriptors [256,n]); the order of operations (and of np.random draws) is the
        reference's: pad the averages, then build the leaves (inference.py:119-130)."""
        dev = _device(device)
        kp = _as_cuda_f32(keypoints3d, dev)
        num_3d = kp.shape[0]
        avg, avg_s = pad_features3d_random(avg_descriptors3d, avg_scores3d, num_3d, dev)
        _, leaf_s, leaves_pm = build_features3d_leaves(clt_descriptors3d, clt_scores3d, idxs, num_3d, num_leaf, dev, rng,
                                                       return_point_major=True)
        return cls(kp, avg.t().contiguous(), leaves_pm, num_leaf, avg_s, leaf_s)

    @classmethod
    def load(cls, avg_anno_3d_path, clt_anno_3d_path, idxs_path, num_leaf, device=None, rng=np.random):
        """inference.py:113-130 from the three files of an SfM model directory."""
        avg_data, clt_data, idxs = np.load(avg_anno_3d_path), np.load(clt_anno_3d_path), np.load(idxs_path)
        return cls.from_arrays(clt_data["keypoints3d"], avg_data["descriptors3d"], avg_data["scores3d"],
                               clt_data["descriptors3d"], clt_data["scores3d"], idxs, num_leaf, device, rng)

    def inputs(self, batch=1):
        """The DB half of the matcher's input dict (GATs_SuperGlue.py:209-218), stride-0 views over the batch."""
        e = lambda t: t[None].expand(batch, *t.shape)
        return {"keypoints3d": e(self.keypoints3d), "descriptors3d_db": e(self.descriptors3d_db),
                "descriptors2d_db": e(self.descriptors2d_db)}

    def register(self, model):
        """Hand the point-major tensors to a ``GATsSuperGlue`` so that ``forward`` on ``self.inputs(B)`` finds its static-DB
        cache filled (no transposition pass at all)."""
        matcher = getattr(model, "matcher", model)
        matcher.register_static_db(self.descriptors3d_db, self.descriptors2d_db, self.avg_pm, self.leaves_pm, self.num_leaf)
        return self
