"""Per-sequence driver loop on the device + the cm-degree evaluator  (SURVEY.md 8(f) row f1).

Mirrors ``inference_core`` (inference.py:96-172) for the part that is this repo's path -- frames in, poses and the
``Evaluator`` summary out -- and ``Evaluator`` itself (src/evaluators/cmd_evaluator.py:3-61) plus ``query_pose_error``
(src/utils/eval_utils.py:45-63).  What changes against the reference loop:

* the object DB is built once and stays resident (``onepose_b200.objectdb.ObjectDB``), instead of the per-frame
  ``pack_data`` upload of 72 MiB (inference.py:80-94,144-145);
* matches -> gather -> PnP -> pose errors stay on the GPU for a whole block of frames (the reference crosses to the host
  twice per frame, inference.py:147-155); one device->host copy of [F] booleans at ``summarize()``.

``Evaluator`` keeps the reference's per-pose numpy interface (``evaluate(pose_pred, pose_gt)``, ``summarize()``) and adds
``evaluate_batch`` for poses that are CUDA tensors.
"""
import numpy as np
import torch

from ._lib import call, ptr, stream_ptr


def pose_errors(pose_pred, pose_gt):
    """[F,3,4] (or [F,4,4]) float64 CUDA tensors -> (t_err_cm [F], r_err_deg [F]) on the device (op_pose_errors)."""
    p = pose_pred[:, :3, :4].to(torch.float64).contiguous()
    g = pose_gt[:, :3, :4].to(device=p.device, dtype=torch.float64).contiguous()
    if not p.is_cuda:
        raise RuntimeError("pose_errors: CUDA tensors only (use query_pose_error for a single numpy pose)")
    F = p.shape[0]
    t = torch.empty(F, dtype=torch.float64, device=p.device)
    r = torch.empty(F, dtype=torch.float64, device=p.device)
    with torch.cuda.device(p.device):
        call("op_pose_errors", ptr(p), ptr(g), F, ptr(t), ptr(r), stream_ptr())
    return t, r


def query_pose_error(pose_pred, pose_gt):
    """eval_utils.py:45-63 -- (angular_distance [deg], translation_distance [cm]) of one numpy pose pair."""
    pose_pred, pose_gt = np.asarray(pose_pred), np.asarray(pose_gt)
    if pose_pred.shape[0] == 4:
        pose_pred = pose_pred[:3]
    if pose_gt.shape[0] == 4:
        pose_gt = pose_gt[:3]
    translation_distance = np.linalg.norm(pose_pred[:, 3] - pose_gt[:, 3]) * 100
    rotation_diff = np.dot(pose_pred[:, :3], pose_gt[:, :3].T)
    trace = np.trace(rotation_diff)
    trace = trace if trace <= 3 else 3
    angular_distance = np.rad2deg(np.arccos((trace - 1.0) / 2.0))
    return angular_distance, translation_distance


class Evaluator:
    """cmd_evaluator.py:3-61: 1 cm-1 deg, 3 cm-3 deg, 5 cm-5 deg success rates over the poses seen since the last summary."""

    def __init__(self):
        self.cmd1, self.cmd3, self.cmd5, self.cmd7, self.add = [], [], [], [], []
        self._pending = []           # device-side [F,3] bool blocks from evaluate_batch

    def evaluate(self, pose_pred, pose_gt):
        if pose_pred is None:                                            # cmd_evaluator.py:37-41
            for lst in (self.cmd5, self.cmd1, self.cmd3, self.cmd7):
                lst.append(False)
            return
        r, t = query_pose_error(pose_pred, pose_gt)
        self.cmd1.append(bool(t < 1 and r < 1))
        self.cmd3.append(bool(t < 3 and r < 3))
        self.cmd5.append(bool(t < 5 and r < 5))

    def evaluate_batch(self, poses_pred, poses_gt):
        """poses [F,3,4] / [F,4,4] CUDA tensors; nothing is copied to the host until ``summarize``."""
        t, r = pose_errors(poses_pred, poses_gt)
        self._pending.append(torch.stack([(t < 1) & (r < 1), (t < 3) & (r < 3), (t < 5) & (r < 5)], 1))

    def summarize(self, verbose=True):
        if self._pending:
            ok = torch.cat(self._pending).cpu().numpy()
            self.cmd1 += ok[:, 0].tolist(); self.cmd3 += ok[:, 1].tolist(); self.cmd5 += ok[:, 2].tolist()
            self._pending = []
        cmd1, cmd3, cmd5 = np.mean(self.cmd1), np.mean(self.cmd3), np.mean(self.cmd5)
        if verbose:
            print("1 cm 1 degree metric: {}".format(cmd1))
            print("3 cm 3 degree metric: {}".format(cmd3))
            print("5 cm 5 degree metric: {}".format(cmd5))
        self.cmd1, self.cmd3, self.cmd5, self.cmd7 = [], [], [], []
        return {"cmd1": cmd1, "cmd3": cmd3, "cmd5": cmd5}


_COPY_STREAMS = {}


def _copy_stream(dev):
    dev = torch.device(dev)
    st = _COPY_STREAMS.get(dev.index)
    if st is None:
        st = _COPY_STREAMS[dev.index] = torch.cuda.Stream(device=dev)
    return st


class _PinnedStaging:
    """Two alternating sets of pinned host buffers: pageable inputs are staged through them so that the uploads of run_sequence stay
    asynchronous (a non_blocking copy from pageable memory silently serialises).  A set is reused only after the uploads that read it
    have completed (its event)."""

    def __init__(self):
        self.slots = [dict(bufs={}, event=None), dict(bufs={}, event=None)]
        self.turn = 0

    def begin(self):
        slot = self.slots[self.turn]
        self.turn ^= 1
        if slot["event"] is not None:
            slot["event"].synchronize()
        return slot

    @staticmethod
    def stage(slot, name, t):
        t = torch.as_tensor(t)
        if t.is_cuda or t.is_pinned():
            return t
        buf = slot["bufs"].get(name)
        if buf is None or buf.shape != t.shape or buf.dtype != t.dtype:
            buf = slot["bufs"][name] = torch.empty(t.shape, dtype=t.dtype).pin_memory()
        buf.copy_(t)
        return buf


@torch.no_grad()
def run_sequence(model, db, keypoints2d, descriptors2d, K, poses_gt=None, batch=64, scale=1000.0, evaluator=None, counts2d=None):
    """The loop of inference_core (inference.py:132-160) for one object and F query frames, in blocks of ``batch`` frames.

    keypoints2d [F,N2,2], descriptors2d [F,256,N2] (the extractor's outputs, any float dtype, host or device), K [3,3] or
    [F,3,3], poses_gt [F,3,4] / [F,4,4] or None.  Returns dict(pose [F,3,4] float64, pose_homo [F,4,4], n_matches [F],
    n_inliers [F]) as CUDA tensors and, with ground truth, 'eval' = Evaluator.summarize().
    counts2d [F] (optional): per-frame key-point counts of a padded block (``data['count2d']`` of GATsSuperGlue.forward)."""
    matcher = getattr(model, "matcher", model)
    dev = db.keypoints3d.device
    db.register(matcher)
    F = keypoints2d.shape[0]
    if F == 0:                                                            # an empty sequence: nothing to match, nothing to evaluate
        res = dict(pose=torch.zeros(0, 3, 4, dtype=torch.float64, device=dev), n_matches=torch.zeros(0, dtype=torch.int32, device=dev),
                   n_inliers=torch.zeros(0, dtype=torch.int32, device=dev), pose_homo=torch.zeros(0, 4, 4, dtype=torch.float64, device=dev))
        if poses_gt is not None:
            res["eval"] = {}
        return res
    K = torch.as_tensor(K, dtype=torch.float64)
    K = (K[None].expand(F, 3, 3) if K.dim() == 2 else K).to(dev)
    evaluator = evaluator if evaluator is not None else Evaluator()
    out = {k: [] for k in ("pose", "n_matches", "n_inliers")}
    # A three-stage pipeline over the blocks: block i+1 is uploaded on a copy stream while block i is matched on the caller's stream,
    # and gather + PnP + pose errors of block i run on the matcher's pose stream (GATsSuperGlue.pnp_overlap) next to the matcher of
    # block i+1.  Pageable host inputs are staged through pinned buffers (pinned or device inputs are used as they are).
    staging = _PinnedStaging()
    cnt_dev = None if counts2d is None else torch.as_tensor(counts2d).to(device=dev, dtype=torch.int32)
    main = torch.cuda.current_stream(dev)
    copy = _copy_stream(dev)
    overlap_was = getattr(matcher, "pnp_overlap", None)
    if overlap_was is not None:
        matcher.pnp_overlap = True

    def upload(f0):
        f1 = min(F, f0 + batch)
        slot = staging.begin()
        with torch.cuda.stream(copy):
            kp = staging.stage(slot, "kp", keypoints2d[f0:f1]).to(dev, non_blocking=True)
            ds = staging.stage(slot, "ds", descriptors2d[f0:f1]).to(dev, non_blocking=True)
            gt = None if poses_gt is None else staging.stage(slot, "gt", poses_gt[f0:f1]).to(dev, non_blocking=True)
            slot["event"] = copy.record_event()
            return f0, f1, kp, ds, gt, slot["event"]

    copy.wait_stream(main)
    nxt = upload(0) if F > 0 else None
    preds = []
    try:
        while nxt is not None:
            f0, f1, kp, ds, gt, ready = nxt
            main.wait_event(ready)
            for t in (kp, ds):
                t.record_stream(main)
            nxt = upload(f1) if f1 < F else None
            data = db.inputs(f1 - f0)
            data["keypoints2d"], data["descriptors2d_query"] = kp, ds
            data["query_intrinsic"] = K[f0:f1]
            data["pnp_scale"] = scale
            if cnt_dev is not None:
                data["count2d"] = cnt_dev[f0:f1]
            pred, _ = matcher(data)
            ev = pred.get("pose_event")
            side = matcher.pose_stream(dev) if ev is not None else main
            if getattr(matcher, "_graphs_on", False):
                # CUDA-graph serving mode: the tensors of pred are the graph's static outputs, reused by the second next call.  The
                # block's results are copied out behind its PnP, and the event the matcher waits for before it reuses the slot is moved
                # behind the copies (pred["pose_event"] IS the slot's event object).
                with torch.cuda.stream(side):
                    pred = dict(pred, **{k: pred[k].clone() for k in out})
                    if ev is not None:
                        ev.record(side)
            preds.append(pred)
            if gt is not None:
                with torch.cuda.stream(side):                           # behind this block's PnP
                    gt.record_stream(side)
                    evaluator.evaluate_batch(pred["pose"], gt)
    finally:
        if overlap_was is not None:
            matcher.pnp_overlap = overlap_was
    if preds and preds[-1].get("pose_event") is not None:               # everything on the pose stream is ordered before what follows
        main.wait_stream(matcher.pose_stream(dev))
    for pred in preds:
        for k in out:
            out[k].append(pred[k])
    res = {k: torch.cat(v) for k, v in out.items()}
    homo = torch.zeros(F, 4, 4, dtype=torch.float64, device=dev)
    homo[:, :3] = res["pose"]; homo[:, 3, 3] = 1.0                        # eval_utils.py:35-36
    res["pose_homo"] = homo
    if poses_gt is not None:
        res["eval"] = evaluator.summarize(verbose=False)
    return res
