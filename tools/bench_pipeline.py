"""Images -> poses on one GPU: SuperPoint front-end (row f3) feeding the matcher + PnP (rows a-d) without leaving the device.

    python tools/bench_pipeline.py [--frames 64] [--iters 5]

64 synthetic 512x512 frames per step (pinned host uint8 images uploaded inside the timed region), max_keypoints = 2048 = the headline
N2d, object DB of the headline shape resident.  Random-initialised extractor weights: the key-points are real NMS maxima but carry no
object signal, so few matches survive and PnP is cheaper than in bench.py (noted in the output); the extractor and matcher times
are the headline ones."""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from onepose_b200 import synth  # noqa: E402
from onepose_b200.matcher import GATsSuperGlue  # noqa: E402
from onepose_b200.superpoint import SuperPoint  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=64)
    ap.add_argument("--iters", type=int, default=5)
    a = ap.parse_args()
    F, N2, N3, L = a.frames, 2048, 8192, 8
    dev = torch.device("cuda")
    torch.manual_seed(0)
    sp = SuperPoint({'descriptor_dim': 256, 'nms_radius': 3, 'max_keypoints': N2, 'keypoints_threshold': 0.6}).eval().to(dev)
    matcher = GATsSuperGlue(dict(synth.DEFAULT_HPARAMS)).to(dev).eval()
    matcher.load_state_dict(synth.make_weights(0, damped=True), strict=True)
    obj = synth.make_object(0, N3, L)
    e = lambda x: torch.as_tensor(x).to(dev)[None].expand(F, *np.shape(x))
    db = {"keypoints3d": e(obj.keypoints3d), "descriptors3d_db": e(obj.descriptors3d_db), "descriptors2d_db": e(obj.descriptors2d_db)}
    K = torch.from_numpy(synth.make_frame(obj, 0, 64).K)[None].to(dev)
    rs = np.random.RandomState(0)
    imgs = rs.randint(0, 256, (F, 1, 512, 512), dtype=np.uint8)
    for f in range(F):                      # frames differ in how much of the image is textured, so the key-point counts differ
        imgs[f, :, (f % 4 + 1) * 128:] = 128
    img_pin = torch.from_numpy(imgs).pin_memory()
    pose_pin = torch.empty(F, 3, 4, dtype=torch.float64).pin_memory()
    caps = ((torch.arange(F, device=dev) % 4 + 1) * (N2 // 4)).to(torch.int32)
    ev = lambda: torch.cuda.Event(enable_timing=True)

    def step(marks=None):
        img = img_pin.to(dev, non_blocking=True).float() / 255.0                  # NormalizedDataset: grayscale / 255
        if marks is not None: marks.append(ev()); marks[-1].record()
        out = sp.forward_batched(img)
        if marks is not None: marks.append(ev()); marks[-1].record()
        # the front-end's padded block goes to the matcher as it is: count2d carries every frame's own key-point count (ragged batch).
        # A random-weight extractor saturates max_keypoints on any image, so the counts are capped per frame (512 .. 2048) to make the
        # block as ragged as real frames are
        out["count"] = torch.minimum(out["count"], caps)
        data = dict(db, keypoints2d=out["keypoints"], descriptors2d_query=out["descriptors_pm"].transpose(1, 2), query_intrinsic=K,
                    count2d=out["count"])
        pred, _ = matcher(data)
        pose_pin.copy_(pred["pose"], non_blocking=True)
        if marks is not None: marks.append(ev()); marks[-1].record()
        return out, pred

    for _ in range(3):
        out, pred = step()
    torch.cuda.synchronize()
    s, t = ev(), ev()
    s.record()
    for _ in range(a.iters):
        step()
    t.record(); torch.cuda.synchronize()
    ms = s.elapsed_time(t) / a.iters
    marks = [ev()]; marks[0].record()
    step(marks); torch.cuda.synchronize()
    parts = [marks[i].elapsed_time(marks[i + 1]) for i in range(3)]
    print(json.dumps({"frames": F, "ms_per_step": round(ms, 3), "frames_per_s": round(F / ms * 1e3, 1),
                      "upload_ms": round(parts[0], 3), "superpoint_ms": round(parts[1], 3), "matcher_pnp_ms": round(parts[2], 3),
                      "keypoints_per_frame": float(out["count"].float().mean()), "keypoints_min": int(out["count"].min()), "keypoints_max": int(out["count"].max()),
                      "mean_matches": float(pred["n_matches"].float().mean()),
                      "h2d_bytes_per_step": img_pin.numel(), "note": "random extractor weights: matches carry no object signal, PnP is cheap"}))


if __name__ == "__main__":
    main()
