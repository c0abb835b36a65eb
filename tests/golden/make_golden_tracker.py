"""Golden fixture for SURVEY.md 8(f) row f4 from the UNMODIFIED reference functions, run in the build container.

    python tests/golden/make_golden_tracker.py          (needs /root/reference, cv2; CPU only)

tracker.npz: outputs of src/tracker/tracking_utils.py AngleAxisRotatePoint / SnavelyReprojectionErrorV2 / project and of
BATracker.apply_triangulation (src/tracker/ba_tracker.py:267-273; the module imports DeepLM, absent here, so the method's own
source is exec'd unmodified) on the seeded arrays of ``oracle.tracker_oracle.tracker_inputs`` (only outputs are stored).
"""
import ast
import os
import sys

sys.dont_write_bytecode = True
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference")
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import cv2
import numpy as np
import torch

from oracle import tracker_oracle as tro


def reference_method(path, cls, name):
    src = open(path).read()
    for node in ast.walk(ast.parse(src)):
        if isinstance(node, ast.ClassDef) and node.name == cls:
            for fn in node.body:
                if isinstance(fn, ast.FunctionDef) and fn.name == name:
                    ns = {"np": np, "cv2": cv2}
                    exec(compile(ast.Module([fn], []), path, "exec"), ns)
                    return ns[name]
    raise KeyError(name)


def main():
    import src.tracker.tracking_utils as tu
    out = {}
    for seed in (0, 1):
        d = tro.tracker_inputs(seed)
        pts = torch.tensor(d["points0"][d["pt_idx"]]); cams = torch.tensor(d["cams0"][d["cam_idx"], :6]); feats = torch.tensor(d["features"])
        out[f"rot_{seed}"] = tu.AngleAxisRotatePoint(cams[:, :3], pts).numpy()
        out[f"res_{seed}"] = tu.SnavelyReprojectionErrorV2(pts, cams, feats).numpy()
        out[f"res3d_{seed}"] = tu.SnavelyReprojectionErrorV2(pts[:, None], cams[:, None], feats).numpy()
    d = tro.tracker_inputs(0)
    K = np.array([[600.0, 0, 256], [0, 600, 256], [0, 0, 1]])
    RT = np.concatenate([cv2.Rodrigues(d["cams_gt"][1, :3])[0], d["cams_gt"][1, 3:6, None]], 1)
    xy, depth = tu.project(d["points_gt"].copy(), K, RT, need_depth=True)
    out["proj_xy"], out["proj_depth"] = xy, depth
    tri = reference_method("/root/reference/src/tracker/ba_tracker.py", "BATracker", "apply_triangulation")
    Ks, Ts, xs = [], [], []
    for c in (1, 2):
        Kc = np.array([[d["cams_gt"][c, 6], 0, d["cams_gt"][c, 7]], [0, d["cams_gt"][c, 6], d["cams_gt"][c, 8]], [0, 0, 1]])
        Tcw = np.eye(4); Tcw[:3, :3] = cv2.Rodrigues(d["cams_gt"][c, :3])[0]; Tcw[:3, 3] = d["cams_gt"][c, 3:6]
        Twc = np.linalg.inv(Tcw)                              # the reference inverts its argument (:268-269)
        Ks.append(Kc); Ts.append(Twc)
        xs.append(tu.project(d["points_gt"].copy(), Kc, Tcw[:3]) + np.random.RandomState(c).normal(0, 0.3, (len(d["points_gt"]), 2)))
    out["tri_X"] = tri(None, Ks[0], Ks[1], Ts[0], Ts[1], xs[0], xs[1])
    out["tri_x1"], out["tri_x2"] = xs
    np.savez_compressed(os.path.join(HERE, "tracker.npz"), **out)
    print("wrote tracker.npz", os.path.getsize(os.path.join(HERE, "tracker.npz")))


if __name__ == "__main__":
    main()
