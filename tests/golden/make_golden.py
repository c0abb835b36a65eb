"""Generate the golden fixtures by running the UNMODIFIED reference in the build container.

    python tests/golden/make_golden.py        (needs /root/reference and cv2; CPU only)

* matcher_*.npz  : outputs of the reference ``GATsSuperGlue.forward``
  (/root/reference/src/models/GATsSPG_architectures/GATs_SuperGlue.py:203-278) on the seeded inputs
  of ``onepose_b200.synth`` with weights ``synth.make_weights(seed, damped)`` loaded through
  ``load_state_dict(strict=True)``.  Only seeds + outputs are stored.
* layer_*.npz    : outputs of the reference sub-modules ``GraphAttentionLayer`` (GATs.py:62-104)
  and ``AttentionPropagation`` (GATs_SuperGlue.py:123-132) on seeded inputs.
* pnp_*.npz      : outputs of the reference ``ransac_PnP`` (src/utils/eval_utils.py:18-42, i.e.
  cv2.solvePnPRansac of this image, cv2 4.13.0) on ``synth.make_pnp_problem`` inputs, plus
  per-sample ``cv2.solvePnP(EPNP)`` poses on all-inlier 5-point samples.
* db_builder.npz : outputs of the reference's DB-side helpers (src/utils/data_utils.py:143-205 pad_features3d_random /
  build_features3d_leaves under np.random.seed(5), src/sfm/postprocess/feature_process.py:297-319 mean_descriptors --
  the module imports h5py, absent here, so the function's own source is exec'd unmodified), of
  src/evaluators/cmd_evaluator.py Evaluator and of eval_utils.query_pose_error on ``oracle.objectdb_oracle.db_builder_inputs``.
The reference has no tests or golden vectors of its own; these fixtures, made from
the live reference, are what pins the oracle.
"""
import io
import contextlib
import os
import sys

sys.dont_write_bytecode = True
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference")
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import numpy as np
import torch

from onepose_b200 import synth

MATCHER_CASES = {
    # name: (weights seed, damped, object id, n2, n3, leaves, frames, hparam overrides)
    "matcher_random": (0, False, 0, 96, 160, 8, 2, {}),
    "matcher_damped": (1, True, 1, 128, 192, 8, 2, {}),
    "matcher_ragged": (2, True, 2, 33, 47, 3, 1, {}),
    "matcher_lowthr": (0, False, 3, 64, 96, 4, 1, {"match_threshold": 0.0}),
    "matcher_flags_lt": (3, True, 4, 40, 56, 4, 1, {"with_linear_transform": True}),
    "matcher_flags_noself": (4, True, 5, 40, 56, 4, 1, {"include_self": False}),
    "matcher_flags_add": (5, True, 6, 40, 56, 4, 1, {"additional": True, "with_linear_transform": True}),
}

PNP_CASES = {
    # name: (seed, ncorr, inlier_ratio, scale)
    "pnp_300": (0, 300, 0.8, 1000.0),
    "pnp_1000": (1, 1000, 0.5, 1000.0),
    "pnp_2000": (2, 2000, 0.3, 1.0),
    "pnp_60": (3, 60, 0.9, 1000.0),
}


def matcher_inputs(case):
    wseed, damped, oid, n2, n3, leaves, frames, over = MATCHER_CASES[case]
    hp = dict(synth.DEFAULT_HPARAMS); hp.update(over)
    w = synth.make_weights(wseed, damped=damped)
    obj = synth.make_object(oid, n3, leaves)
    fr = [synth.make_frame(obj, f, n2) for f in range(frames)]
    return hp, w, obj, fr, synth.pack_batch(obj, fr)


def all_outlier_problem(n, seed=0):
    """Random 2D points against random 3D points: no pose explains them."""
    rng = np.random.default_rng(seed)
    return rng.uniform(0, 512, (n, 2)), rng.uniform(-0.1, 0.1, (n, 3))


def make_pnp_edge(ransac_PnP):
    """Edge cases of the wrapper (SURVEY.md appendix B caveats), all through the unmodified reference ``ransac_PnP``:
    n = 0, 3 (cv2.error -> identity), n = 4 (cv2's P3P branch; eight more seeded problems in ``p3p_*``), n = 5 (direct EPnP),
    all-outlier inputs of 6 and 50 points (cv2 reports failure: ``inliers == []``; the pose the wrapper then builds comes from
    the rvec/tvec Mats that cv2 allocated but never wrote: three calls on the same input return different poses, so only the
    inlier list is stored), and a
    low-inlier problem on which the adaptive iteration bound stays above 4096 (cv2 runs to iteration 10000)."""
    edge = {}
    for n in (0, 3, 4, 5):
        pr = synth.make_pnp_problem(10 + n, max(n, 1), 1.0)
        with contextlib.redirect_stdout(io.StringIO()):
            pose, _, inl = ransac_PnP(pr["K"], pr["pts2d"][:n], pr["pts3d"][:n], scale=1000)
        edge[f"pose_{n}"] = pose; edge[f"inl_{n}"] = np.asarray(inl).ravel()
        print("edge n=%d inliers=%d" % (n, len(inl)))
    p3p = []
    for seed in range(8):
        pr = synth.make_pnp_problem(900 + seed, 4, 1.0, px_noise=0.5 if seed % 2 else 0.0)
        pose, _, inl = ransac_PnP(pr["K"], pr["pts2d"], pr["pts3d"], scale=1000)
        assert list(np.asarray(inl).ravel()) == [0, 1, 2, 3]
        p3p.append(pose)
    edge["p3p_poses"] = np.stack(p3p)
    for n in (6, 50):
        p2, p3 = all_outlier_problem(n)
        poses = []
        for _ in range(3):
            pose, _, inl = ransac_PnP(synth.INTRINSIC, p2, p3, scale=1000)
            poses.append(pose)
        edge[f"inl_out{n}"] = np.asarray(inl).ravel()          # the pose is not stored: it is not a function of the inputs
        varies = not all(np.array_equal(poses[0], q, equal_nan=True) for q in poses[1:])
        print("all-outlier n=%d inliers=%d identity=%s pose varies between calls=%s" % (n, len(inl), np.array_equal(poses[0], np.eye(4)[:3]), varies))
    pr = synth.make_pnp_problem(4, 400, 0.2)
    pose, _, inl = ransac_PnP(pr["K"], pr["pts2d"], pr["pts3d"], scale=1000)
    edge["pose_lowinl"] = pose; edge["inl_lowinl"] = np.asarray(inl).ravel()
    print("low-inlier 400 x 0.2: inliers", len(inl))
    np.savez_compressed(os.path.join(HERE, "pnp_edge.npz"), **edge)


def main():
    from src.models.GATsSPG_architectures.GATs_SuperGlue import GATsSuperGlue, AttentionPropagation
    from src.models.GATsSPG_architectures.GATs import GraphAttentionLayer
    from src.utils.eval_utils import ransac_PnP
    import cv2

    torch.set_grad_enabled(False)
    for case in MATCHER_CASES:
        hp, w, obj, fr, data = matcher_inputs(case)
        ref = GATsSuperGlue(hp).eval()
        ref.load_state_dict(w, strict=True)
        pred, conf = ref(data)
        np.savez_compressed(
            os.path.join(HERE, case + ".npz"), conf=conf.numpy(),
            matches0=pred["matches0"].numpy(), matches1=pred["matches1"].numpy(),
            matching_scores0=pred["matching_scores0"].numpy(), matching_scores1=pred["matching_scores1"].numpy())
        print(case, "conf", tuple(conf.shape), "matches", int((pred["matches0"] > -1).sum()))

    # sub-module fixtures: one GATs layer (shipped flags) and one AttentionPropagation
    rng = np.random.default_rng(77)
    w = synth.make_weights(6)
    leaves = torch.from_numpy(rng.standard_normal((2, 24, 5, 256)).astype(np.float32))
    h3 = torch.from_numpy(rng.standard_normal((2, 24, 256)).astype(np.float32))
    gat = GraphAttentionLayer(256, 256, 0.6, 0.2, True, True, False, False)
    gat.load_state_dict({"W": w["gnn.layers.0.W"], "a": w["gnn.layers.0.a"]})
    out_gat = gat(leaves.reshape(2, 24 * 5, 256), h3)
    x = torch.from_numpy(rng.standard_normal((2, 256, 40)).astype(np.float32))
    src = torch.from_numpy(rng.standard_normal((2, 256, 72)).astype(np.float32))
    ap = AttentionPropagation(256, 4)
    ap.load_state_dict({k[len("gnn.layers.1."):]: v for k, v in w.items() if k.startswith("gnn.layers.1.")})
    out_ap = ap(x, src)
    np.savez_compressed(os.path.join(HERE, "layer_units.npz"), gats_out=out_gat.numpy(), attn_out=out_ap.numpy())

    for case, (seed, n, ratio, scale) in PNP_CASES.items():
        pr = synth.make_pnp_problem(seed, n, ratio)
        pose, pose_h, inl = ransac_PnP(pr["K"], pr["pts2d"], pr["pts3d"], scale=scale)
        # per-sample EPnP poses on all-inlier minimal samples (well-posed => library-independent)
        X32 = (pr["pts3d"] * scale).astype(np.float32); uv32 = pr["pts2d"].astype(np.float32)
        ids = np.nonzero(pr["is_inlier"])[0]
        srng = np.random.default_rng(seed)
        samples = np.stack([srng.choice(ids, 5, replace=False) for _ in range(16)]).astype(np.int32)
        rts = []
        for s in samples:
            ok, rv, tv = cv2.solvePnP(X32[s], uv32[s], pr["K"], np.zeros((8, 1)), flags=cv2.SOLVEPNP_EPNP)
            rts.append(np.concatenate([rv.ravel(), tv.ravel()]))
        np.savez_compressed(os.path.join(HERE, case + ".npz"), pose=pose, inliers=np.asarray(inl).ravel(),
                            samples=samples, sample_rt=np.stack(rts))
        print(case, "inliers", len(inl))

    make_pnp_edge(ransac_PnP)

    # DB builder + evaluator
    import ast
    from src.utils import data_utils
    from src.utils.eval_utils import query_pose_error
    from src.evaluators.cmd_evaluator import Evaluator
    from oracle.objectdb_oracle import db_builder_inputs
    src_path = "/root/reference/src/sfm/postprocess/feature_process.py"
    tree = ast.parse(open(src_path).read())
    ns = {"np": np}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in ("mean_descriptors", "mean_scores"):
            exec(compile(ast.Module([node], []), src_path, "exec"), ns)
    inp = db_builder_inputs()
    out = {}
    out["avg"] = ns["mean_descriptors"](inp["desc"], inp["idxs"])
    out["avg_scores"] = ns["mean_scores"](inp["scores"], inp["idxs"])
    clt = inp["desc"].transpose(1, 0)                               # on-disk layout [256, n] (save_3d_anno)
    for tag, n_target, num_leaf in (("pad", 14, 4), ("trunc", 9, 4), ("wide", 12, 16)):
        np.random.seed(5)
        d, s_ = data_utils.pad_features3d_random(out["avg"].transpose(1, 0), out["avg_scores"], n_target)
        dl, sl = data_utils.build_features3d_leaves(clt, inp["scores"], inp["idxs"], n_target, num_leaf)
        out[f"{tag}_avg"] = d.numpy(); out[f"{tag}_avg_scores"] = s_.numpy()
        out[f"{tag}_leaves"] = dl.numpy(); out[f"{tag}_leaf_scores"] = sl.numpy()
    ev = Evaluator()
    errs = []
    for p_, g_ in zip(inp["poses_pred"], inp["poses_gt"]):
        ev.evaluate(p_, g_)
        errs.append(query_pose_error(p_, g_))
    with contextlib.redirect_stdout(io.StringIO()):
        summ = ev.summarize()
    out["pose_errs"] = np.asarray(errs); out["cmd"] = np.asarray([summ["cmd1"], summ["cmd3"], summ["cmd5"]])
    np.savez_compressed(os.path.join(HERE, "db_builder.npz"), **out)
    print("db_builder", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    if "pnp_edge" in sys.argv[1:]:
        from src.utils.eval_utils import ransac_PnP as _ref_ransac_PnP
        make_pnp_edge(_ref_ransac_PnP)
    else:
        main()
