"""GPU parity tests of the object-DB builder, the registered-DB fast path and the device-side driver loop / evaluator
 against the reference-generated fixture and against the plain matcher path."""
import os

import numpy as np
import pytest
import torch

from onepose_b200 import synth
from oracle import objectdb_oracle as orc
from tests import helpers

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(helpers.GOLDEN, "db_builder.npz"))
CASES = {"pad": (14, 4), "trunc": (9, 4), "wide": (12, 16)}


def test_segment_means_bit_exact():
    from onepose_b200 import objectdb
    inp = orc.db_builder_inputs()
    avg = objectdb.mean_descriptors(inp["desc"], inp["idxs"])
    assert np.array_equal(avg.cpu().numpy(), G["avg"])
    assert np.array_equal(objectdb.mean_scores(inp["scores"], inp["idxs"]).cpu().numpy(), G["avg_scores"])


@pytest.mark.parametrize("tag", list(CASES))
def test_leaf_builder_bit_exact(tag):
    from onepose_b200 import objectdb
    n_target, num_leaf = CASES[tag]
    inp = orc.db_builder_inputs()
    np.random.seed(5)
    d, s = objectdb.pad_features3d_random(G["avg"].T, G["avg_scores"], n_target)
    dl, sl, pm = objectdb.build_features3d_leaves(inp["desc"].T, inp["scores"], inp["idxs"], n_target, num_leaf, return_point_major=True)
    assert np.array_equal(d.cpu().numpy(), G[f"{tag}_avg"]) and np.array_equal(s.cpu().numpy(), G[f"{tag}_avg_scores"])
    assert np.array_equal(dl.cpu().numpy(), G[f"{tag}_leaves"]) and np.array_equal(sl.cpu().numpy(), G[f"{tag}_leaf_scores"])
    assert pm.is_contiguous() and dl.data_ptr() == pm.data_ptr()               # the channel-major result is a view


def _object_files(tmp_path, obj, n_leaves_max):
    """An SfM model directory in the reference's on-disk format from a synthetic object: every point gets 1..n_leaves_max
    real leaves (its first ones), averages are the means of those."""
    from onepose_b200 import objectdb
    rs = np.random.RandomState(3)
    N3, L = obj.keypoints3d.shape[0], obj.num_leaf
    leaves = obj.descriptors2d_db.T.reshape(N3, L, 256)
    extra = rs.standard_normal((N3, max(0, n_leaves_max - L), 256)).astype(np.float32) / 16      # more leaves than num_leaf for some points
    leaves = np.concatenate([leaves, extra], 1)
    idxs = rs.randint(1, n_leaves_max + 1, size=N3)
    clt = np.concatenate([leaves[n, :idxs[n]] for n in range(N3)]).astype(np.float32)            # [sum, 256]
    clt_scores = rs.rand(clt.shape[0], 1).astype(np.float32)
    avg = objectdb.mean_descriptors(clt, idxs).cpu().numpy()
    avg_scores = objectdb.mean_scores(clt_scores, idxs).cpu().numpy()
    p = {k: str(tmp_path / k) for k in ("anno_3d_average.npz", "anno_3d_collect.npz", "idxs.npy")}
    objectdb.save_3d_anno(obj.keypoints3d, avg, avg_scores, p["anno_3d_average.npz"])
    objectdb.save_3d_anno(obj.keypoints3d, clt, clt_scores, p["anno_3d_collect.npz"])
    np.save(p["idxs.npy"], idxs)
    return p, idxs, clt, clt_scores, avg, avg_scores


def test_object_db_from_disk_and_registered_fast_path(tmp_path):
    """ObjectDB.load == the reference's load sequence (inference.py:113-130) and a registered DB gives the same bits as the
    same tensors handed over as plain contiguous channel-major inputs."""
    from onepose_b200 import objectdb
    from onepose_b200.matcher import GATsSuperGlue
    obj = synth.make_object(7, 150, 6)
    p, idxs, clt, clt_scores, avg, avg_scores = _object_files(tmp_path, obj, 9)
    np.random.seed(21)
    db = objectdb.ObjectDB.load(p["anno_3d_average.npz"], p["anno_3d_collect.npz"], p["idxs.npy"], num_leaf=6, device="cuda")
    np.random.seed(21)
    d_ref, _ = orc.pad_features3d_random(avg.T, avg_scores, 150)
    l_ref, ls_ref = orc.build_features3d_leaves(clt.T, clt_scores, idxs, 150, 6)
    assert np.array_equal(db.descriptors3d_db.cpu().numpy(), d_ref) and np.array_equal(db.descriptors2d_db.cpu().numpy(), l_ref)
    assert np.array_equal(db.leaf_scores.cpu().numpy(), ls_ref)
    frames = [synth.make_frame(obj, f, 100) for f in range(3)]
    q = synth.pack_batch(obj, frames)
    m = GATsSuperGlue(dict(synth.DEFAULT_HPARAMS)).cuda().eval()
    m.load_state_dict(synth.make_weights(0, damped=True), strict=True)
    plain = {"keypoints2d": q["keypoints2d"].cuda(), "descriptors2d_query": q["descriptors2d_query"].cuda(),
             "keypoints3d": db.keypoints3d[None].expand(3, -1, -1),
             "descriptors3d_db": db.descriptors3d_db.contiguous()[None].expand(3, -1, -1),
             "descriptors2d_db": db.descriptors2d_db.contiguous()[None].expand(3, -1, -1)}
    _, conf_plain = m(plain)
    conf_plain = conf_plain.clone()
    db.register(m)
    fast = dict(db.inputs(3), keypoints2d=plain["keypoints2d"], descriptors2d_query=plain["descriptors2d_query"])
    _, conf_fast = m(fast)
    assert torch.equal(conf_plain, conf_fast)
    assert m._db_cache["registered"][1]["leaves"].data_ptr() == db.leaves_pm.data_ptr()        # no copy was made


def test_run_sequence_and_device_evaluator():
    """The driver loop over blocks of frames == frame-by-frame calls of the module + ransac_PnP + the numpy Evaluator."""
    from onepose_b200 import objectdb
    from onepose_b200.evaluator import Evaluator, pose_errors, run_sequence
    from onepose_b200.matcher import GATsSuperGlue
    from onepose_b200.pnp import ransac_PnP
    obj = synth.make_object(1, 400, 8)
    frames = [synth.make_frame(obj, f, 256) for f in range(5)]
    q = synth.pack_batch(obj, frames)
    db = objectdb.ObjectDB(torch.from_numpy(obj.keypoints3d).cuda(), q["descriptors3d_db"][0].t().contiguous().cuda(),
                           q["descriptors2d_db"][0].t().contiguous().cuda(), 8)
    m = GATsSuperGlue(dict(synth.DEFAULT_HPARAMS)).cuda().eval()
    m.load_state_dict(synth.make_weights(0, damped=True), strict=True)
    gt = np.stack([f.pose_gt for f in frames])
    res = run_sequence(m, db, q["keypoints2d"], q["descriptors2d_query"], frames[0].K, poses_gt=gt, batch=2)
    assert res["pose"].shape == (5, 3, 4) and res["pose_homo"].shape == (5, 4, 4)
    ev = Evaluator()
    for f in range(5):                                                          # the reference's loop, inference.py:144-160
        data = dict(db.inputs(1), keypoints2d=q["keypoints2d"][f:f + 1].cuda(), descriptors2d_query=q["descriptors2d_query"][f:f + 1].cuda())
        pred, _ = m(data)
        matches = pred["matches0"].cpu().numpy(); valid = matches > -1
        mk2, mk3 = q["keypoints2d"][f].numpy()[valid], obj.keypoints3d[matches[valid]]
        pose, pose_homo, inl = ransac_PnP(frames[f].K, mk2, mk3, scale=1000)
        np.testing.assert_allclose(res["pose"][f].cpu().numpy(), pose, rtol=0, atol=1e-9)
        assert int(res["n_matches"][f]) == int(valid.sum()) and int(res["n_inliers"][f]) == len(inl)
        ev.evaluate(pose_homo, gt[f])
    assert res["eval"] == ev.summarize(verbose=False)
    # the same loop in the CUDA-graph serving mode (static outputs copied out per block): same poses
    m.enable_graphs(True)
    res_g = run_sequence(m, db, q["keypoints2d"], q["descriptors2d_query"], frames[0].K, poses_gt=gt, batch=1)
    m.enable_graphs(False)
    assert torch.equal(res_g["pose"], run_sequence(m, db, q["keypoints2d"], q["descriptors2d_query"], frames[0].K, batch=1)["pose"])
    assert torch.equal(res_g["n_matches"], res["n_matches"]) and res_g["eval"] == res["eval"]
    np.testing.assert_allclose(res_g["pose"].cpu().numpy(), res["pose"].cpu().numpy(), rtol=0, atol=1e-9)
    # an empty sequence returns empty results; a padded block with per-frame counts equals the truncated frames
    empty = run_sequence(m, db, q["keypoints2d"][:0], q["descriptors2d_query"][:0], frames[0].K)
    assert empty["pose"].shape == (0, 3, 4) and empty["n_matches"].numel() == 0
    cnt = torch.tensor([256, 200, 131, 256, 64], dtype=torch.int32)
    res_c = run_sequence(m, db, q["keypoints2d"], q["descriptors2d_query"], frames[0].K, batch=3, counts2d=cnt)
    for f in range(5):
        c = int(cnt[f])
        data = dict(db.inputs(1), keypoints2d=q["keypoints2d"][f:f + 1, :c].cuda(), descriptors2d_query=q["descriptors2d_query"][f:f + 1, :, :c].cuda(),
                    query_intrinsic=torch.from_numpy(frames[f].K)[None])
        pred, _ = m(data)
        assert int(res_c["n_matches"][f]) == int(pred["n_matches"][0])
        np.testing.assert_allclose(res_c["pose"][f].cpu().numpy(), pred["pose"][0].cpu().numpy(), rtol=0, atol=1e-6)
    # device-side errors against the reference arithmetic on the fixture poses
    inp = orc.db_builder_inputs()
    t, r = pose_errors(torch.from_numpy(inp["poses_pred"]).cuda(), torch.from_numpy(inp["poses_gt"]).cuda())
    np.testing.assert_allclose(torch.stack([r, t], 1).cpu().numpy(), G["pose_errs"], rtol=1e-9, atol=1e-9)
    ev2 = Evaluator()
    ev2.evaluate_batch(torch.from_numpy(inp["poses_pred"]).cuda(), torch.from_numpy(inp["poses_gt"]).cuda())
    s = ev2.summarize(verbose=False)
    assert [s["cmd1"], s["cmd3"], s["cmd5"]] == G["cmd"].tolist()
