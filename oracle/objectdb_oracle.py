"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the reference's object-DB builder and evaluator (SURVEY.md 8(f) rows
f1 / f2).  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module; the product
(onepose_b200/objectdb.py, evaluator.py) never does.

Pinned by tests/golden/db_builder.npz, produced by tests/golden/make_golden.py from the UNMODIFIED reference functions
(data_utils.pad_features3d_random / build_features3d_leaves, feature_process.mean_descriptors / mean_scores,
cmd_evaluator.Evaluator, eval_utils.query_pose_error).
"""
import numpy as np


def pad_features3d_random(descriptors, scores, n_target_shape):
    """src/utils/data_utils.py:143-160."""
    descriptors = np.asarray(descriptors, np.float32); scores = np.asarray(scores, np.float32)
    dim = descriptors.shape[0]
    n_pad = n_target_shape - descriptors.shape[1]
    if n_pad < 0:
        return descriptors[:, :n_target_shape], scores[:n_target_shape, :]
    return (np.concatenate([descriptors, np.ones((dim, n_pad), np.float32)], axis=-1),
            np.concatenate([scores, np.zeros((n_pad, 1), np.float32)], axis=0))


def build_features3d_leaves(descriptors, scores, idxs, n_target_shape, num_leaf, rng=np.random):
    """src/utils/data_utils.py:163-205 (np.random.permutation once per 3D point, in point order)."""
    descriptors = np.asarray(descriptors, np.float32); scores = np.asarray(scores, np.float32)
    dim = descriptors.shape[0]
    orig_num = idxs.shape[0]
    n_pad = n_target_shape - orig_num
    descriptors_dustbin = np.concatenate([descriptors, np.ones((dim, 1), np.float32)], axis=1)      # :176
    scores_dustbin = np.concatenate([scores, np.zeros((1, 1), np.float32)], axis=0)                # :177
    dustbin_id = descriptors_dustbin.shape[1] - 1
    upper_idxs = np.cumsum(idxs, axis=0)
    lower_idxs = np.insert(upper_idxs[:-1], 0, 0)
    aff = []
    for start, end in zip(lower_idxs, upper_idxs):                                                  # :183-192
        if num_leaf > end - start:
            ids = np.arange(start, end).tolist()
            ids += [dustbin_id] * (num_leaf - (end - start))
            aff.append(rng.permutation(np.array(ids)))
        else:
            aff.append(rng.permutation(np.arange(start, end))[:num_leaf])
    aff = np.concatenate(aff, axis=0)
    descriptors = descriptors_dustbin[:, aff]
    scores = scores_dustbin[aff, :]
    if n_pad < 0:                                                                                   # :199-204
        return descriptors[:, :num_leaf * n_target_shape], scores[:num_leaf * n_target_shape, :]
    return (np.concatenate([descriptors, np.ones((dim, n_pad * num_leaf), np.float32)], axis=-1),
            np.concatenate([scores, np.zeros((n_pad * num_leaf, 1), np.float32)], axis=0))


def mean_descriptors(descriptors, idxs):
    """src/sfm/postprocess/feature_process.py:297-305."""
    cumsum_idxs = np.cumsum(idxs)
    pre = np.insert(np.cumsum(idxs)[:-1], 0, 0)
    return np.concatenate([np.mean(descriptors[s:e], axis=0).reshape(1, -1) for s, e in zip(pre, cumsum_idxs)], axis=0)


def query_pose_error(pose_pred, pose_gt):
    """src/utils/eval_utils.py:45-63 -> (angular_distance deg, translation_distance cm)."""
    if pose_pred.shape[0] == 4:
        pose_pred = pose_pred[:3]
    if pose_gt.shape[0] == 4:
        pose_gt = pose_gt[:3]
    t = np.linalg.norm(pose_pred[:, 3] - pose_gt[:, 3]) * 100
    trace = np.trace(np.dot(pose_pred[:, :3], pose_gt[:, :3].T))
    trace = trace if trace <= 3 else 3
    return np.rad2deg(np.arccos((trace - 1.0) / 2.0)), t


def evaluator_summary(poses_pred, poses_gt):
    """src/evaluators/cmd_evaluator.py:35-61: success rates {cmd1, cmd3, cmd5} of a list of pose pairs."""
    ok = {1: [], 3: [], 5: []}
    for p, g in zip(poses_pred, poses_gt):
        r, t = query_pose_error(np.asarray(p), np.asarray(g))
        for k in ok:
            ok[k].append(bool(t < k and r < k))
    return {"cmd1": np.mean(ok[1]), "cmd3": np.mean(ok[3]), "cmd5": np.mean(ok[5])}


def db_builder_inputs(seed=11, n_points=12, max_leaves=10):
    """Seeded inputs shared by the golden generator and the tests (only seeds + outputs are stored)."""
    rs = np.random.RandomState(seed)
    idxs = rs.randint(1, max_leaves + 1, size=n_points)
    total = int(idxs.sum())
    desc = rs.standard_normal((total, 256)).astype(np.float32)                # [sum L_n, 256] collected leaf descriptors
    scores = rs.rand(total, 1).astype(np.float32)
    xyz = (rs.rand(n_points, 3).astype(np.float32) - 0.5) * 0.2
    poses_gt, poses_pred = [], []
    for i in range(16):
        rv = rs.standard_normal(3) * 0.5
        th = np.linalg.norm(rv); k = rv / th
        Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
        R = np.eye(3) + np.sin(th) * Kx + (1 - np.cos(th)) * Kx @ Kx
        g = np.concatenate([R, np.array([[0.02], [-0.01], [0.5]])], 1)
        dv = rs.standard_normal(3) * (0.002 * 4 ** (i % 4)); dth = np.linalg.norm(dv); dk = dv / dth
        Dx = np.array([[0, -dk[2], dk[1]], [dk[2], 0, -dk[0]], [-dk[1], dk[0], 0]])
        dR = np.eye(3) + np.sin(dth) * Dx + (1 - np.cos(dth)) * Dx @ Dx
        p = np.concatenate([dR @ R, g[:, 3:] + rs.standard_normal((3, 1)) * (0.002 * 3 ** (i % 4))], 1)
        poses_gt.append(g); poses_pred.append(p)
    return dict(idxs=idxs, desc=desc, scores=scores, xyz=xyz, poses_gt=np.stack(poses_gt), poses_pred=np.stack(poses_pred))
