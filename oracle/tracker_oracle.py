"""TEST INFRASTRUCTURE ONLY -- numpy / torch-CPU restatement of the reference's tracker geometry
(/root/reference/src/tracker/tracking_utils.py, ba_tracker.py).  Only tests/ and bench legs may import this file.

Pinned (tests/test_tracker_oracle.py, fixture tests/golden/tracker.npz made by tests/golden/make_golden_tracker.py from the
unmodified reference functions): AngleAxisRotatePoint, SnavelyReprojectionErrorV2, project, apply_triangulation.
The bundle-adjustment solver is NOT pinned ("parity unpinned"): the reference delegates to DeepLM's Solve
(github.com/hjwdzh/DeepLM, imported at ba_tracker.py:4, not vendored, no version pinned), which is absent here.  ``ba_lm`` below
restates textbook Levenberg-Marquardt (Marquardt diagonal scaling, x0.1 / x10 damping schedule) on the reference's residual
with autograd Jacobians and a dense solve -- the same algorithm the CUDA path runs with analytic Jacobians and a Schur complement.
"""
import numpy as np
import torch


def tracker_inputs(seed=0, n_cams=6, n_pts=60, obs_per_pt=4, noise_px=0.5, perturb=0.01):
    """Seeded synthetic window: object points in a 0.2 m cube, cameras ~0.6 m away looking at it (BAL vectors), observations
    of each point in ``obs_per_pt`` cameras with pixel noise, and a perturbed initial estimate."""
    rs = np.random.RandomState(seed)
    pts = rs.uniform(-0.1, 0.1, (n_pts, 3))
    cams = np.zeros((n_cams, 9))
    for c in range(n_cams):
        cams[c, :3] = rs.uniform(-0.4, 0.4, 3)
        cams[c, 3:6] = [rs.uniform(-0.05, 0.05), rs.uniform(-0.05, 0.05), rs.uniform(0.5, 0.7)]
        cams[c, 6:] = [rs.uniform(500, 700), 256 + rs.uniform(-5, 5), 256 + rs.uniform(-5, 5)]
    cams[0, :3] = 0.0                                       # one exact zero rotation: the reference's first-order branch
    pt_idx, cam_idx = [], []
    for p in range(n_pts):
        for c in rs.choice(n_cams, size=min(obs_per_pt, n_cams), replace=False):
            pt_idx.append(p); cam_idx.append(c)
    pt_idx, cam_idx = np.array(pt_idx), np.array(cam_idx)
    perm = rs.permutation(len(pt_idx))
    pt_idx, cam_idx = pt_idx[perm], cam_idx[perm]
    feats = np.concatenate([np.zeros((len(pt_idx), 2)), cams[cam_idx, 6:]], axis=1)
    feats[:, :2] = snavely_residual(pts[pt_idx], cams[cam_idx, :6], feats) + rs.normal(0, noise_px, (len(pt_idx), 2))
    pts0 = pts + rs.normal(0, perturb, pts.shape)
    cams0 = cams.copy(); cams0[:, :6] += rs.normal(0, perturb, (n_cams, 6))
    return dict(points_gt=pts, cams_gt=cams, points0=pts0, cams0=cams0, features=feats, pt_idx=pt_idx, cam_idx=cam_idx)


def angle_axis_rotate_point(w, p):
    """tracking_utils.py:91-140 (numpy, float64)."""
    w, p = np.asarray(w, np.float64), np.asarray(p, np.float64)
    t2 = (w * w).sum(1)
    mask = (t2 > 0).astype(np.float64)
    t = np.sqrt(t2 + (1 - mask))
    c, s, ti = np.cos(t), np.sin(t), 1.0 / t
    k = w * ti[:, None]
    cross = np.cross(k, p)
    tmp = (k * p).sum(1) * (1 - c)
    res1 = p * c[:, None] + cross * s[:, None] + k * tmp[:, None]
    res2 = p + np.cross(w, p)
    return res1 * mask[:, None] + res2 * (1 - mask[:, None])


def snavely_residual(points_ob, cameras_ob, features):
    """tracking_utils.py:143-169."""
    q = angle_axis_rotate_point(cameras_ob[:, :3], points_ob) + cameras_ob[:, 3:6]
    px = features[:, 2] * (q[:, 0] / q[:, 2]) + features[:, 3]
    py = features[:, 2] * (q[:, 1] / q[:, 2]) + features[:, 4]
    return np.stack([px - features[:, 0], py - features[:, 1]], 1)


def project(xyz, K, RT):
    """tracking_utils.py:74-89 -> (xy, depth)."""
    q = xyz @ RT[:, :3].T + RT[:, 3:].T
    h = q @ K.T
    return h[:, :2] / h[:, 2:], q[:, 2]


def triangulate(P1, P2, x1, x2):
    """ba_tracker.py:250-265 (_triangulatePy; the same DLT cv2.triangulatePoints solves, :267-273)."""
    out = []
    for p1, p2 in zip(x1, x2):
        A = np.stack([p1[0] * P1[2] - P1[0], p1[1] * P1[2] - P1[1], p2[0] * P2[2] - P2[0], p2[1] * P2[2] - P2[1]])
        _, _, vh = np.linalg.svd(A)
        X = vh[3]
        out.append(X[:3] / X[3])
    return np.array(out)


def _residual_torch(pts, cams, feats):
    w, p = cams[:, :3], pts
    t2 = (w * w).sum(1)
    mask = (t2 > 0).double()
    t = torch.sqrt(t2 + (1 - mask))
    k = w / t[:, None]
    res1 = p * torch.cos(t)[:, None] + torch.cross(k, p, dim=1) * torch.sin(t)[:, None] + k * ((k * p).sum(1) * (1 - torch.cos(t)))[:, None]
    res2 = p + torch.cross(w, p, dim=1)
    q = res1 * mask[:, None] + res2 * (1 - mask[:, None]) + cams[:, 3:6]
    return torch.stack([feats[:, 2] * q[:, 0] / q[:, 2] + feats[:, 3] - feats[:, 0], feats[:, 2] * q[:, 1] / q[:, 2] + feats[:, 4] - feats[:, 1]], 1)


def jacobians(points_ob, cameras_ob, features):
    """Per-observation Jacobians by autograd of the reference's residual: (r [n,2], Jc [n,2,6], Jp [n,2,3])."""
    pts = torch.tensor(points_ob, dtype=torch.float64, requires_grad=True)
    cams = torch.tensor(cameras_ob[:, :6], dtype=torch.float64, requires_grad=True)
    feats = torch.tensor(features, dtype=torch.float64)
    r = _residual_torch(pts, cams, feats)
    Jc, Jp = [], []
    for i in range(2):
        gc, gp = torch.autograd.grad(r[:, i].sum(), [cams, pts], retain_graph=True)
        Jc.append(gc.numpy()); Jp.append(gp.numpy())
    return r.detach().numpy(), np.stack(Jc, 1), np.stack(Jp, 1)


def ba_lm(points, cams6, features, pt_idx, cam_idx, num_iterations=5, num_success=5, lambda0=1e-4):
    """Dense Levenberg-Marquardt on the Snavely residual; returns (points, cams6, log of (cost, accepted) per trial)."""
    pts, cm = np.array(points, np.float64), np.array(cams6[:, :6], np.float64)
    C, P = cm.shape[0], pts.shape[0]
    lam, nsucc, log = lambda0, 0, []
    for _ in range(num_iterations):
        r, Jc, Jp = jacobians(pts[pt_idx], cm[cam_idx], features)
        cost = float((r * r).sum())
        n = len(pt_idx)
        J = np.zeros((2 * n, 6 * C + 3 * P))
        for o in range(n):
            J[2 * o:2 * o + 2, 6 * cam_idx[o]:6 * cam_idx[o] + 6] = Jc[o]
            J[2 * o:2 * o + 2, 6 * C + 3 * pt_idx[o]:6 * C + 3 * pt_idx[o] + 3] = Jp[o]
        H = J.T @ J
        g = J.T @ r.reshape(-1)
        H[np.diag_indices_from(H)] += lam * np.diag(H) + 1e-12
        d = -np.linalg.solve(H, g)
        cm_new, pts_new = cm + d[:6 * C].reshape(C, 6), pts + d[6 * C:].reshape(P, 3)
        rn = snavely_residual(pts_new[pt_idx], cm_new[cam_idx], features)
        new_cost = float((rn * rn).sum())
        ok = new_cost < cost
        log.append((cost, new_cost, ok))
        if ok:
            cm, pts, lam, nsucc = cm_new, pts_new, max(lam * 0.1, 1e-12), nsucc + 1
        else:
            lam = min(lam * 10, 1e12)
        if nsucc >= num_success:
            break
    return pts, cm, log
