"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the post-network stages of the reference's SuperPoint
(/root/reference/src/models/extractors/SuperPoint/superpoint.py).  Only tests/, __graft_entry__.smoke() and bench.py's CPU
legs may import this file; the product path (onepose_b200/superpoint.py -> csrc/superpoint.cu) never does.

Pinned against outputs of the unmodified reference functions run in the build container: tests/golden/superpoint.npz, made
by tests/golden/make_golden_superpoint.py (tests/test_superpoint_oracle.py).
"""
import numpy as np


def sp_inputs(seed=3, h=15, w=20):
    """Seeded synthetic detector logits [1,65,h,w] and dense descriptors [1,256,h,w] (numpy RandomState: the same arrays on
    every machine).  The dustbin channel is raised on about half of the cells so that the 0.005 threshold bites."""
    rs = np.random.RandomState(seed)
    logits = (rs.standard_normal((1, 65, h, w)) * 2.5).astype(np.float32)
    logits[:, 64] += (rs.uniform(size=(1, h, w)) < 0.5).astype(np.float32) * 9.0
    dense = rs.standard_normal((1, 256, h, w)).astype(np.float32)
    return logits, dense


def sp_image(seed=5, H=64, W=96):
    """Seeded grayscale test image [1,1,H,W] in [0,1]: blobs and edges on noise."""
    rs = np.random.RandomState(seed)
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
    img = rs.uniform(0, 0.15, size=(H, W)).astype(np.float32)
    for _ in range(12):
        cy, cx, r = rs.uniform(0, H), rs.uniform(0, W), rs.uniform(2, 9)
        img += (((yy - cy) ** 2 + (xx - cx) ** 2) < r * r).astype(np.float32) * np.float32(rs.uniform(0.2, 0.8))
    return np.clip(img, 0, 1)[None, None].astype(np.float32)


def softmax_shuffle(logits):
    """superpoint.py:207-211: softmax over the 65 channels, drop the dustbin, [b,64,h,w] -> [b,8h,8w]."""
    x = logits.astype(np.float32)
    e = np.exp(x - x.max(axis=1, keepdims=True))
    p = (e / e.sum(axis=1, keepdims=True, dtype=np.float32))[:, :-1]
    b, _, h, w = p.shape
    p = p.transpose(0, 2, 3, 1).reshape(b, h, w, 8, 8)
    return p.transpose(0, 1, 3, 2, 4).reshape(b, h * 8, w * 8)


def _max_pool(x, r):
    """max_pool2d(kernel 2r+1, stride 1, padding r): -inf outside."""
    b, H, W = x.shape
    pad = np.full((b, H + 2 * r, W + 2 * r), -np.inf, dtype=x.dtype)
    pad[:, r:r + H, r:r + W] = x
    out = np.full_like(x, -np.inf)
    for dy in range(2 * r + 1):
        for dx in range(2 * r + 1):
            out = np.maximum(out, pad[:, dy:dy + H, dx:dx + W])
    return out


def simple_nms(scores, nms_radius):
    """superpoint.py:47-69."""
    zeros = np.zeros_like(scores)
    max_mask = scores == _max_pool(scores, nms_radius)
    for _ in range(2):
        supp_mask = _max_pool(max_mask.astype(np.float32), nms_radius) > 0
        supp_scores = np.where(supp_mask, zeros, scores)
        new_max_mask = supp_scores == _max_pool(supp_scores, nms_radius)
        max_mask = max_mask | (new_max_mask & ~supp_mask)
    return np.where(max_mask, scores, zeros)


def select(nms_scores, threshold, border, max_keypoints):
    """superpoint.py:214-228 for one image [H,W]: nonzero(s > thr) in row-major order, remove_borders (:72-83),
    top_k_keypoints (:86-93; score descending, ties by position) and the flip to (x, y).  -> (kpts [n,2] fp32, scores [n])."""
    H, W = nms_scores.shape
    k = np.argwhere(nms_scores > threshold)
    s = nms_scores[k[:, 0], k[:, 1]]
    m = (k[:, 0] >= border) & (k[:, 0] < H - border) & (k[:, 1] >= border) & (k[:, 1] < W - border)
    k, s = k[m], s[m]
    if max_keypoints >= 0 and max_keypoints < len(k):
        order = np.lexsort((np.arange(len(s)), -s.astype(np.float64)))[:max_keypoints]
        k, s = k[order], s[order]
    return k[:, ::-1].astype(np.float32), s


def sample_descriptors(keypoints, dense, s=8, align_corners=True):
    """superpoint.py:96-113 for one image: keypoints [n,2] (x,y), dense [256,h,w] (already L2-normalised over channels, :235-236)
    -> [256,n] bilinear samples (zero padding), L2-normalised."""
    c, h, w = dense.shape
    f = np.float32
    kp = keypoints.astype(f) - f(s / 2) + f(0.5)
    kp = kp / np.array([w * s - s / 2 - 0.5, h * s - s / 2 - 0.5], dtype=f)[None]
    kp = kp * f(2) - f(1)
    if align_corners:
        ix, iy = (kp[:, 0] + 1) / 2 * (w - 1), (kp[:, 1] + 1) / 2 * (h - 1)
    else:
        ix, iy = ((kp[:, 0] + 1) * w - 1) / 2, ((kp[:, 1] + 1) * h - 1) / 2
    x0, y0 = np.floor(ix).astype(np.int64), np.floor(iy).astype(np.int64)
    wx1, wy1 = (ix - x0).astype(f), (iy - y0).astype(f)
    out = np.zeros((c, len(kp)), dtype=f)
    for dy, dx, wgt in ((0, 0, (1 - wx1) * (1 - wy1)), (0, 1, wx1 * (1 - wy1)), (1, 0, (1 - wx1) * wy1), (1, 1, wx1 * wy1)):
        y, x = y0 + dy, x0 + dx
        ok = (y >= 0) & (y < h) & (x >= 0) & (x < w)
        v = dense[:, np.clip(y, 0, h - 1), np.clip(x, 0, w - 1)] * ok[None].astype(f)
        out += v * wgt[None].astype(f)
    n = np.maximum(np.sqrt((out.astype(f) ** 2).sum(0, dtype=f)), f(1e-12))
    return out / n[None]


def normalize_dense(dense):
    """F.normalize(descriptors, p=2, dim=1) on [b,256,h,w]."""
    n = np.maximum(np.sqrt((dense ** 2).sum(1, keepdims=True, dtype=np.float32)), np.float32(1e-12))
    return dense / n


def post_network(logits, dense, config, align_corners=True):
    """Everything of SuperPoint.forward after the convolutions, for a batch; lists as the reference returns."""
    nms = simple_nms(softmax_shuffle(logits), config["nms_radius"])
    nd = normalize_dense(dense)
    kpts, scores, descs = [], [], []
    for b in range(nms.shape[0]):
        k, s = select(nms[b], config["keypoint_threshold"], config["remove_borders"], config["max_keypoints"])
        kpts.append(k); scores.append(s); descs.append(sample_descriptors(k, nd[b], 8, align_corners))
    return {"keypoints": kpts, "scores": scores, "descriptors": descs}
