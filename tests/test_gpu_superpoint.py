"""GPU parity of SURVEY.md 8(f) row f3: csrc/superpoint.cu through the C ABI (onepose_b200.superpoint.SuperPoint) against
the golden outputs of the unmodified reference (tests/golden/superpoint.npz), the numpy oracle, and -- at full size -- plain
torch restatements of the same stages on the GPU plus size-independent properties."""
import os

import numpy as np
import pytest
import torch

from oracle import superpoint_oracle as spo
from tests.test_superpoint_oracle import G, SFM_CONF, TOPK_CONF, compare_keypoint_lists

pytestmark = pytest.mark.gpu


def _model(conf, align_corners=True, seed=7):
    from onepose_b200.superpoint import SuperPoint
    torch.manual_seed(seed)                       # CPU initialisation in the reference's layer order: the fixture's weights
    return SuperPoint(conf, align_corners=align_corners).eval().cuda()


def _torch_nms(scores, r):
    mp = lambda x: torch.nn.functional.max_pool2d(x, kernel_size=2 * r + 1, stride=1, padding=r)
    zeros = torch.zeros_like(scores)
    mask = scores == mp(scores)
    for _ in range(2):
        supp = mp(mask.float()) > 0
        ss = torch.where(supp, zeros, scores)
        mask = mask | ((ss == mp(ss)) & ~supp)
    return torch.where(mask, scores, zeros)


def test_score_map_vs_oracle():
    m = _model({})
    for seed, h, w in ((3, 15, 20), (4, 7, 33), (5, 64, 64)):
        logits, _ = spo.sp_inputs(seed, h, w)
        got = m.score_map(torch.from_numpy(logits).cuda()).cpu().numpy()
        np.testing.assert_allclose(got, spo.softmax_shuffle(logits), rtol=3e-6, atol=1e-30)


@pytest.mark.parametrize("r", [0, 1, 3, 4])
def test_nms_golden_bit_exact(r):
    m = _model({})
    logits, _ = spo.sp_inputs()
    scores = torch.from_numpy(spo.softmax_shuffle(logits)).cuda()
    assert np.array_equal(m.nms(scores, r).cpu().numpy(), G[f"stage_nms_r{r}"])


@pytest.mark.parametrize("B,H,W,r", [(3, 512, 512, 3), (2, 480, 640, 4), (1, 37, 91, 2), (2, 64, 40, 9)])
def test_nms_full_size_bit_exact(B, H, W, r):
    m = _model({})
    g = torch.Generator(device="cuda").manual_seed(B * 1000 + H)
    scores = torch.rand(B, H, W, device="cuda", generator=g)
    scores[:, ::7, ::5] = 0.25                                # plateaus: equal values inside one window
    scores[0, : H // 3] = 0.0
    assert torch.equal(m.nms(scores, r), _torch_nms(scores, r))


@pytest.mark.parametrize("name,thr,border,mk", [("a", 0.005, 4, -1), ("b", 0.05, 0, 25), ("c", 0.005, 10, 60)])
def test_select_golden_bit_exact(name, thr, border, mk):
    m = _model({'keypoint_threshold': thr, 'remove_borders': border, 'max_keypoints': mk})
    k, s, c = m.select(torch.from_numpy(G["stage_nms_r3"]).cuda())
    n = int(c[0])
    assert np.array_equal(k[0, :n].cpu().numpy(), G[f"stage_sel_{name}_kpts"])
    assert np.array_equal(s[0, :n].cpu().numpy(), G[f"stage_sel_{name}_scores"])
    assert int(m._overflow.max()) == 0


def test_select_full_size_topk():
    m = _model(SFM_CONF)
    g = torch.Generator(device="cuda").manual_seed(1)
    nms = _torch_nms(torch.rand(4, 512, 512, device="cuda", generator=g), 3)
    nms[3] = 0.0                                                # an image without a single keypoint
    nms[3, 100, 200] = 0.5
    k, s, c = m.select(nms)
    for b in range(4):
        kk = torch.nonzero(nms[b] > 0.005)
        ok = (kk[:, 0] >= 4) & (kk[:, 0] < 508) & (kk[:, 1] >= 4) & (kk[:, 1] < 508)
        kk = kk[ok]
        ss = nms[b][tuple(kk.t())]
        if len(kk) > 4096:
            ss, idx = torch.topk(ss, 4096)
            kk = kk[idx]
        n = int(c[b])
        assert n == len(kk)
        assert torch.equal(s[b, :n], ss)                          # random fp32 scores: no ties
        assert torch.equal(k[b, :n], torch.flip(kk, [1]).float())
    assert int(c[3]) == 1 and k[3, 0].tolist() == [200.0, 100.0]
    assert int(c[0]) == 4096


@pytest.mark.parametrize("ac", [0, 1])
def test_sample_golden(ac):
    m = _model({}, align_corners=bool(ac))
    _, dense = spo.sp_inputs()
    kp = torch.from_numpy(G["stage_sel_a_kpts"]).cuda()[None].contiguous()
    cnt = torch.tensor([kp.shape[1]], dtype=torch.int32, device="cuda")
    d = m.sample(torch.from_numpy(dense).cuda(), kp, cnt)[0].t().cpu().numpy()
    assert np.abs(d - G[f"stage_desc_ac{ac}"]).max() <= 2e-6


@pytest.mark.parametrize("ac", [False, True])
def test_sample_full_size_vs_grid_sample(ac):
    m = _model({}, align_corners=ac)
    g = torch.Generator(device="cuda").manual_seed(2)
    B, h, w, S = 3, 64, 64, 4096
    dense = torch.randn(B, 256, h, w, device="cuda", generator=g)
    kp = torch.stack([torch.randint(0, w * 8, (B, S), device="cuda", generator=g),
                      torch.randint(0, h * 8, (B, S), device="cuda", generator=g)], -1).float()
    cnt = torch.tensor([S, 1000, 0], dtype=torch.int32, device="cuda")
    got = m.sample(dense, kp, cnt)
    nd = torch.nn.functional.normalize(dense, p=2, dim=1)
    q = kp - 8 / 2 + 0.5
    q = q / torch.tensor([w * 8 - 8 / 2 - 0.5, h * 8 - 8 / 2 - 0.5], device="cuda")[None, None]
    q = q * 2 - 1
    ref = torch.nn.functional.grid_sample(nd, q.view(B, 1, -1, 2), mode="bilinear", align_corners=ac)
    ref = torch.nn.functional.normalize(ref.reshape(B, 256, -1), p=2, dim=1).transpose(1, 2)
    assert (got[0] - ref[0]).abs().max() <= 2e-6
    assert (got[1, :1000] - ref[1, :1000]).abs().max() <= 2e-6
    assert torch.count_nonzero(got[1, 1000:]) == 0 and torch.count_nonzero(got[2]) == 0


@pytest.mark.parametrize("tag,conf", [("sfm", SFM_CONF), ("topk", TOPK_CONF)])
def test_post_network_on_reference_maps(tag, conf):
    """The reference's own detector logits and dense map through this repo's kernels -> the reference's forward outputs."""
    m = _model(conf, align_corners=False)
    logits, dense = torch.from_numpy(G["net_logits"]).cuda(), torch.from_numpy(G["net_dense"]).cuda()
    k, s, c = m.select(m.nms(m.score_map(logits)))
    d = m.sample(dense, k, c)
    n = int(c[0])
    compare_keypoint_lists(k[0, :n].cpu().numpy(), s[0, :n].cpu().numpy(), d[0, :n].t().cpu().numpy(),
                           G[f"net_{tag}_kpts"], G[f"net_{tag}_scores"], G[f"net_{tag}_desc"], topk=tag == "topk", rtol=3e-6)


def test_network_dense_maps_vs_reference():
    """cuDNN fp32 convolutions (TF32 off) with the fixture's seeded weights reproduce the reference's CPU maps."""
    m = _model(SFM_CONF)
    logits, dense = m.dense_maps(torch.from_numpy(spo.sp_image()).cuda())
    np.testing.assert_allclose(logits.cpu().numpy(), G["net_logits"], rtol=0, atol=2e-5)
    np.testing.assert_allclose(dense.cpu().numpy(), G["net_dense"], rtol=0, atol=2e-5)


def test_forward_matches_reference_lists():
    """Whole forward, reference output format.  Random-initialised weights give a nearly flat score map, so a handful of local
    maxima may flip under the ~1e-6 differences of the convolutions; everything else must agree."""
    m = _model(SFM_CONF, align_corners=False)
    out = m(torch.from_numpy(spo.sp_image()).cuda())
    assert set(out) == {"keypoints", "scores", "descriptors"} and len(out["keypoints"]) == 1
    k, s, d = out["keypoints"][0].cpu().numpy(), out["scores"][0].cpu().numpy(), out["descriptors"][0].cpu().numpy()
    assert d.shape == (256, len(k)) and s.shape == (len(k),)
    a = {tuple(p): i for i, p in enumerate(k.tolist())}
    b = {tuple(p): i for i, p in enumerate(G["net_sfm_kpts"].tolist())}
    common = set(a) & set(b)
    assert len(common) >= 0.97 * len(b) and abs(len(a) - len(b)) <= 0.03 * len(b)
    for p in common:
        assert abs(s[a[p]] - G["net_sfm_scores"][b[p]]) <= 1e-5 * G["net_sfm_scores"][b[p]]
        assert np.abs(d[:, a[p]] - G["net_sfm_desc"][:, b[p]]).max() <= 1e-4


def test_forward_batched_equals_per_image_and_properties():
    conf = {'nms_radius': 3, 'max_keypoints': 512, 'keypoint_threshold': 0.005, 'remove_borders': 4}
    m = _model(conf)
    imgs = torch.from_numpy(np.concatenate([spo.sp_image(s, 128, 160) for s in (1, 2, 3)])).cuda()
    full = m.forward_batched(imgs)
    assert full["keypoints"].shape == (3, 512, 2) and full["descriptors_pm"].shape == (3, 512, 256)
    logits, dense = m.dense_maps(imgs)             # batch and single-image runs are compared on the same maps

    def post(lg, dn):
        k, s, c = m.select(m.nms(m.score_map(lg)))
        return {"keypoints": k, "scores": s, "count": c, "descriptors_pm": m.sample(dn, k, c)}
    out = post(logits, dense)
    for b in range(3):
        one = post(logits[b:b + 1], dense[b:b + 1])
        n = int(out["count"][b])
        assert n == int(one["count"][0])
        assert torch.equal(out["keypoints"][b, :n], one["keypoints"][0, :n])
        assert torch.equal(out["scores"][b, :n], one["scores"][0, :n])
        assert (out["descriptors_pm"][b, :n] - one["descriptors_pm"][0, :n]).abs().max() <= 1e-5
        k = out["keypoints"][b, :n]
        assert k[:, 0].min() >= 4 and k[:, 0].max() < 156 and k[:, 1].min() >= 4 and k[:, 1].max() < 124
        cheb = (k[:, None] - k[None]).abs().max(-1).values + torch.eye(n, device="cuda") * 1e3
        assert cheb.min() > 3                                           # NMS: no two keypoints within the radius
        assert torch.all(out["scores"][b, :n] > 0.005)
        if n == 512:
            assert torch.all(out["scores"][b, 1:n] <= out["scores"][b, : n - 1])
        nrm = out["descriptors_pm"][b, :n].norm(dim=1)
        assert (nrm - 1).abs().max() <= 1e-5
        assert torch.count_nonzero(out["descriptors_pm"][b, n:]) == 0


def test_feeds_the_matcher():
    """extractor -> matcher hand-over: a block of frames goes from images to poses without leaving the device."""
    from onepose_b200 import synth
    from onepose_b200.matcher import GATsSuperGlue
    conf = {'nms_radius': 3, 'max_keypoints': 256, 'keypoint_threshold': 0.005, 'remove_borders': 4}
    sp = _model(conf)
    imgs = torch.from_numpy(np.concatenate([spo.sp_image(s, 256, 256) for s in (1, 2)])).cuda()
    out = sp.forward_batched(imgs)
    assert int(out["count"].min()) == 256
    obj = synth.make_object(0, 128, 4)
    matcher = GATsSuperGlue(dict(synth.DEFAULT_HPARAMS)).cuda()
    matcher.load_state_dict(synth.make_weights(1, damped=True), strict=True)
    e = lambda a: torch.as_tensor(a).cuda()[None].expand(2, *np.shape(a))
    data = {"keypoints2d": out["keypoints"], "descriptors2d_query": out["descriptors_pm"].transpose(1, 2),
            "keypoints3d": e(obj.keypoints3d), "descriptors3d_db": e(obj.descriptors3d_db), "descriptors2d_db": e(obj.descriptors2d_db)}
    pred, conf_m = matcher(data)
    assert conf_m.shape[0] == 2 and conf_m.numel() == 2 * 128 * 256 and torch.isfinite(conf_m).all()


def test_cpu_input_fails_loudly():
    from onepose_b200.superpoint import SuperPoint
    with pytest.raises(RuntimeError, match="CUDA"):
        SuperPoint({})(torch.zeros(1, 1, 64, 64))
