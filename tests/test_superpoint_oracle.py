"""CPU: the numpy restatement of SuperPoint's post-network stages (oracle/superpoint_oracle.py) against the outputs of the
unmodified reference (tests/golden/superpoint.npz, made by tests/golden/make_golden_superpoint.py)."""
import os

import numpy as np
import pytest

from oracle import superpoint_oracle as spo

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "superpoint.npz"))
SFM_CONF = {'nms_radius': 3, 'keypoint_threshold': 0.005, 'max_keypoints': 4096, 'remove_borders': 4}
TOPK_CONF = {'nms_radius': 4, 'max_keypoints': 40, 'keypoint_threshold': 0.016, 'remove_borders': 6}


def compare_keypoint_lists(k, s, d, k_ref, s_ref, d_ref, topk, rtol=2e-6, dtol=2e-6):
    """Same keypoints, scores within fp32 softmax rounding, descriptors within dtol.  Below the top-k cut the list is in
    nonzero order and must agree exactly; in the top-k case the order is by score, so keypoints whose scores agree to
    rounding may swap places and the comparison is made per keypoint (position -> score, descriptor)."""
    assert k.shape == k_ref.shape
    if not topk:
        assert np.array_equal(k, k_ref)
        np.testing.assert_allclose(s, s_ref, rtol=rtol, atol=0)
        assert np.abs(d - d_ref).max() <= dtol
        return
    assert np.all(np.diff(s) <= 0) and np.all(np.diff(s_ref) <= np.abs(s_ref[1:]) * rtol)
    np.testing.assert_allclose(s, s_ref, rtol=rtol, atol=0)          # the sorted score lists agree
    a = {tuple(p): i for i, p in enumerate(k.tolist())}
    b = {tuple(p): i for i, p in enumerate(k_ref.tolist())}
    common = set(a) & set(b)
    cut = s_ref[-1]
    for p in set(a) ^ set(b):                                          # only ties with the cut-off score may differ
        sc = s[a[p]] if p in a else s_ref[b[p]]
        assert abs(sc - cut) <= rtol * cut, (p, sc, cut)
    assert len(common) >= len(a) - 2
    for p in common:
        assert abs(s[a[p]] - s_ref[b[p]]) <= rtol * s_ref[b[p]]
        assert np.abs(d[:, a[p]] - d_ref[:, b[p]]).max() <= dtol


@pytest.mark.parametrize("r", [0, 1, 3, 4])
def test_simple_nms_bit_exact(r):
    logits, _ = spo.sp_inputs()
    scores = spo.softmax_shuffle(logits)
    assert np.array_equal(spo.simple_nms(scores, r), G[f"stage_nms_r{r}"])


@pytest.mark.parametrize("name,thr,border,mk", [("a", 0.005, 4, -1), ("b", 0.05, 0, 25), ("c", 0.005, 10, 60)])
def test_select_bit_exact(name, thr, border, mk):
    k, s = spo.select(G["stage_nms_r3"][0], thr, border, mk)
    assert np.array_equal(k, G[f"stage_sel_{name}_kpts"])
    assert np.array_equal(s, G[f"stage_sel_{name}_scores"])
    if mk >= 0:
        assert len(k) <= mk


@pytest.mark.parametrize("ac", [0, 1])
def test_sample_descriptors(ac):
    _, dense = spo.sp_inputs()
    d = spo.sample_descriptors(G["stage_sel_a_kpts"], spo.normalize_dense(dense)[0], 8, bool(ac))
    ref = G[f"stage_desc_ac{ac}"]
    assert d.shape == ref.shape
    assert np.abs(d - ref).max() <= 2e-6          # fp32 rounding of the four-tap blend; unit-norm vectors


@pytest.mark.parametrize("tag,conf", [("sfm", SFM_CONF), ("topk", TOPK_CONF)])
def test_post_network_on_reference_maps(tag, conf):
    """The reference's own logits / dense map -> the reference's forward outputs (this container's torch: align_corners=False)."""
    out = spo.post_network(G["net_logits"], G["net_dense"], conf, align_corners=False)
    compare_keypoint_lists(out["keypoints"][0], out["scores"][0], out["descriptors"][0],
                           G[f"net_{tag}_kpts"], G[f"net_{tag}_scores"], G[f"net_{tag}_desc"], topk=conf["max_keypoints"] == 40)


def test_sfm_config_key_quirk():
    """extract_features.py:9-17 spells the key 'keypoints_threshold'; the model reads 'keypoint_threshold' (superpoint.py:136),
    so the 0.005 default stays in force -- the host mirror must behave the same."""
    from onepose_b200.superpoint import SuperPoint
    assert {**SuperPoint.default_config, 'keypoints_threshold': 0.6}['keypoint_threshold'] == 0.005


def test_state_dict_names_are_the_reference_checkpoint_s():
    """superpoint_v1.pth's names and shapes (superpoint.py:147-166) load with strict=True."""
    import torch
    from onepose_b200.superpoint import SuperPoint
    m = SuperPoint({'descriptor_dim': 256, 'nms_radius': 3, 'max_keypoints': 4096, 'keypoints_threshold': 0.6})
    shapes = {"conv1a": (64, 1, 3), "conv1b": (64, 64, 3), "conv2a": (64, 64, 3), "conv2b": (64, 64, 3), "conv3a": (128, 64, 3),
              "conv3b": (128, 128, 3), "conv4a": (128, 128, 3), "conv4b": (128, 128, 3), "convPa": (256, 128, 3), "convPb": (65, 256, 1),
              "convDa": (256, 128, 3), "convDb": (256, 256, 1)}
    sd = {}
    for n, (co, ci, k) in shapes.items():
        sd[n + ".weight"] = torch.zeros(co, ci, k, k); sd[n + ".bias"] = torch.zeros(co)
    m.load_state_dict(sd, strict=True)
    assert m.config['keypoint_threshold'] == 0.005 and m.config['nms_radius'] == 3
    with pytest.raises(ValueError):
        SuperPoint({'max_keypoints': 0})
