"""GPU parity of the tensor-core convolution stack (csrc/conv_tc.cu, row f3) against torch's fp64 convolution, the library
fp32 path of the same module and the golden maps of the unmodified reference."""
import numpy as np
import pytest
import torch

from oracle import superpoint_oracle as spo
from tests.test_superpoint_oracle import G, SFM_CONF

pytestmark = pytest.mark.gpu


def _planes(x_nchw):
    """NCHW fp32 -> NHWC (hi, lo) planes of 64*x."""
    from onepose_b200._lib import call, ptr, stream_ptr
    x = x_nchw.permute(0, 2, 3, 1).contiguous().float()
    hi, lo = torch.empty_like(x, dtype=torch.float16), torch.empty_like(x, dtype=torch.float16)
    call("op_split_planes", ptr(x), ptr(hi), ptr(lo), x.numel(), stream_ptr())
    return hi, lo


def _unplanes(hi, lo):
    return ((hi.float() + lo.float()) / 64.0).permute(0, 3, 1, 2)


def _conv_tc(x, w, bias, relu, f32_stride=0, pool=False):
    from onepose_b200._lib import call, ptr, stream_ptr
    B, ci, H, W = x.shape
    co, _, k, _ = w.shape
    cp = 64 if co == 64 else -(-co // 128) * 128
    wt = torch.zeros(k * k, cp, ci, device="cuda")
    wt[:, :co] = w.permute(2, 3, 0, 1).reshape(k * k, co, ci)
    wh, wl = torch.empty_like(wt, dtype=torch.float16), torch.empty_like(wt, dtype=torch.float16)
    call("op_split_planes", ptr(wt), ptr(wh), ptr(wl), wt.numel(), stream_ptr())
    xh, xl = _planes(x)
    if f32_stride:
        y = torch.full((B, H, W, f32_stride), float("nan"), device="cuda")
        call("op_conv_tc", ptr(xh), ptr(xl), B, H, W, ci, ptr(wh), ptr(wl), cp, ptr(bias), co, k * k, int(relu), 0, None, None, ptr(y), f32_stride, stream_ptr())
        return y.permute(0, 3, 1, 2)
    Ho, Wo = (H // 2, W // 2) if pool else (H, W)
    yh = torch.full((B, Ho, Wo, co), float("nan"), dtype=torch.float16, device="cuda")
    yl = torch.full_like(yh, float("nan"))
    call("op_conv_tc", ptr(xh), ptr(xl), B, H, W, ci, ptr(wh), ptr(wl), cp, ptr(bias), co, k * k, int(relu), int(pool), ptr(yh), ptr(yl), None, 0,
         stream_ptr())
    return _unplanes(yh, yl)


@pytest.mark.parametrize("B,H,W,ci,co,k,relu,f32,pool", [
    # W % 8 == 0 and 3x3: the halo kernel (8 x 16 pixel tiles)
    (2, 16, 128, 64, 64, 3, True, 0, False),
    (1, 9, 256, 64, 64, 3, True, 0, False),       # ragged in y
    (2, 24, 80, 64, 128, 3, True, 0, False),
    (1, 13, 64, 128, 128, 3, True, 0, False),
    (1, 40, 40, 128, 128, 3, False, 0, False),    # no ReLU
    (1, 64, 64, 128, 256, 3, True, 0, False),     # two channel tiles
    (2, 32, 64, 64, 64, 3, True, 0, True),        # fused 2x2 max-pool
    (1, 22, 40, 64, 128, 3, True, 0, True),       # pooled, ragged in y
    (1, 16, 16, 128, 128, 3, True, 0, True),
    (3, 48, 8, 64, 64, 3, True, 0, False),        # a single tile column
    # everything else: the tap-by-tap kernel (128-pixel row tiles)
    (2, 8, 12, 128, 256, 3, True, 0, False),      # image narrower than a tile
    (1, 20, 100, 64, 64, 3, True, 0, False),      # 64 x 2 tiles, ragged in x
    (1, 5, 300, 64, 64, 3, False, 0, False),      # 128 x 1 tiles, ragged in x
    (1, 12, 36, 128, 128, 3, True, 0, False),     # 32 x 4 tiles
    (2, 16, 64, 256, 65, 1, False, 72, False),    # detector head's 1x1, fp32 out, padded channel count
    (1, 8, 32, 256, 256, 1, False, 256, False),   # descriptor head's 1x1
    (1, 7, 130, 64, 64, 1, True, 0, False),       # 1x1 with plane output
])
def test_conv_tc_vs_fp64(B, H, W, ci, co, k, relu, f32, pool):
    g = torch.Generator(device="cuda").manual_seed(H * 1000 + W + ci)
    x = torch.relu(torch.randn(B, ci, H, W, device="cuda", generator=g)) * 3.0
    w = torch.randn(co, ci, k, k, device="cuda", generator=g) / (ci * k * k) ** 0.5
    bias = torch.randn(co, device="cuda", generator=g)
    got = _conv_tc(x, w, bias, relu, f32, pool)
    xq = _unplanes(*_planes(x))                                      # the planes' value of x (2^-22 relative from x)
    ref = torch.nn.functional.conv2d(xq.double(), w.double(), bias.double(), padding=k // 2)
    ref = torch.relu(ref) if relu else ref
    ref = torch.nn.functional.max_pool2d(ref, 2, 2) if pool else ref
    assert got.shape[2:] == ref.shape[2:]
    assert torch.isfinite(got[:, :co]).all()
    err = (got[:, :co].double() - ref).abs().max().item()
    # fp32 TMEM accumulation over up to 9 * Cin / 16 * 3 chained MMAs (the tensor core truncates when it aligns addends): a few 1e-6 of
    # the largest value, the same class as a library fp32 convolution
    assert err <= 8e-6 * ref.abs().max().item(), err
    if f32 > co:
        assert torch.count_nonzero(got[:, co:]) == 0                # the padding columns of a row are written as zeros


def test_conv1a_and_pool():
    from onepose_b200._lib import call, ptr, stream_ptr
    g = torch.Generator(device="cuda").manual_seed(3)
    img = torch.rand(2, 1, 40, 72, device="cuda", generator=g)
    w = torch.randn(64, 1, 3, 3, device="cuda", generator=g) / 3
    b = torch.randn(64, device="cuda", generator=g) / 4
    yh = torch.empty(2, 40, 72, 64, dtype=torch.float16, device="cuda"); yl = torch.empty_like(yh)
    call("op_conv1a_planes", ptr(img), ptr(w.reshape(64, 9).contiguous()), ptr(b), 2, 40, 72, ptr(yh), ptr(yl), stream_ptr())
    ref = torch.relu(torch.nn.functional.conv2d(img.double(), w.double(), b.double(), padding=1))
    y = _unplanes(yh, yl)
    assert (y.double() - ref).abs().max() <= 2e-6 * ref.abs().max()
    ph = torch.empty(2, 20, 36, 64, dtype=torch.float16, device="cuda"); pl = torch.empty_like(ph)
    call("op_pool2_planes", ptr(yh), ptr(yl), 2, 40, 72, 64, ptr(ph), ptr(pl), stream_ptr())
    assert torch.equal(_unplanes(ph, pl), torch.nn.functional.max_pool2d(y, 2, 2))        # exact: a selection, not arithmetic


def _model(conf, **kw):
    from onepose_b200.superpoint import SuperPoint
    torch.manual_seed(7)
    return SuperPoint(conf, **kw).eval().cuda()


def test_network_tc_vs_reference_maps():
    """The whole stack with the fixture's seeded weights on the fixture's image: tensor-core maps against the reference's CPU maps."""
    m = _model(SFM_CONF)
    logits, dense = m.dense_maps_tc(torch.from_numpy(spo.sp_image()).cuda())
    assert logits.shape == (1, 8, 12, 72) and dense.shape == (1, 8, 12, 256)
    np.testing.assert_allclose(logits[..., :65].permute(0, 3, 1, 2).cpu().numpy(), G["net_logits"], rtol=0, atol=2e-5)
    np.testing.assert_allclose(dense.permute(0, 3, 1, 2).cpu().numpy(), G["net_dense"], rtol=0, atol=2e-5)


def test_network_tc_vs_library_full_size():
    m = _model(SFM_CONF)
    g = torch.Generator(device="cuda").manual_seed(11)
    img = torch.rand(3, 1, 512, 512, device="cuda", generator=g)
    lt, dt = m.dense_maps_tc(img)
    lc, dc = m.dense_maps(img)
    assert (lt[..., :65].permute(0, 3, 1, 2) - lc).abs().max() <= 2e-5 * lc.abs().max().clamp_min(1.0)
    assert (dt.permute(0, 3, 1, 2) - dc).abs().max() <= 2e-5 * dc.abs().max().clamp_min(1.0)


def test_forward_tc_equals_forward_library_off_ties():
    """Same keypoints / descriptors from both convolution paths, except where two candidates tie to rounding."""
    conf = {'nms_radius': 3, 'max_keypoints': 1024, 'keypoint_threshold': 0.005, 'remove_borders': 4}
    a, b = _model(conf, conv_impl="tc"), _model(conf, conv_impl="cudnn")
    img = torch.from_numpy(np.concatenate([spo.sp_image(s, 256, 320) for s in (1, 2)])).cuda()
    oa, ob = a.forward_batched(img), b.forward_batched(img)
    for i in range(2):
        ka = {tuple(p): j for j, p in enumerate(oa["keypoints"][i, :int(oa["count"][i])].tolist())}
        kb = {tuple(p): j for j, p in enumerate(ob["keypoints"][i, :int(ob["count"][i])].tolist())}
        common = set(ka) & set(kb)
        assert len(common) >= 0.95 * len(kb)
        ia, ib = [ka[p] for p in common], [kb[p] for p in common]
        assert (oa["descriptors_pm"][i, ia] - ob["descriptors_pm"][i, ib]).abs().max() <= 1e-4
        assert ((oa["scores"][i, ia] - ob["scores"][i, ib]).abs() <= 1e-5 * ob["scores"][i, ib]).all()
