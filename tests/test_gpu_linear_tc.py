"""GPU parity tests of the tcgen05 split-precision projections (op_linear_tc and its plane producers) against
float64 restatements of the reference ops (Conv1d k=1, elu+1, InstanceNorm1d+ReLU, residual;
GATs_SuperGlue.py:88-147), and of the whole matcher against the exact-fp32 CUDA-core path."""
import numpy as np
import pytest
import torch

from onepose_b200 import synth

pytestmark = pytest.mark.gpu


def _planes(x):
    from onepose_b200._lib import call, ptr, stream_ptr
    x = x.contiguous()
    hi = torch.empty(x.shape, dtype=torch.float16, device=x.device)
    lo = torch.empty_like(hi)
    call("op_split_planes", ptr(x), ptr(hi), ptr(lo), x.numel(), stream_ptr())
    return hi, lo


def _unsplit(p):
    return (p[0].double() + p[1].double()) / 64.0


def _linear_tc(A1, W, bias, *, A2=None, W2=None, want_c=True, want_planes=False, flags=0, elu_cols=0, C_init=None, ksum=None, zscale_num=1.0):
    from onepose_b200._lib import call, lib, ptr, stream_ptr
    B, rows, K1 = A1.shape
    Cout = W.shape[-2]
    a1 = _planes(A1); w = _planes(W)
    a2 = _planes(A2) if A2 is not None else None
    w2 = _planes(W2) if W2 is not None else None
    C = (C_init.clone() if C_init is not None else torch.full((B, rows, Cout), float("nan"), device="cuda")) if want_c else None
    Cp = (torch.zeros(B, rows, Cout, dtype=torch.float16, device="cuda"), torch.zeros(B, rows, Cout, dtype=torch.float16, device="cuda")) if want_planes else None
    parts = lib.op_linear_tc_stat_parts(rows)
    stats = torch.full((B, parts, Cout, 2), float("nan"), device="cuda") if flags & (4 | 16 | 256) else None
    call("op_linear_tc", ptr(a1[0]), ptr(a1[1]), K1, ptr(a2[0]) if a2 else None, ptr(a2[1]) if a2 else None,
         0 if A2 is None else A2.shape[2], ptr(w[0]), ptr(w[1]), ptr(w2[0]) if w2 else None, ptr(w2[1]) if w2 else None, ptr(bias), ptr(C), ptr(Cp[0]) if Cp else None,
         ptr(Cp[1]) if Cp else None, B, rows, Cout, flags, elu_cols, ptr(stats), ptr(ksum), float(zscale_num), stream_ptr())
    torch.cuda.synchronize()
    return C, Cp, stats


def _rel(x, ref):
    return ((x.double() - ref).abs().max() / ref.abs().max()).item()


def test_split_planes_roundtrip():
    g = torch.Generator().manual_seed(1)
    x = (torch.randn(4096, generator=g) * torch.logspace(-4, 2, 4096)).cuda()
    back = _unsplit(_planes(x))
    err = (back - x.double()).abs()
    assert float((err / x.double().abs().clamp_min(1e-3)).max()) < 2e-6        # 22 significant bits above fp16's subnormal floor


@pytest.mark.parametrize("shape", [(1, 128, 256, 256), (2, 300, 256, 512), (3, 77, 512, 256), (1, 2048, 256, 256), (2, 1000, 512, 512)])
def test_linear_tc_plain(shape):
    B, rows, K, Cout = shape
    g = torch.Generator().manual_seed(sum(shape))
    A = torch.randn(B, rows, K, generator=g).cuda() * 0.3
    W = (torch.randn(Cout, K, generator=g) / K ** 0.5).cuda()
    bias = torch.randn(Cout, generator=g).cuda() * 0.1
    ref = A.double() @ W.double().t() + bias.double()
    C, Cp, _ = _linear_tc(A, W, bias, want_planes=True)
    assert _rel(C, ref) < 5e-6
    assert _rel(_unsplit(Cp), ref) < 5e-6
    C2, _, _ = _linear_tc(A, W, bias)
    assert torch.equal(C, C2)                                                   # deterministic


def test_linear_tc_two_sources_elu_stats_residual():
    B, rows = 2, 333
    g = torch.Generator().manual_seed(5)
    A1 = torch.randn(B, rows, 256, generator=g).cuda() * 0.2
    A2 = torch.randn(B, rows, 256, generator=g).cuda() * 0.5
    W = (torch.randn(512, 512, generator=g) / 512 ** 0.5).cuda()
    bias = torch.randn(512, generator=g).cuda()
    lin = torch.cat([A1, A2], 2).double() @ W.double().t()
    # concatenated K + InstanceNorm statistics (MLP layer 0, :135-147)
    C, _, stats = _linear_tc(A1, W, bias, A2=A2, flags=4)
    assert _rel(C, lin + bias.double()) < 5e-6
    from onepose_b200._lib import call, lib, ptr, stream_ptr
    mean = torch.empty(B, 512, device="cuda"); rstd = torch.empty(B, 512, device="cuda")
    call("op_instnorm_finalize_tc", ptr(stats), ptr(bias), B, lib.op_linear_tc_stat_parts(rows), 512, rows, 1e-5, ptr(mean), ptr(rstd), stream_ptr())
    h = lin + bias.double()
    np.testing.assert_allclose(mean.cpu().numpy(), h.mean(1).cpu().numpy(), rtol=0, atol=2e-6)
    np.testing.assert_allclose(rstd.cpu().numpy(), (1.0 / torch.sqrt(h.var(1, unbiased=False) + 1e-5)).cpu().numpy(), rtol=3e-6)
    # InstanceNorm + ReLU producer of the next GEMM's planes
    hn = (torch.zeros(B, rows, 512, dtype=torch.float16, device="cuda"), torch.zeros(B, rows, 512, dtype=torch.float16, device="cuda"))
    call("op_norm_relu_split", ptr(C), ptr(mean), ptr(rstd), ptr(hn[0]), ptr(hn[1]), B, rows, 512, stream_ptr())
    ref_n = torch.relu((h - h.mean(1, keepdim=True)) / torch.sqrt(h.var(1, unbiased=False, keepdim=True) + 1e-5))
    assert float((_unsplit(hn) - ref_n).abs().max()) < 2e-5
    # elu+1 on the first 256 output columns only (K | V projection, :90-91)
    Ce, _, _ = _linear_tc(A1, W[:, :256].contiguous(), bias, flags=1, elu_cols=256)
    l2 = A1.double() @ W[:, :256].double().t() + bias.double()
    ref_e = torch.cat([torch.nn.functional.elu(l2[..., :256]) + 1, l2[..., 256:]], 2)
    assert _rel(Ce, ref_e) < 5e-6
    # residual add in place (:78,83)
    x0 = torch.randn(B, rows, 512, generator=g).cuda()
    Cr, Crp, _ = _linear_tc(A1, W[:, :256].contiguous(), bias, flags=2, C_init=x0, want_planes=True)
    assert _rel(Cr, x0.double() + l2) < 5e-6
    assert _rel(_unsplit(Crp), x0.double() + l2) < 5e-6


@pytest.mark.parametrize("damped", [False, True])
def test_matcher_tensor_core_path_vs_cuda_core_path(damped):
    from onepose_b200.matcher import GATsSuperGlue
    from tests import helpers
    hp = dict(synth.DEFAULT_HPARAMS); hp["match_threshold"] = 0.0 if not damped else 0.2
    obj = synth.make_object(1, 700, 8)
    frames = [synth.make_frame(obj, f, 450) for f in range(2)]
    data = {k: v.cuda() for k, v in synth.pack_batch(obj, frames).items()}
    m = GATsSuperGlue(hp).cuda().eval()
    m.load_state_dict(synth.make_weights(0, damped=damped), strict=True)
    pred_tc, conf_tc = m(data)
    conf_tc = conf_tc.clone(); m0_tc = pred_tc["matches0_batched"].clone()
    m.use_tensor_cores = False
    pred_cc, conf_cc = m(data)
    assert helpers.conf_rel_err(conf_tc.cpu(), conf_cc.cpu()) < helpers.CONF_RTOL
    for b in range(2):
        helpers.assert_matches_equal_off_ties(m0_tc[b].cpu(), pred_cc["matches0_batched"][b].cpu(), conf_cc[b].cpu(), axis=2)


@pytest.mark.parametrize("n2,n3,shared", [(450, 700, True), (512, 512, False)])
def test_two_stream_schedule_gives_the_same_bits_as_one_stream(n2, n3, shared):
    """The 2D / 3D streams of a layer run on two CUDA streams; scratch buffers are per stream (also when N2 == N3) and the
    cross-layer hand-over is ordered by events, so the result may not depend on the schedule."""
    from onepose_b200.matcher import GATsSuperGlue
    obj = synth.make_object(2, n3, 8)
    frames = [synth.make_frame(obj, f, n2) for f in range(3)]
    data = {k: v.cuda() for k, v in synth.pack_batch(obj, frames).items()}
    if not shared:                                   # per-frame DB copies: no shared-prefix shortcut
        data = {k: v.contiguous() for k, v in data.items()}
    m = GATsSuperGlue(dict(synth.DEFAULT_HPARAMS)).cuda().eval()
    m.load_state_dict(synth.make_weights(0, damped=True), strict=True)
    outs = []
    for two in (True, False, True):
        m.two_streams = two
        pred, conf = m(data)
        outs.append((conf.clone(), pred["matches0_batched"].clone(), pred["matching_scores0_batched"].clone()))
    for o in outs[1:]:
        assert all(torch.equal(a, b) for a, b in zip(outs[0], o))


@pytest.mark.parametrize("shape", [(1, 512), (2, 700), (3, 33), (1, 2048)])
def test_atb_tc_and_kv_post_vs_float64_linear_attention(shape):
    """G = K'^T X on MN-major UMMA operands, then KV' / Ksum / P_b against a float64 restatement of :88-99,119."""
    from onepose_b200._lib import call, lib, ptr, stream_ptr
    B, M = shape
    g = torch.Generator().manual_seed(M)
    X = (torch.randn(B, M, 256, generator=g) * 0.2).cuda()
    Kp = (torch.rand(B, M, 256, generator=g) * 2).cuda()              # elu+1 > 0
    nch = lib.op_atb_chunks(M)
    part = torch.full((B, nch, 256, 256), float("nan"), device="cuda")
    kp, xp = _planes(Kp), _planes(X)
    call("op_atb_tc", ptr(kp[0]), ptr(kp[1]), ptr(xp[0]), ptr(xp[1]), B, M, ptr(part), stream_ptr())
    torch.cuda.synchronize()
    G = part.double().sum(1)
    G_ref = Kp.double().transpose(1, 2) @ X.double()
    assert _rel(G, G_ref) < 5e-6
    # post: Ksum partials as OPB_LIN_COLSUM would write them (one part per 32 rows here), KV', P_b
    Wv = (torch.randn(256, 256, generator=g) / 16).cuda(); bv = (torch.randn(256, generator=g) * 0.1).cuda()
    n_out = 256 if M % 2 else 512                                       # merge alone / merge folded into MLP layer 0
    Wm = (torch.randn(n_out, 256, generator=g) / 16).cuda()
    nparts = (M + 31) // 32
    pad = torch.zeros(B, nparts * 32, 256, device="cuda"); pad[:, :M] = Kp
    ks_part = torch.stack([pad.view(B, nparts, 32, 256).sum(2), torch.zeros(B, nparts, 256, device="cuda")], -1).contiguous()
    ksum = torch.empty(B, 256, device="cuda"); kv = torch.empty(B, 4, 64, 64, device="cuda")
    Ph = torch.empty(B, n_out, 256, dtype=torch.float16, device="cuda"); Pl = torch.empty_like(Ph)
    WvT = Wv.view(4, 64, 256).transpose(1, 2).contiguous()              # [h, c, q]
    WmT = Wm.view(n_out, 4, 64).permute(1, 2, 0).contiguous()           # [h, q, o]
    call("op_kv_post", ptr(part), nch, ptr(ks_part), nparts, ptr(WvT), ptr(bv), ptr(WmT), n_out, B, 1.0 / M, ptr(ksum), ptr(kv), ptr(Ph), ptr(Pl), stream_ptr())
    torch.cuda.synchronize()
    V = X.double() @ Wv.double().t() + bv.double()
    Kh = Kp.double().view(B, M, 4, 64); Vh = V.view(B, M, 4, 64)
    KV_ref = torch.einsum("bmhd,bmhq->bhdq", Kh, Vh) / M
    ks_ref = Kp.double().sum(1)
    assert _rel(ksum, ks_ref) < 2e-6
    assert _rel(kv, KV_ref) < 1e-5
    P_ref = torch.einsum("bhdq,ohq->bohd", KV_ref, Wm.double().view(n_out, 4, 64)).reshape(B, n_out, 256)
    assert _rel(_unsplit((Ph, Pl)), P_ref) < 1e-5


def test_q_projection_with_zscale_and_per_frame_merge_equals_attention_readout():
    """OPB_LIN_ZSCALE + OPB_LIN_WBATCHED + OPB_LIN_COLSUM against the reference read-out (:97-98,119)."""
    B, N = 2, 300
    g = torch.Generator().manual_seed(9)
    X = (torch.randn(B, N, 256, generator=g) * 0.2).cuda()
    Wq = (torch.randn(256, 256, generator=g) / 16).cuda(); bq = (torch.randn(256, generator=g) * 0.1).cuda()
    ksum = (torch.rand(B, 256, generator=g) * 50 + 1).cuda()
    P = (torch.randn(B, 256, 256, generator=g) / 16).cuda(); bm = (torch.randn(256, generator=g) * 0.1).cuda()
    _, Qp, st = _linear_tc(X, Wq, bq, want_c=False, want_planes=True, flags=1 | 16 | 32, elu_cols=256, ksum=ksum, zscale_num=777.0)
    q1 = torch.nn.functional.elu(X.double() @ Wq.double().t() + bq.double()) + 1
    den = (q1.view(B, N, 4, 64) * ksum.double().view(B, 1, 4, 64)).sum(-1, keepdim=True) + 1e-6
    # COLSUM sums the stored value (after the ZSCALE division) over the valid rows
    q2 = (q1.view(B, N, 4, 64) * 777.0 / den).reshape(B, N, 256)
    assert _rel(_unsplit(Qp), q2) < 5e-6
    np.testing.assert_allclose(st[..., 0].double().sum(1).cpu().numpy(), q2.sum(1).cpu().numpy(), rtol=1e-5)
    msg, _, _ = _linear_tc(q2.float(), P, bm, flags=64)
    ref = torch.einsum("bnk,bok->bno", q2, P.double()) + bm.double()
    assert _rel(msg, ref) < 5e-6


@pytest.mark.parametrize("shape", [(2, 333, True), (1, 2048, True), (3, 50, False)])
def test_mlp_with_folded_merge_and_analytic_instancenorm_mean(shape):
    """merge (:119) + MLP (:135-147) + residual (:78) the way the matcher runs them: column sums -> op_mlp_center_bias ->
    layer 0 = W0a x + P'_b Q'' with BIASB | CSTATS_RELU (P'_b = W0b P_b per frame) -> op_instnorm_scale_weight_planes ->
    layer 1 with RESIDUAL | COLSUM | WBATCHED; against merge + Conv1d + InstanceNorm1d + ReLU + Conv1d in float64."""
    from onepose_b200._lib import call, lib, ptr, stream_ptr
    B, N, batched = shape
    g = torch.Generator().manual_seed(31 + N)
    x = (torch.randn(B, N, 256, generator=g) * 0.3 + torch.randn(B, 1, 256, generator=g) * 0.2).cuda()
    q = (torch.rand(B, N, 256, generator=g) * 2).cuda()                     # Q'' > 0
    Pb = (torch.randn(B if batched else 1, 256, 256, generator=g) / 16).cuda(); bm = torch.randn(256, generator=g).cuda()
    W0 = (torch.randn(512, 512, generator=g) / 512 ** 0.5).cuda(); b0 = torch.randn(512, generator=g).cuda()
    W1 = (torch.randn(256, 512, generator=g) / 512 ** 0.5).cuda(); b1 = (torch.randn(256, generator=g) * 0.1).cuda()
    W0a = W0[:, :256].contiguous()
    Pp = (W0[:, 256:].double() @ Pb.double()).float().contiguous()           # P'_b [B or 1, 512, 256]
    parts = lib.op_linear_tc_stat_parts(N)
    xs = torch.full((B, parts, 256, 2), float("nan"), device="cuda"); qs = torch.full_like(xs, float("nan"))
    call("op_colsum_partial", ptr(x), B, N, 256, ptr(xs), stream_ptr())
    call("op_colsum_partial", ptr(q), B, N, 256, ptr(qs), stream_ptr())
    np.testing.assert_allclose(xs[..., 0].double().sum(1).cpu().numpy(), x.double().sum(1).cpu().numpy(), rtol=1e-5, atol=1e-4)
    W0aT = W0a.t().contiguous()
    pp = _planes(Pp)
    cb = torch.empty(B, 512, device="cuda")
    call("op_mlp_center_bias", ptr(xs), parts, ptr(qs), parts, ptr(W0aT), ptr(pp[0]), ptr(pp[1]), int(batched), B, N, ptr(cb), stream_ptr())
    msg = q.double() @ Pb.double().transpose(1, 2) + bm.double()             # the reference's merged message
    h = torch.cat([x.double(), msg], 2) @ W0.double().t() + b0.double()
    mean_lin = x.double().mean(1) @ W0a.double().t() + torch.einsum("bcj,bj->bc", Pp.double().expand(B, 512, 256), q.double().mean(1))
    np.testing.assert_allclose(cb.double().cpu().numpy(), -mean_lin.cpu().numpy(), rtol=1e-5, atol=1e-5)
    # layer 0: centred pre-activation (bm and b0 cancel in InstanceNorm: the kernel never adds them), statistics, ReLU, planes out
    a1, a2, w, w2 = _planes(x), _planes(q), _planes(W0a), pp
    hn = (torch.empty(B, N, 512, dtype=torch.float16, device="cuda"), torch.empty(B, N, 512, dtype=torch.float16, device="cuda"))
    st = torch.full((B, parts, 512, 2), float("nan"), device="cuda")
    call("op_linear_tc", ptr(a1[0]), ptr(a1[1]), 256, ptr(a2[0]), ptr(a2[1]), 256, ptr(w[0]), ptr(w[1]), ptr(w2[0]), ptr(w2[1]), ptr(cb), None,
         ptr(hn[0]), ptr(hn[1]), B, N, 512, 128 | 256 | (512 if batched else 0), 0, ptr(st), None, 1.0, stream_ptr())
    torch.cuda.synchronize()
    d = h - h.mean(1, keepdim=True)
    assert ((_unsplit(hn) - d.clamp_min(0)).abs().max() / d.abs().max()).item() < 5e-6
    np.testing.assert_allclose(st[..., 1].double().sum(1).cpu().numpy(), (d * d).sum(1).cpu().numpy(), rtol=2e-5)
    assert float(st[..., 0].double().sum(1).abs().max()) < 1e-3 * N ** 0.5 + 1e-2      # residual mean is rounding noise
    mean = torch.empty(B, 512, device="cuda"); rstd = torch.empty(B, 512, device="cuda")
    call("op_instnorm_finalize_tc", ptr(st), None, B, parts, 512, N, 1e-5, ptr(mean), ptr(rstd), stream_ptr())
    rstd_ref = 1.0 / torch.sqrt(d.var(1, unbiased=False) + 1e-5)
    np.testing.assert_allclose(rstd.double().cpu().numpy(), rstd_ref.cpu().numpy(), rtol=2e-5)
    W1h = torch.empty(B, 256, 512, dtype=torch.float16, device="cuda"); W1l = torch.empty_like(W1h)
    call("op_scale_weight_planes", ptr(W1), ptr(rstd), B, 256, 512, ptr(W1h), ptr(W1l), stream_ptr())
    assert _rel(_unsplit((W1h, W1l)), W1.double()[None] * rstd.double()[:, None, :]) < 2e-6
    W2h = torch.zeros_like(W1h); W2l = torch.zeros_like(W1l); rstd2 = torch.zeros_like(rstd)       # the fused form of the two
    call("op_instnorm_scale_weight_planes", ptr(st), B, parts, N, 1e-5, ptr(W1), 256, 512, ptr(W2h), ptr(W2l), ptr(rstd2), stream_ptr())
    np.testing.assert_allclose(rstd2.cpu().numpy(), rstd.cpu().numpy(), rtol=1e-6)
    assert _rel(_unsplit((W2h, W2l)), _unsplit((W1h, W1l))) < 1e-6
    # layer 1 through the raw entry point (per-frame weight planes are already split)
    xo = x.clone(); xo_p = (torch.empty(B, N, 256, dtype=torch.float16, device="cuda"), torch.empty(B, N, 256, dtype=torch.float16, device="cuda"))
    call("op_linear_tc", ptr(hn[0]), ptr(hn[1]), 512, None, None, 0, ptr(W2h), ptr(W2l), None, None, ptr(b1), ptr(xo), ptr(xo_p[0]), ptr(xo_p[1]),
         B, N, 256, 2 | 16 | 64, 0, ptr(xs), None, 1.0, stream_ptr())
    torch.cuda.synchronize()
    ref = x.double() + torch.relu(d * rstd_ref[:, None, :]) @ W1.double().t() + b1.double()
    assert _rel(xo, ref) < 1e-5
    assert _rel(_unsplit(xo_p), ref) < 1e-5
    np.testing.assert_allclose(xs[..., 0].double().sum(1).cpu().numpy(), ref.sum(1).cpu().numpy(), rtol=1e-4, atol=1e-3)


def test_linear_tc_second_weight_equals_concatenated_weight():
    """A1 . W^T + A2 . W2^T (own tensor map for the A2 part) against the single [W | W2] matrix, bit for bit."""
    B, N = 2, 200
    g = torch.Generator().manual_seed(77)
    A1 = (torch.randn(B, N, 256, generator=g) * 0.3).cuda(); A2 = (torch.randn(B, N, 256, generator=g) * 0.3).cuda()
    W = (torch.randn(512, 256, generator=g) / 16).cuda(); W2 = (torch.randn(512, 256, generator=g) / 16).cuda()
    bias = torch.randn(512, generator=g).cuda()
    C1, _, _ = _linear_tc(A1, W, bias, A2=A2, W2=W2)
    C2, _, _ = _linear_tc(A1, torch.cat([W, W2], 1).contiguous(), bias, A2=A2)
    assert torch.equal(C1, C2)
    W2b = torch.stack([W2, -W2]).contiguous()
    C3, _, _ = _linear_tc(A1, W, bias, A2=A2, W2=W2b, flags=512)
    ref = A1.double() @ W.double().t() + torch.einsum("bnk,bok->bno", A2.double(), W2b.double()) + bias.double()
    assert _rel(C3, ref) < 5e-6


def test_residual_onto_the_output_planes_in_place():
    """OPB_LIN_RESIDUAL without an fp32 output: the old value is read from the output planes (hi + lo) and the sum written back."""
    from onepose_b200._lib import call, ptr, stream_ptr
    B, N = 2, 333
    g = torch.Generator().manual_seed(12)
    A = (torch.randn(B, N, 512, generator=g) * 0.3).cuda(); W = (torch.randn(256, 512, generator=g) / 512 ** 0.5).cuda()
    bias = (torch.randn(256, generator=g) * 0.1).cuda(); x0 = torch.randn(B, N, 256, generator=g).cuda()
    a, w, xp = _planes(A), _planes(W), _planes(x0)
    x0q = _unsplit(xp)                                        # what the planes actually hold (22 bits)
    call("op_linear_tc", ptr(a[0]), ptr(a[1]), 512, None, None, 0, ptr(w[0]), ptr(w[1]), None, None, ptr(bias), None, ptr(xp[0]), ptr(xp[1]),
         B, N, 256, 2, 0, None, None, 1.0, stream_ptr())
    torch.cuda.synchronize()
    ref = x0q + A.double() @ W.double().t() + bias.double()
    assert _rel(_unsplit(xp), ref) < 5e-6


@pytest.mark.parametrize("inc,add,shared,n3", [(1, 0, True, 100), (1, 1, True, 64), (0, 0, False, 37), (1, 0, False, 8)])
def test_gats_on_planes_vs_fp32_kernel(inc, add, shared, n3):
    """op_gats_leaf_aggregate_planes (in place on the planes, column sums out) against the fp32 kernel (GATs.py:62-123)."""
    from onepose_b200._lib import call, lib, ptr, stream_ptr
    B, L = 3, 6
    g = torch.Generator().manual_seed(n3)
    leaves = torch.randn(1 if shared else B, n3, L, 256, generator=g).cuda()
    h3 = torch.randn(B, n3, 256, generator=g).cuda()
    v1 = (torch.randn(256, generator=g) / 16).cuda(); v2 = (torch.randn(256, generator=g) / 16).cuda()
    hp = _planes(h3)
    h3q = _unsplit(hp).float()                                # the fp32 kernel sees exactly what the planes hold
    ref = torch.empty_like(h3q)
    bstride = 0 if shared else n3 * L * 256
    call("op_gats_leaf_aggregate", ptr(leaves), bstride, ptr(h3q), ptr(h3q) if (add and inc) else None, ptr(v1), ptr(v2), ptr(ref), B, n3, L,
         0.2, inc, 1, stream_ptr())
    parts = lib.op_gats_planes_parts(n3)
    cs = torch.full((B, parts, 256, 2), float("nan"), device="cuda")
    call("op_gats_leaf_aggregate_planes", ptr(leaves), bstride, ptr(hp[0]), ptr(hp[1]), ptr(v1), ptr(v2), ptr(cs), B, n3, L, 0.2, inc,
         int(add and inc), stream_ptr())
    torch.cuda.synchronize()
    assert _rel(_unsplit(hp), ref.double()) < 2e-6
    np.testing.assert_allclose(cs[..., 0].double().sum(1).cpu().numpy(), ref.double().sum(1).cpu().numpy(), rtol=1e-5, atol=1e-4)
    assert float(cs[..., 1].abs().max()) == 0.0


@pytest.mark.parametrize("flags", [dict(with_linear_transform=True), dict(include_self=False), dict(additional=True),
                                   dict(with_linear_transform=True, additional=True)])
def test_matcher_gats_variants_tensor_core_vs_cuda_core(flags):
    """The non-shipped GATs flags (GATs.py:23-31) through both paths: planes-only streams without the linear transform,
    an fp32 copy of the streams with it."""
    from onepose_b200.matcher import GATsSuperGlue
    from tests import helpers
    hp = dict(synth.DEFAULT_HPARAMS); hp.update(flags)
    obj = synth.make_object(4, 300, 5)
    frames = [synth.make_frame(obj, f, 200) for f in range(2)]
    data = {k: v.cuda() for k, v in synth.pack_batch(obj, frames).items()}
    m = GATsSuperGlue(hp).cuda().eval()
    m.load_state_dict(synth.make_weights(3, damped=True), strict=True)
    pred_tc, conf_tc = m(data)
    conf_tc = conf_tc.clone(); m0_tc = pred_tc["matches0_batched"].clone()
    m.use_tensor_cores = False
    pred_cc, conf_cc = m(data)
    assert helpers.conf_rel_err(conf_tc.cpu(), conf_cc.cpu()) < helpers.CONF_RTOL
    for b in range(2):
        helpers.assert_matches_equal_off_ties(m0_tc[b].cpu(), pred_cc["matches0_batched"][b].cpu(), conf_cc[b].cpu(), axis=2)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(2, 200), (3, 2048), (1, 33)])
def test_channel_major_to_planes_equals_the_three_kernel_path(shape):
    """op_channel_major_to_planes = op_channel_major_to_point_major + op_split_planes + op_colsum_partial, bit for bit."""
    from onepose_b200._lib import call, lib, ptr, stream_ptr
    B, N = shape
    g = torch.Generator().manual_seed(N)
    x = torch.randn(B, 256, N, generator=g).cuda()
    parts = lib.op_linear_tc_stat_parts(N)
    pm = torch.empty(B, N, 256, device="cuda")
    call("op_channel_major_to_point_major", ptr(x), ptr(pm), B, 256, N, x.stride(0), stream_ptr())
    h0 = torch.empty(B, N, 256, dtype=torch.float16, device="cuda"); l0 = torch.empty_like(h0)
    call("op_split_planes", ptr(pm), ptr(h0), ptr(l0), pm.numel(), stream_ptr())
    cs0 = torch.full((B, parts, 256, 2), float("nan"), device="cuda")
    call("op_colsum_partial", ptr(pm), B, N, 256, ptr(cs0), stream_ptr())
    h1 = torch.empty_like(h0); l1 = torch.empty_like(h0)
    cs1 = torch.full((B, parts, 256, 2), float("nan"), device="cuda")
    call("op_channel_major_to_planes", ptr(x), x.stride(0), ptr(h1), ptr(l1), ptr(cs1), B, N, stream_ptr())
    torch.cuda.synchronize()
    assert torch.equal(h0, h1) and torch.equal(l0, l1) and torch.equal(cs0, cs1)


def test_final_projection_with_fused_normalize_writes_similarity_planes():
    """OPB_LIN_UNIT_PLANES8: final_proj (GATs_SuperGlue.py:241) + F.normalize (:245-246) in the GEMM epilogue, emitted as the operand
    planes of op_sim_dual_softmax_planes: hi = fp16(64 u), lo = (64 u - hi) as e5m2, hi again as e4m3."""
    from onepose_b200._lib import call, ptr, stream_ptr
    g = torch.Generator().manual_seed(11)
    B, N = 2, 333
    x = torch.randn(B, N, 256, generator=g).cuda()
    x[1, 5] = 0.0
    W = (torch.randn(256, 256, generator=g) / 16).cuda()
    bias = (torch.randn(256, generator=g) * 0.1).cuda()
    split = lambda t: (torch.empty_like(t, dtype=torch.float16), torch.empty_like(t, dtype=torch.float16))
    xp, wp = split(x), split(W)
    call("op_split_planes", ptr(x), ptr(xp[0]), ptr(xp[1]), x.numel(), stream_ptr())
    call("op_split_planes", ptr(W), ptr(wp[0]), ptr(wp[1]), W.numel(), stream_ptr())
    uh = torch.zeros(B, N, 256, dtype=torch.float16, device="cuda")
    ub = torch.zeros(2, B, N, 256, dtype=torch.uint8, device="cuda")
    call("op_linear_tc", ptr(xp[0]), ptr(xp[1]), 256, None, None, 0, ptr(wp[0]), ptr(wp[1]), None, None, ptr(bias), None, ptr(uh), ptr(ub),
         B, N, 256, 16384, 0, None, None, 1.0, stream_ptr())
    torch.cuda.synchronize()
    ref = torch.nn.functional.normalize(x.double() @ W.double().t() + bias.double(), dim=2) * 64.0
    hi = uh.double()
    assert float((hi - ref).abs().max()) <= 2.0 ** -11 * 64 * 1.01                   # fp16 rounding of |64 u| <= 64
    lo = ub[0].view(torch.float8_e5m2).double()
    assert float((hi + lo - ref).abs().max()) < 6e-4                                  # e5m2 lo: 2 mantissa bits of a remainder <= 2^-12 * 16
    hx = ub[1].view(torch.float8_e4m3fn).double()
    assert float((hx - hi).abs().max()) <= 2.0 ** -4 * 64 * 1.01                     # e4m3 copy of hi: 3 mantissa bits
    # unit rows: |hi + lo| / 64 = 1 (the all-zero input row has the bias as its projection, so it is a unit row too)
    nrm = ((hi + lo) / 64.0).norm(dim=2)
    assert float((nrm - 1).abs().max()) < 1e-4
    # the planes feed the similarity op: same conf as the fp32-rows entry on the unnormalised projection
    m = (x.double() @ W.double().t() + bias.double()).float()
    from tests.test_gpu_matcher import _run_sim
    rows = _run_sim("op_sim_dual_softmax_rows", m[:, :200].contiguous(), m[:, 200:].contiguous())
    conf = torch.empty(B, 200, 133, device="cuda")
    i0 = torch.empty(B, 200, dtype=torch.int64, device="cuda"); i1 = torch.empty(B, 133, dtype=torch.int64, device="cuda")
    s0 = torch.empty(B, 200, device="cuda"); s1 = torch.empty(B, 133, device="cuda")
    from onepose_b200._lib import lib
    nbytes = lib.op_sim_dual_softmax_workspace_bytes(B, 200, 133)
    ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    a_h, a_b = uh[:, :200].contiguous(), ub[:, :, :200].contiguous()
    b_h, b_b = uh[:, 200:].contiguous(), ub[:, :, 200:].contiguous()
    call("op_sim_dual_softmax_planes", ptr(a_h), ptr(a_b), ptr(b_h), ptr(b_b), B, 200, 133, 0.07, 0.2, ptr(conf), ptr(i0), ptr(i1), ptr(s0), ptr(s1),
         ptr(ws), nbytes, stream_ptr())
    torch.cuda.synchronize()
    rel = ((conf - rows[0]).abs() / rows[0].clamp_min(1e-30)).max().item()
    assert rel < 2e-4, rel
