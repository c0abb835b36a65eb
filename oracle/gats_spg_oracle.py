"""CPU oracle for the GATsSPG 2D-3D matcher hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``onepose_b200/`` may import this file; it is the
checker used by ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py``.

It is a functional restatement (plain torch CPU ops on point-major ``[B, N, C]`` arrays, no
``nn.Module``) of the algorithm in the reference files

* ``src/models/GATsSPG_architectures/GATs_SuperGlue.py:67-99,112-147,203-278``
* ``src/models/GATsSPG_architectures/GATs.py:62-123``

Parity pin: ``tests/golden/make_golden.py`` imports the unmodified reference from
``/root/reference`` in the build container, runs it on the seeded inputs of
``onepose_b200.synth`` and stores its outputs under ``tests/golden/``;
``tests/test_oracle_golden.py`` checks this file against those fixtures (and against the live
reference whenever ``/root/reference`` exists).  The reference itself ships no tests or golden
vectors, so the fixtures generated from the live reference are the pin.

Weights are a flat ``dict`` keyed by the reference ``state_dict`` names
(``gnn.layers.{i}.W``, ``gnn.layers.{i}.attn.proj.{0,1,2}.weight`` ...).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

LAYER_KINDS = ("GATs", "self", "cross") * 4  # GATs_SuperGlue.py:184


def _pw(x, weight, bias):
    """1x1 Conv1d on a point-major array: x [B,N,Ci], weight [Co,Ci,1] -> [B,N,Co].

    Restates nn.Conv1d(kernel_size=1) as used at GATs_SuperGlue.py:108-109,140-142,194-199.
    """
    return torch.matmul(x, weight[:, :, 0].t()) + bias


def gats_leaf_aggregate(leaves, h3, W, a, *, alpha=0.2, include_self=True,
                        with_linear_transform=False, additional=False, concat=True):
    """Leaf-aggregation graph attention (GATs.py:62-123).

    leaves [B, N3, L, D] (leaf j of point n = column n*L+j of descriptors2d_db, GATs.py:76),
    h3 [B, N3, D].  Returns [B, N3, D].
    """
    B, N3, L, D = leaves.shape
    wh_leaf = torch.matmul(leaves, W)                     # GATs.py:68
    wh_3d = torch.matmul(h3, W)                           # GATs.py:69
    d_out = W.shape[1]
    s_leaf = torch.matmul(wh_leaf, a[:d_out])             # GATs.py:113  [B,N3,L,1]
    t_3d = torch.matmul(wh_3d, a[d_out:])                 # GATs.py:115  [B,N3,1]
    if include_self:
        s_all = torch.cat([t_3d.unsqueeze(2), s_leaf], dim=2)   # GATs.py:117-120 (self term uses a[256:])
    else:
        s_all = s_leaf
    e = F.leaky_relu(t_3d.unsqueeze(2) + s_all, alpha)    # GATs.py:122-123
    att = torch.softmax(e, dim=2)                         # GATs.py:73  [B,N3,(1+)L,1]
    if include_self:
        feats = torch.cat([(wh_3d if with_linear_transform else h3).unsqueeze(2),
                           wh_leaf if with_linear_transform else leaves], dim=2)
        out = (att * feats).sum(dim=2)                    # GATs.py:88-91
        if additional:
            out = out + h3                                # GATs.py:93-94
    else:
        if with_linear_transform:
            out = (att * wh_leaf).sum(dim=2) / 2.0 + wh_3d    # GATs.py:97
        else:
            out = (att * leaves).sum(dim=2) / 2.0 + h3        # GATs.py:99
    return F.elu(out) if concat else out                  # GATs.py:101-104


def linear_attention(q, k, v):
    """elu+1 linear attention (GATs_SuperGlue.py:88-99) on point-major heads.

    q [B,N,dh,H], k,v [B,M,dh,H]  ->  [B,N,dh,H]
    """
    eps = 1e-6
    q = F.elu(q) + 1
    k = F.elu(k) + 1
    m = v.shape[1]
    v = v / m
    kv = torch.einsum("bmdh,bmqh->bqdh", k, v)            # :96
    z = 1 / (torch.einsum("bndh,bdh->bnh", q, k.sum(1)) + eps)   # :97
    return torch.einsum("bndh,bqdh,bnh->bnqh", q, kv, z) * m      # :98


def attention_propagation(x, src, w, prefix, num_heads=4):
    """AttentionPropagation.forward (GATs_SuperGlue.py:112-120,130-132,135-147).

    x [B,N,D] attends to src [B,M,D]; returns the residual delta [B,N,D].
    """
    B, N, D = x.shape
    dh = D // num_heads
    g = lambda name: w[prefix + name]
    # channel c of the conv output is head (c % H), dim (c // H): .view(b, dim, heads, -1) at :116
    q = _pw(x, g("attn.proj.0.weight"), g("attn.proj.0.bias")).view(B, N, dh, num_heads)
    k = _pw(src, g("attn.proj.1.weight"), g("attn.proj.1.bias")).view(B, -1, dh, num_heads)
    v = _pw(src, g("attn.proj.2.weight"), g("attn.proj.2.bias")).view(B, -1, dh, num_heads)
    msg = linear_attention(q, k, v).reshape(B, N, D)
    msg = _pw(msg, g("attn.merge.weight"), g("attn.merge.bias"))       # :120
    h = _pw(torch.cat([x, msg], dim=2), g("mlp.0.weight"), g("mlp.0.bias"))   # :132,141
    # InstanceNorm1d(512): per sample, per channel over the point axis, biased var, eps 1e-5 (:144-145)
    mean = h.mean(dim=1, keepdim=True)
    var = h.var(dim=1, unbiased=False, keepdim=True)
    h = torch.relu((h - mean) / torch.sqrt(var + 1e-5))
    return _pw(h, g("mlp.3.weight"), g("mlp.3.bias"))


def gnn_forward(desc2d, desc3d, leaves, w, hparams):
    """AttentionalGNN.forward (GATs_SuperGlue.py:67-85).  All arrays point-major."""
    for i, kind in enumerate(LAYER_KINDS):
        p = f"gnn.layers.{i}."
        if kind == "GATs":
            desc3d = gats_leaf_aggregate(
                leaves, desc3d, w[p + "W"], w[p + "a"], alpha=0.2,
                include_self=hparams["include_self"],
                with_linear_transform=hparams["with_linear_transform"],
                additional=hparams["additional"])
        else:
            src0, src1 = (desc3d, desc2d) if kind == "cross" else (desc2d, desc3d)   # :76,81
            d0 = attention_propagation(desc2d, src0, w, p)
            d1 = attention_propagation(desc3d, src1, w, p)
            desc2d, desc3d = desc2d + d0, desc3d + d1
    return desc2d, desc3d


def dual_softmax_select(m2, m3, scale_factor, match_threshold):
    """Similarity, dual-softmax, mutual-NN + threshold (GATs_SuperGlue.py:251-267).

    m2 [B,N2,D], m3 [B,N3,D] unit rows.  Returns batched indices/scores and conf [B,N2,N3].
    """
    scores = torch.matmul(m2, m3.transpose(1, 2)) / scale_factor
    conf = torch.softmax(scores, 1) * torch.softmax(scores, 2)
    max0, max1 = conf.max(2), conf.max(1)
    i0, i1 = max0.indices, max1.indices
    ar0 = torch.arange(i0.shape[1])[None]
    ar1 = torch.arange(i1.shape[1])[None]
    mutual0 = ar0 == i1.gather(1, i0)
    mutual1 = ar1 == i0.gather(1, i1)
    zero = conf.new_tensor(0)
    ms0 = torch.where(mutual0, max0.values, zero)
    ms1 = torch.where(mutual1, ms0.gather(1, i1), zero)
    valid0 = mutual0 & (ms0 > match_threshold)
    valid1 = mutual1 & valid0.gather(1, i1)
    i0 = torch.where(valid0, i0, i0.new_tensor(-1))
    i1 = torch.where(valid1, i1, i1.new_tensor(-1))
    return i0, i1, ms0, ms1, conf


def to_point_major(data, dtype=torch.float32):
    """Public channel-major dict (GATs_SuperGlue.py:209-218) -> point-major arrays."""
    d2 = data["descriptors2d_query"].to(dtype).transpose(1, 2).contiguous()     # [B,N2,D]
    d3 = data["descriptors3d_db"].to(dtype).transpose(1, 2).contiguous()        # [B,N3,D]
    lv = data["descriptors2d_db"].to(dtype).transpose(1, 2).contiguous()        # [B,N3*L,D]
    B, N3, D = d3.shape
    L = lv.shape[1] // N3                                                      # GATs.py:65
    return d2, d3, lv.view(B, N3, L, D)


def matcher_forward(data, w, hparams, dtype=torch.float32, return_batched=False):
    """GATsSuperGlue.forward (GATs_SuperGlue.py:203-278) restated.

    Returns ``(pred, conf_matrix)``; ``pred`` holds batch element 0 only, as the reference does
    (:269-274).  Empty inputs return the single early-out dict (:223-231).
    ``dtype=torch.float64`` gives the accuracy yard-stick used in the precision tests.
    """
    k2, k3 = data["keypoints2d"], data["keypoints3d"]
    if k2.shape[1] == 0 or k3.shape[1] == 0:
        return {
            "matches0": torch.full(k2.shape[:-1], -1, dtype=torch.int)[0],
            "matches1": torch.full(k3.shape[:-1], -1, dtype=torch.int)[0],
            "matching_scores0": torch.zeros(k2.shape[:-1])[0],
            "matching_scores1": torch.zeros(k3.shape[:-1])[0],
            "skip_train": True,
        }
    if hparams["match_type"] != "softmax":
        raise NotImplementedError
    w = {k: v.to(dtype) for k, v in w.items() if v.is_floating_point()}
    d2, d3, leaves = to_point_major(data, dtype)
    d2, d3 = gnn_forward(d2, d3, leaves, w, hparams)
    m2 = F.normalize(_pw(d2, w["final_proj.weight"], w["final_proj.bias"]), p=2, dim=2)   # :241-246
    m3 = F.normalize(_pw(d3, w["final_proj.weight"], w["final_proj.bias"]), p=2, dim=2)
    i0, i1, ms0, ms1, conf = dual_softmax_select(
        m2, m3, hparams["scale_factor"], hparams["match_threshold"])
    pred = {"matches0": i0[0], "matches1": i1[0],
            "matching_scores0": ms0[0], "matching_scores1": ms1[0]}
    if return_batched:
        pred.update(matches0_batched=i0, matches1_batched=i1,
                    matching_scores0_batched=ms0, matching_scores1_batched=ms1,
                    mdesc2d=m2, mdesc3d=m3)
    return pred, conf


def gather_matches(matches0, scores0, kpts2d, kpts3d):
    """Valid-match gather of the caller (inference.py:147-152; eval_utils.py:72-77)."""
    valid = matches0 > -1
    return kpts2d[valid], kpts3d[matches0[valid]], scores0[valid]
