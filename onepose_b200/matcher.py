"""Drop-in ``GATsSuperGlue`` / ``LitModelGATsSPG`` running on the sm_100a CUDA library.

Mirrors the reference interface
(``src/models/GATsSPG_architectures/GATs_SuperGlue.py:162-278``,
``src/models/GATsSPG_lightning_model.py:15-37``): same constructor hparams, same parameter names
and shapes (a reference ``state_dict`` loads with ``strict=True``, including the never-used
``kenc_2d`` / ``kenc_3d`` / ``bin_score``), same ``forward(data) -> (pred, conf_matrix)`` and the same
single-dict early-out for empty inputs.  The modules below only *hold* parameters; every arithmetic
step of ``forward`` is a call into ``libonepose_b200.so`` through the C ABI of
``include/onepose_b200.h``.  There is no CPU path: tensors must live on a CUDA device.
"""
from __future__ import annotations

import copy

import os

import torch
import torch.nn as nn

from . import _lib
from ._lib import call, lib, ptr, stream_ptr, timeline

GNN_LAYER_KINDS = ("GATs", "self", "cross") * 4      # GATs_SuperGlue.py:184
_D = 256
_HEADS = 4


def _mlp_holder(channels):
    """Parameter holder with the module indices of the reference ``MLP()`` (Conv, InstanceNorm, ReLU, ...)."""
    mods = []
    for i in range(1, len(channels)):
        mods.append(nn.Conv1d(channels[i - 1], channels[i], kernel_size=1, bias=True))
        if i < len(channels) - 1:
            mods.append(nn.InstanceNorm1d(channels[i]))
            mods.append(nn.ReLU())
    return nn.Sequential(*mods)


class _KeypointEncoderParams(nn.Module):
    # constructed by the reference (GATs_SuperGlue.py:172-182) but never called in forward
    def __init__(self, inp_dim, feature_dim, layers):
        super().__init__()
        self.encoder = _mlp_holder([inp_dim] + list(layers) + [feature_dim])
        nn.init.constant_(self.encoder[-1].bias, 0.0)


class _GATsParams(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.W = nn.Parameter(torch.empty(dim, dim))
        nn.init.xavier_normal_(self.W.data, gain=1.414)
        self.a = nn.Parameter(torch.empty(2 * dim, 1))
        nn.init.xavier_normal_(self.a.data, gain=1.414)


class _AttnParams(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.merge = nn.Conv1d(dim, dim, kernel_size=1)
        # three copies of merge, no extra RNG draws (as :110), so seeded construction matches the reference
        self.proj = nn.ModuleList([copy.deepcopy(self.merge) for _ in range(3)])


class _AttnPropParams(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.attn = _AttnParams(dim)
        self.mlp = _mlp_holder([2 * dim, 2 * dim, dim])
        nn.init.constant_(self.mlp[-1].bias, 0.0)


class _GnnParams(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.layers = nn.ModuleList(
            [_GATsParams(dim) if i % 3 == 0 else _AttnPropParams(dim) for i in range(len(GNN_LAYER_KINDS))])


def _head_major_perm(device):
    """new channel h*64+d  <-  reference channel d*4+h   (the `.view(b, 64, 4, n)` split, :116)."""
    c = torch.arange(_D, device=device)
    return (c % (_D // _HEADS)) * _HEADS + c // (_D // _HEADS)


class _Workspace:
    """Device buffers reused across calls, keyed by shape (caller-owned memory of the C ABI)."""

    def __init__(self):
        self._bufs = {}

    def get(self, name, shape, dtype, device):
        key = (name, tuple(shape), dtype, str(device))
        t = self._bufs.get(key)
        if t is None:
            # keep at most two shapes per name (the per-frame and the shared single-frame use of a buffer)
            same = [k for k in self._bufs if k[0] == name]
            for k in same[:-1]:
                del self._bufs[k]
            t = torch.empty(shape, dtype=dtype, device=device)
            self._bufs[key] = t
        return t


class GATsSuperGlue(nn.Module):
    """B200 drop-in for the reference ``GATsSuperGlue`` (GATs_SuperGlue.py:162-278)."""

    def __init__(self, hparams):
        super().__init__()
        self.hparams = hparams
        self.match_type = hparams["match_type"]
        D = hparams["descriptor_dim"]
        if D != _D:
            raise ValueError(f"descriptor_dim {D}: the sm_100a kernels are built for {_D}")
        self.kenc_2d = _KeypointEncoderParams(3, D, hparams["keypoints_encoder"])
        self.kenc_3d = _KeypointEncoderParams(4, D, hparams["keypoints_encoder"])
        self.gnn = _GnnParams(D)
        self.final_proj = nn.Conv1d(D, D, kernel_size=1, bias=True)
        self.register_parameter("bin_score", nn.Parameter(torch.tensor(1.0)))
        self._packed = None          # device weights in kernel layout
        self._packed_key = None
        self._ws = _Workspace()
        self._db_cache = {}          # static object DB in point-major layout
        self._trace = None           # debugging aid: list that receives (layer, x2, x3) clones
        # projections on the tcgen05 tensor cores (split-precision planes); False = exact-fp32 CUDA-core kernels,
        # kept as the on-device cross-check of the tensor-core path (tests), not as a fallback
        self.use_tensor_cores = True
        self.two_streams = True      # 2D / 3D streams of a layer on two CUDA streams (tensor-core path)
        self.fused_final_normalize = True   # final_proj + F.normalize write the similarity kernel's operand planes (False: fp32 rows, for A/B runs and tests)
        self._side = {}
        # Pipelines only (off by default: the drop-in contract returns a pose that is ready on the caller's stream).  True: gather +
        # PnP of a call run on the module's pose stream, so that the (latency-bound, few-CTA) RANSAC of one block of frames overlaps
        # the matcher of the next; ``pred["pose_event"]`` is recorded behind them and ``wait_pose(pred)`` makes the current stream
        # wait for it.  Work that consumes the pose can also be enqueued on ``pose_stream(device)`` directly.
        self.pnp_overlap = False
        self._pose_streams = {}
        # CUDA graphs (enable_graphs): one captured graph per (device, shapes, object DB, weights) so that a forward is a few host calls
        # instead of ~150 launches from Python -- see _forward_graphed
        self._graphs_on = False
        self._graphs = {}
        self.pnp_refine = None       # None = cv2's result (parity mode); "lm" = optional Levenberg-Marquardt polish (pnp.py)

    # ------------------------------------------------------------------ weights in kernel layout
    def _weights_key(self):
        return tuple((p.data_ptr(), p._version) for p in self.parameters())

    def _pack_weights(self, device):
        key = (str(device),) + self._weights_key()
        if self._packed is not None and self._packed_key == key:
            return self._packed
        perm = _head_major_perm(device)
        f = lambda t: t.detach().to(device=device, dtype=torch.float32).contiguous()
        packed = []
        for i, kind in enumerate(GNN_LAYER_KINDS):
            lay = self.gnn.layers[i]
            if kind == "GATs":
                W = lay.W.detach().to(device=device, dtype=torch.float64)
                a = lay.a.detach().to(device=device, dtype=torch.float64)[:, 0]
                packed.append(dict(
                    kind=kind,
                    v1=(W @ a[:_D]).float().contiguous(), v2=(W @ a[_D:]).float().contiguous(),   # W collapsed through a
                    a1=a[:_D].float().contiguous(), a2=a[_D:].float().contiguous(),
                    Wt=f(lay.W.t())))                          # h @ W  ==  linear with weight W^T
            else:
                at, mlp = lay.attn, lay.mlp
                wq, wk, wv = (f(at.proj[j].weight[:, :, 0])[perm] for j in range(3))
                bq, bk, bv = (f(at.proj[j].bias)[perm] for j in range(3))
                packed.append(dict(
                    kind=kind, Wq=wq.contiguous(), bq=bq.contiguous(), Wk=wk.contiguous(), bk=bk.contiguous(),
                    Wv=wv.contiguous(), bv=bv.contiguous(),
                    WvT=wv.view(_HEADS, 64, _D).transpose(1, 2).contiguous(),            # [h, c, q] for op_kv_post
                    WmT=f(at.merge.weight[:, :, 0])[:, perm].view(_D, _HEADS, 64).permute(1, 2, 0).contiguous(),   # [h, q, o]
                    Wkv=torch.cat([wk, wv], 0).contiguous(), bkv=torch.cat([bk, bv], 0).contiguous(),
                    Wm=f(at.merge.weight[:, :, 0])[:, perm].contiguous(), bm=f(at.merge.bias),
                    W0=f(mlp[0].weight[:, :, 0]), b0=f(mlp[0].bias),
                    W1=f(mlp[3].weight[:, :, 0]), b1=f(mlp[3].bias)))
                # merge folded into MLP layer 0 (tensor-core path): h = W0a x + (W0b Wm) attn + const, and the constant
                # (W0b bm + b0) is removed by InstanceNorm1d -- op_kv_post builds P'_b = KV' (W0b Wm)^T per frame
                lw = packed[-1]
                w0 = mlp[0].weight[:, :, 0].detach().to(device=device, dtype=torch.float64)
                w0m = (w0[:, _D:] @ at.merge.weight[:, :, 0].detach().to(device=device, dtype=torch.float64))[:, perm]
                lw["W0mT"] = w0m.float().view(2 * _D, _HEADS, 64).permute(1, 2, 0).contiguous()       # [h, q, c]
                lw["W0a"] = f(w0[:, :_D])
                lw["W0aT"] = f(w0[:, :_D].t())                                                      # [k, c]
        self._packed = dict(layers=packed, Wf=f(self.final_proj.weight[:, :, 0]), bf=f(self.final_proj.bias))
        # split-precision fp16 planes of every projection weight for the tcgen05 path (linear_tc.cuh)
        with torch.cuda.device(device):
            for lw in packed:
                for k in ("Wq", "Wk", "W0a"):
                    if k in lw:
                        lw[k + "_p"] = self._split_new(lw[k])
            self._packed["Wf_p"] = self._split_new(self._packed["Wf"])
            # identity block of the in-GEMM residual add (_propagate_tc): built here, on the caller's stream, BEFORE forward forks
            # the side stream -- both streams' mlp1 launches read these planes, and a lazily built copy would be written on
            # whichever stream got there first with nothing ordering the other stream behind it
            self._packed["eye_p"] = self._split_new(torch.eye(_D, dtype=torch.float32, device=device))
        self._packed_key = key
        # cached W-transformed leaves depend on the weights; the point-major DB copies themselves (incl. a DB registered by
        # ObjectDB.register before the first forward) do not
        for _, entry in self._db_cache.values():
            entry["wh_leaf"] = {}
        return self._packed

    # ------------------------------------------------------------------ layout helpers
    @staticmethod
    def _to_point_major(x, out):
        """x [B, C, N] (any strides) -> out [B, N, C] fp32 through the CUDA transpose."""
        B, C, N = x.shape
        if x.stride(2) != 1 or x.stride(1) != N:
            x = x[:1].contiguous() if x.stride(0) == 0 else x.contiguous()
        bstride = x.stride(0) if x.shape[0] > 1 else 0
        call("op_channel_major_to_point_major", ptr(x), ptr(out), out.shape[0], C, N, bstride, stream_ptr())
        return out

    @staticmethod
    def _split(x, hi, lo):
        """fp32 tensor -> (hi, lo) fp16 planes of 64*x (op_split_planes)."""
        call("op_split_planes", ptr(x), ptr(hi), ptr(lo), x.numel(), stream_ptr())
        return hi, lo

    @classmethod
    def _split_new(cls, x):
        x = x.contiguous()
        return cls._split(x, torch.empty(x.shape, dtype=torch.float16, device=x.device),
                          torch.empty(x.shape, dtype=torch.float16, device=x.device))

    def _side_stream(self, dev):
        s = self._side.get(dev.index)
        if s is None:
            s = self._side[dev.index] = torch.cuda.Stream(device=dev)
        return s

    def pose_stream(self, dev):
        dev = torch.device(dev)
        s = self._pose_streams.get(dev.index)
        if s is None:
            # high priority: the few, long-running PnP blocks take the first SM resources that come free between the matcher's kernels
            # (OPB_POSE_PRIORITY=0 for A/B runs)
            s = self._pose_streams[dev.index] = torch.cuda.Stream(device=dev, priority=int(os.environ.get("OPB_POSE_PRIORITY", "-1")))
        return s

    @staticmethod
    def wait_pose(pred):
        """With ``pnp_overlap``: makes the current stream wait for the pose of this ``pred`` (no host synchronisation)."""
        ev = pred.get("pose_event")
        if ev is not None:
            cur = torch.cuda.current_stream(pred["pose"].device)
            cur.wait_event(ev)
            for k in _POSE_KEYS:
                if k in pred:
                    pred[k].record_stream(cur)
        return pred

    def _planes(self, name, shape, dev):
        return (self._ws.get(name + ".hi", shape, torch.float16, dev), self._ws.get(name + ".lo", shape, torch.float16, dev))

    def register_static_db(self, descriptors3d_db, descriptors2d_db, avg_pm, leaves_pm, num_leaf):
        """An object DB that already exists in the internal point-major layout (``onepose_b200.objectdb.ObjectDB``):
        ``forward`` calls whose DB tensors are (batch-expanded views of) descriptors3d_db [256,N3] / descriptors2d_db
        [256,N3*L] use avg_pm [N3,256] / leaves_pm [N3*L,256] directly."""
        N3 = avg_pm.shape[0]
        if leaves_pm.shape[0] != N3 * num_leaf or avg_pm.shape[1] != _D or leaves_pm.shape[1] != _D:
            raise ValueError("register_static_db: inconsistent DB shapes")
        entry = dict(d3=avg_pm[None].float().contiguous(), leaves=leaves_pm[None].float().contiguous(), shared=True, wh_leaf={},
                     _src=(descriptors3d_db, descriptors2d_db))
        self._db_cache["registered"] = (self._registered_key(descriptors3d_db, descriptors2d_db, N3, N3 * num_leaf) +
                                        (avg_pm._version, leaves_pm._version), entry)
        entry["_pm_src"] = (avg_pm, leaves_pm)

    @staticmethod
    def _registered_key(d3, lv, n3, nl):
        return (d3.data_ptr(), lv.data_ptr(), n3, nl, d3._version, lv._version)

    def _static_db(self, d3, lv, L):
        """Point-major copy of the object DB, cached while the caller keeps handing the same tensors."""
        shared = d3.shape[0] == 1 or (d3.stride(0) == 0 and lv.stride(0) == 0)
        reg = self._db_cache.get("registered")
        if reg is not None and shared and d3.stride(1) == 1 and lv.stride(1) == 1 and \
                reg[0] == self._registered_key(d3, lv, d3.shape[2], lv.shape[2]) + tuple(t._version for t in reg[1]["_pm_src"]):
            return reg[1]
        Bd = 1 if shared else d3.shape[0]
        key = (d3.data_ptr(), lv.data_ptr(), tuple(d3.shape), tuple(lv.shape), d3._version, lv._version, Bd)
        hit = self._db_cache.get("db")
        if hit is not None and hit[0] == key:
            return hit[1]
        N3 = d3.shape[2]
        d3_pm = torch.empty(Bd, N3, _D, dtype=torch.float32, device=d3.device)
        lv_pm = torch.empty(Bd, N3 * L, _D, dtype=torch.float32, device=d3.device)
        self._to_point_major(d3[:Bd], d3_pm)
        self._to_point_major(lv[:Bd], lv_pm)
        # the source tensors are kept alive so their addresses cannot be recycled under the cache key
        entry = dict(d3=d3_pm, leaves=lv_pm, shared=shared, wh_leaf={}, _src=(d3, lv))
        self._db_cache["db"] = (key, entry)
        return entry

    # ------------------------------------------------------------------ layer launchers
    def _linear(self, A1, W, bias, C, B, rows, *, A2=None, flags=0, elu_cols=0, mean=None, rstd=None, stats=None, tag=""):
        K1 = A1.shape[-1]
        K2 = 0 if A2 is None else A2.shape[-1]
        timeline.work = 2.0 * B * rows * (K1 + K2) * W.shape[0]           # algorithmic flops of this call
        call("op_linear", ptr(A1), K1, ptr(A2), K2, ptr(W), ptr(bias), ptr(C), B, rows, W.shape[0], flags,
                            elu_cols, ptr(mean), ptr(rstd), ptr(stats), stream_ptr(), tag=tag)

    def _kv(self, src, B, M, lw, tag):
        """K/V projection of the source stream + the KV / Ksum reduction (GATs_SuperGlue.py:90-97)."""
        dev = src.device
        kv = self._ws.get(f"kv{M}", (B, M, 2 * _D), torch.float32, dev)
        self._linear(src, lw["Wkv"], lw["bkv"], kv, B, M, flags=1, elu_cols=_D, tag="kv")      # OPB_LIN_ELU1 on K
        KV = self._ws.get("KV" + tag, (B, _HEADS, 64, 64), torch.float32, dev)
        Ks = self._ws.get("Ks" + tag, (B, _HEADS, 64), torch.float32, dev)
        nbytes = lib.op_kv_reduce_workspace_bytes(B, M)
        w = self._ws.get(f"kvws{M}", (nbytes,), torch.uint8, dev)
        call("op_kv_reduce", ptr(kv), ptr(KV), ptr(Ks), B, M, ptr(w), nbytes, stream_ptr())
        return KV, Ks

    def _propagate(self, x, B, N, M, KV, Ks, lw):
        """x-side of AttentionPropagation (GATs_SuperGlue.py:112-147); x [B,N,256] is updated in place."""
        dev = x.device
        q = self._ws.get(f"q{N}", (B, N, _D), torch.float32, dev)
        self._linear(x, lw["Wq"], lw["bq"], q, B, N, flags=1, elu_cols=_D, tag="q")
        att = self._ws.get(f"att{N}", (B, N, _D), torch.float32, dev)
        call("op_attn_apply", ptr(q), ptr(KV), ptr(Ks), ptr(att), B, N, M, stream_ptr())
        msg = q                                                     # q is dead: reuse as the merge output
        self._linear(att, lw["Wm"], lw["bm"], msg, B, N, tag="merge")
        h = self._ws.get(f"h{N}", (B, N, 2 * _D), torch.float32, dev)
        tiles = lib.op_linear_row_tiles(N)
        stats = self._ws.get(f"stats{N}", (B, tiles, 2 * _D, 2), torch.float64, dev)
        self._linear(x, lw["W0"], lw["b0"], h, B, N, A2=msg, flags=4, stats=stats, tag="mlp0")  # OPB_LIN_STATS
        mean = self._ws.get("mean", (B, 2 * _D), torch.float32, dev)
        rstd = self._ws.get("rstd", (B, 2 * _D), torch.float32, dev)
        call("op_instnorm_finalize", ptr(stats), B, tiles, 2 * _D, N, 1e-5, ptr(mean), ptr(rstd), stream_ptr())
        self._linear(h, lw["W1"], lw["b1"], x, B, N, flags=2 | 8, mean=mean, rstd=rstd, tag="mlp1")   # NORM_RELU in, RESIDUAL out

    # ------------------------------------------------------------------ tcgen05 launchers (linear_tc.cuh)
    def _linear_tc(self, A1, W, bias, B, rows, *, A2=None, W2=None, C=None, Cp=None, flags=0, elu_cols=0, stats=None, ksum=None, zscale_num=1.0, tag="",
                   work_k=None, counts=None):
        """A1/A2/W/W2/Cp are (hi, lo) plane pairs; C an fp32 tensor; at least one of C / Cp.  work_k: the K of the reference's
        own projection when the GEMM carries extra columns (the identity block of the in-GEMM residual add)."""
        K1 = A1[0].shape[-1]
        K2 = 0 if A2 is None else A2[0].shape[-1]
        Cout = W[0].shape[-2]
        timeline.work = 2.0 * B * rows * (work_k or (K1 + K2)) * Cout
        # HBM side of the same call: split-precision planes are 4 bytes per element (hi + lo fp16); a broadcast operand and shared
        # weights are read once (they stay in L2), per-frame weights once per frame
        a1b = 1 if flags & 2048 else B
        a2b = 1 if flags & 4096 else B
        timeline.bytes = 4.0 * rows * (a1b * K1 + a2b * K2) + 4.0 * B * rows * Cout * ((1 if Cp else 0) + (1 if C is not None else 0)) + \
            4.0 * Cout * (K1 + K2) * (B if flags & (64 | 512) else 1)
        head = (ptr(A1[0]), ptr(A1[1]), K1, ptr(A2[0]) if A2 else None, ptr(A2[1]) if A2 else None, K2,
                ptr(W[0]), ptr(W[1]), ptr(W2[0]) if W2 else None, ptr(W2[1]) if W2 else None, ptr(bias), ptr(C), ptr(Cp[0]) if Cp else None, ptr(Cp[1]) if Cp else None,
                B, rows, Cout, flags, elu_cols, ptr(stats), ptr(ksum), float(zscale_num))
        if counts is None:
            call("op_linear_tc", *head, stream_ptr(), tag=tag)
        else:
            call("op_linear_tc_ragged", *head, ptr(counts), stream_ptr(), tag=tag)

    def _kv_tc(self, src_p, B, M, lw, tag, counts=None):
        """Source side of linear attention (:90-97) re-associated for the tensor cores (atb_tc.cu): K' projection with
        column sums, G = K'^T X, then per frame Ksum and P'_b = W0b Wm blockdiag(KV')^T [512,256] (merge and the message half
        of MLP layer 0 in one matrix).  Returns (ksum, P'_b planes)."""
        dev = src_p[0].device
        kp = self._planes("kp" + tag, (B, M, _D), dev)
        parts = lib.op_linear_tc_stat_parts(M)
        ks_part = self._ws.get("kspart" + tag, (B, parts, _D, 2), torch.float32, dev)
        # counts (ragged batch): K' rows past a frame's count are written as zeros and left out of Ksum, so they add nothing to G either;
        # the 1/M of :93 and the M of :98 cancel, both stay the padded M
        self._linear_tc(src_p, lw["Wk_p"], lw["bk"], B, M, Cp=kp, flags=1 | 16, elu_cols=_D, stats=ks_part, tag="k", counts=counts)   # ELU1 | COLSUM
        nch = lib.op_atb_chunks(M)
        part = self._ws.get("gpart" + tag, (B, nch, _D, _D), torch.float32, dev)
        timeline.work = 2.0 * B * M * _D * _D
        timeline.bytes = 2 * 4.0 * B * M * _D + 4.0 * B * nch * _D * _D        # K' and X planes in, fp32 chunk partials out
        call("op_atb_tc", ptr(kp[0]), ptr(kp[1]), ptr(src_p[0]), ptr(src_p[1]), B, M, ptr(part), stream_ptr())
        ksum = self._ws.get("ksum" + tag, (B, _D), torch.float32, dev)
        P = self._planes("P" + tag, (B, 2 * _D, _D), dev)
        call("op_kv_post", ptr(part), nch, ptr(ks_part), parts, ptr(lw["WvT"]), ptr(lw["bv"]), ptr(lw["W0mT"]), 2 * _D, B, 1.0 / M, ptr(ksum), None,
             ptr(P[0]), ptr(P[1]), stream_ptr())
        return ksum, P

    def _propagate_tc(self, st, M, ksum, P, lw, shared_src=False, x_in=None):
        """x-side of AttentionPropagation on the tensor cores; the stream state st (planes st.p, column sums st.xs, and the
        fp32 copy st.x when one is kept) is updated in place.
        With Q'' = Q' M / (Q'_h . Ksum_h + 1e-6)  (:97-98):  message = Q'' P_b^T + bm  (:119), and the MLP's first
        layer sees it only through W0b, so  h = W0a x + P'_b Q'' + const  -- the merge projection is never run.
        InstanceNorm + ReLU (:144-146) without a normalisation pass: the mean is analytic (column sums of x and Q'' through
        W0a / P'_b), layer 0 writes relu(h - mean) as planes, 1/std is folded into per-frame layer-1 weights.
        st.ch names the stream ("2" / "3"): its scratch buffers are its own, the two may run concurrently.
        x_in: planes [1, N, 256] of a stream that is still the same for every frame (the shared 3D prefix at its first per-frame
        layer): the GEMMs read it as a broadcast operand (OPB_LIN_A1_BCAST / A2_BCAST) and write the per-frame result to st.p, so
        the B-fold copy of the prefix is never made."""
        B, N, ch, x_p = st.B, st.N, st.ch, (x_in if x_in is not None else st.p)
        cnt = st.counts                  # ragged batch: valid rows per frame of THIS stream (None: all N)
        b1, b2 = (2048, 4096) if x_in is not None else (0, 0)
        dev = x_p[0].device
        parts = lib.op_linear_tc_stat_parts(N)
        if shared_src:          # one source (ksum [1,256], P'_b [1,512,256]) for all B frames: P' is an ordinary weight
            ksum = self._ws.get("ksum_bc", (B, _D), torch.float32, dev).copy_(ksum.expand(B, _D))
            P, pflag = (P[0][0], P[1][0]), 0
        else:
            pflag = 512                                                                                           # W2BATCHED
        q_p = self._planes("q" + ch, (B, N, _D), dev)
        qs = self._ws.get("qpart" + ch, (B, parts, _D, 2), torch.float32, dev)
        self._linear_tc(x_p, lw["Wq_p"], lw["bq"], B, N, Cp=q_p, flags=1 | 16 | 32 | b1, elu_cols=_D, stats=qs, ksum=ksum, zscale_num=M,
                        tag="q", counts=cnt)                                                                      # ELU1 | COLSUM | ZSCALE
        cb = self._ws.get("cbias" + ch, (B, 2 * _D), torch.float32, dev)
        if cnt is None:
            call("op_mlp_center_bias", ptr(st.xs), st.parts, ptr(qs), parts, ptr(lw["W0aT"]), ptr(P[0]), ptr(P[1]), int(pflag != 0), B, N, ptr(cb),
                 stream_ptr())
        else:
            call("op_mlp_center_bias_ragged", ptr(st.xs), st.parts, ptr(qs), parts, ptr(lw["W0aT"]), ptr(P[0]), ptr(P[1]), int(pflag != 0), B, N, ptr(cnt),
                 ptr(cb), stream_ptr())
        hn_p = self._planes("hn" + ch, (B, N, 2 * _D), dev)
        stats = self._ws.get("tcstats" + ch, (B, parts, 2 * _D, 2), torch.float32, dev)
        self._linear_tc(x_p, lw["W0a_p"], cb, B, N, A2=q_p, W2=P, Cp=hn_p, flags=128 | 256 | pflag | b1, stats=stats, tag="mlp0", counts=cnt)   # BIASB | CSTATS_RELU
        W1s = self._planes("W1s" + ch, (B, _D, 2 * _D), dev)
        if cnt is None:
            call("op_instnorm_scale_weight_planes", ptr(stats), B, parts, N, 1e-5, ptr(lw["W1"]), _D, 2 * _D, ptr(W1s[0]), ptr(W1s[1]), None,
                 stream_ptr())
        else:
            call("op_instnorm_scale_weight_planes_ragged", ptr(stats), B, parts, N, ptr(cnt), 1e-5, ptr(lw["W1"]), _D, 2 * _D, ptr(W1s[0]), ptr(W1s[1]), None,
                 stream_ptr())
        if st.x is None:
            # residual add (:78,83) inside the GEMM: x_new = [hn | x] [W1 diag(rstd) | I]^T.  The old value streams through the
            # prefetched TMA pipeline as a second operand instead of being fetched by the epilogue (exposed latency); the
            # identity's planes are exact (64, 0), so the tensor core reproduces hi + lo of x to fp32 rounding.  In place: a
            # row tile is read completely (all MMAs retired) before its epilogue stores, and Cout = 256 is one column tile.
            self._linear_tc(hn_p, W1s, lw["b1"], B, N, A2=x_p, W2=self._packed["eye_p"], Cp=st.p, flags=16 | 64 | 1024 | 8192 | b2, stats=st.xs,
                            tag="mlp1", work_k=2 * _D, counts=cnt)                                                          # COLSUM | WBATCHED | W2_EXACT_HI
        else:   # an fp32 copy is kept (with_linear_transform): residual add onto it in the epilogue
            self._linear_tc(hn_p, W1s, lw["b1"], B, N, C=st.x, Cp=x_p, flags=2 | 16 | 64, stats=st.xs, tag="mlp1", counts=cnt)  # RESIDUAL | COLSUM | WBATCHED
        st.parts = parts

    def _new_stream(self, ch, B, N, dev, keep_fp32, name=None):
        """Buffers of one descriptor stream of the tensor-core path: planes, column-sum partials (sized for either
        producer: the projection epilogue or the GATs kernel) and, only when something still reads it, an fp32 copy."""
        st = _StreamState()
        st.ch, st.B, st.N = ch, B, N
        name = name or ch
        st.p = self._planes("x" + name, (B, N, _D), dev)
        maxparts = max(lib.op_linear_tc_stat_parts(N), lib.op_gats_planes_parts(N))
        st.xs = self._ws.get("xs" + name, (B, maxparts, _D, 2), torch.float32, dev)
        st.parts = 0
        st.x = self._ws.get("xf" + name, (B, N, _D), torch.float32, dev) if keep_fp32 else None
        st.counts = None
        return st

    def _load_stream(self, st, x):
        """st <- fp32 x [B, N, 256]: planes + column sums (+ the fp32 copy)."""
        call("op_split_planes", ptr(x), ptr(st.p[0]), ptr(st.p[1]), x.numel(), stream_ptr())
        call("op_colsum_partial", ptr(x), st.B, st.N, _D, ptr(st.xs), stream_ptr())
        st.parts = lib.op_linear_tc_stat_parts(st.N)
        if st.x is not None and st.x.data_ptr() != x.data_ptr():
            st.x.copy_(x)

    @staticmethod
    def _stream_fp32(st):
        return st.x if st.x is not None else (st.p[0].float() + st.p[1].float()) / 64.0

    def _gats_tc(self, st, db, L, lw, li):
        """GATs layer on a tensor-core stream state: in place on the planes when no linear transform is configured,
        else through the fp32 kernel and a re-split."""
        hp = self.hparams
        inc, add = hp["include_self"], hp["additional"]
        if st.x is None:
            leaves = db["leaves"]
            timeline.work = 4.0 * _D * (leaves.shape[0] * st.N * L + 2 * st.B * st.N)
            call("op_gats_leaf_aggregate_planes", ptr(leaves), 0 if db["shared"] else st.N * L * _D, ptr(st.p[0]), ptr(st.p[1]),
                 ptr(lw["v1"]), ptr(lw["v2"]), ptr(st.xs), st.B, st.N, L, 0.2, int(inc), int(add and inc), stream_ptr())
            st.parts = lib.op_gats_planes_parts(st.N)
        else:
            self._gats(st.x, db, st.B, st.N, L, lw, li)
            self._load_stream(st, st.x)

    def _colsum(self, x, B, N, name):
        """Per-frame column sums of a plain fp32 tensor, in the partial layout of OPB_LIN_COLSUM."""
        part = self._ws.get(name, (B, lib.op_linear_tc_stat_parts(N), _D, 2), torch.float32, x.device)
        call("op_colsum_partial", ptr(x), B, N, _D, ptr(part), stream_ptr())
        return part

    def _gats(self, x3, db, B, N3, L, lw, li):
        hp = self.hparams
        lt, inc, add = hp["with_linear_transform"], hp["include_self"], hp["additional"]
        leaves = db["leaves"]
        Bl = leaves.shape[0]
        leaf_bstride = 0 if db["shared"] else N3 * L * _D
        h3_add = x3 if (add and inc) else None
        if lt:
            wh_leaf = db["wh_leaf"].get(li)
            if wh_leaf is None:                                      # static DB x static weights: once per object
                wh_leaf = torch.empty_like(leaves)
                self._linear(leaves, lw["Wt"], None, wh_leaf, Bl, N3 * L, tag="gats_w")
                db["wh_leaf"][li] = wh_leaf
            wh3 = self._ws.get("wh3", (B, N3, _D), torch.float32, x3.device)
            self._linear(x3, lw["Wt"], None, wh3, B, N3, tag="gats_w")
            leaf_src, h3L, v1, v2 = wh_leaf, wh3, lw["a1"], lw["a2"]
        else:
            leaf_src, h3L, v1, v2 = leaves, x3, lw["v1"], lw["v2"]
        timeline.work = 4.0 * _D * (Bl * N3 * L + 2 * B * N3)               # algorithmic bytes: leaves once, h3 in/out
        call("op_gats_leaf_aggregate", ptr(leaf_src), leaf_bstride, ptr(h3L), ptr(h3_add), ptr(v1), ptr(v2), ptr(x3),
                                         B, N3, L, 0.2, int(inc), 1, stream_ptr())

    # ------------------------------------------------------------------ the 12 GNN layers + final_proj (:232-241)
    def _encode_simt(self, x2, db, pk, B, N2, N3, L, m2, m3, dev):
        """Exact-fp32 CUDA-core path (cross-check of the tensor-core path in the tests)."""
        x3 = self._ws.get("x3", (B, N3, _D), torch.float32, dev)
        x3.copy_(db["d3"] if not db["shared"] else db["d3"].expand(B, N3, _D))
        for li, lw in enumerate(pk["layers"]):
            kind = lw["kind"]
            if kind == "GATs":
                self._gats(x3, db, B, N3, L, lw, li)
            else:
                # both deltas of a layer are computed from the layer's input (:77-78,82-83): reduce the source
                # side of both calls first, then update the two streams in place.
                if kind == "cross":
                    KVa, Ksa = self._kv(x3, B, N3, lw, "a")       # desc2d attends to desc3d
                    KVb, Ksb = self._kv(x2, B, N2, lw, "b")       # desc3d attends to desc2d
                    Ma, Mb = N3, N2
                else:
                    KVa, Ksa = self._kv(x2, B, N2, lw, "a")
                    KVb, Ksb = self._kv(x3, B, N3, lw, "b")
                    Ma, Mb = N2, N3
                self._propagate(x2, B, N2, Ma, KVa, Ksa, lw)
                self._propagate(x3, B, N3, Mb, KVb, Ksb, lw)
            if self._trace is not None:
                self._trace.append((li, x2.clone(), x3.clone()))
        self._linear(x2, pk["Wf"], pk["bf"], m2, B, N2, tag="final")
        self._linear(x3, pk["Wf"], pk["bf"], m3, B, N3, tag="final")

    def _encode_tc(self, x2, db, pk, B, N2, N3, L, m2, m3, dev, x2_cm=None, count2d=None):
        """Tensor-core path.  The streams live as split-precision planes + per-frame column sums; an fp32 copy is kept
        only when the GATs layers need one (with_linear_transform)."""
        keep = bool(self.hparams["with_linear_transform"])
        s2 = self._new_stream("2", B, N2, dev, keep)
        s2.counts = count2d
        if x2 is None:
            x = x2_cm
            if x.dtype != torch.float32 or x.stride(2) != 1 or x.stride(1) != N2:
                x = (x[:1] if x.stride(0) == 0 else x).float().contiguous()
            if count2d is None:
                call("op_channel_major_to_planes", ptr(x), x.stride(0) if x.shape[0] > 1 else 0, ptr(s2.p[0]), ptr(s2.p[1]), ptr(s2.xs), B, N2, stream_ptr())
            else:
                call("op_channel_major_to_planes_ragged", ptr(x), x.stride(0) if x.shape[0] > 1 else 0, ptr(s2.p[0]), ptr(s2.p[1]), ptr(s2.xs), B, N2,
                     ptr(count2d), stream_ptr())
            s2.parts = lib.op_linear_tc_stat_parts(N2)
        else:
            self._load_stream(s2, x2)
        s3 = self._new_stream("3", B, N3, dev, keep)
        # With one object DB shared by the whole batch the 3D stream does not depend on the query frame until it first
        # receives a message from the 2D stream (the x-side update of the first 'cross' layer): up to there it is
        # computed once (batch of 1) and then broadcast -- the same bits per frame, 1/B of the work.
        single3 = bool(db["shared"])
        if single3:
            s3s = self._new_stream("3", 1, N3, dev, keep, name="3s")
            self._load_stream(s3s, db["d3"])
        else:
            self._load_stream(s3, db["d3"])
        main = torch.cuda.current_stream(dev)
        side = self._side_stream(dev) if (self.two_streams and self._trace is None) else main
        side.wait_stream(main)
        ev2 = ev3 = None                 # end of the last cross-layer update of the 2D / 3D stream

        def broadcast3(planes=True):
            if planes:
                s3.p[0].copy_(s3s.p[0].expand(B, N3, _D)); s3.p[1].copy_(s3s.p[1].expand(B, N3, _D))
            s3.xs.copy_(s3s.xs.expand_as(s3.xs))
            s3.parts = s3s.parts
            if s3.x is not None:
                s3.x.copy_(s3s.x.expand(B, N3, _D))

        for li, lw in enumerate(pk["layers"]):
            kind = lw["kind"]
            cur3 = s3s if single3 else s3
            if kind == "GATs":
                self._gats_tc(cur3, db, L, lw, li)
            else:
                # The 2D stream runs on a side CUDA stream, the 3D stream on the caller's: within a layer the two
                # updates only meet at the source-side reductions of a 'cross' layer (events below), so the short
                # launches and the tail of every persistent GEMM of one stream overlap the other's work.
                # Both deltas of a layer are computed from the layer's input (:77-78,82-83).
                if ev3 is not None:
                    side.wait_event(ev3)                          # the 3D stream is done with last layer's "2" reductions
                if ev2 is not None:
                    main.wait_event(ev2)
                ev2 = ev3 = None
                if kind == "cross":
                    k3 = self._kv_tc(cur3.p, cur3.B, N3, lw, "3")     # desc2d attends to desc3d
                    e3 = main.record_event()
                    with torch.cuda.stream(side):
                        k2 = self._kv_tc(s2.p, B, N2, lw, "2", counts=count2d)        # desc3d attends to desc2d
                        e2 = side.record_event()
                        side.wait_event(e3)
                        self._propagate_tc(s2, N3, *k3, lw, shared_src=single3)
                        ev2 = side.record_event()
                    main.wait_event(e2)
                    if single3:                                   # from here on the 3D stream is per frame
                        lazy = s3.x is None                       # planes-only stream: the prefix is read as a broadcast operand
                        broadcast3(planes=not lazy)
                        single3 = False
                        self._propagate_tc(s3, N2, *k2, lw, x_in=s3s.p if lazy else None)
                    else:
                        self._propagate_tc(s3, N2, *k2, lw)
                    ev3 = main.record_event()
                else:
                    with torch.cuda.stream(side):
                        k2 = self._kv_tc(s2.p, B, N2, lw, "2", counts=count2d)
                        self._propagate_tc(s2, N2, *k2, lw)
                    k3 = self._kv_tc(cur3.p, cur3.B, N3, lw, "3")
                    self._propagate_tc(cur3, N3, *k3, lw)
            if self._trace is not None:
                cur3 = s3s if single3 else s3
                self._trace.append((li, self._stream_fp32(s2).clone(), self._stream_fp32(cur3).expand(B, N3, _D).clone()))
        main.wait_stream(side)
        if single3:                                               # no cross layer at all (not the shipped layer list)
            broadcast3()
        if m2 is None:
            # final_proj (:241) with F.normalize (:245-246) in its epilogue, written as the similarity kernel's operand planes
            u2 = (self._ws.get("u2h", (B, N2, _D), torch.float16, dev), self._ws.get("u2b", (2, B, N2, _D), torch.uint8, dev))
            u3 = (self._ws.get("u3h", (B, N3, _D), torch.float16, dev), self._ws.get("u3b", (2, B, N3, _D), torch.uint8, dev))
            self._linear_tc(s2.p, pk["Wf_p"], pk["bf"], B, N2, Cp=u2, flags=16384, tag="final")          # OPB_LIN_UNIT_PLANES8
            self._linear_tc(s3.p, pk["Wf_p"], pk["bf"], B, N3, Cp=u3, flags=16384, tag="final")
            return u2, u3
        self._linear_tc(s2.p, pk["Wf_p"], pk["bf"], B, N2, C=m2, tag="final")          # final_proj, :241
        self._linear_tc(s3.p, pk["Wf_p"], pk["bf"], B, N3, C=m3, tag="final")
        return None

    # ------------------------------------------------------------------ forward
    @torch.no_grad()
    def forward(self, data):
        """Same contract as the reference forward (GATs_SuperGlue.py:203-278).

        Extra (non-reference) keys of ``pred``: ``matches0_batched`` [B,N2], ``matches1_batched`` [B,N3],
        ``matching_scores{0,1}_batched``; and, when ``data`` carries ``'query_intrinsic'`` (as in
        eval_utils.py:68) or ``'K'``, the products of the caller's gather + PnP (inference.py:147-155):
        ``mkpts2d``, ``mkpts3d``, ``mconf``, ``pose`` [B,3,4], ``inliers`` -- see ``onepose_b200.pnp``.

        Ragged batches (an extra input key, the reference contract untouched): ``data['count2d']`` [B] int gives the number of valid
        query keypoints of every frame of a batch padded to a common N2 (``SuperPoint.forward_batched`` pads that way).  Frame b's
        results then equal the reference run on its first ``count2d[b]`` keypoints alone (inference.py:140-146 runs B = 1 with
        whatever count SuperPoint found); padding rows come back as conf rows of zeros and ``matches0 = -1``.
        """
        kpts2d, kpts3d = data["keypoints2d"].float(), data["keypoints3d"].float()
        desc2d = data["descriptors2d_query"].float()
        desc3d, leaves = data["descriptors3d_db"].float(), data["descriptors2d_db"].float()

        if kpts2d.shape[1] == 0 or kpts3d.shape[1] == 0:            # :223-231, a single dict
            shape0, shape1 = kpts2d.shape[:-1], kpts3d.shape[:-1]
            return {
                "matches0": kpts2d.new_full(shape0, -1, dtype=torch.int)[0],
                "matches1": kpts3d.new_full(shape1, -1, dtype=torch.int)[0],
                "matching_scores0": kpts2d.new_zeros(shape0)[0],
                "matching_scores1": kpts3d.new_zeros(shape1)[0],
                "skip_train": True,
            }
        if self.match_type != "softmax":
            raise NotImplementedError                                 # :275-276
        if not desc2d.is_cuda:
            raise RuntimeError("onepose_b200.GATsSuperGlue runs on CUDA only (no CPU fallback): move the inputs "
                               "to a B200 device")
        dev = desc2d.device
        count2d = data.get("count2d")
        if count2d is not None:
            count2d = torch.as_tensor(count2d).to(device=dev, dtype=torch.int32).reshape(-1).contiguous()
            if count2d.numel() != desc2d.shape[0]:
                raise ValueError(f"count2d has {count2d.numel()} entries for a batch of {desc2d.shape[0]} frames")
            if not self.use_tensor_cores or self.hparams["with_linear_transform"] or not self.fused_final_normalize:
                raise NotImplementedError("count2d (ragged batches) exists on the production path only: tensor cores, planes-only streams")
        with torch.cuda.device(dev):
            return self._forward_cuda(data, kpts2d, kpts3d, desc2d, desc3d, leaves, dev, count2d)

    def _forward_cuda(self, data, kpts2d, kpts3d, desc2d, desc3d, leaves, dev, count2d=None):
        K = data.get("query_intrinsic", data.get("K"))
        scale = float(data.get("pnp_scale", 1000.0))
        if self._graphs_on and self.use_tensor_cores and self._trace is None and not timeline.enabled:
            return self._forward_graphed(kpts2d, kpts3d, desc2d, desc3d, leaves, K, scale, dev, count2d)
        pred, conf, i0, s0 = self._match(kpts2d, kpts3d, desc2d, desc3d, leaves, dev, count2d)
        if K is not None:
            from . import pnp
            if self.pnp_overlap:
                main, ps = torch.cuda.current_stream(dev), self.pose_stream(dev)
                ps.wait_stream(main)                                  # the matches are complete
                with torch.cuda.stream(ps):
                    pred.update(pnp.pose_from_matches(i0, s0, kpts2d, kpts3d, K, scale=scale, refine=self.pnp_refine))
                    for t in (i0, s0, kpts2d, kpts3d):                # read on the pose stream after the caller may have dropped them
                        t.record_stream(ps)
                    pred["pose_event"] = ps.record_event()
            else:
                pred.update(pnp.pose_from_matches(i0, s0, kpts2d, kpts3d, K, scale=scale, refine=self.pnp_refine))
        return pred, conf

    def _match(self, kpts2d, kpts3d, desc2d, desc3d, leaves, dev, count2d=None):
        """The matcher proper (:232-274) on the current stream: returns (pred, conf, matches0 [B,N2], scores0 [B,N2])."""
        B, D, N2 = desc2d.shape
        N3 = desc3d.shape[2]
        L = int(leaves.shape[2] / N3)                                 # GATs.py:65
        if D != _D or desc3d.shape[1] != _D or L < 1:
            raise ValueError(f"bad descriptor shapes {tuple(desc2d.shape)} {tuple(desc3d.shape)} {tuple(leaves.shape)}")
        pk = self._pack_weights(dev)
        db = self._static_db(desc3d, leaves[:, :, :N3 * L], L)
        # the final projections leave as unit-row operand planes of the similarity kernel (production), or as fp32 rows (the exact-fp32
        # cross-check path and the A/B switches of the similarity kernel: OPB_SIM_CG=1, OPB_SIM_MMA=f16, OPB_LINEAR_CG=1, OPB_LIN_DBG)
        unit_planes = self.use_tensor_cores and self.fused_final_normalize and not any(
            os.environ.get(k) for k in ("OPB_SIM_CG", "OPB_SIM_MMA", "OPB_LINEAR_CG", "OPB_LIN_DBG"))
        if count2d is not None and not unit_planes:
            raise NotImplementedError("count2d (ragged batches) needs the production kernels: unset the OPB_SIM_* / OPB_LIN* A/B switches")
        m2 = m3 = None
        if not unit_planes:
            m2 = self._ws.get("m2", (B, N2, _D), torch.float32, dev)
            m3 = self._ws.get("m3", (B, N3, _D), torch.float32, dev)
        planes = None
        if self.use_tensor_cores and not self.hparams["with_linear_transform"]:
            # planes-only streams: the query descriptors go from the public channel-major layout straight into planes + column sums
            planes = self._encode_tc(None, db, pk, B, N2, N3, L, m2, m3, dev, x2_cm=desc2d, count2d=count2d)
        else:
            x2 = self._to_point_major(desc2d, self._ws.get("x2", (B, N2, _D), torch.float32, dev))
            if self.use_tensor_cores:
                planes = self._encode_tc(x2, db, pk, B, N2, N3, L, m2, m3, dev)
            else:
                self._encode_simt(x2, db, pk, B, N2, N3, L, m2, m3, dev)
        if not self.use_tensor_cores:                                 # the tensor-core op normalises while it splits the operands
            call("op_l2_normalize_rows", ptr(m2), ptr(m2), B * N2, _D, stream_ptr())
            call("op_l2_normalize_rows", ptr(m3), ptr(m3), B * N3, _D, stream_ptr())

        conf = torch.empty(B, N2, N3, dtype=torch.float32, device=dev)      # returned to the caller
        i0 = torch.empty(B, N2, dtype=torch.int64, device=dev)
        i1 = torch.empty(B, N3, dtype=torch.int64, device=dev)
        s0 = torch.empty(B, N2, dtype=torch.float32, device=dev)
        s1 = torch.empty(B, N3, dtype=torch.float32, device=dev)
        nbytes = lib.op_sim_dual_softmax_workspace_bytes(B, N2, N3)
        w = self._ws.get("simws", (nbytes,), torch.uint8, dev)
        tail = (float(self.hparams["scale_factor"]), float(self.hparams["match_threshold"]), ptr(conf), ptr(i0), ptr(i1), ptr(s0),
                ptr(s1), ptr(w), nbytes, stream_ptr())
        timeline.work = B * (4.0 * _D * (N2 + N3) + 4.0 * N2 * N3 + 24.0 * (N2 + N3))   # bytes, SURVEY.md 8(d)
        if planes is not None:
            (u2h, u2b), (u3h, u3b) = planes
            if count2d is None:
                call("op_sim_dual_softmax_planes", ptr(u2h), ptr(u2b), ptr(u3h), ptr(u3b), B, N2, N3, *tail)
            else:
                call("op_sim_dual_softmax_planes_ragged", ptr(u2h), ptr(u2b), ptr(u3h), ptr(u3b), B, N2, N3, ptr(count2d), *tail)
        else:
            call("op_sim_dual_softmax_rows" if self.use_tensor_cores else "op_sim_dual_softmax", ptr(m2), ptr(m3), B, N2, N3, *tail)
        pred = {
            "matches0": i0[0], "matches1": i1[0],                     # batch element 0 only, :269-274
            "matching_scores0": s0[0], "matching_scores1": s1[0],
            "matches0_batched": i0, "matches1_batched": i1,
            "matching_scores0_batched": s0, "matching_scores1_batched": s1,
        }
        return pred, conf, i0, s0

    # ------------------------------------------------------------------ CUDA graphs
    def enable_graphs(self, on=True):
        """Serving mode: ``forward`` replays one captured CUDA graph per (device, B, N2, N3, L, object DB, weights, intrinsics layout)
        instead of enqueuing ~150 launches from Python ("CUDA graphs per (shape, B) to keep 8 GPUs fed").  The inputs of a
        call are copied into the graph's static buffers; the tensors of ``pred`` / ``conf`` are the graph's static outputs: two slots
        are used in turn, so the results of a call stay valid until the second next call with the same key.  The object DB tensors are
        captured by address (they are static: keep them alive and unmodified, as ``ObjectDB`` does).  With ``pnp_overlap`` the gather runs
        at the end of the matcher graph and the RANSAC is a second graph on the pose stream, overlapping the next call's matcher."""
        self._graphs_on = bool(on)
        if not on:
            self._graphs.clear()
        return self

    def _forward_graphed(self, kpts2d, kpts3d, desc2d, desc3d, leaves, K, scale, dev, count2d=None):
        from . import pnp
        B, _, N2 = desc2d.shape
        Kt = None
        if K is not None:
            Kt = (K if isinstance(K, torch.Tensor) else torch.as_tensor(K)).to(device=dev, dtype=torch.float64).reshape(-1, 3, 3)
        key = (dev.index, B, N2, tuple(desc3d.shape), tuple(leaves.shape), tuple(kpts3d.shape), desc3d.data_ptr(), leaves.data_ptr(), kpts3d.data_ptr(),
               desc3d._version, leaves._version, kpts3d._version, desc3d.stride(0), leaves.stride(0), kpts3d.stride(0),
               None if Kt is None else Kt.shape[0], scale, self.pnp_overlap, self.pnp_refine, self.two_streams, self.fused_final_normalize,
               float(self.hparams["match_threshold"]), float(self.hparams["scale_factor"]), count2d is not None) + self._weights_key()
        ent = self._graphs.get(key)
        if ent is None:
            if len(self._graphs) >= 2:                                # static outputs (conf!) are large: keep at most two shapes alive
                self._graphs.pop(next(iter(self._graphs)))
            ent = self._graphs[key] = self._capture(kpts2d, kpts3d, desc2d, desc3d, leaves, Kt, scale, dev, count2d)
        slot = ent["slots"][ent["turn"]]
        ent["turn"] ^= 1
        main = torch.cuda.current_stream(dev)
        if slot["pose_event"] is not None:
            main.wait_event(slot["pose_event"])                       # the RANSAC that last read this slot's gather buffers has finished
        slot["desc2d"].copy_(desc2d, non_blocking=True)
        slot["kpts2d"].copy_(kpts2d, non_blocking=True)
        if count2d is not None:
            slot["count2d"].copy_(count2d, non_blocking=True)
        if Kt is not None:
            slot["K"].copy_(Kt, non_blocking=True)
        slot["match_graph"].replay()
        pred = dict(slot["pred"])
        if Kt is not None:
            if self.pnp_overlap:
                ps = self.pose_stream(dev)
                ps.wait_stream(main)
                with torch.cuda.stream(ps):
                    slot["pnp_graph"].replay()
                    slot["pose_event"] = pred["pose_event"] = ps.record_event()
            else:
                slot["pnp_graph"].replay()
        return pred, slot["conf"]

    def _capture(self, kpts2d, kpts3d, desc2d, desc3d, leaves, Kt, scale, dev, count2d=None):
        from . import pnp
        # one eager call first: weights packed, DB cached, every workspace buffer allocated, every kernel's attributes configured
        _, _, i0, s0 = self._match(kpts2d, kpts3d, desc2d, desc3d, leaves, dev, count2d)
        if Kt is not None:
            pnp.pose_from_matches(i0, s0, kpts2d, kpts3d, Kt, scale=scale, refine=self.pnp_refine)
        torch.cuda.synchronize(dev)
        slots = []
        for _ in range(2):
            slot = dict(desc2d=desc2d.detach().clone().contiguous(), kpts2d=kpts2d.detach().clone().contiguous(),
                        K=None if Kt is None else Kt.detach().clone().contiguous(), pose_event=None, pnp_graph=None,
                        count2d=None if count2d is None else count2d.detach().clone())
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                pred, conf, i0, s0 = self._match(slot["kpts2d"], kpts3d, slot["desc2d"], desc3d, leaves, dev, slot["count2d"])
                gathered = pnp.gather_matches_device(i0, s0, slot["kpts2d"], kpts3d, scale) if Kt is not None else None
            slot.update(match_graph=g, conf=conf)
            if Kt is not None:
                g2 = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g2, pool=g.pool()):
                    pred.update(pnp.pose_from_gathered(gathered, slot["K"], scale, refine=self.pnp_refine))
                slot["pnp_graph"] = g2
            slot["pred"] = pred
            slots.append(slot)
        torch.cuda.synchronize(dev)
        return dict(slots=slots, turn=0, keep=(desc3d, leaves, kpts3d))


_POSE_KEYS = ("mkpts2d", "mkpts3d", "mconf", "n_matches", "match_query_idx", "pose", "inliers", "n_inliers", "ransac_iters")


class _StreamState:
    """One descriptor stream (2D query / 3D object) of the tensor-core path: see GATsSuperGlue._new_stream."""
    __slots__ = ("ch", "B", "N", "p", "xs", "parts", "x", "counts")


class LitModelGATsSPG(nn.Module):
    """Lightning-free look-alike of ``LitModelGATsSPG`` (GATsSPG_lightning_model.py:15-37).

    Keeps what the inference callers use (inference.py:55-58,146): ``.matcher``, ``forward(x) =
    self.matcher(x)``, ``.cuda() / .eval() / .freeze()``, ``load_from_checkpoint`` (reads the
    ``state_dict`` and ``hyper_parameters`` of a Lightning checkpoint).  Training, the SuperPoint
    extractor and logging are out of scope.
    """

    def __init__(self, *args, **kwargs):
        super().__init__()
        hp = dict(args[0]) if args and isinstance(args[0], dict) else {}
        hp.update(kwargs)
        from .synth import DEFAULT_HPARAMS
        self.hparams = {**DEFAULT_HPARAMS, **hp}
        self.matcher = GATsSuperGlue(hparams=self.hparams)

    def forward(self, x):
        return self.matcher(x)

    def freeze(self):
        for p in self.parameters():
            p.requires_grad = False
        return self.eval()

    @classmethod
    def load_from_checkpoint(cls, checkpoint_path, map_location="cpu", **kwargs):
        ckpt = torch.load(checkpoint_path, map_location=map_location, weights_only=False)
        hp = dict(ckpt.get("hyper_parameters", {}))
        hp.update(kwargs)
        model = cls(**hp)
        sd = ckpt.get("state_dict", ckpt)
        own = {k: v for k, v in sd.items() if k.startswith("matcher.")}     # extractor.* is out of scope
        model.load_state_dict(own, strict=True)
        return model
