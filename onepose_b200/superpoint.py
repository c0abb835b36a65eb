"""SuperPoint front-end  (SURVEY.md 8(f) row f3).

Mirrors ``SuperPoint`` of src/models/extractors/SuperPoint/superpoint.py:120-244: same constructor config, same parameter
names (a reference ``state_dict`` -- superpoint_v1.pth -- loads with ``strict=True``), same output dict
``{'keypoints': [..], 'scores': [..], 'descriptors': [..]}`` of per-image tensors.

What runs where:

* the twelve convolutions run on the tensor cores at fp32-class accuracy (``csrc/conv_tc.cu``: tcgen05 implicit GEMMs over
  NHWC split-precision fp16 planes, three MMAs per product, fp32 TMEM accumulation; the 1-channel first layer and the 2x2
  max-pools are small CUDA-core kernels).  ``conv_impl='cudnn'`` (or env ``OPB_SP_CONV=cudnn``) runs them through the library
  instead (fp32, TF32 off) -- kept as the measurement baseline and the parity cross-check, not as a fallback;
* everything after them is this repo's CUDA (``csrc/superpoint.cu``): channel softmax + 8x8 depth-to-space, the five
  max-pool stages of ``simple_nms``, threshold + border removal + ordered compaction + top-k in one kernel per image,
  descriptor normalisation into a pixel-major map, bilinear sampling + L2 normalisation.  A whole batch of frames is
  processed at once and nothing crosses to the host (the reference loops over images in Python with ``torch.nonzero`` host
  syncs, superpoint.py:214-241).

``forward_batched`` returns padded device tensors + counts (what ``GATsSuperGlue`` wants for a block of frames);
``forward`` cuts them into the reference's lists.  CUDA only; no CPU fallback.
"""
import os

import torch
from torch import nn

from ._lib import call, lib, ptr, stream_ptr

_MAX_CAND = 16384


class SuperPoint(nn.Module):
    default_config = {
        'descriptor_dim': 256,
        'nms_radius': 4,
        'keypoint_threshold': 0.005,
        'max_keypoints': -1,
        'remove_borders': 4,
    }

    def __init__(self, config, align_corners=True, conv_impl=None):
        """``config`` as the reference's (superpoint.py:133-139; unknown keys are carried but ignored, which is how the
        'keypoints_threshold' key of extract_features.py:9-17 leaves the 0.005 default in force).  ``align_corners``: the
        reference picks grid_sample's align_corners from the torch version string (:106); True is what its pinned
        torch 1.8 gives."""
        super().__init__()
        self.config = {**self.default_config, **config}
        self.align_corners = bool(align_corners)
        self.conv_impl = conv_impl or os.environ.get("OPB_SP_CONV", "tc")
        if self.conv_impl not in ("tc", "cudnn"):
            raise ValueError("conv_impl must be 'tc' or 'cudnn'")
        self._packed = None
        c1, c2, c3, c4, c5 = 64, 64, 128, 128, 256
        self.relu = nn.ReLU(inplace=True)
        self.pool = nn.MaxPool2d(kernel_size=2, stride=2)
        self.conv1a = nn.Conv2d(1, c1, kernel_size=3, stride=1, padding=1)
        self.conv1b = nn.Conv2d(c1, c1, kernel_size=3, stride=1, padding=1)
        self.conv2a = nn.Conv2d(c1, c2, kernel_size=3, stride=1, padding=1)
        self.conv2b = nn.Conv2d(c2, c2, kernel_size=3, stride=1, padding=1)
        self.conv3a = nn.Conv2d(c2, c3, kernel_size=3, stride=1, padding=1)
        self.conv3b = nn.Conv2d(c3, c3, kernel_size=3, stride=1, padding=1)
        self.conv4a = nn.Conv2d(c3, c4, kernel_size=3, stride=1, padding=1)
        self.conv4b = nn.Conv2d(c4, c4, kernel_size=3, stride=1, padding=1)
        self.convPa = nn.Conv2d(c4, c5, kernel_size=3, stride=1, padding=1)
        self.convPb = nn.Conv2d(c5, 65, kernel_size=1, stride=1, padding=0)
        self.convDa = nn.Conv2d(c4, c5, kernel_size=3, stride=1, padding=1)
        self.convDb = nn.Conv2d(c5, self.config['descriptor_dim'], kernel_size=1, stride=1, padding=0)
        if self.config['descriptor_dim'] != 256:
            raise ValueError("onepose_b200.SuperPoint: descriptor_dim must be 256 (the matcher's width)")
        mk = self.config['max_keypoints']
        if mk == 0 or mk < -1:
            raise ValueError('"max_keypoints" must be positive or "-1"')          # superpoint.py:168-170
        if mk > _MAX_CAND:
            raise ValueError('"max_keypoints" above %d is not supported' % _MAX_CAND)

    # ------------------------------------------------------------------------------------------ network (library convs)
    @torch.no_grad()
    def dense_maps(self, inp):
        """superpoint.py:177-206,231-234: image [B,1,H,W] -> (detector logits [B,65,H/8,W/8], descriptor map [B,256,H/8,W/8])."""
        tf32 = torch.backends.cudnn.allow_tf32
        torch.backends.cudnn.allow_tf32 = False
        try:
            x = self.relu(self.conv1a(inp))
            x = self.relu(self.conv1b(x))
            x = self.pool(x)
            x = self.relu(self.conv2a(x))
            x = self.relu(self.conv2b(x))
            x = self.pool(x)
            x = self.relu(self.conv3a(x))
            x = self.relu(self.conv3b(x))
            x = self.pool(x)
            x = self.relu(self.conv4a(x))
            x = self.relu(self.conv4b(x))
            logits = self.convPb(self.relu(self.convPa(x)))
            dense = self.convDb(self.relu(self.convDa(x)))
        finally:
            torch.backends.cudnn.allow_tf32 = tf32
        return logits, dense

    # ------------------------------------------------------------------------------------------ network on the tensor cores
    _TC_LAYERS = ("conv1b", "conv2a", "conv2b", "conv3a", "conv3b", "conv4a", "conv4b", "convPa", "convPb", "convDa", "convDb")

    def _pack(self, dev):
        """Weight planes [taps, Cout_pad, Cin] of 64*w (op_split_planes), cached until a parameter changes."""
        key = (dev, tuple((p.data_ptr(), p._version) for p in self.parameters()))
        if self._packed is not None and self._packed[0] == key:
            return self._packed[1]
        packed = {}
        for name in self._TC_LAYERS:
            conv = getattr(self, name)
            w = conv.weight.detach().float()
            co, ci, k, _ = w.shape
            cp = 64 if co == 64 else -(-co // 128) * 128
            wt = torch.zeros(k * k, cp, ci, dtype=torch.float32, device=dev)
            wt[:, :co] = w.permute(2, 3, 0, 1).reshape(k * k, co, ci)
            hi, lo = torch.empty_like(wt, dtype=torch.float16), torch.empty_like(wt, dtype=torch.float16)
            call("op_split_planes", ptr(wt), ptr(hi), ptr(lo), wt.numel(), stream_ptr())
            packed[name] = (hi, lo, conv.bias.detach().float().contiguous(), co, cp, ci, k * k)
        packed["conv1a"] = (self.conv1a.weight.detach().float().reshape(64, 9).contiguous(), self.conv1a.bias.detach().float().contiguous())
        self._packed = (key, packed)
        return packed

    @torch.no_grad()
    def dense_maps_tc(self, inp):
        """The network of ``dense_maps`` on this repo's kernels.  image [B,1,H,W] -> (detector logits [B,H/8,W/8,72] fp32,
        pixel-major with columns 65..71 zero, descriptor map [B,H/8,W/8,256] fp32 pixel-major)."""
        self._require_cuda(inp, "dense_maps_tc")
        B, one, H, W = inp.shape
        if one != 1 or H % 8 or W % 8:
            raise ValueError("onepose_b200.SuperPoint: images must be [B,1,H,W] with H and W multiples of 8")
        dev = inp.device
        img = inp.float().contiguous()
        f16 = dict(dtype=torch.float16, device=dev)

        def planes(h, w, c):
            return torch.empty(B, h, w, c, **f16), torch.empty(B, h, w, c, **f16)

        with torch.cuda.device(dev):
            pk = self._pack(dev)
            st = stream_ptr()

            def conv(name, x, h, w, relu=True, f32_stride=0, pool=False):
                hi, lo, bias, co, cp, ci, taps = pk[name]
                if f32_stride:
                    y32 = torch.empty(B, h, w, f32_stride, dtype=torch.float32, device=dev)
                    call("op_conv_tc", ptr(x[0]), ptr(x[1]), B, h, w, ci, ptr(hi), ptr(lo), cp, ptr(bias), co, taps, int(relu), 0, None, None,
                         ptr(y32), f32_stride, st, tag=name)
                    return y32
                y = planes(h // 2, w // 2, co) if pool else planes(h, w, co)
                call("op_conv_tc", ptr(x[0]), ptr(x[1]), B, h, w, ci, ptr(hi), ptr(lo), cp, ptr(bias), co, taps, int(relu), int(pool), ptr(y[0]),
                     ptr(y[1]), None, 0, st, tag=name)
                return y

            def pool(x, h, w, c):
                y = planes(h // 2, w // 2, c)
                call("op_pool2_planes", ptr(x[0]), ptr(x[1]), B, h, w, c, ptr(y[0]), ptr(y[1]), st, tag="pool")
                return y

            x = planes(H, W, 64)
            call("op_conv1a_planes", ptr(img), ptr(pk["conv1a"][0]), ptr(pk["conv1a"][1]), B, H, W, ptr(x[0]), ptr(x[1]), st, tag="conv1a")
            fused = os.environ.get("OPB_CONV_HALO", "1") != "0"        # MaxPool2d(2,2) inside the convolution's epilogue

            def conv_pool(name, x, h, w, c):
                return conv(name, x, h, w, pool=True) if fused else pool(conv(name, x, h, w), h, w, c)

            x = conv_pool("conv1b", x, H, W, 64); H, W = H // 2, W // 2
            x = conv("conv2a", x, H, W)
            x = conv_pool("conv2b", x, H, W, 64); H, W = H // 2, W // 2
            x = conv("conv3a", x, H, W)
            x = conv_pool("conv3b", x, H, W, 128); H, W = H // 2, W // 2
            x = conv("conv4a", x, H, W)
            x = conv("conv4b", x, H, W)
            logits = conv("convPb", conv("convPa", x, H, W), H, W, relu=False, f32_stride=72)
            dense = conv("convDb", conv("convDa", x, H, W), H, W, relu=False, f32_stride=256)
        return logits, dense

    # ------------------------------------------------------------------------------------------ this repo's kernels
    @staticmethod
    def _require_cuda(t, what):
        if not t.is_cuda:
            raise RuntimeError("onepose_b200.SuperPoint.%s: CUDA tensors only (no CPU fallback)" % what)

    def score_map(self, logits, pixel_major=False):
        """:207-211 -- [B,65,h,w] (or pixel-major [B,h,w,stride]) -> [B,8h,8w]."""
        self._require_cuda(logits, "score_map")
        logits = logits.float().contiguous()
        if pixel_major:
            B, h, w, stride = logits.shape
        else:
            (B, _, h, w), stride = logits.shape, 0
        out = torch.empty(B, h * 8, w * 8, dtype=torch.float32, device=logits.device)
        with torch.cuda.device(logits.device):
            call("op_sp_softmax_shuffle", ptr(logits), B, h, w, stride, ptr(out), stream_ptr())
        return out

    def nms(self, scores, radius=None):
        """simple_nms (:47-69) on [B,H,W]."""
        self._require_cuda(scores, "nms")
        scores = scores.float().contiguous()
        B, H, W = scores.shape
        radius = self.config['nms_radius'] if radius is None else radius
        nbytes = lib.op_sp_nms_workspace_bytes(B, H, W)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=scores.device)
        out = torch.empty_like(scores)
        with torch.cuda.device(scores.device):
            call("op_sp_nms", ptr(scores), B, H, W, radius, ptr(out), ptr(ws), nbytes, stream_ptr())
        return out

    def select(self, nms_scores):
        """:214-228 -> (kpts [B,S,2] (x,y), scores [B,S], count [B] int32), S = max_keypoints (16384 when unlimited)."""
        self._require_cuda(nms_scores, "select")
        B, H, W = nms_scores.shape
        mk = self.config['max_keypoints']
        S = mk if mk >= 0 else _MAX_CAND
        dev = nms_scores.device
        kpts = torch.zeros(B, S, 2, dtype=torch.float32, device=dev)
        sc = torch.zeros(B, S, dtype=torch.float32, device=dev)
        count = torch.empty(B, dtype=torch.int32, device=dev)
        self._overflow = torch.empty(B, dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            call("op_sp_select", ptr(nms_scores), B, H, W, float(self.config['keypoint_threshold']), int(self.config['remove_borders']),
                 mk, S, ptr(kpts), ptr(sc), ptr(count), ptr(self._overflow), stream_ptr())
        return kpts, sc, count

    def sample(self, dense, kpts, count, s=8, pixel_major=False):
        """:235-241 -> [B,S,256] point-major, L2-normalised (rows >= count are zero).  dense [B,256,h,w], or [B,h,w,256] with
        ``pixel_major`` (normalised in place)."""
        self._require_cuda(dense, "sample")
        dense = dense.float().contiguous()
        if pixel_major:
            B, h, w, C = dense.shape
        else:
            B, C, h, w = dense.shape
        S = kpts.shape[1]
        out = torch.empty(B, S, C, dtype=torch.float32, device=dense.device)
        with torch.cuda.device(dense.device):
            if pixel_major:
                pm = dense
                call("op_l2_normalize_rows", ptr(pm), ptr(pm), B * h * w, C, stream_ptr())
            else:
                pm = torch.empty(B, h * w, C, dtype=torch.float32, device=dense.device)
                call("op_sp_desc_prepare", ptr(dense), B, h * w, ptr(pm), stream_ptr())
            call("op_sp_sample", ptr(pm), ptr(kpts), ptr(count), B, S, h, w, s, int(self.align_corners), ptr(out), stream_ptr())
        return out

    @torch.no_grad()
    def forward_batched(self, inp):
        """image batch [B,1,H,W] (H, W multiples of 8) -> dict of padded device tensors: keypoints [B,S,2], scores [B,S],
        descriptors_pm [B,S,256] (point-major), count [B] int32.  No host synchronisation."""
        self._require_cuda(inp, "forward")
        tc = self.conv_impl == "tc"
        logits, dense = self.dense_maps_tc(inp) if tc else self.dense_maps(inp)
        kpts, sc, count = self.select(self.nms(self.score_map(logits, pixel_major=tc)))
        desc = self.sample(dense, kpts, count, pixel_major=tc)
        return {"keypoints": kpts, "scores": sc, "descriptors_pm": desc, "count": count}

    @torch.no_grad()
    def forward(self, inp):
        """The reference's output (superpoint.py:242-246): lists over the batch of keypoints [n,2], scores [n] and
        descriptors [256,n]; one device->host copy of the counts."""
        out = self.forward_batched(inp)
        counts = out["count"].tolist()
        if int(self._overflow.max()) != 0:
            raise RuntimeError("onepose_b200.SuperPoint: more than %d NMS survivors in one image" % _MAX_CAND)
        return {
            'keypoints': [out["keypoints"][b, :n] for b, n in enumerate(counts)],
            'scores': tuple(out["scores"][b, :n] for b, n in enumerate(counts)),
            'descriptors': [out["descriptors_pm"][b, :n].t() for b, n in enumerate(counts)],
        }
