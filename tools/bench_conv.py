"""Micro-benchmark of op_conv_tc on one layer shape (row f3):  python tools/bench_conv.py [--layer conv1b] [--frames 64]
Environment switches: OPB_CONV_HALO=0 (tap-by-tap kernel), OPB_CONV_DBG (timing experiments, see conv_tc.cu)."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from onepose_b200._lib import call, ptr, stream_ptr  # noqa: E402

LAYERS = {"conv1b": (512, 64, 64, 3, 1), "conv2a": (256, 64, 64, 3, 0), "conv2b": (256, 64, 64, 3, 1), "conv3a": (128, 64, 128, 3, 0),
          "conv3b": (128, 128, 128, 3, 1), "conv4a": (64, 128, 128, 3, 0), "convPa": (64, 128, 256, 3, 0), "convDb": (64, 256, 256, 1, 0)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layer", default="conv1b")
    ap.add_argument("--frames", type=int, default=64)
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    out = {}
    for name in a.layer.split(","):
        S, ci, co, k, pool = LAYERS[name]
        pool = pool if os.environ.get("OPB_CONV_HALO", "1") != "0" else 0
        B = a.frames
        cp = 64 if co == 64 else -(-co // 128) * 128
        g = torch.Generator(device="cuda").manual_seed(0)
        xh = (torch.rand(B, S, S, ci, device="cuda", generator=g) * 64).half(); xl = torch.zeros_like(xh)
        wh = (torch.randn(k * k, cp, ci, device="cuda", generator=g)).half(); wl = torch.zeros_like(wh)
        bias = torch.zeros(co, device="cuda")
        So = S // 2 if pool else S
        yh = torch.empty(B, So, So, co, dtype=torch.float16, device="cuda"); yl = torch.empty_like(yh)
        y32 = torch.empty(B, S, S, co, device="cuda") if k == 1 else None
        fn = lambda: call("op_conv_tc", ptr(xh), ptr(xl), B, S, S, ci, ptr(wh), ptr(wl), cp, ptr(bias), co, k * k, 1, pool if k == 3 else 0,
                          None if k == 1 else ptr(yh), None if k == 1 else ptr(yl), ptr(y32), co if k == 1 else 0, stream_ptr())
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(a.iters):
            fn()
        e.record(); torch.cuda.synchronize()
        ms = s.elapsed_time(e) / a.iters
        flop = 2.0 * B * S * S * co * ci * k * k * 3
        out[name] = {"ms": round(ms, 4), "TFLOPs_issued": round(flop / ms / 1e9, 1)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
