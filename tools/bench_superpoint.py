"""Row f3 measurement: SuperPoint front-end on a block of 512x512 frames (extract_features.py:20-22 preprocessing size).

    python tools/bench_superpoint.py [--frames 64] [--iters 10] [--cpu-frames 1]

Times, with CUDA events on the launching stream: the convolution stack (library), each post-network stage (this repo's
kernels) with its algorithmic HBM bytes, and the whole ``forward_batched``.  The CPU leg times the numpy oracle of the
post-network stages on ``--cpu-frames`` frames (a port; the reference cannot travel to the GPU box)."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from onepose_b200.superpoint import SuperPoint  # noqa: E402


def timed(fn, iters):
    fn(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        out = fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / iters, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=64)
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--cpu-frames", type=int, default=1)
    a = ap.parse_args()
    B, H = a.frames, a.size
    conf = {'descriptor_dim': 256, 'nms_radius': 3, 'max_keypoints': 4096, 'keypoints_threshold': 0.6}
    torch.manual_seed(0)
    m = SuperPoint(conf).eval().cuda()
    g = torch.Generator(device="cuda").manual_seed(0)
    img = torch.rand(B, 1, H, H, device="cuda", generator=g)
    h = H // 8
    res = {"frames": B, "size": H}
    with torch.no_grad():
        t_net, (logits, dense) = timed(lambda: m.dense_maps(img), a.iters)
        t_tc, (lg_tc, dn_tc) = timed(lambda: m.dense_maps_tc(img), a.iters)
        res["conv_stack_tc_ms"] = round(t_tc, 3)
        t_sm, scores = timed(lambda: m.score_map(logits), a.iters)
        t_nms, nms = timed(lambda: m.nms(scores), a.iters)
        t_sel, (k, s, c) = timed(lambda: m.select(nms), a.iters)
        t_smp, d = timed(lambda: m.sample(dense, k, c), a.iters)
        t_all, _ = timed(lambda: m.forward_batched(img), a.iters)
        from onepose_b200._lib import timeline
        timeline.enabled = True; timeline.records = []
        m.dense_maps_tc(img); torch.cuda.synchronize()
        timeline.enabled = False
        layers = {}
        for name, tag, s0, e0, _ in timeline.records:
            layers[tag or name] = round(layers.get(tag or name, 0.0) + s0.elapsed_time(e0), 4)
        res["conv_tc_layers_ms"] = layers
    px = B * H * H
    stages = {
        "conv_stack(cuDNN)": (t_net, None),
        "softmax_shuffle": (t_sm, B * h * h * 65 * 4 + px * 4),
        "nms(5 launches)": (t_nms, px * (4 + 1) + px * (1 + 1) + px * (4 + 1 + 1 + 1) + px * (1 + 1) + px * (4 + 1 + 1 + 1 + 4)),
        "select": (t_sel, px * 4 + B * 4096 * 12),
        "desc_prepare+sample": (t_smp, 2 * B * h * h * 256 * 4 + int(c.sum()) * (4 * 256 * 4 + 256 * 4)),
    }
    res["stages"] = {n: {"ms": round(t, 4), **({"alg_GB_s": round(b / t / 1e6, 1)} if b else {})} for n, (t, b) in stages.items()}
    res["forward_batched_ms"] = round(t_all, 3)
    res["frames_per_s"] = round(B / t_all * 1e3, 1)
    res["post_network_ms"] = round(t_sm + t_nms + t_sel + t_smp, 3)
    res["keypoints_per_frame"] = float(c.float().mean())
    if a.cpu_frames > 0:
        from oracle import superpoint_oracle as spo
        lg, dn = logits[:a.cpu_frames].cpu().numpy(), dense[:a.cpu_frames].cpu().numpy()
        t0 = time.perf_counter()
        spo.post_network(lg, dn, m.config)
        res["cpu_port_post_network_ms_per_frame"] = round((time.perf_counter() - t0) * 1e3 / a.cpu_frames, 2)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
