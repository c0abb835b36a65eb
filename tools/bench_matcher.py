"""A/B timing of the whole matcher step on the headline workload (no per-launch events).
   python tools/bench_matcher.py [steps]      env: FRAMES, N2, N3"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from onepose_b200 import synth                      # noqa: E402
from onepose_b200.matcher import GATsSuperGlue      # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
F, N2, N3 = int(os.environ.get("FRAMES", 64)), int(os.environ.get("N2", 2048)), int(os.environ.get("N3", 8192))
dev = torch.device("cuda", 0)
m = GATsSuperGlue(dict(synth.DEFAULT_HPARAMS)).to(dev).eval()
m.load_state_dict(synth.make_weights(0, damped=True), strict=True)
obj = synth.make_object(0, N3, 8)
frames = [synth.make_frame(obj, f, N2) for f in range(F)]
host = synth.pack_batch(obj, frames)
data = {k: host[k][:1].to(dev).expand(F, *host[k].shape[1:]) for k in ("keypoints3d", "descriptors3d_db", "descriptors2d_db")}
data.update({k: host[k].to(dev) for k in ("keypoints2d", "descriptors2d_query")})
data["query_intrinsic"] = torch.from_numpy(frames[0].K)[None].to(dev)


def run(label, **attrs):
    for k, v in attrs.items():
        setattr(m, k, v)
    for _ in range(3):
        m(data)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(steps):
        m(data)
    e.record()
    torch.cuda.synchronize()
    print(f"{label:30s} {s.elapsed_time(e) / steps:8.3f} ms/step  {F * steps / s.elapsed_time(e) * 1e3:8.1f} frames/s", flush=True)


run("two streams", two_streams=True)
run("one stream", two_streams=False)
run("two streams", two_streams=True)
run("one stream", two_streams=False)

# whole step as one CUDA graph (launch gaps removed): capture once, replay
try:
    m.two_streams = True
    g = torch.cuda.CUDAGraph()
    s_cap = torch.cuda.Stream()
    s_cap.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s_cap):
        m(data)
    torch.cuda.current_stream().wait_stream(s_cap)
    torch.cuda.synchronize()
    with torch.cuda.graph(g):
        out = m(data)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(steps):
        g.replay()
    e.record()
    torch.cuda.synchronize()
    print(f"{'CUDA graph replay':30s} {s.elapsed_time(e) / steps:8.3f} ms/step  {F * steps / s.elapsed_time(e) * 1e3:8.1f} frames/s", flush=True)
    print("poses finite:", bool(torch.isfinite(out[0]["pose"]).all()), "inliers mean:", float(out[0]["n_inliers"].float().mean()))
except Exception as ex:                                  # noqa: BLE001
    print("graph capture failed:", type(ex).__name__, str(ex)[:300])
