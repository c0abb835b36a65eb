"""What the PnP overlap costs the matcher: CUDA-graph steps of the headline workload (a) matcher only, (b) matcher + PnP on the same
stream, (c) matcher + PnP on the pose stream (bench.py's default).   python tools/bench_pnp_overlap.py [steps]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from onepose_b200 import synth                      # noqa: E402
from onepose_b200.matcher import GATsSuperGlue      # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
F, N2, N3 = 64, 2048, 8192
dev = torch.device("cuda", 0)
m = GATsSuperGlue(dict(synth.DEFAULT_HPARAMS)).to(dev).eval()
m.load_state_dict(synth.make_weights(0, damped=True), strict=True)
m.enable_graphs(True)
obj = synth.make_object(0, N3, 8)
frames = [synth.make_frame(obj, f, N2) for f in range(F)]
host = synth.pack_batch(obj, frames)
base = {k: host[k][:1].to(dev).expand(F, *host[k].shape[1:]) for k in ("keypoints3d", "descriptors3d_db", "descriptors2d_db")}
base.update({k: host[k].to(dev) for k in ("keypoints2d", "descriptors2d_query")})
K = torch.from_numpy(frames[0].K)[None].to(dev)


def run(label, with_pnp, overlap):
    m.pnp_overlap = overlap
    data = dict(base, query_intrinsic=K) if with_pnp else dict(base)
    for _ in range(4):
        pred, _ = m(data)
    if overlap:
        torch.cuda.current_stream().wait_stream(m.pose_stream(dev))
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(steps):
        pred, _ = m(data)
    if overlap:
        torch.cuda.current_stream().wait_stream(m.pose_stream(dev))
    e.record()
    torch.cuda.synchronize()
    print(f"{label:44s} {s.elapsed_time(e) / steps:8.3f} ms/step", flush=True)


for _ in range(4):
    run("matcher only", False, False)
    run("matcher + PnP, same stream", True, False)
    run("matcher + PnP on the pose stream (overlap)", True, True)
