"""One deterministic step for ncu: 3 warm-up forwards + 1 forward of the headline workload on one CUDA stream.
   Launches matching k_linear_tc|k_sim_tc per forward: 66 + 2; the last seven of a forward are
   q(3D), mlp0(3D), mlp1(3D), final(2D), final(3D), k_sim_tc<1>, k_sim_tc<2>."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from onepose_b200 import synth                      # noqa: E402
from onepose_b200.matcher import GATsSuperGlue      # noqa: E402

F, N2, N3 = 64, 2048, 8192
dev = torch.device("cuda", 0)
m = GATsSuperGlue(dict(synth.DEFAULT_HPARAMS)).to(dev).eval()
m.load_state_dict(synth.make_weights(0, damped=True), strict=True)
m.two_streams = False
obj = synth.make_object(0, N3, 8)
frames = [synth.make_frame(obj, f, N2) for f in range(F)]
host = synth.pack_batch(obj, frames)
data = {k: host[k][:1].to(dev).expand(F, *host[k].shape[1:]) for k in ("keypoints3d", "descriptors3d_db", "descriptors2d_db")}
data.update({k: host[k].to(dev) for k in ("keypoints2d", "descriptors2d_query")})
data["query_intrinsic"] = torch.from_numpy(frames[0].K)[None].to(dev)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 4):
    m(data)
torch.cuda.synchronize()
print("done")
