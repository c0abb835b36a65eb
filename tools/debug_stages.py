"""Developer aid: per-layer difference between the CUDA path and the oracle (run on a GPU box)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from onepose_b200 import synth
from onepose_b200.matcher import GATsSuperGlue
from oracle import gats_spg_oracle as orc

torch.set_grad_enabled(False)
damped = "--damped" in sys.argv
hp = dict(synth.DEFAULT_HPARAMS)
w = synth.make_weights(0, damped=damped)
obj = synth.make_object(0, 160, 8)
frames = [synth.make_frame(obj, f, 96) for f in range(2)]
data = synth.pack_batch(obj, frames)
d2, d3, leaves = orc.to_point_major(data)
ref = []
for i, kind in enumerate(orc.LAYER_KINDS):
    p = f"gnn.layers.{i}."
    if kind == "GATs":
        d3 = orc.gats_leaf_aggregate(leaves, d3, w[p + "W"], w[p + "a"])
    else:
        s0, s1 = (d3, d2) if kind == "cross" else (d2, d3)
        d2, d3 = d2 + orc.attention_propagation(d2, s0, w, p), d3 + orc.attention_propagation(d3, s1, w, p)
    ref.append((d2.clone(), d3.clone()))
m = GATsSuperGlue(hp).cuda().eval(); m.load_state_dict(w)
m._trace = []
pred, conf = m({k: v.cuda() for k, v in data.items()})
for (li, x2, x3), (r2, r3) in zip(m._trace, ref):
    e2 = (x2.cpu() - r2).abs().max().item() / r2.abs().max().item()
    e3 = (x3.cpu() - r3).abs().max().item() / r3.abs().max().item()
    print(f"layer {li:2d} {orc.LAYER_KINDS[li]:5s} rel-max err x2 {e2:.3e} x3 {e3:.3e}")
_, co = orc.matcher_forward(data, w, hp)
print("conf rel err", ((conf.cpu().double() - co.double()).abs() / co.double().clamp_min(1e-30)).max().item())
