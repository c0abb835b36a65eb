"""Diagnostic: SuperPoint block + counts -> matcher; ragged batch vs frame-by-frame vs the float64 oracle."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from onepose_b200 import synth
from onepose_b200.matcher import GATsSuperGlue
from onepose_b200.superpoint import SuperPoint
from oracle import gats_spg_oracle as orc
from tests import helpers

torch.manual_seed(3)
sp = SuperPoint({'descriptor_dim': 256, 'nms_radius': 3, 'max_keypoints': 512, 'keypoints_threshold': 0.6}).eval().cuda()
rs = np.random.RandomState(1)
imgs = rs.randint(0, 256, (3, 1, 160, 192)).astype(np.float32) / 255.0
out = sp.forward_batched(torch.from_numpy(imgs).cuda())
out["count"] = torch.minimum(out["count"], torch.tensor([512, 300, 77], dtype=out["count"].dtype, device="cuda"))
counts = out["count"].tolist()
obj = synth.make_object(0, 512, 4)
db = synth.pack_batch(obj, [synth.make_frame(obj, 0, 8)] * 3)
for damped in (True, False):
    hp = dict(synth.DEFAULT_HPARAMS); hp["match_threshold"] = 0.0
    w = synth.make_weights(0, damped=damped)
    m = GATsSuperGlue(hp).cuda().eval(); m.load_state_dict(w, strict=True)
    data = {k: db[k].cuda() for k in ("keypoints3d", "descriptors3d_db", "descriptors2d_db")}
    data.update(keypoints2d=out["keypoints"], descriptors2d_query=out["descriptors_pm"].transpose(1, 2), count2d=out["count"])
    pred, conf = m(data)
    conf = conf.clone()
    for b, c in enumerate(counts):
        one = {k: v[b:b + 1] for k, v in data.items() if k != "count2d"}
        one["keypoints2d"] = one["keypoints2d"][:, :c]
        one["descriptors2d_query"] = one["descriptors2d_query"][:, :, :c].contiguous()
        p1, c1 = m(one)
        cpu = {k: v.cpu() for k, v in one.items()}
        _, co64 = orc.matcher_forward(cpu, w, hp, dtype=torch.float64)
        _, co32 = orc.matcher_forward(cpu, w, hp)
        e = helpers.conf_rel_err
        print(f"damped={damped} b={b} c={c}: ragged-vs-single {e(conf[b,:c].cpu(), c1[0].cpu()):.2e}  ragged-vs-f64 {e(conf[b,:c].cpu(), co64[0]):.2e}  "
              f"single-vs-f64 {e(c1[0].cpu(), co64[0]):.2e}  oracle32-vs-f64 {e(co32[0], co64[0]):.2e}  bits-equal {bool(torch.equal(conf[b,:c], c1[0]))}")
