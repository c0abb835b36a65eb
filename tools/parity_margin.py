"""conf_matrix error of the CUDA path against the committed golden fixtures and the oracle (how much of the 1e-3 bound is used).
   OPB_SIM_MMA=f16 python tools/parity_margin.py    # three fp16 MMAs per product in the similarity kernel, for comparison"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from onepose_b200 import synth
from onepose_b200.matcher import GATsSuperGlue
from oracle import gats_spg_oracle as orc
from tests import helpers
from tests.golden import make_golden as mg


def model(hp, w):
    m = GATsSuperGlue(hp).cuda().eval(); m.load_state_dict(w, strict=True); return m


for case in mg.MATCHER_CASES:
    hp, w, obj, fr, data = mg.matcher_inputs(case)
    g = np.load(os.path.join(helpers.GOLDEN, case + ".npz"))
    _, conf = model(hp, w)({k: v.cuda() for k, v in data.items()})
    print(f"golden {case}: conf rel err {float(helpers.conf_rel_err(conf.cpu(), g['conf'])):.2e}")
for damped in (False, True):
    hp = dict(synth.DEFAULT_HPARAMS)
    w = synth.make_weights(0, damped=damped)
    obj = synth.make_object(0, 4096, 8)
    data = synth.pack_batch(obj, [synth.make_frame(obj, 0, 1024)])
    _, co = orc.matcher_forward(data, w, hp, dtype=torch.float64)
    _, conf = model(hp, w)({k: v.cuda() for k, v in data.items()})
    print(f"configs[0] 1024 x 4096 x 8, damped={damped}: conf rel err vs float64 oracle {float(helpers.conf_rel_err(conf.cpu(), co)):.2e}")

# the shapes the bench numbers are quoted on: BASELINE.json configs[1] (headline, two frames) and configs[2] (large object, one frame)
for name, (oid, n2, n3, L, nf) in {"configs[1] 2048 x 8192 x 8": (0, 2048, 8192, 8, 2), "configs[2] 4096 x 32768 x 16": (1, 4096, 32768, 16, 1)}.items():
    for damped in (False, True):
        hp = dict(synth.DEFAULT_HPARAMS)
        w = synth.make_weights(0, damped=damped)
        obj = synth.make_object(oid, n3, L)
        data = synth.pack_batch(obj, [synth.make_frame(obj, f, n2) for f in range(nf)])
        _, conf = model(hp, w)({k: v.cuda() for k, v in data.items()})
        conf = conf.cpu()
        _, c32 = orc.matcher_forward(data, w, hp)
        e32 = float(helpers.conf_rel_err(conf, c32))
        _, c64 = orc.matcher_forward(data, w, hp, dtype=torch.float64)
        print(f"{name}, damped={damped}: conf rel err vs fp32 oracle {e32:.2e}, vs float64 oracle {float(helpers.conf_rel_err(conf, c64)):.2e}; "
              f"fp32 oracle vs float64 oracle {float(helpers.conf_rel_err(c32, c64)):.2e}", flush=True)
        del c32, c64, conf
