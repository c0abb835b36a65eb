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
