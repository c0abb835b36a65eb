"""CPU tests: the oracle restatement against the golden fixtures made from the live reference."""
import os

import numpy as np
import pytest
import torch

from onepose_b200 import synth
from oracle import gats_spg_oracle as orc
from tests import helpers

mg = helpers.load_make_golden()


@pytest.mark.parametrize("case", list(mg.MATCHER_CASES))
def test_matcher_oracle_vs_golden(case):
    hp, w, obj, fr, data = mg.matcher_inputs(case)
    g = np.load(os.path.join(helpers.GOLDEN, case + ".npz"))
    with torch.no_grad():
        pred, conf = orc.matcher_forward(data, w, hp)
    assert conf.shape == g["conf"].shape and conf.dtype == torch.float32
    assert helpers.conf_rel_err(conf, g["conf"]) < helpers.CONF_RTOL
    helpers.assert_matches_equal_off_ties(pred["matches0"], g["matches0"], g["conf"][0], axis=2)
    helpers.assert_matches_equal_off_ties(pred["matches1"], g["matches1"], g["conf"][0], axis=1)
    np.testing.assert_allclose(pred["matching_scores0"].numpy(), g["matching_scores0"], rtol=1e-3, atol=1e-30)
    assert pred["matches0"].dtype == torch.int64


def test_layer_units_vs_golden():
    g = np.load(os.path.join(helpers.GOLDEN, "layer_units.npz"))
    rng = np.random.default_rng(77)
    w = synth.make_weights(6)
    leaves = torch.from_numpy(rng.standard_normal((2, 24, 5, 256)).astype(np.float32))
    h3 = torch.from_numpy(rng.standard_normal((2, 24, 256)).astype(np.float32))
    out = orc.gats_leaf_aggregate(leaves, h3, w["gnn.layers.0.W"], w["gnn.layers.0.a"])
    np.testing.assert_allclose(out.numpy(), g["gats_out"], rtol=2e-5, atol=2e-5)
    x = torch.from_numpy(rng.standard_normal((2, 256, 40)).astype(np.float32))
    src = torch.from_numpy(rng.standard_normal((2, 256, 72)).astype(np.float32))
    d = orc.attention_propagation(x.transpose(1, 2), src.transpose(1, 2), w, "gnn.layers.1.")
    np.testing.assert_allclose(d.transpose(1, 2).numpy(), g["attn_out"], rtol=1e-4, atol=2e-5)


def test_empty_input_early_out():
    # GATs_SuperGlue.py:223-231: a single dict (not a tuple) with skip_train
    hp = dict(synth.DEFAULT_HPARAMS)
    data = {"keypoints2d": torch.zeros(1, 0, 2), "keypoints3d": torch.zeros(1, 7, 3),
            "descriptors2d_query": torch.zeros(1, 256, 0), "descriptors3d_db": torch.zeros(1, 256, 7),
            "descriptors2d_db": torch.zeros(1, 256, 14)}
    out = orc.matcher_forward(data, synth.make_weights(0), hp)
    assert isinstance(out, dict) and out["skip_train"] is True
    assert out["matches0"].shape == (0,) and out["matches1"].shape == (7,)
    assert (out["matches1"] == -1).all() and out["matches0"].dtype == torch.int32


@pytest.mark.skipif(not os.path.isdir("/root/reference/src"), reason="live reference only in the build container")
def test_oracle_vs_live_reference():
    import sys
    sys.path.insert(0, "/root/reference")
    from src.models.GATsSPG_architectures.GATs_SuperGlue import GATsSuperGlue
    hp = dict(synth.DEFAULT_HPARAMS)
    w = synth.make_weights(11, damped=True)
    ref = GATsSuperGlue(hp).eval()
    ref.load_state_dict(w, strict=True)
    obj = synth.make_object(9, 300, 8)
    data = synth.pack_batch(obj, [synth.make_frame(obj, f, 200) for f in range(2)])
    with torch.no_grad():
        pr, cr = ref(data)
        po, co = orc.matcher_forward(data, w, hp)
    assert helpers.conf_rel_err(co, cr) < helpers.CONF_RTOL
    assert (pr["matches0"] == po["matches0"]).all() and (pr["matches1"] == po["matches1"]).all()
    assert int((pr["matches0"] > -1).sum()) > 0
