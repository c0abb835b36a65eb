"""Shared comparison utilities for the parity tests."""
import importlib.util
import os
import sys

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# tolerance stated by BASELINE.json north_star: conf_matrix within 1e-3 relative error
CONF_RTOL = 1e-3
# conf entries below this magnitude are compared absolutely (fp32 cannot resolve them relatively
# once exp() underflows into denormals; reference entries span ~30 orders of magnitude)
CONF_FLOOR = 1e-30


def load_make_golden():
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(GOLDEN, "make_golden.py"))
    mod = importlib.util.module_from_spec(spec)
    # make_golden inserts /root/reference in sys.path; harmless when absent
    spec.loader.exec_module(mod)
    return mod


def conf_rel_err(conf, conf_ref):
    c = torch.as_tensor(np.asarray(conf)).double()
    r = torch.as_tensor(np.asarray(conf_ref)).double()
    rel = (c - r).abs() / r.clamp_min(CONF_FLOOR)
    return rel.max().item()


def assert_matches_equal_off_ties(m, m_ref, conf_ref, axis, tie_rtol=CONF_RTOL):
    """Index vectors must be identical except where the oracle's top-2 gap is a near-tie.

    ``axis=2``: matches0 (per query row, best 3D column); ``axis=1``: matches1.
    conf_ref is [N2, N3] (batch element 0).
    """
    m = np.asarray(m); m_ref = np.asarray(m_ref)
    bad = np.nonzero(m != m_ref)[0]
    if len(bad) == 0:
        return 0
    c = torch.as_tensor(np.asarray(conf_ref)).double()
    c = c if axis == 2 else c.t()
    excused = 0
    for i in bad:
        top2 = torch.topk(c[i], 2).values
        near_tie = (top2[0] - top2[1]) <= tie_rtol * top2[0]
        # a flipped argmax elsewhere can also flip the mutual check of this entry
        assert near_tie or _partner_tie(c, m, m_ref, i, tie_rtol), f"index {i}: {m[i]} vs {m_ref[i]} not a tie"
        excused += 1
    return excused


def _partner_tie(c, m, m_ref, i, tie_rtol):
    j = max(int(m[i]), int(m_ref[i]))
    if j < 0:
        return False
    col = c[:, j]
    top2 = torch.topk(col, 2).values
    return bool((top2[0] - top2[1]) <= tie_rtol * top2[0])
