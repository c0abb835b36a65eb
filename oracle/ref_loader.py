"""Loader of the staged reference implementation (``oracle/_ref/``, see ``build_oracle.build_ref``).

TEST INFRASTRUCTURE ONLY: used by the CPU legs of ``bench.py`` (``--impl reference``, ``cpu_baseline``) and by tests.
``load()`` returns ``(GATsSuperGlue, ransac_PnP)`` -- the reference's own module class
(src/models/GATsSPG_architectures/GATs_SuperGlue.py:162-278) and PnP wrapper (src/utils/eval_utils.py:18-42) -- or
``None`` when the files were not staged (then the callers fall back to the restatement and report ``kind: "port"``).
"""
import importlib
import os
import sys

from . import build_oracle


def load():
    if not build_oracle.ref_available():
        return None
    root = build_oracle.REF_DIR
    if root not in sys.path:
        sys.path.insert(0, root)
    for name in [m for m in sys.modules if m == "src" or m.startswith("src.")]:      # a different tree's `src` package
        if not getattr(sys.modules[name], "__file__", "").startswith(root):
            del sys.modules[name]
    g = importlib.import_module("src.models.GATsSPG_architectures.GATs_SuperGlue")
    e = importlib.import_module("src.utils.eval_utils")
    return g.GATsSuperGlue, e.ransac_PnP
