"""Build recipe for the compiled / staged parts of the oracle (test infrastructure, never shipped in the product).

``python -m oracle.build_oracle`` does two things:

* ``build()``     compiles ``oracle/pnp_oracle.c`` into ``oracle/_build/libpnp_oracle.so`` with gcc.
  ``-ffp-contract=off`` keeps the float32 reprojection-error arithmetic free of FMA contraction, as in the x86
  OpenCV build it restates.
* ``build_ref()`` stages the reference's OWN implementation of the hot path under ``oracle/_ref/`` when
  ``/root/reference`` is present (the build container): the three pure-Python files the path consists of
  (``GATs_SuperGlue.py``, ``GATs.py``, ``eval_utils.py``) are copied unmodified, from where they lie, into a
  package skeleton that mirrors the reference tree.  ``oracle/_ref/`` is git-ignored (no reference source
  enters the history) but travels to the GPU box with the working tree, so that ``bench.py --impl reference``
  and the ``cpu_baseline`` leg time the reference itself there (``kind: "reference"``) instead of the port.
  The reference has no compiled sources; nothing is built.
"""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "pnp_oracle.c")
OUT_DIR = os.path.join(HERE, "_build")
OUT = os.path.join(OUT_DIR, "libpnp_oracle.so")
REF_ROOT = os.environ.get("OPB_REFERENCE_ROOT", "/root/reference")
REF_DIR = os.path.join(HERE, "_ref")
REF_FILES = ("src/models/GATsSPG_architectures/GATs_SuperGlue.py", "src/models/GATsSPG_architectures/GATs.py",
             "src/utils/eval_utils.py")


def build(force=False):
    os.makedirs(OUT_DIR, exist_ok=True)
    if not force and os.path.exists(OUT) and os.path.getmtime(OUT) >= os.path.getmtime(SRC):
        return OUT
    cmd = ["gcc", "-O2", "-ffp-contract=off", "-fno-fast-math", "-shared", "-fPIC", "-o", OUT, SRC, "-lm"]
    subprocess.run(cmd, check=True)
    return OUT


def build_ref():
    """Stage the reference's hot-path files under oracle/_ref/ (no-op when the reference tree is absent)."""
    if not all(os.path.exists(os.path.join(REF_ROOT, f)) for f in REF_FILES):
        return REF_DIR if ref_available() else None
    for f in REF_FILES:
        dst = os.path.join(REF_DIR, f)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(REF_ROOT, f), dst)
        d = os.path.dirname(dst)
        while os.path.abspath(d) != os.path.abspath(REF_DIR):          # package skeleton for the relative imports
            open(os.path.join(d, "__init__.py"), "a").close()
            d = os.path.dirname(d)
    return REF_DIR


def ref_available():
    return all(os.path.exists(os.path.join(REF_DIR, f)) for f in REF_FILES)


if __name__ == "__main__":
    print(build(force=True))
    print(build_ref())
