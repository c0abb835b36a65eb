"""Build recipe for the C part of the oracle (test infrastructure, never shipped in the product).

``python -m oracle.build_oracle`` compiles ``oracle/pnp_oracle.c`` into
``oracle/_build/libpnp_oracle.so`` with gcc.  ``-ffp-contract=off`` keeps the float32
reprojection-error arithmetic free of FMA contraction, as in the x86 OpenCV build it restates.
The reference itself is pure Python (no C sources to compile), so there is no ``oracle/_ref``.
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "pnp_oracle.c")
OUT_DIR = os.path.join(HERE, "_build")
OUT = os.path.join(OUT_DIR, "libpnp_oracle.so")


def build(force=False):
    os.makedirs(OUT_DIR, exist_ok=True)
    if not force and os.path.exists(OUT) and os.path.getmtime(OUT) >= os.path.getmtime(SRC):
        return OUT
    cmd = ["gcc", "-O2", "-ffp-contract=off", "-fno-fast-math", "-shared", "-fPIC", "-o", OUT, SRC, "-lm"]
    subprocess.run(cmd, check=True)
    return OUT


if __name__ == "__main__":
    print(build(force=True))
