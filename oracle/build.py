"""Build the CPU oracle (TEST INFRASTRUCTURE ONLY) into oracle/_build/libnlos_oracle.so.

gcc only; -ffp-contract=off so that culling arithmetic is single IEEE ops.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT_DIR = os.path.join(HERE, "_build")
LIB = os.path.join(OUT_DIR, "libnlos_oracle.so")
SRCS = [os.path.join(HERE, "nlos_oracle.c"), os.path.join(HERE, "nlos_oracle_body.inc")]


def build(force: bool = False) -> str:
    os.makedirs(OUT_DIR, exist_ok=True)
    if not force and os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(s) for s in SRCS):
        return LIB
    cmd = ["gcc", "-O2", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off", "-fno-fast-math",
           "-march=x86-64-v2", "-Wall", "-Wno-unused-function", "-o", LIB, SRCS[0], "-lm"]
    subprocess.check_call(cmd, cwd=HERE)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
