"""Compile the CPU oracle (oracle/nfb_oracle.c) with gcc into oracle/_build/libnfb_oracle.so.

TEST INFRASTRUCTURE ONLY — see the header of nfb_oracle.c.
-ffp-contract=off keeps gcc from fusing a*b+c on its own; the only fused
multiply-adds are the explicit fmaf() calls that mirror nvcc's contractions of
the reference kernels.
"""
import hashlib
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "nfb_oracle.c")
OUT_DIR = os.path.join(HERE, "_build")
LIB = os.path.join(OUT_DIR, "libnfb_oracle.so")
STAMP = os.path.join(OUT_DIR, "libnfb_oracle.sha1")

FLAGS = ["-O2", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-shared", "-fPIC",
         "-fvisibility=hidden", "-std=gnu11", "-Wall", "-Wno-unknown-pragmas", "-Wno-maybe-uninitialized"]


def _digest() -> str:
    h = hashlib.sha1()
    with open(SRC, "rb") as f:
        h.update(f.read())
    h.update(" ".join(FLAGS).encode())
    return h.hexdigest()


def build(force: bool = False) -> str:
    os.makedirs(OUT_DIR, exist_ok=True)
    dig = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(STAMP):
        with open(STAMP) as f:
            if f.read().strip() == dig:
                return LIB
    cmd = ["gcc", *FLAGS, SRC, "-o", LIB, "-lm"]
    subprocess.run(cmd, check=True)
    with open(STAMP, "w") as f:
        f.write(dig)
    return LIB


if __name__ == "__main__":
    print(build(force=True))
