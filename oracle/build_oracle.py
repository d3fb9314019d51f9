#!/usr/bin/env python
"""TEST INFRASTRUCTURE ONLY -- compiles oracle/h3dgs_oracle.c (the CPU restatement) with gcc into
oracle/_build/liborc.so.  -ffp-contract=off: the FFMA contractions of the reference build are written
out with fmaf() in the source; the compiler must not add its own."""
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "h3dgs_oracle.c")
OUT_DIR = os.path.join(HERE, "_build")
LIB = os.path.join(OUT_DIR, "liborc.so")
# -mfma makes fmaf() one instruction instead of a libm call; every x86-64 server CPU since 2013 has it.
FLAGS = ["-O2", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-fPIC", "-shared", "-fvisibility=hidden",
         "-Wall", "-Wno-comment"]


def _cpu_flags():
    try:
        txt = open("/proc/cpuinfo").read()
        return ["-mfma"] if " fma " in txt else []
    except OSError:
        return []


def build(force=False):
    os.makedirs(OUT_DIR, exist_ok=True)
    flags = FLAGS + _cpu_flags()
    dg = hashlib.sha256(open(SRC, "rb").read() + " ".join(flags).encode()).hexdigest()
    stamp = LIB + ".sha256"
    if not force and os.path.exists(LIB) and os.path.exists(stamp) and open(stamp).read() == dg:
        return LIB
    cmd = ["gcc"] + flags + ["-o", LIB, SRC, "-lm"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("gcc failed:\n" + r.stdout + r.stderr)
    with open(stamp, "w") as fh:
        fh.write(dg)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
