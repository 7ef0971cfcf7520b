#!/usr/bin/env python3
"""build_oracle.py — TEST INFRASTRUCTURE: compile oracle/sph_oracle.c into oracle/liboracle.so.

Flags mirror the reference's kernel build (no FMA contraction, no fast-math):
gcc -std=gnu99 -O2 -ffp-contract=off -fno-fast-math -fopenmp.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liboracle.so")


def build(force=False, verbose=False):
    src = os.path.join(HERE, "sph_oracle.c")
    hdr = os.path.join(HERE, "..", "include", "sph_capi.h")
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= max(os.path.getmtime(src), os.path.getmtime(hdr)):
        return LIB
    cmd = ["gcc", "-std=gnu99", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-fPIC", "-shared",
           "-Wall", "-Wextra", "-I", os.path.join(HERE, "..", "include"), src, "-o", LIB, "-lm"]
    if verbose:
        print("[build_oracle]", " ".join(cmd))
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
