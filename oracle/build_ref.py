#!/usr/bin/env python3
"""build_ref.py — TEST INFRASTRUCTURE (oracle/): compile the reference's own
sources into oracle/_ref/ (git-ignored, travels to the GPU box like our .so).

  oracle/_ref/libsphref.so  /root/reference/src/sphFluid.cl compiled AS C++ through
                            ref_shim/cl_shim.hpp + ref_shim/ref_driver.cpp
  oracle/_ref/libowgen.so   /root/reference/src/owHelper.cpp (scene generator) compiled
                            unchanged + ref_shim/owgen_driver.cpp

The kernel text is read where it lies and streamed through two mechanical
rewrites into a temporary file that is deleted after the compile — no reference
source is kept in the repo:
  1. `(float4)( … )` / `(int4)( … )` vector literals -> `float4( … )` / `int4( … )`
  2. the hard-coded `& 0xffff;` hash mask (sphFluid.cl:229,377) -> `& cl_shim_hash_mask;`
     (run-time parameter, default 0xffff; needed for the synthetic >65 536-cell configs)

Does nothing (returns False) when /root/reference is absent, e.g. on the GPU box.
"""
import os
import re
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("SPH_REFERENCE_DIR", "/root/reference")
OUT = os.path.join(HERE, "_ref")
CXXFLAGS = ["-std=c++17", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-fPIC", "-shared", "-w"]


def _newer(target, *srcs):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in srcs if os.path.exists(s))


def build(force=False, verbose=True):
    cl = os.path.join(REF, "src", "sphFluid.cl")
    helper = os.path.join(REF, "src", "owHelper.cpp")
    if not os.path.exists(cl):
        if verbose:
            print("[build_ref] %s absent: keeping whatever is in oracle/_ref/" % cl)
        return False
    os.makedirs(OUT, exist_ok=True)
    shim = os.path.join(HERE, "ref_shim")
    inc = os.path.join(HERE, "..", "include")
    lib = os.path.join(OUT, "libsphref.so")
    deps = [cl, os.path.join(shim, "cl_shim.hpp"), os.path.join(shim, "ref_driver.cpp"),
            os.path.join(inc, "sph_capi.h"), __file__]
    if force or not _newer(lib, *deps):
        text = open(cl, encoding="latin-1").read()
        text, n1 = re.subn(r"\((float4|int4)\)\s*\(", r"\1(", text)
        text, n2 = re.subn(r"&\s*0xffff;", "& cl_shim_hash_mask;", text)
        assert n1 == 37 and n2 == 2, (n1, n2)  # 36 float4 + 1 int4 literals; 2 mask sites
        with tempfile.NamedTemporaryFile("w", suffix=".inc", dir=OUT, delete=False, encoding="latin-1") as f:
            f.write(text)
            tmp = f.name
        try:
            cmd = ["g++"] + CXXFLAGS + ["-I", shim, "-I", inc, "-I", os.path.join(REF, "src"),
                                        '-DSPH_CL_INC="%s"' % tmp,
                                        os.path.join(shim, "ref_driver.cpp"), "-o", lib]
            if verbose:
                print("[build_ref]", " ".join(cmd))
            subprocess.check_call(cmd)
        finally:
            os.unlink(tmp)
    gen = os.path.join(OUT, "libowgen.so")
    if force or not _newer(gen, helper, os.path.join(shim, "owgen_driver.cpp"), __file__):
        cmd = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-w",
               "-I", os.path.join(REF, "src"), os.path.join(shim, "owgen_driver.cpp"), helper, "-o", gen]
        if verbose:
            print("[build_ref]", " ".join(cmd))
        subprocess.check_call(cmd)
    return True


if __name__ == "__main__":
    ok = build(force="--force" in sys.argv)
    sys.exit(0 if ok or not os.path.exists(REF) else 1)
