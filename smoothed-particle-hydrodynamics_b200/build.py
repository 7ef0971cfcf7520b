"""Build libsph_b200.so (CUDA sm_100a) in-tree with nvcc.

Numerics flags are part of the parity contract (DESIGN.md "Numerics"):
  -fmad=false -prec-div=true -prec-sqrt=true -ftz=false
so every float op is a single IEEE round-to-nearest operation, as in the reference's
-cl-opt-disable kernels.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
INC = os.path.join(PKG, "..", "include")
LIB = os.path.join(PKG, "libsph_b200.so")
OBJ = os.path.join(PKG, "build")
SOURCES = ["sph_sort.cu", "sph_neighbors.cu", "sph_pcisph.cu", "sph_lanes.cu", "sph_membranes.cu", "sph_slab.cu", "sph_noalloc.cu", "sph_capi.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
         "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math", "-I", CSRC, "-I", INC]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, ptxas_info=False):
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers += [os.path.join(INC, "sph_capi.h"), os.path.abspath(__file__)]
    objs, jobs = [], []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(OBJ, s.replace(".cu", ".o"))
        objs.append(obj)
        if force or _stale(obj, [src] + headers):
            cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if ptxas_info else []) + ["-c", src, "-o", obj]
            jobs.append(cmd)
    def run(cmd):
        if verbose:
            print("[build]", " ".join(cmd))
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0 or ptxas_info:
            sys.stderr.write(r.stdout + r.stderr)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed: " + " ".join(cmd))
    with ThreadPoolExecutor(max_workers=8) as ex:
        list(ex.map(run, jobs))
    if jobs or force or _stale(LIB, objs):
        run([NVCC, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lcudart_static", "-ldl", "-lrt", "-lpthread"])
    return LIB


def build_variant(name, defines, verbose=False):
    """Experiment builds: the library with extra -D macros -> build_exp/libsph_<name>.so (git-ignored; selected at run
    time with SPH_LIB=<path>, see capi.load_product). Not part of the product build."""
    out_dir = os.path.join(PKG, "build_exp")
    os.makedirs(out_dir, exist_ok=True)
    objs = []
    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("nvcc failed: " + " ".join(cmd))
    cmds = []
    for s in SOURCES:
        obj = os.path.join(out_dir, name + "_" + s.replace(".cu", ".o"))
        objs.append(obj)
        cmds.append([NVCC] + FLAGS + ["-D" + d for d in defines] + ["-c", os.path.join(CSRC, s), "-o", obj])
    with ThreadPoolExecutor(max_workers=8) as ex:
        list(ex.map(run, cmds))
    lib = os.path.join(out_dir, "libsph_%s.so" % name)
    run([NVCC, "-shared", "-o", lib] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lcudart_static", "-ldl", "-lrt", "-lpthread"])
    for o in objs:
        os.remove(o)
    return lib


HOST = os.path.join(PKG, "host")
HOST_LIB = os.path.join(PKG, "libow_b200.so")
HOST_BIN = os.path.join(PKG, "sibernetic_b200")
HOST_SOURCES = ["owOpenCLSolver.cpp", "owHelper.cpp", "owPhysicsFluidSimulator.cpp", "PyramidalSimulation.cpp", "ow_capi.cpp"]


def build_host(force=False, verbose=False):
    """The drop-in C++ classes (owPhysicsFluidSimulator / owOpenCLSolver / owHelper / PyramidalSimulation)
    over the C-ABI -> libow_b200.so, and the headless CLI driver -> sibernetic_b200."""
    import sysconfig
    build()
    srcs = [os.path.join(HOST, f) for f in HOST_SOURCES]
    hdrs = [os.path.join(INC, "ow", f) for f in os.listdir(os.path.join(INC, "ow"))] + [os.path.join(INC, "sph_capi.h")]
    pyinc = sysconfig.get_paths()["include"]
    pyver = "python%d.%d" % sys.version_info[:2]
    common = ["g++", "-std=c++17", "-O2", "-fPIC", "-Wall", "-I", INC, "-I", pyinc]
    link = ["-L", PKG, "-lsph_b200", "-Wl,-rpath,$ORIGIN", "-l" + pyver, "-ldl", "-lm", "-lpthread"]
    def run(cmd):
        if verbose:
            print("[build]", " ".join(cmd))
        subprocess.check_call(cmd)
    if force or _stale(HOST_LIB, srcs + hdrs + [LIB]):
        run(common + ["-shared"] + srcs + ["-o", HOST_LIB] + link)
    if force or _stale(HOST_BIN, [os.path.join(HOST, "main.cpp"), HOST_LIB] + hdrs):
        run(common + [os.path.join(HOST, "main.cpp"), "-o", HOST_BIN, "-L", PKG, "-low_b200"] + link)
    return HOST_LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True, ptxas_info="-v" in sys.argv))
    print(build_host(force="--force" in sys.argv, verbose=True))
