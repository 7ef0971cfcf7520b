"""CPU tests: the C-ABI library loads and exports every symbol include/sph_capi.h declares
(no compute without a GPU), and the product refuses to run without a device."""
import ctypes
import os
import re

import pytest

import common
from common import capi

HEADER = os.path.join(common.ROOT, "include", "sph_capi.h")


def declared_symbols(part="all"):
    """part: 'core' = the reference-facing ABI (all three libraries), 'all' = + instrumentation (product only)."""
    text = open(HEADER).read()
    core, _, instr = text.partition("---- instrumentation")
    text = core if part == "core" else text
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sph_[a-z_]+)\s*\(", text)))


def _build_product():
    import importlib
    b = importlib.import_module(common.PKG + ".build")
    return b.build()


def test_header_symbols_exported_by_all_libraries():
    core, full = declared_symbols("core"), declared_symbols("all")
    assert "sph_create" in core and "sph_step" in core and len(core) >= 14
    assert "sph_run_steps" in full and "sph_run_steps" not in core
    libs = [(_build_product(), full), (common.oracle_lib().path, core)]
    if common.ref_lib() is not None:
        libs.append((common.REF_LIB, core))
    for path, syms in libs:
        L = ctypes.CDLL(path)
        for s in syms:
            assert hasattr(L, s), "%s does not export %s" % (path, s)


def test_product_constants_without_gpu():
    P = capi.SphLib(_build_product())
    assert P.backend == "cuda-sm_100a"
    assert P.default_params().as_bytes().hex() == common.digests()["params_default_hex"]
    assert P.default_params(box=(36, 24, 54)).as_bytes().hex() == common.digests()["params_boxA_hex"]


def test_product_fails_loudly_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    P = capi.SphLib(_build_product())
    cfg = common.config_a(P)
    with pytest.raises(capi.SphError) as e:
        common.create(P, cfg)
    assert "no CUDA device" in str(e.value) or "cuda" in str(e.value).lower()
