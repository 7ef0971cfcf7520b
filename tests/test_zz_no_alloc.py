"""GPU test of the allocation-free entry point sph_cell_start_noalloc (include/sph_capi.h; no counterpart in the
reference). Through the C-ABI, on two calls in a row with different
inputs: its output must equal, bit for bit (integer work: tolerance 0, like every other test here), what the existing
entry points produce — SPH_STAGE_HASH_PARTICLES + SPH_STAGE_SORT, then SPH_BUF_GRID_CELL_INDEX_FIXED_UP.
Collected last (file name): a failure here hides no other test under `pytest -x`."""
import ctypes as C

import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu


def _bind(P):
    L = P.lib
    L.sph_cell_start_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    L.sph_noalloc_last_error.restype = C.c_char_p
    L.sph_device_alloc.argtypes = [C.c_size_t]
    L.sph_device_alloc.restype = C.c_void_p
    L.sph_device_free.argtypes = [C.c_void_p]
    L.sph_device_free.restype = None
    L.sph_device_copy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
    return L


def _existing(P, cfg):
    ctx = common.create(P, cfg)
    ctx.stage("HASH_PARTICLES")
    ctx.stage("SORT")
    out = ctx.read("gridCellIndexFixedUp")
    ctx.close()
    return out


def test_noalloc_cell_start_matches_existing_entry_points_twice():
    P = common.product_lib()
    L = _bind(P)
    # two different scenes, the second with FEWER cells and particles than the first: stale counts of the first call
    # in the shared static histogram would show
    cfgs = [common.config_a(P)]
    pos, vel = common.synthetic_block(16, 12, 20, (24, 20, 48), seed=11)
    p2 = common.fill_counts(P.default_params(box=(12, 10, 24)), pos)
    p2.hash_mask = 0xFFFFFFFF
    cfgs.append(dict(params=p2, position=pos, velocity=vel))
    cfgs.append(cfgs[0])   # and the first again, after the smaller one
    for k, cfg in enumerate(cfgs):
        want = _existing(P, cfg)
        p, pos = cfg["params"], np.ascontiguousarray(cfg["position"], np.float32)
        n, words = len(pos), int(p.grid_cell_count) + 1
        assert len(want) == words
        d_pos, d_out = L.sph_device_alloc(pos.nbytes), L.sph_device_alloc(4 * words)   # the CALLER's buffers
        assert d_pos and d_out
        try:
            assert L.sph_device_copy(d_pos, pos.ctypes.data_as(C.c_void_p), pos.nbytes) == 0
            rc = L.sph_cell_start_noalloc(C.byref(p), d_pos, n, d_out, None)
            assert rc == 0, L.sph_noalloc_last_error().decode()
            got = np.empty(words, np.uint32)
            assert L.sph_device_copy(got.ctypes.data_as(C.c_void_p), d_out, got.nbytes) == 0   # synchronises
        finally:
            L.sph_device_free(d_pos)
            L.sph_device_free(d_out)
        assert got[-1] == n and got[0] == 0
        assert np.array_equal(got, want), "call %d: %d cells differ" % (k, int((got != want).sum()))


def test_noalloc_rejects_grids_beyond_the_static_scratch():
    P = common.product_lib()
    L = _bind(P)
    p = P.default_params(box=(400, 400, 400))   # 401^3 = 64 M cells > 2^20
    assert p.grid_cell_count + 1 > L.sph_noalloc_max_cells()
    d = L.sph_device_alloc(64)
    try:
        rc = L.sph_cell_start_noalloc(C.byref(p), d, 4, d, None)
    finally:
        L.sph_device_free(d)
    assert rc == -5   # SPH_E_SIZE
    assert "exceeds the static scratch" in L.sph_noalloc_last_error().decode()
