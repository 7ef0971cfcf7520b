"""ctypes binding of include/sph_capi.h.

Generic over the library file: the product (`libsph_b200.so`, CUDA sm_100a) and the
two CPU checkers under oracle/ export the same ABI, so tests drive all three
through this one class.  The product path never falls back: `load_product()`
raises if the CUDA library is missing.
"""
import ctypes as C
import os

import numpy as np

MAX_NEIGHBOR_COUNT = 32
MAX_MEMBRANES = 7
LIQUID, ELASTIC, BOUNDARY = 1, 2, 3

STAGES = [
    "CLEAR_BUFFERS", "HASH_PARTICLES", "SORT", "SORT_POST_PASS", "INDEXX", "INDEX_POST_PASS",
    "FIND_NEIGHBORS", "COMPUTE_DENSITY", "COMPUTE_FORCES_INIT_PRESSURE", "COMPUTE_ELASTIC_FORCES",
    "PREDICT_POSITIONS", "PREDICT_DENSITY", "CORRECT_PRESSURE", "COMPUTE_PRESSURE_FORCE_ACCELERATION",
    "INTEGRATE", "CLEAR_MEMBRANE_BUFFERS", "MEMBRANE_INTERACTION", "MEMBRANE_FINALIZE",
]
STAGE = {n: i for i, n in enumerate(STAGES)}

# name -> (id, numpy dtype, trailing shape) in the reference layout
BUFFERS = {
    "position": (0, np.float32, (4,)),
    "velocity": (1, np.float32, (4,)),
    "particleIndex": (2, np.uint32, (2,)),
    "particleIndexBack": (3, np.uint32, ()),
    "sortedPosition": (4, np.float32, (4,)),
    "sortedVelocity": (5, np.float32, (4,)),
    "gridCellIndex": (6, np.uint32, ()),
    "gridCellIndexFixedUp": (7, np.uint32, ()),
    "neighborMap": (8, np.float32, (2,)),
    "rho": (9, np.float32, ()),
    "pressure": (10, np.float32, ()),
    "acceleration": (11, np.float32, (4,)),
    "elasticConnectionsData": (12, np.float32, (4,)),
    "membraneData": (13, np.int32, ()),
    "particleMembranesList": (14, np.int32, ()),
    "muscle_activation_signal": (15, np.float32, ()),
}


class SphParams(C.Structure):
    _fields_ = [
        ("particle_count", C.c_int32), ("num_elastic", C.c_int32), ("num_liquid", C.c_int32),
        ("num_boundary", C.c_int32), ("num_membranes", C.c_int32), ("muscle_count", C.c_int32),
        ("elastic_offset", C.c_int32), ("max_iteration", C.c_int32),
        ("xmin", C.c_float), ("xmax", C.c_float), ("ymin", C.c_float), ("ymax", C.c_float),
        ("zmin", C.c_float), ("zmax", C.c_float),
        ("grid_cells_x", C.c_int32), ("grid_cells_y", C.c_int32), ("grid_cells_z", C.c_int32),
        ("grid_cell_count", C.c_int32), ("hash_mask", C.c_uint32),
        ("h", C.c_float), ("hash_grid_cell_size", C.c_float), ("hash_grid_cell_size_inv", C.c_float),
        ("simulation_scale", C.c_float), ("simulation_scale_inv", C.c_float), ("r0", C.c_float),
        ("rho0", C.c_float), ("mass", C.c_float), ("time_step", C.c_float), ("viscosity", C.c_float),
        ("stiffness", C.c_float), ("damping", C.c_float),
        ("gravity_x", C.c_float), ("gravity_y", C.c_float), ("gravity_z", C.c_float),
        ("delta", C.c_float),
        ("wpoly6_coefficient", C.c_double), ("grad_wspiky_coefficient", C.c_double),
        ("del2_wviscosity_coefficient", C.c_double),
    ]

    def as_bytes(self):
        return bytes(memoryview(self))

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class SphError(RuntimeError):
    pass


class SphLib:
    """One loaded library exporting include/sph_capi.h."""

    def __init__(self, path):
        if not os.path.exists(path):
            raise SphError("library not found: %s" % path)
        self.path = path
        self.lib = L = C.CDLL(path, mode=C.RTLD_LOCAL)
        P = C.POINTER
        L.sph_params_default.argtypes = [P(SphParams)]
        L.sph_params_set_box.argtypes = [P(SphParams), C.c_double, C.c_double, C.c_double]
        L.sph_create.restype = C.c_void_p
        L.sph_create.argtypes = [P(SphParams), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.sph_destroy.argtypes = [C.c_void_p]
        L.sph_destroy.restype = None
        L.sph_stage.argtypes = [C.c_void_p, C.c_int]
        L.sph_step.argtypes = [C.c_void_p]
        L.sph_sync.argtypes = [C.c_void_p]
        L.sph_set_muscles.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.sph_read.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
        L.sph_write.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
        L.sph_buffer_bytes.argtypes = [C.c_void_p, C.c_int]
        L.sph_buffer_bytes.restype = C.c_size_t
        L.sph_get_params.argtypes = [C.c_void_p, P(SphParams)]
        L.sph_last_error.restype = C.c_char_p
        L.sph_backend.restype = C.c_char_p
        self.has_instrumentation = hasattr(L, "sph_run_steps")
        if hasattr(L, "sph_read_begin"):   # product only: pipelined read-back
            L.sph_read_begin.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
            L.sph_read_end.argtypes = [C.c_void_p]
        if self.has_instrumentation:  # CUDA library only
            L.sph_run_steps.argtypes = [C.c_void_p, C.c_int, P(C.c_float)]
            L.sph_profile_enable.argtypes = [C.c_void_p, C.c_int]
            L.sph_profile_report.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
            L.sph_launch_count.argtypes = [C.c_void_p]
            L.sph_launch_count.restype = C.c_longlong
            L.sph_host_alloc.argtypes = [C.c_size_t]
            L.sph_host_alloc.restype = C.c_void_p
            L.sph_host_free.argtypes = [C.c_void_p]
            L.sph_host_free.restype = None
            L.sph_timer_start.argtypes = [C.c_void_p]
            L.sph_out_of_grid_count.argtypes = [C.c_void_p]
            L.sph_out_of_grid_count.restype = C.c_longlong
            L.sph_empty_chain_ms.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, P(C.c_float)]
            L.sph_timer_stop.argtypes = [C.c_void_p, P(C.c_float)]

    @property
    def backend(self):
        return self.lib.sph_backend().decode()

    def last_error(self):
        return self.lib.sph_last_error().decode()

    def check(self, rc, what):
        if rc != 0:
            raise SphError("%s failed (%d): %s" % (what, rc, self.last_error()))

    def default_params(self, box=None):
        p = SphParams()
        self.check(self.lib.sph_params_default(C.byref(p)), "sph_params_default")
        if box is not None:
            self.check(self.lib.sph_params_set_box(C.byref(p), *[float(b) for b in box]), "sph_params_set_box")
        return p

    def pinned_array(self, shape, dtype=np.float32):
        """numpy view of page-locked host memory (freed with the library; keep the lib alive)."""
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        ptr = self.lib.sph_host_alloc(n)
        if not ptr:
            raise SphError("sph_host_alloc failed: " + self.last_error())
        buf = (C.c_char * n).from_address(ptr)
        return np.frombuffer(buf, dtype=dtype).reshape(shape)

    def create(self, params, position, velocity, elastic=None, membranes=None, pml=None, device=0):
        return SphContext(self, params, position, velocity, elastic, membranes, pml, device)


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class SphContext:
    """Mirror of owOpenCLSolver (reference owOpenCLSolver.h:28-124) over the C-ABI."""

    def __init__(self, lib, params, position, velocity, elastic=None, membranes=None, pml=None, device=0):
        self.L = lib
        pos = np.ascontiguousarray(position, dtype=np.float32).reshape(-1, 4)
        vel = np.ascontiguousarray(velocity, dtype=np.float32).reshape(-1, 4)
        assert pos.shape == vel.shape and pos.shape[0] == params.particle_count
        el = None if elastic is None else np.ascontiguousarray(elastic, dtype=np.float32)
        mem = None if membranes is None else np.ascontiguousarray(membranes, dtype=np.int32)
        pm = None if pml is None else np.ascontiguousarray(pml, dtype=np.int32)
        self.params = params
        self.N = params.particle_count
        self.h = lib.lib.sph_create(C.byref(params), _ptr(pos), _ptr(vel), _ptr(el), _ptr(mem), _ptr(pm), device)
        if not self.h:
            raise SphError("sph_create failed: " + lib.last_error())

    def close(self):
        if self.h:
            self.L.lib.sph_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def stage(self, name_or_id):
        sid = STAGE[name_or_id] if isinstance(name_or_id, str) else int(name_or_id)
        self.L.check(self.L.lib.sph_stage(self.h, sid), "sph_stage(%s)" % name_or_id)

    def step(self, sync=True):
        self.L.check(self.L.lib.sph_step(self.h), "sph_step")
        if sync:
            self.sync()

    def step_staged(self):
        """The 18 reference stages in simulationStep order (owPhysicsFluidSimulator.cpp:88-113)."""
        for s in STAGES[:10]:
            self.stage(s)
        for _ in range(self.params.max_iteration):
            for s in STAGES[10:14]:
                self.stage(s)
        for s in STAGES[14:]:
            self.stage(s)

    def sync(self):
        self.L.check(self.L.lib.sph_sync(self.h), "sph_sync")

    def run_steps(self, steps):
        """K fused steps timed on the device (CUDA events on the context's stream); returns ms."""
        ms = C.c_float(0)
        self.L.check(self.L.lib.sph_run_steps(self.h, int(steps), C.byref(ms)), "sph_run_steps")
        return float(ms.value)

    def stream_timer_start(self):
        self.L.check(self.L.lib.sph_timer_start(self.h), "sph_timer_start")
        return True

    def stream_timer_stop(self, _token=None):
        ms = C.c_float(0)
        self.L.check(self.L.lib.sph_timer_stop(self.h, C.byref(ms)), "sph_timer_stop")
        return float(ms.value)

    def empty_chain_ms(self, nodes, blocks=240, threads=256, replays=200):
        ms = C.c_float(0)
        self.L.check(self.L.lib.sph_empty_chain_ms(self.h, int(nodes), int(blocks), int(threads), int(replays), C.byref(ms)), "sph_empty_chain_ms")
        return float(ms.value)

    def debug_counters(self):
        out = (C.c_uint * 4)()
        self.L.check(self.L.lib.sph_debug_counters(self.h, out), "sph_debug_counters")
        return list(out)

    def out_of_grid_count(self):
        return int(self.L.lib.sph_out_of_grid_count(self.h))

    def profile(self, on):
        self.L.check(self.L.lib.sph_profile_enable(self.h, 1 if on else 0), "sph_profile_enable")

    def profile_report(self):
        import json
        buf = C.create_string_buffer(8192)
        self.L.check(self.L.lib.sph_profile_report(self.h, buf, len(buf)), "sph_profile_report")
        return json.loads(buf.value.decode())

    def launch_count(self):
        return int(self.L.lib.sph_launch_count(self.h))

    def set_muscles(self, signal):
        s = np.ascontiguousarray(signal, dtype=np.float32)
        self.L.check(self.L.lib.sph_set_muscles(self.h, _ptr(s), int(s.size)), "sph_set_muscles")

    def buffer_bytes(self, name):
        return self.L.lib.sph_buffer_bytes(self.h, BUFFERS[name][0])

    def read(self, name, count=None):
        """Read a buffer in the reference layout; `count` = leading elements (default: all)."""
        bid, dt, tail = BUFFERS[name]
        nbytes = self.buffer_bytes(name)
        item = np.dtype(dt).itemsize * int(np.prod(tail, dtype=np.int64)) if tail else np.dtype(dt).itemsize
        n = nbytes // item if count is None else int(count)
        out = np.empty((n,) + tail, dtype=dt)
        self.L.check(self.L.lib.sph_read(self.h, bid, _ptr(out), out.nbytes), "sph_read(%s)" % name)
        return out

    def read_into(self, name, out):
        bid = BUFFERS[name][0]
        self.L.check(self.L.lib.sph_read(self.h, bid, _ptr(out), out.nbytes), "sph_read(%s)" % name)
        return out

    def read_begin(self, name, out):
        """Start a device-to-host copy of position / velocity into `out` (pinned) that overlaps the next step."""
        self.L.check(self.L.lib.sph_read_begin(self.h, BUFFERS[name][0], _ptr(out), out.nbytes), "sph_read_begin(%s)" % name)

    def read_end(self):
        self.L.check(self.L.lib.sph_read_end(self.h), "sph_read_end")

    def write(self, name, arr):
        bid, dt, _ = BUFFERS[name]
        a = np.ascontiguousarray(arr, dtype=dt)
        self.L.check(self.L.lib.sph_write(self.h, bid, _ptr(a), a.nbytes), "sph_write(%s)" % name)


_PKG = os.path.dirname(os.path.abspath(__file__))
PRODUCT_LIB = os.path.join(_PKG, "libsph_b200.so")


def load_product():
    """The CUDA library. Fails loudly if it has not been built — there is no CPU fallback."""
    exp = os.environ.get("SPH_LIB")   # experiment builds (build.build_variant); never set in the product
    if exp:
        return SphLib(exp)
    if not os.path.exists(PRODUCT_LIB):
        raise SphError("CUDA library %s is missing: run `python -c 'import __graft_entry__ as g; g.build()'`" % PRODUCT_LIB)
    return SphLib(PRODUCT_LIB)
