"""ctypes front end of libpmps.so — the host-side mirror of the reference's per-step engine interface.

The reference's seam is the set of engine-class methods `void f(MpsParticleSystem*, MpsParticle*, MpsBucket*)` called
from mainLoopOfSimulation (reference src/main.cpp:230-468).  `Engine` exposes one method per reference method with the
reference's names (snake_case aliases of the C ABI in include/pmps.h): `updateParticlesID`, `calcViscosity`, ...
so that a parity test reads like the reference's main loop.  State is the reference-layout `HostState`; everything is
computed by the CUDA library — there is no CPU path here, a missing library or GPU raises.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

from .params import ParticleSystem, PmpsParams, to_c_params
from .state import HostState, PmpsHostArrays, PmpsScalars

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PMPS_LIB", os.path.join(HERE, "libpmps.so"))   # PMPS_LIB: an experimental build of the same library (A/B timing)

STAGES = ["update_buckets", "calc_volume_fraction", "calc_viscosity_interaction", "calc_wall_viscosity_interaction", "calc_viscosity", "add_gravity", "predict", "check_out_domain", "collisions", "wall_closest_point",
          "wall_npcd", "correction_matrix", "calc_normals", "calc_wall_normals", "pnd_neigh_npcd", "pnd_diffusive", "wall_pnd_diffusive", "update_pnd",
          "mean_pnd_wall_dummy_surface", "update_bc", "calc_press", "extrapolate_press", "press_gradient",
          "wall_press_gradient", "prediction_press_gradient", "prediction_wall_press_gradient", "wall_viscosity", "correct", "shifting", "wall_shifting", "update_wall_dummy_vel",
          "initial_pnd", "mean_wall_pnd", "mean_pnd", "mean_neigh_fluid_pnd", "init_step0", "synchronize"]

# reference method name -> stage call (src/main.cpp:256-456)
REFERENCE_NAMES = {
    "updateParticlesID": "update_buckets", "calcVolumeFraction": "calc_volume_fraction",
    "calcViscosityInteractionVal": "calc_viscosity_interaction", "calcWallSlipViscosityInteractionVal": "calc_wall_viscosity_interaction",
    "calcWallNoSlipViscosityInteractionVal": "calc_wall_viscosity_interaction", "calcViscosity": "calc_viscosity", "addGravity": "add_gravity",
    "predictionVelocityPosition": "predict", "checkParticleOutDomain": "check_out_domain",
    "checkParticleCollisions": "collisions", "closestPointPNDBoundaryAABB": "wall_closest_point", "calcWallNPCD": "wall_npcd",
    "correctionMatrix": "correction_matrix", "calcPndnNeighNPCD": "pnd_neigh_npcd", "calcPndDiffusiveTerm": "pnd_diffusive",
    "calcWallNoSlipPndDiffusiveTerm": "wall_pnd_diffusive", "calcWallSlipPndDiffusiveTerm": "wall_pnd_diffusive",
    "updatePnd": "update_pnd", "meanPndParticlesWallDummySurface": "mean_pnd_wall_dummy_surface",
    "updateParticleBC": "update_bc", "calcNormalParticles": "calc_normals", "calcWallNormalParticles": "calc_wall_normals", "calcPress": "calc_press", "extrapolatePressParticlesWallDummy": "extrapolate_press",
    "predictionPressGradient": "prediction_press_gradient", "predictionWallPressGradient": "prediction_wall_press_gradient",
    "calcPressGradient": "press_gradient", "calcWallPressGradient": "wall_press_gradient",
    "calcWallNoSlipViscosity": "wall_viscosity", "calcWallSlipViscosity": "wall_viscosity",
    "correctionVelocityPosition": "correct", "calcConcAndConcGradient": "shifting",
    "calcWallConcAndConcGradient": "wall_shifting", "updateVelocityParticlesWallDummy": "update_wall_dummy_vel",
    "setInitialPndNumberOfNeighNPCD": "initial_pnd", "meanWallPnd": "mean_wall_pnd", "meanPnd": "mean_pnd",
    "meanNeighFluidPnd": "mean_neigh_fluid_pnd",
}


class PmpsError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libpmps error %d: %s" % (code, msg))
        self.code = code


_lib = None


def load_library() -> ctypes.CDLL:
    """Loads polymps_b200/libpmps.so; raises (never falls back) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise ImportError("%s not built: run `python -c 'import __graft_entry__ as g; g.build()'` or `make -C polymps_b200/csrc`" % LIB_PATH)
    L = ctypes.CDLL(LIB_PATH)
    L.pmps_last_error.restype = ctypes.c_char_p
    L.pmps_last_error.argtypes = [ctypes.c_void_p]
    L.pmps_create.argtypes = [ctypes.POINTER(PmpsParams), ctypes.c_int, ctypes.POINTER(ctypes.c_void_p)]
    L.pmps_stream.restype = ctypes.c_void_p
    L.pmps_stream.argtypes = [ctypes.c_void_p]
    for s in STAGES + ["destroy"]:
        getattr(L, "pmps_" + s).argtypes = [ctypes.c_void_p]
    L.pmps_step.argtypes = [ctypes.c_void_p, ctypes.c_int]
    L.pmps_upload.argtypes = [ctypes.c_void_p, ctypes.POINTER(PmpsHostArrays)]
    L.pmps_step_host.argtypes = [ctypes.c_void_p, ctypes.POINTER(PmpsHostArrays), ctypes.POINTER(PmpsHostArrays)]
    L.pmps_upload_device.argtypes = [ctypes.c_void_p, ctypes.POINTER(PmpsHostArrays)]
    L.pmps_download.argtypes = [ctypes.c_void_p, ctypes.POINTER(PmpsHostArrays)]
    L.pmps_get_scalars.argtypes = [ctypes.c_void_p, ctypes.POINTER(PmpsScalars)]
    L.pmps_set_velmax.argtypes = [ctypes.c_void_p, ctypes.c_double]
    L.pmps_set_uniform_vel_wall.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_double)]
    L.pmps_debug_neighbors.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int32), ctypes.c_int]
    L.pmps_debug_buckets.argtypes = [ctypes.c_void_p] + [ctypes.POINTER(ctypes.c_int32)] * 3
    L.pmps_profile_enable.argtypes = [ctypes.c_void_p, ctypes.c_int]
    L.pmps_profile_get.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.POINTER(ctypes.c_char_p)),
                                   ctypes.POINTER(ctypes.POINTER(ctypes.c_float)), ctypes.POINTER(ctypes.POINTER(ctypes.c_int))]
    L.pmps_launch_count.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_longlong)]
    L.pmps_get_counters.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_longlong)]
    L.pmps_register_host.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t]
    L.pmps_unregister_host.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    if os.environ.get("PMPS_LIB") is None or hasattr(L, "pmps_set_stale_list_replay"):   # (an older experimental build named by PMPS_LIB may lack them)
        L.pmps_set_stale_list_replay.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.pmps_debug_set_stale.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32)]
    if os.environ.get("PMPS_LIB") is None or hasattr(L, "pmps_observe_extent"):
        L.pmps_observe_extent.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_longlong)]
        L.pmps_observe_box.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_longlong), ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]
    if os.environ.get("PMPS_LIB") is None or hasattr(L, "pmps_write_vtu"):
        L.pmps_write_vtu.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_uint]
        L.pmps_debug_format_f6.argtypes = [ctypes.POINTER(ctypes.c_double), ctypes.c_int, ctypes.c_char_p, ctypes.c_size_t]
    if os.environ.get("PMPS_LIB") is None or hasattr(L, "pmps_load_grid"):
        L.pmps_load_grid.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.POINTER(ctypes.c_int)]
    L.pmps_set_mesh.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_double), ctypes.c_int,
                                ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_double), ctypes.c_int,
                                ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double),
                                ctypes.POINTER(ctypes.c_int32)]
    _lib = L
    return L


def _dp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


def _ip(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int32))


class Engine:
    """One simulation on one GPU.  `ps` must have had set_parameters() and allocate_buckets() applied (the reference's
    setParameters / MpsBucket::allocateMemory); `meshes` are host PolygonMesh objects with their wall tables."""

    def __init__(self, ps: ParticleSystem, meshes=(), device: int = 0, shifting_jacobi: bool = True):
        self.lib = load_library()
        self.ps = ps
        self.P = to_c_params(ps, shifting_jacobi)
        self.h = ctypes.c_void_p()
        rc = self.lib.pmps_create(ctypes.byref(self.P), int(device), ctypes.byref(self.h))
        if rc != 0:
            raise PmpsError(rc, (self.lib.pmps_last_error(None) or b"").decode())
        self._keep = []
        self._pinned = {}   # address -> the numpy array (kept alive while registered: the contract of pmps_register_host)
        self._n0_value = 0
        for k, m in enumerate(meshes):
            self.set_mesh(k, m)

    # ---- plumbing
    def _chk(self, rc):
        if rc < 0:
            raise PmpsError(rc, (self.lib.pmps_last_error(self.h) or b"").decode())
        return rc

    def close(self):
        if self.h:
            self.lib.pmps_destroy(self.h)   # unregisters what register_host pinned
            self.h = ctypes.c_void_p()
            self._pinned = {}

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __getattr__(self, name):
        stage = REFERENCE_NAMES.get(name, name)
        if stage in STAGES:
            fn = getattr(self.lib, "pmps_" + stage)
            return lambda: self._chk(fn(self.h))
        raise AttributeError(name)

    def set_mesh(self, mesh_id, m):
        V = np.ascontiguousarray(m.NV, dtype=np.float64); F = np.ascontiguousarray(m.NF, dtype=np.int32)
        N = np.ascontiguousarray(m.NNormals, dtype=np.float64)
        xs = np.ascontiguousarray(m.xDataPND, dtype=np.float64); xn = np.ascontiguousarray(m.xDataNeigh, dtype=np.float64)
        Z = np.ascontiguousarray(m.Zij, dtype=np.float64); nw = np.ascontiguousarray(m.nNeighWall, dtype=np.int32)
        self._chk(self.lib.pmps_set_mesh(self.h, mesh_id, len(V), _dp(V), len(F), _ip(F), _dp(N), len(xs), _dp(xs), _dp(xn), _dp(Z), _ip(nw)))

    # ---- transfers
    def upload(self, state: HostState, fields=None):
        c = state.c_struct(fields)
        self._n0_value = state.numParticlesZero
        self._chk(self.lib.pmps_upload(self.h, ctypes.byref(c)))

    def load_grid(self, path: str) -> int:
        """The reference's .grid particle file straight into the device layout (pmps_load_grid); returns the particle count."""
        n = ctypes.c_int()
        self._chk(self.lib.pmps_load_grid(self.h, os.fsencode(path), ctypes.byref(n)))
        self._n0_value = int(n.value)
        return int(n.value)

    def download(self, state: HostState, fields=None):
        c = state.c_struct(fields)
        self._chk(self.lib.pmps_download(self.h, ctypes.byref(c)))
        state.numParticles = c.numParticles
        return state

    def register_host(self, state: HostState, fields=None):
        """Page-lock the named arrays of `state` for asynchronous copies (pmps_register_host).  The engine keeps a reference to
        every registered array until close(), which is the lifetime the C contract asks for."""
        from .state import ALL_FIELDS
        for f in (fields or ALL_FIELDS):
            a = getattr(state, f)
            if a.nbytes < (1 << 16) or a.ctypes.data in self._pinned:
                continue
            self._chk(self.lib.pmps_register_host(self.h, ctypes.c_void_p(a.ctypes.data), a.nbytes))
            self._pinned[a.ctypes.data] = a

    def unregister_host(self, state: HostState, fields=None):
        from .state import ALL_FIELDS
        for f in (fields or ALL_FIELDS):
            a = getattr(state, f)
            if a.ctypes.data in self._pinned:
                self._chk(self.lib.pmps_unregister_host(self.h, ctypes.c_void_p(a.ctypes.data)))
                del self._pinned[a.ctypes.data]

    def upload_device(self, c: PmpsHostArrays):
        self._chk(self.lib.pmps_upload_device(self.h, ctypes.byref(c)))

    # ---- stepping
    def step(self, n: int = 1):
        self._chk(self.lib.pmps_step(self.h, int(n)))

    def step_host(self, state, up_fields=None, down_fields=None):
        """One host-driven step (pmps_step_host): upload `up_fields` of `state`, step, download `down_fields` into it, copies overlapped."""
        a = state.c_struct(up_fields); b = state.c_struct(down_fields)
        self._chk(self.lib.pmps_step_host(self.h, ctypes.byref(a), ctypes.byref(b)))
        state.numParticles = b.numParticles

    def scalars(self, check: bool = True) -> PmpsScalars:
        s = PmpsScalars()
        rc = self.lib.pmps_get_scalars(self.h, ctypes.byref(s))
        if check:
            self._chk(rc)
        return s

    def set_time(self, t: float):
        self._chk(self.lib.pmps_set_time(self.h, ctypes.c_double(float(t))))

    def set_velmax(self, v: float):
        self._chk(self.lib.pmps_set_velmax(self.h, float(v)))

    def set_uniform_vel_wall(self, v):
        a = (ctypes.c_double * 3)(*[float(x) for x in v])
        self._chk(self.lib.pmps_set_uniform_vel_wall(self.h, a))

    @property
    def stream(self) -> int:
        return int(self.lib.pmps_stream(self.h) or 0)

    def set_stale_list_replay(self, on: bool):
        self._chk(self.lib.pmps_set_stale_list_replay(self.h, 1 if on else 0))

    # ---- integrated observables, reduced on the device (pmps_observe_*)
    def observe_extent(self, particle_type: int):
        """(min xyz, max xyz, count) of the particles of one type; surge front of a dam break = max x of the fluid."""
        mn = (ctypes.c_double * 3)(); mx = (ctypes.c_double * 3)(); n = ctypes.c_longlong()
        self._chk(self.lib.pmps_observe_extent(self.h, int(particle_type), mn, mx, ctypes.byref(n)))
        return np.array(mn[:]), np.array(mx[:]), int(n.value)

    def observe_box(self, particle_type: int, lo, hi):
        """(count, pressure sum, pressure max) of the particles of one type inside the box [lo, hi]; probe pressure = sum / count."""
        a = (ctypes.c_double * 3)(*[float(v) for v in lo]); b = (ctypes.c_double * 3)(*[float(v) for v in hi])
        n = ctypes.c_longlong(); s = ctypes.c_double(); m = ctypes.c_double()
        self._chk(self.lib.pmps_observe_box(self.h, int(particle_type), a, b, ctypes.byref(n), ctypes.byref(s), ctypes.byref(m)))
        return int(n.value), float(s.value), float(m.value)

    # ---- output (pmps_write_vtu): the reference's ASCII .vtu file from the device state
    VTU_PND, VTU_NEIGH, VTU_DEVIATION, VTU_CONCENTRATION, VTU_NONNEWTONIAN, VTU_AUXILIAR, VTU_FREE_SURFACE_ONLY = 1, 2, 4, 8, 16, 32, 64

    def write_vtu(self, path: str, flags: int = 1):
        """MpsInputOutput::writeVtuAscii / writeVtuAsciiFreeSurface (src/MpsInputOutput.cpp:1313-1875) for the state on the device."""
        self._chk(self.lib.pmps_write_vtu(self.h, os.fsencode(path), int(flags)))

    @staticmethod
    def vtu_flags(options: dict) -> int:
        """flags.output_VTU of an input document -> pmps_write_vtu flags (src/MpsInputOutput.cpp:150-170)."""
        o = options["flags"]["output_VTU"]
        f = 0
        for key, bit in (("pnd", 1), ("neigh", 2), ("deviation", 4), ("concentration", 8), ("non_newtonian", 16), ("auxiliar", 32), ("freeSurfWall", 64)):
            if o.get(key, False):
                f |= bit
        return f

    # ---- parity hooks
    def set_stale(self, host_idx, bucket):
        """Tests: array elements `host_idx` are listed in reference buckets `bucket` until the next step start (pmps_debug_set_stale)."""
        a = np.ascontiguousarray(host_idx, dtype=np.int32); b = np.ascontiguousarray(bucket, dtype=np.int32)
        self._chk(self.lib.pmps_debug_set_stale(self.h, len(a), _ip(a), _ip(b)))

    def neighbors(self, i: int, kind: int, cap: int = 1024) -> np.ndarray:
        out = np.zeros(cap, dtype=np.int32)
        n = self._chk(self.lib.pmps_debug_neighbors(self.h, int(i), int(kind), _ip(out), cap))
        return out[:min(n, cap)].copy()

    def buckets(self):
        n0 = self._n0_value
        nB = self.ps.numBucketsXYZ
        b = np.zeros(n0, dtype=np.int32); st = np.zeros(nB + 1, dtype=np.int32); mem = np.zeros(n0, dtype=np.int32)
        self._chk(self.lib.pmps_debug_buckets(self.h, _ip(b), _ip(st), _ip(mem)))
        return b, st, mem

    def remember_capacity(self, n0):
        self._n0_value = int(n0)

    # ---- profiling
    def profile(self, on: bool):
        self._chk(self.lib.pmps_profile_enable(self.h, 1 if on else 0))

    def profile_get(self):
        n = ctypes.c_int(); names = ctypes.POINTER(ctypes.c_char_p)(); ms = ctypes.POINTER(ctypes.c_float)(); ln = ctypes.POINTER(ctypes.c_int)()
        self._chk(self.lib.pmps_profile_get(self.h, ctypes.byref(n), ctypes.byref(names), ctypes.byref(ms), ctypes.byref(ln)))
        return [(names[k].decode(), float(ms[k]), int(ln[k])) for k in range(n.value)]

    def counters(self) -> dict:
        """Work counters of the temporal caches: bucket-order rebuilds, hit-list refilters, fused steps, guard violation flag."""
        v = (ctypes.c_longlong * 4)()
        self._chk(self.lib.pmps_get_counters(self.h, v))
        return {"bucket_rebuilds": int(v[0]), "hit_refilters": int(v[1]), "steps": int(v[2]), "guard_violated": int(v[3])}

    def launch_count(self) -> int:
        v = ctypes.c_longlong()
        self._chk(self.lib.pmps_launch_count(self.h, ctypes.byref(v)))
        return int(v.value)


def make_engine(case, device: int = 0, shifting_jacobi: bool = True):
    """Case (polymps_b200.cases.Case) -> (ParticleSystem, meshes, HostState, Engine) with the state uploaded."""
    from .mesh import PolygonMesh
    ps = ParticleSystem.from_json(case.options).set_parameters()
    ps.allocate_buckets(case.pos)
    meshes = []
    if ps.wallType == 1:
        meshes = [PolygonMesh.from_triangles(case.tris).init_wall_tables(int(ps.dim), ps.weightType, ps.partDist, ps.reL, ps.reS)]
    st = HostState.from_case(case, ps)
    eng = Engine(ps, meshes, device=device, shifting_jacobi=shifting_jacobi)
    eng.remember_capacity(st.numParticlesZero)
    eng.upload(st)
    return ps, meshes, st, eng
