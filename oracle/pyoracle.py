"""oracle/pyoracle.py — ctypes front end of the CPU restatement (oracle/_build/libmpsoracle.so).  TEST INFRASTRUCTURE:
only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module."""
from __future__ import annotations

import ctypes
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from polymps_b200.params import PmpsParams, to_c_params  # noqa: E402  (struct layouts are the product's C ABI PODs)
from polymps_b200.state import HostState, PmpsHostArrays  # noqa: E402

LIB_PATH = os.path.join(HERE, "_build", "libmpsoracle.so")


def build(force: bool = False) -> str:
    src = [os.path.join(HERE, f) for f in ("mps_oracle.c", "mps_oracle.h")]
    if force or not os.path.isfile(LIB_PATH) or any(os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in src):
        subprocess.check_call(["make", "-C", HERE, "port"], stdout=subprocess.DEVNULL)
    return LIB_PATH


class OBuckets(ctypes.Structure):
    _fields_ = [("first", ctypes.POINTER(ctypes.c_int32)), ("last", ctypes.POINTER(ctypes.c_int32)),
                ("next", ctypes.POINTER(ctypes.c_int32))]


class OMesh(ctypes.Structure):
    _fields_ = [("nV", ctypes.c_int32), ("nF", ctypes.c_int32), ("nNodes", ctypes.c_int32), ("nTable", ctypes.c_int32),
                ("V", ctypes.POINTER(ctypes.c_double)), ("F", ctypes.POINTER(ctypes.c_int32)), ("N", ctypes.POINTER(ctypes.c_double)),
                ("bbMin", ctypes.POINTER(ctypes.c_double)), ("bbMax", ctypes.POINTER(ctypes.c_double)), ("prim", ctypes.POINTER(ctypes.c_int32)),
                ("xDataPND", ctypes.POINTER(ctypes.c_double)), ("xDataNeigh", ctypes.POINTER(ctypes.c_double)),
                ("Zij", ctypes.POINTER(ctypes.c_double)), ("nNeighWall", ctypes.POINTER(ctypes.c_int32))]


def _dp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


def _ip(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int32))


class Oracle:
    """One simulation on the CPU oracle: params (ParticleSystem), HostState, buckets, meshes."""

    STAGES_PAB = ["vel_divergence", "wall_vel_divergence", "prediction_press_gradient", "calc_volume_fraction", "calc_viscosity", "collisions", "wall_npcd", "correction_matrix", "pnd_neigh_npcd", "pnd_diffusive",
                  "wall_pnd_diffusive", "mean_pnd_wall_dummy_surface", "extrapolate_press", "press_gradient", "wall_viscosity",
                  "shifting", "wall_shifting", "update_wall_dummy_vel", "initial_pnd", "mean_wall_pnd", "mean_pnd",
                  "mean_neigh_fluid_pnd", "update_bc", "calc_normals", "calc_wall_normals", "calc_wall_viscosity_interaction"]
    STAGES_PA = ["add_gravity", "predict", "update_pnd", "calc_press_eos"]

    def __init__(self, ps, state: HostState, meshes=(), shifting_jacobi=True, threads=1):
        self.lib = ctypes.CDLL(build())
        L = self.lib
        L.oracle_correct.restype = ctypes.c_double
        L.oracle_calc_viscosity_interaction.restype = ctypes.c_double
        L.oracle_wall_press_gradient.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_double]
        L.oracle_prediction_wall_press_gradient.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_double]
        L.oracle_set_threads(int(threads))
        self.ps = ps
        self.P = to_c_params(ps, shifting_jacobi)
        self.state = state
        nB = ps.numBucketsXYZ
        self.first = np.full(nB, -1, dtype=np.int32)
        self.last = np.full(nB, -1, dtype=np.int32)
        self.next = np.full(state.numParticlesZero, -1, dtype=np.int32)
        self.B = OBuckets(_ip(self.first), _ip(self.last), _ip(self.next))
        self._keep = []
        self.meshes = (OMesh * max(1, len(meshes)))()
        for k, m in enumerate(meshes):
            V = np.ascontiguousarray(m.NV, dtype=np.float64); F = np.ascontiguousarray(m.NF, dtype=np.int32)
            Nn = np.ascontiguousarray(m.NNormals, dtype=np.float64)
            nn = L.oracle_tree_nodes(len(F))
            bmin = np.zeros(3 * nn); bmax = np.zeros(3 * nn); prim = np.zeros(nn, dtype=np.int32)
            xs = np.ascontiguousarray(m.xDataPND); xn = np.ascontiguousarray(m.xDataNeigh); Z = np.ascontiguousarray(m.Zij)
            nw = np.ascontiguousarray(m.nNeighWall, dtype=np.int32)
            self._keep += [V, F, Nn, bmin, bmax, prim, xs, xn, Z, nw]
            om = self.meshes[k]
            om.nV, om.nF, om.nNodes, om.nTable = len(V), len(F), nn, len(xs)
            om.V, om.F, om.N = _dp(V), _ip(F), _dp(Nn)
            om.bbMin, om.bbMax, om.prim = _dp(bmin), _dp(bmax), _ip(prim)
            om.xDataPND, om.xDataNeigh, om.Zij, om.nNeighWall = _dp(xs), _dp(xn), _dp(Z), _ip(nw)
            L.oracle_build_tree(ctypes.byref(om))
        self.solverIter = 0
        self.solverError = 0.0

    # -- plumbing
    def _A(self):
        self._c = self.state.c_struct()
        return ctypes.byref(self._c)

    def _sync(self):
        self.state.numParticles = self._c.numParticles

    def __getattr__(self, name):
        if name in Oracle.STAGES_PAB:
            def f():
                getattr(self.lib, "oracle_" + name)(ctypes.byref(self.P), self._A(), ctypes.byref(self.B)); self._sync()
            return f
        if name in Oracle.STAGES_PA:
            def f():
                getattr(self.lib, "oracle_" + name)(ctypes.byref(self.P), self._A()); self._sync()
            return f
        raise AttributeError(name)

    def calc_viscosity_interaction(self):
        self.CFLvisc = float(self.lib.oracle_calc_viscosity_interaction(ctypes.byref(self.P), self._A(), ctypes.byref(self.B))); self._sync()

    def update_buckets(self):
        self.lib.oracle_update_buckets(ctypes.byref(self.P), self._A(), ctypes.byref(self.B))

    def check_out_domain(self):
        r = self.lib.oracle_check_out_domain(ctypes.byref(self.P), self._A()); self._sync(); return r

    def wall_closest_point(self):
        self.lib.oracle_wall_closest_point(ctypes.byref(self.P), self._A(), self.meshes); self._sync()

    def prediction_wall_press_gradient(self):
        self.lib.oracle_prediction_wall_press_gradient(ctypes.byref(self.P), self._A(), ctypes.byref(self.B), ctypes.c_double(self.P.velMax)); self._sync()

    def wall_press_gradient(self):
        self.lib.oracle_wall_press_gradient(ctypes.byref(self.P), self._A(), ctypes.byref(self.B), ctypes.c_double(self.P.velMax))

    def correct(self):
        self.P.velMax = self.lib.oracle_correct(ctypes.byref(self.P), self._A()); return self.P.velMax

    def solve_ppe(self):
        err = ctypes.c_double(0.0)
        self.lib.oracle_solve_ppe.restype = ctypes.c_int
        self.solverIter = self.lib.oracle_solve_ppe(ctypes.byref(self.P), self._A(), ctypes.byref(self.B), ctypes.byref(err))
        self.solverError = err.value

    def calc_press(self):
        if self.P.mpsType in (0, 1):
            self.calc_press_eos()
        else:
            if self.P.mpsType == 3:   # IMPLICIT_PND_DIVU, src/MpsPressure.cpp:44-56
                self.vel_divergence()
                if self.P.wallType == 1:
                    self.wall_vel_divergence()
            self.solve_ppe()

    def init_step0(self):
        self.lib.oracle_init_step0(ctypes.byref(self.P), self._A(), ctypes.byref(self.B), self.meshes); self._sync()

    def step(self, n=1):
        it = ctypes.c_int(0); err = ctypes.c_double(0.0)
        for _ in range(n):
            self.lib.oracle_step(ctypes.byref(self.P), self._A(), ctypes.byref(self.B), self.meshes, ctypes.byref(it), ctypes.byref(err))
            self._sync()
        self.solverIter, self.solverError = it.value, err.value

    def neighbors(self, i, kind, cap=512):
        out = np.zeros(cap, dtype=np.int32)
        self.lib.oracle_neighbors.restype = ctypes.c_int
        n = self.lib.oracle_neighbors(ctypes.byref(self.P), self._A(), ctypes.byref(self.B), int(i), int(kind), _ip(out), cap)
        return out[:min(n, cap)].copy()

    def closest_point(self, mesh_id, p):
        p = np.ascontiguousarray(p, dtype=np.float64); c = np.zeros(3); sq = ctypes.c_double(); f = ctypes.c_int()
        self.lib.oracle_closest_point(ctypes.byref(self.meshes[mesh_id]), _dp(p), ctypes.byref(sq), ctypes.byref(f), _dp(c))
        return sq.value, f.value, c
