"""pmps_observe_extent_static (include/pmps.h): the extent reduction with its scratch in a file-scope device array, for callers
that may not allocate.  No counterpart in the reference; checked here against the
existing entry point pmps_observe_extent through the C ABI, on two calls in a row with the state moved in between, so that a
result left in the shared array by the first call would show in the second.  The reduction is min / max / count: bit-exact.
Collected last on purpose (the name)."""
import ctypes

import numpy as np
import pytest

from polymps_b200 import cases
from polymps_b200.engine import Engine
from polymps_b200.mesh import PolygonMesh
from polymps_b200.params import ParticleSystem
from polymps_b200.state import HostState

pytestmark = pytest.mark.gpu


def _extent(lib, fn, h, ptype):
    mn = (ctypes.c_double * 3)(); mx = (ctypes.c_double * 3)(); n = ctypes.c_longlong(-1)
    rc = fn(h, int(ptype), mn, mx, ctypes.byref(n))
    assert rc == 0, (rc, lib.pmps_last_error(h))
    return np.array(mn[:]), np.array(mx[:]), int(n.value)


def test_static_scratch_extent_equals_the_allocating_entry_point_twice_in_a_row():
    case = cases.dam3d(16, 8, 10, jitter=0.05, name="noalloc")
    ps = ParticleSystem.from_json(case.options).set_parameters(); ps.allocate_buckets(case.pos)
    st = HostState.from_case(case, ps)
    meshes = [PolygonMesh.from_triangles(case.tris).init_wall_tables(int(ps.dim), ps.weightType, ps.partDist, ps.reL, ps.reS)]
    eng = Engine(ps, meshes); eng.remember_capacity(st.numParticlesZero); eng.upload(st)
    lib = eng.lib
    sig = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_longlong)]
    lib.pmps_observe_extent_static.argtypes = sig; lib.pmps_observe_extent_static.restype = ctypes.c_int
    lib.pmps_observe_extent.argtypes = sig; lib.pmps_observe_extent.restype = ctypes.c_int
    seen = []
    for call in range(2):
        for ptype in (ps.fluid, ps.wall):   # the tank is a polygon wall: no wall particles, the empty reduction is a case too
            got = _extent(lib, lib.pmps_observe_extent_static, eng.h, ptype)
            want = _extent(lib, lib.pmps_observe_extent, eng.h, ptype)
            again = _extent(lib, lib.pmps_observe_extent_static, eng.h, ptype)
            for g in (got, again):
                assert g[2] == want[2], (call, ptype, g[2], want[2])
                assert np.array_equal(g[0], want[0]) and np.array_equal(g[1], want[1]), (call, ptype, g, want)
            # against the host arrays too, so that both being wrong the same way cannot pass
            host = eng.download(st.copy())
            sel = host.particleType == ptype; xyz = host.pos.reshape(-1, 3)[sel]
            assert got[2] == int(sel.sum()) and (got[2] > 0) == (ptype == ps.fluid)
            if got[2]:
                assert np.array_equal(got[0], xyz.min(axis=0)) and np.array_equal(got[1], xyz.max(axis=0))
            seen.append(got)
        if call == 0:   # move the particles: the second round of calls must see the new state, not the array's old contents
            eng.init_step0(); eng.step(3)
    assert not all(np.array_equal(seen[0][k], seen[2][k]) for k in (0, 1)), "the fluid did not move between the two calls"
    eng.close()


def test_static_scratch_rejects_bad_arguments():
    case = cases.dam3d(8, 4, 6, jitter=0.0, name="noalloc_args")
    ps = ParticleSystem.from_json(case.options).set_parameters(); ps.allocate_buckets(case.pos)
    meshes = [PolygonMesh.from_triangles(case.tris).init_wall_tables(int(ps.dim), ps.weightType, ps.partDist, ps.reL, ps.reS)]
    eng = Engine(ps, meshes)
    lib = eng.lib
    lib.pmps_observe_extent_static.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_longlong)]
    lib.pmps_observe_extent_static.restype = ctypes.c_int
    mn = (ctypes.c_double * 3)(); mx = (ctypes.c_double * 3)()
    assert lib.pmps_observe_extent_static(None, 0, mn, mx, None) != 0
    assert lib.pmps_observe_extent_static(eng.h, 0, None, mx, None) != 0
    assert lib.pmps_observe_extent_static(eng.h, 0, mn, mx, None) != 0          # no state uploaded yet
    assert b"no particle state" in lib.pmps_last_error(eng.h)
    eng.close()
