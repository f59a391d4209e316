"""the reference's .grid particle file ingested straight into the device layout (pmps_load_grid) gives the state
MpsInputOutput::readMpsParticleFile + pmps_upload give (src/MpsInputOutput.cpp:645-857) — and the run that follows is the same run."""
import os
import tempfile

import numpy as np
import pytest

from common import compare_states, natural_scales
from polymps_b200 import cases
from polymps_b200.engine import Engine
from polymps_b200.mesh import PolygonMesh
from polymps_b200.params import ParticleSystem
from polymps_b200.state import HostState

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("kind", ["wc3d", "nn2d"])
def test_grid_file_ingestion_equals_host_reader_plus_upload(kind):
    case = cases.dam3d(12, 6, 8, jitter=0.05, name="grid") if kind == "wc3d" else cases.dam2d_erosion(16, 16, 6, jitter=0.03, name="grid")
    ps = ParticleSystem.from_json(case.options).set_parameters(); ps.allocate_buckets(case.pos)
    meshes = [PolygonMesh.from_triangles(case.tris).init_wall_tables(int(ps.dim), ps.weightType, ps.partDist, ps.reL, ps.reS)] if ps.wallType == 1 else []
    root = tempfile.mkdtemp(prefix="pmps_grid_")
    path = os.path.join(root, "case.grid")
    cases.write_grid(path, case.ptype, case.pos, case.vel, case.press)
    want = HostState.from_case(case, ps)
    a = Engine(ps, meshes); a.remember_capacity(want.numParticlesZero); a.upload(want)
    b = Engine(ps, meshes); n = b.load_grid(path)
    assert n == case.n
    ga = want.copy(); gb = want.copy()
    for f in ("pos", "vel", "press", "RHO", "MEU", "distParticleWall2"):
        getattr(gb, f)[:] = -1.0
    gb.particleType[:] = -7; gb.PTYPE[:] = -7
    a.download(ga); b.download(gb)
    for f in ("particleType", "PTYPE", "pos", "vel", "press", "RHO", "MEU", "distParticleWall2", "pndi", "acc", "particleBC", "numNeigh"):
        assert np.array_equal(getattr(ga, f), getattr(gb, f)), f
    a.init_step0(); a.step(3); b.init_step0(); b.step(3)
    a.download(ga); b.download(gb)
    bad = compare_states(gb, ga, rtol=0.0, skip=())
    assert not bad, bad
    a.close(); b.close()
