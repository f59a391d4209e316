"""Non-Newtonian rheology with POLYGON walls (the remainder round 1 refused): MpsViscosity::calcWallSlipViscosityInteractionVal /
calcWallNoSlipViscosityInteractionVal (src/MpsViscosity.cpp:703-1526; option set of input/damErosion3D_WC.json) as the stage call
pmps_calc_wall_viscosity_interaction, against the oracle that tests/test_oracle_variants.py pins bit-exactly to the live reference on the variants
nn3d_poly (no-slip) and nn3d_poly_slip.  Whole steps of the same variants run in test_gpu_parity.py::test_option_variants_against_oracle."""
import numpy as np
import pytest

from common import VARIANTS, compare_states, natural_scales
from polymps_b200.engine import Engine
from polymps_b200.mesh import PolygonMesh
from polymps_b200.params import ParticleSystem
from polymps_b200.state import HostState
import pyoracle

pytestmark = pytest.mark.gpu
RHEO = ("Cv", "II", "MEU", "MEU_Y", "Inertia", "p_rheo_new", "p_smooth", "VF", "RHO", "S11", "S12", "S13", "S22", "S23", "S33")


@pytest.mark.parametrize("name", ["nn3d_poly", "nn3d_poly_slip"])
def test_wall_rheology_stage_against_oracle(name):
    case = VARIANTS[name][1]()
    ps = ParticleSystem.from_json(case.options).set_parameters(); ps.allocate_buckets(case.pos)
    meshes = [PolygonMesh.from_triangles(case.tris).init_wall_tables(int(ps.dim), ps.weightType, ps.partDist, ps.reL, ps.reS)]
    st0 = HostState.from_case(case, ps)
    want = st0.copy(); orc = pyoracle.Oracle(ps, want, meshes, shifting_jacobi=True); orc.init_step0(); orc.step(2)
    # the third step up to the second rheology site (src/main.cpp:256-386)
    orc.update_buckets(); orc.calc_volume_fraction(); orc.calc_viscosity_interaction(); orc.calc_viscosity(); orc.add_gravity(); orc.predict(); orc.check_out_domain()
    orc.collisions(); orc.wall_closest_point(); orc.wall_npcd(); orc.pnd_neigh_npcd(); orc.pnd_diffusive(); orc.wall_pnd_diffusive(); orc.update_pnd()
    orc.mean_pnd_wall_dummy_surface(); orc.update_bc(); orc.calc_press(); orc.press_gradient(); orc.wall_press_gradient()
    pre = want.copy()
    near = (pre.particleNearWall[:pre.numParticles] != 0) & (pre.particleType[:pre.numParticles] == ps.fluid)
    assert near.sum() > 100 and ((pre.PTYPE[:pre.numParticles] == 2) & near).sum() > 50   # bed particles at the walls: the model branch is exercised
    eng = Engine(ps, meshes, device=0, shifting_jacobi=True); eng.remember_capacity(st0.numParticlesZero)
    eng.upload(pre); eng.set_velmax(orc.P.velMax); eng.set_time(orc.P.timeCurrent); eng.update_buckets()
    got = pre.copy()
    for stage in ("calc_volume_fraction", "calc_viscosity_interaction", "calc_wall_viscosity_interaction"):
        getattr(orc, stage)(); getattr(eng, stage)()
        eng.download(got)
        bad = compare_states(got, want, rtol=1e-10, fields=RHEO, scales=natural_scales(ps))
        assert not bad, "%s %s: %s" % (name, stage, bad)
    # the wall pass changed something: the near-wall particles' strain-rate invariant and viscosity are not those of the base pass
    chk = want.copy(); o2 = pyoracle.Oracle(ps, chk, meshes, shifting_jacobi=True)
    chk_pre = pre.copy()
    for f in RHEO + ("pos", "vel", "press", "mirrorParticlePos", "particleNearWall", "PTYPE", "particleType"):
        getattr(chk, f)[:] = getattr(chk_pre, f)
    o2.P.timeCurrent = orc.P.timeCurrent; o2.update_buckets(); o2.calc_volume_fraction(); o2.calc_viscosity_interaction()
    n = want.numParticles
    assert np.abs(chk.MEU[:n] - want.MEU[:n])[near].max() > 1e-6 * np.abs(want.MEU[:n]).max()
    assert np.array_equal(chk.MEU[:n][~near], want.MEU[:n][~near])
    eng.close()
