"""Option variants of the kernel family (weight / gradient / repulsion / PND / free-surface / slip settings, tests/common.py
VARIANTS): the UNMODIFIED reference is run right here on a small case (oracle/_ref/ref_driver, one thread), and every stage of
two steps is replayed through the oracle from the reference's own pre-state: bit-identical.  This is what lets the GPU tests use
the oracle as the checker for these settings (tests/test_gpu_parity.py::test_option_variants_against_oracle).  Needs the reference
tree at build time: skipped where oracle/_ref is absent (the GPU box)."""
import numpy as np
import pytest

from common import stage_calls, VARIANTS, stage_sequence, dumped_steps, state_from, compare_states
from polymps_b200.params import ParticleSystem
from polymps_b200.mesh import PolygonMesh
import pyoracle
import refrun
from make_golden import reference_dumps
from test_oracle_golden import _oracle_for


@pytest.mark.parametrize("name", sorted(VARIANTS))
def test_oracle_matches_live_reference_on_option_variant(name):
    if not refrun.have_ref():
        pytest.skip("oracle/_ref/ref_driver not built (needs /root/reference)")
    case = VARIANTS[name][0]()
    order, dumps, params = reference_dumps(case, (1, 2))
    ps = ParticleSystem.from_json(case.options).set_parameters(); ps.allocate_buckets(case.pos)
    meshes = []
    if ps.wallType == 1:
        meshes = [PolygonMesh.from_triangles(case.tris).init_wall_tables(int(ps.dim), ps.weightType, ps.partDist, ps.reL, ps.reS)]
    n = 0
    for step in dumped_steps(order):
        for stage, prev_name, cur_name in stage_sequence(order, step):
            prev, cur = dumps[prev_name], dumps[cur_name]
            st = state_from(prev)
            orc = _oracle_for(ps, st, meshes, prev)
            for call in stage_calls(stage, ps):
                getattr(orc, call)()
            ppe = stage == "calcPress" and ps.mpsType >= 2   # Eigen's dot products are SIMD-blocked: tolerance on the converged pressure
            bad = compare_states(st, cur, rtol=1e-9 if ppe else 0.0, live_only=False)
            assert not bad, "%s step %d stage %s: %s" % (name, step, stage, bad)
            if stage == "correctionVelocityPosition":
                assert orc.P.velMax == float(cur["scalars"][0])
            n += 1
    assert n >= 20


@pytest.mark.parametrize("name", sorted(n for n in VARIANTS if "per" in n))
def test_periodic_domain_limits_equal_the_references_and_particles_wrap(name):
    """polymps_b200/params.py, periodic branch of calcDomainLimits (src/MpsBucket.cpp:142-289) against the live reference's own
    parameter block; and the variant does what it is for: particles re-enter through the opposite face in the dumped steps."""
    if not refrun.have_ref():
        pytest.skip("oracle/_ref/ref_driver not built (needs /root/reference)")
    case = VARIANTS[name][0]()
    order, dumps, params = reference_dumps(case, (1, 2))
    ps = ParticleSystem.from_json(case.options).set_parameters(); ps.allocate_buckets(case.pos)
    for k in ("numBucketsX", "numBucketsY", "numBucketsZ", "bucketSide", "invBucketSide", "domainMinX", "domainMaxX", "domainMinY", "domainMaxY", "domainMinZ", "domainMaxZ"):
        assert float(params[k]) == float(getattr(ps, k)), (k, params[k], getattr(ps, k))
    ax = "XYZ"[ps.periodicAxis]
    assert float(params["physDomMin" + ax]) == ps.physDomMin and float(params["physDomMax" + ax]) == ps.physDomMax
    assert [float(v) for v in params["periodicLength"]] == [float(v) for v in ps.periodicLength] and ps.periodicLen == ps.periodicLength[ps.periodicAxis]
    wrapped = 0
    for step in dumped_steps(order):
        for stage, prev_name, cur_name in stage_sequence(order, step):
            if stage.startswith("checkParticleOutDomain"):
                a = np.asarray(dumps[prev_name]["pos"]).reshape(-1, 3)[:, ps.periodicAxis]; b = np.asarray(dumps[cur_name]["pos"]).reshape(-1, 3)[:, ps.periodicAxis]
                wrapped += int((np.abs(a - b) > 0.5 * ps.periodicLen).sum())
    assert wrapped >= 3, wrapped
