"""Pins the CPU oracle (oracle/mps_oracle.c) to the reference: every stage of every dumped step of the committed
golden fixtures (raw dumps of the UNMODIFIED reference, tests/golden/make_golden.py) is replayed through the oracle
from the reference's own pre-state and must come out BIT-IDENTICAL (shifting in the reference's 1-thread Gauss-Seidel
mode).  Only the PPE solve is compared with a tolerance: Eigen's dot products are SIMD-blocked, the oracle's are not."""
import numpy as np
import pytest

from common import GOLDEN_ORACLE, stage_calls, GOLDEN, STAGE_MAP, golden, stage_sequence, dumped_steps, state_from, compare_states
import pyoracle


def _oracle_for(ps, st, meshes, prev, jacobi=False):
    orc = pyoracle.Oracle(ps, st, meshes, shifting_jacobi=jacobi)
    orc.first[:] = prev["firstParticleInBucket"]; orc.last[:] = prev["lastParticleInBucket"]; orc.next[:] = prev["nextParticleInSameBucket"]
    orc.P.velMax = float(prev["scalars"][0])
    orc.P.timeCurrent = float(prev["scalars"][2])   # (the 3D rheology blends its normal stress over the first second, src/MpsViscosity.cpp:523)
    return orc


@pytest.mark.parametrize("name", GOLDEN)
def test_oracle_stage_by_stage_bit_identical(name):
    order, dumps, meta, ps, meshes = golden(name)
    n_checked = 0
    for step in dumped_steps(order):
        for stage, prev_name, cur_name in stage_sequence(order, step):
            prev, cur = dumps[prev_name], dumps[cur_name]
            st = state_from(prev)
            orc = _oracle_for(ps, st, meshes, prev)
            for call in STAGE_MAP[stage]:
                getattr(orc, call)()
            ppe = stage == "calcPress" and ps.mpsType >= 2
            bad = compare_states(st, cur, rtol=1e-9 if ppe else 0.0, live_only=False)
            if stage == "updateParticlesID":
                for a, b in (("firstParticleInBucket", orc.first), ("lastParticleInBucket", orc.last), ("nextParticleInSameBucket", orc.next)):
                    assert np.array_equal(cur[a], b), a
            if stage == "correctionVelocityPosition":
                assert orc.P.velMax == float(cur["scalars"][0])
            assert st.numParticles == int(cur["numParticles"][0])
            assert not bad, "%s step %d stage %s: %s" % (name, step, stage, bad)
            n_checked += 1
    assert n_checked >= 20


@pytest.mark.parametrize("name", GOLDEN_ORACLE)
def test_oracle_round2_features_bit_identical_to_committed_reference_dumps(name):
    """Periodic domains and the polygon-wall rheology against reference dumps that travel with the repository (the live-reference runs of
    tests/test_oracle_variants.py need /root/reference at build time): every stage of the dumped steps, bit for bit."""
    order, dumps, meta, ps, meshes = golden(name)
    n_checked = 0; wrapped = 0
    for step in dumped_steps(order):
        for stage, prev_name, cur_name in stage_sequence(order, step):
            prev, cur = dumps[prev_name], dumps[cur_name]
            st = state_from(prev)
            orc = _oracle_for(ps, st, meshes, prev)
            for call in stage_calls(stage, ps):
                getattr(orc, call)()
            bad = compare_states(st, cur, rtol=0.0, live_only=False)
            assert not bad, "%s step %d stage %s: %s" % (name, step, stage, bad)
            if stage == "updateParticlesID":   # the aliased border layers included (copyDataBetweenBuckets)
                for a, b in (("firstParticleInBucket", orc.first), ("lastParticleInBucket", orc.last), ("nextParticleInSameBucket", orc.next)):
                    assert np.array_equal(cur[a], b), a
            if ps.domainTypeBC == 1 and stage.startswith("checkParticleOutDomain"):
                a = np.asarray(prev["pos"]).reshape(-1, 3)[:, ps.periodicAxis]; b = np.asarray(cur["pos"]).reshape(-1, 3)[:, ps.periodicAxis]
                wrapped += int((np.abs(a - b) > 0.5 * ps.periodicLen).sum())
            n_checked += 1
    assert n_checked >= 15
    if ps.domainTypeBC == 1:
        assert wrapped >= 3, wrapped
    else:
        assert ps.fluidType == 1 and ps.wallType == 1


@pytest.mark.parametrize("name", GOLDEN)
def test_oracle_step0_matches_reference(name):
    order, dumps, meta, ps, meshes = golden(name)
    st = state_from(dumps["pre_init"])
    orc = pyoracle.Oracle(ps, st, meshes, shifting_jacobi=False)
    # the reference has already run the bucket build, the out-of-domain check and the wall pass before pre_init
    orc.first[:] = dumps["pre_init"]["firstParticleInBucket"]; orc.last[:] = dumps["pre_init"]["lastParticleInBucket"]
    orc.next[:] = dumps["pre_init"]["nextParticleInSameBucket"]
    orc.initial_pnd()
    if ps.wallType == 1:
        orc.mean_wall_pnd()
    orc.mean_pnd(); orc.mean_neigh_fluid_pnd(); orc.update_bc(); orc.calc_press()
    bad = compare_states(st, dumps["init"], rtol=1e-9 if ps.mpsType >= 2 else 0.0, live_only=False)
    assert not bad, bad


@pytest.mark.parametrize("name", ["w3", "w2"])
def test_oracle_whole_step_chain(name):
    """oracle_step (the whole-step driver) from the reference's pre-state reproduces the reference's end-of-step state."""
    order, dumps, meta, ps, meshes = golden(name)
    step = dumped_steps(order)[0]
    pre, end = dumps["s%04d_pre" % step], dumps["s%04d_end" % step]
    st = state_from(pre)
    orc = _oracle_for(ps, st, meshes, pre)
    orc.step(1)
    bad = compare_states(st, end, rtol=0.0, live_only=False)
    assert not bad, bad
    assert orc.P.velMax == float(end["scalars"][0])


def test_jacobi_shifting_differs_only_slightly():
    """The order-independent (Jacobi) definition of the shifting sweep, which the CUDA path implements, stays within
    the displacement scale of one shift of the reference's sequential sweep."""
    order, dumps, meta, ps, meshes = golden("w3j")
    step = dumped_steps(order)[0]
    seq = {s: (p, c) for s, p, c in stage_sequence(order, step)}
    prev = dumps[seq["calcConcAndConcGradient"][0]]
    a = state_from(prev); b = state_from(prev)
    _oracle_for(ps, a, meshes, prev, jacobi=False).shifting()
    _oracle_for(ps, b, meshes, prev, jacobi=True).shifting()
    assert np.array_equal(a.concentration, b.concentration)
    shift = np.abs(a.pos - prev["pos"]).max()
    assert shift > 0
    assert np.abs(a.pos - b.pos).max() < 0.2 * shift
