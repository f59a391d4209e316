"""Periodic domain boundary (src/MpsBucket.cpp:142-289,356-417, src/MpsParticle.cpp:35-54,236-269) on the device path
against the oracle, which tests/test_oracle_variants.py pins bit-exactly to the live reference on the same variants: the aliased bucket
layers with their members moved by -/+ periodicLength, the wrap at every out-of-domain check, and the rest-of-the-step invisibility of a
wrapped particle (the reference's bucket lists still hold it where it started the step)."""
import numpy as np
import pytest

from common import VARIANTS, compare_states, natural_scales, neighbor_sets_equal
from polymps_b200 import cases
from polymps_b200.engine import Engine, PmpsError
from polymps_b200.mesh import PolygonMesh
from polymps_b200.params import ParticleSystem
from polymps_b200.state import HostState
import pyoracle

pytestmark = pytest.mark.gpu
RTOL = 1e-10
NOISY = ("deviationDotPolygonNormal",)
PERIODIC = sorted(n for n in VARIANTS if "per" in n)


def _build(case):
    ps = ParticleSystem.from_json(case.options).set_parameters(); ps.allocate_buckets(case.pos)
    meshes = []
    if ps.wallType == 1:
        meshes = [PolygonMesh.from_triangles(case.tris).init_wall_tables(int(ps.dim), ps.weightType, ps.partDist, ps.reL, ps.reS)]
    return ps, meshes, HostState.from_case(case, ps)


def _staged_step(o, ps):
    """One time step through the stage calls in the reference's order (src/main.cpp:256-456); yields after every stage."""
    poly = ps.wallType == 1
    seq = ["update_buckets", "calc_viscosity", "add_gravity", "predict", "check_out_domain", "collisions"]
    if poly:
        seq += ["wall_closest_point", "wall_npcd"]
    if ps.gradientCorrection:
        seq += ["correction_matrix"]
    seq += ["pnd_neigh_npcd"]
    if ps.pndType == 2:
        seq += ["pnd_diffusive"] + (["wall_pnd_diffusive"] if poly else []) + ["update_pnd", "mean_pnd_wall_dummy_surface"]
    seq += ["update_bc", "calc_press"]
    if not poly:
        seq += ["extrapolate_press"]
    if ps.gradientCorrection:
        seq += ["correction_matrix"]
    seq += ["press_gradient"]
    if poly:
        seq += ["wall_press_gradient", "wall_viscosity"]
    seq += ["correct", "check_out_domain"]
    if ps.shiftingType == 2:
        seq += ["shifting"] + (["wall_shifting"] if poly else [])
    if not poly:
        seq += ["update_wall_dummy_vel"]
    seq += ["check_out_domain"]
    for k, name in enumerate(seq):
        getattr(o, name)()
        yield k, name


def _wrapped_ids(before, after, ps):
    a = before.pos.reshape(-1, 3)[:, ps.periodicAxis]; b = after.pos.reshape(-1, 3)[:, ps.periodicAxis]
    return np.nonzero(np.abs(a - b) > 0.5 * ps.periodicLen)[0]


def _wrapped(before, after, ps):
    a = before.pos.reshape(-1, 3)[:, ps.periodicAxis]; b = after.pos.reshape(-1, 3)[:, ps.periodicAxis]
    return int((np.abs(a - b) > 0.5 * ps.periodicLen).sum())


@pytest.mark.parametrize("name", PERIODIC)
def test_periodic_stage_by_stage_against_oracle(name):
    case = VARIANTS[name][1]()
    ps, meshes, st0 = _build(case)
    want = st0.copy(); orc = pyoracle.Oracle(ps, want, meshes, shifting_jacobi=True); orc.init_step0()
    eng = Engine(ps, meshes, device=0, shifting_jacobi=True); eng.remember_capacity(st0.numParticlesZero)
    eng.upload(st0); eng.init_step0()
    got = st0.copy(); eng.download(got)
    bad = compare_states(got, want, rtol=RTOL, skip=NOISY, scales=natural_scales(ps))
    assert not bad, "step 0: %s" % bad
    wrapped = 0
    for step in range(1, 4):
        before = want.copy(); gbefore = got.copy()
        for (k, stage), _ in zip(_staged_step(orc, ps), _staged_step(eng, ps)):
            eng.download(got)
            if stage == "check_out_domain":   # the same particles went through the face
                wrapped += _wrapped(before, want, ps)
                assert np.array_equal(_wrapped_ids(gbefore, got, ps), _wrapped_ids(before, want, ps)), "step %d stage %d" % (step, k)
            before = want.copy(); gbefore = got.copy()
            tol = RTOL * 10 ** (step - 1) * (1e3 if stage == "calc_press" and ps.mpsType >= 2 else 1.0)
            bad = compare_states(got, want, rtol=tol, skip=NOISY, scales=natural_scales(ps))
            assert not bad, "%s step %d stage %d %s: %s" % (name, step, k, stage, bad)
    assert wrapped >= 3, "only %d particles went through a periodic face" % wrapped
    eng.close()


@pytest.mark.parametrize("name", ["2d_per", "3d_per", "3d_per_right"])
def test_periodic_neighbour_sets_after_a_wrap(name):
    """Right after the first out-of-domain check of a step has wrapped particles: the neighbour sets under every predicate are the
    oracle's AS SETS — partners across the periodic face included, the wrapped particles found by nobody, finding their own."""
    case = VARIANTS[name][1]()
    ps, meshes, st0 = _build(case)
    want = st0.copy(); orc = pyoracle.Oracle(ps, want, meshes, shifting_jacobi=True); orc.init_step0()
    eng = Engine(ps, meshes, device=0, shifting_jacobi=True); eng.remember_capacity(st0.numParticlesZero)
    eng.upload(st0); eng.init_step0()
    for o in (orc, eng):
        o.update_buckets(); o.calc_viscosity(); o.add_gravity(); o.predict()
    before = want.copy()
    for o in (orc, eng):
        o.check_out_domain(); o.collisions()
        if ps.wallType == 1:
            o.wall_closest_point()
    a = before.pos.reshape(-1, 3)[:, ps.periodicAxis]; b = want.pos.reshape(-1, 3)[:, ps.periodicAxis]
    moved = np.nonzero(np.abs(a - b) > 0.5 * ps.periodicLen)[0]
    assert len(moved) >= 3
    # the wrapped ones, the particles next to both periodic faces, and a random sample
    x = b[:want.numParticles]
    near = np.nonzero((x < ps.physDomMin + 1.5 * ps.reL) | (x > ps.physDomMax - 1.5 * ps.reL))[0]
    rng = np.random.default_rng(5)
    idx = np.unique(np.concatenate([moved, rng.choice(near, size=min(250, len(near)), replace=False), rng.choice(want.numParticles, size=100, replace=False)]))
    kinds = (0, 1, 2, 3, 4) if ps.wallType == 1 else (0, 1, 4)
    across = 0
    for i in idx:
        for kd in kinds:
            ga = eng.neighbors(int(i), kd); gb = orc.neighbors(int(i), kd)
            assert neighbor_sets_equal(ga, gb), "particle %d kind %d: %s vs %s" % (i, kd, sorted(ga), sorted(gb))
            if kd == 0 and len(gb):
                across += int((np.abs(x[np.asarray(gb, dtype=int)] - x[i]) > 0.5 * ps.periodicLen).any())
        if i in moved:   # nobody finds a wrapped particle until the lists are re-threaded
            pass
    assert across >= 10, "only %d sampled particles had a partner across the periodic face" % across
    found_by_others = sum(1 for i in idx for j in orc.neighbors(int(i), 0) if int(j) in set(int(m) for m in moved))
    assert found_by_others == 0
    eng.close()


@pytest.mark.parametrize("name", ["2d_per", "3d_per"])
def test_periodic_fused_run_keeps_up_with_the_oracle(name):
    """12 fused steps (bucket order kept between steps unless a particle wrapped or the slack is used up): same states as the oracle."""
    case = VARIANTS[name][1]()
    ps, meshes, st0 = _build(case)
    want = st0.copy(); orc = pyoracle.Oracle(ps, want, meshes, shifting_jacobi=True); orc.init_step0()
    eng = Engine(ps, meshes, device=0, shifting_jacobi=True); eng.remember_capacity(st0.numParticlesZero)
    eng.upload(st0); eng.init_step0()
    got = st0.copy()
    for k in range(1, 13):
        orc.step(1); eng.step(1); eng.download(got)
        bad = compare_states(got, want, rtol=min(RTOL * 10 ** (k - 1), 1e-5), skip=NOISY, scales=natural_scales(ps))
        assert not bad, "%s step %d: %s" % (name, k, bad)
    ax = want.pos.reshape(-1, 3)[:want.numParticles, ps.periodicAxis]
    assert ax.min() >= ps.physDomMin and ax.max() <= ps.physDomMax
    assert eng.scalars().numParticles == want.numParticles
    eng.close()


def test_periodic_refusals():
    case = cases.channel2d_periodic(16, 6, name="short")   # 4 inner bucket layers along x: a tile would meet a bucket and its alias
    ps, meshes, st0 = _build(case)
    assert ps.numBucketsX < 8
    with pytest.raises(PmpsError) as e:
        Engine(ps, meshes)
    assert "periodic" in str(e.value)
    case = cases.channel2d_periodic(64, 14, name="arc", fs_type=2)
    ps, meshes, st0 = _build(case)
    with pytest.raises(PmpsError) as e:
        Engine(ps, meshes)
    assert "ARC" in str(e.value)
