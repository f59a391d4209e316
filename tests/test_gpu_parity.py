"""GPU parity tests (run with -m gpu on the B200 box).  Everything goes through the C ABI (libpmps.so via
polymps_b200.engine); the checker is (a) the committed golden dumps of the unmodified reference and (b) the CPU oracle
on seeded inputs.  Tolerances: integer / index / flag fields and neighbour SETS bit-exact; floating-point fields
1e-10 relative to the field's max magnitude (BASELINE.json north_star)."""
import os

import numpy as np
import pytest

from common import (GOLDEN, STAGE_MAP, VARIANTS, SHIFTING_STAGES, golden, stage_sequence, dumped_steps, state_from, compare_states,
                    neighbor_sets_equal, natural_scales)
from polymps_b200 import cases
from polymps_b200.params import ParticleSystem
from polymps_b200.mesh import PolygonMesh
from polymps_b200.state import HostState

pytestmark = pytest.mark.gpu
RTOL = 1e-10
# sign of a dot product that is pure rounding noise for symmetric neighbourhoods: compared separately
NOISY = ("deviationDotPolygonNormal",)


def _engine(ps, meshes, n0):
    from polymps_b200.engine import Engine
    e = Engine(ps, meshes, device=0, shifting_jacobi=True)
    e.remember_capacity(n0)
    return e


def _oracle(ps, st, meshes, prev=None, jacobi=True):
    import pyoracle
    orc = pyoracle.Oracle(ps, st, meshes, shifting_jacobi=jacobi)
    if prev is not None:
        orc.first[:] = prev["firstParticleInBucket"]; orc.last[:] = prev["lastParticleInBucket"]; orc.next[:] = prev["nextParticleInSameBucket"]
        orc.P.velMax = float(prev["scalars"][0])
    return orc


def _check_noisy_sign(got, want_state):
    """deviationDotPolygonNormal is sign(npcd . n): must agree wherever the dot product is not rounding noise."""
    n = got.numParticles
    d = want_state.npcdDeviation.reshape(-1, 3)[:n]; pn = want_state.polygonNormal.reshape(-1, 3)[:n]
    dot = (d * pn).sum(axis=1)
    sure = np.abs(dot) > 1e-12 * max(np.abs(d).max(), 1e-300)
    assert np.array_equal(got.deviationDotPolygonNormal[:n][sure], want_state.deviationDotPolygonNormal[:n][sure])


def _stale_elements(prev, ps):
    """Array elements the reference's bucket lists hold somewhere their content no longer is: after checkParticleOutDomain copied
    the last live particle over a hole (src/MpsParticle.cpp:298-396) the lists still show the hole's old bucket until the next
    updateParticlesID.  -> (host indices, listed buckets) of the elements listed in another bucket than the one their position is in:
    a neighbour loop finds such an element only when its own 27 buckets include the LISTED one (hid_visible, pmps_device.cuh)."""
    first = np.asarray(prev["firstParticleInBucket"]); nxt = np.asarray(prev["nextParticleInSameBucket"])
    n = int(prev["numParticles"][0]); pos = np.asarray(prev["pos"]).reshape(-1, 3); typ = np.asarray(prev["particleType"])
    nX, nY = ps.numBucketsX, ps.numBucketsY
    idx, cells = [], []
    for b in np.nonzero(first >= 0)[0]:
        j = int(first[b])
        while j != -1:
            if j < n and typ[j] != ps.ghost:
                c = np.floor((pos[j] - np.array([ps.domainMinX, ps.domainMinY, ps.domainMinZ])) * ps.invBucketSide + (ps.epsilonZero if ps.dim == 2 else 0.0)).astype(int)
                lb = np.array([b % nX, (b // nX) % nY, b // (nX * nY)])
                if ps.dim == 2:
                    c[2] = 0
                if np.any(c != lb):   # (also true for a particle that crossed a bucket face since the step start: the rule below is the reference's for it too)
                    idx.append(j); cells.append(int(b))
            j = int(nxt[j])
    return idx, cells


@pytest.mark.parametrize("name", GOLDEN)
def test_stage_by_stage_against_reference_dumps(name):
    """Every stage call from the reference's state before it against the reference's state after it.  A stage starts from an upload
    + pmps_update_buckets of the dumped state, i.e. from bucket lists that are up to date — except in a step in which a particle left
    the domain (golden o3): there the reference's lists are stale for the element that received the moved particle, and the test
    hands the device the same knowledge (pmps_debug_set_stale, derived from the dumped lists).  The product path's own bookkeeping
    of that (k_ood_fix) is checked by test_chained_stages_with_particles_leaving and the whole-step variant `ood`."""
    order, dumps, meta, ps, meshes = golden(name)
    n0 = int(dumps["pre_init"]["numParticles"][1])
    eng = _engine(ps, meshes, n0)
    checked = 0
    for step in dumped_steps(order):
        for stage, prev_name, cur_name in stage_sequence(order, step):
            prev, cur = dumps[prev_name], dumps[cur_name]
            st = state_from(prev)
            eng.upload(st)
            eng.set_velmax(float(prev["scalars"][0]))
            eng.update_buckets()
            if stage != "updateParticlesID":
                if name == "o3":
                    eng.set_stale(*_stale_elements(prev, ps))
                for call in STAGE_MAP[stage]:
                    getattr(eng, call)()
            got = state_from(prev)
            eng.download(got)
            if stage in SHIFTING_STAGES:   # the reference's sweep is sequential: the checker is the Jacobi oracle
                want = state_from(prev); orc = _oracle(ps, want, meshes, prev, jacobi=True)
                for call in STAGE_MAP[stage]:
                    getattr(orc, call)()
            else:
                want = state_from(cur)
            assert got.numParticles == want.numParticles
            bad = compare_states(got, want, rtol=RTOL, skip=NOISY, scales=natural_scales(ps))
            assert not bad, "%s step %d stage %s: %s" % (name, step, stage, bad)
            _check_noisy_sign(got, want)
            if stage == "correctionVelocityPosition":
                assert abs(eng.scalars().velMax - float(cur["scalars"][0])) <= 1e-14 * max(1.0, float(cur["scalars"][0]))
            checked += 1
    assert checked >= 20
    eng.close()


def test_chained_stages_with_particles_leaving():
    """golden o3, steps in which particles leave the domain, WITHOUT re-uploading between the stages: the device
    carries its own record of which array element the reference's stale bucket lists hold where (k_ood_fix -> hid), so the stages
    after a removal — PND, pressure, gradient, correction — see the neighbour sets the reference saw.  Chained up to the shifting
    (whose reference sweep is sequential); numParticles, every integer field and the array order exact."""
    order, dumps, meta, ps, meshes = golden("o3")
    n0 = int(dumps["pre_init"]["numParticles"][1])
    eng = _engine(ps, meshes, n0)
    removed = 0
    for step in dumped_steps(order):
        seq = stage_sequence(order, step)
        pre = dumps[seq[0][1]]
        eng.upload(state_from(pre)); eng.set_velmax(float(pre["scalars"][0]))
        k = 0
        for stage, prev_name, cur_name in seq:
            if stage in SHIFTING_STAGES:
                break
            for call in (["update_buckets"] if stage == "updateParticlesID" else STAGE_MAP[stage]):
                getattr(eng, call)()
            cur = dumps[cur_name]
            got = state_from(pre); eng.download(got); want = state_from(cur)
            assert got.numParticles == want.numParticles, (step, stage)
            k += 1
            bad = compare_states(got, want, rtol=RTOL * 10, skip=NOISY, scales=natural_scales(ps))
            assert not bad, "o3 step %d after %d chained stages (%s): %s" % (step, k, stage, bad)
            for f in ("particleType", "particleBC", "numNeigh", "nearMeshType"):
                assert np.array_equal(getattr(got, f)[:n0], getattr(want, f)[:n0]), (step, stage, f)
        removed += int(pre["numParticles"][0]) - got.numParticles
    assert removed > 0, "no particle left the domain in the dumped steps"
    eng.close()


@pytest.mark.parametrize("name", GOLDEN)
def test_bucket_membership_matches_reference_lists(name):
    """Bucket assignment and in-bucket order are bit-exact: same bucket id per particle, same members per bucket in the
    reference's list order (ascending index)."""
    order, dumps, meta, ps, meshes = golden(name)
    step = dumped_steps(order)[-1]
    pre = dumps["s%04d_pre" % step]; cur = dumps[stage_sequence(order, step)[0][2]]
    n0 = int(pre["numParticles"][1])
    eng = _engine(ps, meshes, n0)
    eng.upload(state_from(pre)); eng.update_buckets()
    b, start, members = eng.buckets()
    first, nxt = cur["firstParticleInBucket"], cur["nextParticleInSameBucket"]
    for cell in range(ps.numBucketsXYZ):
        ref = []; j = int(first[cell])
        while j != -1:
            ref.append(j); j = int(nxt[j])
        assert list(members[start[cell]:start[cell + 1]]) == ref, "bucket %d" % cell
        for j in ref:
            assert b[j] == cell
    eng.close()


def _seeded_case(kind):
    if kind == "w3":
        return cases.dam3d(20, 10, 14, jitter=0.05, name="w3s")
    if kind == "w3lattice":
        return cases.dam3d(16, 8, 12, name="w3l")
    if kind == "w2":
        return cases.dam2d_particle(30, 40, name="w2s")
    if kind == "i3":
        return cases.dam3d(16, 8, 12, incompressible=True, jitter=0.03, name="i3s")
    if kind == "i2":
        return cases.dam2d_particle(30, 40, incompressible=True, name="i2s")
    if kind == "e3":   # explicit MPS equation of state
        return cases.dam3d(16, 8, 12, jitter=0.04, name="e3s", mps_type=0)
    if kind == "b3":   # incompressible with the BiCGSTAB solver
        return cases.dam3d(16, 8, 12, incompressible=True, jitter=0.03, name="b3s", solver=1)
    if kind == "n2":   # non-Newtonian erosional dam break (BASELINE configs[0] option set)
        return cases.dam2d_erosion(24, 24, 8, jitter=0.03, name="n2s")
    raise KeyError(kind)


def _build(case):
    ps = ParticleSystem.from_json(case.options).set_parameters()
    ps.allocate_buckets(case.pos)
    meshes = []
    if ps.wallType == 1:
        meshes = [PolygonMesh.from_triangles(case.tris).init_wall_tables(int(ps.dim), ps.weightType, ps.partDist, ps.reL, ps.reS)]
    return ps, meshes, HostState.from_case(case, ps)


@pytest.mark.parametrize("kind", ["w3", "w3lattice", "w2", "i3", "i2", "n2", "b3", "e3"])
def test_whole_steps_against_oracle(kind):
    """pmps_init_step0 + pmps_step against the oracle's drivers on a seeded case a few thousand particles large."""
    case = _seeded_case(kind)
    ps, meshes, st0 = _build(case)
    want = st0.copy(); orc = _oracle(ps, want, meshes, jacobi=True)
    orc.init_step0()
    eng = _engine(ps, meshes, st0.numParticlesZero)
    eng.upload(st0); eng.init_step0()
    got = st0.copy(); eng.download(got)
    bad = compare_states(got, want, rtol=RTOL, skip=NOISY, scales=natural_scales(ps))
    assert not bad, "step 0: %s" % bad
    for k in range(1, 4):
        orc.step(1); eng.step(1)
        eng.download(got)
        bad = compare_states(got, want, rtol=RTOL * 10 ** (k - 1), skip=NOISY, scales=natural_scales(ps))
        assert not bad, "step %d: %s" % (k, bad)
        s = eng.scalars()
        assert abs(s.velMax - orc.P.velMax) <= 1e-9 * max(orc.P.velMax, 1e-30)
        assert s.numParticles == want.numParticles
        if ps.mpsType >= 2:   # Krylov solve: same stopping rule, iteration counts may differ by rounding only
            assert abs(s.solverIter - orc.solverIter) <= 3 and s.solverError < 4e-16
    eng.close()


@pytest.mark.parametrize("kind", ["w3", "w2"])
def test_neighbour_sets_bit_exact(kind):
    """Neighbour sets under every predicate of the reference (P(reL), P(reS), M(reS), M(reL), collision) are identical
    AS SETS to the oracle's bucket walk, after positions have moved away from the bucket build (mid-step)."""
    case = _seeded_case(kind)
    ps, meshes, st0 = _build(case)
    want = st0.copy(); orc = _oracle(ps, want, meshes, jacobi=True)
    orc.init_step0(); orc.step(2)
    # go to the middle of step 3: after prediction + wall pass the mirror positions are live
    orc.update_buckets(); orc.calc_viscosity(); orc.add_gravity(); orc.predict(); orc.check_out_domain(); orc.collisions()
    eng = _engine(ps, meshes, st0.numParticlesZero)
    pre = want.copy()
    if ps.wallType == 1:
        orc.wall_closest_point()
    # device: same bucket build (start-of-step positions), then the same moves
    eng.upload(st0); eng.init_step0(); eng.step(2)
    eng.update_buckets(); eng.calc_viscosity(); eng.add_gravity(); eng.predict(); eng.check_out_domain(); eng.collisions()
    if ps.wallType == 1:
        eng.wall_closest_point()
    rng = np.random.default_rng(7)
    idx = rng.choice(want.numParticles, size=min(400, want.numParticles), replace=False)
    kinds = (0, 1, 2, 3, 4) if ps.wallType == 1 else (0, 1, 4)
    for i in idx:
        for kd in kinds:
            a = eng.neighbors(int(i), kd); b = orc.neighbors(int(i), kd)
            assert neighbor_sets_equal(a, b), "particle %d kind %d: %s vs %s" % (i, kd, sorted(a), sorted(b))
    del pre
    eng.close()


def _observables(st, ps, probe_x, probe_z):
    """Integrated observables of a dam break (north_star): surge-front position = largest x of a fluid particle, and a
    pressure probe = mean pressure of the fluid particles within 2 lo of the point (probe_x, *, probe_z) on the far wall side."""
    n = st.numParticles
    pos = st.vec("pos")[:n]; fluid = st.particleType[:n] == ps.fluid
    front = float(pos[fluid, 0].max())
    near = fluid & (np.abs(pos[:, 0] - probe_x) < 2.0 * ps.partDist) & (np.abs(pos[:, 2] - probe_z) < 2.0 * ps.partDist)
    probe = float(st.press[:n][near].mean()) if near.any() else 0.0
    return front, probe, int(near.sum())


@pytest.mark.gpu
def test_integrated_observables_over_a_run():
    """300 steps of a 3D WC polygon-wall dam break (the bucket order is rebuilt many times on the way): surge-front position
    within 1e-9 relative and pressure-probe history within 1e-6 of the reference scale rho*g*H, sampled every 25 steps,
    against the oracle on the same input.  (Rounding differences of ~1e-16 per step are amplified by the stiff equation of
    state, hence the looser pressure tolerance; positions integrate them and stay far tighter.)"""
    case = cases.dam3d(16, 8, 12, jitter=0.05, name="obs")
    ps, meshes, st0 = _build(case)
    want = st0.copy(); orc = _oracle(ps, want, meshes, jacobi=True); orc.init_step0()
    eng = _engine(ps, meshes, st0.numParticlesZero); eng.upload(st0); eng.init_step0()
    got = st0.copy()
    px, pz = 1.5 * ps.partDist, 1.5 * ps.partDist      # probe at the foot of the wall behind the column
    pscale = ps.DNS_FL1 * 9.81 * case.meta["H"]
    hist = []
    for k in range(12):
        orc.step(25); eng.step(25); eng.download(got, ["pos", "press", "particleType"])
        fg, pg, ng = _observables(got, ps, px, pz); fw, pw, nw = _observables(want, ps, px, pz)
        hist.append((fg, fw, pg, pw))
        # the same two observables reduced on the device (pmps_observe_*): no array download needed for a history
        mn, mx, nf = eng.observe_extent(ps.fluid)
        big = 1e9
        nb, psum, pmax = eng.observe_box(ps.fluid, [px - 2.0 * ps.partDist, -big, pz - 2.0 * ps.partDist], [px + 2.0 * ps.partDist, big, pz + 2.0 * ps.partDist])
        assert mx[0] == fg and nf == int((got.particleType[:got.numParticles] == ps.fluid).sum())
        assert nb == ng and abs(psum / nb - pg) <= 1e-12 * max(abs(pg), pscale)
        assert ng == nw and ng > 0
        assert abs(fg - fw) <= 1e-9 * abs(fw), "step %d: surge front %r vs %r" % (25 * (k + 1), fg, fw)
        assert abs(pg - pw) <= 1e-6 * pscale, "step %d: probe pressure %r vs %r (scale %g)" % (25 * (k + 1), pg, pw, pscale)
    assert hist[-1][1] > hist[0][1], "the front did not advance: %s" % hist
    assert eng.scalars().numParticles == want.numParticles
    eng.close()


def test_errors_are_loud():
    from polymps_b200.engine import Engine, PmpsError
    case = cases.dam3d(6, 5, 6, name="err")
    ps, meshes, st = _build(case)
    eng = Engine(ps, meshes)
    with pytest.raises(PmpsError):
        eng.calc_viscosity()          # nothing uploaded
    eng.upload(st)
    with pytest.raises(PmpsError):
        eng.calc_viscosity()          # buckets not built
    eng.close()


@pytest.mark.gpu
def test_host_driven_loop_keeps_lists_and_results():
    """upload -> step -> download every step (the e2e pattern of bench.py): sending back the positions just received must not
    force a re-sort (the displacement trackers decide, pmps_transfer.cu), and the run must equal the device-resident run bit for
    bit (same slot order, same candidate lists => same summation order).  A changed particle type does force the rebuild."""
    case = cases.dam3d(16, 8, 12, jitter=0.04, name="e2e")
    ps, meshes, st0 = _build(case)
    a = _engine(ps, meshes, st0.numParticlesZero); a.upload(st0); a.init_step0(); a.step(8)
    sa = st0.copy(); a.download(sa)
    b = _engine(ps, meshes, st0.numParticlesZero); b.upload(st0); b.init_step0()
    sb = st0.copy(); b.download(sb)
    l0 = b.launch_count()
    for _ in range(8):
        b.upload(sb); b.step(1); b.download(sb)
    per_step = (b.launch_count() - l0) / 8.0
    assert np.array_equal(sa.pos, sb.pos) and np.array_equal(sa.vel, sb.vel) and np.array_equal(sa.press, sb.press)
    # a forced rebuild per step launches the sort + list build (>= 12 kernels) on top of the step and the transfers
    b2 = _engine(ps, meshes, st0.numParticlesZero); b2.upload(st0); b2.init_step0(); s2 = st0.copy(); b2.download(s2)
    os.environ["PMPS_HARD_UPLOAD"] = "1"
    try:
        l0 = b2.launch_count()
        for _ in range(8):
            b2.upload(s2); b2.step(1); b2.download(s2)
        hard = (b2.launch_count() - l0) / 8.0
    finally:
        del os.environ["PMPS_HARD_UPLOAD"]
    assert per_step < hard - 8, (per_step, hard)
    assert not compare_states(s2, sa, rtol=1e-9, skip=NOISY, scales=natural_scales(ps))
    # the same loop through pmps_step_host (copies overlapped with the step on a second stream): bit-identical again
    b3 = _engine(ps, meshes, st0.numParticlesZero); b3.upload(st0); b3.init_step0(); s3 = st0.copy(); b3.download(s3)
    for _ in range(8):
        b3.step_host(s3)
    assert np.array_equal(sa.pos, s3.pos) and np.array_equal(sa.vel, s3.vel) and np.array_equal(sa.press, s3.press)
    assert np.array_equal(sa.pndi, s3.pndi) and np.array_equal(sa.particleBC, s3.particleBC) and np.array_equal(sa.mirrorParticlePos, s3.mirrorParticlePos)
    b3.close()
    # a particle turned into a ghost on the host: the upload notices the changed type word and the next step rebuilds
    sb.particleType[5] = ps.ghost
    b.upload(sb); b.step(1); b.download(sb)
    assert sb.particleType[5] == ps.ghost and b.scalars().numParticles == sb.numParticles
    for e in (a, b, b2):
        e.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(VARIANTS))
def test_option_variants_against_oracle(name):
    """Weight / gradient / repulsion / PND / free-surface / slip variants of the kernel family: init_step0 + 3 fused steps against
    the oracle (pinned bit-exactly to the live reference on the same variants by tests/test_oracle_variants.py)."""
    case = VARIANTS[name][1]()
    ps, meshes, st0 = _build(case)
    want = st0.copy(); orc = _oracle(ps, want, meshes, jacobi=True)
    orc.init_step0()
    eng = _engine(ps, meshes, st0.numParticlesZero)
    eng.upload(st0); eng.init_step0()
    got = st0.copy(); eng.download(got)
    bad = compare_states(got, want, rtol=RTOL, skip=NOISY, scales=natural_scales(ps))
    assert not bad, "step 0: %s" % bad
    for k in range(1, 4):
        orc.step(1); eng.step(1)
        eng.download(got)
        bad = compare_states(got, want, rtol=RTOL * 10 ** (k - 1), skip=NOISY, scales=natural_scales(ps))
        assert not bad, "%s step %d: %s" % (name, k, bad)
        assert abs(eng.scalars().velMax - orc.P.velMax) <= 1e-9 * max(orc.P.velMax, 1e-30)
    eng.close()


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["w2", "i3"])
def test_step_host_equals_resident_run(kind):
    """pmps_step_host on particle walls (velocities final only after the dummy-velocity pass) and on the incompressible scheme
    (pressure final after the Poisson solve): bit-identical to the device-resident run."""
    case = _seeded_case(kind)
    ps, meshes, st0 = _build(case)
    a = _engine(ps, meshes, st0.numParticlesZero); a.upload(st0); a.init_step0(); a.step(5)
    sa = st0.copy(); a.download(sa)
    b = _engine(ps, meshes, st0.numParticlesZero); b.upload(st0); b.init_step0(); sb = st0.copy(); b.download(sb)
    for _ in range(5):
        b.step_host(sb)
    for f in ("pos", "vel", "press", "pndi", "particleBC", "pressurePPE"):
        assert np.array_equal(getattr(sa, f), getattr(sb, f)), f
    a.close(); b.close()
