"""Shared helpers of the test-suite: golden fixtures, case construction, field comparison."""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(HERE, "golden")):
    if p not in sys.path:
        sys.path.insert(0, p)

from polymps_b200.params import ParticleSystem  # noqa: E402
from polymps_b200.mesh import PolygonMesh  # noqa: E402
from polymps_b200.state import HostState, ALL_FIELDS, I32, U8  # noqa: E402
from make_golden import load_golden  # noqa: E402

GOLDEN = ["w3", "w3j", "i3", "w2", "i2", "n2", "b3", "e3", "o3"]
# reference dumps that pin the ORACLE only (tests/test_oracle_golden.py): periodic domains (2D, 3D, and the search from inside an aliased border
# layer) and the non-Newtonian rheology with polygon walls.  The device path is checked against the oracle on the same option sets
# (tests/test_periodic_gpu.py, tests/test_wall_rheology_gpu.py): a stage replay from a dump would need the reference's stale bucket lists.
GOLDEN_ORACLE = ["p2", "p3", "pe", "q3"]

# ref_driver stage name -> stage calls (oracle method names == engine method names)
STAGE_MAP = {
    "updateParticlesID": ["update_buckets"], "nonNewtonian1": ["calc_volume_fraction", "calc_viscosity_interaction"],
    "nonNewtonian2": ["calc_volume_fraction", "calc_viscosity_interaction"], "calcViscosity": ["calc_viscosity"], "addGravity": ["add_gravity"],
    "predictionVelocityPosition": ["predict"], "checkParticleOutDomain1": ["check_out_domain"], "collisions": ["collisions"],
    "wallClosestPoint": ["wall_closest_point"], "calcWallNPCD": ["wall_npcd"], "correctionMatrix1": ["correction_matrix"],
    "calcPndnNeighNPCD": ["pnd_neigh_npcd"], "calcPndDiffusiveTerm": ["pnd_diffusive"], "calcWallPndDiffusiveTerm": ["wall_pnd_diffusive"],
    "updatePnd": ["update_pnd"], "meanPndParticlesWallDummySurface": ["mean_pnd_wall_dummy_surface"], "updateParticleBC": ["update_bc"],
    "calcPress": ["calc_press"], "extrapolatePress": ["extrapolate_press"], "correctionMatrix2": ["correction_matrix"],
    "calcPressGradient": ["press_gradient"], "calcWallPressGradient": ["wall_press_gradient"], "calcWallViscosity": ["wall_viscosity"],
    "correctionVelocityPosition": ["correct"], "checkParticleOutDomain2": ["check_out_domain"],
    "calcShifting": ["shifting"], "calcConcAndConcGradient": ["shifting"], "calcWallConcAndConcGradient": ["wall_shifting"],
    "updateVelocityParticlesWallDummy": ["update_wall_dummy_vel"], "checkParticleOutDomain3": ["check_out_domain"],
}
# stages whose call list depends on the wall type
def stage_calls(stage, ps):
    if stage == "predictionPressGradient":   # relax_fact < 1: src/main.cpp:270-276
        return ["prediction_press_gradient"] + (["prediction_wall_press_gradient"] if ps.wallType == 1 else [])
    if stage == "meanPnd":   # pnd type 1: src/main.cpp:345-350
        return (["mean_wall_pnd"] if ps.wallType == 1 else []) + ["mean_pnd"]
    if stage in ("nonNewtonian1", "nonNewtonian2"):   # src/main.cpp:259-262, :387-399 (the polygon-wall variants only at the second site)
        return ["calc_volume_fraction", "calc_viscosity_interaction"] + (["calc_wall_viscosity_interaction"] if stage == "nonNewtonian2" and ps.wallType == 1 else [])
    if stage == "normals":   # ARC free-surface detection: src/main.cpp:351-357
        return ["calc_normals"] + (["calc_wall_normals"] if ps.wallType == 1 else [])
    return STAGE_MAP[stage]


SHIFTING_STAGES = {"calcShifting", "calcConcAndConcGradient", "calcWallConcAndConcGradient"}
# fields the reference dumps do not carry, or that are scratch in one of the implementations
SKIP_FIELDS = {"dvelCollision"}


def golden(name):
    """-> (order, dumps, meta, ps, meshes) for a committed fixture."""
    order, dumps, meta = load_golden(os.path.join(HERE, "golden", name + ".npz"))
    ps = ParticleSystem.from_json(meta["options"]).set_parameters()
    ps.allocate_buckets(np.asarray(meta["case"]["pos"]).reshape(-1, 3))
    meshes = []
    if ps.wallType == 1:
        meshes = [PolygonMesh.from_triangles(meta["case"]["tris"]).init_wall_tables(int(ps.dim), ps.weightType, ps.partDist, ps.reL, ps.reS)]
    return order, dumps, meta, ps, meshes


def stage_sequence(order, step):
    """[(stage name, previous dump name, this dump name)] of one dumped step, in execution order."""
    pre = "s%04d_pre" % step
    names = [o for o in order if o.startswith("s%04d_" % step) and o[6:8].isdigit()]
    out = []; prev = pre
    for n in names:
        out.append((n[9:], prev, n)); prev = n
    return out


def dumped_steps(order):
    return sorted({int(o[1:5]) for o in order if o.startswith("s") and o.endswith("_pre")})


def state_from(d) -> HostState:
    return HostState.from_dump(d)


def natural_scales(ps):
    """Magnitudes below which a field is rounding noise of the formula that produces it (used as the floor of the
    relative comparison).  The weakly-compressible equation of state is stiff: p = rho c0^2/gamma ((n/n0)^gamma - 1), so a
    1e-16 relative change of n moves p by rho c0^2 * 1e-16 however small p itself is; accelerations inherit it through
    grad p / rho ~ p / (rho lo)."""
    p0 = ps.DNS_FL1 * ps.soundSpeed ** 2 / ps.gamma
    return {"press": p0, "pressurePPE": p0, "acc": p0 / (ps.DNS_FL1 * ps.partDist), "accStar": p0 / (ps.DNS_FL1 * ps.partDist)}


def compare_states(got: HostState, want, rtol=0.0, fields=None, skip=(), live_only=True, scales=None):
    """List of (field, n_bad, max_abs_err, max_rel_err) for every field that differs by more than
    rtol * max(max|want|, scales[field]).  rtol=0 means bit-identical.  `want` is a HostState or a dump dict."""
    bad = []
    scales = scales or {}
    n = got.numParticles
    for f in (fields or ALL_FIELDS):
        if f in SKIP_FIELDS or f in skip:
            continue
        y = want.get(f) if isinstance(want, dict) else getattr(want, f)
        if y is None:
            continue
        x = getattr(got, f)
        k = x.size // got.numParticlesZero
        if live_only:
            x = x[:n * k]; y = np.asarray(y)[:n * k]
        if np.array_equal(x, y):
            continue
        if f in I32 or f in U8:
            bad.append((f, int(np.count_nonzero(x != y)), float(np.abs(x.astype(np.int64) - y.astype(np.int64)).max()), float("inf")))
            continue
        with np.errstate(invalid="ignore"):
            err = np.abs(x - y)
        if not np.all(np.isfinite(x) == np.isfinite(y)):
            bad.append((f, -1, float("nan"), float("nan"))); continue
        scale = max(float(np.abs(y[np.isfinite(y)]).max()) if y.size else 0.0, float(scales.get(f, 0.0)), 1e-300)
        e = float(np.nanmax(err)) if err.size else 0.0
        if e <= rtol * scale:
            continue
        bad.append((f, int(np.count_nonzero(err > rtol * scale)), e, e / scale))
    return bad


def neighbor_sets_equal(a, b):
    return sorted(int(v) for v in a) == sorted(int(v) for v in b)


# Option variants of the same kernel family (weight types, gradient types, slip condition, repulsion models, PND
# types, free-surface tests).  name -> (builder of a SMALL case for the live reference run, builder of the seeded GPU case)
def _v3(**kw):
    from polymps_b200 import cases
    return (lambda: cases.dam3d(6, 5, 5, jitter=0.05, name="var", **kw), lambda: cases.dam3d(14, 8, 10, jitter=0.04, name="var", **kw))


def _v2(**kw):
    from polymps_b200 import cases
    return (lambda: cases.dam2d_particle(8, 10, tank_nx=18, tank_nz=16, name="var", **kw), lambda: cases.dam2d_particle(24, 30, name="var", **kw))


def _jittered(build, amp=0.04, seed=7):
    """Fluid particles moved off the lattice: the dynamic collision model is discontinuous at zero approach speed (a separating
    pair is pushed apart by a finite background pressure, an approaching one is not, src/MpsParticleCollision.cpp:212-241), so on
    a perfect lattice — neighbours side by side with relative velocities of rounding size — the branch is decided by noise."""
    def f():
        c = build(); rng = np.random.default_rng(seed); lo = c.options["numerical"]["particle_dist"]
        fl = c.ptype == 0
        c.pos[fl, :2] += rng.uniform(-amp * lo, amp * lo, size=(int(fl.sum()), 2))
        return c
    return f


def _splash(a, b):
    from polymps_b200 import cases
    return (lambda: cases.dam3d_splash(*a, name="var"), lambda: cases.dam3d_splash(*b, name="var"))


def _per2(**kw):
    from polymps_b200 import cases
    return (lambda: cases.channel2d_periodic(24, 6, hump=3, name="var", **kw), lambda: cases.channel2d_periodic(64, 14, hump=6, name="var", **kw))


def _per3(**kw):
    from polymps_b200 import cases
    return (lambda: cases.channel3d_periodic(5, 22, 5, jitter=0.05, name="var", **kw), lambda: cases.channel3d_periodic(8, 30, 8, jitter=0.04, name="var", **kw))


VARIANTS = {
    # periodic domain (one direction): alias bucket layers + shifted pair geometry + the wrap at every out-of-domain check
    # (src/MpsBucket.cpp:142-289,356-417, src/MpsParticle.cpp:35-54,236-269); the drifting channel wraps particles from the first step on
    "2d_per": _per2(), "2d_per_right": _per2(drift=-1.0, window=-0.45), "2d_per_inc": _per2(incompressible=True), "2d_per_dpc": _per2(collision=1),
    # (the larger block also for the live reference run: in its first step a near-wall particle crosses the face during the shifting and
    # takes the wall-shifting pass from INSIDE the aliased border layer — its search then runs over the end of a bucket row)
    # non-Newtonian rheology with polygon walls (option set of input/damErosion3D_WC.json): the wall variants of the viscosity model, src/MpsViscosity.cpp:703-1526
    "nn3d_poly": (lambda: __import__("polymps_b200.cases", fromlist=["x"]).dam3d_erosion(5, 4, 5, 3, jitter=0.03, name="var"),
                  lambda: __import__("polymps_b200.cases", fromlist=["x"]).dam3d_erosion(10, 6, 10, 4, jitter=0.03, name="var")),
    "nn3d_poly_slip": (lambda: __import__("polymps_b200.cases", fromlist=["x"]).dam3d_erosion(5, 4, 5, 3, jitter=0.03, name="var", slip=1),
                       lambda: __import__("polymps_b200.cases", fromlist=["x"]).dam3d_erosion(10, 6, 10, 4, jitter=0.03, name="var", slip=1)),
    "3d_per_edge": (lambda: __import__("polymps_b200.cases", fromlist=["x"]).channel3d_periodic(8, 30, 8, jitter=0.04, name="var"),) * 2,
    # the alias layers under more of the kernel family: velocity shifting, mean PND, correction matrix + gradient type 0, free slip, DPC, E-MPS
    "2d_per_shift1": _per2(shifting=1), "2d_per_pnd1": _per2(pnd_type=1), "3d_per_gradcorr": _per3(grad_type=0, grad_corr=True, pnd_type=0),
    "3d_per_slip": _per3(slip=1), "3d_per_dpc": tuple(_jittered(b) for b in _per3(collision=1)), "3d_per_emps": _per3(mps_type=0),
    "3d_per": _per3(), "3d_per_right": _per3(drift=-1.0, window=-0.45), "3d_per_inc": _per3(incompressible=True),
    "ood": _splash((6, 5, 5), (14, 8, 10)),   # particles leave the domain (swap-with-last compaction, src/MpsParticle.cpp:286-396)
    "slip": _v3(slip=1), "grad0": _v3(grad_type=0), "grad1": _v3(grad_type=1), "grad2": _v3(grad_type=2),
    "gradcorr": _v3(grad_type=3, grad_corr=True), "rep0": _v3(rep_type=0), "rep1": _v3(rep_type=1), "rep3": _v3(rep_type=3),
    "w1": _v3(weight_type=1), "w2": _v3(weight_type=2), "w3": _v3(weight_type=3), "w4": _v3(weight_type=4),
    "pnd0": _v3(pnd_type=0), "pnd1": _v3(pnd_type=1), "fs0": _v3(fs_type=0), "noshift": _v3(shifting=0), "shift1": _v3(shifting=1), "2d_shift1": _v2(shifting=1),
    "divu": _v3(incompressible=True, mps_type=3), "divu_slip": _v3(incompressible=True, mps_type=3, slip=1),
    "2d_divu": _v2(incompressible=True, mps_type=3), "divu_nocorr": _v3(incompressible=True, mps_type=3, grad_corr=False, grad_type=3),
    "dpc": _v3(collision=1), "2d_dpc": tuple(_jittered(b) for b in _v2(collision=1)),
    "relax": _v3(relax=0.7), "relax_grad0": _v3(relax=0.5, grad_type=0), "2d_relax": _v2(relax=0.6),
    "arc": _v3(fs_type=2), "2d_arc": _v2(fs_type=2), "arc_pnd0": _v3(fs_type=2, pnd_type=0),   # ARC free-surface detection + particle normals
    "2d_grad3": _v2(grad_type=3, grad_corr=False), "2d_w3": _v2(weight_type=3), "2d_fs1": _v2(fs_type=1), "2d_pnd0": _v2(pnd_type=0),
}
