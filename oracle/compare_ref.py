"""oracle/compare_ref.py — TEST INFRASTRUCTURE: replay a ref_driver dump directory through the C oracle stage by
stage (each stage starts from the reference's own pre-state) and report the first field that is not bit-identical."""
from __future__ import annotations
import glob, os, re, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__)); sys.path.insert(0, HERE); sys.path.insert(0, os.path.dirname(HERE))
import refrun, pyoracle
from polymps_b200.params import ParticleSystem
from polymps_b200.mesh import PolygonMesh
from polymps_b200.state import HostState, ALL_FIELDS

STAGE_MAP = {   # ref_driver stage name -> oracle calls
    "updateParticlesID": ["update_buckets"], "calcViscosity": ["calc_viscosity"], "addGravity": ["add_gravity"],
    "predictionVelocityPosition": ["predict"], "checkParticleOutDomain1": ["check_out_domain"], "collisions": ["collisions"],
    "wallClosestPoint": ["wall_closest_point"], "calcWallNPCD": ["wall_npcd"], "correctionMatrix1": ["correction_matrix"],
    "calcPndnNeighNPCD": ["pnd_neigh_npcd"], "calcPndDiffusiveTerm": ["pnd_diffusive"], "calcWallPndDiffusiveTerm": ["wall_pnd_diffusive"],
    "updatePnd": ["update_pnd"], "meanPndParticlesWallDummySurface": ["mean_pnd_wall_dummy_surface"], "updateParticleBC": ["update_bc"],
    "calcPress": ["calc_press"], "extrapolatePress": ["extrapolate_press"], "correctionMatrix2": ["correction_matrix"],
    "calcPressGradient": ["press_gradient"], "calcWallPressGradient": ["wall_press_gradient"], "calcWallViscosity": ["wall_viscosity"],
    "correctionVelocityPosition": ["correct"], "checkParticleOutDomain2": ["check_out_domain"],
    "calcConcAndConcGradient": ["shifting"], "calcWallConcAndConcGradient": ["wall_shifting"],
    "updateVelocityParticlesWallDummy": ["update_wall_dummy_vel"], "checkParticleOutDomain3": ["check_out_domain"],
}
SKIP = {"dvelCollision"}   # scratch arrays the dumps do not carry
TOL_FIELDS = {"press", "pressurePPE", "acc", "vel", "pos"}   # only after a PPE solve (Eigen's SIMD dots): tolerance


def load_case(json_path, dump_dir):
    ps = ParticleSystem.from_json(json_path).set_parameters()
    pre = refrun.load_dump(os.path.join(dump_dir, "pre_init.bin"))
    ps.allocate_buckets(pre["pos"].reshape(-1, 3))
    meshes = []
    if ps.wallType == 1:
        root = os.path.dirname(os.path.dirname(json_path))
        meshes = [PolygonMesh.from_stl(os.path.join(root, ps.meshRigidFilename)).init_wall_tables(int(ps.dim), ps.weightType, ps.partDist, ps.reL, ps.reS)]
    return ps, meshes


def diff_states(a: HostState, d: dict, exact=True, tol=0.0):
    bad = []
    for f in ALL_FIELDS:
        if f in SKIP or f not in d:
            continue
        x = getattr(a, f); y = d[f]
        if np.array_equal(x, y):
            continue
        with np.errstate(invalid="ignore"):
            err = np.abs(x.astype(np.float64) - y.astype(np.float64))
        scale = max(float(np.abs(y).max()), 1e-300)
        if not exact and err.max() <= tol * scale:
            continue
        bad.append((f, int(np.count_nonzero(x != y)), float(err.max()), float(err.max() / scale)))
    return bad


def replay(json_path, dump_dir, step, jacobi=False, verbose=True):
    ps, meshes = load_case(json_path, dump_dir)
    files = sorted(glob.glob(os.path.join(dump_dir, "s%04d_[0-9][0-9]_*.bin" % step)))
    prev = refrun.load_dump(os.path.join(dump_dir, "s%04d_pre.bin" % step))
    ok = True
    for fn in files:
        nm = re.match(r"s\d+_\d+_(.*)\.bin", os.path.basename(fn)).group(1)
        cur = refrun.load_dump(fn)
        st = HostState.from_dump(prev)
        orc = pyoracle.Oracle(ps, st, meshes, shifting_jacobi=jacobi)
        orc.first[:] = prev["firstParticleInBucket"]; orc.last[:] = prev["lastParticleInBucket"]; orc.next[:] = prev["nextParticleInSameBucket"]
        orc.P.velMax = float(prev["scalars"][0])
        for call in STAGE_MAP[nm]:
            getattr(orc, call)()
        bad = diff_states(st, cur, exact=not (nm == "calcPress" and ps.mpsType >= 2), tol=1e-9)
        if nm == "updateParticlesID":
            for a, b in (("firstParticleInBucket", orc.first), ("lastParticleInBucket", orc.last), ("nextParticleInSameBucket", orc.next)):
                if not np.array_equal(cur[a], b): bad.append((a, int(np.count_nonzero(cur[a] != b)), 0, 0))
        if nm == "correctionVelocityPosition" and orc.P.velMax != float(cur["scalars"][0]):
            bad.append(("velMax", 1, abs(orc.P.velMax - float(cur["scalars"][0])), 0))
        if st.numParticles != int(cur["numParticles"][0]): bad.append(("numParticles", st.numParticles, int(cur["numParticles"][0]), 0))
        if verbose: print("%-36s %s" % (nm, "bit-identical" if not bad else "MISMATCH %s" % bad))
        ok = ok and not bad
        prev = cur
    return ok


if __name__ == "__main__":
    sys.exit(0 if replay(sys.argv[1], sys.argv[2], int(sys.argv[3]), jacobi=False) else 1)
