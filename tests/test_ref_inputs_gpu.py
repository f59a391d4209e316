"""The reference's OWN input documents (BASELINE.json configurations) through both binaries.

tests/golden/ref_inputs.json holds the eight documents as parsed from /root/reference/input/*.json (generator:
tests/golden/make_ref_inputs.py).  Each is used VERBATIM — every option, lo, dt, c0, domain, relax_pnd as the file has it; only
`pathNames` point at a grid / STL synthesized here at the file's own particle distance (the shipped grids are stripped LFS blobs)
and final_time / iter_output are shortened — by the stock reference binary (1 thread) and by the drop-in (unmodified main.cpp +
pmps_shim.cpp -> libpmps.so), and the .vtu files are compared.

Shifting: every WC configuration uses shifting type 2, which in the 1-thread reference is a Gauss-Seidel sweep and on any parallel
implementation a Jacobi one.  So the drop-in is checked twice: (a) against the CPU oracle in JACOBI mode on
the same input, to the 6 printed decimals — the tight check; (b) against the stock binary with a tolerance MEASURED on the same case
as 1.5 x the difference between the oracle's Gauss-Seidel and Jacobi modes (the oracle's Gauss-Seidel mode reproduces the stock
binary to the printed decimals, asserted here too).  Without shifting (INC configurations) (b) is at the printed decimals as well."""
import json
import os
import re
import shutil
import subprocess
import tempfile

import numpy as np
import pytest

from common import ROOT
from polymps_b200 import cases
from polymps_b200.params import ParticleSystem
from polymps_b200.mesh import PolygonMesh
from polymps_b200.state import HostState

REF_BIN = os.path.join(ROOT, "oracle", "_ref", "polymps_ref")
GPU_BIN = os.path.join(ROOT, "polymps_b200", "host", "_build", "polymps_gpu")
DOCS = json.load(open(os.path.join(ROOT, "tests", "golden", "ref_inputs.json")))
STEPS = 12


def build_case(name):
    doc = json.loads(json.dumps(DOCS[name])); lo = doc["numerical"]["particle_dist"]
    if name == "dam1610_BC":          # periodic in x, limits from the border particles (boundary.limit 0): the drifting channel reaches the face within the run
        c = cases.channel2d_periodic(64, 14, hump=6, lo=lo, name="vb")
    elif name == "damErosion3D_WC":   # non-Newtonian rheology with polygon walls
        c = cases.dam3d_erosion(10, 6, 10, 4, lo=lo, jitter=0.03, name="vb")
    elif name.startswith("InputDam"):
        c = cases.dam3d(12, 6, 8, lo=lo, jitter=0.04, name="vb")
    elif name.startswith("dam1610"):
        c = cases.dam2d_particle(20, 24, lo=lo, name="vb")
    elif name.startswith("2D_dam"):
        c = cases.dam2d_erosion(16, 16, 6, lo=lo, jitter=0.03, name="vb")
    else:
        c = cases.slosh3d(24, 6, 8, lo=lo, jitter=0.03, name="vb")
    c.options = doc   # the reference's document, verbatim; write_case redirects pathNames only
    num = doc["numerical"]; num["final_time"] = num["time_step"] * (STEPS + 0.5); num["iter_output"] = STEPS // 2
    return c


def _vtu(path):
    t = open(path).read(); out = {}
    for m in re.finditer(r"<DataArray type='(\w+)' Name='(\w+)'[^>]*>\s*(.*?)\s*</DataArray>", t, flags=re.S):
        out[m.group(2)] = np.array(m.group(3).split(), dtype=np.float64)
    return out


def _run(binary, root, name):
    r = subprocess.run([binary], input=name + "\n", cwd=root, env=dict(os.environ, OMP_NUM_THREADS="1"), capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, (binary, r.returncode, r.stdout[-2000:], r.stderr[-2000:])
    d = os.path.join(root, "output", name)
    files = sorted(f for f in os.listdir(d) if f.endswith(".vtu"))
    return [_vtu(os.path.join(d, f)) for f in files]


def _oracle(case, root, jacobi):
    import pyoracle
    opts = json.load(open(os.path.join(root, "input", case.name + ".json")))
    ps = ParticleSystem.from_json(opts).set_parameters(); ps.allocate_buckets(case.pos)
    if ps.numOfMeshs > 1 or ps.fluidType == 1 and False:
        return None, ps
    meshes = [PolygonMesh.from_triangles(case.tris).init_wall_tables(int(ps.dim), ps.weightType, ps.partDist, ps.reL, ps.reS)] if ps.wallType == 1 else []
    st = HostState.from_case(case, ps)
    orc = pyoracle.Oracle(ps, st, meshes, shifting_jacobi=jacobi); orc.init_step0(); orc.step(STEPS)
    return st, ps


PRINT = 2.1e-6   # one unit of the 6th printed decimal (+ rounding)
# .vtu array name (src/MpsInputOutput.cpp:1338-1570) -> state field
VTU_FIELD = {"Position": "pos", "Velocity": "vel", "pressure": "press", "BC": "particleBC", "pnd": "pndi", "pndSmall": "pndSmall", "nNeigh": "numNeigh",
             "npcdDeviation": "npcdDeviation", "polygonNormal": "polygonNormal", "concentration": "concentration", "GradCi": "gradConcentration",
             "RHO": "RHO", "MEU": "MEU", "PTYPE": "PTYPE", "ParticleType": "particleType", "nearMeshType": "nearMeshType",
             "ConcVol": "Cv", "MEUy": "MEU_Y", "II": "II", "p_smooth": "p_smooth", "Inertia": "Inertia", "Normal_Stress": "p_rheo_new", "VF": "VF"}
INT_FIELDS = ("BC", "nNeigh", "PTYPE", "ParticleType", "nearMeshType")


def _field(st, k, n):
    x = getattr(st, VTU_FIELD[k]); w = x.size // st.numParticlesZero
    return np.asarray(x[:n * w], dtype=np.float64)


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(DOCS))
def test_reference_documents_verbatim_through_both_binaries(name, capsys):
    if not (os.path.isfile(REF_BIN) and os.path.isfile(GPU_BIN)):
        pytest.skip("drop-in binaries not built (they need the reference tree at build time)")
    case = build_case(name)
    root = tempfile.mkdtemp(prefix="pmps_verbatim_")
    try:
        cases.write_case(case, root)
        ref = _run(REF_BIN, root, case.name)
        shutil.move(os.path.join(root, "output", case.name), os.path.join(root, "output", "ref"))
        gpu = _run(GPU_BIN, root, case.name)
        assert len(ref) == len(gpu) and len(ref) >= 3
        a, b = ref[-1], gpu[-1]
        assert a.keys() == b.keys() and a["Position"].shape == b["Position"].shape
        shifting = DOCS[name]["numerical"]["particle_shifting"]["type"] != 0
        tol = {k: PRINT for k in a}
        gs, ps = _oracle(case, root, jacobi=False)
        if gs is not None:   # (single-mesh cases) the oracle's Gauss-Seidel mode IS the stock binary, to the printed decimals
            n = gs.numParticles
            for k in a:   # every array of the file (values are printed as float32, 6 decimals)
                if k in VTU_FIELD and hasattr(gs, VTU_FIELD[k]):
                    w = _field(gs, k, n); err = float(np.abs(a[k] - w).max())
                    assert err <= (0.0 if k in INT_FIELDS else PRINT + 1e-6 * float(np.abs(w).max())), (name, "stock binary vs Gauss-Seidel oracle", k, err)
        if shifting and gs is not None:
            jc, _ = _oracle(case, root, jacobi=True)
            # (a) tight: drop-in == Jacobi oracle to the printed decimals, every field the file holds
            for k in a:
                if k in INT_FIELDS:
                    assert np.array_equal(b[k], _field(jc, k, n).astype(np.float64)), k
                elif k in VTU_FIELD and hasattr(jc, VTU_FIELD[k]):
                    w = _field(jc, k, n); err = float(np.abs(b[k] - w).max())
                    assert err <= PRINT + 1e-6 * float(np.abs(w).max()), (name, "drop-in vs Jacobi oracle", k, err)
            # (b) against the stock binary: the measured Gauss-Seidel / Jacobi difference of this very case and field, x 1.5
            for k in a:
                if k in VTU_FIELD and k not in INT_FIELDS and hasattr(jc, VTU_FIELD[k]):
                    tol[k] = PRINT + 1.5 * float(np.abs(_field(gs, k, n) - _field(jc, k, n)).max())
            with capsys.disabled():
                print("\n%s: Gauss-Seidel vs Jacobi shifting after %d steps: pos %.2e lo, vel %.2e, press %.2e" % (name, STEPS, (tol["Position"] - PRINT) / 1.5 / ps.partDist, (tol["Velocity"] - PRINT) / 1.5, (tol["pressure"] - PRINT) / 1.5))
        elif shifting:   # three-mesh cases without an oracle run here: stock binary only, loose bound of a few shifts on the positions
            lo = DOCS[name]["numerical"]["particle_dist"]
            tol = {k: None for k in a}; tol["Position"] = 0.05 * lo
        for k in a:
            if k in ("connectivity", "offsets", "types") or tol.get(k, PRINT) is None:
                continue
            if k in INT_FIELDS:   # free-surface flags / neighbour counts flip where the two sweeps leave a particle on either side of a threshold
                assert (a[k] != b[k]).mean() <= (0.05 if shifting else 0.0), (k, float((a[k] != b[k]).mean()))
                continue
            err = float(np.abs(a[k] - b[k]).max())
            assert err <= tol.get(k, PRINT) + 1e-8 * float(np.abs(a[k]).max()), (name, k, err, tol.get(k, PRINT))
    finally:
        shutil.rmtree(root, ignore_errors=True)
