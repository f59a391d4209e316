"""Known-answer tests of the host-side mirror: polymps_b200/params.py = MpsInputOutput::readInputFile +
MpsParticle::setParameters + MpsBucket::allocateMemory, polymps_b200/mesh.py = PolygonMesh::initWijnNeigh + normals.  The answers are
the reference's own: every golden fixture carries the parameter block `ref_driver` dumped after the reference's setParameters
(tests/golden/make_golden.py, `params_json`) and the mesh + wall tables it built (`mesh` dump).  Runs without a GPU."""
import numpy as np
import pytest

from common import GOLDEN, golden


def _close(a, b, rel=1e-14):
    return a == b or abs(a - b) <= rel * max(abs(a), abs(b))


@pytest.mark.parametrize("name", GOLDEN)
def test_parameter_block_equals_the_references(name):
    order, dumps, meta, ps, meshes = golden(name)
    ref = meta["params"]
    checked, bad = 0, []
    # (the reference reads the rheology block only when the input file has one and otherwise leaves those members unset: not compared
    # for a Newtonian case, where nothing reads them)
    rheo = {"PHI_1", "PHI_2", "PHI_WAL", "PHI_BED", "VF_min", "VF_max", "cohes", "MEU0", "DG", "I0", "mm", "Cd", "N", "MEU_Y", "Fraction_method", "visc_max"}
    for k, want in ref.items():
        if not hasattr(ps, k) or isinstance(want, (list, dict, str)) or (k in rheo and int(ps.fluidType) == 0):
            continue
        got = getattr(ps, k)
        if isinstance(got, (bool, np.bool_)):
            got = int(got)
        if isinstance(want, bool):
            want = int(want)
        try:
            ok = _close(float(got), float(want))
        except (TypeError, ValueError):
            continue
        checked += 1
        if not ok:
            bad.append((k, got, want))
    assert checked >= 60, "only %d members compared" % checked
    assert not bad, bad


@pytest.mark.parametrize("name", [g for g in GOLDEN if g not in ("w2", "i2", "n2")])
def test_wall_tables_and_normals_equal_the_references(name):
    order, dumps, meta, ps, meshes = golden(name)
    d = dumps["mesh"]; m = meshes[0]
    assert np.array_equal(np.asarray(m.NF, dtype=np.int64).ravel(), np.asarray(d["mesh0_F"], dtype=np.int64).ravel())
    np.testing.assert_allclose(np.asarray(m.NV).ravel(), np.asarray(d["mesh0_V"]).ravel(), rtol=0, atol=0)
    np.testing.assert_allclose(np.asarray(m.NNormals).ravel(), np.asarray(d["mesh0_N"]).ravel(), rtol=1e-15, atol=1e-15)
    np.testing.assert_allclose(np.asarray(m.xDataPND), np.asarray(d["mesh0_xDataPND"]), rtol=1e-15, atol=0)
    np.testing.assert_allclose(np.asarray(m.xDataNeigh), np.asarray(d["mesh0_xDataNeigh"]), rtol=1e-15, atol=0)
    np.testing.assert_allclose(np.asarray(m.Zij), np.asarray(d["mesh0_Zij"]), rtol=1e-13, atol=1e-13)
    assert np.array_equal(np.asarray(m.nNeighWall, dtype=np.int64), np.asarray(d["mesh0_nNeighWall"], dtype=np.int64))
