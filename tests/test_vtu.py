"""Device-fed VTU output: pmps_write_vtu writes the bytes of MpsInputOutput::writeVtuAscii /
writeVtuAsciiFreeSurface (src/MpsInputOutput.cpp:1313-1875).  Three checks: the fixed-point formatter against printf's "%f" (CPU);
the file against a rendering of the reference's layout from the downloaded arrays (GPU); and — the one that involves the reference's
own code — the unmodified main.cpp + shim run twice on the same case, once with the reference's writer on the host arrays
(PMPS_VTU_HOST=1), once with the device writer: the files are identical byte for byte."""
import ctypes
import filecmp
import os
import shutil
import subprocess
import tempfile

import numpy as np
import pytest

from common import ROOT
from polymps_b200 import cases
from polymps_b200.engine import Engine, load_library
from polymps_b200.mesh import PolygonMesh
from polymps_b200.params import ParticleSystem
from polymps_b200.state import HostState

GPU_BIN = os.path.join(ROOT, "polymps_b200", "host", "_build", "polymps_gpu")


def test_fixed_point_formatter_prints_what_printf_prints():
    lib = load_library()
    rng = np.random.default_rng(1)
    vals = [0.0, -0.0, 0.5, 1.0 / 128, 3.0 / 128, 0.0000005, 0.0000015, 0.9999995, 0.99999949999, 1.9999995, 123456.7890125, -1e-9,
            -0.0000005, 2.5e-7, 7.5e-7, 1e15, 3.9e15, 5e15, 1e300, float("inf"), -float("inf"), 1e-320, 0.1, 0.2, 0.3, 1234567.0000005]
    vals += list(rng.uniform(-2, 2, 100000)) + list(rng.uniform(-1e6, 1e6, 100000)) + list(np.float32(rng.uniform(-100, 100, 100000)).astype(np.float64))
    vals += [k / 128.0 for k in range(-400, 400)] + [k / 128.0 + 1000.0 for k in range(256)]   # exact ties of the sixth decimal: round half even
    for k in range(1, 200, 2):
        vals += [np.nextafter(k / 128.0, 0), np.nextafter(k / 128.0, 10)]                          # and their neighbours
    v = np.array(vals, dtype=np.float64)
    buf = ctypes.create_string_buffer(30 * len(v) + 1000)
    n = lib.pmps_debug_format_f6(v.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), len(v), buf, len(buf))
    assert n > 0
    got = buf.value.decode().split(" ")[:-1]
    want = ["%f" % x for x in v]
    bad = [(x, g, w) for x, g, w in zip(v, got, want) if g != w]
    assert not bad, bad[:5]


def _render(st, ps, flags):
    """The reference's layout (src/MpsInputOutput.cpp:1313-1875) from host arrays."""
    n = st.numParticles
    fs = bool(flags & 64)
    sel = [i for i in range(n) if (not fs) or st.particleBC[i] == ps.surface or st.particleType[i] == ps.wall]
    m = len(sel)
    f32 = lambda a: np.asarray(a, dtype=np.float64).astype(np.float32).astype(np.float64)
    def F1(name, a):
        a = f32(a); return "        <DataArray type='Float32' Name='%s' format='ascii'>\n" % name + "".join("%f " % a[i] for i in sel) + "\n        </DataArray>\n"
    def F3(name, a):
        a = f32(a).reshape(-1, 3); return "        <DataArray type='Float32' Name='%s' NumberOfComponents='3' format='ascii'>\n" % name + "".join("%f %f %f " % tuple(a[i]) for i in sel) + "\n        </DataArray>\n"
    def I1(name, a, typ="Int32"):
        return "        <DataArray type='%s' Name='%s' format='ascii'>\n" % (typ, name) + "".join("%d " % a[i] for i in sel) + "\n        </DataArray>\n"
    pos = st.pos.reshape(-1, 3)
    t = "<?xml version='1.0' encoding='UTF-8'?>\n<VTKFile xmlns='VTK' byte_order='LittleEndian' version='0.1' type='UnstructuredGrid'>\n  <UnstructuredGrid>\n"
    t += "    <Piece NumberOfCells='%d' NumberOfPoints='%d'>\n      <Points>\n" % (m, m)
    t += "        <DataArray type='Float32' Name='Position' NumberOfComponents='3' format='ascii'>\n          " + "".join("%f %f %f " % tuple(pos[i]) for i in sel) + "\n        </DataArray>\n"
    t += "      </Points>\n      <PointData>\n"
    t += F3("Velocity", st.vel) + F1("pressure", st.press) + I1("BC", st.particleBC)
    if flags & 1:
        t += F1("pnd", st.pndi) + F1("pndSmall", st.pndSmall)
    if flags & 2:
        t += I1("nNeigh", st.numNeigh)
    if flags & 4:
        t += F3("npcdDeviation", st.npcdDeviation) + F3("polygonNormal", st.polygonNormal)
    if flags & 8:
        t += F1("concentration", st.concentration) + F3("GradCi", st.gradConcentration)
    if flags & 16:
        t += F1("RHO", st.RHO) + F1("ConcVol", st.Cv) + F1("MEU", st.MEU) + F1("MEUy", st.MEU_Y) + F1("II", st.II) + I1("PTYPE", st.PTYPE) + F1("p_smooth", st.p_smooth)
        if not fs:
            t += F1("Inertia", st.Inertia) + F1("Normal_Stress", st.p_rheo_new) + F1("VF", st.VF)
    if flags & 32:
        if not fs:
            t += I1("ParticleType", st.particleType)
        if ps.wallType == 1:
            t += I1("nearMeshType", st.nearMeshType)
    t += "      </PointData>\n      <Cells>\n"
    t += I1("connectivity", {i: k for k, i in enumerate(sel)}) + I1("offsets", {i: k + 1 for k, i in enumerate(sel)}) + I1("types", {i: 1 for i in sel}, "UInt8")
    t += "      </Cells>\n    </Piece>\n  </UnstructuredGrid>\n</VTKFile>\n"
    return t


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["wc3d", "wc3d_fs", "nn2d", "nn2d_fs"])
def test_device_writer_against_the_reference_layout(kind):
    if kind.startswith("wc3d"):
        case = cases.dam3d(12, 6, 8, jitter=0.05, name="vtu"); flags = 1 | 2 | 4 | 8 | 32
    else:
        case = cases.dam2d_erosion(16, 16, 6, jitter=0.03, name="vtu"); flags = 1 | 2 | 8 | 16 | 32
    if kind.endswith("_fs"):
        flags |= 64
    ps = ParticleSystem.from_json(case.options).set_parameters(); ps.allocate_buckets(case.pos)
    meshes = [PolygonMesh.from_triangles(case.tris).init_wall_tables(int(ps.dim), ps.weightType, ps.partDist, ps.reL, ps.reS)] if ps.wallType == 1 else []
    st = HostState.from_case(case, ps)
    eng = Engine(ps, meshes); eng.remember_capacity(st.numParticlesZero); eng.upload(st); eng.init_step0(); eng.step(5)
    root = tempfile.mkdtemp(prefix="pmps_vtu_")
    try:
        path = os.path.join(root, "output00001.vtu")
        eng.write_vtu(path, flags)
        got = open(path).read()
        host = eng.download(st.copy())
        want = _render(host, ps, flags)
        assert len(got) == len(want) and got == want, next((k, got[max(0, k - 60):k + 60], want[max(0, k - 60):k + 60]) for k in range(min(len(got), len(want))) if got[k] != want[k])
        if kind.endswith("_fs"):
            assert 0 < got.count("NumberOfCells") and int(got.split("NumberOfCells='")[1].split("'")[0]) < host.numParticles
    finally:
        shutil.rmtree(root, ignore_errors=True); eng.close()


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["wc3d", "wc3d_fs", "nn2d"])
def test_device_writer_and_reference_writer_put_out_the_same_bytes(kind):
    if not os.path.isfile(GPU_BIN):
        pytest.skip("drop-in binary not built (needs the reference tree at build time)")
    if kind.startswith("wc3d"):
        case = cases.dam3d(12, 6, 8, jitter=0.05, name="vtu", shifting=2)
        case.options["flags"]["output_VTU"].update(pnd=True, neigh=True, deviation=True, concentration=True, auxiliar=True, freeSurfWall=kind.endswith("_fs"))
    else:
        case = cases.dam2d_erosion(16, 16, 6, jitter=0.03, name="vtu")
        case.options["flags"]["output_VTU"].update(pnd=True, neigh=True, non_newtonian=True, auxiliar=True)
    num = case.options["numerical"]
    num["final_time"] = num["time_step"] * 10.5; num["iter_output"] = 5
    root = tempfile.mkdtemp(prefix="pmps_vtu2_")
    try:
        cases.write_case(case, root)
        outs = {}
        for mode in ("host", "device"):
            env = dict(os.environ, OMP_NUM_THREADS="1")
            if mode == "host":
                env["PMPS_VTU_HOST"] = "1"
            r = subprocess.run([GPU_BIN], input=case.name + "\n", cwd=root, env=env, capture_output=True, text=True, timeout=600)
            assert r.returncode == 0, (mode, r.stdout[-1500:], r.stderr[-1500:])
            os.rename(os.path.join(root, "output", case.name), os.path.join(root, "output", mode))
            outs[mode] = sorted(f for f in os.listdir(os.path.join(root, "output", mode)) if f.endswith(".vtu"))
        assert outs["host"] == outs["device"] and len(outs["host"]) >= 3
        for f in outs["host"]:
            assert filecmp.cmp(os.path.join(root, "output", "host", f), os.path.join(root, "output", "device", f), shallow=False), f
    finally:
        shutil.rmtree(root, ignore_errors=True)
