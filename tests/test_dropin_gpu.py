"""The literal drop-in: the reference's UNMODIFIED main.cpp and host code linked with polymps_b200/host/pmps_shim.cpp (hot-path
methods forwarded to libpmps.so) against the stock reference binary, on the same synthesized case directory, compared through
the files both write (.vtu, ASCII %lf: 6 decimals).  Both binaries are built only where the reference tree exists
(oracle/Makefile `ref`, polymps_b200/host/Makefile) and travel to the GPU box as files; the test skips when they are absent.
Shifting is off in these cases: the reference's 1-thread shifting loop is an order-dependent Gauss-Seidel sweep."""
import os
import re
import shutil
import subprocess
import tempfile

import numpy as np
import pytest

from common import ROOT
from polymps_b200 import cases

REF_BIN = os.path.join(ROOT, "oracle", "_ref", "polymps_ref")
GPU_BIN = os.path.join(ROOT, "polymps_b200", "host", "_build", "polymps_gpu")


def _vtu(path):
    t = open(path).read()
    out = {}
    for m in re.finditer(r"<DataArray type='(\w+)' Name='(\w+)'[^>]*>\s*(.*?)\s*</DataArray>", t, flags=re.S):
        out[m.group(2)] = np.array(m.group(3).split(), dtype=np.float64)
    return out


def _run(binary, root, name, threads=1):
    env = dict(os.environ, OMP_NUM_THREADS=str(threads))
    r = subprocess.run([binary], input=name + "\n", cwd=root, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (binary, r.returncode, r.stdout[-2000:], r.stderr[-2000:])
    return r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["wc3d", "inc3d", "wc2d", "nn2d", "slosh3d", "arc3d", "per2d", "per3d", "nn3d"])
def test_unmodified_main_with_shim_matches_stock_reference(kind):
    if not (os.path.isfile(REF_BIN) and os.path.isfile(GPU_BIN)):
        pytest.skip("drop-in binaries not built (they need the reference tree at build time)")
    if kind == "wc3d":
        case = cases.dam3d(14, 7, 10, jitter=0.05, name="dropin", shifting=0)
    elif kind == "arc3d":   # ARC free-surface detection (free_surface_threshold.type 2): particle normals + arc test on the device
        case = cases.dam3d(14, 7, 10, jitter=0.05, name="dropin", shifting=0, fs_type=2)
    elif kind == "per2d":   # periodic domain (domain.boundary.type 1), particle walls: the drifting channel wraps particles from the first step on
        case = cases.channel2d_periodic(64, 14, hump=6, name="dropin", shifting=0)
    elif kind == "per3d":   # periodic in y, polygon walls
        case = cases.channel3d_periodic(8, 30, 8, jitter=0.04, name="dropin", shifting=0)
    elif kind == "nn3d":   # option set of input/damErosion3D_WC.json: non-Newtonian rheology with POLYGON walls (the wall variants of the viscosity model)
        case = cases.dam3d_erosion(10, 6, 10, 4, jitter=0.03, name="dropin", shifting=0)
        case.options["flags"]["output_VTU"].update(non_newtonian=True, auxiliar=True)
    elif kind == "inc3d":
        case = cases.dam3d(12, 6, 8, jitter=0.03, incompressible=True, name="dropin")
    elif kind == "slosh3d":   # BASELINE configs[3]: three meshes, the tank is the forced mesh the host rotates every step
        case = cases.slosh3d(30, 6, 8, jitter=0.03, name="dropin", shifting=0)
    elif kind == "nn2d":   # BASELINE configs[0]: 2D erosional dam break, non-Newtonian rheology (rheology arrays written to the .vtu)
        case = cases.dam2d_erosion(20, 20, 8, jitter=0.03, name="dropin", shifting=0)
        case.options["flags"]["output_VTU"].update(non_newtonian=True, auxiliar=True)
    else:
        case = cases.dam2d_particle(24, 30, name="dropin", shifting=0)
    num = case.options["numerical"]
    num["final_time"] = num["time_step"] * 30.5; num["iter_output"] = 10
    if kind == "per3d":   # 20 steps: in step 21 of this case a near-wall particle sits on a discrete choice that amplifies the 1e-13 differences of two
        num["final_time"] = num["time_step"] * 20.5; num["iter_output"] = 5   # summation orders a thousandfold (every stage of steps 19-22 agrees to 1e-11 from the same input)
    root = tempfile.mkdtemp(prefix="pmps_dropin_")
    try:
        cases.write_case(case, root)
        _run(REF_BIN, root, case.name)
        os.rename(os.path.join(root, "output", case.name), os.path.join(root, "output", "ref"))
        out = _run(GPU_BIN, root, case.name)
        assert "End PolyMPS" in out
        files = sorted(f for f in os.listdir(os.path.join(root, "output", "ref")) if f.endswith(".vtu"))
        assert len(files) >= 4, files   # iterations 0, 10, 20, 30
        for f in files:
            a = _vtu(os.path.join(root, "output", "ref", f)); b = _vtu(os.path.join(root, "output", case.name, f))
            assert a.keys() == b.keys() and "Position" in a and "pressure" in a
            for k in a:
                assert a[k].shape == b[k].shape, (f, k)
                if k in ("BC", "connectivity", "offsets", "types"):
                    assert np.array_equal(a[k], b[k]), (f, k)
                else:   # 6 printed decimals: one unit of the last digit + 1e-8 relative
                    tol = 2.1e-6 + 1e-8 * np.abs(a[k])
                    bad = np.abs(a[k] - b[k]) > tol
                    assert not bad.any(), (f, k, int(bad.sum()), float(np.abs(a[k] - b[k]).max()))
    finally:
        shutil.rmtree(root, ignore_errors=True)
