"""Generates the committed golden fixtures by running the UNMODIFIED reference (oracle/_ref/ref_driver, built by
`make -C oracle ref` from /root/reference) on small synthesized cases, single-threaded, and packing its raw stage-by-stage
dumps into one compressed .npz per case.  Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

Each npz holds  "<dump file stem>/<array name>"  entries (dump files: pre_init, init, s0001_pre, s0001_NN_<stage>, ...),
plus "case/*" (the synthesized input: ptype, pos, vel, press, tris) and "case/options_json".
"""
from __future__ import annotations

import glob
import json
import os
import shutil
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
from polymps_b200 import cases  # noqa: E402
import refrun  # noqa: E402


def golden_cases():
    """name -> (Case, steps to dump).  Small on purpose: the whole set must stay around a megabyte."""
    return {
        # 3D weakly-compressible, polygon tank (InputDamWC option set); exact lattice => equidistant-wall ties
        "w3": (cases.dam3d(7, 5, 6, name="w3"), (1, 2)),
        # same with jittered positions and a few steps of evolution
        "w3j": (cases.dam3d(6, 5, 6, jitter=0.08, name="w3j"), (3,)),
        # 3D incompressible (InputDamINC option set): gradient type 0 + correction matrix, PPE + CG
        "i3": (cases.dam3d(6, 5, 5, incompressible=True, jitter=0.05, name="i3"), (1, 2)),
        # 2D particle walls (dam1610_h300_lo0p0050_{WC,INC} option sets)
        "w2": (cases.dam2d_particle(8, 10, tank_nx=18, tank_nz=16, name="w2"), (1, 3)),
        "i2": (cases.dam2d_particle(8, 10, tank_nx=18, tank_nz=16, incompressible=True, name="i2"), (1, 2)),
        # 3D explicit MPS (mps_type 0: E-MPS equation of state, src/MpsPressure.cpp:62-108), otherwise the WC option set
        "e3": (cases.dam3d(6, 5, 5, jitter=0.05, name="e3", mps_type=0), (1, 2)),
        # 3D incompressible with the BiCGSTAB solver (solver_type 1)
        "b3": (cases.dam3d(6, 5, 5, incompressible=True, jitter=0.05, name="b3", solver=1), (1, 2)),
        # 2D erosional dam break, non-Newtonian two-phase rheology (2D_dam_fraccarollo_lo02p50e-03_WC option set, BASELINE configs[0])
        "n2": (cases.dam2d_erosion(8, 8, 4, tank_nx=20, tank_nz=15, jitter=0.05, name="n2"), (1, 2, 6)),
        # particles leave the domain in the out-of-domain checks of steps 3-7 (dumped: 4 and 5), the last live particle among them
        # (serial swap-with-last compaction, src/MpsParticle.cpp:286-396)
        "o3": (cases.dam3d_splash(7, 5, 6, name="o3"), (4, 5)),
        # ---- oracle-only fixtures (tests/test_oracle_golden.py, GOLDEN_ORACLE in tests/common.py): features of round 2 pinned where the
        # reference tree is absent.  Periodic domain, 2D particle walls and 3D polygon walls: particles wrap in the first dumped step;
        "p2": (cases.channel2d_periodic(24, 6, hump=3, name="p2"), (1, 2)),
        "p3": (cases.channel3d_periodic(5, 22, 5, jitter=0.05, name="p3"), (1, 2)),
        # the larger periodic block: in step 1 a near-wall particle crosses the face during the shifting and takes the wall-shifting pass from
        # inside the aliased border layer (the shift is that of the bucket actually read, oracle periodic_shift);
        "pe": (cases.channel3d_periodic(8, 30, 8, jitter=0.04, name="pe"), (1,)),
        # non-Newtonian rheology with polygon walls (option set of input/damErosion3D_WC.json): the wall variants of the viscosity model
        "q3": (cases.dam3d_erosion(5, 4, 5, 3, jitter=0.03, name="q3"), (1, 2)),
    }


def chrono_key(stem):
    if stem == "mesh":
        return (-3, 0)
    if stem == "pre_init":
        return (-2, 0)
    if stem == "init":
        return (-1, 0)
    if stem == "final":
        return (10 ** 9, 0)
    step = int(stem[1:5]); rest = stem[6:]
    return (step, -1 if rest == "pre" else 10 ** 6 if rest == "end" else int(rest[:2]))


def load_golden(path):
    """-> (ordered list of dump names, {name: full dict of arrays}) with the delta encoding undone."""
    z = np.load(path)
    order = bytes(z["order"]).decode().split("\n")
    by = {}
    for k in z.files:
        if "/" in k:
            stem, arr = k.split("/", 1)
            by.setdefault(stem, {})[arr] = z[k]
    dumps = {}; cur = {}
    for stem in order:
        cur = dict(cur); cur.update(by.get(stem, {}))
        dumps[stem] = cur
    meta = {"params": json.loads(bytes(z["params_json"]).decode()), "options": json.loads(bytes(z["case/options_json"]).decode()),
            "case": by.get("case", {})}
    return order, dumps, meta


def reference_dumps(case, steps):
    """Runs the unmodified reference on `case` right now (needs oracle/_ref) -> (order, dumps, params) like load_golden."""
    root = tempfile.mkdtemp(prefix="refdump_")
    try:
        cases.write_case(case, root)
        dump = os.path.join(root, "dump")
        refrun.run_reference(root, case.name, max(steps), dump, dump_steps=steps, threads=1)
        stems = [os.path.splitext(os.path.basename(fn))[0] for fn in glob.glob(os.path.join(dump, "*.bin"))]
        stems.sort(key=chrono_key)
        dumps = {stem: refrun.load_dump(os.path.join(dump, stem + ".bin")) for stem in stems}
        with open(os.path.join(dump, "params.json")) as fh:
            params = json.load(fh)
        return stems, dumps, params
    finally:
        shutil.rmtree(root, ignore_errors=True)


def main():
    if not refrun.have_ref():
        sys.exit("oracle/_ref/ref_driver missing: run `make -C oracle ref` (needs /root/reference)")
    only = set(sys.argv[1:])   # optional: names of the fixtures to (re)generate
    for name, (case, steps) in golden_cases().items():
        if only and name not in only:
            continue
        root = tempfile.mkdtemp(prefix="golden_")
        try:
            cases.write_case(case, root)
            dump = os.path.join(root, "dump")
            refrun.run_reference(root, case.name, max(steps), dump, dump_steps=steps, threads=1)
            out = {}
            stems = [os.path.splitext(os.path.basename(fn))[0] for fn in glob.glob(os.path.join(dump, "*.bin"))]
            stems.sort(key=chrono_key)
            prev = {}
            for stem in stems:   # delta encoding: an array is stored only where it differs from the previous dump
                for k, v in refrun.load_dump(os.path.join(dump, stem + ".bin")).items():
                    if k in prev and prev[k].shape == v.shape and np.array_equal(prev[k], v, equal_nan=True):
                        continue
                    out["%s/%s" % (stem, k)] = v
                    prev[k] = v
            out["order"] = np.frombuffer("\n".join(stems).encode(), dtype=np.uint8)
            with open(os.path.join(dump, "params.json")) as fh:
                out["params_json"] = np.frombuffer(fh.read().encode(), dtype=np.uint8)
            out["case/ptype"] = case.ptype; out["case/pos"] = case.pos; out["case/vel"] = case.vel; out["case/press"] = case.press
            if case.tris is not None:
                out["case/tris"] = case.tris
            out["case/options_json"] = np.frombuffer(json.dumps(case.options).encode(), dtype=np.uint8)
            path = os.path.join(HERE, name + ".npz")
            np.savez_compressed(path, **out)
            print("%-4s N=%d  %d arrays  %.0f kB" % (name, case.n, len(out), os.path.getsize(path) / 1e3))
        finally:
            shutil.rmtree(root, ignore_errors=True)


if __name__ == "__main__":
    main()
