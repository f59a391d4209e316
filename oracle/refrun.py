"""oracle/refrun.py — TEST INFRASTRUCTURE: run the unmodified reference (oracle/_ref/ref_driver) on a synthesized
case and load its raw dumps.  Only usable where oracle/_ref has been built (needs /root/reference at build time);
tests that depend on it are skipped otherwise and rely on the committed fixtures under tests/golden/ instead."""
from __future__ import annotations

import os
import struct
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DRIVER = os.path.join(HERE, "_ref", "ref_driver")
_DT = {0: np.float64, 1: np.int32, 2: np.uint8}


def have_ref() -> bool:
    return os.path.isfile(REF_DRIVER) and os.access(REF_DRIVER, os.X_OK)


def load_dump(path: str) -> dict:
    out = {}
    with open(path, "rb") as fh:
        b = fh.read()
    o = 0
    while o < len(b):
        (l,) = struct.unpack_from("<I", b, o); o += 4
        name = b[o:o + l].decode(); o += l
        dt = b[o]; o += 1
        (n,) = struct.unpack_from("<Q", b, o); o += 8
        t = np.dtype(_DT[dt])
        out[name] = np.frombuffer(b, dtype=t, count=n, offset=o).copy(); o += n * t.itemsize
    return out


def run_reference(case_root: str, name: str, nsteps: int, dump_dir: str, dump_steps=(), threads: int = 1,
                  timing: bool = False) -> str:
    """Runs ref_driver with cwd=case_root. Returns stdout."""
    env = dict(os.environ, OMP_NUM_THREADS=str(threads))
    steps = ",".join(str(s) for s in dump_steps) if dump_steps else "none"
    cmd = [REF_DRIVER, name, str(nsteps), dump_dir, steps] + (["time"] if timing else [])
    r = subprocess.run(cmd, cwd=case_root, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("ref_driver failed (%d):\n%s" % (r.returncode, r.stdout[-4000:]))
    return r.stdout
