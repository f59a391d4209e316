"""Slab decomposition on the device: two ranks sharing cuda:0 (gloo, host-staged messages) against the single-GPU engine.
The multi-GPU NCCL run of the same script is `gpurun --gpus 2 -- python -m torch.distributed.run ... tests/slab_parity.py`."""
import os
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.gpu
@pytest.mark.parametrize("world,scheme", [(2, "wc"), (3, "wc"), (2, "inc"), (2, "splash"), (2, "arc"), (2, "arc_inc"), (2, "pnd1")])
def test_slab_run_matches_single_engine(world, scheme):
    port = 29700 + (os.getpid() % 1000) + world + {"wc": 0, "inc": 7, "splash": 13, "arc": 19, "arc_inc": 23, "pnd1": 29}[scheme]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(HERE, "slab_parity.py"), "--backend", "gloo", "--one-gpu", "--steps", "40"]
    if scheme == "splash":   # particles leave through the open top: out-of-domain removal on slab ranks
        cmd += ["--splash", "4.0", "--drift", "0.5"]
    if scheme.startswith("arc"):   # ARC free-surface detection on slab ranks (normals + arc test from the halo copies' positions)
        cmd += ["--arc"] + (["--inc", "--steps", "20"] if scheme == "arc_inc" else [])
    if scheme == "pnd1":
        cmd += ["--pnd1"]
    if scheme == "inc":   # PPE assembly + CG sharded over the slabs (all-reduced dot products, halo refresh of p per iteration)
        cmd += ["--inc", "--steps", "20"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    sys.stdout.write(r.stdout[-4000:]); sys.stderr.write(r.stderr[-4000:])
    assert r.returncode == 0 and "SLAB PARITY OK" in r.stdout
