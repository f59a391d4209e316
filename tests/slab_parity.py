#!/usr/bin/env python
"""Slab-decomposed run vs the single-GPU engine on the same case (launched under torch.distributed.run).

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 tests/slab_parity.py --backend nccl
    ... --backend gloo --one-gpu      # both ranks on cuda:0, messages staged through the host (single-GPU boxes)

Every rank builds the same small 3D weakly-compressible polygon-wall dam break with an initial drift towards +x (so that
bucket rebuilds and migration across the cuts happen within a few dozen steps), runs `--steps` steps as one slab of it,
and the owned particles' fields are compared by global id with a single-engine run of the whole case on rank 0.
The single-engine path is itself pinned to the oracle by tests/test_gpu_parity.py; differences here come only from the
summation order of neighbour sums (slot order differs), i.e. ~1e-15 relative.
"""
import argparse
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from polymps_b200 import cases  # noqa: E402
from polymps_b200.engine import make_engine  # noqa: E402
from polymps_b200.mesh import PolygonMesh  # noqa: E402
from polymps_b200.params import ParticleSystem  # noqa: E402
from polymps_b200.slab import SlabEngine, choose_cuts  # noqa: E402

FIELDS = ["pos", "vel", "press", "pndi", "particleBC", "numNeigh", "pndSmall", "concentration", "particleType", "pressurePPE"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default="nccl")
    ap.add_argument("--one-gpu", action="store_true")
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--nx", type=int, default=40); ap.add_argument("--ny", type=int, default=10); ap.add_argument("--nz", type=int, default=14)
    ap.add_argument("--drift", type=float, default=1.5)
    ap.add_argument("--rtol", type=float, default=1e-9)
    ap.add_argument("--inc", action="store_true", help="incompressible scheme: PPE assembly + CG sharded over the slabs")
    a = ap.parse_args()
    rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = 0 if a.one_gpu else int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if a.backend == "nccl":
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        dist.init_process_group("gloo")
    case = cases.dam3d(a.nx, a.ny, a.nz, jitter=0.05, name="slabparity", incompressible=a.inc)
    case.vel[:, 0] = a.drift
    ps = ParticleSystem.from_json(case.options).set_parameters(); ps.allocate_buckets(case.pos)
    meshes = [PolygonMesh.from_triangles(case.tris).init_wall_tables(int(ps.dim), ps.weightType, ps.partDist, ps.reL, ps.reS)]
    col = ((case.pos[:, 0] - ps.domainMinX) * ps.invBucketSide).astype(np.int64)
    cuts = choose_cuts(np.bincount(col, minlength=ps.numBucketsX), world)
    cuts_x = [ps.domainMinX + c * ps.bucketSide for c in cuts]
    mine = np.nonzero((case.pos[:, 0] >= cuts_x[rank]) & (case.pos[:, 0] < cuts_x[rank + 1]))[0]
    loc = {"ptype": case.ptype[mine], "pos": case.pos[mine], "vel": case.vel[mine], "press": case.press[mine], "gid": mine.astype(np.int64)}
    se = SlabEngine(ps, meshes, loc, cuts_x, rank, world, device=local, capacity=case.n + 1000)
    se.init_step0(); se.step(a.steps)
    got = se.download_owned(FIELDS)
    sc = se.scalars()
    stats = dict(se.dom.stats); stats.update(n_own=se.n_own, velMax=sc.velMax, launches=se.eng.launch_count(), solverIter=sc.solverIter)
    allgot = [None] * world; allstats = [None] * world
    dist.all_gather_object(allgot, got); dist.all_gather_object(allstats, stats)
    ok = True
    if rank == 0:
        ps1, m1, st, eng = make_engine(case, device=local)
        eng.init_step0(); eng.step(a.steps); eng.download(st); s1 = eng.scalars()
        gid = np.concatenate([g["gid"] for g in allgot])
        assert len(gid) == case.n and len(np.unique(gid)) == case.n, "owned sets do not partition the particles: %d ids, %d unique, %d particles" % (len(gid), len(np.unique(gid)), case.n)
        print("slab parity: %d particles, %d ranks, %d steps; per rank: %s" % (case.n, world, a.steps, allstats))
        for f in FIELDS:
            v = np.concatenate([g[f] for g in allgot]); w = getattr(st, f)
            w = w.reshape(-1, 3)[gid] if v.ndim == 2 else w[gid]
            if v.dtype.kind in "iu":
                bad = int((v != w).sum())
                print("  %-16s int mismatches %d" % (f, bad)); ok &= bad <= max(1, case.n // 2000)
            else:
                scale = max(float(np.abs(w).max()), 1e-300)
                err = float(np.abs(v - w).max()) / scale
                print("  %-16s max |diff| / max |ref| = %.3e" % (f, err)); ok &= err <= a.rtol
        dv = abs(allstats[0]["velMax"] - s1.velMax) / max(s1.velMax, 1e-300)
        if a.inc:
            print("  CG iterations of the last step: slabs %d, single engine %d" % (allstats[0]["solverIter"], s1.solverIter))
            ok &= abs(allstats[0]["solverIter"] - s1.solverIter) <= 3 and s1.solverIter > 3
        print("  velMax %.12g vs %.12g (rel %.2e); migrants %d, repartitions %d" % (allstats[0]["velMax"], s1.velMax, dv, sum(s["migrants_out"] for s in allstats), allstats[0]["repartitions"]))
        ok &= dv <= a.rtol
        if world > 1:
            ok &= sum(s["migrants_out"] for s in allstats) > 0 and allstats[0]["repartitions"] > 2
        print("SLAB PARITY " + ("OK" if ok else "FAILED"))
        eng.close()
    flag = torch.tensor([1.0 if ok else 0.0]); 
    if a.backend == "nccl":
        flag = flag.cuda()
    dist.broadcast(flag, 0)
    se.close(); dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1.0 else 1)


if __name__ == "__main__":
    main()
