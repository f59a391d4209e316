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
    ap.add_argument("--splash", type=float, default=0.0, help="upward velocity of the top two particle layers under a low domain lid: they leave the domain (out-of-domain removal on slabs)")
    ap.add_argument("--inc", action="store_true", help="incompressible scheme: PPE assembly + CG sharded over the slabs")
    ap.add_argument("--pnd1", action="store_true", help="pnd type 1 (mean PND over the neighbours): one more halo refresh inside the phase")
    ap.add_argument("--arc", action="store_true", help="ARC free-surface detection (free_surface_threshold.type 2): particle normals + arc test on slab ranks")
    a = ap.parse_args()
    rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = 0 if a.one_gpu else int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if a.backend == "nccl":
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        dist.init_process_group("gloo")
    tank = None
    if a.splash > 0.0:   # open-top tank barely higher than the column: what flies up leaves [domainMin + bucketSide, domainMax - bucketSide]
        tank = ((1.61 / 0.6) * a.nx * 0.01, a.ny * 0.01, (a.nz + 1.5) * 0.01)
    case = cases.dam3d(a.nx, a.ny, a.nz, jitter=0.05, name="slabparity", incompressible=a.inc, tank=tank, **({"fs_type": 2} if a.arc else {}), **({"pnd_type": 1} if a.pnd1 else {}))
    case.vel[:, 0] = a.drift
    if a.splash > 0.0:
        # A patch of the two top layers, away from the END of the particle array (x fastest, then y, then z: the array ends with the
        # far rows of the top layer).  The reference removes a particle by copying the LAST live one over it inside a serial loop and
        # does not re-test the copy in the same call (src/MpsParticle.cpp:286-396): when the last particle is itself outside, it
        # survives until the next check.  The single engine replays that; a slab rank removes every particle at the check that finds
        # it (array order has no meaning across ranks).  The patch keeps the two behaviours identical.
        sel = (case.pos[:, 2] > (a.nz - 2) * 0.01) & (case.pos[:, 0] < 0.25 * a.nx * 0.01) & (case.pos[:, 1] < 0.5 * a.ny * 0.01)
        case.vel[sel, 2] = a.splash
    ps = ParticleSystem.from_json(case.options).set_parameters(); ps.allocate_buckets(case.pos)
    meshes = [PolygonMesh.from_triangles(case.tris).init_wall_tables(int(ps.dim), ps.weightType, ps.partDist, ps.reL, ps.reS)]
    col = ((case.pos[:, 0] - ps.domainMinX) * ps.invBucketSide).astype(np.int64)
    cuts = choose_cuts(np.bincount(col, minlength=ps.numBucketsX), world)
    cuts_x = [ps.domainMinX + c * ps.bucketSide for c in cuts]
    mine = np.nonzero((case.pos[:, 0] >= cuts_x[rank]) & (case.pos[:, 0] < cuts_x[rank + 1]))[0]
    loc = {"ptype": case.ptype[mine], "pos": case.pos[mine], "vel": case.vel[mine], "press": case.press[mine], "gid": mine.astype(np.int64)}
    se = SlabEngine(ps, meshes, loc, cuts_x, rank, world, device=local, capacity=case.n + 1000, cuts=cuts)
    se.init_step0(); se.step(a.steps)
    got = se.download_owned(FIELDS)
    sc = se.scalars()
    stats = dict(se.dom.stats); stats.update(cuts=list(se.dom.cuts), n_own=se.n_own, velMax=sc.velMax, launches=se.eng.launch_count(), solverIter=sc.solverIter)
    allgot = [None] * world; allstats = [None] * world
    dist.all_gather_object(allgot, got); dist.all_gather_object(allstats, stats)
    ok = True
    if rank == 0:
        ps1, m1, st, eng = make_engine(case, device=local)
        if a.splash > 0.0:   # a slab rank's table order means nothing across ranks: it cannot replay the reference's stale bucket lists after a removal
            eng.set_stale_list_replay(False)
        eng.init_step0(); eng.step(a.steps); eng.download(st); s1 = eng.scalars()
        gid = np.concatenate([g["gid"] for g in allgot])
        alive = np.concatenate([g["particleType"] for g in allgot]) != ps.ghost
        n_alive = int(alive.sum())
        assert len(np.unique(gid)) == len(gid), "a particle is owned twice"
        if a.splash > 0.0 and (n_alive < case.n or s1.numParticles < case.n):   # the single engine compacts its arrays when a particle leaves (swap with the last): compare the survivors by position order instead of id
            assert n_alive == s1.numParticles, "slabs keep %d particles, single engine %d of %d" % (n_alive, s1.numParticles, case.n)
            print("slab parity (splash): %d of %d particles left the domain on both" % (case.n - n_alive, case.n))
            sp = np.concatenate([g["pos"] for g in allgot])[alive]; wp = st.pos.reshape(-1, 3)[:s1.numParticles]
            from scipy.spatial import cKDTree
            dd, kw = cKDTree(wp).query(sp)   # nearest single-engine survivor of every slab survivor
            assert len(np.unique(kw)) == len(kw), "two slab particles map onto the same single-engine particle"
            err = float(dd.max()) / float(np.abs(wp).max())
            pv = np.concatenate([g["press"] for g in allgot])[alive]; pw = st.press[:s1.numParticles][kw]
            perr = float(np.abs(pv - pw).max()) / max(float(np.abs(pw).max()), 1e-300)
            print("  survivors matched by position: pos %.3e (%d beyond 1e-9), press %.3e" % (err, int((dd > 1e-9 * np.abs(wp).max()).sum()), perr))
            ok = err <= a.rtol and perr <= 1e-7
            FIELDS[:] = []
        else:
            assert len(gid) == case.n and n_alive == case.n, "owned sets do not partition the particles: %d ids, %d alive, %d particles" % (len(gid), n_alive, case.n)
        print("slab parity: %d particles, %d ranks, %d steps; per rank: %s" % (case.n, world, a.steps, allstats))
        for f in FIELDS:
            v = np.concatenate([g[f] for g in allgot]); w = getattr(st, f)
            w = w.reshape(-1, 3)[gid] if v.ndim == 2 else w[gid]
            if v.dtype.kind in "iu":
                # integer fields must be EXACT.  The only excuse is a threshold tie: the slabs sum a particle's neighbours in another
                # order (1e-16 relative), so a comparison sitting on its threshold to rounding may fall either way — each such
                # particle is printed with its margin, anything else fails.
                idx = np.nonzero(v != w)[0]
                print("  %-16s int mismatches %d" % (f, len(idx)))
                for q in idx:
                    g = int(gid[q]); why = None
                    if f == "particleBC":
                        m1 = abs(st.pndSmall[g] - ps.betaPnd) / ps.betaPnd; m2 = abs(st.npcdDeviation2[g] - ps.delta2) / ps.delta2
                        m3 = abs(st.numNeigh[g] - ps.betaNeigh)
                        if min(m1, m2) < 1e-9: why = "threshold tie: |pndSmall - beta n0|/beta n0 = %.1e, |npcd^2 - delta^2|/delta^2 = %.1e" % (m1, m2)
                    elif f == "numNeigh":
                        P3 = st.pos.reshape(-1, 3)[:s1.numParticles]; d = np.linalg.norm(P3 - P3[g], axis=1)
                        near = np.abs(d - ps.reL) < 1e-9 * ps.reL
                        if near.any(): why = "a neighbour sits on the cut-off: | |r| - reL | / reL = %.1e" % float(np.abs(d[near] - ps.reL).min() / ps.reL)
                    print("    gid %d: slabs %d, single engine %d -- %s" % (g, int(v[q]), int(w[q]), why or "UNEXPLAINED"))
                    ok &= why is not None
            else:
                scale = max(float(np.abs(w).max()), 1e-300)
                err = float(np.abs(v - w).max()) / scale
                print("  %-16s max |diff| / max |ref| = %.3e" % (f, err)); ok &= err <= a.rtol
        dv = abs(allstats[0]["velMax"] - s1.velMax) / max(s1.velMax, 1e-300)
        if a.inc:
            print("  CG iterations of the last step: slabs %d, single engine %d" % (allstats[0]["solverIter"], s1.solverIter))
            ok &= abs(allstats[0]["solverIter"] - s1.solverIter) <= 3 and s1.solverIter > 3
        print("  velMax %.12g vs %.12g (rel %.2e); migrants %d, repartitions %d, cuts %s -> %s (%d moves), owned %s" % (allstats[0]["velMax"], s1.velMax, dv, sum(s["migrants_out"] for s in allstats), allstats[0]["repartitions"], list(cuts), allstats[0]["cuts"], allstats[0].get("cut_moves", 0), [s["n_own"] for s in allstats]))
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
