"""Periodic channel at benchmark size: ms per fused step, wraps per step, re-sorts (the cost of the wrap handling at scale)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from polymps_b200 import cases
from polymps_b200.engine import Engine
from polymps_b200.mesh import PolygonMesh
from polymps_b200.params import ParticleSystem
from polymps_b200.state import HostState
nx, ny, nz = (int(v) for v in (sys.argv[1:4] if len(sys.argv) > 3 else (160, 400, 160)))
case = cases.channel3d_periodic(nx, ny, nz, jitter=0.04, name="perbench", drift=float(sys.argv[4]) if len(sys.argv) > 4 else 1.0)
ps = ParticleSystem.from_json(case.options).set_parameters(); ps.allocate_buckets(case.pos)
meshes = [PolygonMesh.from_triangles(case.tris).init_wall_tables(int(ps.dim), ps.weightType, ps.partDist, ps.reL, ps.reS)]
st = HostState.from_case(case, ps)
eng = Engine(ps, meshes); eng.remember_capacity(st.numParticlesZero); eng.upload(st); eng.init_step0(); eng.step(3)
c0 = eng.counters(); torch.cuda.synchronize(); t = time.time(); K = 10
eng.step(K); eng.synchronize(); dt = time.time() - t; c1 = eng.counters()
print(json.dumps({"particles": case.n, "buckets": [ps.numBucketsX, ps.numBucketsY, ps.numBucketsZ], "ms_per_step": round(1e3 * dt / K, 2),
                  "rebuilds_per_step": (c1["bucket_rebuilds"] - c0["bucket_rebuilds"]) / K, "scalars_ok": eng.scalars().numParticles == case.n}))
