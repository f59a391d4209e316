"""The benchmark block (10.0 M particles, moving initial state) over several hundred steps: no guard violation, particle count kept,
front position / extent history from the device reductions, rebuild and refilter counters, ms per step over the whole run."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from polymps_b200 import cases
from polymps_b200.engine import Engine
from polymps_b200.mesh import PolygonMesh
from polymps_b200.params import ParticleSystem
from polymps_b200.state import HostState
k = int(sys.argv[1]) if len(sys.argv) > 1 else 171; steps = int(sys.argv[2]) if len(sys.argv) > 2 else 400
case = cases.dam3d_block(k, flow=True)
ps = ParticleSystem.from_json(case.options).set_parameters(); ps.allocate_buckets(case.pos)
meshes = [PolygonMesh.from_triangles(case.tris).init_wall_tables(int(ps.dim), ps.weightType, ps.partDist, ps.reL, ps.reS)]
st = HostState.from_case(case, ps); n = st.numParticles
eng = Engine(ps, meshes); eng.remember_capacity(st.numParticlesZero); eng.upload(st); del st
eng.init_step0(); eng.step(5); eng.synchronize()
c0 = eng.counters(); hist = []; t0 = time.time()
for s in range(0, steps, 20):
    eng.step(20)
    mn, mx, cnt = eng.observe_extent(ps.fluid); sc = eng.scalars()
    hist.append({"step": s + 20, "front_x_over_lo": round(float(mx[0]) / ps.partDist, 4), "top_z_over_lo": round(float(mx[2]) / ps.partDist, 4), "fluid": cnt, "velMax": round(sc.velMax, 6), "maxDisp_over_lo": round(sc.maxDisplacement / ps.partDist, 4)})
eng.synchronize(); dt = time.time() - t0; c1 = eng.counters()
print(json.dumps({"particles": n, "steps": steps, "ms_per_step_wall": round(1e3 * dt / steps, 2), "bucket_rebuilds": c1["bucket_rebuilds"] - c0["bucket_rebuilds"], "hit_refilters": c1["hit_refilters"] - c0["hit_refilters"],
                  "guard_violated": c1["guard_violated"], "particles_kept": hist[-1]["fluid"] == n, "history": hist}))
