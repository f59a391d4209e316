"""Time pmps_write_vtu on the benchmark block (default output switches of the block's input: pnd on) with all host cores and with one."""
import json, os, sys, time, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from polymps_b200.engine import Engine
k = int(sys.argv[1]) if len(sys.argv) > 1 else 171
if True:
    from polymps_b200 import cases
    from polymps_b200.mesh import PolygonMesh
    from polymps_b200.params import ParticleSystem
    from polymps_b200.state import HostState
    case = cases.dam3d_block(k, flow=True)
    ps = ParticleSystem.from_json(case.options).set_parameters(); ps.allocate_buckets(case.pos)
    meshes = [PolygonMesh.from_triangles(case.tris).init_wall_tables(int(ps.dim), ps.weightType, ps.partDist, ps.reL, ps.reS)]
    st = HostState.from_case(case, ps); n = st.numParticles
    eng = Engine(ps, meshes); eng.remember_capacity(st.numParticlesZero); eng.upload(st); eng.init_step0(); eng.step(2)
root = tempfile.mkdtemp(prefix="pmps_vtu_t_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)   # (memory-backed: the formatter, not the disk)
out = {"particles": n, "flags": "pnd"}
for threads in (0, 1):
    if threads: os.environ["PMPS_VTU_THREADS"] = str(threads)
    path = os.path.join(root, "t%d.vtu" % threads)
    t = time.time(); eng.write_vtu(path, Engine.VTU_PND); dt = time.time() - t
    out["threads_%s" % (threads or "all")] = {"seconds": round(dt, 3), "MB": round(os.path.getsize(path) / 1e6, 1), "MB_per_s": round(os.path.getsize(path) / 1e6 / dt, 1)}
    os.remove(path)
out["host_cores"] = os.cpu_count()
import shutil; shutil.rmtree(root, ignore_errors=True)
print(json.dumps(out))
