#!/usr/bin/env python
"""bench.py — particle-steps/s of the 3D weakly-compressible dam break (BASELINE.json metric) on N B200s.

    python bench.py --gpus N --steps K --warmup W            # our CUDA path (one process per GPU under torchrun for N>1)
    python bench.py --impl reference --steps K --warmup W    # the reference's own CPU code on the host cores

A "step" is one pass of the per-timestep hot path (reference src/main.cpp:256-456) over the synthetic dam-break block of
SURVEY.md §8d (InputDamWC option set: 3D, polygon walls, WC-MPS, re 2.1/2.1, gradient type 3, shifting 2, diffusive PND,
Lennard-Jones wall repulsion, NPCD free surface, PC collisions, no-slip).  One JSON line is printed by rank 0.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "particle-steps/sec (3D WC dam break)"
UNIT = "particle-steps/s"
# Algorithmic bytes per particle per launch (FP64 SoA, every reference array a kernel must read or write counted once;
# the candidate list is NOT algorithmic traffic).  DESIGN.md §6 derives each line; SURVEY.md §8d gives the step total.
ALGO_BYTES = {
    # fused groups of the whole-step path (names = LAUNCH names in polymps_b200/csrc/pmps_stages.inc)
    "F1.viscosity+predict": 288, "F2.wall+collision": 225, "F2.collision": 104, "F3.pnd+diffusive+wall": 217, "F3.pnd+diffusive": 217,
    "F3.pnd+wall": 160, "mean_pnd_surface": 72, "E4.bc+eos": 48, "F5.gradient+wall+correct": 313, "F5.gradient+correct": 313,
    "shift.concentration": 64, "shift.gradient": 113, "shift.wall_gradient": 0, "sort.permute8": 308,   # G0: the gather of every plane into the new bucket order
    "F5.gradient": 165, "W5.wall+correct": 148,
    # incompressible scheme (--scheme inc; SURVEY.md 8d "Same, INC"): ELL row of A^T 40 x 12 B + count/diag + x, y
    "ppe.spmv": 500, "ppe.assemble": 580, "ppe.update_xr": 56, "ppe.update_p": 32, "ppe.residual0": 516,
    # unfused stage kernels (PMPS_STAGED=1 / stage API)
    "viscosity": 116, "collision": 104, "pnd_neigh_npcd": 160, "pnd_diffusive": 124, "press_gradient": 96, "wall.closest_point": 121,
    "predict": 172, "correct": 148, "update_bc": 48,
}
TRAFFIC_JSON = os.path.join(ROOT, "profiles", "r02_traffic.json")
STEP_ALGO_BYTES = 1732   # sum over the fused groups G0..G7 + E1..E5 of the 3D WC polygon option set


def build_id():
    """sha256 over the kernel sources the library is built from: an ncu capture applies to a run only when both carry the same id."""
    import hashlib
    d = os.path.join(ROOT, "polymps_b200", "csrc"); h = hashlib.sha256()
    for f in sorted(os.listdir(d)):
        if f.endswith((".cu", ".cuh", ".inc", ".h")) or f == "Makefile":
            h.update(f.encode()); h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:16]


def ncu_capture(kernel, n_particles):
    """(traffic bytes per launch, fp64 pipe active %) of `kernel` from profiles/r02_traffic.json — an `ncu --set full` capture —
    when it was taken on THIS build (same build_id) at this problem size; (None, None) otherwise: a citation is not a measurement."""
    try:
        d = json.load(open(TRAFFIC_JSON))
    except Exception:
        return None, None
    if d.get("build_id") != build_id() or abs(d.get("particles", 0) - n_particles) > 0.02 * n_particles:
        return None, None
    e = d.get("kernels", {}).get(kernel)
    return (e.get("dram_bytes"), e.get("fp64_active_pct")) if e else (None, None)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        try:
            d = json.load(open(p))
            return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)", float(d.get("sm_max_mhz", 1965.0))
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        self.index = index; self.rows = []; self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True); self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15); self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if f[3 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons), "samples": len(sm)}


def block_case(k, jitter=0.0, inc=False, flow=True):
    from polymps_b200 import cases
    return cases.dam3d_block(k, flow=flow, jitter=jitter, name="block%d" % k, incompressible=inc)


STATE_NOTE = {True: "MOVING initial state: vortex cells of ~40 lo, peak speed = expectMaxVelocity 2 sqrt(gH) = 0.086 lo per step (cases.flow_field), both arms",
              False: "column at rest (hydrostatic start)"}


# ------------------------------------------------------------------------------------------------ CPU arm
def run_reference_cpu(k, steps, warmup, threads, flow=True):
    """Times the reference's own CPU implementation on a bounded sample (block of 2k*k*k particles) of the workload.
    kind "reference": oracle/_ref/ref_driver = the unmodified reference sources driven through main.cpp's call order;
    kind "port": the C oracle when the reference binary is not present."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import refrun
    case = block_case(k, flow=flow)
    n = case.n
    if refrun.have_ref():
        from polymps_b200 import cases
        root = tempfile.mkdtemp(prefix="pmps_ref_")
        try:
            cases.write_case(case, root)
            # ref_driver times its own step loop; run warmup+steps and a warmup-only pass to subtract
            out = refrun.run_reference(root, case.name, warmup + steps, os.path.join(root, "d"), threads=threads, timing=True)
            tot = float(out.split("loop_seconds=")[1].split()[0])
            t_w = 0.0
            if warmup > 0:
                out_w = refrun.run_reference(root, case.name, warmup, os.path.join(root, "d"), threads=threads, timing=True)
                t_w = float(out_w.split("loop_seconds=")[1].split()[0])
            sec = max(tot - t_w, 1e-9)
        finally:
            shutil.rmtree(root, ignore_errors=True)
        kind = "reference"
    else:
        import pyoracle
        from polymps_b200.params import ParticleSystem
        from polymps_b200.mesh import PolygonMesh
        from polymps_b200.state import HostState
        ps = ParticleSystem.from_json(case.options).set_parameters(); ps.allocate_buckets(case.pos)
        meshes = [PolygonMesh.from_triangles(case.tris).init_wall_tables(int(ps.dim), ps.weightType, ps.partDist, ps.reL, ps.reS)]
        st = HostState.from_case(case, ps)
        orc = pyoracle.Oracle(ps, st, meshes, shifting_jacobi=True, threads=threads)
        orc.init_step0(); orc.step(warmup)
        t0 = time.time(); orc.step(steps); sec = time.time() - t0
        kind = "port"
    return {"value": n * steps / sec, "unit": UNIT, "cores": threads, "kind": kind,
            "sample": "%d steps (after %d warm-up) of a %dx%dx%d-particle dam-break block (%d particles), same option set, %s" % (steps, warmup, 2 * k, k, k, n, "moving initial state" if flow else "at rest")}, sec


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    flow = args.state == "flow"
    cb, sec = run_reference_cpu(args.ref_k, args.steps, args.warmup, threads, flow=flow)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * sec / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": {"workload": "synthetic 3D WC dam-break block %dx%dx%d = %d particles per GPU, STL polygon tank (10 triangles), InputDamWC option set" % (2 * args.k, args.k, args.k, 2 * args.k ** 3),
                                            "state": STATE_NOTE[flow],
                                            "cpu_sample": "each step runs a %d-particle block of the same recipe (CPU throughput is size-insensitive, SURVEY 8d; BASELINE.md section 3: k = 80)" % (2 * args.ref_k ** 3)},
            "cpu_baseline": cb, "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ GPU arm
E2E_UP = ["pos", "vel", "press", "pndi", "particleBC", "mirrorParticlePos"]      # state a host-driven step must send
E2E_DOWN = ["pos", "vel", "press", "pndi", "particleBC", "mirrorParticlePos"]    # ... and read back (reference layout)


def field_bytes(st, names):
    return int(sum(getattr(st, f).nbytes for f in names))


def roofline_from_profile(prof, profile_steps, n_particles, ms_step, traffic):
    """prof = [(launch name, total ms, launches)] of `profile_steps` profiled steps of ONE rank holding n_particles."""
    prof_ms = {nm: t_ms / profile_steps for nm, t_ms, _ in prof}
    step_prof_ms = sum(prof_ms.values())
    # dominant kernel = the neighbour-pass (or sort/elementwise) launch group with the largest device time among those that carry
    # algorithmic traffic; the candidate filters (list.hits / list.tile_build) move no reference array: they are overhead
    # of this design and count only in the step-level figure below
    top = max([r for r in prof if ALGO_BYTES.get(r[0], 0) > 0] or prof, key=lambda r: r[1])
    top_name, top_avg_ms = top[0], top[1] / max(top[2], 1)
    peak, peak_src, _ = measured_peaks()
    algo = ALGO_BYTES.get(top_name, 0) * n_particles
    achieved = algo / (top_avg_ms * 1e-3) / 1e9
    cap_traffic, fp64 = ncu_capture(top_name, n_particles)
    if traffic is None:
        traffic = cap_traffic
    return {"bound": "hbm", "kernel": top_name, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": traffic, "fp64_active": None if fp64 is None else fp64 / 100.0,
            "traffic_source": ("profiles/r02_traffic.json (ncu --set full of this build, id %s)" % build_id()) if cap_traffic is not None else None,
            "peak_source": peak_src, "kernel_ms": top_avg_ms, "kernel_share_of_step": top[1] / profile_steps / max(step_prof_ms, 1e-9),
            "algorithmic_bytes_per_particle": ALGO_BYTES.get(top_name, 0),
            "step": {"algorithmic_bytes_per_particle_step": STEP_ALGO_BYTES, "achieved_GBps": STEP_ALGO_BYTES * n_particles / (ms_step * 1e-3) / 1e9,
                     "frac": STEP_ALGO_BYTES * n_particles / (ms_step * 1e-3) / 1e9 / peak},
            "kernels_ms_per_step": {k: round(v, 4) for k, v in sorted(prof_ms.items(), key=lambda kv: -kv[1])},
            "kernels_frac_of_hbm_peak": {k: round(ALGO_BYTES[k] * n_particles / (t_ms / max(c, 1) * 1e-3) / 1e9 / peak, 4) for k, t_ms, c in prof if ALGO_BYTES.get(k, 0) > 0 and t_ms > 0}}


def main_gpu_slab(args, rank, world, local):
    """N > 1: ONE dam-break block of N x (the single-GPU particle count), cut into x-slabs, one rank per GPU (polymps_b200/slab.py):
    halo refresh after every phase of the step and migration at bucket rebuilds over NCCL send/recv, three scalars all-reduced per step."""
    import torch
    import torch.distributed as dist
    from polymps_b200 import cases
    from polymps_b200.mesh import PolygonMesh
    from polymps_b200.params import ParticleSystem
    from polymps_b200.slab import SlabEngine, choose_cuts, TABLE_FIELDS, field_width, field_dtype

    # weak scaling (default): N x the single-GPU particle count; --total-k fixes the WHOLE block instead (strong scaling: the named
    # sizes of BASELINE.json configs[4], k = 292 -> 49.8 M, 464 -> 199.8 M particles over the N ranks)
    kN = args.total_k if args.total_k > 0 else int(round(args.k * world ** (1.0 / 3.0)))
    case = cases.dam3d_block_options(kN, incompressible=args.scheme == "inc")
    ps = ParticleSystem.from_json(case.options).set_parameters(); ps.allocate_buckets(None)
    meshes = [PolygonMesh.from_triangles(case.tris).init_wall_tables(int(ps.dim), ps.weightType, ps.partDist, ps.reL, ps.reS)]
    nx, ny, nz = 2 * kN, kN, kN; lo = ps.partDist
    xi = (np.arange(nx) + 0.5) * lo
    col = ((xi - ps.domainMinX) * ps.invBucketSide).astype(np.int64)
    cuts = choose_cuts(np.bincount(col, minlength=ps.numBucketsX) * (ny * nz), world)
    cuts_x = [ps.domainMinX + c * ps.bucketSide for c in cuts]
    mine = np.nonzero((xi >= cuts_x[rank]) & (xi < cuts_x[rank + 1]))[0]
    flow = args.state == "flow"
    loc = cases.dam3d_block_columns(kN, int(mine[0]), int(mine[-1]) + 1, lo=lo, flow=flow)
    halo_est = int(2 * 3.6 * ny * nz)
    se = SlabEngine(ps, meshes, loc, cuts_x, rank, world, device=local, capacity=len(loc["ptype"]) + 6 * halo_est + 200000, cuts=cuts)
    del loc
    eng = se.eng; stream = se.stream
    se.init_step0(); se.step(args.warmup); eng.synchronize()

    def barrier():
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()

    sampler = ClockSampler(local)
    l0 = eng.launch_count(); st0 = dict(se.dom.stats); c0 = eng.counters()
    barrier(); sampler.start()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(stream); se.step(args.steps); e1.record(stream)
    eng.synchronize(); barrier()
    clocks = sampler.stop()
    ms_local = e0.elapsed_time(e1)
    launches = eng.launch_count() - l0; c1 = eng.counters()
    st1 = dict(se.dom.stats)
    n_own = se.n_own
    t = torch.tensor([ms_local], dtype=torch.float64, device="cuda"); cnt = torch.tensor([float(n_own)], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX); dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    ms = float(t.item()); total = float(cnt.item())
    value = total * args.steps / (ms * 1e-3)
    sc = eng.scalars()

    eng.profile(True); se.step(args.profile_steps); prof = eng.profile_get(); eng.profile(False)
    roofline = roofline_from_profile(prof, args.profile_steps, se.dom.n_live, ms / args.steps, args.traffic)
    roofline["note"] = "rank 0's kernels; a rank's launches run over its owned particles + halo copies (%d + %d)" % (se.n_own, se.dom.n_live - se.n_own)

    # end to end with HOST buffers: every step sends the carried state of the rank's table from pinned host memory, steps, reads it back
    e2e_fields = [f for f in E2E_UP]
    host = {f: torch.empty((se.backend.n0, 3) if field_width(f) == 3 else (se.backend.n0,), dtype=field_dtype(f)).pin_memory() for f in e2e_fields}
    e2e_steps = max(1, args.e2e_steps)

    def roundtrip(first):
        import time
        tm = os.environ.get("PMPS_SLAB_TIMING") is not None and rank == 0
        def tick(label, _t=[None]):
            if not tm:
                return
            torch.cuda.synchronize(); now = time.perf_counter()
            if _t[0] is not None:
                print("[e2e] %-10s %.1f ms" % (label, 1e3 * (now - _t[0])), flush=True)
            _t[0] = now
        tick("start")
        with torch.cuda.stream(stream):
            if first:
                T = se.backend.download_table(e2e_fields)
                for f in e2e_fields:
                    host[f].copy_(T[f], non_blocking=True)
                eng.synchronize()
            T = {f: host[f].to(se.dev, non_blocking=True) for f in e2e_fields}
            tick("h2d")
            se.backend.upload_table(T, se.dom.n_live, se.dom.n_own, same_rows=True)   # the rows the engine handed out: no re-partition, no re-sort unless somebody moved too far
        tick("upload")
        se.step(1)
        tick("step")
        with torch.cuda.stream(stream):
            T = se.backend.download_table(e2e_fields)
            tick("download")
            for f in e2e_fields:
                host[f].copy_(T[f], non_blocking=True)
            eng.synchronize()
        tick("d2h")
    roundtrip(True)
    barrier(); e0.record(stream)
    for _ in range(e2e_steps):
        roundtrip(False)
    e1.record(stream); eng.synchronize(); barrier()
    te = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda"); dist.all_reduce(te, op=dist.ReduceOp.MAX)
    nb = torch.tensor([float(sum(host[f].numel() * host[f].element_size() for f in e2e_fields))], dtype=torch.float64, device="cuda"); dist.all_reduce(nb)
    e2e = {"value": total * e2e_steps / (float(te.item()) * 1e-3), "unit": UNIT, "steps": e2e_steps, "h2d_bytes_per_step": int(nb.item()), "d2h_bytes_per_step": int(nb.item()),
           "api": "per rank and step: pinned host table -> pmps_upload_values_device (same rows: partition, bucket order and lists survive, the displacement trackers decide), the phased step with halo refreshes and mid-step re-partitions, pmps_download_device -> pinned host"}
    per_rank = [None] * world
    dist.all_gather_object(per_rank, {"owned": n_own, "halo": se.dom.n_live - n_own, "ms_per_step": ms_local / args.steps,
                                      "repartitions": st1["repartitions"] - st0["repartitions"], "migrants_out": st1["migrants_out"] - st0["migrants_out"],
                                      "bucket_rebuilds": c1["bucket_rebuilds"] - c0["bucket_rebuilds"], "hit_refilters": c1["hit_refilters"] - c0["hit_refilters"],
                                      "halo_refresh_MB_per_step": (st1["refresh_bytes"] - st0["refresh_bytes"]) / args.steps / 1e6})
    inc = args.scheme == "inc"
    line = {"metric": METRIC if not inc else "particle-steps/sec (3D INC dam break: PPE assembly + Jacobi-CG)", "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong" if args.total_k > 0 else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "synthetic 3D %s dam-break block %dx%dx%d = %d particles over %d GPUs (%.2fM per GPU), STL polygon tank (10 triangles), %s option set" % ("INC" if inc else "WC", nx, ny, nz, int(total), world, total / world / 1e6, "InputDamINC" if inc else "InputDamWC"),
                       "state": STATE_NOTE[flow], "cg_iterations_last_step": sc.solverIter if inc else None,
                       "bucket_rebuilds": c1["bucket_rebuilds"] - c0["bucket_rebuilds"], "hit_refilters": c1["hit_refilters"] - c0["hit_refilters"],
                       "particles_total": int(total), "build_id": build_id(), "parallelism": "%d x-slabs cut on bucket columns, one rank per GPU; halo refresh after every phase + migration at rebuilds over NCCL send/recv; 3 scalars MAX-all-reduced per step" % world,
                       "l2": "state per GPU (%.1f GB) is far larger than the 126 MB L2; no flush needed" % (se.dom.n_live * 1500 / 1e9),
                       "host_binding": args.cpu_binding, "ranks": per_rank, "max_displacement_over_lo": sc.maxDisplacement / ps.partDist},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline}
    if rank == 0:
        print(json.dumps(line), flush=True)
    torch.cuda.synchronize()
    host.clear(); del roundtrip   # pinned buffers were used on the engine's stream: release them before the stream goes away
    import gc; gc.collect()
    se.close()
    dist.destroy_process_group()


def bind_to_gpu_numa_node(local):
    """Pins this process to the CPU cores NVML names as local to its GPU, BEFORE any pinned host buffer is allocated: the pages of the
    end-to-end leg's host tables then sit on the NUMA node whose PCIe root the GPU hangs off (first touch), instead of wherever the
    launcher happened to start the rank.  Returns a short description for the JSON line (None when NVML has nothing to say)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = [64 * w + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1]
        cpus = [c for c in cpus if c in os.sched_getaffinity(0)]
        if cpus:
            os.sched_setaffinity(0, cpus)
            return "%d cores local to GPU %d (NVML)" % (len(cpus), local)
    except Exception:
        pass
    return None


def main_gpu(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    args.cpu_binding = bind_to_gpu_numa_node(local) if world > 1 else None   # (one GPU: the CPU baseline leg wants all cores)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU path; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        return main_gpu_slab(args, rank, world, local)
    from polymps_b200.engine import make_engine

    flow = args.state == "flow"
    case = block_case(args.k, inc=args.scheme == "inc", flow=flow)
    ps, meshes, st, eng = make_engine(case, device=local)
    n = st.numParticles
    stream = torch.cuda.ExternalStream(eng.stream, device=torch.device("cuda", local))
    eng.init_step0()
    eng.step(args.warmup)
    eng.synchronize(); eng.scalars()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    l0 = eng.launch_count(); c0 = eng.counters()
    barrier(); sampler.start()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(stream); eng.step(args.steps); e1.record(stream)
    eng.synchronize(); barrier()
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1)
    launches = eng.launch_count() - l0; c1 = eng.counters()
    sc = eng.scalars()
    total_particles = float(sc.numParticles)
    value = total_particles * args.steps / (ms * 1e-3)

    # per-kernel device time (CUDA events around every launch on the step stream; separate pass: it serialises)
    cp0 = eng.counters()
    eng.profile(True); eng.step(args.profile_steps); prof = eng.profile_get(); eng.profile(False)
    cp1 = eng.counters()
    roofline = roofline_from_profile(prof, args.profile_steps, sc.numParticles, ms / args.steps, args.traffic)
    roofline["profile_pass"] = {"steps": args.profile_steps, "bucket_rebuilds": cp1["bucket_rebuilds"] - cp0["bucket_rebuilds"], "hit_refilters": cp1["hit_refilters"] - cp0["hit_refilters"]}

    # end to end through the C ABI with HOST buffers: every step uploads the carried state, steps, and reads it back
    e2e_steps = max(1, args.e2e_steps)
    eng.download(st)   # host copy of the current state
    eng.register_host(st, sorted(set(E2E_UP + E2E_DOWN)))   # explicit opt-in: the library never pins caller memory by itself
    eng.step_host(st, E2E_UP, E2E_DOWN)   # warm: first touch
    ce0 = eng.counters()
    barrier()
    e0.record(stream)
    for _ in range(e2e_steps):
        eng.step_host(st, E2E_UP, E2E_DOWN)
    e1.record(stream); eng.synchronize(); barrier()
    ms_e = e0.elapsed_time(e1)
    ce1 = eng.counters()
    sc2 = eng.scalars()
    e2e = {"value": total_particles * e2e_steps / (ms_e * 1e-3), "unit": UNIT, "steps": e2e_steps,
           "h2d_bytes_per_step": field_bytes(st, E2E_UP), "d2h_bytes_per_step": field_bytes(st, E2E_DOWN),
           "bucket_rebuilds": ce1["bucket_rebuilds"] - ce0["bucket_rebuilds"], "hit_refilters": ce1["hit_refilters"] - ce0["hit_refilters"],
           "api": "pmps_step_host(in, out) per step: upload of the carried state, one step, download — the copies the step does not wait for ride on a second stream; host arrays registered once with pmps_register_host"}
    eng.close()

    # the same block AT REST (round 1's headline regime: no rebuild ever falls into the timed region), for comparison
    at_rest = None
    if flow and args.rest_steps > 0:
        case.vel[:] = 0.0
        ps_r, _, st_r, eng_r = make_engine(case, device=local)
        stream_r = torch.cuda.ExternalStream(eng_r.stream, device=torch.device("cuda", local))
        eng_r.init_step0(); eng_r.step(args.warmup); eng_r.synchronize(); r0 = eng_r.counters()
        barrier(); e0.record(stream_r); eng_r.step(args.rest_steps); e1.record(stream_r); eng_r.synchronize(); barrier()
        r1 = eng_r.counters(); ms_r = e0.elapsed_time(e1)
        at_rest = {"value": float(eng_r.scalars().numParticles) * args.rest_steps / (ms_r * 1e-3), "ms_per_step": ms_r / args.rest_steps, "steps": args.rest_steps,
                   "bucket_rebuilds": r1["bucket_rebuilds"] - r0["bucket_rebuilds"], "hit_refilters": r1["hit_refilters"] - r0["hit_refilters"]}
        eng_r.close(); del st_r

    line = {"metric": METRIC if args.scheme == "wc" else "particle-steps/sec (3D INC dam break: PPE assembly + Jacobi-CG)", "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "synthetic 3D %s dam-break block %dx%dx%d = %d particles per GPU, STL polygon tank (10 triangles), %s option set" % (args.scheme.upper(), 2 * args.k, args.k, args.k, n, "InputDamWC" if args.scheme == "wc" else "InputDamINC"),
                       "state": STATE_NOTE[flow],
                       "bucket_rebuilds": c1["bucket_rebuilds"] - c0["bucket_rebuilds"], "hit_refilters": c1["hit_refilters"] - c0["hit_refilters"],
                       "at_rest_value": None if at_rest is None else at_rest["value"], "at_rest": at_rest,
                       "cg_iterations_last_step": sc2.solverIter if args.scheme == "inc" else None,
                       "particles_per_gpu": n, "parallelism": "1 GPU",
                       "l2": "state per GPU (%.1f GB) is far larger than the 126 MB L2; no flush needed" % (n * 1500 / 1e9),
                       "max_displacement_over_lo": sc2.maxDisplacement / ps.partDist, "max_bucket_count": sc2.maxBucketCount, "vel_max": sc2.velMax,
                       "build_id": build_id()},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline}
    if rank == 0 and world == 1 and not args.no_cpu:
        cb, _ = run_reference_cpu(args.cpu_k, args.cpu_steps, 2, os.cpu_count() or 1, flow=flow)
        if args.cpu1_k > 0:   # the deterministic 1-thread figure next to the all-cores one
            c1t, _ = run_reference_cpu(args.cpu1_k, 2, 1, 1, flow=flow)
            cb["one_thread"] = {"value": c1t["value"], "sample": c1t["sample"]}
        line["cpu_baseline"] = cb
    if rank == 0:
        print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40, help="timed steps (40: several bucket rebuilds / re-partitions fall inside the timed region)")
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scheme", default="wc", choices=["wc", "inc"], help="wc = the BASELINE.json metric (default); inc = the incompressible configuration (#3), single GPU, reported under profiles/")
    ap.add_argument("--k", type=int, default=171, help="block edge: 2k x k x k particles per GPU (171 -> 10.0M)")
    ap.add_argument("--total-k", type=int, default=0, help="N > 1: block edge of the WHOLE block (strong scaling at a named size: 292 -> 49.8M, 464 -> 199.8M particles); 0 = weak scaling from --k")
    ap.add_argument("--state", default="flow", choices=["flow", "rest"], help="flow (default): the moving initial state both arms start from; rest: the column at rest")
    ap.add_argument("--rest-steps", type=int, default=20, help="timed steps of the at-rest comparison run (config.at_rest_value); 0 skips it")
    ap.add_argument("--ref-k", type=int, default=80, help="block edge of the CPU sample for --impl reference (80 -> 1.02 M particles, BASELINE.md section 3)")
    ap.add_argument("--cpu-k", type=int, default=80, help="block edge of the cpu_baseline sample")
    ap.add_argument("--cpu-steps", type=int, default=10)
    ap.add_argument("--cpu1-k", type=int, default=40, help="block edge of the 1-thread CPU figure (0 skips it)")
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--profile-steps", type=int, default=6, help="steps of the per-kernel timing pass (serialised by its events; long enough to see the rebuild / refilter mix of the timed region)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--traffic", type=float, default=None, help="dram bytes per launch of the top kernel from an ncu --set full capture (profiles/)")
    args = ap.parse_args()
    if args.impl == "reference":
        main_reference(args)
    else:
        main_gpu(args)


if __name__ == "__main__":
    main()
