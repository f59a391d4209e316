#!/usr/bin/env python
"""profiles/r02_traffic.json from an `ncu --set full` capture of bench.py at the benchmark size: per launch group of the step the DRAM
bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) and the FP64-pipe / issue utilisation, keyed by the hash of the
kernel sources the capture ran (bench.build_id) — bench.py quotes `roofline.traffic` / `fp64_active` only when that hash matches.

    python profiles/make_traffic.py gpurun_out/prof.ncu-rep <particles> [out.json]
"""
import csv, io, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
NAMES = [   # substring of the demangled kernel name -> LAUNCH name of pmps_stages.inc (first match wins)
    ("BViscosity, ", "F1.viscosity+predict"), ("BPndNeighNpcd, ", "F3.pnd+diffusive"), ("BWallNpcd", "W3a.wall_npcd"),
    ("BWallPndDiffusiveT<0>", "W3b.wall_pnd_diffusive"), ("BMeanPnd<0>", "mean_pnd_surface"), ("BWallPressGradientT<0>, ", "W5.wall+correct"),
    ("BPressGradientT<0>>", "F5.gradient"), ("BWallGradConcentration", "shift.wall_gradient"), ("BGradConcentration", "shift.gradient"),
    ("BConcentration", "shift.concentration"), ("k_hl_build", "list.build"), ("k_collide", "F2.collision"), ("k_wall_closest", "wall.closest_point"),
]


def main():
    rep, particles = sys.argv[1], int(sys.argv[2])
    out = sys.argv[3] if len(sys.argv) > 3 else os.path.join(ROOT, "profiles", "r02_traffic.json")
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--kernel-name-base", "demangled"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt))); hdr = rows[0]
    col = {k: hdr.index(k) for k in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum",
                                     "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active")}
    units = rows[1]
    def scale(u):
        return {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1.0)
    acc = {}
    for r in rows[2:]:
        name = r[col["Kernel Name"]]
        label = next((l for s, l in NAMES if s in name), None)
        f = lambda k: float(r[col[k]].replace(",", "") or 0)
        ms = f("gpu__time_duration.sum") * {"ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3}.get(units[col["gpu__time_duration.sum"]], 1.0)
        if label is None or ms < 0.05:   # (launches that return at once: a list check that found the list still valid)
            continue
        e = acc.setdefault(label, {"launches": 0, "dram_bytes": 0.0, "fp64": 0.0, "issue": 0.0, "ms": 0.0})
        e["launches"] += 1; e["ms"] += ms
        e["dram_bytes"] += f("dram__bytes_read.sum") * scale(units[col["dram__bytes_read.sum"]]) + f("dram__bytes_write.sum") * scale(units[col["dram__bytes_write.sum"]])
        e["fp64"] += f("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"); e["issue"] += f("smsp__issue_active.avg.pct_of_peak_sustained_active")
    import bench
    doc = {"build_id": bench.build_id(), "particles": particles, "source": os.path.basename(rep),
           "note": "ncu --set full --clock-control none: per-launch averages; durations under ncu are cold-cache and serialised",
           "kernels": {k: {"dram_bytes": v["dram_bytes"] / v["launches"], "fp64_active_pct": v["fp64"] / v["launches"], "issue_active_pct": v["issue"] / v["launches"],
                           "ms_under_ncu": v["ms"] / v["launches"], "launches": v["launches"]} for k, v in acc.items()}}
    json.dump(doc, open(out, "w"), indent=1)
    for k, v in doc["kernels"].items():
        print("%-26s %7.2f GB  fp64 %5.1f %%  issue %5.1f %%  %.2f ms x %d" % (k, v["dram_bytes"] / 1e9, v["fp64_active_pct"], v["issue_active_pct"], v["ms_under_ncu"], v["launches"]))


if __name__ == "__main__":
    main()
