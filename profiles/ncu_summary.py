#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed): per-kernel headline metrics, stall mix and — for one kernel —
instruction / stall-sample shares per SOURCE LINE (joins ncu's SASS page with nvdisasm's line table of libpmps.so).

    python profiles/ncu_summary.py gpurun_out/prof.ncu-rep [--lines <substring of the mangled kernel name> [--ncu-name <substring of the demangled name>]] [--top N] [--lib <the libpmps.so the capture ran>]
"""
import collections
import csv
import io
import re
import subprocess
import sys
import os
import glob
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEAD = [
    "gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_adu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_bytes.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__cycles_elapsed.max",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
]


def ncu(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def raw(rep):
    rows = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "raw", "--csv", "--kernel-name-base", "demangled"]))))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print("=====", r[hdr.index("Kernel Name")][:140])
        for k in HEAD:
            if k in hdr:
                print("  %-70s %s %s" % (k, r[hdr.index(k)], units[hdr.index(k)]))
        st = [(float(r[i].replace(",", "") or 0), k) for i, k in enumerate(hdr) if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio")]
        tot = sum(v for v, _ in st) or 1.0
        print("  stall mix: " + ", ".join("%s %.0f%%" % (k[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")], 100 * v / tot) for v, k in sorted(st, reverse=True)[:8]))


def lines(rep, pat, top, ncupat):
    tmp = tempfile.mkdtemp()
    # the walker translation units all come from pmps_walk_g.cu: extracted from the library their cubins overwrite one another, so the
    # object files of the build (one directory each) are taken where they exist
    objs = sorted(glob.glob(os.path.join(ROOT, "polymps_b200", "csrc", "_build", "*.o"))) if "--lib" not in sys.argv else []
    srcs = objs or [sys.argv[sys.argv.index("--lib") + 1] if "--lib" in sys.argv else os.path.join(ROOT, "polymps_b200", "libpmps.so")]
    for k, src in enumerate(srcs):
        d = os.path.join(tmp, "e%03d" % k); os.makedirs(d)
        subprocess.run(["cuobjdump", "-xelf", "all", src], cwd=d, capture_output=True)
    seq = None
    for cub in glob.glob(os.path.join(tmp, "*", "*.cubin")):
        txt = subprocess.run(["nvdisasm", "-g", "-c", cub], capture_output=True, text=True).stdout
        cur, name, acc = None, None, {}
        for ln in txt.splitlines():
            m = re.match(r"\.text\.(\S+):", ln)
            if m:
                name = m.group(1); acc[name] = []; continue
            if name is None:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
            if m:
                cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
            if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
                acc[name].append(cur)
        for nm, s in acc.items():
            if pat in nm and s:
                seq, kname = s, nm
        if seq:
            break
    if not seq:
        print("no kernel matching", pat); return
    out = ncu(["-i", rep, "--page", "source", "--csv", "-k", "regex:.", "--kernel-name-base", "mangled"])
    # several kernels may be in the report: split on the "Kernel Name" rows
    blocks, curb = [], None
    for r in csv.reader(io.StringIO(out)):
        if r and r[0] == "Kernel Name":
            curb = {"name": r[1], "rows": []}; blocks.append(curb); continue
        if curb is not None:
            curb["rows"].append(r)
    def weight(b):   # several launches of the kernel may be in the report (some return at once): take the one that did the most work
        try:
            ie_ = b["rows"][0].index("Instructions Executed"); return sum(int(r[ie_]) for r in b["rows"][1:])
        except Exception:
            return 0
    blocks = sorted([b for b in blocks if ncupat in b["name"]], key=weight, reverse=True)[:1]
    for b in blocks:
        hdr, data = b["rows"][0], b["rows"][1:]
        ie, isamp = hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
        if len(data) != len(seq):
            print("warning: %d SASS rows in the report vs %d in the library (rebuilt since the capture?)" % (len(data), len(seq)))
        agg, st = collections.Counter(), collections.Counter()
        for loc, r in zip(seq, data):
            agg[loc] += int(r[ie]); st[loc] += int(r[isamp])
        ti, ts = sum(agg.values()) or 1, sum(st.values()) or 1
        print("===== %s: %d warp instructions, %d stall samples" % (b["name"][:100], ti, ts))
        for loc, v in sorted(agg.items(), key=lambda kv: -(kv[1] / ti + st[kv[0]] / ts))[:top]:
            print("  %-28s %5.1f%% inst  %5.1f%% samples" % ("%s:%d" % loc if loc else "?", 100 * v / ti, 100 * st[loc] / ts))
        break


if __name__ == "__main__":
    rep = sys.argv[1]
    if "--lines" in sys.argv:
        top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 30
        pat = sys.argv[sys.argv.index("--lines") + 1]
        lines(rep, pat, top, sys.argv[sys.argv.index("--ncu-name") + 1] if "--ncu-name" in sys.argv else pat)
    else:
        raw(rep)
