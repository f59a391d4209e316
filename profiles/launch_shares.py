#!/usr/bin/env python
"""Summarise an ncu launch list (`--metrics gpu__time_duration.sum --csv`): time and share per kernel name.

    python profiles/launch_shares.py gpurun_out/launches.csv > profiles/rNN_launch_shares.txt
Per-launch times under ncu are cold-cache and serialised: compare SHARES with bench.py's kernels_ms_per_step, not absolutes.
"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[0].isdigit()]
tot = collections.OrderedDict()
for r in rows:
    full = r[4]
    m = re.search(r"k_tv<.*?Fused<(.*?)>>\(", full)
    if m:   # a walker pass: name it by its bodies
        name = "k_tv<" + re.sub(r"pmps_set_\w+::", "", m.group(1)) + ">"
    else:
        name = re.sub(r"\(.*", "", full); name = re.sub(r"pmps_set_\w+::", "", name)
        name = re.sub(r"<.*", "<...>", name) if len(name) > 90 else name
    t = tot.setdefault(name, [0, 0.0]); t[0] += 1; t[1] += float(r[14].replace(",", ""))
total = sum(v[1] for v in tot.values())
print("# %d launches, %.3f ms total (whole command incl. upload, 0-th step, warm-up, profile pass, e2e pass)" % (len(rows), total / 1e6))
print("%-72s %8s %12s %7s" % ("kernel", "launches", "total ms", "share"))
for k, v in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print("%-72s %8d %12.3f %6.1f%%" % (k[:72], v[0], v[1] / 1e6, 100 * v[1] / total))
