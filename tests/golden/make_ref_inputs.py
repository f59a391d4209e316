"""Packs the reference's own input documents for the BASELINE.json configurations into ONE fixture, tests/golden/ref_inputs.json
(the GPU box has no /root/reference).  Every document is stored as parsed from the reference file — same keys, same values, comments
stripped (the reference parses with ignore_comments, src/MpsInputOutput.cpp:121-127) — so that tests/test_ref_inputs_gpu.py can feed
the reference's OWN option sets, with only `pathNames` redirected to a synthesized grid / STL and the run length shortened, to both
the stock binary and the drop-in.  Run in the build container:   python tests/golden/make_ref_inputs.py"""
import json
import os

HERE = os.path.dirname(os.path.abspath(__file__))
REF_INPUT = "/root/reference/input"
NAMES = ["InputDamWC", "InputDamINC", "dam1610_h300_lo0p0050_WC", "dam1610_h300_lo0p0050_INC", "2D_dam_fraccarollo_lo02p50e-03_WC",
         "sloshing_3D_bulian_lo0p00930", "dam1610_BC", "damErosion3D_WC"]


def strip_comments(t):
    out = []; i = 0; n = len(t); ins = False
    while i < n:
        c = t[i]
        if ins:
            out.append(c)
            if c == "\\":
                out.append(t[i + 1]); i += 2; continue
            if c == '"':
                ins = False
            i += 1; continue
        if c == '"':
            ins = True; out.append(c); i += 1; continue
        if t.startswith("//", i):
            while i < n and t[i] != "\n":
                i += 1
            continue
        if t.startswith("/*", i):
            i = t.index("*/", i) + 2; continue
        out.append(c); i += 1
    return "".join(out)


def main():
    docs = {nm: json.loads(strip_comments(open(os.path.join(REF_INPUT, nm + ".json")).read())) for nm in NAMES}
    with open(os.path.join(HERE, "ref_inputs.json"), "w") as fh:
        json.dump(docs, fh, separators=(",", ":"), sort_keys=True)
    print("%d documents, %d bytes" % (len(docs), os.path.getsize(os.path.join(HERE, "ref_inputs.json"))))


if __name__ == "__main__":
    main()
