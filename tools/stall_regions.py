#!/usr/bin/env python
"""Where the warps of one kernel spend their time, by code region: joins an ncu source-page CSV
(`ncu -i X.ncu-rep --page source --csv`, per-SASS-instruction stall samples) with the line table of the
same cubin (tools/hotlines.py) and groups the samples into the walk's node loop, the rest of the walk and
everything else (ray generation, attribute fetch, shading, colour pipeline).

    python tools/stall_regions.py <lib.so> <mangled kernel name> <ncu source csv> [out.json]
"""
import csv
import json
import os
import sys
from collections import defaultdict

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import hotlines as H  # noqa: E402

STALLS = ["stall_long_sb", "stall_wait", "stall_no_inst", "stall_not_selected", "stall_selected", "stall_math",
          "stall_dispatch", "stall_branch_resolving", "stall_short_sb", "stall_lg", "stall_mio"]


def region(src, fn_of_line):
    f, line = src
    if f == "sm_32_intrinsics.hpp":  # __ldg: node / triangle record fetches of the walk
        return "walk: node loop + record fetches"
    if f == "trace.cuh":
        return fn_of_line(line)
    return "ray generation + attribute fetch + shading + colour"


def trace_regions(path):
    """Line ranges of trace.cuh: slab_fast / pick / the node loop of walk_fast_body vs the rest."""
    lines = open(path).read().splitlines()
    start = next(i for i, l in enumerate(lines) if "while (!(cur & kLeafFlag))" in l) + 1
    end = next(i for i, l in enumerate(lines) if "if (cur == kDone) break;" in l) + 1
    slab0 = next(i for i, l in enumerate(lines) if "struct FastRay" in l) + 1
    slab1 = next(i for i, l in enumerate(lines) if "template <bool COUNT, bool SPHERES>" in l and i > slab0) + 1

    def fn(line):
        if start <= line <= end or slab0 <= line <= slab1:
            return "walk: node loop + record fetches"
        return "walk: leaves, pops, entry / exit"

    return fn


def main():
    so, kernel, src_csv = sys.argv[1:4]
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    fn = trace_regions(os.path.join(root, "rayca_b200", "csrc", "trace.cuh"))
    table = H.line_table(so, kernel)
    rows = list(csv.reader(open(src_csv)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr, data = rows[hdr_i], rows[hdr_i + 1:]
    if len(table) != len(data):
        raise SystemExit(f"instruction count mismatch: nvdisasm {len(table)}, ncu {len(data)}")
    col = {n: hdr.index(n) for n in STALLS + ["# Samples", "Instructions Executed"]}
    agg = defaultdict(lambda: defaultdict(float))
    for (src, _txt), r in zip(table, data):
        g = region(src, fn)
        agg[g]["samples"] += float(r[col["# Samples"]] or 0)
        agg[g]["executed"] += float(r[col["Instructions Executed"]] or 0)
        agg[g]["sass"] += 1
        for n in STALLS:
            agg[g][n] += float(r[col[n]] or 0)
    ts = sum(a["samples"] for a in agg.values())
    te = sum(a["executed"] for a in agg.values())
    out = {"kernel": kernel, "sass_instructions": len(table), "regions": {}}
    for g, a in sorted(agg.items(), key=lambda kv: -kv[1]["samples"]):
        out["regions"][g] = {"sass_instructions": int(a["sass"]),
                             "warp_time_pct": round(100 * a["samples"] / ts, 1),
                             "executed_instructions_pct": round(100 * a["executed"] / te, 1),
                             "stall_pct_of_region": {n[6:]: round(100 * a[n] / max(a["samples"], 1), 1) for n in STALLS}}
    txt = json.dumps(out, indent=1)
    print(txt)
    if len(sys.argv) > 4:
        open(sys.argv[4], "w").write(txt + "\n")


if __name__ == "__main__":
    main()
