#!/usr/bin/env python
"""Per-source-line profile of one kernel: joins an ncu source-page CSV (per-SASS-instruction samples and
executed counts, `ncu -i X.ncu-rep --page source --csv`) with `nvdisasm --print-line-info` of the same
cubin (instruction order is identical), and prints the hottest file:line entries.

    python tools/hotlines.py <lib.so> <mangled kernel name> <ncu source csv> [top N]
"""
import csv
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict


def line_table(so, kernel):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, capture_output=True)
    cub = [f for f in os.listdir(tmp) if f.startswith("rayca_cuda.") and f.endswith(".cubin")][0]
    txt = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(tmp, cub)], capture_output=True, text=True).stdout
    out, cur, on = [], ("?", 0), False
    for ln in txt.splitlines():
        if ln.startswith(".text."):
            on = ln.strip() == f".text.{kernel}:"
            continue
        if not on:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            if "inlined at" not in ln:
                cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
            out.append((cur, ln.split("*/", 1)[1].strip()))
    return out


def main():
    so, kernel, src_csv = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    lt = line_table(so, kernel)
    rows = list(csv.reader(open(src_csv)))
    hdr, data = rows[1], rows[2:]
    i_s, i_e = hdr.index("# Samples"), hdr.index("Instructions Executed")
    i_t = hdr.index("Thread Instructions Executed")
    print(f"sass instructions: nvdisasm {len(lt)}, ncu {len(data)}")
    n = min(len(lt), len(data))
    agg = defaultdict(lambda: [0, 0, 0, 0])
    for k in range(n):
        a = agg[lt[k][0]]
        a[0] += int(data[k][i_s]); a[1] += int(data[k][i_e]); a[2] += int(data[k][i_t]); a[3] += 1
    ts = sum(v[0] for v in agg.values()); te = sum(v[1] for v in agg.values())
    byfile = defaultdict(lambda: [0, 0])
    for (f, l), v in agg.items():
        byfile[f][0] += v[0]; byfile[f][1] += v[1]
    for f, v in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
        print(f"{f:14s} samples {100*v[0]/ts:5.1f}%  executed {100*v[1]/te:5.1f}%")
    print("file:line        samples%  exec%  simt  sass")
    for (f, l), v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{f}:{l:<5d} {100*v[0]/ts:7.2f} {100*v[1]/te:7.2f} {v[2]/max(v[1],1):5.1f} {v[3]:5d}")


if __name__ == "__main__":
    main()
