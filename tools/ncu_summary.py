#!/usr/bin/env python
"""Prints (and optionally writes as JSON) the judged subset of one `ncu --set full` report.

    python tools/ncu_summary.py <report.ncu-rep> [out.json] [note]
"""
import csv
import json
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "l1tex__t_sector_hit_rate.pct", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_bytes.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed"]


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    res = []
    for vals in rows[2:]:
        d = {"Kernel Name": vals[hdr.index("Kernel Name")]}
        for i, h in enumerate(hdr):
            if h in WANT or ("warps_issue_stalled" in h and h.endswith("per_issue_active.ratio") and "not_issued" not in h):
                d[h] = f"{vals[i]} {units[i]}".strip()
        res.append(d)
    doc = {"report": rep.split("/")[-1], "note": sys.argv[3] if len(sys.argv) > 3 else "", "launches": res}
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    try:  # dram__bytes_read.sum + dram__bytes_write.sum of the first launch: bench.py's roofline.traffic
        first = res[0]
        doc["dram_bytes_per_launch"] = int(sum(float(first[k].split()[0]) * scale[first[k].split()[1]]
                                               for k in ("dram__bytes_read.sum", "dram__bytes_write.sum")))
    except Exception:
        pass
    print(json.dumps(doc, indent=1))
    if len(sys.argv) > 2 and sys.argv[2] != "-":
        json.dump(doc, open(sys.argv[2], "w"), indent=1)


if __name__ == "__main__":
    main()
