#!/usr/bin/env python
"""Turn an .ncu-rep (ncu --set full) into the small CSV kept under profiles/:  python profiles/summarise.py in.ncu-rep out.csv"""
import csv
import subprocess
import sys

KEEP = ["gpu__time_duration", "dram__bytes", "dram__throughput", "gpu__dram_throughput", "issue_active", "inst_executed_pipe",
        "pipe_alu", "pipe_fma", "smsp__inst_executed.sum", "launch__", "warps_active", "issue_stalled", "sm__cycles_active.avg",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_bytes.sum", "sm__throughput"]
DROP = [".max", ".min", "tensor", "fp64", "tmem", "tma", "ipa", "tex", "workid"]


def main(src, dst):
    raw = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["metric", "unit"] + ["launch %d: %s" % (i, r[hdr.index("Kernel Name")]) for i, r in enumerate(rows[2:])])
        for j, h in enumerate(hdr):
            if any(k in h for k in KEEP) and not any(x in h for x in DROP):
                w.writerow([h, units[j]] + [r[j] for r in rows[2:]])


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
