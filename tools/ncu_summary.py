#!/usr/bin/env python
"""Summarise an .ncu-rep (raw page) per kernel: time, DRAM bytes, occupancy, issue rate, top stalls."""
import csv
import subprocess
import sys


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[0]
    col = {h: i for i, h in enumerate(hdr)}
    want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "launch__registers_per_thread", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct"]
    for r in rows[2:]:
        print("== %s" % r[col["Kernel Name"]][:90])
        for wname in want:
            if wname in col:
                print("   %-62s %s %s" % (wname, r[col[wname]], rows[1][col[wname]]))
        st = []
        for h, i in col.items():
            if "issue_stalled" in h and h.endswith("per_issue_active.ratio"):
                try:
                    st.append((float(r[i].replace(",", "")), h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")))
                except ValueError:
                    pass
        print("   stalls: " + ", ".join("%s=%.2f" % (n, v) for v, n in sorted(st, reverse=True)[:7]))


if __name__ == "__main__":
    main()
