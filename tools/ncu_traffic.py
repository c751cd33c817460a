#!/usr/bin/env python
"""Writes profiles/traffic.json from an `ncu --set full` report of tools/profile_step.py: DRAM bytes
(dram__bytes_read.sum + dram__bytes_write.sum) per unit of work for the kernels bench.py reports.
usage: python tools/ncu_traffic.py <report.ncu-rep> <units: k-mer records> <units: FASTA bytes> <out.json>"""
import csv
import json
import subprocess
import sys


def to_bytes(v, unit):
    v = float(v.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]


def main():
    rep, n_rec, n_bytes, out_path = sys.argv[1], float(sys.argv[2]), float(sys.argv[3]), sys.argv[4]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    col = {h: i for i, h in enumerate(rows[0])}
    units = rows[1]
    names = {"m3_pass1_kernel": n_rec, "m3_pass2_kernel": n_rec, "m3_pass3_kernel": n_rec, "m3_hist_kernel": n_rec,
             "fb_summ_kernel": n_bytes, "fb_emit_kernel": n_bytes}
    res, detail = {}, {}
    for r in rows[2:]:
        kn = r[col["Kernel Name"]]
        for key, u in names.items():
            if key in kn and key not in res:
                rd = to_bytes(r[col["dram__bytes_read.sum"]], units[col["dram__bytes_read.sum"]])
                wr = to_bytes(r[col["dram__bytes_write.sum"]], units[col["dram__bytes_write.sum"]])
                res[key] = (rd + wr) / u
                detail[key] = {"dram_read_bytes": rd, "dram_write_bytes": wr, "units": u,
                               "gpu_time_us": r[col["gpu__time_duration.sum"]]}
    json.dump({"source": rep, "how": "ncu --set full --clock-control none, one launch each (the second, warm build of "
               "tools/profile_step.py): %.4g k-mer records, %.4g FASTA bytes" % (n_rec, n_bytes),
               "dram_bytes_per_unit": res, "launches": detail},
              open(out_path, "w"), indent=1)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
