#!/bin/bash
# round 2, call 31 (1 GPU): launch list of one k = 31 build of 500 genomes (one rank's share of config 4 on 2 GPUs)
mkdir -p gpurun_out
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_k31_500genomes.csv python tools/profile_step.py 500 31 > gpurun_out/r2_31_ncu.log 2>&1
echo "exit $?"
python - <<PY
import csv
rows = [r for r in csv.reader(open("gpurun_out/r02_launches_k31_500genomes.csv")) if len(r) > 5]
col = {h: i for i, h in enumerate(rows[0])}
d = rows[1:]
for r in d[len(d) // 2:]: print("%-70s %8.3f" % (r[col["Kernel Name"]][:70], float(r[col["Metric Value"]].replace(",", "")) / 1e6))
PY
