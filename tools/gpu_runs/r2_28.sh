#!/bin/bash
# round 2, call 28 (1 GPU): rows kernel with 8 lanes per bucket -- long-k tests, bench k = 31, launch list
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "long or k31 or scale or msdlong or 17 or 31" > gpurun_out/r2_28_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_28_pytest.log
tail -n 3 gpurun_out/r2_28_pytest.log | cut -c1-300
timeout 600 python bench.py --k 31 --steps 3 --warmup 3 --no-cpu-baseline --no-h2d-ceiling > gpurun_out/r2_28_k31.log 2> gpurun_out/r2_28_k31.err; echo "exit $?" >> gpurun_out/r2_28_k31.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_k31_final.csv python tools/profile_step.py 100 31 > gpurun_out/r2_28_ncu.log 2>&1
python - <<PY
import json, csv
for l in open("gpurun_out/r2_28_k31.log"):
    if l.startswith("{"):
        d = json.loads(l)
        k = d["roofline"]["all_kernels"][0]
        print("k31 ms/step %.3f | e2e %.3f | parity %s | %s %.3f ms frac %.3f | ms_sort %.3f" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["result"]["parity"], k["kernel"][:22], k["ms_per_launch"], k["frac"], d["phases_ms"]["ms_sort"]))
    elif "exit" in l and "exit 0" not in l: print(l.strip()[:200])
rows = [r for r in csv.reader(open("gpurun_out/r02_launches_k31_final.csv")) if len(r) > 5]
col = {h: i for i, h in enumerate(rows[0])}
d = rows[1:]
for r in d[len(d) // 2:]:
    if "ml_" in r[col["Kernel Name"]]: print("%-70s %8.3f" % (r[col["Kernel Name"]][:70], float(r[col["Metric Value"]].replace(",", "")) / 1e6))
PY
