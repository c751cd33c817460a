#!/bin/bash
# round 2, call 22 (1 GPU): k = 31 bucket sort launched per bucket-size class (occupancy 4 -> 6 / 8 CTAs per SM for the small classes)
mkdir -p gpurun_out
B="python bench.py --k 31 --steps 3 --warmup 3 --no-cpu-baseline --no-h2d-ceiling"
for c in 1 2 3; do
PDG_ML_CLASSES=$c timeout 600 $B > gpurun_out/r2_22_c$c.log 2> gpurun_out/r2_22_c$c.err; echo "exit $?" >> gpurun_out/r2_22_c$c.log
done
PDG_ML_CLASSES=3 timeout 900 python -m pytest tests -m gpu -x -q -k "long or k31 or scale or msdlong or 17 or 31" > gpurun_out/r2_22_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_22_pytest.log
tail -n 3 gpurun_out/r2_22_pytest.log | cut -c1-300
python - <<PY
import json
for c in "123":
    f = "c" + c
    for l in open("gpurun_out/r2_22_%s.log" % f):
        if l.startswith("{"):
            d = json.loads(l)
            k = d["roofline"]["all_kernels"][0]
            print("%-10s ms/step %.3f | e2e %.3f | parity %s | %s %.3f ms | ms_sort %.3f" % (f, d["ms_per_step"], d["e2e"]["ms_per_step"], d["result"]["parity"], k["kernel"][:22], k["ms_per_launch"], d["phases_ms"]["ms_sort"]))
        elif "exit" in l and "exit 0" not in l: print(f, l.strip()[:200])
PY
