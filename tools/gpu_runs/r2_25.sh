#!/bin/bash
# round 2, call 25 (1 GPU): k > 16 bucket sort writes position column + codes + head bits (rows from the bits) -- tests, bench
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_25_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_25_pytest.log
tail -n 4 gpurun_out/r2_25_pytest.log | cut -c1-300
B="python bench.py --k 31 --steps 3 --warmup 3 --no-cpu-baseline --no-h2d-ceiling"
for c in 3 2 1; do
PDG_ML_CLASSES=$c timeout 600 $B > gpurun_out/r2_25_c$c.log 2> gpurun_out/r2_25_c$c.err; echo "exit $?" >> gpurun_out/r2_25_c$c.log
done
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_25_k11.log 2> gpurun_out/r2_25_k11.err; echo "exit $?" >> gpurun_out/r2_25_k11.log
python - <<PY
import json
for f in ["c3", "c2", "c1", "k11"]:
    for l in open("gpurun_out/r2_25_%s.log" % f):
        if l.startswith("{"):
            d = json.loads(l)
            k = d["roofline"]["all_kernels"][0]
            print("%-6s ms/step %.3f | e2e %.3f | parity %s | %s %.3f ms frac %.3f | ms_sort %.3f" % (f, d["ms_per_step"], d["e2e"]["ms_per_step"], d["result"]["parity"], k["kernel"][:22], k["ms_per_launch"], k["frac"], d["phases_ms"]["ms_sort"]))
        elif "exit" in l and "exit 0" not in l: print(f, l.strip()[:200])
PY
