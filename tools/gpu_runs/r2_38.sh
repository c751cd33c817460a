#!/bin/bash
# round 2, call 38 (1 GPU): rows kernel with four rows per lane and trip -- long-k tests, bench k = 31
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_scale.py -m gpu -x -q -k "long_k" > gpurun_out/r2_38_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_38_pytest.log
tail -n 2 gpurun_out/r2_38_pytest.log | cut -c1-200
timeout 300 python bench.py --k 31 --steps 3 --warmup 3 --no-cpu-baseline --no-h2d-ceiling > gpurun_out/r2_38_k31.log 2> gpurun_out/r2_38_k31.err; echo "exit $?" >> gpurun_out/r2_38_k31.log
python - <<PY
import json
for l in open("gpurun_out/r2_38_k31.log"):
    if l.startswith("{"):
        d = json.loads(l)
        print("k31 ms/step %.3f | e2e %.3f | parity %s | ms_total %.3f ms_sort %.3f" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["result"]["parity"], d["phases_ms"]["ms_total"], d["phases_ms"]["ms_sort"]))
    elif "exit" in l and "exit 0" not in l: print(l.strip()[:200])
PY
