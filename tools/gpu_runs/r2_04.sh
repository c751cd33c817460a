#!/bin/bash
# round 2, call 4 (1 GPU): partition core v2/v3, transposed histogram, filter() -- GPU tests, A/B of the variants, bench
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_04_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_04_pytest.log
timeout 900 python tools/tune_msd3.py --run 100 > gpurun_out/r2_04_tune.log 2>&1; echo "exit $?" >> gpurun_out/r2_04_tune.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_04_bench.log 2> gpurun_out/r2_04_bench.err; echo "exit $?" >> gpurun_out/r2_04_bench.log
tail -n 4 gpurun_out/r2_04_pytest.log | cut -c1-300; cat gpurun_out/r2_04_tune.log; tail -n 3 gpurun_out/r2_04_bench.err
python - <<PY
import json
for l in open("gpurun_out/r2_04_bench.log"):
    if l.startswith("{"):
        d = json.loads(l)
        print("ms/step %.3f | e2e %.3f | parity %s" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["result"]["parity"]))
        for k in d["roofline"]["all_kernels"]: print("  %-62s %.3f ms  frac %.3f" % (k["kernel"][:62], k["ms_per_launch"], k["frac"]))
    else: print(l.strip()[:200])
PY
