#!/bin/bash
# round 2, call 6 (1 GPU): product = v3 core + pass 3 at 4 CTAs/SM + merge-tree exchange; tests, bench, ncu launch list,
# ncu --set full with source lines for the timed kernels (numbers printed under ncu are never bench values)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_06_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_06_pytest.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r2_06_bench.log 2> gpurun_out/r2_06_bench.err; echo "exit $?" >> gpurun_out/r2_06_bench.log
timeout 300 python tools/profile_step.py 100 > gpurun_out/r2_06_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_k11.csv python tools/profile_step.py 100 > gpurun_out/r2_06_ncu1.log 2>&1
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'m3_pass|m3_hist|fb_summ|fb_emit' --launch-skip 6 -c 6 -f -o gpurun_out/r02_full_k11 python tools/profile_step.py 100 > gpurun_out/r2_06_ncu2.log 2>&1
ls -la gpurun_out/*.ncu-rep 2>/dev/null
tail -n 4 gpurun_out/r2_06_pytest.log | cut -c1-300; tail -n 3 gpurun_out/r2_06_bench.err | cut -c1-300
python - <<PY
import json
for l in open("gpurun_out/r2_06_bench.log"):
    if l.startswith("{"):
        d = json.loads(l)
        print("ms/step %.3f | e2e %.3f (ceiling %.2f) | parity %s" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["e2e"]["h2d_ceiling_ms"], d["result"]["parity"]))
        for k in d["roofline"]["all_kernels"]: print("    %-62s %.3f ms  frac %.3f" % (k["kernel"][:62], k["ms_per_launch"], k["frac"]))
PY
tail -n 2 gpurun_out/r2_06_ncu1.log gpurun_out/r2_06_ncu2.log | cut -c1-300
