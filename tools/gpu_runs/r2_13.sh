#!/bin/bash
# round 2, call 13 (1 GPU): pass 3 walking its bucket backwards -- tests, A/B, bench (k = 11 / k = 31), DRAM traffic
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_13_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_13_pytest.log
timeout 900 python tools/tune_msd3.py --run 100 > gpurun_out/r2_13_tune.log 2>&1; echo "exit $?" >> gpurun_out/r2_13_tune.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_13_bench.log 2> gpurun_out/r2_13_bench.err; echo "exit $?" >> gpurun_out/r2_13_bench.log
timeout 900 python bench.py --k 31 --steps 3 --warmup 3 --no-cpu-baseline --no-h2d-ceiling > gpurun_out/r2_13_bench_k31.log 2> gpurun_out/r2_13_bench_k31.err; echo "exit $?" >> gpurun_out/r2_13_bench_k31.log
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'m3_pass|m3_hist|fb_summ|fb_emit' --launch-skip 6 -c 6 -f -o gpurun_out/r02_full_k11_p3rev python tools/profile_step.py 100 > gpurun_out/r2_13_ncu2.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_k11_p3rev.csv python tools/profile_step.py 100 > gpurun_out/r2_13_ncu1.log 2>&1
tail -n 4 gpurun_out/r2_13_pytest.log | cut -c1-300; cat gpurun_out/r2_13_tune.log
python - <<PY
import json
for f in ["bench", "bench_k31"]:
    for l in open("gpurun_out/r2_13_%s.log" % f):
        if l.startswith("{"):
            d = json.loads(l)
            print("%-9s ms/step %.3f | e2e %.3f | parity %s" % (f, d["ms_per_step"], d["e2e"]["ms_per_step"], d["result"]["parity"]))
            for k in d["roofline"]["all_kernels"]: print("    %-66s %.3f ms  frac %.3f" % (k["kernel"][:66], k["ms_per_launch"], k["frac"]))
        elif "exit" in l and "exit 0" not in l: print(f, l.strip()[:200])
PY
