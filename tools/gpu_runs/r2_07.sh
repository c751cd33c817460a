#!/bin/bash
# round 2, call 7 (1 GPU): TMA prefetch in passes 2-3 -- tests, A/B (resident kernel times), bench; e2e A/B of the three
# libraries (the e2e step got 0.3 ms slower than round 1 while the resident step got faster)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_07_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_07_pytest.log
timeout 900 python tools/tune_msd3.py --run 100 > gpurun_out/r2_07_tune.log 2>&1; echo "exit $?" >> gpurun_out/r2_07_tune.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_07_bench.log 2> gpurun_out/r2_07_bench.err; echo "exit $?" >> gpurun_out/r2_07_bench.log
for v in r1core notma tma; do
  PDG_LIBPDGPU=$PWD/prairiedog_b200/build/variants/libpdgpu_$v.so timeout 900 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-h2d-ceiling > gpurun_out/r2_07_e2e_$v.log 2> gpurun_out/r2_07_e2e_$v.err; echo "exit $?" >> gpurun_out/r2_07_e2e_$v.log
done
tail -n 4 gpurun_out/r2_07_pytest.log | cut -c1-300; cat gpurun_out/r2_07_tune.log; tail -n 3 gpurun_out/r2_07_bench.err | cut -c1-300
python - <<PY
import json
for f in ["bench", "e2e_r1core", "e2e_notma", "e2e_tma"]:
    for l in open("gpurun_out/r2_07_%s.log" % f):
        if l.startswith("{"):
            d = json.loads(l)
            print("%-11s ms/step %.3f | e2e %.3f | parity %s | %s" % (f, d["ms_per_step"], d["e2e"]["ms_per_step"], d["result"]["parity"], {k: round(v, 2) for k, v in d["phases_ms"].items() if k.startswith("e2e")}))
            if f == "bench":
                for k in d["roofline"]["all_kernels"]: print("    %-62s %.3f ms  frac %.3f" % (k["kernel"][:62], k["ms_per_launch"], k["frac"]))
        elif "exit" in l and "exit 0" not in l: print(f, l.strip()[:200])
PY
