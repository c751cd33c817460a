#!/bin/bash
# round 2, call 5 (1 GPU): byte-wide count table (core v4) -- tests, A/B, bench; how the k = 31 sort scales with genomes
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_05_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_05_pytest.log
timeout 900 python tools/tune_msd3.py --run 100 > gpurun_out/r2_05_tune.log 2>&1; echo "exit $?" >> gpurun_out/r2_05_tune.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_05_bench.log 2> gpurun_out/r2_05_bench.err; echo "exit $?" >> gpurun_out/r2_05_bench.log
for G in 100 200 300 500; do
  timeout 900 python bench.py --k 31 --genomes $G --steps 2 --warmup 2 --no-cpu-baseline --no-h2d-ceiling > gpurun_out/r2_05_k31_g$G.log 2> gpurun_out/r2_05_k31_g$G.err; echo "exit $?" >> gpurun_out/r2_05_k31_g$G.log
done
tail -n 4 gpurun_out/r2_05_pytest.log | cut -c1-300; cat gpurun_out/r2_05_tune.log; tail -n 3 gpurun_out/r2_05_bench.err
python - <<PY
import json
for f in ["bench", "k31_g100", "k31_g200", "k31_g300", "k31_g500"]:
    for l in open("gpurun_out/r2_05_%s.log" % f):
        if l.startswith("{"):
            d = json.loads(l)
            print("%-9s ms/step %.3f | e2e %.3f | parity %s | phases %s" % (f, d["ms_per_step"], d["e2e"]["ms_per_step"], d["result"]["parity"], {k: round(v, 2) for k, v in d["phases_ms"].items() if k in ("ms_sort", "ms_nodes", "ms_total", "e2e_ms_finalize", "e2e_sort_groups", "e2e_ms_h2d_copies")}))
            if f == "bench":
                for k in d["roofline"]["all_kernels"]: print("    %-62s %.3f ms  frac %.3f" % (k["kernel"][:62], k["ms_per_launch"], k["frac"]))
        elif "exit" in l and "exit 0" not in l: print(f, l.strip()[:200])
PY
tail -n 3 gpurun_out/r2_05_k31_g500.err | cut -c1-300
