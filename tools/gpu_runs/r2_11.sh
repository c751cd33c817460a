#!/bin/bash
# round 2, call 11 (1 GPU): side array only in pipelined builds -- tests, bench; sweep of the pipeline group size (e2e)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_11_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_11_pytest.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r2_11_bench.log 2> gpurun_out/r2_11_bench.err; echo "exit $?" >> gpurun_out/r2_11_bench.log
for MB in 16 24 48 64; do
  PDG_FB_GROUP_BYTES=$((MB << 20)) timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-h2d-ceiling > gpurun_out/r2_11_group$MB.log 2> gpurun_out/r2_11_group$MB.err; echo "exit $?" >> gpurun_out/r2_11_group$MB.log
done
timeout 300 python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/r2_11_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/r2_11_smoke.log
tail -n 4 gpurun_out/r2_11_pytest.log | cut -c1-300; tail -n 2 gpurun_out/r2_11_smoke.log
python - <<PY
import json
for f in ["bench", "group16", "group24", "group48", "group64"]:
    for l in open("gpurun_out/r2_11_%s.log" % f):
        if l.startswith("{"):
            d = json.loads(l)
            print("%-8s ms/step %.3f | e2e %.3f | parity %s | groups %s finalize %.2f" % (f, d["ms_per_step"], d["e2e"]["ms_per_step"], d["result"]["parity"], d["phases_ms"]["e2e_sort_groups"], d["phases_ms"]["e2e_ms_finalize"]))
            if f == "bench":
                for k in d["roofline"]["all_kernels"]: print("    %-66s %.3f ms  frac %.3f" % (k["kernel"][:66], k["ms_per_launch"], k["frac"]))
                print("    cpu_baseline", d.get("cpu_baseline", {}).get("value"), d.get("cpu_baseline", {}).get("port", {}).get("one_thread", {}).get("value"))
        elif "exit" in l and "exit 0" not in l: print(f, l.strip()[:200])
PY
