#!/bin/bash
# round 2, call 21 (1 GPU): k = 31 bucket sort with one body of code for all bucket sizes (instruction cache) -- tests, then A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "long or k31 or scale or msdlong or 17 or 31" > gpurun_out/r2_21_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_21_pytest.log
tail -n 3 gpurun_out/r2_21_pytest.log | cut -c1-300
B="python bench.py --k 31 --steps 3 --warmup 3 --no-cpu-baseline --no-h2d-ceiling"
V=$PWD/prairiedog_b200/build/variants
for v in mltpl mlrt mlrt5 mlrt3; do for pipe in 0 1; do
PDG_LIBPDGPU=$V/libpdgpu_$v.so PDG_ML_PIPE=$pipe timeout 600 $B > gpurun_out/r2_21_${v}_p$pipe.log 2> gpurun_out/r2_21_${v}_p$pipe.err; echo "exit $?" >> gpurun_out/r2_21_${v}_p$pipe.log
done; done
python - <<PY
import json
for v in ["mltpl", "mlrt", "mlrt5", "mlrt3"]:
  for p in "01":
    f = "%s_p%s" % (v, p)
    for l in open("gpurun_out/r2_21_%s.log" % f):
        if l.startswith("{"):
            d = json.loads(l)
            k = d["roofline"]["all_kernels"][0]
            print("%-10s ms/step %.3f | e2e %.3f | parity %s | %s %.3f ms | ms_sort %.3f" % (f, d["ms_per_step"], d["e2e"]["ms_per_step"], d["result"]["parity"], k["kernel"][:22], k["ms_per_launch"], d["phases_ms"]["ms_sort"]))
        elif "exit" in l and "exit 0" not in l: print(f, l.strip()[:200])
PY
