#!/bin/bash
# round 2, call 23 (1 GPU): k = 31 bucket sort per class, loads serial / batched, small class at 8 / 6 CTAs per SM
mkdir -p gpurun_out
B="python bench.py --k 31 --steps 3 --warmup 3 --no-cpu-baseline --no-h2d-ceiling"
V=$PWD/prairiedog_b200/build/variants
for v in mlser mllb4 mllb4a6 mlsera6; do for c in 2 3; do
PDG_LIBPDGPU=$V/libpdgpu_$v.so PDG_ML_CLASSES=$c timeout 600 $B > gpurun_out/r2_23_${v}_c$c.log 2> gpurun_out/r2_23_${v}_c$c.err; echo "exit $?" >> gpurun_out/r2_23_${v}_c$c.log
done; done
python - <<PY
import json
for v in ["mlser", "mllb4", "mllb4a6", "mlsera6"]:
  for c in "23":
    f = "%s_c%s" % (v, c)
    for l in open("gpurun_out/r2_23_%s.log" % f):
        if l.startswith("{"):
            d = json.loads(l)
            k = d["roofline"]["all_kernels"][0]
            print("%-12s ms/step %.3f | e2e %.3f | parity %s | %s %.3f ms | ms_sort %.3f" % (f, d["ms_per_step"], d["e2e"]["ms_per_step"], d["result"]["parity"], k["kernel"][:22], k["ms_per_launch"], d["phases_ms"]["ms_sort"]))
        elif "exit" in l and "exit 0" not in l: print(f, l.strip()[:200])
PY
