#!/bin/bash
# round 2, call 17 (1 GPU): k = 31 bucket sort with the three-buckets-ahead loop (PDG_ML_PIPE) and the L2 fetch granularity
mkdir -p gpurun_out
B="python bench.py --k 31 --steps 3 --warmup 3 --no-cpu-baseline --no-h2d-ceiling"
PDG_ML_PIPE=0 timeout 600 $B > gpurun_out/r2_17_pipe0.log 2> gpurun_out/r2_17_pipe0.err; echo "exit $?" >> gpurun_out/r2_17_pipe0.log
PDG_ML_PIPE=1 timeout 600 $B > gpurun_out/r2_17_pipe1.log 2> gpurun_out/r2_17_pipe1.err; echo "exit $?" >> gpurun_out/r2_17_pipe1.log
PDG_ML_PIPE=0 PDG_L2_FETCH=32 timeout 600 $B > gpurun_out/r2_17_pipe0_f32.log 2> gpurun_out/r2_17_pipe0_f32.err; echo "exit $?" >> gpurun_out/r2_17_pipe0_f32.log
PDG_ML_PIPE=1 PDG_L2_FETCH=32 timeout 600 $B > gpurun_out/r2_17_pipe1_f32.log 2> gpurun_out/r2_17_pipe1_f32.err; echo "exit $?" >> gpurun_out/r2_17_pipe1_f32.log
timeout 900 python -m pytest tests -m gpu -x -q -k "long or k31 or scale or msdlong or 17 or 31" > gpurun_out/r2_17_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_17_pytest.log
tail -n 3 gpurun_out/r2_17_pytest.log | cut -c1-300
python - <<PY
import json
for f in ["pipe0", "pipe1", "pipe0_f32", "pipe1_f32"]:
    for l in open("gpurun_out/r2_17_%s.log" % f):
        if l.startswith("{"):
            d = json.loads(l)
            k = d["roofline"]["all_kernels"][0]
            print("%-10s ms/step %.3f | e2e %.3f | parity %s | %s %.3f ms | ms_sort %.3f" % (f, d["ms_per_step"], d["e2e"]["ms_per_step"], d["result"]["parity"], k["kernel"][:22], k["ms_per_launch"], d["phases_ms"]["ms_sort"]))
        elif "exit" in l and "exit 0" not in l: print(f, l.strip()[:200])
PY
