#!/bin/bash
# round 2, call 12 (1 GPU): early pass 3 per super-group + concatenation (pipelined builds) -- tests, then the e2e step
# without it (PDG_M3_NO_EARLY=1) and with 1 / 2 / 3 / 4 / 6 groups per super-group
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_12_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_12_pytest.log
PDG_M3_NO_EARLY=1 timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-h2d-ceiling > gpurun_out/r2_12_noearly.log 2> gpurun_out/r2_12_noearly.err; echo "exit $?" >> gpurun_out/r2_12_noearly.log
for SG in 1 2 3 4 6; do
  PDG_M3_SUPER=$SG timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-h2d-ceiling > gpurun_out/r2_12_sg$SG.log 2> gpurun_out/r2_12_sg$SG.err; echo "exit $?" >> gpurun_out/r2_12_sg$SG.log
done
PDG_M3_SUPER=2 timeout 600 python bench.py --k 31 --steps 3 --warmup 3 --no-cpu-baseline --no-h2d-ceiling > gpurun_out/r2_12_k31_sg2.log 2> gpurun_out/r2_12_k31_sg2.err; echo "exit $?" >> gpurun_out/r2_12_k31_sg2.log
PDG_M3_NO_EARLY=1 timeout 600 python bench.py --k 31 --steps 3 --warmup 3 --no-cpu-baseline --no-h2d-ceiling > gpurun_out/r2_12_k31_noearly.log 2> gpurun_out/r2_12_k31_noearly.err; echo "exit $?" >> gpurun_out/r2_12_k31_noearly.log
tail -n 6 gpurun_out/r2_12_pytest.log | cut -c1-300
python - <<PY
import json
for f in ["noearly", "sg1", "sg2", "sg3", "sg4", "sg6", "k31_noearly", "k31_sg2"]:
    for l in open("gpurun_out/r2_12_%s.log" % f):
        if l.startswith("{"):
            d = json.loads(l)
            print("%-12s ms/step %.3f | e2e %.3f | parity %s | ingest_call %.2f h2d %.2f finalize %.2f" % (f, d["ms_per_step"], d["e2e"]["ms_per_step"], d["result"]["parity"], d["phases_ms"]["e2e_ms_ingest_call"], d["phases_ms"]["e2e_ms_h2d_copies"], d["phases_ms"]["e2e_ms_finalize"]))
        elif "exit" in l and "exit 0" not in l: print(f, l.strip()[:200])
PY
tail -n 3 gpurun_out/r2_12_sg2.err | cut -c1-300
