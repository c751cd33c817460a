#!/bin/bash
# round 2, call 14 (1 GPU): digit split 7|8|7 against 7|7|8 -- A/B of the kernels, then tests + bench with the 7|8|7 library
mkdir -p gpurun_out
V=prairiedog_b200/build/variants
timeout 900 python tools/tune_msd3.py --run 100 > gpurun_out/r2_14_tune.log 2>&1; echo "exit $?" >> gpurun_out/r2_14_tune.log
PDG_LIBPDGPU=$PWD/$V/libpdgpu_split78.so timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_14_pytest78.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_14_pytest78.log
for v in split87 split78; do
PDG_LIBPDGPU=$PWD/$V/libpdgpu_$v.so timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_14_bench_$v.log 2> gpurun_out/r2_14_bench_$v.err; echo "exit $?" >> gpurun_out/r2_14_bench_$v.log
done
PDG_LIBPDGPU=$PWD/$V/libpdgpu_split78.so timeout 900 python bench.py --k 31 --steps 3 --warmup 3 --no-cpu-baseline --no-h2d-ceiling > gpurun_out/r2_14_bench_k31_split78.log 2> gpurun_out/r2_14_bench_k31_split78.err; echo "exit $?" >> gpurun_out/r2_14_bench_k31_split78.log
tail -n 4 gpurun_out/r2_14_pytest78.log | cut -c1-300; cat gpurun_out/r2_14_tune.log
python - <<PY
import json
for f in ["bench_split87", "bench_split78", "bench_k31_split78"]:
    for l in open("gpurun_out/r2_14_%s.log" % f):
        if l.startswith("{"):
            d = json.loads(l)
            print("%-18s ms/step %.3f | e2e %.3f | parity %s" % (f, d["ms_per_step"], d["e2e"]["ms_per_step"], d["result"]["parity"]))
            for k in d["roofline"]["all_kernels"]: print("    %-66s %.3f ms  frac %.3f" % (k["kernel"][:66], k["ms_per_launch"], k["frac"]))
        elif "exit" in l and "exit 0" not in l: print(f, l.strip()[:200])
PY
