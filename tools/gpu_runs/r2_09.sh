#!/bin/bash
# round 2, call 9 (1 GPU): pass 1 with its codes in registers (A/B), bench, then the round's ncu evidence for both
# pass-1 variants: launch list (gpu__time_duration) and --set full with source lines of the timed kernels
mkdir -p gpurun_out
timeout 900 python tools/tune_msd3.py --run 100 > gpurun_out/r2_09_tune.log 2>&1; echo "exit $?" >> gpurun_out/r2_09_tune.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_09_bench.log 2> gpurun_out/r2_09_bench.err; echo "exit $?" >> gpurun_out/r2_09_bench.log
timeout 900 python bench.py --k 31 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2_09_bench_k31.log 2> gpurun_out/r2_09_bench_k31.err; echo "exit $?" >> gpurun_out/r2_09_bench_k31.log
for v in p1regs p1recompute; do
  export PDG_LIBPDGPU=$PWD/prairiedog_b200/build/variants/libpdgpu_$v.so
  timeout 300 python tools/profile_step.py 100 > gpurun_out/r2_09_plain_$v.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_k11_$v.csv python tools/profile_step.py 100 > gpurun_out/r2_09_ncu1_$v.log 2>&1
  timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'m3_pass|m3_hist|fb_summ|fb_emit' --launch-skip 6 -c 6 -f -o gpurun_out/r02_full_k11_$v python tools/profile_step.py 100 > gpurun_out/r2_09_ncu2_$v.log 2>&1
done
unset PDG_LIBPDGPU
timeout 300 python tools/profile_step.py 100 31 > gpurun_out/r2_09_plain_k31.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_k31.csv python tools/profile_step.py 100 31 > gpurun_out/r2_09_ncu1_k31.log 2>&1
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'ml_bucket_sort' --launch-skip 1 -c 1 -f -o gpurun_out/r02_full_k31_bucket python tools/profile_step.py 100 31 > gpurun_out/r2_09_ncu2_k31.log 2>&1
ls -la gpurun_out/*.ncu-rep; cat gpurun_out/r2_09_tune.log
python - <<PY
import json
for f in ["bench", "bench_k31"]:
    for l in open("gpurun_out/r2_09_%s.log" % f):
        if l.startswith("{"):
            d = json.loads(l)
            print("%-9s ms/step %.3f | e2e %.3f | parity %s" % (f, d["ms_per_step"], d["e2e"]["ms_per_step"], d["result"]["parity"]))
            for k in d["roofline"]["all_kernels"]: print("    %-66s %.3f ms  frac %.3f" % (k["kernel"][:66], k["ms_per_launch"], k["frac"]))
        elif "exit" in l and "exit 0" not in l: print(f, l.strip()[:200])
PY
