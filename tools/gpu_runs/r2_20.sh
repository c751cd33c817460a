#!/bin/bash
# round 2, call 20 (1 GPU): what bounds the k = 31 bucket sort -- the kernel without its gathers, without its network, without both
mkdir -p gpurun_out
V=$PWD/prairiedog_b200/build/variants
for v in mlbase mlfake mlnosort mlfakenosort; do
PDG_LIBPDGPU=$V/libpdgpu_$v.so timeout 600 python tools/profile_step.py 100 31 > gpurun_out/r2_20_$v.log 2>&1; echo "exit $?" >> gpurun_out/r2_20_$v.log
python - <<PY
import re
t = open("gpurun_out/r2_20_$v.log").read()
m = re.search(r"'ms_sort_kernel_avg': ([0-9.]+)", t)
print("$v", "bucket sort kernel ms:", m.group(1) if m else t[-300:])
PY
done
