#!/bin/bash
# round 2, call 29 (1 GPU): the new test over every bucket size class of msdlong.cu (also with 1 and 2 launch classes)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_scale.py -m gpu -x -q > gpurun_out/r2_29_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_29_pytest.log
tail -n 3 gpurun_out/r2_29_pytest.log | cut -c1-300
for c in 1 2; do
PDG_ML_CLASSES=$c timeout 600 python -m pytest tests/test_gpu_scale.py -m gpu -x -q -k "every_bucket or long_k" > gpurun_out/r2_29_pytest_c$c.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_29_pytest_c$c.log
tail -n 2 gpurun_out/r2_29_pytest_c$c.log | cut -c1-300
done
