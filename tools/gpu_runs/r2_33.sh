#!/bin/bash
# round 2, call 33 (1 GPU): per-tier bucket lists -- long-k tests; ncu --set full of the 4-warp CTA sort at 300 genomes
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_scale.py -m gpu -x -q > gpurun_out/r2_33_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_33_pytest.log
tail -n 3 gpurun_out/r2_33_pytest.log | cut -c1-300
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'ml_mid_bucket' --launch-skip 2 -c 1 -f -o gpurun_out/r02_full_k31_mid python tools/profile_step.py 300 31 > gpurun_out/r2_33_ncu.log 2>&1
echo "ncu exit $?"; ls -la gpurun_out/r02_full_k31_mid.ncu-rep
