#!/bin/bash
# round 2, call 36 (1 GPU): last check of the committed tree -- all GPU tests, smoke
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_36_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_36_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_36_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/r2_36_smoke.log
tail -n 2 gpurun_out/r2_36_pytest.log | cut -c1-200; tail -n 2 gpurun_out/r2_36_smoke.log | cut -c1-200
