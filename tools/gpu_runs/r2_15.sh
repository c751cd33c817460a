#!/bin/bash
# round 2, call 15 (1 GPU): what the driver runs at round end, with the final tree -- tests, smoke, both bench arms
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_15_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_15_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_15_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/r2_15_smoke.log
timeout 900 python bench.py --impl reference > gpurun_out/r2_15_bench_ref.log 2> gpurun_out/r2_15_bench_ref.err; echo "exit $?" >> gpurun_out/r2_15_bench_ref.log
timeout 900 python bench.py > gpurun_out/r2_15_bench.log 2> gpurun_out/r2_15_bench.err; echo "exit $?" >> gpurun_out/r2_15_bench.log
timeout 900 python bench.py --k 31 --steps 3 --warmup 3 --no-cpu-baseline --no-h2d-ceiling > gpurun_out/r2_15_bench_k31.log 2> gpurun_out/r2_15_bench_k31.err; echo "exit $?" >> gpurun_out/r2_15_bench_k31.log
tail -n 3 gpurun_out/r2_15_pytest.log | cut -c1-300; tail -n 3 gpurun_out/r2_15_smoke.log | cut -c1-300
for f in bench_ref bench bench_k31; do tail -n 2 gpurun_out/r2_15_$f.log | cut -c1-1500; done
