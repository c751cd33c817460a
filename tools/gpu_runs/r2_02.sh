#!/bin/bash
# round 2, call 2 (1 GPU): new tests, bench through the parity gate, config-5 queries with oracle check + strings
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_02_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_02_pytest.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r2_02_bench.log 2> gpurun_out/r2_02_bench.err; echo "exit $?" >> gpurun_out/r2_02_bench.log
timeout 900 python tools/bench_queries.py --genomes 100 --k 11 --strings > gpurun_out/r2_02_q11.log 2>&1; echo "exit $?" >> gpurun_out/r2_02_q11.log
timeout 900 python tools/bench_queries.py --genomes 40 --k 31 --strings > gpurun_out/r2_02_q31.log 2>&1; echo "exit $?" >> gpurun_out/r2_02_q31.log
tail -n 15 gpurun_out/r2_02_pytest.log | cut -c1-300; tail -c 1500 gpurun_out/r2_02_bench.log; tail -n 5 gpurun_out/r2_02_bench.err; tail -n 3 gpurun_out/r2_02_q11.log | cut -c1-1500; tail -n 3 gpurun_out/r2_02_q31.log | cut -c1-1500
