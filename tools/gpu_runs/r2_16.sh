#!/bin/bash
# round 2, call 16 (1 GPU): pdg_codes_checksum_static -- its test alone in a fresh process and after all the others (compute-sanitizer is closed on this pool: that line only reports so)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2_16_alone.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_16_alone.log
timeout 900 compute-sanitizer --tool memcheck python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2_16_memcheck.log 2>&1; echo "exit $?" >> gpurun_out/r2_16_memcheck.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_16_all.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_16_all.log
for f in alone memcheck all; do echo "== $f"; tail -n 6 gpurun_out/r2_16_$f.log | cut -c1-300; done
