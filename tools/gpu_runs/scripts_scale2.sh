#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi_rank.py -m gpu -x -q > gpurun_out/pytest_multi.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_multi.log
for n in 2 8; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) bench.py --gpus $n --k 31 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/scale_k31_n$n.log 2>&1; echo "exit $?" >> gpurun_out/scale_k31_n$n.log
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29700 bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/scale_n8.log 2>&1; echo "exit $?" >> gpurun_out/scale_n8.log
for f in pytest_multi scale_k31_n2 scale_k31_n8 scale_n8; do echo "== $f"; tail -n 2 gpurun_out/$f.log | cut -c1-300; done
