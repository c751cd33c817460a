#!/bin/bash
# round 2, call 3 (2 GPUs): N > 1 tests (build + collective query), bench N=2 with the parity gate, config 4 at N=2,
# config-5 queries over 2 ranks
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_gpu_multi_rank.py tests/test_gpu_query_strings.py -m gpu -q > gpurun_out/r2_03_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_03_pytest.log
timeout 600 $TR --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2_03_bench_n2.log 2> gpurun_out/r2_03_bench_n2.err; echo "exit $?" >> gpurun_out/r2_03_bench_n2.log
timeout 900 $TR --master-port 29512 bench.py --gpus 2 --k 31 --steps 3 --warmup 3 > gpurun_out/r2_03_bench_n2_k31.log 2> gpurun_out/r2_03_bench_n2_k31.err; echo "exit $?" >> gpurun_out/r2_03_bench_n2_k31.log
timeout 1500 $TR --master-port 29513 bench.py --gpus 2 --config 4 --steps 3 --warmup 3 > gpurun_out/r2_03_config4_n2.log 2> gpurun_out/r2_03_config4_n2.err; echo "exit $?" >> gpurun_out/r2_03_config4_n2.log
timeout 900 $TR --master-port 29514 tools/bench_queries.py --genomes 200 --k 11 > gpurun_out/r2_03_q11_n2.log 2>&1; echo "exit $?" >> gpurun_out/r2_03_q11_n2.log
timeout 1500 $TR --master-port 29515 tools/bench_queries.py --genomes 1000 --k 31 --strings > gpurun_out/r2_03_q31_n2.log 2>&1; echo "exit $?" >> gpurun_out/r2_03_q31_n2.log
nvidia-smi --query-gpu=index,memory.used,memory.total --format=csv > gpurun_out/r2_03_smi.log 2>&1
tail -n 6 gpurun_out/r2_03_pytest.log | cut -c1-400
for f in bench_n2 bench_n2_k31 config4_n2; do echo "== $f"; tail -n 4 gpurun_out/r2_03_$f.err | cut -c1-600; python - <<PY
import json
for l in open("gpurun_out/r2_03_$f.log"):
    if l.startswith("{"):
        d = json.loads(l)
        print("ms/step %.2f value %.2f Gbp/s | e2e %.2f ms %.2f Gbp/s ceiling %s ms | parity %s | %s" % (d["ms_per_step"], d["value"] / 1e9, d["e2e"]["ms_per_step"], d["e2e"]["value"] / 1e9, d["e2e"].get("h2d_ceiling_ms"), d["result"]["parity"], d["result"].get("golden") or d["result"].get("why")))
        print("phases", {k: round(v, 2) for k, v in d["phases_ms"].items()})
    else:
        print(l.strip()[:300])
PY
done
for f in q11_n2 q31_n2; do echo "== $f"; tail -n 3 gpurun_out/r2_03_$f.log | cut -c1-1800; done
