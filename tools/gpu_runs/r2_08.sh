#!/bin/bash
# round 2, call 8 (2 GPUs): merge-tree owner side of the node exchange -- multi-rank tests, k = 31 bench with the merge
# and with round 1's LSD sort (PDG_EXCHANGE_LSD=1), config 4 at N = 2, config-5 queries over 2 ranks (device merge)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_gpu_multi_rank.py -m gpu -q > gpurun_out/r2_08_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_08_pytest.log
timeout 900 $TR --master-port 29521 bench.py --gpus 2 --k 31 --steps 3 --warmup 3 > gpurun_out/r2_08_k31_merge.log 2> gpurun_out/r2_08_k31_merge.err; echo "exit $?" >> gpurun_out/r2_08_k31_merge.log
PDG_EXCHANGE_LSD=1 timeout 900 $TR --master-port 29522 bench.py --gpus 2 --k 31 --steps 3 --warmup 3 --no-h2d-ceiling > gpurun_out/r2_08_k31_lsd.log 2> gpurun_out/r2_08_k31_lsd.err; echo "exit $?" >> gpurun_out/r2_08_k31_lsd.log
timeout 600 $TR --master-port 29523 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2_08_k11.log 2> gpurun_out/r2_08_k11.err; echo "exit $?" >> gpurun_out/r2_08_k11.log
timeout 1500 $TR --master-port 29524 bench.py --gpus 2 --config 4 --steps 3 --warmup 3 > gpurun_out/r2_08_config4.log 2> gpurun_out/r2_08_config4.err; echo "exit $?" >> gpurun_out/r2_08_config4.log
timeout 900 $TR --master-port 29525 tools/bench_queries.py --genomes 200 --k 11 > gpurun_out/r2_08_q11.log 2>&1; echo "exit $?" >> gpurun_out/r2_08_q11.log
timeout 1500 $TR --master-port 29526 tools/bench_queries.py --genomes 1000 --k 31 --strings > gpurun_out/r2_08_q31.log 2>&1; echo "exit $?" >> gpurun_out/r2_08_q31.log
tail -n 4 gpurun_out/r2_08_pytest.log | cut -c1-300
python - <<PY
import json
for f in ["k31_merge", "k31_lsd", "k11", "config4"]:
    for l in open("gpurun_out/r2_08_%s.log" % f):
        if l.startswith("{"):
            d = json.loads(l)
            print("%-10s ms/step %.2f value %.2f Gbp/s | e2e %.2f ms | parity %s | %s" % (f, d["ms_per_step"], d["value"] / 1e9, d["e2e"]["ms_per_step"], d["result"]["parity"], {k: round(v, 2) for k, v in d["phases_ms"].items() if k in ("ms_sort", "ms_nodes", "ms_total")}))
        elif "exit" in l and "exit 0" not in l: print(f, l.strip()[:200])
PY
for f in q11 q31; do tail -n 2 gpurun_out/r2_08_$f.log | cut -c1-900; done
