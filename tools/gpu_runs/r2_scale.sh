#!/bin/bash
# round 2: the BASELINE.json configs on N GPUs of one node (N = $1: 2, 4 or 8), every line parity-gated by bench.py /
# oracle-checked by bench_queries.py.  gpurun --gpus N -- 'bash tools/gpu_runs/r2_scale.sh N'
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi topo -m > gpurun_out/r2_scale_n${N}_topo.log 2>&1
timeout 600 $TR --master-port 29531 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2_scale_n${N}_config3_k11.log 2> gpurun_out/r2_scale_n${N}_config3_k11.err; echo "exit $?" >> gpurun_out/r2_scale_n${N}_config3_k11.log
timeout 900 $TR --master-port 29532 bench.py --gpus $N --k 31 --steps 3 --warmup 3 > gpurun_out/r2_scale_n${N}_config3_k31.log 2> gpurun_out/r2_scale_n${N}_config3_k31.err; echo "exit $?" >> gpurun_out/r2_scale_n${N}_config3_k31.log
timeout 1500 $TR --master-port 29533 bench.py --gpus $N --config 4 --steps 3 --warmup 3 > gpurun_out/r2_scale_n${N}_config4.log 2> gpurun_out/r2_scale_n${N}_config4.err; echo "exit $?" >> gpurun_out/r2_scale_n${N}_config4.log
timeout 1500 $TR --master-port 29534 tools/bench_queries.py --genomes 1000 --k 31 --strings > gpurun_out/r2_scale_n${N}_config5_k31.log 2>&1; echo "exit $?" >> gpurun_out/r2_scale_n${N}_config5_k31.log
timeout 900 $TR --master-port 29535 tools/bench_queries.py --genomes $((100 * N)) --k 11 > gpurun_out/r2_scale_n${N}_config5_k11.log 2>&1; echo "exit $?" >> gpurun_out/r2_scale_n${N}_config5_k11.log
python - <<PY
import json
N = $N
for f in ["config3_k11", "config3_k31", "config4"]:
    for l in open("gpurun_out/r2_scale_n%d_%s.log" % (N, f)):
        if l.startswith("{"):
            d = json.loads(l)
            print("%-12s N=%d ms/step %.2f value %.2f Gbp/s | e2e %.2f ms %.2f Gbp/s (copy-only %.2f ms) | parity %s | %s | clocks %s" % (f, d["n_gpus"], d["ms_per_step"], d["value"] / 1e9, d["e2e"]["ms_per_step"], d["e2e"]["value"] / 1e9, d["e2e"]["h2d_ceiling_ms"] or 0, d["result"]["parity"], {k: round(v, 2) for k, v in d["phases_ms"].items() if k in ("ms_sort", "ms_nodes", "ms_total", "e2e_ms_h2d_copies")}, d["clocks"]["reasons"]))
        elif "exit" in l and "exit 0" not in l: print(f, l.strip()[:200])
for f in ["config5_k31", "config5_k11"]:
    for l in open("gpurun_out/r2_scale_n%d_%s.log" % (N, f)):
        if l.startswith("{"):
            d = json.loads(l)
            print("%-12s N=%d %.0f queries/s  %.1f ms/batch (local kernels %.1f) hits %d missing %d oracle_checked %d (%d hits compared)" % (f, d["n_gpus"], d["value"], d["ms_per_batch"], d["ms_per_batch_local_kernels"], d["hits"], d["positives_missing_own_hit"], d["oracle_checked"], d["oracle_compared_hits"]))
        elif "exit" in l and "exit 0" not in l: print(f, l.strip()[:300])
PY
