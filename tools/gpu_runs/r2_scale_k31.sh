#!/bin/bash
# round 2: the k = 31 lines of r2_scale.sh only (config 3 at k = 31, config 4) on N GPUs -- rerun after the bucket-sort rewrite
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29532 bench.py --gpus $N --k 31 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2_k31_n${N}_config3_k31.log 2> gpurun_out/r2_k31_n${N}_config3_k31.err; echo "exit $?" >> gpurun_out/r2_k31_n${N}_config3_k31.log
timeout 1500 $TR --master-port 29533 bench.py --gpus $N --config 4 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2_k31_n${N}_config4.log 2> gpurun_out/r2_k31_n${N}_config4.err; echo "exit $?" >> gpurun_out/r2_k31_n${N}_config4.log
python - <<PY
import json
N = $N
for f in ["config3_k31", "config4"]:
    for l in open("gpurun_out/r2_k31_n%d_%s.log" % (N, f)):
        if l.startswith("{"):
            d = json.loads(l)
            print("%-12s N=%d ms/step %.2f value %.2f Gbp/s | e2e %.2f ms %.2f Gbp/s | parity %s | %s" % (f, d["n_gpus"], d["ms_per_step"], d["value"] / 1e9, d["e2e"]["ms_per_step"], d["e2e"]["value"] / 1e9, d["result"]["parity"], {k: round(v, 2) for k, v in d["phases_ms"].items() if k in ("ms_sort", "ms_nodes", "ms_total")}))
        elif "exit" in l and "exit 0" not in l: print(f, l.strip()[:200])
PY
