#!/bin/bash
# round 2, call 34 (2 GPUs): CTA sort with per-tier lists, positions in shared memory, rolled searches -- long-k tests, bench k = 31, config 4 on 2 GPUs
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_scale.py -m gpu -x -q > gpurun_out/r2_34_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_34_pytest.log
tail -n 3 gpurun_out/r2_34_pytest.log | cut -c1-300
timeout 600 python bench.py --k 31 --steps 3 --warmup 3 --no-cpu-baseline --no-h2d-ceiling > gpurun_out/r2_34_k31.log 2> gpurun_out/r2_34_k31.err; echo "exit $?" >> gpurun_out/r2_34_k31.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 1500 $TR --master-port 29533 bench.py --gpus 2 --config 4 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2_34_n2_config4.log 2> gpurun_out/r2_34_n2_config4.err; echo "exit $?" >> gpurun_out/r2_34_n2_config4.log
python - <<PY
import json
for f in ["k31", "n2_config4"]:
    for l in open("gpurun_out/r2_34_%s.log" % f):
        if l.startswith("{"):
            d = json.loads(l)
            print("%-12s N=%d ms/step %.2f value %.2f Gbp/s | e2e %.2f ms | parity %s | %s" % (f, d["n_gpus"], d["ms_per_step"], d["value"] / 1e9, d["e2e"]["ms_per_step"], d["result"]["parity"], {k: round(v, 2) for k, v in d["phases_ms"].items() if k in ("ms_sort", "ms_nodes", "ms_total")}))
        elif "exit" in l and "exit 0" not in l: print(f, l.strip()[:200])
PY
