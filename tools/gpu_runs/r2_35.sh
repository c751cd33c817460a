#!/bin/bash
# round 2, call 35 (2 GPUs): CTA sort with unrolled searches; positions in shared memory or read twice -- config 4 on 2 GPUs
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_scale.py -m gpu -x -q -k "long_k" > gpurun_out/r2_35_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_35_pytest.log
tail -n 2 gpurun_out/r2_35_pytest.log | cut -c1-300
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
V=$PWD/prairiedog_b200/build/variants
for v in mlspos mlnospos; do
PDG_LIBPDGPU=$V/libpdgpu_$v.so timeout 1500 $TR --master-port 29533 bench.py --gpus 2 --config 4 --steps 3 --warmup 3 --no-cpu-baseline --no-h2d-ceiling > gpurun_out/r2_35_$v.log 2> gpurun_out/r2_35_$v.err; echo "exit $?" >> gpurun_out/r2_35_$v.log
done
python - <<PY
import json
for f in ["mlspos", "mlnospos"]:
    for l in open("gpurun_out/r2_35_%s.log" % f):
        if l.startswith("{"):
            d = json.loads(l)
            print("%-10s N=%d ms/step %.2f value %.2f Gbp/s | e2e %.2f ms | parity %s | %s" % (f, d["n_gpus"], d["ms_per_step"], d["value"] / 1e9, d["e2e"]["ms_per_step"], d["result"]["parity"], {k: round(v, 2) for k, v in d["phases_ms"].items() if k in ("ms_sort", "ms_nodes", "ms_total")}))
        elif "exit" in l and "exit 0" not in l: print(f, l.strip()[:200])
PY
