#!/bin/bash
# round 2, final 8-GPU run (r02f build): native N=8 / 4 / 2 / 1, reference arm under torchrun at N=8, cfg5 N=8
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 $TR --nproc-per-node 8 --master-port 29911 bench.py --gpus 8 --steps 40 --warmup 5 > gpurun_out/c38_bench_n8.json 2> gpurun_out/c38_bench_n8.err; echo "n8 rc=$?"
timeout 600 $TR --nproc-per-node 8 --master-port 29912 bench.py --impl reference --gpus 8 --steps 20 --warmup 5 > gpurun_out/c38_bench_ref_n8.json 2> gpurun_out/c38_bench_ref_n8.err; echo "ref n8 rc=$?"
timeout 600 $TR --nproc-per-node 4 --master-port 29913 bench.py --gpus 4 --steps 40 --warmup 5 > gpurun_out/c38_bench_n4.json 2> gpurun_out/c38_bench_n4.err; echo "n4 rc=$?"
timeout 600 $TR --nproc-per-node 2 --master-port 29914 bench.py --gpus 2 --steps 40 --warmup 5 > gpurun_out/c38_bench_n2.json 2> gpurun_out/c38_bench_n2.err; echo "n2 rc=$?"
timeout 600 python bench.py --gpus 1 --steps 40 --warmup 5 --no-cpu-baseline > gpurun_out/c38_bench_n1.json 2> gpurun_out/c38_bench_n1.err; echo "n1 rc=$?"
timeout 900 $TR --nproc-per-node 8 --master-port 29915 bench.py --gpus 8 --workload cfg5_large --steps 10 --warmup 3 > gpurun_out/c38_bench_cfg5_n8.json 2> gpurun_out/c38_bench_cfg5_n8.err; echo "cfg5 n8 rc=$?"
python - <<'PY'
import json
for f in ("c38_bench_n1", "c38_bench_n2", "c38_bench_n4", "c38_bench_n8", "c38_bench_ref_n8", "c38_bench_cfg5_n8"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["value"], d["state_check"])
    except Exception as e: print(f, "unreadable", e)
PY
