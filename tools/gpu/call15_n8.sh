#!/bin/bash
# round 2, GPU call 15 (8 GPUs): after local zero_grad / hoisted bias corrections / spread iso-reg
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 $TR --nproc-per-node 8 --master-port 29711 bench.py --gpus 8 --steps 40 --warmup 5 > gpurun_out/c20_bench_n8.json 2> gpurun_out/c20_bench_n8.err; echo "n8 rc=$?"
timeout 600 $TR --nproc-per-node 8 --master-port 29712 bench.py --gpus 8 --steps 40 --warmup 5 --peer-mode nomc > gpurun_out/c20_bench_n8_nomc.json 2> gpurun_out/c20_bench_n8_nomc.err; echo "n8 nomc rc=$?"
timeout 600 $TR --nproc-per-node 4 --master-port 29713 bench.py --gpus 4 --steps 40 --warmup 5 > gpurun_out/c20_bench_n4.json 2> gpurun_out/c20_bench_n4.err; echo "n4 rc=$?"
timeout 600 $TR --nproc-per-node 2 --master-port 29714 bench.py --gpus 2 --steps 40 --warmup 5 > gpurun_out/c20_bench_n2.json 2> gpurun_out/c20_bench_n2.err; echo "n2 rc=$?"
timeout 900 $TR --nproc-per-node 8 --master-port 29715 bench.py --gpus 8 --workload cfg5_large --steps 10 --warmup 3 > gpurun_out/c20_bench_cfg5_n8.json 2> gpurun_out/c20_bench_cfg5_n8.err; echo "cfg5 n8 rc=$?"
python - <<'PY'
import json
for f in ("c20_bench_n8", "c20_bench_n8_nomc", "c20_bench_n4", "c20_bench_n2", "c20_bench_cfg5_n8"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["value"], d["state_check"])
        print("    adam_peer", (d["rooflines_all"].get("adam_peer") or {}).get("avg_ms"))
    except Exception as e: print(f, "unreadable", e)
PY
