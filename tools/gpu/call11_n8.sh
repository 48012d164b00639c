#!/bin/bash
# round 2, GPU call 11 (8 GPUs): exchange modes at N=8, bench N=8 / N=4 at cfg3, bench N=8 at cfg5
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 $TR --nproc-per-node 8 --master-port 29711 tools/peer_modes.py cfg3_replica 60 > gpurun_out/c11_modes_n8.json 2> gpurun_out/c11_modes_n8.err; echo "modes rc=$?"; cat gpurun_out/c11_modes_n8.json; grep -v Warning gpurun_out/c11_modes_n8.err | tail -3 | cut -c1-300
timeout 600 $TR --nproc-per-node 8 --master-port 29712 bench.py --gpus 8 --steps 40 --warmup 5 > gpurun_out/c11_bench_n8.json 2> gpurun_out/c11_bench_n8.err; echo "bench n8 rc=$?"
timeout 600 $TR --nproc-per-node 8 --master-port 29713 bench.py --gpus 8 --steps 40 --warmup 5 --peer-mode nomc > gpurun_out/c11_bench_n8_nomc.json 2> gpurun_out/c11_bench_n8_nomc.err; echo "bench n8 nomc rc=$?"
timeout 600 $TR --nproc-per-node 4 --master-port 29714 bench.py --gpus 4 --steps 40 --warmup 5 > gpurun_out/c11_bench_n4.json 2> gpurun_out/c11_bench_n4.err; echo "bench n4 rc=$?"
timeout 900 $TR --nproc-per-node 8 --master-port 29715 bench.py --gpus 8 --workload cfg5_large --steps 10 --warmup 3 > gpurun_out/c11_bench_cfg5_n8.json 2> gpurun_out/c11_bench_cfg5_n8.err; echo "bench cfg5 n8 rc=$?"; grep -v Warning gpurun_out/c11_bench_cfg5_n8.err | tail -3 | cut -c1-300
python - <<'PY'
import json
for f in ("c11_bench_n8", "c11_bench_n8_nomc", "c11_bench_n4", "c11_bench_cfg5_n8"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["value"], d["state_check"], d["config"]["parallelism"][-60:])
        for k, v in d["rooflines_all"].items(): print("   ", k, v["avg_ms"], v["frac"], v["achieved"], v["unit"])
    except Exception as e: print(f, "unreadable", e)
PY
