#!/bin/bash
# round 2, GPU call 7 (2 GPUs): lean window over peer memory — NCCL vs plain P2P vs NVLS multicast; bench N=2
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 $TR --nproc-per-node 2 --master-port 29611 tools/peer_modes.py cfg3_replica 40 > gpurun_out/c07_modes_n2.json 2> gpurun_out/c07_modes_n2.err; echo "modes rc=$?"; cat gpurun_out/c07_modes_n2.json; tail -5 gpurun_out/c07_modes_n2.err | cut -c1-300
timeout 600 $TR --nproc-per-node 2 --master-port 29612 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/c07_bench_n2.json 2> gpurun_out/c07_bench_n2.err; echo "bench rc=$?"; tail -3 gpurun_out/c07_bench_n2.err | cut -c1-300
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/c07_bench_n2.json").read().strip().splitlines()[-1])
    print(d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["value"], d["state_check"], d["config"]["parallelism"])
    for k, v in d["rooflines_all"].items(): print(k, v["avg_ms"], v["frac"], v["achieved"], v["unit"])
except Exception as e: print("unreadable", e)
PY
