#!/bin/bash
# round 2, 2 GPUs: peer path with the finalize kernel beside the first barrier (b2gs_peer_barrier_status)
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2"
timeout 300 $TR --master-port 29921 tools/peer_check.py cfg1_plumbing > gpurun_out/c33_peer_eager.log 2>&1; echo "eager rc=$?"; grep "^rank" gpurun_out/c33_peer_eager.log
timeout 300 $TR --master-port 29922 tools/peer_check.py cfg1_plumbing graph > gpurun_out/c33_peer_graph.log 2>&1; echo "graph rc=$?"; grep "^rank" gpurun_out/c33_peer_graph.log
timeout 300 $TR --master-port 29923 tools/peer_check.py cfg3_replica graph > gpurun_out/c33_peer_graph3.log 2>&1; echo "graph cfg3 rc=$?"; grep "^rank" gpurun_out/c33_peer_graph3.log
timeout 600 $TR --master-port 29924 bench.py --gpus 2 --steps 40 --warmup 5 > gpurun_out/c33_bench_n2.json 2> gpurun_out/c33_bench_n2.err; echo "n2 rc=$?"
python -c "
import json
d = json.loads(open('gpurun_out/c33_bench_n2.json').read().strip().splitlines()[-1])
print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['state_check'])"
