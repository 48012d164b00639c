#!/bin/bash
# round 2, 2 GPUs: peer window with a binning overflow on rank 0 only (deferred finalize + b2gs_peer_barrier_status)
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2"
timeout 300 $TR --master-port 29931 tools/peer_check.py cfg1_plumbing graph > gpurun_out/c48_peer_overflow.log 2>&1; echo "rc=$?"; grep "^rank" gpurun_out/c48_peer_overflow.log; tail -n 3 gpurun_out/c48_peer_overflow.log | cut -c1-300
