#!/bin/bash
# round 2: secondary configs + tracking kernel split on the last tree (r02g)
mkdir -p gpurun_out
python tools/track_kernels.py > gpurun_out/c50_track.json 2> gpurun_out/c50_track.err; echo "track rc=$?"; cat gpurun_out/c50_track.json
timeout 900 python tools/secondary_configs.py > gpurun_out/c50_secondary.log 2>&1; echo "secondary rc=$?"; grep -v "^W\|warn" gpurun_out/c50_secondary.log | cut -c1-400 | tail -12
