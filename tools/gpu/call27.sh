#!/bin/bash
# round 2: ncu full capture of the one-view chain after the render_bwd rewrites + kernel durations
mkdir -p gpurun_out
python tools/prof_window.py 3 > gpurun_out/c27_plain.log 2>&1; echo "plain rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'render_|preprocess_|tile_|scatter_' --launch-skip 12 -c 7 -o gpurun_out/c27_view python tools/prof_window.py 3 > gpurun_out/c27_ncu_full.log 2>&1; echo "ncu full rc=$?"
bash tools/gpu/kern_times.sh > gpurun_out/c27_kern_times.txt 2>&1; cat gpurun_out/c27_kern_times.txt
