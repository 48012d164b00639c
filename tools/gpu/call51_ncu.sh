#!/bin/bash
# round 2: ncu full capture + launch list of the LAST tree (r02g)
mkdir -p gpurun_out
python tools/prof_window.py 3 > gpurun_out/c51_plain.log 2>&1; echo "plain rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'render_|preprocess_|tile_|scatter_|adam_kernel|window_|pose_step' --launch-skip 12 -c 12 -o gpurun_out/c51_view python tools/prof_window.py 3 > gpurun_out/c51_ncu_full.log 2>&1; echo "ncu full rc=$?"
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/c51_plain_bench.log 2>&1; echo "plain bench rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/c51_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/c51_ncu.log 2>&1; echo "ncu list rc=$?"
