#!/bin/bash
# round 2: tracking loop (cfg2) kernel split + secondary configs with the current kernels
mkdir -p gpurun_out
python tools/track_kernels.py > gpurun_out/c35_track.json 2> gpurun_out/c35_track.err; echo "track rc=$?"; cat gpurun_out/c35_track.json
timeout 900 python tools/secondary_configs.py > gpurun_out/c35_secondary.log 2>&1; echo "secondary rc=$?"; grep -v "^W\|warn" gpurun_out/c35_secondary.log | cut -c1-400 | tail -12
