#!/bin/bash
# round 2: cfg5 (4M, SH3, 1080p, 16 keyframes) step time against the number of forked view streams
mkdir -p gpurun_out
for vs in 1 2 4 8 16; do
  python bench.py --workload cfg5_large --steps 8 --warmup 3 --no-cpu-baseline --view-streams $vs > gpurun_out/c44_cfg5_vs$vs.json 2> gpurun_out/c44_cfg5_vs$vs.err
  python -c "
import json
d = json.loads(open('gpurun_out/c44_cfg5_vs$vs.json').read().strip().splitlines()[-1])
print('cfg5 view_streams $vs', d['value'], d['ms_per_step'], d['e2e']['value'])"
done
for vs in 2 4; do
  python bench.py --steps 40 --warmup 5 --no-cpu-baseline --view-streams $vs > gpurun_out/c44_cfg3_vs$vs.json 2> gpurun_out/c44_cfg3_vs$vs.err
  python -c "
import json
d = json.loads(open('gpurun_out/c44_cfg3_vs$vs.json').read().strip().splitlines()[-1])
print('cfg3 view_streams $vs', d['value'], d['ms_per_step'], d['e2e']['value'])"
done
