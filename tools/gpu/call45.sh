#!/bin/bash
# round 2: automatic view-stream choice (cfg3 -> 8 streams, cfg5 -> 1)
mkdir -p gpurun_out
python bench.py --steps 40 --warmup 5 --no-cpu-baseline > gpurun_out/c45_cfg3.json 2> gpurun_out/c45_cfg3.err; echo "cfg3 rc=$?"
python bench.py --workload cfg5_large --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/c45_cfg5.json 2> gpurun_out/c45_cfg5.err; echo "cfg5 rc=$?"
python -c "
import json
for f in ('c45_cfg3', 'c45_cfg5'):
    d = json.loads(open('gpurun_out/%s.json' % f).read().strip().splitlines()[-1])
    print(f, d['value'], d['ms_per_step'], d['e2e']['value'], d['config']['launch'][-60:])"
