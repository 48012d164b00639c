#!/bin/bash
# round 2: GPU suite twice (flakiness) + smoke + default bench (with the CPU baseline leg) on the final tree
mkdir -p gpurun_out
for i in 1 2; do timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c40_pytest_$i.log 2>&1; echo "pytest $i rc=$?"; tail -n 1 gpurun_out/c40_pytest_$i.log; grep "FAILED" gpurun_out/c40_pytest_$i.log | cut -c1-200; done
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c40_smoke.log 2>&1; echo "smoke rc=$?"
( time python bench.py ) > gpurun_out/c40_bench_default.json 2> gpurun_out/c40_bench_default.err; echo "bench rc=$?"; grep real gpurun_out/c40_bench_default.err
python -c "
import json
d = json.loads(open('gpurun_out/c40_bench_default.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['steps'], d['warmup'], d['cpu_baseline'], d['roofline']['ncu'], d['gpu_launches'])"
