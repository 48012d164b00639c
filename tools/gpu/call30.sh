#!/bin/bash
# round 2: GPU suite three times (flakiness check of the loop-parity bounds) + bench
mkdir -p gpurun_out
for i in 1 2 3; do timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c30_pytest_$i.log 2>&1; echo "pytest $i rc=$?"; tail -n 1 gpurun_out/c30_pytest_$i.log; grep "tracking 100\|FAILED" gpurun_out/c30_pytest_$i.log | cut -c1-200; done
timeout 600 python -m pytest tests/test_loop_parity_gpu.py -q -s -m gpu 2>&1 | grep "tracking\|passed\|failed" | cut -c1-250
python bench.py --steps 40 --warmup 5 --no-cpu-baseline > gpurun_out/c30_bench_native.json 2> gpurun_out/c30_bench_native.err; echo "native rc=$?"
python -c "
import json
d = json.loads(open('gpurun_out/c30_bench_native.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['state_check'])"
