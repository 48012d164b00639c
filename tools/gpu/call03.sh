#!/bin/bash
# round 2, GPU call 3: LPT tile order, GT prefetch, hoisted loads, faster finalize / pose step;
# parity experiment (graph vs eager, reversed-order reference twin); RAW-instantiation parity test.
mkdir -p gpurun_out
python -m pytest tests/test_loop_parity_gpu.py -q -s > gpurun_out/c03_parity.log 2>&1; echo "parity rc=$?"
grep -E "step [0-9]|off-fraction|kf [0-9]|densify stats|tracking:|passed|failed" gpurun_out/c03_parity.log | head -60
python -m pytest tests -m gpu -q -x --deselect tests/test_loop_parity_gpu.py > gpurun_out/c03_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/c03_pytest.log
python tools/chain_time.py 1 8 > gpurun_out/c03_chain.json 2> gpurun_out/c03_chain.err; echo "chain rc=$?"; cat gpurun_out/c03_chain.json
python bench.py --steps 20 --warmup 5 > gpurun_out/c03_bench_native.json 2> gpurun_out/c03_bench_native.err; echo "native rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/c03_bench_native.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["state_check"], d["gpu_launches"])
print(json.dumps(d["roofline"])[:900])
for k, v in d["rooflines_all"].items(): print(k, v["avg_ms"], v["frac"], v["achieved"], v["unit"])
PY
