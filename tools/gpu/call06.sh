#!/bin/bash
# round 2, GPU call 6: bulk-async (TMA) staging of compacted splat records in both render kernels
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_rasterizer_gpu.py -q -x > gpurun_out/c06_raster.log 2>&1; echo "raster rc=$?"; tail -12 gpurun_out/c06_raster.log | cut -c1-400
timeout 900 python -m pytest tests -m gpu -q -x --deselect tests/test_rasterizer_gpu.py -s > gpurun_out/c06_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/c06_pytest.log | cut -c1-400
grep -E "step [0-9]|densify stats|tracking" gpurun_out/c06_pytest.log | cut -c1-250 | head -30
timeout 600 python tools/chain_time.py 1 8 > gpurun_out/c06_chain.json 2> gpurun_out/c06_chain.err; echo "chain rc=$?"; cat gpurun_out/c06_chain.json
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/c06_bench_native.json 2> gpurun_out/c06_bench_native.err; echo "native rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/c06_bench_native.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["state_check"], d["gpu_launches"], d["single_view_fwd_bwd_iters_s"])
for k, v in d["rooflines_all"].items(): print(k, v["avg_ms"], v["frac"], v["achieved"], v["unit"])
PY
