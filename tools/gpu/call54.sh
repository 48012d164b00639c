#!/bin/bash
# round 2: 384-entry staging batches in render_bwd only (A/B)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_rasterizer_gpu.py -m gpu -x -q > gpurun_out/c54_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 1 gpurun_out/c54_pytest.log
python bench.py --steps 40 --warmup 5 --no-cpu-baseline > gpurun_out/c54_bench_native.json 2> gpurun_out/c54_bench_native.err; echo "native rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/c54_bench_native.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"])
for k in ("render_bwd", "render_fwd"): print(k, d["rooflines_all"][k]["avg_ms"])
PY
