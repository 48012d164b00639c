#!/bin/bash
# round 2, GPU call 5: parity tests with the 1-ulp twin, densify kernels vs oracle, cfg5 bench (N=1, both arms)
mkdir -p gpurun_out
python -m pytest tests/test_loop_parity_gpu.py -q -s > gpurun_out/c05_parity.log 2>&1; echo "parity rc=$?"
grep -E "step [0-9]|off-fraction|kf [0-9]|densify stats|tracking|passed|failed|Error" gpurun_out/c05_parity.log | cut -c1-300 | head -70
python -m pytest tests/test_loop_gpu.py -q -x -k "densify or lean or graph_mapping" > gpurun_out/c05_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/c05_pytest.log | cut -c1-300
python bench.py --workload cfg5_large --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/c05_bench_cfg5_native.json 2> gpurun_out/c05_bench_cfg5_native.err; echo "cfg5 native rc=$?"; tail -3 gpurun_out/c05_bench_cfg5_native.err | cut -c1-300
python bench.py --workload cfg5_large --impl reference --steps 3 --warmup 3 > gpurun_out/c05_bench_cfg5_ref.json 2> gpurun_out/c05_bench_cfg5_ref.err; echo "cfg5 ref rc=$?"; tail -3 gpurun_out/c05_bench_cfg5_ref.err | cut -c1-300
python - <<'PY'
import json
for f in ("gpurun_out/c05_bench_cfg5_native.json", "gpurun_out/c05_bench_cfg5_ref.json"):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["value"], d["ms_per_step"], d["e2e"]["value"], d["state_check"])
        for k, v in (d.get("rooflines_all") or {}).items(): print("  ", k, v["avg_ms"], v["frac"], v["achieved"], v["unit"])
    except Exception as e: print(f, "unreadable", e)
PY
