#!/bin/bash
# round 2, GPU call 13: software-pipelined views (stagger) on/off at N=1, distributed iso-reg, peer-adam changes
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_loop_gpu.py tests/test_loop_parity_gpu.py tests/test_canary_gpu.py -q -x > gpurun_out/c13_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/c13_pytest.log | cut -c1-300
for S in 1 0; do
  B2GS_STAGGER=$S timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/c13_bench_stagger$S.json 2> gpurun_out/c13_bench_stagger$S.err; echo "stagger=$S rc=$?"
done
python - <<'PY'
import json
for s in (1, 0):
    try:
        d = json.loads(open(f"gpurun_out/c13_bench_stagger{s}.json").read().strip().splitlines()[-1])
        print("stagger", s, d["value"], d["ms_per_step"], d["e2e"]["value"], d["state_check"])
    except Exception as e: print(s, "unreadable", e)
PY
