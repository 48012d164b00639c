#!/bin/bash
# round 2, final single-GPU evidence run (r02f): GPU suite, smoke, both bench arms, ncu full capture + launch list
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/c37_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/c37_pytest.log | cut -c1-200
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c37_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/c37_smoke.log
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/c37_bench_ref.json 2> gpurun_out/c37_bench_ref.err; echo "ref rc=$?"
python bench.py --steps 20 --warmup 5 > gpurun_out/c37_bench_native.json 2> gpurun_out/c37_bench_native.err; echo "native rc=$?"
python - <<'PY'
import json
for f in ("c37_bench_native", "c37_bench_ref"):
    d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, d["value"], d["ms_per_step"], d["e2e"]["value"], d["state_check"], d.get("gpu_launches"))
PY
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'render_|preprocess_|tile_|scatter_|adam_kernel|window_|pose_step' --launch-skip 12 -c 12 -o gpurun_out/c37_view python tools/prof_window.py 3 > gpurun_out/c37_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/c37_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/c37_ncu.log 2>&1; echo "ncu list rc=$?"
