#!/bin/bash
# round 2, GPU call 2: lean window — tests, chain timing, bench, ncu full capture of one view.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -s > gpurun_out/c02_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c02_pytest.log
tail -4 gpurun_out/c02_pytest.log
grep -E "step [0-9]|off-fraction|kf [0-9]|densify stats|tracking:" gpurun_out/c02_pytest.log | head -40
python tools/chain_time.py 1 8 > gpurun_out/c02_chain.json 2> gpurun_out/c02_chain.err; echo "chain rc=$?"; cat gpurun_out/c02_chain.json; tail -3 gpurun_out/c02_chain.err
python bench.py --steps 20 --warmup 5 > gpurun_out/c02_bench_native.json 2> gpurun_out/c02_bench_native.err; echo "native rc=$?"
head -c 1200 gpurun_out/c02_bench_native.json; echo; tail -3 gpurun_out/c02_bench_native.err
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/c02_bench_ref.json 2> gpurun_out/c02_bench_ref.err; echo "ref rc=$?"
head -c 400 gpurun_out/c02_bench_ref.json; echo
python - <<'PY'
import json
for f in ("gpurun_out/c02_bench_native.json", "gpurun_out/c02_bench_ref.json"):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["value"], d["e2e"]["value"], d["state_check"])
    except Exception as e: print(f, "unreadable", e)
PY
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'render_|preprocess_|tile_|scatter_|adam_kernel|window_|pose_step' -c 26 -o gpurun_out/c02_view python tools/prof_window.py 2 > gpurun_out/c02_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/c02_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/c02_ncu.log 2>&1; echo "ncu list rc=$?"
