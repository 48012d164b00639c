#!/bin/bash
# round 2, last check of the final tree (r02g): GPU suite, smoke, both bench arms with their default flags
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c46_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 1 gpurun_out/c46_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c46_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 1 gpurun_out/c46_smoke.log
python bench.py --impl reference > gpurun_out/c46_bench_ref.json 2> gpurun_out/c46_bench_ref.err; echo "ref rc=$?"
python bench.py > gpurun_out/c46_bench_native.json 2> gpurun_out/c46_bench_native.err; echo "native rc=$?"
python - <<'PY'
import json
for f in ("c46_bench_native", "c46_bench_ref"):
    d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, d["value"], d["ms_per_step"], d["e2e"]["value"], d["state_check"], d.get("gpu_launches"))
d = json.loads(open("gpurun_out/c46_bench_native.json").read().strip().splitlines()[-1])
for k, r in d["rooflines_all"].items(): print(k, r.get("avg_ms"), r.get("frac"))
print(d["roofline"]["ncu"], d["cpu_baseline"])
PY
