#!/bin/bash
# round 2: SH staging buffer of preprocess_bwd as dynamic shared memory (only when the path runs)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c52_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/c52_pytest.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c52_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 1 gpurun_out/c52_smoke.log
python bench.py --steps 40 --warmup 5 --no-cpu-baseline > gpurun_out/c52_bench_native.json 2> gpurun_out/c52_bench_native.err; echo "native rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/c52_bench_native.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["state_check"], d.get("gpu_launches"))
for k, r in d["rooflines_all"].items():
    print(k, r.get("avg_ms"), r.get("frac"))
PY
python bench.py --workload cfg5_large --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/c52_bench_cfg5_n1.json 2> gpurun_out/c52_bench_cfg5_n1.err; echo "cfg5 rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/c52_bench_cfg5_n1.json").read().strip().splitlines()[-1])
print("cfg5", d["value"], d["ms_per_step"], d["state_check"])
print("cfg5 preprocess_bwd", d["rooflines_all"]["preprocess_bwd"]["avg_ms"], d["rooflines_all"]["preprocess_bwd"]["frac"])
PY
