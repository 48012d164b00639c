#!/bin/bash
# per-kernel durations of the one-view lean window chain under ncu (cold cache, serialised) + graph step time
mkdir -p gpurun_out
python tools/chain_time.py 1 2>/dev/null | python -c "import sys,json; d=json.load(sys.stdin); print('graph step ms', d['views1_lean']['ms_per_step'])"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/kt.csv python tools/prof_window.py 6 > /dev/null 2>&1
python - <<'PY'
import csv, collections, re
rows = [l for l in open("gpurun_out/kt.csv") if not l.startswith("==")]
agg = collections.defaultdict(list)
for r in csv.DictReader(rows):
    if r.get("Metric Name") != "gpu__time_duration.sum": continue
    v = float(r["Metric Value"].replace(",", "")); u = r["Metric Unit"]
    v = v / 1e3 if u == "ns" else (v * 1e3 if u == "ms" else v)
    agg[re.sub(r"\(.*", "", r["Kernel Name"])[:44]].append(v)
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    v2 = v[len(v) // 2:]      # later launches: warmed-up sizes
    print(f"{k:46s} n={len(v):3d} avg_us={sum(v2)/len(v2):8.2f} min={min(v):8.2f}")
PY
