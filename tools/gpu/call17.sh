#!/bin/bash
# round 2, GPU call 17: forward render variant (8x4 block, one pixel per lane, two splats per trip)
mkdir -p gpurun_out
B2GS_FWD_PAIR=1 timeout 900 python -m pytest tests/test_rasterizer_gpu.py tests/test_loop_gpu.py tests/test_canary_gpu.py -q -x > gpurun_out/c17_pytest.log 2>&1; echo "pytest (pair) rc=$?"; tail -3 gpurun_out/c17_pytest.log | cut -c1-300
for S in 1 0; do
  B2GS_FWD_PAIR=$S timeout 600 python tools/chain_time.py 1 > gpurun_out/c17_chain_pair$S.json 2>/dev/null; echo "chain pair=$S rc=$?"
  python -c "import json; d=json.load(open('gpurun_out/c17_chain_pair$S.json')); print('PAIR', $S, d['views1_lean'], {k: v for k, v in d['chain_us'].items() if 'render' in k})"
  B2GS_FWD_PAIR=$S timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/c17_bench_pair$S.json 2>/dev/null; echo "bench pair=$S rc=$?"
  python -c "import json; d=json.loads(open('gpurun_out/c17_bench_pair$S.json').read().strip().splitlines()[-1]); print('PAIR', $S, d['value'], d['ms_per_step'], d['e2e']['value'], d['rooflines_all']['render_fwd']['avg_ms'], {k: d['config'][k] for k in d['config'] if 'fwd_' in k})"
done
