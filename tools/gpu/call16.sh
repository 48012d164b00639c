#!/bin/bash
# round 2, GPU call 16: programmatic dependent launch on/off — full GPU suite with it on, chain + bench both ways
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/c16_pytest.log 2>&1; echo "pytest (PDL on) rc=$?"; tail -3 gpurun_out/c16_pytest.log | cut -c1-300
for S in 1 0; do
  B2GS_PDL=$S timeout 600 python tools/chain_time.py 1 > gpurun_out/c16_chain_pdl$S.json 2>/dev/null; echo "chain pdl=$S rc=$?"
  python -c "import json; d=json.load(open('gpurun_out/c16_chain_pdl$S.json')); print('PDL', $S, d['views1_lean'], d['views1_autograd'])"
  B2GS_PDL=$S timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/c16_bench_pdl$S.json 2>/dev/null; echo "bench pdl=$S rc=$?"
  python -c "import json; d=json.loads(open('gpurun_out/c16_bench_pdl$S.json').read().strip().splitlines()[-1]); print('PDL', $S, d['value'], d['ms_per_step'], d['e2e']['value'], d['single_view_fwd_bwd_iters_s'])"
done
timeout 600 python tools/secondary_configs.py gpurun_out/c16_secondary.json 2>/dev/null | cut -c1-420
