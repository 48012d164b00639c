#!/bin/bash
# round 2, GPU call 8: full GPU suite on the final kernels (lean tracking loop, golden tests), secondary configs
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x -s > gpurun_out/c08_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/c08_pytest.log | cut -c1-400
grep -E "step [0-9]|densify stats|tracking|off-fraction" gpurun_out/c08_pytest.log | cut -c1-260 | head -40
timeout 900 python tools/secondary_configs.py gpurun_out/c08_secondary.json > gpurun_out/c08_secondary.log 2>&1; echo "secondary rc=$?"; cut -c1-600 gpurun_out/c08_secondary.log
