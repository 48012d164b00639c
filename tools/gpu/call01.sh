#!/bin/bash
# round 2, GPU call 1: full GPU test suite (incl. the new loop parity tests), both bench arms,
# compute-sanitizer on smoke(), ncu launch list.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader > gpurun_out/c01_gpu.txt
python -m pytest tests -m gpu -x -q -s > gpurun_out/c01_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c01_pytest.log
tail -5 gpurun_out/c01_pytest.log
python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/c01_bench_ref.json 2> gpurun_out/c01_bench_ref.err; echo "ref rc=$?"
python bench.py --steps 20 --warmup 5 > gpurun_out/c01_bench_native.json 2> gpurun_out/c01_bench_native.err; echo "native rc=$?"
tail -c 1500 gpurun_out/c01_bench_ref.json; tail -c 3000 gpurun_out/c01_bench_native.json
tail -5 gpurun_out/c01_bench_native.err
timeout 600 compute-sanitizer --tool memcheck --log-file gpurun_out/c01_memcheck.log python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c01_memcheck.out 2>&1; echo "memcheck rc=$?"
timeout 900 compute-sanitizer --tool racecheck --log-file gpurun_out/c01_racecheck.log python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c01_racecheck.out 2>&1; echo "racecheck rc=$?"
tail -3 gpurun_out/c01_memcheck.log gpurun_out/c01_racecheck.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/c01_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/c01_ncu.log 2>&1; echo "ncu rc=$?"
