#!/bin/bash
# round-end style pass: GPU parity tests, smoke, bench line, timings of the other configurations, final profiles
mkdir -p gpurun_out; rm -f gpurun_out/cfg_full.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -c 600 gpurun_out/bench.log
for c in "C4 192" "C5 128" "C3 128" "C1 4096"; do timeout 600 python scripts/bench_configs.py $c >> gpurun_out/cfg_full.log 2>&1; done; cut -c1-260 gpurun_out/cfg_full.log
bash scripts/gpu_prof_jit.sh
echo done
