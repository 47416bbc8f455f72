#!/bin/bash
# re-entry pass: parity tests, bench line, C4 generic-kernel capture (64^3)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/bench.log
for c in C4 C5 C3; do n=64; [ $c = C3 ] && n=32; timeout 300 python scripts/bench_configs.py $c $n >> gpurun_out/cfg_small.log 2>&1; done; cat gpurun_out/cfg_small.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:generic_assembly -s 2 -c 1 -f -o gpurun_out/prof_c4_n64 \
    python scripts/bench_configs.py C4 64 > gpurun_out/ncu_c4.log 2>&1; echo "ncu rc=$?"
echo done
