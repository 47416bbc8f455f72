#!/bin/bash
# first GPU pass: parity tests, bench line, ncu launch list + one full capture of the dominant kernel
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.log | cut -c1-1500
tail -3 gpurun_out/bench.err
timeout 300 python bench.py --steps 3 --warmup 3 --n 128 > gpurun_out/bench128.log 2>> gpurun_out/bench.err &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_n128.csv \
    python bench.py --steps 3 --warmup 3 --n 128 > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:thermal_hex -s 3 -c 1 -f -o gpurun_out/prof_thermal_n128 \
    python bench.py --steps 3 --warmup 3 --n 128 > gpurun_out/ncu_full.log 2>&1
echo done
