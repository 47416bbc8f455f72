#!/bin/bash
# batched-kernel pass: parity tests, full-size timings of C3/C4/C5, ncu capture of the batched kernel (C4 64^3, C5 64^3)
mkdir -p gpurun_out; rm -f gpurun_out/cfg_full.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for c in "C4 64" "C5 64" "C3 32" "C4 192" "C5 128" "C3 128"; do timeout 600 python scripts/bench_configs.py $c >> gpurun_out/cfg_full.log 2>&1; done; cat gpurun_out/cfg_full.log
if [ "$1" != "noncu" ]; then
for c in C4 C5 C3; do n=64; [ $c = C3 ] && n=32
timeout 600 ncu --set full --clock-control none --import-source on -k regex:batched_assembly -s 2 -c 1 -f -o gpurun_out/prof_bat_${c} \
    python scripts/bench_configs.py $c $n > gpurun_out/ncu_${c}.log 2>&1; echo "ncu $c rc=$?"
done
fi
echo done
