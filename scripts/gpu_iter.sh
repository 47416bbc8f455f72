#!/bin/bash
# iteration pass: (optional) RED microbench, parity tests, bench line, ncu launch list + full capture of the dominant kernel
mkdir -p gpurun_out
if [ -x tools/red_microbench ] && [ "$1" = "red" ]; then timeout 120 tools/red_microbench > gpurun_out/red_microbench.txt 2>&1; cat gpurun_out/red_microbench.txt; fi
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/bench.log').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('value','ms_per_step','gpu_launches','clocks')}); print(d['roofline']); print(d['roofline_fp64']); print(d['e2e']); print(d['cpu_baseline'])
except Exception as e: print('bench parse failed',e); print(open('gpurun_out/bench.err').read()[-2000:])
PY
# profiles are taken only after the same command exited 0 without ncu
timeout 300 python bench.py --steps 3 --warmup 3 --kernel-only > gpurun_out/bench_ko.log 2>> gpurun_out/bench.err &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 3 --warmup 3 --kernel-only > gpurun_out/ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:thermal_hex -s 3 -c 1 -f -o gpurun_out/prof_thermal_n256 \
    python bench.py --steps 3 --warmup 3 --kernel-only > gpurun_out/ncu_full.log 2>&1
echo done
