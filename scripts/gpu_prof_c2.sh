#!/bin/bash
# launch list + full capture of the dominant kernel at the bench configuration (C2, 256^3); taken after the plain run exited 0
mkdir -p gpurun_out
timeout 300 python bench.py --steps 3 --warmup 3 --kernel-only > gpurun_out/bench_ko.log 2> gpurun_out/bench_ko.err || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 3 --warmup 3 --kernel-only > gpurun_out/ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:thermal_hex -s 3 -c 1 -f -o gpurun_out/prof_thermal_n256 \
    python bench.py --steps 3 --warmup 3 --kernel-only > gpurun_out/ncu_full.log 2>&1; echo "full capture rc=$?"
echo done
