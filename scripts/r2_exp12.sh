#!/bin/bash
# round-2 experiment 12 (1 GPU): slot-list dictionary of the tile plan (parity, timing, DRAM traffic), C5 timing repeat
mkdir -p gpurun_out
export MRH_NO_REBUILD=1
T=gpurun_out/r2_e12_timing.txt; : > $T
P='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["ms_per_step"],4), round(d.get("volume_kernel_ms",0),4), d.get("path"))'
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "specialised_kernel_variants or default_path or ragged or mid_size_c2 or large_mesh or host_buffer or numbering or tiled_scatter_2d or set_nodes" > gpurun_out/r2_e12_parity.log 2>&1; echo "tile parity rc=$?"; tail -3 gpurun_out/r2_e12_parity.log
run() { echo "$1" >> $T; shift; env "$@" 2>&1 | tail -1 | python -c "$P" >> $T 2>&1; }
run "C2 default (slot-list dictionary)" timeout 120 python bench.py --steps 10 --warmup 3 --kernel-only
run "C2 again" timeout 120 python bench.py --steps 10 --warmup 3 --kernel-only
run "C5 kernel-only" timeout 180 python bench.py --config C5 --steps 5 --warmup 3 --kernel-only
run "C5 kernel-only again" timeout 180 python bench.py --config C5 --steps 10 --warmup 3 --kernel-only
cat $T
timeout 400 ncu --set full --clock-control none --import-source on -k regex:thermal_hex -s 3 -c 1 -f -o gpurun_out/r2_prof_c2d python bench.py --steps 3 --warmup 3 --kernel-only > gpurun_out/r2_ncu_c2d.log 2>&1; echo "ncu c2 rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_bench_c2.csv python bench.py --steps 2 --warmup 1 --kernel-only > gpurun_out/r2_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 900 python bench.py > gpurun_out/r2_bench6.log 2> gpurun_out/r2_bench6.err; echo "bench rc=$?"; tail -c 500 gpurun_out/r2_bench6.log
