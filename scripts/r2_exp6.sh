#!/bin/bash
# round-2 experiment 6 (1 GPU): new defaults (bulk list staging, volume-only batched kernels, shuffle-free Q2 fragments), timings, profiles
mkdir -p gpurun_out
export MRH_NO_REBUILD=1
T=gpurun_out/r2_e6_timing.txt; : > $T
P='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["ms_per_step"],4), round(d.get("volume_kernel_ms",0),4), d.get("path"))'
timeout 500 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "C3 or cdr or Q2 or boundary or Neumann or weak or Flux or side or variants or default_path or ragged" > gpurun_out/r2_e6_parity.log 2>&1; echo "parity subset rc=$?"; tail -3 gpurun_out/r2_e6_parity.log
run() { echo "$1" >> $T; shift; env "$@" 2>&1 | tail -1 | python -c "$P" >> $T 2>&1; }
run "C1 default" timeout 120 python bench.py --config C1 --steps 10 --warmup 3 --kernel-only
run "C1 MINB2D=2" MRH_JIT_MINB2D=2 timeout 120 python bench.py --config C1 --steps 10 --warmup 3 --kernel-only
run "C2 default (bulk)" timeout 120 python bench.py --steps 10 --warmup 3 --kernel-only
run "C2 nobulk" MRH_WS_BULK=0 timeout 120 python bench.py --steps 10 --warmup 3 --kernel-only
for v in 2 4; do run "C2 QUNROLL=$v" MRH_JIT_DEFINE=MRH_WS_QUNROLL=$v timeout 120 python bench.py --steps 10 --warmup 3 --kernel-only; done
run "C2 CREG=208 SREG=88" MRH_WS_CREG=208 MRH_WS_SREG=88 timeout 120 python bench.py --steps 10 --warmup 3 --kernel-only
run "C3 default" timeout 180 python bench.py --config C3 --steps 5 --warmup 3 --kernel-only
run "C3 gather" MRH_BAT_DEFINE=MRH_BAT_FRAG_GATHER timeout 180 python bench.py --config C3 --steps 5 --warmup 3 --kernel-only
run "C4 default" timeout 180 python bench.py --config C4 --steps 5 --warmup 3 --kernel-only
run "C4 atoms=1" MRH_BAT_ATOMS=1 timeout 180 python bench.py --config C4 --steps 5 --warmup 3 --kernel-only
run "C5 default" timeout 180 python bench.py --config C5 --steps 5 --warmup 3 --kernel-only
cat $T
timeout 400 ncu --set full --clock-control none --import-source on -k regex:thermal_hex -s 3 -c 1 -f -o gpurun_out/r2_prof_c2 python bench.py --steps 3 --warmup 3 --kernel-only > gpurun_out/r2_ncu_c2.log 2>&1; echo "ncu c2 rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:batched -s 3 -c 1 -f -o gpurun_out/r2_prof_c3 python bench.py --config C3 --n 64 --steps 3 --warmup 3 --kernel-only > gpurun_out/r2_ncu_c3.log 2>&1; echo "ncu c3 rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:batched -s 3 -c 1 -f -o gpurun_out/r2_prof_c4 python bench.py --config C4 --n 64 --steps 3 --warmup 3 --kernel-only > gpurun_out/r2_ncu_c4.log 2>&1; echo "ncu c4 rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_bench_c2.csv python bench.py --steps 2 --warmup 1 --kernel-only > gpurun_out/r2_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 500 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "default_path or tiled_scatter_2d or (C3 and 8)" > gpurun_out/r2_sanitizer_memcheck.txt 2>&1; echo "memcheck rc=$?"; tail -4 gpurun_out/r2_sanitizer_memcheck.txt
timeout 900 python bench.py > gpurun_out/r2_bench2.log 2> gpurun_out/r2_bench2.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/r2_bench2.log
