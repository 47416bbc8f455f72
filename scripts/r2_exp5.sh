#!/bin/bash
# round-2 experiment 5 (1 GPU): timings of the kernel variants (A/B inside one build), Q2 parity with the DMMA-built fragments
mkdir -p gpurun_out
export MRH_NO_REBUILD=1
T=gpurun_out/r2_e5_timing.txt; : > $T
P='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["ms_per_step"],4), round(d.get("volume_kernel_ms",0),4), d.get("path"))'
timeout 400 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "C3 or cdr or Q2 or parameter or non_affine or foreign" > gpurun_out/r2_e5_q2.log 2>&1; echo "q2 + parameter parity rc=$?"; tail -3 gpurun_out/r2_e5_q2.log
run() { echo "$1" >> $T; shift; env "$@" 2>&1 | tail -1 | python -c "$P" >> $T 2>&1; }
for v in "1 3" "1 4" "1 5" "0 3"; do set -- $v
  run "C1 TILED2D=$1 MINB2D=$2" MRH_TILED2D=$1 MRH_JIT_MINB2D=$2 timeout 120 python bench.py --config C1 --steps 10 --warmup 3 --kernel-only
done
for v in "MRH_JIT_EXTRA=0" "MRH_WS_BULK"; do
  run "C2 define $v" MRH_JIT_DEFINE=$v timeout 120 python bench.py --steps 10 --warmup 3 --kernel-only
done
for v in 2 4; do
  run "C2 QUNROLL=$v" MRH_JIT_DEFINE=MRH_WS_QUNROLL=$v timeout 120 python bench.py --steps 10 --warmup 3 --kernel-only
done
for v in "216 64" "224 48" "208 88"; do set -- $v
  run "C2 WSK CREG=$1 SREG=$2" MRH_WSK=1 MRH_WSK_CREG=$1 MRH_WSK_SREG=$2 timeout 90 python bench.py --steps 10 --warmup 3 --kernel-only
done
run "C3 frag=dmma atoms=1" timeout 180 python bench.py --config C3 --steps 5 --warmup 3 --kernel-only
run "C3 frag=gather atoms=1" MRH_BAT_DEFINE=MRH_BAT_FRAG_GATHER timeout 180 python bench.py --config C3 --steps 5 --warmup 3 --kernel-only
run "C3 frag=dmma atoms=0" MRH_BAT_ATOMS=0 timeout 180 python bench.py --config C3 --steps 5 --warmup 3 --kernel-only
run "C4 atoms=1" timeout 180 python bench.py --config C4 --steps 5 --warmup 3 --kernel-only
run "C4 atoms=0" MRH_BAT_ATOMS=0 timeout 180 python bench.py --config C4 --steps 5 --warmup 3 --kernel-only
run "C5" timeout 180 python bench.py --config C5 --steps 5 --warmup 3 --kernel-only
cat $T
# ncu of the new default C1 and C3 kernels (one launch each)
timeout 300 ncu --set full --clock-control none --import-source on -k regex:thermal_hex -s 3 -c 1 -f -o gpurun_out/r2_prof_c1 python bench.py --config C1 --steps 3 --warmup 3 --kernel-only > gpurun_out/r2_ncu_c1.log 2>&1; echo "ncu c1 rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:batched -s 3 -c 1 -f -o gpurun_out/r2_prof_c3 python bench.py --config C3 --n 64 --steps 3 --warmup 3 --kernel-only > gpurun_out/r2_ncu_c3.log 2>&1; echo "ncu c3 rc=$?"
