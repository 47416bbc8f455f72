#!/bin/bash
# round-2 experiment 2: K-in-registers thermal kernel (parity + timing variants), ncu of ws8 / wsk / ws, new tests, first full bench line
mkdir -p gpurun_out
export MRH_NO_REBUILD=1     # the library built in the container travels with the snapshot
timeout 900 python -m pytest tests/test_gpu_newton.py tests/test_facade_cpp.py -m gpu -q > gpurun_out/r2_e2_newton.log 2>&1; echo "newton+facade rc=$?"; tail -6 gpurun_out/r2_e2_newton.log
MRH_WSK=1 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "thermal or C2 or ragged or default_path or set_nodes or host_buffer" > gpurun_out/r2_e2_wsk.log 2>&1; echo "wsk parity rc=$?"; tail -4 gpurun_out/r2_e2_wsk.log
for v in "216 80" "224 64" "232 48"; do
  set -- $v
  echo "WSK CREG=$1 SREG=$2"
  MRH_WSK=1 MRH_WSK_CREG=$1 MRH_WSK_SREG=$2 timeout 300 python bench.py --steps 10 --warmup 3 --kernel-only 2>&1 | tail -1
done | tee gpurun_out/r2_e2_timing.txt
for k in wsk ws8 ws; do
  env=""; [ $k = wsk ] && env="MRH_WSK=1"; [ $k = ws8 ] && env="MRH_WS8=1"
  env $env timeout 900 ncu --set full --clock-control none --import-source on -k regex:thermal_hex -s 3 -c 1 -f -o gpurun_out/r2_prof_$k \
      python bench.py --steps 3 --warmup 3 --kernel-only > gpurun_out/r2_ncu_$k.log 2>&1; echo "ncu $k rc=$?"
done
timeout 1500 python bench.py > gpurun_out/r2_bench1.log 2> gpurun_out/r2_bench1.err; echo "bench rc=$?"; tail -c 2500 gpurun_out/r2_bench1.log
timeout 200 python bench.py --impl reference > gpurun_out/r2_bench1_ref.log 2>&1; tail -c 600 gpurun_out/r2_bench1_ref.log
