#!/bin/bash
# round-2 experiment 1: new host-side pieces (facade, Newton) + the lane-per-point thermal kernel (parity, then timing variants)
mkdir -p gpurun_out
export MRH_NO_REBUILD=1     # the library built in the container travels with the snapshot
timeout 900 python -m pytest tests/test_facade_cpp.py tests/test_gpu_newton.py -m gpu -q -x > gpurun_out/r2_e1_new.log 2>&1; echo "new tests rc=$?"; tail -25 gpurun_out/r2_e1_new.log
MRH_WS8=1 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "thermal or C2 or ragged or large_mesh or default_path or opcode or set_nodes or host_buffer" > gpurun_out/r2_e1_ws8.log 2>&1; echo "ws8 parity rc=$?"; tail -8 gpurun_out/r2_e1_ws8.log
for v in "0 3 136 96" "1 3 136 96" "1 3 144 80" "1 2 200 96" "1 2 168 96"; do
  set -- $v
  echo "WS8=$1 NCG=$2 CREG=$3 SREG=$4"
  MRH_WS8=$1 MRH_WS8_NCG=$2 MRH_WS8_CREG=$3 MRH_WS8_SREG=$4 timeout 300 python bench.py --steps 10 --warmup 3 --kernel-only 2>&1 | tail -1
done | tee gpurun_out/r2_e1_timing.txt
