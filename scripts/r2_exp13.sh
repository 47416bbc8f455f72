#!/bin/bash
# round-2 experiment 13 (1 GPU): explicit JAC / TR unswitching of the batched point phase (C5 regression), split batched TUs
mkdir -p gpurun_out
export MRH_NO_REBUILD=1
T=gpurun_out/r2_e13_timing.txt; : > $T
P='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["ms_per_step"],4), round(d.get("volume_kernel_ms",0),4), d.get("path"))'
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "test_cuda_matches_oracle or batched_kernel or nvrtc_specialised_batched or boundary or baseline_configs" > gpurun_out/r2_e13_parity.log 2>&1; echo "batched parity rc=$?"; tail -3 gpurun_out/r2_e13_parity.log
run() { echo "$1" >> $T; shift; env "$@" 2>&1 | tail -1 | python -c "$P" >> $T 2>&1; }
run "C5 (templated point phase)" timeout 180 python bench.py --config C5 --steps 10 --warmup 3 --kernel-only
run "C5 again" timeout 180 python bench.py --config C5 --steps 10 --warmup 3 --kernel-only
run "C4" timeout 180 python bench.py --config C4 --steps 5 --warmup 3 --kernel-only
run "C3" timeout 180 python bench.py --config C3 --steps 5 --warmup 3 --kernel-only
cat $T
