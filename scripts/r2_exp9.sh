#!/bin/bash
# round-2 experiment 9 (1 GPU): helmholtz on the device, table-driven atoms (parity + timing), final defaults, full suite, bench lines
mkdir -p gpurun_out
export MRH_NO_REBUILD=1
T=gpurun_out/r2_e9_timing.txt; : > $T
P='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["ms_per_step"],4), round(d.get("volume_kernel_ms",0),4), d.get("path"))'
timeout 400 python -m pytest tests/test_gpu_parity.py tests/test_gpu_newton.py -m gpu -q -x -k "helmholtz" > gpurun_out/r2_e9_helmholtz.log 2>&1; echo "helmholtz rc=$?"; tail -3 gpurun_out/r2_e9_helmholtz.log
MRH_BAT_ATOMS=1 timeout 500 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "nvrtc or batched_kernel_matches or jacobian_is_the_derivative" > gpurun_out/r2_e9_atoms.log 2>&1; echo "atoms parity rc=$?"; tail -3 gpurun_out/r2_e9_atoms.log
run() { echo "$1" >> $T; shift; env "$@" 2>&1 | tail -1 | python -c "$P" >> $T 2>&1; }
run "C1 default (128 regs, no pipeline)" timeout 120 python bench.py --config C1 --steps 10 --warmup 3 --kernel-only
run "C1 MINB2D=5" MRH_JIT_MINB2D=5 timeout 120 python bench.py --config C1 --steps 10 --warmup 3 --kernel-only
run "C2 default (bulk, qunroll 2, 208/88)" timeout 120 python bench.py --steps 10 --warmup 3 --kernel-only
run "C3 atoms=0" timeout 180 python bench.py --config C3 --steps 5 --warmup 3 --kernel-only
run "C3 atoms=1 (table)" MRH_BAT_ATOMS=1 timeout 180 python bench.py --config C3 --steps 5 --warmup 3 --kernel-only
run "C4 atoms=0" timeout 180 python bench.py --config C4 --steps 5 --warmup 3 --kernel-only
run "C4 atoms=1 (table)" MRH_BAT_ATOMS=1 timeout 180 python bench.py --config C4 --steps 5 --warmup 3 --kernel-only
cat $T
MRH_BAT_ATOMS=1 timeout 400 ncu --set full --clock-control none --import-source on -k regex:batched -s 3 -c 1 -f -o gpurun_out/r2_prof_c4_atoms2 python bench.py --config C4 --n 64 --steps 3 --warmup 3 --kernel-only > gpurun_out/r2_ncu_c4a2.log 2>&1; echo "ncu c4 atoms rc=$?"
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_e9_pytest.log 2>&1; echo "full gpu suite rc=$?"; tail -5 gpurun_out/r2_e9_pytest.log
timeout 900 python bench.py > gpurun_out/r2_bench4.log 2> gpurun_out/r2_bench4.err; echo "bench rc=$?"; tail -c 600 gpurun_out/r2_bench4.log
