#!/bin/bash
# round-2 experiment 11 (1 GPU): whole GPU suite + smoke + bench lines with the final defaults
mkdir -p gpurun_out
export MRH_NO_REBUILD=1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_e11_pytest.log 2>&1; echo "full gpu suite rc=$?"; tail -5 gpurun_out/r2_e11_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
T=gpurun_out/r2_e11_timing.txt; : > $T
P='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["ms_per_step"],4), round(d.get("volume_kernel_ms",0),4), d.get("path"))'
run() { echo "$1" >> $T; shift; env "$@" 2>&1 | tail -1 | python -c "$P" >> $T 2>&1; }
run "C4 default (atoms auto)" timeout 180 python bench.py --config C4 --steps 5 --warmup 3 --kernel-only
run "C3 default (atoms auto)" timeout 180 python bench.py --config C3 --steps 5 --warmup 3 --kernel-only
cat $T
timeout 900 python bench.py > gpurun_out/r2_bench5.log 2> gpurun_out/r2_bench5.err; echo "bench rc=$?"; tail -c 700 gpurun_out/r2_bench5.log
timeout 300 python bench.py --impl reference > gpurun_out/r2_bench5_ref.log 2>&1; tail -c 300 gpurun_out/r2_bench5_ref.log
timeout 400 ncu --set full --clock-control none --import-source on -k regex:batched -s 3 -c 1 -f -o gpurun_out/r2_prof_c3b python bench.py --config C3 --n 64 --steps 3 --warmup 3 --kernel-only > gpurun_out/r2_ncu_c3b.log 2>&1; echo "ncu c3 rc=$?"
