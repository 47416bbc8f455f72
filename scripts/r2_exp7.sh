#!/bin/bash
# round-2 experiment 7 (1 GPU): pipelined 2-D loop, register splits with two points per trip, atoms profile, full suite, bench lines
mkdir -p gpurun_out
export MRH_NO_REBUILD=1
T=gpurun_out/r2_e7_timing.txt; : > $T
P='import sys,json; d=json.loads(sys.stdin.read()); print(round(d["ms_per_step"],4), round(d.get("volume_kernel_ms",0),4), d.get("path"))'
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "tiled_scatter_2d or thermal_quad or default_path or ragged" > gpurun_out/r2_e7_parity.log 2>&1; echo "parity subset rc=$?"; tail -3 gpurun_out/r2_e7_parity.log
run() { echo "$1" >> $T; shift; env "$@" 2>&1 | tail -1 | python -c "$P" >> $T 2>&1; }
run "C1 default (pipelined)" timeout 120 python bench.py --config C1 --steps 10 --warmup 3 --kernel-only
run "C1 pipelined MINB2D=4" MRH_JIT_MINB2D=4 timeout 120 python bench.py --config C1 --steps 10 --warmup 3 --kernel-only
run "C1 nopipe" MRH_JIT_DEFINE=MRH_Q2D_NOPIPE timeout 120 python bench.py --config C1 --steps 10 --warmup 3 --kernel-only
run "C1 nopipe MINB2D=4" MRH_JIT_DEFINE=MRH_Q2D_NOPIPE MRH_JIT_MINB2D=4 timeout 120 python bench.py --config C1 --steps 10 --warmup 3 --kernel-only
run "C1 no-jit" MRH_BENCH_NOJIT=1 timeout 120 python bench.py --config C1 --steps 10 --warmup 3 --kernel-only
run "C2 default (bulk, qunroll 2, 200/96)" timeout 120 python bench.py --steps 10 --warmup 3 --kernel-only
for v in "208 88" "216 64" "192 112"; do set -- $v
  run "C2 qunroll 2 CREG=$1 SREG=$2" MRH_WS_CREG=$1 MRH_WS_SREG=$2 timeout 120 python bench.py --steps 10 --warmup 3 --kernel-only
done
run "C2 qunroll 1" MRH_WS_QUNROLL=1 timeout 120 python bench.py --steps 10 --warmup 3 --kernel-only
run "C3 default" timeout 180 python bench.py --config C3 --steps 5 --warmup 3 --kernel-only
run "C4 default" timeout 180 python bench.py --config C4 --steps 5 --warmup 3 --kernel-only
run "C5 default" timeout 180 python bench.py --config C5 --steps 5 --warmup 3 --kernel-only
cat $T
MRH_BAT_ATOMS=1 timeout 400 ncu --set full --clock-control none --import-source on -k regex:batched -s 3 -c 1 -f -o gpurun_out/r2_prof_c4_atoms python bench.py --config C4 --n 64 --steps 3 --warmup 3 --kernel-only > gpurun_out/r2_ncu_c4a.log 2>&1; echo "ncu c4 atoms rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:thermal_hex -s 3 -c 1 -f -o gpurun_out/r2_prof_c1d python bench.py --config C1 --steps 3 --warmup 3 --kernel-only > gpurun_out/r2_ncu_c1d.log 2>&1; echo "ncu c1 rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:thermal_hex -s 3 -c 1 -f -o gpurun_out/r2_prof_c2 python bench.py --steps 3 --warmup 3 --kernel-only > gpurun_out/r2_ncu_c2.log 2>&1; echo "ncu c2 rc=$?"
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_e7_pytest.log 2>&1; echo "full gpu suite rc=$?"; tail -5 gpurun_out/r2_e7_pytest.log
timeout 900 python bench.py > gpurun_out/r2_bench3.log 2> gpurun_out/r2_bench3.err; echo "bench rc=$?"; tail -c 600 gpurun_out/r2_bench3.log
timeout 300 python bench.py --impl reference > gpurun_out/r2_bench3_ref.log 2>&1; tail -c 400 gpurun_out/r2_bench3_ref.log
