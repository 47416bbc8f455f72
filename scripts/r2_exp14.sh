#!/bin/bash
# round-2 final check (1 GPU): whole GPU suite, smoke, bench lines (own arm + reference arm), launch list
mkdir -p gpurun_out
export MRH_NO_REBUILD=1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_e14_pytest.log 2>&1; echo "full gpu suite rc=$?"; tail -4 gpurun_out/r2_e14_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/r2_bench7.log 2> gpurun_out/r2_bench7.err; echo "bench rc=$?"; tail -c 400 gpurun_out/r2_bench7.log
timeout 300 python bench.py --impl reference > gpurun_out/r2_bench7_ref.log 2>&1; tail -c 300 gpurun_out/r2_bench7_ref.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_bench_c2.csv python bench.py --steps 2 --warmup 1 --kernel-only > gpurun_out/r2_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
