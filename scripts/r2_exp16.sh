#!/bin/bash
# round-2 last run (1 GPU): the new cdr transient device test, then the whole GPU suite
mkdir -p gpurun_out
export MRH_NO_REBUILD=1
timeout 200 python -m pytest tests/test_gpu_newton.py -m gpu -q -x -k "cdr_transient" > gpurun_out/r2_e16_cdr.log 2>&1; echo "cdr transient rc=$?"; tail -4 gpurun_out/r2_e16_cdr.log
timeout 560 python -m pytest tests -m gpu -x -q > gpurun_out/r2_e16_pytest.log 2>&1; echo "full gpu suite rc=$?"; tail -4 gpurun_out/r2_e16_pytest.log
