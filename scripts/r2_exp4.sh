#!/bin/bash
# round-2 experiment 4 (1 GPU): new kernels first (short timeouts), then the whole GPU suite, then timings
mkdir -p gpurun_out
export MRH_NO_REBUILD=1
T=gpurun_out/r2_e4_timing.txt; : > $T
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "tiled_scatter_2d or thermal_quad" > gpurun_out/r2_e4_c1tiled.log 2>&1; echo "c1 tiled parity rc=$?"; tail -3 gpurun_out/r2_e4_c1tiled.log
for v in "1 3" "1 4" "1 5" "0 3"; do set -- $v
  echo "C1 TILED2D=$1 MINB2D=$2" >> $T
  MRH_TILED2D=$1 MRH_JIT_MINB2D=$2 timeout 120 python bench.py --config C1 --steps 10 --warmup 3 --kernel-only 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['config'].get('volume_kernel'))" >> $T 2>&1
done
MRH_WS_BULK=1 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "ragged_meshes or default_path or mid_size_c2 or host_buffer" > gpurun_out/r2_e4_bulk.log 2>&1; echo "bulk parity rc=$?"; tail -3 gpurun_out/r2_e4_bulk.log
for v in "MRH_JIT_EXTRA=0" "MRH_WS_BULK"; do
  echo "C2 define $v" >> $T
  MRH_JIT_DEFINE=$v timeout 120 python bench.py --steps 10 --warmup 3 --kernel-only 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['config'].get('volume_kernel'))" >> $T 2>&1
done
MRH_WSK=1 timeout 120 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "default_path or mid_size_c2 or ragged_meshes" > gpurun_out/r2_e4_wsk.log 2>&1; rc=$?; echo "wsk parity rc=$rc"; tail -3 gpurun_out/r2_e4_wsk.log
if [ $rc = 0 ]; then
  for v in "216 64" "224 48" "208 88"; do set -- $v
    echo "C2 WSK CREG=$1 SREG=$2" >> $T
    MRH_WSK=1 MRH_WSK_CREG=$1 MRH_WSK_SREG=$2 timeout 90 python bench.py --steps 10 --warmup 3 --kernel-only 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['config'].get('volume_kernel'))" >> $T 2>&1
  done
fi
for a in 1 0; do for c in C4 C3; do
  echo "$c ATOMS=$a" >> $T
  MRH_BAT_ATOMS=$a timeout 180 python bench.py --config $c --steps 5 --warmup 3 --kernel-only 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['config'].get('volume_kernel'))" >> $T 2>&1
done; done
cat $T
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_e4_pytest.log 2>&1; echo "full gpu suite rc=$?"; tail -5 gpurun_out/r2_e4_pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
