#!/bin/bash
# A/B of the NVRTC used for the run-time specialised kernels: CUDA toolkit 12.9 (default) against the PyTorch-bundled 12.8
mkdir -p gpurun_out
NV128=$(python -c "import nvidia.cuda_nvrtc, os; print(os.path.join(os.path.dirname(nvidia.cuda_nvrtc.__file__), 'lib', 'libnvrtc.so.12'))")
for lib in default $NV128; do
  [ $lib = default ] && unset MRH_NVRTC_LIB || export MRH_NVRTC_LIB=$lib
  echo "== NVRTC: $lib"
  python bench.py --steps 10 --warmup 3 --kernel-only 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('C2', d['ms_per_step'], d['value'])"
  python scripts/bench_configs.py C4 192 | cut -c1-200
  python scripts/bench_configs.py C3 128 | cut -c1-200
  python scripts/bench_configs.py C1 4096 | cut -c1-200
done
