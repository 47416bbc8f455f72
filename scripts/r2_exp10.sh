#!/bin/bash
# round-2 experiment 10 (N = 4 or 8 GPUs): engine + NCCL parity on the 2x2x1 / 2x2x2 grids, then the weak + strong bench line
mkdir -p gpurun_out
export MRH_NO_REBUILD=1
N=${1:-8}
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q -x > gpurun_out/r2_e10_multi_$N.log 2>&1; echo "multi rc=$?"; tail -4 gpurun_out/r2_e10_multi_$N.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2_e10_bench_$N.log 2> gpurun_out/r2_e10_bench_$N.err
echo "bench $N rc=$?"; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_e10_bench_$N.log').read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ('n_gpus','value','ms_per_step','gpu_launches')}); print('strong', d.get('strong')); print('e2e', d.get('e2e',{}).get('value'))
except Exception as e:
    print('parse failed', e); print(open('gpurun_out/r2_e10_bench_$N.err').read()[-2500:])
PY
