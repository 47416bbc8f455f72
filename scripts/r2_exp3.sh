#!/bin/bash
# round-2 experiment 3 (2+ GPUs): engine + NCCL exchange against the single-GPU engine, then weak / strong scaling lines
mkdir -p gpurun_out
export MRH_NO_REBUILD=1
N=${1:-2}
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -x > gpurun_out/r2_e3_multi_$N.log 2>&1; echo "multi rc=$?"; tail -8 gpurun_out/r2_e3_multi_$N.log
for n in 1 $N; do
  if [ $n = 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 --no-other > gpurun_out/r2_e3_bench_$n.log 2> gpurun_out/r2_e3_bench_$n.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/r2_e3_bench_$n.log 2> gpurun_out/r2_e3_bench_$n.err
  fi
  echo "bench $n rc=$?"; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_e3_bench_$n.log').read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ('n_gpus','value','ms_per_step','gpu_launches','scaling')}); print('strong', d.get('strong')); print('e2e', d.get('e2e',{}).get('value'))
except Exception as e:
    print('parse failed', e); print(open('gpurun_out/r2_e3_bench_$n.err').read()[-1500:])
PY
done
