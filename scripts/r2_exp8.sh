#!/bin/bash
# round-2 experiment 8 (N GPUs): engine + NCCL parity, SMs reserved for the overlapped exchange, weak / strong lines
mkdir -p gpurun_out
export MRH_NO_REBUILD=1
N=${1:-2}
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -x > gpurun_out/r2_e8_multi_$N.log 2>&1; echo "multi rc=$?"; tail -4 gpurun_out/r2_e8_multi_$N.log
for sm in 4 2 1 0; do
  MRH_XCH_SMS=$sm timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$sm bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2_e8_bench_${N}_sm$sm.log 2> gpurun_out/r2_e8_bench_${N}_sm$sm.err
  echo "xch sms=$sm rc=$?"; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_e8_bench_${N}_sm$sm.log').read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ('n_gpus','value','ms_per_step','gpu_launches')}); s=d.get('strong') or {}; print('strong', s.get('ms_per_step'), s.get('value')); print('e2e', d.get('e2e',{}).get('value'))
except Exception as e:
    print('parse failed', e); print(open('gpurun_out/r2_e8_bench_${N}_sm$sm.err').read()[-1500:])
PY
done
