#!/bin/bash
# round-2 experiment 15 (4 GPUs): SMs reserved for the overlapped exchange under STRONG scaling (256^3 cut 2x2x1, 3 neighbours per rank)
mkdir -p gpurun_out
export MRH_NO_REBUILD=1
N=${1:-4}
T=gpurun_out/r2_e15_timing.txt; : > $T
for sm in 0 2 4 8; do
  echo "strong N=$N MRH_XCH_SMS=$sm" >> $T
  MRH_XCH_SMS=$sm timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2956$sm bench.py --gpus $N --scaling strong --steps 20 --warmup 5 --kernel-only 2> gpurun_out/r2_e15_sm$sm.err | tail -1 >> $T
done
echo "strong N=$N no fused exchange (serial, after the launch)" >> $T
MRH_XCH_SMS=0 MRH_BENCH_FLAGS=128 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29569 bench.py --gpus $N --scaling strong --steps 20 --warmup 5 --kernel-only 2>> gpurun_out/r2_e15_sm0.err | tail -1 >> $T
cat $T
