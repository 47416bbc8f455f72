#!/bin/bash
# ncu full captures of the batched kernel: NVRTC-specialised (C4 64^3, C3 32^3) and ahead-of-time (C5 64^3)
mkdir -p gpurun_out
for c in C4 C3 C5; do n=64; [ $c = C3 ] && n=32; k=mrh_jit_batched; [ $c = C5 ] && k=batched_assembly
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 2 -c 1 -f -o gpurun_out/prof_fin_${c} \
    python scripts/bench_configs.py $c $n > gpurun_out/ncu_fin_${c}.log 2>&1; echo "ncu $c rc=$?"
done
echo done
