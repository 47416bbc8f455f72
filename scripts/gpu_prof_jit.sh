#!/bin/bash
# ncu full captures of the NVRTC-specialised batched kernel (C4 64^3, C3 32^3) and the launch list of one C4 assembly
mkdir -p gpurun_out
for c in C4 C3; do n=64; [ $c = C3 ] && n=32
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mrh_jit_batched -s 2 -c 1 -f -o gpurun_out/prof_jit_${c} \
    python scripts/bench_configs.py $c $n > gpurun_out/ncu_jit_${c}.log 2>&1; echo "ncu $c rc=$?"
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_c4.csv python scripts/bench_configs.py C4 64 > gpurun_out/ncu_l.log 2>&1
echo done
