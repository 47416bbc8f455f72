#!/bin/bash
# A/B of the launch bounds (register budget) of the NVRTC-specialised batched kernel
for b in "" "128,4" "160,2" "128,3"; do
  [ -z "$b" ] && unset MRH_BAT_BOUNDS || export MRH_BAT_BOUNDS=$b
  echo "== MRH_BAT_BOUNDS=${b:-default}"
  python scripts/bench_configs.py C3 128 | cut -c1-190
  python scripts/bench_configs.py C4 192 | cut -c1-190
done
