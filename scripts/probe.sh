#!/bin/bash
# compile-only probe of the NVRTC kernels (no GPU): refresh embedded headers, rebuild the probe, print resource usage of one entry
python -c "
import sys; sys.path.insert(0,'.')
from mrhyde_b200 import build; build.write_embedded_headers()"
g++ -O1 -std=c++17 -I/usr/local/cuda/include tools/jit_probe.cpp -o /tmp/jit_probe -L/usr/local/cuda/lib64 -lnvrtc -ldl -Wl,-rpath,/usr/local/cuda/lib64 || exit 1
/tmp/jit_probe > /tmp/jit_probe.log 2>&1; grep "rc=" /tmp/jit_probe.log; grep -B2 -A8 "error" /tmp/jit_probe.log | head -40
grep -A3 "entry function '${1:-kws8}'" /tmp/jit_probe.log
cuobjdump -sass /tmp/jit_probe.cubin | awk "/Function : ${1:-kws8}\$/,/\.\.\.\.\.\.\.\.\.\./" > /tmp/probe.sass
echo "lines $(wc -l < /tmp/probe.sass) local $(grep -c 'STL\|LDL' /tmp/probe.sass) shfl $(grep -c SHFL /tmp/probe.sass) fp64 $(grep -c 'DFMA\|DMUL\|DADD' /tmp/probe.sass) sel $(grep -c 'SEL' /tmp/probe.sass) lds $(grep -c 'LDS' /tmp/probe.sass) sts $(grep -c 'STS' /tmp/probe.sass) mufu $(grep -c MUFU /tmp/probe.sass)"
