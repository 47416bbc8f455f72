#!/bin/bash
# usage: tools/probe.sh  -> registers, spills, static SASS mix of the specialised kernel
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -lineinfo -Xptxas -v -cubin -o /tmp/probe.cubin tools/probe_fast.cu 2>&1 | grep -E "registers|spill|error"
echo "static instructions: $(cuobjdump -sass /tmp/probe.cubin | grep -cE '^\s+/\*[0-9a-f]{4}\*/')"
cuobjdump -sass /tmp/probe.cubin | grep -oE "\b(DFMA|DMUL|DADD|LDL|STL|LDS|STS|LDG|RED|BRA|IMAD|MUFU|LDC|ISETP|FSEL|LOP3|PRMT|LEA|IADD3)[A-Z0-9_.]*" | sed 's/\..*//' | sort | uniq -c | sort -rn | tr '\n' ' '; echo
