"""Summarise an `ncu --page source --csv` export: samples and executed instructions per block of N SASS lines,
plus the hottest individual instructions.  usage: ncu_regions.py src.csv [block=80]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
blk = int(sys.argv[2]) if len(sys.argv) > 2 else 80
hdr = rows[1]; body = rows[2:]
iS, iX, iSrc = hdr.index('# Samples'), hdr.index('Instructions Executed'), hdr.index('Source')
iW = hdr.index('L1 Wavefronts Shared') if 'L1 Wavefronts Shared' in hdr else None
iWi = hdr.index('L1 Wavefronts Shared Ideal') if 'L1 Wavefronts Shared Ideal' in hdr else None
tot = sum(int(r[iS]) for r in body); totx = sum(int(r[iX]) for r in body)
print("instructions", len(body), "samples", tot, "warp-instr executed", totx)
for b in range(0, len(body), blk):
    seg = body[b:b + blk]
    s = sum(int(r[iS]) for r in seg); x = sum(int(r[iX]) for r in seg)
    if s == 0 and x == 0: continue
    wf = sum(int(r[iW] or 0) for r in seg) if iW is not None else 0; wi = sum(int(r[iWi] or 0) for r in seg) if iWi is not None else 0
    ops = {}
    for r in seg:
        op = r[iSrc].split()[1 if r[iSrc].strip().startswith('@') else 0].split('.')[0]
        ops[op] = ops.get(op, 0) + int(r[iX])
    top = sorted(ops.items(), key=lambda kv: -kv[1])[:4]
    print(f"{b:5d}-{b+len(seg):5d} samples {s:7d} ({100*s/tot:5.1f}%) exec {x:10d} ({100*x/totx:5.1f}%)  " + " ".join(f"{k}:{v*100//max(x,1)}%" for k, v in top) + (f"  smem wf {wf} (ideal {wi})" if wf else ""))
print("hottest:")
for r in sorted(body, key=lambda r: -int(r[iS]))[:25]:
    print(f"  {body.index(r):5d} {int(r[iS]):6d} {r[iSrc].strip()[:90]}")
