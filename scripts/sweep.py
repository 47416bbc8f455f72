"""Kernel-time sweep of the C2 volume kernel over engine variants (tuning aid; not the bench)."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
import problems
from mrhyde_b200 import Engine

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
variants = [dict(no_ws=False, creg="200", sreg="96"), dict(no_ws=False, creg="192", sreg="112"), dict(no_ws=False, creg="184", sreg="128"), dict(no_ws=False, creg="176", sreg="144")]
dev = torch.device("cuda", 0)
for v in variants:
    os.environ["MRH_WS_CREG"] = v["creg"]; os.environ["MRH_WS_SREG"] = v["sreg"]
    e = Engine.from_problem(problems.config_c2(n), device=0, no_ws=v["no_ws"])
    s = e.sizes()
    xyz, var = e.dof_coords()
    u = torch.from_numpy(problems.synthetic_state(xyz, var)).to(dev)
    res = torch.zeros(s["nlocal"], dtype=torch.float64, device=dev)
    jac = torch.zeros(s["nnz"], dtype=torch.float64, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    ms = []
    for i in range(6):
        e.assemble_jacres(u, res, jac, stream=st)
        ms.append(e.volume_kernel_ms())
    torch.cuda.synchronize()
    print(json.dumps({"n": n, **v, "path": e.kernel_path(), "ms": [round(x, 3) for x in ms[2:]], "checksum": float(res.abs().sum()), "jsum": float(jac.sum())}), flush=True)
    del e, u, res, jac
    torch.cuda.empty_cache()
