"""Timing of the other BASELINE.json configurations (C1, C3, C4, C5) through the generic fused kernel.
They are parity-test cases, not bench lines; this script records where the generic kernel stands against the
SURVEY.md 8(d) rooflines.  usage: python scripts/bench_configs.py <C1|C3|C4|C5> [n]"""
import os, sys, json, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import problems
from mrhyde_b200 import Engine

FULL = {"C1": (problems.config_c1, 4096, 120.0, 800.0), "C3": (problems.config_c3, 128, 4370.0, 213e3),
        "C4": (problems.config_c4, 192, 2123.0, 36.5e3), "C5": (problems.config_c5, 128, 3701.0, 150e3)}
name = sys.argv[1]
mk, nfull, B, F = FULL[name]
n = int(sys.argv[2]) if len(sys.argv) > 2 else nfull
t0 = time.time()
e = Engine.from_problem(mk(n), device=0)
s = e.sizes()
xyz, var = e.dof_coords()
dev = torch.device("cuda", 0)
u = torch.from_numpy(problems.synthetic_state(xyz, var)).to(dev)
res = torch.zeros(s["nlocal"], dtype=torch.float64, device=dev)
jac = torch.zeros(s["nnz"], dtype=torch.float64, device=dev)
setup = time.time() - t0
st = torch.cuda.current_stream().cuda_stream
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for _ in range(2):
    e.assemble_jacres(u, res, jac, stream=st)
torch.cuda.synchronize()
K = 5
ev0.record()
for _ in range(K):
    e.assemble_jacres(u, res, jac, stream=st)
ev1.record(); torch.cuda.synchronize()
ms = ev0.elapsed_time(ev1) / K
rate = s["nelem"] / (ms * 1e-3)
# size-independent sanity: repeated assembly accumulates (7 assemblies so far), finite values
assert torch.isfinite(res).all() and torch.isfinite(jac).all()
print(json.dumps({"config": name, "n": n, "nelem": s["nelem"], "nlocal": s["nlocal"], "nnz": s["nnz"], "path": e.kernel_path(),
                  "ms_per_assembly": ms, "elements_per_s": rate, "launches": e.launch_count(), "setup_s": round(setup, 1),
                  "fp64_TFs": F * rate / 1e12, "hbm_GBs": B * rate / 1e9, "patterns": s["npatterns"]}), flush=True)
