"""Quick device check of the warp-specialised kernel against the plain tiled kernel (tuning aid)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch, problems
from mrhyde_b200 import Engine
for n in (4, 12, 40):
    out = {}
    for no_ws in (True, False):
        e = Engine.from_problem(problems.config_c2(n), device=0, no_ws=no_ws)
        s = e.sizes(); xyz, var = e.dof_coords()
        u = torch.from_numpy(problems.synthetic_state(xyz, var)).cuda()
        res = torch.zeros(s["nlocal"], dtype=torch.float64, device="cuda"); jac = torch.zeros(s["nnz"], dtype=torch.float64, device="cuda")
        e.assemble_jacres(u, res, jac, stream=torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        out[no_ws] = (res.clone(), jac.clone(), e.kernel_path())
    dr = (out[True][0] - out[False][0]).abs().max().item() / out[True][0].abs().max().item()
    dj = (out[True][1] - out[False][1]).abs().max().item() / out[True][1].abs().max().item()
    print(n, out[True][2], out[False][2], "rel diff res", dr, "jac", dj, flush=True)
