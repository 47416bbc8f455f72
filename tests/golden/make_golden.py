#!/usr/bin/env python3
"""Extract the reference's own golden numbers for the assembly path into tests/golden/reference_golds.json.

Run in the build container (where /root/reference exists):  python tests/golden/make_golden.py
Sources (all under /root/reference/regression):
  discretization/HGRAD/mrhyde.gold            basis values / physical gradients on the unit quad and hex (6 digits)
  thermal/2D_verification, thermal/3D_verification, cdr/2D_manufactured, le/3D_manufactured,
  le/2D_manufactured, navierstokes/channel, thermal/2D_mixed_bcs (Neumann sides), thermal/2D_verification_nonzeroDBC
  (L2-projected Dirichlet data)                printed L2 errors (6 digits) of the solved problems
  thermal/2D_verification_transient            the L2 error after every one of its 20 backward-Euler steps
Only numbers are extracted; no reference source is copied.
"""
import json
import os
import re

REF = "/root/reference/regression"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_golds.json")


def hgrad():
    lines = open(f"{REF}/discretization/HGRAD/mrhyde.gold").read().splitlines()
    vals, grads = [], []
    for ln in lines:
        m = re.match(r"dof (\d+), point (\d+): ([-\d.e+]+)$", ln)
        if m:
            vals.append((int(m.group(1)), int(m.group(2)), float(m.group(3))))
        m = re.match(r"dof (\d+), point (\d+) grad: \(([^)]*)\)", ln)
        if m:
            grads.append((int(m.group(1)), int(m.group(2)), [float(x) for x in m.group(3).split(",")]))
    # file order: quad projection (16), quad Laplace (16), hex projection (64), hex Laplace (64)
    assert len(vals) == 16 + 64 and len(grads) == 16 + 64, (len(vals), len(grads))
    return {"quad_basis": vals[:16], "quad_grad": grads[:16], "hex_basis": vals[16:], "hex_grad": grads[16:]}


def l2(path):
    out = {}
    for ln in open(f"{REF}/{path}/mrhyde.gold"):
        m = re.search(r"L2 norm of the error for (\S+) = ([-\d.e+]+)", ln)
        if m:
            out[m.group(1)] = float(m.group(2))
    return out


def l2_series(path):
    """transient cases print one line per recorded time: [(time, error), ...] per variable"""
    out = {}
    for ln in open(f"{REF}/{path}/mrhyde.gold"):
        m = re.search(r"L2 norm of the error for (\S+) = ([-\d.e+]+)\s+\(time = ([-\d.e+]+)\)", ln)
        if m:
            out.setdefault(m.group(1), []).append((float(m.group(3)), float(m.group(2))))
    return out


if __name__ == "__main__":
    data = {"hgrad": hgrad(), "l2": {}}
    for p in ["thermal/2D_verification", "thermal/3D_verification", "cdr/2D_manufactured", "le/3D_manufactured",
              "le/2D_manufactured", "navierstokes/channel", "thermal/2D_mixed_bcs", "thermal/2D_verification_nonzeroDBC",
              "stokes/2D_verification_pspg", "helmholtz/manufactured_solution"]:
        data["l2"][p] = l2(p)
    data["l2_series"] = {"thermal/2D_verification_transient": l2_series("thermal/2D_verification_transient"),
                         "cdr/2D_transient": l2_series("cdr/2D_transient")}
    json.dump(data, open(OUT, "w"), indent=1)
    print("wrote", OUT, {k: v for k, v in data["l2"].items()})
