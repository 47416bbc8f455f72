"""Shared parity checker: one engine library (the CUDA product on a GPU box, or the CPU stepping harness of the
kernels in tests/emu) against the oracle on the same seeded inputs.

Bars: CRS rowptr/colind, element LIDs, offsets, fixed-DOF flags bit-exact; residual and Jacobian
values <= 1e-12 relative, measured per vector/matrix as max|a-b|/max|b| and additionally per row.
"""
import copy
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import problems  # noqa: E402
from oracle.pyoracle import Oracle, config_text  # noqa: E402

TOL = 1e-12
ROW_TOL = 1e-12     # per-row bar of SURVEY.md 8c (same as the global one)
EMU_LIB = os.path.join(HERE, "emu", "libmrhyde_emu.so")


def build_emu():
    src = os.path.join(HERE, "emu", "emu.cpp")
    csrc = os.path.join(ROOT, "mrhyde_b200", "csrc")
    deps = [src] + [os.path.join(csrc, f) for f in os.listdir(csrc)]
    if (not os.path.exists(EMU_LIB)) or any(os.path.getmtime(d) > os.path.getmtime(EMU_LIB) for d in deps):
        subprocess.check_call(["g++", "-O1", "-std=c++17", "-fPIC", "-shared", "-w", "-o", EMU_LIB, src])
    return EMU_LIB


def shrink(p, n3=4, n2=10):
    p = copy.deepcopy(p)
    m = p["mesh"]
    if m["dim"] == 3:
        m["NX"], m["NY"], m["NZ"] = n3, n3 + 1, n3 - 1
    else:
        m["NX"], m["NY"] = n2, n2 - 3
    return p


def cases():
    """(id, problem, kwargs) — kwargs: transient=dict|None"""
    P = problems
    out = []
    for mk in (P.regression_thermal_2d, P.regression_thermal_3d, P.regression_cdr_2d, P.regression_le_2d,
               P.regression_le_3d, P.regression_ns_channel):
        p = shrink(mk())
        out.append((p["name"], p, {}))
    out.append(("C3 cdr Q2 hex", shrink(P.config_c3(3), 3), {}))
    out.append(("C5 ns supg pspg hex", shrink(P.config_c5(4)), {}))
    # thermal Q2 quad, non-default quadrature
    p = shrink(P.regression_thermal_2d()); p["order"] = 2; p["name"] = "thermal 2D Q2"
    out.append((p["name"], p, {}))
    # boundary terms: weak Dirichlet (Nitsche), Neumann, Flux
    p = shrink(P.regression_thermal_3d()); p["extra"] = {"use weak Dirichlet": True}; p["name"] = "thermal 3D weak Dirichlet"
    p["dirichlet"] = {"e": {"all boundaries": "0.1*x+y*z"}}
    out.append((p["name"], p, {}))
    p = shrink(P.regression_thermal_2d()); p["name"] = "thermal 2D Neumann+Flux+Dirichlet"
    p["dirichlet"] = {"e": {"left": "0.0", "bottom": "sin(x)"}}
    p["neumann"] = {"e": {"right": "1.0+y"}}
    p["flux"] = {"e": {"top": "0.5*x"}}
    out.append((p["name"], p, {}))
    p = shrink(P.regression_le_3d()); p["extra"] = {"use weak Dirichlet": True}; p["name"] = "LE 3D weak Dirichlet + Neumann"
    p["dirichlet"] = {v: {"bottom": "0.01*x", "left": "0.0"} for v in ("dx", "dy", "dz")}
    p["neumann"] = {"dz": {"top": "0.3"}, "dx": {"right": "0.1*y"}}
    out.append((p["name"], p, {}))
    p = shrink(P.regression_le_2d()); p["extra"] = {"use weak Dirichlet": True}; p["name"] = "LE 2D weak Dirichlet"
    out.append((p["name"], p, {}))
    p = shrink(P.regression_ns_channel()); p["name"] = "NS 2D Neumann outflow + SUPG"
    p["settings"] = {"usePSPG": True, "useSUPG": True}
    p["neumann"] = {"ux": {"right": "0.2"}, "uy": {"right": "0.1*y"}}
    out.append((p["name"], p, {}))
    # state-dependent coefficient in thermal (forces the generic kernel) and advection
    p = shrink(P.regression_thermal_3d()); p["name"] = "thermal 3D nonlinear diffusion + advection"
    p["functions"] = dict(p["functions"]); p["functions"].update({"thermal diffusion": "1.0+e*e", "advection x": "1.0", "advection y": "x", "advection z": "0.5"})
    p["settings"] = {"include advection": True}
    out.append((p["name"], p, {}))
    p = shrink(P.regression_thermal_3d()); p["name"] = "thermal 3D advection (specialised kernel)"
    p["functions"] = dict(p["functions"]); p["functions"].update({"thermal diffusion": "1.0+x*y", "advection x": "1.0", "advection y": "x", "advection z": "0.5"})
    p["settings"] = {"include advection": True}
    out.append((p["name"], p, {}))
    # transient (backward Euler and a 2-stage DIRK stage 1 with BDF2-like weights)
    be = dict(A=[[1.0]], b=[1.0], bdf=[1.0, -1.0], dt=0.05, time=0.3, stage=0)
    dirk = dict(A=[[0.25, 0.0], [0.5, 0.25]], b=[0.5, 0.5], bdf=[1.5, -2.0, 0.5], dt=0.02, time=0.1, stage=1)
    for nm, ti in (("BE", be), ("DIRK2 stage1", dirk)):
        p = shrink(P.regression_thermal_3d()); p["extra"] = {"transient": True}; p["name"] = "thermal 3D transient " + nm
        p["functions"] = dict(p["functions"]); p["functions"].update({"density": "2.0", "specific heat": "1.0+x", "thermal source": "sin(t)*x"})
        out.append((p["name"], p, {"transient": ti}))
    p = shrink(P.config_c5(4)); p["extra"] = {"transient": True}; p["name"] = "NS 3D transient BE supg/pspg"
    out.append((p["name"], p, {"transient": be}))
    p = shrink(P.regression_cdr_2d()); p["extra"] = {"transient": True, "workset size": 40}; p["name"] = "cdr 2D transient DIRK2"
    out.append((p["name"], p, {"transient": dirk}))
    # stokes: 2-D PSPG (h / 2 nu quirk), 3-D PSPG + LSIC with variable viscosity, 2-D Q2
    p = shrink(P.regression_stokes_2d(6)); p["name"] = "stokes 2D pspg"
    out.append((p["name"], p, {}))
    p = shrink(P.regression_stokes_2d(4)); p["name"] = "stokes 3D pspg+lsic"
    p["mesh"] = shrink(P.config_c5(4))["mesh"]; p["settings"] = {"usePSPG": True, "useLSIC": True}
    p["functions"] = {"source ux": "1.0", "source uz": "x*y", "viscosity": "1.0+0.5*x"}
    p["dirichlet"] = {v: {s: "0.0" for s in ("bottom", "top", "front")} for v in ("ux", "uy", "uz")}
    out.append((p["name"], p, {}))
    p = shrink(P.regression_stokes_2d(5)); p["name"] = "stokes 2D Q2"; p["order"] = 2; p["settings"] = {"useLSIC": True}
    out.append((p["name"], p, {}))
    # helmholtz: the reference's 2-D deck (Robin side on the right), and a 3-D case with every coefficient non-trivial
    p = shrink(P.regression_helmholtz_2d(8)); p["name"] = "helmholtz 2D robin"
    p["functions"] = dict(p["functions"]); p["functions"].update({"robin_alpha_r": "0.5+y", "robin_alpha_i": "0.25*x", "omega2i": "0.3"})
    out.append((p["name"], p, {}))
    p = shrink(P.regression_helmholtz_2d(4)); p["name"] = "helmholtz 3D"
    p["mesh"] = shrink(P.config_c5(4))["mesh"]
    p["functions"] = {"c2r_x": "1.0+x*x", "c2i_x": "0.5*y", "c2r_y": "2.0-z", "c2i_y": "0.1", "c2r_z": "1.5", "c2i_z": "x*y*z", "omega2r": "2.0+x", "omega2i": "0.3*y",
                      "source_r": "sin(2*pi*x)*z", "source_i": "cos(pi*y)", "robin_alpha_r": "1.0+z", "robin_alpha_i": "0.5", "source_r_side": "x+y", "source_i_side": "1.0-z"}
    p["dirichlet"] = {v: {s: "0.0" for s in ("left", "bottom", "back")} for v in ("ureal", "uimag")}
    p["neumann"] = {v: {s: "0.0" for s in ("right", "top")} for v in ("ureal", "uimag")}
    out.append((p["name"], p, {}))
    p = shrink(P.regression_helmholtz_2d(5)); p["name"] = "helmholtz 2D Q2"; p["order"] = 2
    out.append((p["name"], p, {}))
    # Burgers (the reference's own decks are 1-D, so these are oracle-pinned only): plain 2-D with a Neumann flux side,
    # 2-D with entropy viscosity + SUPG (transient: dt enters tau), plain 3-D, 2-D Q2
    bg = dict(name="burgers 2D neumann", mesh=shrink(P.regression_thermal_2d())["mesh"], physics="burgers", order=1,
              functions={"Burgers source": "sin(pi*x)*y", "diffusion": "0.05+0.1*x", "xvel": "1.0+0.5*y", "yvel": "0.5-x"},
              dirichlet={"u": {"left": "0.0", "bottom": "0.0"}}, neumann={"u": {"right": "0.0", "top": "0.0"}}, true={})
    out.append((bg["name"], bg, {}))
    p = copy.deepcopy(bg); p["name"] = "burgers 2D evisc+supg transient"; p["extra"] = {"transient": True}
    p["settings"] = {"use entropy viscosity": True, "use SUPG": True}
    p["functions"].update({"C1": "1.0+x", "C2": "2.0", "supg C": "0.5", "supg C1": "1.0", "supg C2": "2.0+y"})
    out.append((p["name"], p, {"transient": be}))
    p = copy.deepcopy(bg); p["name"] = "burgers 3D"; p["mesh"] = shrink(P.config_c5(4))["mesh"]
    p["functions"]["zvel"] = "0.25+x*y"; p["dirichlet"] = {"u": {"left": "0.0", "bottom": "0.0", "back": "0.0"}}; p["neumann"] = {}
    out.append((p["name"], p, {}))
    p = copy.deepcopy(bg); p["name"] = "burgers 2D Q2"; p["order"] = 2
    out.append((p["name"], p, {}))
    # lumped mass, fix-zero-rows
    p = shrink(P.regression_thermal_2d()); p["extra"] = {"lump mass": True, "fix zero rows": True}; p["name"] = "thermal 2D lump mass"
    out.append((p["name"], p, {}))
    return out


def distorted_nodes(nodes, mesh, amp=0.2, seed=4711):
    """Non-affine test mesh: every node moved by amp*h*(splitmix64 - 0.5) per coordinate; a node on a
    boundary plane moves only inside that plane, so the domain stays the brick and side sets stay planar but their faces
    become general quadrilaterals.  The trilinear cross terms of every element are non-zero afterwards."""
    dim = nodes.shape[1]
    n = [mesh["NX"], mesh["NY"], mesh.get("NZ", 1)]
    lo = [mesh.get("xmin", 0.0), mesh.get("ymin", 0.0), mesh.get("zmin", 0.0)]
    hi = [mesh.get("xmax", 1.0), mesh.get("ymax", 1.0), mesh.get("zmax", 1.0)]
    out = nodes.copy()
    idx = np.arange(len(nodes))
    for d in range(dim):
        h = (hi[d] - lo[d]) / n[d]
        r = problems.splitmix64(seed + d, idx) - 0.5
        inside = (np.abs(nodes[:, d] - lo[d]) > 1e-9 * h) & (np.abs(nodes[:, d] - hi[d]) > 1e-9 * h)
        out[inside, d] += amp * h * r[inside]
    return out


def distorted_cases():
    """every case of cases() on a non-affine mesh"""
    out = []
    for name, p, kw in cases():
        kw = dict(kw); kw["distort"] = 0.2
        out.append((name + " [distorted]", p, kw))
    return out


OPCODE_EXPRS = {
    "pow": "1.0+(x+0.5)^2.5", "div": "1.0/(1.0+x*y)", "lt": "1.0+0.5*(x<0.5)", "gt": "1.0+0.5*(y>0.4)", "lte": "1.0+0.5*(x<=0.5)",
    "gte": "1.0+0.5*(y>=0.4)", "max": "max(x,y)+0.5", "min": "min(x,0.25)+1.0", "exp": "exp(x-y)", "log": "log(2.0+x)",
    "tan": "1.5+tan(0.5*x)", "abs": "abs(x-0.5)+1.0", "sqrt": "sqrt(y+1.0)", "sqrt_neg": "1.0+sqrt(x-0.5)", "sinh": "1.0+sinh(x)",
    "cosh": "cosh(y)", "mean": "mean(1.0+x,2.0*y+1.0)", "time": "1.0+t*x"}
OPCODE_STATE_EXPRS = {
    "pow": "1.0+e^2", "div": "1.0/(1.0+e*e)", "lt": "1.0+(e<0.1)*e", "max": "1.0+max(e,0.0)", "min": "1.0+min(e,0.1)", "exp": "exp(0.3*e)",
    "log": "log(2.0+e*e)", "tan": "1.5+tan(0.2*e)", "abs": "1.0+abs(e)", "sqrt": "sqrt(1.0+e*e)", "sqrt_neg": "1.0+sqrt(e)",
    "sinh": "1.0+sinh(0.3*e)*sinh(0.3*e)", "cosh": "cosh(0.5*e)", "mean": "mean(1.0+e*e,1.0)", "grad": "1.0+grad(e)[x]*grad(e)[x]+grad(e)[z]*grad(e)[z]"}


def opcode_cases():
    """one case per coefficient-language operator: as a state-independent diffusion coefficient (specialised kernels: bytecode
    interpreter and NVRTC-printed) and as a state-dependent one (batched / generic kernels: PAD interpreter)"""
    P = problems
    out = []
    for op, ex in OPCODE_EXPRS.items():
        p = shrink(P.regression_thermal_3d()); p["name"] = "opcode " + op
        p["functions"] = dict(p["functions"]); p["functions"]["thermal diffusion"] = ex
        out.append((p["name"], p, {}))
    for op, ex in OPCODE_STATE_EXPRS.items():
        p = shrink(P.regression_thermal_3d()); p["name"] = "opcode(state) " + op
        p["functions"] = dict(p["functions"]); p["functions"]["thermal diffusion"] = ex
        out.append((p["name"], p, {}))
    # parameter leaves: `kap` is component 0 of a vector parameter, `kap(2)` component 2 (functionManager.cpp:270-355)
    p = shrink(P.regression_thermal_3d()); p["name"] = "opcode parameter"
    p["functions"] = dict(p["functions"]); p["functions"]["thermal diffusion"] = "kap+kap(2)*x*y+amp*z"
    p["parameters"] = {"kap": [1.25, -7.0, 0.5], "amp": 0.75}
    out.append((p["name"], p, {}))
    p = shrink(P.regression_thermal_3d()); p["name"] = "opcode(state) parameter"
    p["functions"] = dict(p["functions"]); p["functions"]["thermal diffusion"] = "kap+amp*e*e"
    p["parameters"] = {"kap": [1.25], "amp": 0.75}
    out.append((p["name"], p, {}))
    p = shrink(P.regression_thermal_2d()); p["name"] = "opcode 2D mix"
    p["functions"] = dict(p["functions"]); p["functions"]["thermal diffusion"] = "1.0+(x<0.5)*sqrt(y)+min(x,y)^2"
    out.append((p["name"], p, {}))
    return out


def config_parity_cases():
    """BASELINE.json configs C2..C5 at SURVEY.md 8d's parity sizes (8^3 and 24^3), entry by entry against the oracle"""
    P = problems
    out = []
    for mk, tag in ((P.config_c2, "C2"), (P.config_c3, "C3"), (P.config_c4, "C4"), (P.config_c5, "C5")):
        for n in (8, 24):
            out.append((f"{tag} {n}^3", mk(n), {"threads": 8}))
    out.append(("C1 40x40", P.config_c1(40), {}))
    return out


def oracle_problem(p):
    """The oracle's config text does not know engine-only keys."""
    q = copy.deepcopy(p)
    for k in ("true", "name", "time", "initial"):
        q.pop(k, None)
    return config_text(**q)


def rel_errors(a, b, rowptr=None):
    scale = np.abs(b).max()
    g = np.abs(a - b).max() / (scale if scale > 0 else 1.0)
    if rowptr is None:
        return g, g
    # per-row: ||d row||_inf / ||row||_inf
    d = np.abs(a - b)
    rows = np.repeat(np.arange(len(rowptr) - 1), np.diff(rowptr))
    dmax = np.zeros(len(rowptr) - 1); bmax = np.zeros(len(rowptr) - 1)
    np.maximum.at(dmax, rows, d); np.maximum.at(bmax, rows, np.abs(b))
    ok = bmax > 0
    return g, (dmax[ok] / bmax[ok]).max() if ok.any() else 0.0


def check_case(p, kw, lib, on_gpu, force_generic=False, jacobian=True, no_jit=False, no_tiles=False, no_ws=False, no_batched=False, jit_batched=False, tiled2d=False):
    from mrhyde_b200 import Engine
    o = Oracle(oracle_problem(p))
    e = Engine.from_problem(p, lib=lib, force_generic=force_generic, no_jit=no_jit, no_tiles=no_tiles, no_ws=no_ws, no_batched=no_batched, jit_batched=jit_batched, tiled2d=tiled2d)
    s = e.sizes()
    assert (s["nlocal"], s["nnz"], s["nelem"]) == (o.nlocal, o.nnz, o.nelem)
    lids, offs, fixed, xyz, var = e.dofmap()
    ol, oo, of, oxyz, ovar = o.dofmap()
    rp, ci = e.crs()
    orp, oci = o.crs()
    for a, b in ((lids, ol), (offs, oo), (fixed, of), (rp, orp), (ci, oci), (var, ovar), (xyz, oxyz)):
        assert np.array_equal(a, b), "indexing is not bit-exact"
    assert np.array_equal(e.bc_codes(), o.bc_codes())
    if kw.get("distort"):
        nd = distorted_nodes(o.mesh()[0], p["mesh"], kw["distort"])
        o.set_nodes(nd); e.set_nodes(nd)
        assert np.array_equal(e.mesh()[0], nd)
    u = problems.synthetic_state(xyz, var)
    ti = kw.get("transient")
    u_prev = u_stage = None
    if ti is not None:
        nsteps, nst = len(ti["bdf"]) - 1, len(ti["b"])
        u_prev = np.stack([problems.synthetic_state(xyz, var, seed=77 + k) * 0.9 for k in range(nsteps)], axis=1).copy()
        u_stage = np.stack([problems.synthetic_state(xyz, var, seed=99 + k) * 1.1 for k in range(nst)], axis=1).copy()
        o.set_time(ti["A"], ti["b"], ti["bdf"], ti["dt"], ti["time"], ti["stage"])
    ores, ojac = o.assemble(u, jacobian=jacobian, u_prev=u_prev, u_stage=u_stage, nthreads=kw.get("threads", 1))
    if on_gpu:
        import torch
        dev = torch.device("cuda:0")
        tu = torch.from_numpy(u).to(dev)
        tp = torch.from_numpy(u_prev).to(dev) if u_prev is not None else None
        ts = torch.from_numpy(u_stage).to(dev) if u_stage is not None else None
        tres = torch.zeros(o.nlocal, dtype=torch.float64, device=dev)
        tjac = torch.zeros(o.nnz, dtype=torch.float64, device=dev)
        if jacobian:
            e.assemble_jacres(tu, tres, tjac, u_prev=tp, u_stage=ts, time_int=ti)
        else:
            e.assemble_res(tu, tres, u_prev=tp, u_stage=ts, time_int=ti)
        torch.cuda.synchronize()
        res, jac = tres.cpu().numpy(), tjac.cpu().numpy()
    else:
        res, jac = np.zeros(o.nlocal), np.zeros(o.nnz)
        if jacobian:
            e.assemble_jacres(u, res, jac, u_prev=u_prev, u_stage=u_stage, time_int=ti)
        else:
            e.assemble_res(u, res, u_prev=u_prev, u_stage=u_stage, time_int=ti)
    er, _ = rel_errors(res, ores)
    assert er <= TOL, f"residual rel err {er:.3e}"
    out = {"res": er, "launches": e.launch_count(), "path": e.kernel_path()}
    if jacobian:
        eg, erow = rel_errors(jac, ojac, rp)
        assert eg <= TOL and erow <= ROW_TOL, f"jacobian rel err {eg:.3e} (per-row {erow:.3e})"
        out.update(jac=eg, jac_row=erow)
    e.close()
    return out


def check_foreign_numbering(p, lib, on_gpu, seed=5, jacobian=True, distort=0.0, **flags):
    """The drop-in path proper: a caller-supplied mesh / numbering / graph (mrh_set_mesh, mrh_set_dofmap, mrh_set_crs,
    mrh_add_side) that is NOT the brick builder's - element order and local DOF ids randomly permuted - against the oracle's
    result mapped through the permutations.  Returns the engine's report (kernel path, number of slot patterns)."""
    import scipy.sparse as sp
    from mrhyde_b200 import Engine
    o = Oracle(oracle_problem(p))
    rng = np.random.default_rng(seed)
    nodes, e2n = o.mesh()
    if distort:
        nodes = distorted_nodes(nodes, p["mesh"], distort)
        o.set_nodes(nodes)
    lids, offs, _, xyz, var = o.dofmap()
    rp, ci = o.crs()
    pe = rng.permutation(o.nelem)                    # new element k is old element pe[k]
    inv_pe = np.empty_like(pe); inv_pe[pe] = np.arange(o.nelem)
    pl = rng.permutation(o.nlocal).astype(np.int32)  # old LID i becomes pl[i]
    pn = rng.permutation(o.nnode).astype(np.int32)   # old node a becomes pn[a]
    new_nodes = np.empty_like(nodes); new_nodes[pn] = nodes
    new_e2n = pn[e2n[pe]]
    new_lids = pl[lids[pe]]
    # CRS graph of the permuted numbering; `src` maps every new slot to the old slot holding the same (row, col) pair
    rows_old = np.repeat(np.arange(o.nlocal), np.diff(rp))
    A = sp.coo_matrix((np.arange(1, o.nnz + 1, dtype=np.float64), (pl[rows_old], pl[ci])), shape=(o.nlocal, o.nlocal)).tocsr()
    A.sort_indices()
    assert A.nnz == o.nnz
    src = A.data.astype(np.int64) - 1
    extra = p.get("extra", {})
    e = Engine(p["physics"], p["mesh"]["dim"], p.get("order", 1), lib=lib, quadrature=extra.get("quadrature", -1),
               side_quadrature=extra.get("side quadrature", -1), use_strong_dbc=extra.get("use strong DBCs", True),
               fix_zero_rows=extra.get("fix zero rows", False), lump_mass=extra.get("lump mass", False), settings=p.get("settings", {}), **flags)
    e.set_mesh(new_nodes, new_e2n)
    e.set_dofmap(new_lids, offs, o.nlocal)
    e.set_crs(A.indptr.astype(np.int64), A.indices.astype(np.int32))
    for name, ls, el in o.sides():
        e.add_side(name, ls, np.sort(inv_pe[el]).astype(np.int32))
    for name, expr in p.get("functions", {}).items():
        e.add_function(name, expr, "ip"); e.add_function(name, expr, "side ip")
    weak = bool(extra.get("use weak Dirichlet", False))
    for var_, sides in p.get("dirichlet", {}).items():
        for side, expr in sides.items():
            e.set_bc(var_, side, "weak_dirichlet" if weak else "dirichlet", expr)
    for kind in ("neumann", "flux"):
        for var_, sides in p.get(kind, {}).items():
            for side, expr in sides.items():
                e.set_bc(var_, side, kind, expr)
    e.finalize()
    fixed_new = e.dofmap(coords=False)[2]
    of = o.dofmap()[2]
    assert np.array_equal(fixed_new[pl], of), "fixed-DOF flags differ under the permutation"
    u_old = problems.synthetic_state(xyz, var)
    u_new = np.empty_like(u_old); u_new[pl] = u_old
    ores, ojac = o.assemble(u_old, jacobian=jacobian)
    want_res = np.empty_like(ores); want_res[pl] = ores
    if on_gpu:
        import torch
        dev = torch.device("cuda:0")
        tu = torch.from_numpy(u_new).to(dev)
        tres = torch.zeros(o.nlocal, dtype=torch.float64, device=dev); tjac = torch.zeros(o.nnz, dtype=torch.float64, device=dev)
        if jacobian:
            e.assemble_jacres(tu, tres, tjac)
        else:
            e.assemble_res(tu, tres)
        torch.cuda.synchronize()
        res, jac = tres.cpu().numpy(), tjac.cpu().numpy()
    else:
        res, jac = np.zeros(o.nlocal), np.zeros(o.nnz)
        if jacobian:
            e.assemble_jacres(u_new, res, jac)
        else:
            e.assemble_res(u_new, res)
    er, _ = rel_errors(res, want_res)
    assert er <= TOL, f"residual rel err {er:.3e}"
    out = {"res": er, "path": e.kernel_path(), "npatterns": e.sizes()["npatterns"], "launches": e.launch_count()}
    if jacobian:
        eg, erow = rel_errors(jac, ojac[src], A.indptr)
        assert eg <= TOL and erow <= ROW_TOL, f"jacobian rel err {eg:.3e} (per-row {erow:.3e})"
        out.update(jac=eg, jac_row=erow)
    e.close()
    return out
