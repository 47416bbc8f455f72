"""Problem definitions shared by the oracle tests, the GPU parity tests, smoke() and bench.py.

Each problem is a plain dict understood both by oracle.pyoracle.config_text() and by
mrhyde_b200.Engine.from_problem().  The regression_* entries restate the reference's regression inputs
(/root/reference/regression/<path>/input.yaml); the config_* entries are BASELINE.json's five configs
 at a caller-chosen mesh size.
"""
import numpy as np

PI2 = "2*pi"


def _mesh(dim, n, hi=(1.0, 1.0, 1.0)):
    m = dict(dim=dim, NX=n[0], NY=n[1], xmin=0.0, xmax=hi[0], ymin=0.0, ymax=hi[1])
    if dim == 3:
        m.update(NZ=n[2], zmin=0.0, zmax=hi[2])
    return m


def regression_thermal_2d(n=40):
    return dict(name="thermal/2D_verification", mesh=_mesh(2, (n, n)), physics="thermal", order=1,
                functions={"thermal source": "8*(pi*pi)*sin(2*pi*x)*sin(2*pi*y)"},
                dirichlet={"e": {"all boundaries": "0.0"}},
                true={"e": "sin(2*pi*x)*sin(2*pi*y)"})


def regression_thermal_2d_mixed_bcs(n=40):
    """regression/thermal/2D_mixed_bcs: Dirichlet left / right, Neumann top / bottom (pins the side integration)"""
    return dict(name="thermal/2D_mixed_bcs", mesh=_mesh(2, (n, n)), physics="thermal", order=1,
                functions={"thermal source": "8*pi*pi*sin(2*pi*x)*sin(2*pi*y)"},      # (the input's unused "thermal Neumann source" is left out)
                dirichlet={"e": {"left": "0.0", "right": "0.0"}},
                neumann={"e": {"top": "2*pi*sin(2*pi*x)*cos(2*pi*y)", "bottom": "-2*pi*sin(2*pi*x)*cos(2*pi*y)"}},
                true={"e": "sin(2*pi*x)*sin(2*pi*y)"})


def regression_helmholtz_2d(n=100):
    """regression/helmholtz/manufactured_solution: complex-valued Helmholtz operator as two HGRAD variables, Dirichlet on three sides,
    Robin condition (sides where ureal is "Neumann") on the right"""
    f = {"source_r_side": "2.0*pi*cos(2*pi*x)*sin(2*pi*y)", "source_i_side": "2.0*pi*cos(2*pi*x)*sin(2*pi*y)", "omega": "1.0",
         "scoeff": "8*pi*pi*(x*x-2*x-1)-1.0", "scoeffi": "8*pi*pi*(x*x+2*x-1)-1.0", "srcoeff": "2.0-2*x", "sicoeff": "-2.0-2*x",
         "source_r": "scoeff*sin(2*pi*x)*sin(2*pi*y) + srcoeff*2*pi*cos(2*pi*x)*sin(2*pi*y)",
         "source_i": "scoeffi*sin(2*pi*x)*sin(2*pi*y) + sicoeff*2*pi*cos(2*pi*x)*sin(2*pi*y)",
         "c2r_x": "x*x-1.0", "c2i_x": "2.0*x", "c2r_y": "x*x-1.0", "c2i_y": "2.0*x", "omega2r": "1.0", "omega2i": "0.0"}
    d = {v: {"left": "0.0", "top": "0.0", "bottom": "0.0"} for v in ("ureal", "uimag")}
    return dict(name="helmholtz/manufactured_solution", mesh=_mesh(2, (n, n)), physics="helmholtz", order=1, functions=f, dirichlet=d,
                neumann={"ureal": {"right": "0.0"}, "uimag": {"right": "0.0"}},
                true={"ureal": "sin(2*pi*x)*sin(2*pi*y)", "uimag": "sin(2*pi*x)*sin(2*pi*y)"})


def regression_thermal_2d_nonzero_dbc(n=40):
    """regression/thermal/2D_verification_nonzeroDBC: Dirichlet data sin(x) sin(y), L2-projected onto the boundary DOFs"""
    return dict(name="thermal/2D_verification_nonzeroDBC", mesh=_mesh(2, (n, n)), physics="thermal", order=1,
                functions={"thermal source": "2.0*sin(x)*sin(y)"}, dirichlet={"e": {"all boundaries": "sin(x)*sin(y)"}},
                true={"e": "sin(x)*sin(y)"})


def regression_thermal_2d_transient(n=40):
    """regression/thermal/2D_verification_transient: backward Euler, 20 steps to t = 1, zero initial and boundary data"""
    return dict(name="thermal/2D_verification_transient", mesh=_mesh(2, (n, n)), physics="thermal", order=1,
                functions={"thermal source": "(8*(pi*pi)*sin(2*pi*t)+2*pi*cos(2*pi*t))*sin(2*pi*x)*sin(2*pi*y)"},
                dirichlet={"e": {"all boundaries": "0.0"}}, extra={"transient": True},
                true={"e": "sin(2*pi*t)*sin(2*pi*x)*sin(2*pi*y)"}, time=dict(final=1.0, steps=20, tableau="BWE", bdf=1))


def regression_cdr_2d_transient(n=40):
    """regression/cdr/2D_transient: a Gaussian bubble advected with velocity (10, 0) and the nonlinear reaction 0.5 c^2, backward
    Euler, 10 steps to t = 0.1; the initial state is the L2 projection of exp(bubble)"""
    return dict(name="cdr/2D_transient", mesh=_mesh(2, (n, n)), physics="cdr", order=1,
                functions={"source": "0.0", "xvel": "10.0", "yvel": "0.0", "reaction": "0.5*c*c", "SUPG tau": "0.0",
                           "bubble": "-10.0*(x-0.2)*(x-0.2) - 10.0*(y-0.5)*(y-0.5)"},
                dirichlet={"c": {"all boundaries": "0.0"}}, extra={"transient": True, "workset size": 100},
                initial={"c": "exp(bubble)"}, true={"c": "0.0"}, time=dict(final=0.1, steps=10, tableau="BWE", bdf=1))


def regression_thermal_3d(n=10):
    return dict(name="thermal/3D_verification", mesh=_mesh(3, (n, n, n)), physics="thermal", order=1,
                functions={"thermal source": "12*(pi*pi)*sin(2*pi*x)*sin(2*pi*y)*sin(2*pi*z)"},
                dirichlet={"e": {"all boundaries": "0.0"}},
                true={"e": "sin(2*pi*x)*sin(2*pi*y)*sin(2*pi*z)"})


def regression_cdr_2d(n=40):
    return dict(name="cdr/2D_manufactured", mesh=_mesh(2, (n, n)), physics="cdr", order=1,
                functions={"source": "(8*(pi*pi)+0.5*sin(2*pi*x)*sin(2*pi*y))*sin(2*pi*x)*sin(2*pi*y) + "
                                     "2.0*2*pi*cos(2*pi*x)*sin(2*pi*y) + 1.0*2*pi*sin(2*pi*x)*cos(2*pi*y)",
                           "xvel": "2.0", "yvel": "1.0", "reaction": "0.5*c*c", "SUPG tau": "0.0"},
                dirichlet={"c": {"all boundaries": "0.0"}}, extra={"workset size": 40},
                true={"c": "sin(2*pi*x)*sin(2*pi*y)"})


def _le_functions(dim):
    f = {"lambda": "1.0", "mu": "1.0", "A": "1.0", "B": "2.0"}
    if dim == 2:
        f.update({
            "dxxx": "(A*pi)*(A*pi)*sin(A*pi*x)*sin(A*pi*y)", "dxxy": "-1.0*(A*pi)*(A*pi)*cos(A*pi*x)*cos(A*pi*y)",
            "dxyy": "(A*pi)*(A*pi)*sin(A*pi*x)*sin(A*pi*y)", "dyxx": "(B*pi)*(B*pi)*sin(B*pi*x)*sin(B*pi*y)",
            "dyxy": "-1.0*(B*pi)*(B*pi)*cos(B*pi*x)*cos(B*pi*y)", "dyyy": "(B*pi)*(B*pi)*sin(B*pi*x)*sin(B*pi*y)",
            "source dx": "(lambda+2.0*mu)*dxxx + mu*(dxyy+dyxy) + lambda*dyxy",
            "source dy": "(lambda+2.0*mu)*dyyy + mu*(dyxx+dxxy) + lambda*dxxy"})
        return f
    f["C"] = "3.0"
    for v, L in (("dx", "A"), ("dy", "B"), ("dz", "C")):
        s = f"({L}*pi)*({L}*pi)*"
        sss = f"sin({L}*pi*x)*sin({L}*pi*y)*sin({L}*pi*z)"
        f[v + "xx"] = s + sss
        f[v + "yy"] = s + sss
        f[v + "zz"] = s + sss
        f[v + "xy"] = "-1.0*" + s + f"cos({L}*pi*x)*cos({L}*pi*y)*sin({L}*pi*z)"
        f[v + "xz"] = "-1.0*" + s + f"cos({L}*pi*x)*sin({L}*pi*y)*cos({L}*pi*z)"
        f[v + "yz"] = "-1.0*" + s + f"sin({L}*pi*x)*cos({L}*pi*y)*cos({L}*pi*z)"
    f["source dx"] = "(lambda+2.0*mu)*dxxx + mu*(dxyy+dyxy+dxzz+dzxz) + lambda*(dyxy+dzxz)"
    f["source dy"] = "(lambda+2.0*mu)*dyyy + mu*(dyxx+dxxy+dyzz+dzyz) + lambda*(dxxy+dzyz)"
    f["source dz"] = "(lambda+2.0*mu)*dzzz + mu*(dzxx+dxxz+dzyy+dyyz) + lambda*(dxxz+dyyz)"
    return f


def regression_le_2d(n=40):
    return dict(name="le/2D_manufactured", mesh=_mesh(2, (n, n)), physics="linearelasticity", order=1,
                functions=_le_functions(2),
                dirichlet={"dx": {"all boundaries": "0.0"}, "dy": {"all boundaries": "0.0"}},
                true={"dx": "sin(A*pi*x)*sin(A*pi*y)", "dy": "sin(B*pi*x)*sin(B*pi*y)"})


def regression_le_3d(n=10):
    return dict(name="le/3D_manufactured", mesh=_mesh(3, (n, n, n)), physics="linearelasticity", order=1,
                functions=_le_functions(3),
                dirichlet={v: {"all boundaries": "0.0"} for v in ("dx", "dy", "dz")},
                true={"dx": "sin(A*pi*x)*sin(A*pi*y)*sin(A*pi*z)", "dy": "sin(B*pi*x)*sin(B*pi*y)*sin(B*pi*z)",
                      "dz": "sin(C*pi*x)*sin(C*pi*y)*sin(C*pi*z)"})


def regression_ns_channel(nx=50, ny=10):
    return dict(name="navierstokes/channel", mesh=_mesh(2, (nx, ny), hi=(5.0, 1.0, 1.0)), physics="navierstokes",
                order=1, functions={"source ux": "1.0"}, settings={"usePSPG": True},
                dirichlet={"ux": {"bottom": "0.0", "top": "0.0"}, "uy": {"bottom": "0.0", "top": "0.0"}},
                true={"ux": "0.5*y*(1.0-y)", "uy": "0.0", "pr": "0.0"})


def regression_stokes_2d(n=4):
    """regression/stokes/2D_verification_pspg (the next HGRAD module)"""
    return dict(name="stokes/2D_verification_pspg", mesh=_mesh(2, (n, n)), physics="stokes", order=1,
                functions={"source ux": "1.0"}, settings={"usePSPG": True}, extra={"workset size": 16},
                dirichlet={"ux": {"bottom": "0.0", "top": "0.0"}, "uy": {"bottom": "0.0", "top": "0.0"}},
                true={"ux": "0.5*y*(1.0-y)", "uy": "0.0", "pr": "0.0"})


# ---- BASELINE.json configs ------------------------------------------------------------
def config_c1(n=40):
    p = regression_thermal_2d(n)
    p["name"] = "C1 thermal 2D Q1"
    return p


def config_c2(n=8):
    p = regression_thermal_3d(n)
    p["name"] = "C2 thermal 3D Q1"
    return p


def config_c3(n=4):
    src = ("(12*(pi*pi)+0.5*sin(2*pi*x)*sin(2*pi*y)*sin(2*pi*z))*sin(2*pi*x)*sin(2*pi*y)*sin(2*pi*z) + "
           "2.0*2*pi*cos(2*pi*x)*sin(2*pi*y)*sin(2*pi*z) + 1.0*2*pi*sin(2*pi*x)*cos(2*pi*y)*sin(2*pi*z) + "
           "0.5*2*pi*sin(2*pi*x)*sin(2*pi*y)*cos(2*pi*z)")
    return dict(name="C3 cdr 3D Q2", mesh=_mesh(3, (n, n, n)), physics="cdr", order=2,
                functions={"source": src, "xvel": "2.0", "yvel": "1.0", "zvel": "0.5", "reaction": "0.5*c*c",
                           "diffusion": "1.0"},
                dirichlet={"c": {"all boundaries": "0.0"}},
                true={"c": "sin(2*pi*x)*sin(2*pi*y)*sin(2*pi*z)"})


def config_c4(n=8):
    p = regression_le_3d(n)
    p["name"] = "C4 linear elasticity 3D Q1"
    return p


def config_c5(n=8):
    d = {v: {s: "0.0" for s in ("bottom", "top", "front", "back")} for v in ("ux", "uy", "uz")}
    return dict(name="C5 navier-stokes 3D Q1 SUPG/PSPG", mesh=_mesh(3, (n, n, n)), physics="navierstokes", order=1,
                functions={"source ux": "1.0", "viscosity": "1.0", "density": "1.0"},
                settings={"useSUPG": True, "usePSPG": True}, dirichlet=d, true={})


def splitmix64(seed, idx):
    """Deterministic perturbation generator of SURVEY.md 8d (vectorised)."""
    z = (np.uint64(seed) + (idx.astype(np.uint64) + np.uint64(1)) * np.uint64(0x9E3779B97F4A7C15))
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    z = z ^ (z >> np.uint64(31))
    return z.astype(np.float64) / 2.0 ** 64


def synthetic_state(dofxyz, dofvar, seed=20261104):
    """u_i = prod_d sin(2 pi x_d) + 1e-3 (splitmix64(seed+var, i) - 0.5)   (SURVEY.md 8d)."""
    with np.errstate(over="ignore"):
        u = np.prod(np.sin(2.0 * np.pi * dofxyz), axis=1)
        idx = np.arange(len(u))
        pert = np.zeros_like(u)
        for v in np.unique(dofvar):
            m = dofvar == v
            pert[m] = splitmix64(seed + int(v), idx[m])
    return u + 1e-3 * (pert - 0.5)
