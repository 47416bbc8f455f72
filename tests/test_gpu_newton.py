"""The caller of the hot path on the device: mrh_nonlinear_solve / mrh_linear_solve / mrh_transient_solve.
The reference's own 6-digit regression golds (tests/golden/reference_golds.json) are reproduced FROM THE GPU SOLUTION: CUDA
assembly + CUDA Krylov solve + Newton update; the oracle only integrates the L2 error of the vector it is handed."""
import json
import os

import numpy as np
import pytest

import parity_util as U
import problems

pytestmark = pytest.mark.gpu
GOLD = json.load(open(os.path.join(U.HERE, "golden", "reference_golds.json")))

CASES = [
    ("thermal/2D_verification", problems.regression_thermal_2d, 1),
    ("thermal/3D_verification", problems.regression_thermal_3d, 1),
    ("cdr/2D_manufactured", problems.regression_cdr_2d, 0),
    ("le/2D_manufactured", problems.regression_le_2d, 1),
    ("le/3D_manufactured", problems.regression_le_3d, 0),
    ("navierstokes/channel", problems.regression_ns_channel, 0),
    ("thermal/2D_mixed_bcs", problems.regression_thermal_2d_mixed_bcs, 1),
    ("stokes/2D_verification_pspg", problems.regression_stokes_2d, 0),
]
VARS = {"thermal": ["e"], "cdr": ["c"], "linearelasticity": ["dx", "dy", "dz"], "navierstokes": ["ux", "pr", "uy", "uz"], "stokes": ["ux", "pr", "uy", "uz"], "helmholtz": ["ureal", "uimag"]}


@pytest.fixture(scope="module")
def lib():
    import torch
    from mrhyde_b200 import lib_path
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return lib_path()


def _buffers(e):
    import torch
    s = e.sizes()
    dev = torch.device("cuda:0")
    return (torch.zeros(s["nlocal"], dtype=torch.float64, device=dev), torch.zeros(s["nlocal"], dtype=torch.float64, device=dev),
            torch.zeros(s["nnz"], dtype=torch.float64, device=dev))


@pytest.mark.parametrize("path,make,method", CASES, ids=[c[0] for c in CASES])
def test_reference_golds_from_the_gpu_solution(lib, path, make, method):
    import torch
    from mrhyde_b200 import Engine
    from oracle.pyoracle import Oracle
    p = make()
    e = Engine.from_problem(p, lib=lib)
    u, res, jac = _buffers(e)
    info = e.nonlinear_solve(u, res, jac, linear_method=method, linear_tol=1e-12, max_linear_iters=20000, krylov_vectors=300,
                             nonlinear_tol=1e-9, max_nonlinear_iters=12)
    torch.cuda.synchronize()
    assert info["status"] == 0 and info["converged"] == 1, info
    assert info["last_linear_relres"] <= 1e-10, info
    o = Oracle(U.oracle_problem(p))
    uh = u.cpu().numpy()
    names = VARS[p["physics"]][: (o.dim if p["physics"] == "linearelasticity" else o.dim + 1 if p["physics"] in ("navierstokes", "stokes") else
                                  2 if p["physics"] == "helmholtz" else 1)]
    for var, gold in GOLD["l2"][path].items():
        if var not in p["true"]:
            continue
        err = o.l2_error(uh, names.index(var), p["true"][var])
        assert err == pytest.approx(gold, rel=2e-5, abs=1e-9), (path, var, err, gold, info)
    e.close()


def test_helmholtz_gold_from_the_gpu_assembly(lib):
    """regression/helmholtz/manufactured_solution: the operator is indefinite and non-symmetric (the reference preconditions it with
    ILUT), which the device's Jacobi-preconditioned GMRES does not solve - so the Newton loop runs here with the ENGINE's residual and
    Jacobian (mrh_assemble_jacres on the device: volume kernel + the Robin side launches) and a sparse direct solve on the host; the
    oracle only integrates the L2 error of the result"""
    import torch
    import scipy.sparse as sp
    import scipy.sparse.linalg as spl
    from mrhyde_b200 import Engine
    from oracle.pyoracle import Oracle
    p = problems.regression_helmholtz_2d()
    e = Engine.from_problem(p, lib=lib)
    u, res, jac = _buffers(e)
    rp, ci = e.crs()
    n = len(rp) - 1
    for it in range(6):
        res.zero_(); jac.zero_()
        e.assemble_jacres(u, res, jac)
        torch.cuda.synchronize()
        r = res.cpu().numpy()
        if np.abs(r).max() < 1e-10:
            break
        J = sp.csr_matrix((jac.cpu().numpy(), ci, rp), shape=(n, n)).tocsc()
        u += torch.from_numpy(spl.spsolve(J, r)).to(u.device)          # (the sign convention of tests/test_oracle_golden.py::newton)
    assert it <= 2, it                                                  # the problem is linear
    o = Oracle(U.oracle_problem(p))
    uh = u.cpu().numpy()
    for k, var in enumerate(("ureal", "uimag")):
        err = o.l2_error(uh, k, p["true"][var])
        assert err == pytest.approx(GOLD["l2"]["helmholtz/manufactured_solution"][var], rel=2e-5), (var, err)
    e.close()


def test_transient_regression_series_from_the_gpu(lib):
    """regression/thermal/2D_verification_transient: mrh_transient_solve one backward-Euler step at a time; the L2 error after
    every step equals the reference's printed series (6 digits)"""
    import torch
    from mrhyde_b200 import Engine
    from oracle.pyoracle import Oracle
    p = problems.regression_thermal_2d_transient()
    e = Engine.from_problem(p, lib=lib)
    o = Oracle(U.oracle_problem(p))
    u, res, jac = _buffers(e)
    nsteps, T = p["time"]["steps"], p["time"]["final"]
    dt = T / nsteps
    gold = GOLD["l2_series"]["thermal/2D_verification_transient"]["e"]
    for k in range(nsteps):
        info = e.transient_solve(u, res, jac, k * dt, (k + 1) * dt, dt, tableau=p["time"]["tableau"], bdf_order=p["time"]["bdf"],
                                 linear_method=1, linear_tol=1e-12, max_linear_iters=5000, nonlinear_tol=1e-9)
        assert info["status"] == 0 and info["time_steps"] == 1, info
        err = o.l2_error(u.cpu().numpy(), 0, p["true"]["e"], time=(k + 1) * dt)
        assert err == pytest.approx(gold[k + 1][1], rel=3e-5), (k, err, gold[k + 1], info)
    e.close()


def test_cdr_transient_series_from_the_gpu(lib):
    """regression/cdr/2D_transient on the device: mrh_set_initial (L2 projection of the Gaussian), Dirichlet DOFs to their data,
    then mrh_transient_solve step by step (state-dependent reaction: batched kernel with the PAD interpreter, Newton + GMRES).
    4-digit agreement with the reference's series (see tests/test_oracle_golden.py::test_cdr_transient_regression_series)."""
    import torch
    from mrhyde_b200 import Engine
    from oracle.pyoracle import Oracle
    p = problems.regression_cdr_2d_transient()
    e = Engine.from_problem(p, lib=lib)
    o = Oracle(U.oracle_problem(p))
    u, res, jac = _buffers(e)
    info = e.set_initial(u, res, jac)
    assert info["status"] == 0, info
    fixed = torch.from_numpy(e.dofmap(coords=False)[2].astype(bool)).to(u.device)
    u[fixed] = 0.0
    gold = GOLD["l2_series"]["cdr/2D_transient"]["c"]
    assert o.l2_error(u.cpu().numpy(), 0, "0.0") == pytest.approx(gold[0][1], rel=5e-4)
    nsteps, dt = p["time"]["steps"], p["time"]["final"] / p["time"]["steps"]
    for k in range(nsteps):
        info = e.transient_solve(u, res, jac, k * dt, (k + 1) * dt, dt, tableau="BWE", bdf_order=1,
                                 linear_method=0, linear_tol=1e-12, max_linear_iters=5000, krylov_vectors=200, nonlinear_tol=1e-10, max_nonlinear_iters=10)
        assert info["status"] == 0 and info["time_steps"] == 1, info
        err = o.l2_error(u.cpu().numpy(), 0, "0.0", time=(k + 1) * dt)
        assert err == pytest.approx(gold[k + 1][1], rel=5e-4), (k, err, gold[k + 1], info)
    e.close()


@pytest.mark.parametrize("method", [0, 1], ids=["gmres", "cg"])
def test_linear_solve_against_a_direct_solver(lib, method):
    import scipy.sparse as sp
    import scipy.sparse.linalg as spl
    import torch
    from mrhyde_b200 import Engine
    p = problems.config_c2(9)
    e = Engine.from_problem(p, lib=lib)
    xyz, var = e.dof_coords()
    u, res, jac = _buffers(e)
    u.copy_(torch.from_numpy(problems.synthetic_state(xyz, var)))
    e.assemble_jacres(u, res, jac)
    x = torch.zeros_like(res)
    info = e.linear_solve(jac, res, x, linear_method=method, linear_tol=1e-13, max_linear_iters=5000, krylov_vectors=60)
    rp, ci = e.crs()
    J = sp.csr_matrix((jac.cpu().numpy(), ci, rp), shape=(len(rp) - 1,) * 2).tocsc()
    want = spl.spsolve(J, res.cpu().numpy())
    assert info["converged"] == 1 and info["linear_iters"] > 5, info
    assert np.abs(x.cpu().numpy() - want).max() <= 1e-10 * np.abs(want).max(), info
    e.close()


def test_newton_converges_quadratically_on_the_nonlinear_configs(lib):
    """C3 (reaction 0.5 c^2) and C5 (convection + SUPG/PSPG) at small sizes: residual drops below 1e-10 of its start"""
    import torch
    from mrhyde_b200 import Engine
    for mk, n in ((problems.config_c3, 4), (problems.config_c5, 5)):
        e = Engine.from_problem(mk(n), lib=lib)
        u, res, jac = _buffers(e)
        info = e.nonlinear_solve(u, res, jac, linear_tol=1e-12, max_linear_iters=20000, krylov_vectors=400, nonlinear_tol=1e-10,
                                 max_nonlinear_iters=15)
        torch.cuda.synchronize()
        assert info["status"] == 0 and 2 <= info["nonlinear_iters"] <= 10, info
        assert info["resnorm_last"] <= 1e-10 * info["resnorm_first"], info
        e.close()


@pytest.mark.parametrize("tableau,order,rate", [("BWE", 1, 1.0), ("DIRK-2,2", 1, 2.0), ("BWE", 2, 2.0)])
def test_transient_solver_order_of_accuracy(lib, tableau, order, rate):
    """thermal 2-D with source = 0, u(x, y, 0) = sin(pi x) sin(pi y), exact nodal decay known for the SEMI-discrete system
    only through a fine-step run: halving dt must reduce the error against a 16x finer run by 2^rate."""
    import torch
    from mrhyde_b200 import Engine
    p = problems.regression_thermal_2d(8)
    p["functions"] = {"thermal source": "0.0", "density": "1.0", "specific heat": "1.0"}
    p["extra"] = {"transient": True}
    e = Engine.from_problem(p, lib=lib)
    xyz, _ = e.dof_coords()
    u0 = np.sin(np.pi * xyz[:, 0]) * np.sin(np.pi * xyz[:, 1])
    T = 0.04

    def run(nsteps):
        u, res, jac = _buffers(e)
        u.copy_(torch.from_numpy(u0))
        info = e.transient_solve(u, res, jac, 0.0, T, T / nsteps, tableau=tableau, bdf_order=order, linear_method=1, linear_tol=1e-13,
                                 max_linear_iters=2000, nonlinear_tol=1e-11)
        assert info["status"] == 0 and info["time_steps"] == nsteps and abs(info["final_time"] - T) < 1e-12, info
        return u.cpu().numpy()
    fine = run(256)
    e4, e8 = np.abs(run(4) - fine).max(), np.abs(run(8) - fine).max()
    assert 0.0 < e8 < e4 < 0.1
    assert np.log2(e4 / e8) == pytest.approx(rate, abs=0.35), (e4, e8)
    assert 0.3 < fine.max() < 0.6                                    # exp(-2 pi^2 0.04) = 0.454 for the continuous problem
    e.close()


# ---- L2-projected Dirichlet / initial data on the device -------------------------------------------------------
def test_nonzero_dirichlet_gold_from_the_gpu(lib):
    """regression/thermal/2D_verification_nonzeroDBC (1.42007e-05): mrh_set_dirichlet projects sin(x) sin(y) onto the boundary DOFs,
    mrh_nonlinear_solve keeps them (fixed rows) - all on the device"""
    import torch
    from mrhyde_b200 import Engine
    from oracle.pyoracle import Oracle
    p = problems.regression_thermal_2d_nonzero_dbc()
    e = Engine.from_problem(p, lib=lib)
    u, res, jac = _buffers(e)
    info = e.set_dirichlet(u, res, jac)
    assert info["status"] == 0 and info["last_linear_relres"] <= 1e-12, info
    fixed = e.dofmap(coords=False)[2].astype(bool)
    u0 = u.cpu().numpy().copy()
    assert np.all(u0[~fixed] == 0.0) and np.abs(u0[fixed]).max() > 0.5
    info = e.nonlinear_solve(u, res, jac, linear_method=1, linear_tol=1e-12, max_linear_iters=5000, nonlinear_tol=1e-9)
    torch.cuda.synchronize()
    assert info["status"] == 0, info
    uh = u.cpu().numpy()
    assert np.array_equal(uh[fixed], u0[fixed])
    o = Oracle(U.oracle_problem(p))
    err = o.l2_error(uh, 0, p["true"]["e"])
    assert err == pytest.approx(GOLD["l2"]["thermal/2D_verification_nonzeroDBC"]["e"], rel=2e-5), err
    e.close()


@pytest.mark.parametrize("mk,n,distort", [("regression_thermal_2d_nonzero_dbc", 12, 0.2), ("config_c2", 6, 0.2), ("config_c4", 4, 0.0), ("config_c3", 3, 0.2)],
                         ids=["thermal-2D", "thermal-3D", "LE-3D", "cdr-Q2"])
def test_dirichlet_projection_system_on_the_device(lib, mk, n, distort):
    import torch
    from mrhyde_b200 import Engine
    from oracle.pyoracle import Oracle
    p = getattr(problems, mk)(n)
    if mk != "regression_thermal_2d_nonzero_dbc":
        for v in p["dirichlet"]:
            p["dirichlet"][v] = {"left": "0.3*y+z", "top": "sin(x)+t", "front": "1.0"}
    o = Oracle(U.oracle_problem(p))
    e = Engine.from_problem(p, lib=lib)
    if distort:
        nd = U.distorted_nodes(o.mesh()[0], p["mesh"], distort)
        o.set_nodes(nd); e.set_nodes(nd)
    orhs, omass = o.assemble_dirichlet(time=0.4)
    u, rhs, mass = _buffers(e)
    e.assemble_projection(0, rhs, mass, time=0.4)
    torch.cuda.synchronize()
    assert U.rel_errors(rhs.cpu().numpy(), orhs)[0] <= U.TOL
    assert U.rel_errors(mass.cpu().numpy(), omass, o.crs()[0]) <= (U.TOL, U.TOL)
    e.close()


def test_initial_projection_on_the_device(lib):
    import torch
    from mrhyde_b200 import Engine
    p = problems.config_c4(5)
    p["initial"] = {"dx": "1.0+2.0*x+y*z", "dz": "x*y*z"}
    e = Engine.from_problem(p, lib=lib)
    u, rhs, mass = _buffers(e)
    info = e.set_initial(u, rhs, mass)
    assert info["status"] == 0, info
    xyz, var = e.dof_coords()
    want = np.where(var == 0, 1.0 + 2.0 * xyz[:, 0] + xyz[:, 1] * xyz[:, 2], np.where(var == 2, xyz[:, 0] * xyz[:, 1] * xyz[:, 2], 0.0))
    np.testing.assert_allclose(u.cpu().numpy(), want, rtol=1e-9, atol=1e-11)
    e.close()
