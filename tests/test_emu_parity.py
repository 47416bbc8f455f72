"""CPU stepping of the product's kernel phase functions (tests/emu) against the oracle — runs without a GPU."""
import pytest

import parity_util as U

CASES = U.cases()


@pytest.fixture(scope="module")
def emu_lib():
    return U.build_emu()


@pytest.mark.parametrize("name,p,kw", CASES, ids=[c[0] for c in CASES])
def test_kernel_phases_match_oracle(emu_lib, name, p, kw):
    U.check_case(p, kw, emu_lib, on_gpu=False)
    U.check_case(p, kw, emu_lib, on_gpu=False, jacobian=False)


@pytest.mark.parametrize("name,p,kw", CASES, ids=[c[0] for c in CASES])
def test_generic_kernel_phases_match_oracle(emu_lib, name, p, kw):
    """the generic kernel alone (no specialised, no batched volume kernel)"""
    r = U.check_case(p, kw, emu_lib, on_gpu=False, force_generic=True, no_batched=True)
    assert r["path"] == "generic", r


@pytest.mark.parametrize("name,p,kw", [c for c in CASES if c[1]["physics"] == "thermal" and c[1]["mesh"]["dim"] == 3],
                         ids=lambda v: v if isinstance(v, str) else "")
def test_generic_kernel_on_specialised_cases(emu_lib, name, p, kw):
    r = U.check_case(p, kw, emu_lib, on_gpu=False, force_generic=True)
    assert r["path"] == "batched", r


# ---- round 2: non-affine meshes, every coefficient operator, BASELINE configs at 8^3, caller-supplied numbering ------------
DISTORTED = U.distorted_cases()
OPCODES = U.opcode_cases()
CONFIGS8 = [c for c in U.config_parity_cases() if "24^3" not in c[0]]


@pytest.mark.parametrize("name,p,kw", DISTORTED, ids=[c[0] for c in DISTORTED])
def test_kernel_phases_on_non_affine_meshes(emu_lib, name, p, kw):
    """every node displaced by 0.2 h (splitmix64 - 0.5): the trilinear cross terms of the Jacobian, the per-point inverse
    and the side normals J t_U x J t_V are non-trivial (on uniform bricks they vanish identically)"""
    U.check_case(p, kw, emu_lib, on_gpu=False)
    U.check_case(p, kw, emu_lib, on_gpu=False, jacobian=False)
    r = U.check_case(p, kw, emu_lib, on_gpu=False, force_generic=True, no_batched=True)
    assert r["path"] == "generic", r


@pytest.mark.parametrize("name,p,kw", OPCODES, ids=[c[0] for c in OPCODES])
def test_every_coefficient_operator_in_a_kernel(emu_lib, name, p, kw):
    U.check_case(p, kw, emu_lib, on_gpu=False)
    U.check_case(p, kw, emu_lib, on_gpu=False, force_generic=True)
    U.check_case(p, kw, emu_lib, on_gpu=False, force_generic=True, no_batched=True)


@pytest.mark.parametrize("name,p,kw", CONFIGS8, ids=[c[0] for c in CONFIGS8])
def test_baseline_configs_at_parity_size(emu_lib, name, p, kw):
    U.check_case(p, kw, emu_lib, on_gpu=False)


FOREIGN = [("C1", "config_c1", 9), ("C2", "config_c2", 5), ("C3", "config_c3", 3), ("C4", "config_c4", 4), ("C5", "config_c5", 4)]


@pytest.mark.parametrize("tag,mk,n", FOREIGN, ids=[f[0] for f in FOREIGN])
def test_caller_supplied_numbering_assembles(emu_lib, tag, mk, n):
    """mrh_set_mesh / set_dofmap / set_crs / add_side with permuted elements, nodes and LIDs on a distorted mesh"""
    import problems
    U.check_foreign_numbering(getattr(problems, mk)(n), emu_lib, False, distort=0.2)
    U.check_foreign_numbering(getattr(problems, mk)(n), emu_lib, False, distort=0.2, jacobian=False)


def test_boundary_cases_with_caller_supplied_numbering(emu_lib):
    for name, p, kw in CASES:
        if "transient" in kw or not (p.get("neumann") or p.get("flux") or p.get("extra", {}).get("use weak Dirichlet")):
            continue
        U.check_foreign_numbering(p, emu_lib, False, distort=0.2)


# ---- L2-projection systems: setDirichlet / setInitial against the oracle and closed forms -----------------------
PROJ = [("thermal 2D", "regression_thermal_2d_nonzero_dbc", 9, 0.0), ("thermal 3D", "config_c2", 4, 0.2), ("LE 3D", "config_c4", 3, 0.2)]


@pytest.mark.parametrize("tag,mk,n,distort", PROJ, ids=[c[0] for c in PROJ])
def test_dirichlet_projection_system_matches_oracle(emu_lib, tag, mk, n, distort):
    import numpy as np
    import problems
    from mrhyde_b200 import Engine
    from oracle.pyoracle import Oracle
    p = getattr(problems, mk)(n)
    if mk != "regression_thermal_2d_nonzero_dbc":
        for v in p["dirichlet"]:
            p["dirichlet"][v] = {"left": "0.3*y+z", "top": "sin(x)+t", "front": "1.0"}
    o = Oracle(U.oracle_problem(p))
    e = Engine.from_problem(p, lib=emu_lib)
    if distort:
        nd = U.distorted_nodes(o.mesh()[0], p["mesh"], distort)
        o.set_nodes(nd); e.set_nodes(nd)
    orhs, omass = o.assemble_dirichlet(time=0.4)
    rhs, mass = np.zeros(o.nlocal), np.zeros(o.nnz)
    e.assemble_projection(0, rhs, mass, time=0.4)
    assert U.rel_errors(rhs, orhs)[0] <= U.TOL and U.rel_errors(mass, omass, o.crs()[0]) <= (U.TOL, U.TOL)


def test_initial_projection_reproduces_a_function_in_the_space(emu_lib):
    """setInitial: the L2 projection of a (bi)linear function is that function; mass rows sum to the nodal volumes"""
    import numpy as np
    import scipy.sparse as sp
    import scipy.sparse.linalg as spl
    import problems
    from mrhyde_b200 import Engine
    p = U.shrink(problems.regression_thermal_2d())
    p["initial"] = {"e": "1.0+2.0*x+y+0.5*x*y"}
    e = Engine.from_problem(p, lib=emu_lib)
    s = e.sizes()
    rhs, mass = np.zeros(s["nlocal"]), np.zeros(s["nnz"])
    e.assemble_projection(1, rhs, mass)
    rp, ci = e.crs()
    M = sp.csr_matrix((mass, ci, rp), shape=(s["nlocal"],) * 2)
    xyz, _ = e.dof_coords()
    want = 1.0 + 2.0 * xyz[:, 0] + xyz[:, 1] + 0.5 * xyz[:, 0] * xyz[:, 1]
    np.testing.assert_allclose(spl.spsolve(M.tocsc(), rhs), want, rtol=1e-11, atol=1e-13)
    assert M.sum() == pytest.approx(1.0, rel=1e-13)                      # area of the unit square
