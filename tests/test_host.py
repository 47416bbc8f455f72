"""Host-side logic of the product library, checked without a GPU: the library loads and exports every symbol the
header declares, the brick builder's indexing is bit-exact against the oracle, expressions parse like the
reference's interpreter, and every GPU entry point fails loudly when there is no device (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

import parity_util as U
import problems

ROOT = U.ROOT


@pytest.fixture(scope="module")
def lib():
    from mrhyde_b200 import build
    return build.build()


def test_library_exports_every_declared_symbol(lib):
    hdr = open(os.path.join(ROOT, "include", "mrhyde_b200.h")).read()
    declared = set(re.findall(r"\b(mrh_[a-z_0-9]+)\s*\(", hdr))
    L = ctypes.CDLL(lib)
    from mrhyde_b200.engine import SYMBOLS
    assert declared == set(SYMBOLS), declared ^ set(SYMBOLS)
    for s in declared:
        assert hasattr(L, s), s


@pytest.mark.parametrize("mk,n", [(problems.config_c1, 17), (problems.config_c2, 6), (problems.config_c3, 3),
                                  (problems.config_c4, 5), (problems.config_c5, 5)])
def test_indexing_bit_exact_vs_oracle(lib, mk, n):
    from mrhyde_b200 import Engine
    from oracle.pyoracle import Oracle
    p = mk(n)
    e = Engine.from_problem(p, lib=lib, host_only=True)
    o = Oracle(U.oracle_problem(p))
    for a, b in zip(e.dofmap(), o.dofmap()):
        assert np.array_equal(a, b)
    for a, b in zip(e.crs(), o.crs()):
        assert np.array_equal(a, b)
    for a, b in zip(e.mesh(), o.mesh()):
        assert np.array_equal(a, b)
    for (n1, l1, e1), (n2, l2, e2) in zip(e.sides(), o.sides()):
        assert (n1, l1) == (n2, l2) and np.array_equal(e1, e2)
    assert np.array_equal(e.bc_codes(), o.bc_codes())


@pytest.mark.parametrize("side,axis,high", [("left", 0, 0), ("right", 0, 1), ("bottom", 1, 0), ("top", 1, 1), ("back", 2, 0), ("front", 2, 1)])
def test_3d_side_set_names_follow_the_stk_cube_factory(lib, side, axis, high):
    """panzer_stk::CubeHexMeshFactory: bottom / top are the y planes, back / front the z planes (the reference's
    regression/maxwell/Multiblock/input_mesh.yaml pairs 'top;bottom' with periodic type 'xz-all')."""
    from mrhyde_b200 import Engine
    p = problems.config_c2(4)
    p["mesh"].update(NX=3, NY=4, NZ=5, xmax=1.0, ymax=2.0, zmax=3.0)
    p["dirichlet"] = {"e": {side: "0.0"}}
    e = Engine.from_problem(p, lib=lib, host_only=True)
    fixed = e.dofmap()[2].astype(bool)
    xyz, _ = e.dof_coords()
    plane = np.isclose(xyz[:, axis], [1.0, 2.0, 3.0][axis] if high else 0.0)
    assert fixed.any() and np.array_equal(fixed, plane)


def test_foreign_dofmap_path_equals_brick_builder(lib):
    """mrh_set_mesh/dofmap/crs/add_side (what a Panzer/Tpetra caller hands over) gives the same tables."""
    from mrhyde_b200 import Engine
    p = problems.config_c4(3)
    a = Engine.from_problem(p, lib=lib, host_only=True)
    b = Engine("linearelasticity", 3, 1, lib=lib)
    nodes, e2n = a.mesh()
    lids, offs, fixed, _, _ = a.dofmap()
    rp, ci = a.crs()
    b.set_mesh(nodes, e2n); b.set_dofmap(lids, offs, a.sizes()["nlocal"]); b.set_crs(rp, ci)
    for name, ls, el in a.sides():
        b.add_side(name, ls, el)
    for name, expr in p["functions"].items():
        b.add_function(name, expr, "ip"); b.add_function(name, expr, "side ip")
    for v, sides in p["dirichlet"].items():
        for sd, ex in sides.items():
            b.set_bc(v, sd, "dirichlet", ex)
    b.finalize_host()
    assert np.array_equal(b.dofmap(coords=False)[2], fixed)
    assert np.array_equal(b.bc_codes(), a.bc_codes())


def test_expression_known_answers(lib):
    """Interpreter quirks of the reference (interpreter.cpp split rules), product compiler vs numpy."""
    from mrhyde_b200 import Engine
    p = problems.regression_thermal_2d(4)
    p["functions"] = dict(p["functions"])
    exprs = {"q1": "8*(pi*pi)*sin(2*pi*x)*sin(2*pi*y)", "q2": "x/2.0*4.0", "q3": "-x+y", "q4": "3.0*x^2", "A": "2.0",
             "f": "A*x+1.0", "q5": "f*f", "q6": "x<0.5", "q7": "exp(sin((x+1.0)*2.0))", "q8": "max(x,y)+min(x,0.25)",
             "q9": "-1.0*(A*pi)*cos(A*pi*x)", "q10": "1.0e-3*x", "q11": "x-y-0.25", "q12": "x/y/2.0", "q13": "abs(x-0.5)+sqrt(y+1.0)",
             "q14": "t*x", "q15": "1.0+sqrt(x-0.5)"}
    p["functions"].update(exprs)
    e = Engine.from_problem(p, lib=lib, host_only=True)
    rng = np.random.default_rng(3)
    X = rng.random((50, 2)) + 0.01
    x, y = X[:, 0], X[:, 1]
    want = {"q1": 8 * np.pi ** 2 * np.sin(2 * np.pi * x) * np.sin(2 * np.pi * y), "q2": x / 2.0 * 4.0, "q3": (0.0 - x) + y,
            "q4": 3.0 * x ** 2, "q5": (2.0 * x + 1.0) ** 2, "q6": (x < 0.5).astype(float), "q7": np.exp(np.sin((x + 1.0) * 2.0)),
            "q8": np.maximum(x, y) + np.minimum(x, 0.25), "q9": -2.0 * np.pi * np.cos(2 * np.pi * x),
            "q10": 1.0 - 3 * x,          # the reference splits the literal at its sign
            "q11": x - y - 0.25, "q12": x / y / 2.0, "q13": np.abs(x - 0.5) + np.sqrt(y + 1.0), "q14": 0.7 * x,
            "q15": 1.0 + np.sqrt(np.maximum(x - 0.5, 0.0))}      # sqrt of a non-positive argument is 0 (functionManager.cpp:1062-1067)
    for k, w in want.items():
        np.testing.assert_allclose(e.eval_function(k, X, time=0.7), w, rtol=1e-14, atol=1e-15, err_msg=k)
    from oracle.pyoracle import Oracle
    o = Oracle(U.oracle_problem(p))
    _, _, _, ip = o.group_basis(0, 16)
    pts = np.stack([ip[0].ravel(), ip[1].ravel()], axis=1)
    for k in want:
        if k == "q14":
            continue
        np.testing.assert_allclose(e.eval_function(k, pts), o.eval_expression(k, 0, 16).ravel(), rtol=1e-14, atol=1e-15, err_msg=k)


def test_parameter_leaves_known_answers(lib):
    """Scalar / vector parameters as leaves `name`, `name(k)`, `name(kk)` (functionManager.cpp:270-355): product compiler vs numpy and
    vs the oracle; a parameter shadows a function of the same name (it is looked up first), a component out of range is an error"""
    from mrhyde_b200 import Engine, MrhError
    p = problems.regression_thermal_2d(4)
    p["functions"] = dict(p["functions"])
    p["parameters"] = {"kap": [1.25, -7.0, 0.5], "amp": 0.75, "long": list(np.arange(12.0) * 0.5)}
    p["functions"].update({"g1": "kap+kap(2)*x*y+amp*y", "g2": "sin(kap(1)*x)+long(11)*y", "amp2": "amp*amp", "g3": "amp2+x", "kap": "100.0", "g4": "kap+1.0"})
    e = Engine.from_problem(p, lib=lib, host_only=True)
    rng = np.random.default_rng(5)
    X = rng.random((40, 2))
    x, y = X[:, 0], X[:, 1]
    want = {"g1": 1.25 + 0.5 * x * y + 0.75 * y, "g2": np.sin(-7.0 * x) + 5.5 * y, "g3": 0.75 * 0.75 + x, "g4": 1.25 + 1.0 + 0 * x}
    for k, w in want.items():
        np.testing.assert_allclose(e.eval_function(k, X), w, rtol=1e-14, atol=1e-15, err_msg=k)
    from oracle.pyoracle import Oracle
    o = Oracle(U.oracle_problem(p))
    _, _, _, ip = o.group_basis(0, 16)
    pts = np.stack([ip[0].ravel(), ip[1].ravel()], axis=1)
    for k in want:
        np.testing.assert_allclose(e.eval_function(k, pts), o.eval_expression(k, 0, 16).ravel(), rtol=1e-14, atol=1e-15, err_msg=k)
    p["functions"]["bad"] = "kap(7)+x"
    p["functions"]["thermal diffusion"] = "bad"
    with pytest.raises(MrhError, match="out of range"):
        Engine.from_problem(p, lib=lib, host_only=True)


def test_errors_are_reported_not_thrown(lib):
    from mrhyde_b200 import Engine, MrhError
    with pytest.raises(MrhError):
        Engine("maxwell", 3, lib=lib)
    p = problems.regression_thermal_2d(4)
    p["functions"] = {"thermal source": "sin(x"}
    with pytest.raises(MrhError, match="decompose"):
        Engine.from_problem(p, lib=lib, host_only=True)
    p["functions"] = {"thermal source": "emax(x)"}
    with pytest.raises(MrhError, match="not supported"):
        Engine.from_problem(p, lib=lib, host_only=True)
    # out-of-range indices are argument errors, not out-of-bounds device accesses
    e = Engine("thermal", 2, lib=lib)
    with pytest.raises(MrhError, match="out of range"):
        e.set_mesh(np.zeros((4, 2)), np.array([[0, 1, 2, 4]], dtype=np.int32))
    e.set_mesh(np.array([[0.0, 0], [1, 0], [1, 1], [0, 1]]), np.array([[0, 1, 2, 3]], dtype=np.int32))
    with pytest.raises(MrhError, match="out of range"):
        e.add_side("left", 3, np.array([1], dtype=np.int32))


def test_no_cpu_fallback(lib):
    """Without a CUDA device every GPU entry point must fail with MRH_ECUDA."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from mrhyde_b200 import Engine, MrhError
    e = Engine.from_problem(problems.config_c2(2), lib=lib, host_only=True)
    with pytest.raises(MrhError, match="no CPU fallback"):
        e.finalize()
    u = np.zeros(27); r = np.zeros(27); j = np.zeros(e.sizes()["nnz"])
    with pytest.raises(MrhError, match="no CPU fallback"):
        e.assemble_jacres(u, r, j)
