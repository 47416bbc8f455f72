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
