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
