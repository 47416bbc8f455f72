"""Allocation-free entry point mrh_assemble_res_host_static (static device scratch instead of cudaMalloc'ed staging):
its output against the existing device-buffer entry point mrh_assemble_res, through the C ABI, on two calls in a row.
Collected last on purpose (file name)."""
import numpy as np
import pytest

import parity_util as U
import problems

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    import torch
    from mrhyde_b200 import lib_path
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return lib_path()


def _reference_residual(e, u):
    """existing entry point: caller-owned device buffers (torch allocates them; the engine does not)"""
    import torch
    dev = torch.device("cuda:0")
    res = torch.zeros(len(u), dtype=torch.float64, device=dev)
    e.assemble_res(torch.from_numpy(u).to(dev), res)
    torch.cuda.synchronize()
    return res.cpu().numpy()


@pytest.mark.parametrize("mk,n", [(problems.config_c2, 20), (problems.config_c4, 9), (problems.config_c1, 40)], ids=["C2-20", "C4-9", "C1-40"])
def test_static_scratch_entry_point_matches_device_buffer_entry_point(lib, mk, n):
    from mrhyde_b200 import Engine
    e = Engine.from_problem(mk(n), lib=lib)
    xyz, var = e.dof_coords()
    u1 = problems.synthetic_state(xyz, var)
    u2 = problems.synthetic_state(xyz, var, seed=99) * 0.7 + 0.1       # second call: other input, the scratch holds call 1's data
    for u in (u1, u2):
        want = _reference_residual(e, u)
        got = np.full(len(u), 123.0)                                     # must be overwritten, not accumulated into
        e.assemble_res_host_static(u, got)
        err, _ = U.rel_errors(got, want)
        assert err <= U.TOL, err
    # the oracle agrees as well (first input)
    from oracle.pyoracle import Oracle
    ores, _ = Oracle(U.oracle_problem(mk(n))).assemble(u1, jacobian=False)
    got = np.zeros(len(u1))
    e.assemble_res_host_static(u1, got)
    assert U.rel_errors(got, ores)[0] <= U.TOL
    e.close()


def test_static_scratch_entry_point_rejects_oversized_meshes(lib):
    from mrhyde_b200 import Engine, MrhError
    p = problems.config_c1(1100)                     # 1101^2 = 1 212 201 DOFs > MRH_STATIC_SCRATCH_MAX_DOFS = 1 048 576
    e = Engine.from_problem(p, lib=lib)
    n = e.sizes()["nlocal"]
    assert n > 1048576
    with pytest.raises(MrhError, match="exceed the static scratch limit"):
        e.assemble_res_host_static(np.zeros(n), np.zeros(n))
    e.close()
