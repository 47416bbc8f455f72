"""GPU parity tests proper: the CUDA library through its C ABI against the oracle (run with -m gpu on a B200)."""
import numpy as np
import pytest

import parity_util as U
import problems

pytestmark = pytest.mark.gpu
CASES = U.cases()


@pytest.fixture(scope="module")
def lib():
    import torch
    from mrhyde_b200 import lib_path
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return lib_path()


@pytest.mark.parametrize("name,p,kw", CASES, ids=[c[0] for c in CASES])
def test_cuda_matches_oracle(lib, name, p, kw):
    U.check_case(p, kw, lib, on_gpu=True)
    U.check_case(p, kw, lib, on_gpu=True, jacobian=False)


@pytest.mark.parametrize("name,p,kw", [c for c in CASES if c[1]["physics"] == "thermal" and c[1]["mesh"]["dim"] == 3],
                         ids=lambda v: v if isinstance(v, str) else "")
def test_cuda_batched_kernel_on_specialised_cases(lib, name, p, kw):
    """thermal 3-D cases through the batched volume kernel instead of the specialised one"""
    r = U.check_case(p, kw, lib, on_gpu=True, force_generic=True)
    assert r["path"] == "batched", r


@pytest.mark.parametrize("name,p,kw", CASES, ids=[c[0] for c in CASES])
def test_cuda_generic_kernel_alone(lib, name, p, kw):
    """every case through the generic kernel only (no specialised, no batched volume kernel)"""
    r = U.check_case(p, kw, lib, on_gpu=True, force_generic=True, no_batched=True)
    assert r["path"] == "generic", r


JIT_BATCHED = ("cdr/2D_manufactured", "le/3D_manufactured", "C3 cdr Q2 hex", "LE 3D weak Dirichlet + Neumann", "LE 2D weak Dirichlet",
               "thermal 3D nonlinear diffusion + advection", "thermal 3D transient BE", "cdr 2D transient DIRK2", "thermal 2D lump mass")


@pytest.mark.parametrize("name,p,kw", [c for c in CASES if c[0] in JIT_BATCHED], ids=lambda v: v if isinstance(v, str) else "")
def test_cuda_nvrtc_specialised_batched_kernel(lib, name, p, kw):
    """the batched volume kernel with its state-independent coefficient programs compiled by NVRTC (forced on small meshes)"""
    r = U.check_case(p, kw, lib, on_gpu=True, force_generic=True, jit_batched=True)
    assert r["path"] == "nvrtc-batched", r


def test_default_path_of_multivariable_configs_is_batched(lib):
    for mk, n in ((problems.config_c3, 3), (problems.config_c4, 5), (problems.config_c5, 4)):
        r = U.check_case(mk(n), {}, lib, on_gpu=True)
        assert r["path"] == "batched", r


@pytest.mark.parametrize("name,p,kw", [c for c in CASES if c[1]["physics"] == "thermal" and c[1]["mesh"]["dim"] == 3],
                         ids=lambda v: v if isinstance(v, str) else "")
@pytest.mark.parametrize("no_jit,no_tiles,no_ws", [(True, False, False), (False, True, False), (True, True, False), (False, False, True)])
def test_cuda_specialised_kernel_variants(lib, name, p, kw, no_jit, no_tiles, no_ws):
    """the specialised kernel without the NVRTC step (bytecode interpreter inside the kernel), with the direct
    instead of the tile-staged scatter, and the tile-staged scatter without warp specialisation"""
    r = U.check_case(p, kw, lib, on_gpu=True, no_jit=no_jit, no_tiles=no_tiles, no_ws=no_ws)
    if r["path"] not in ("generic", "batched"):   # cases with state-dependent coefficients use the batched kernel
        assert ("nvrtc" in r["path"]) == (not no_jit) and ("tiled" in r["path"]) == (not no_tiles), r
        assert ("-ws" in r["path"]) == (not no_jit and not no_tiles and not no_ws), r
    U.check_case(p, kw, lib, on_gpu=True, no_jit=no_jit, no_tiles=no_tiles, no_ws=no_ws, jacobian=False)


@pytest.mark.parametrize("name,p,kw", [c for c in CASES if c[1]["physics"] == "thermal" and c[1]["mesh"]["dim"] == 2],
                         ids=lambda v: v if isinstance(v, str) else "")
@pytest.mark.parametrize("force_generic,no_jit", [(False, False), (False, True), (True, False)])
def test_cuda_thermal_quad_kernel_variants(lib, name, p, kw, force_generic, no_jit):
    """2-D thermal Q1 cases through the specialised quadrilateral kernel (NVRTC-printed coefficients, or the bytecode
    interpreter inside the kernel) and through the generic kernel"""
    r = U.check_case(p, kw, lib, on_gpu=True, force_generic=force_generic, no_jit=no_jit)
    if r["path"] not in ("generic", "batched"):
        assert not force_generic
        assert r["path"] in (("specialised-2d", "specialised-2d+tiled") if no_jit else ("nvrtc-specialised-2d", "nvrtc-specialised-2d+tiled")), r
    U.check_case(p, kw, lib, on_gpu=True, force_generic=force_generic, no_jit=no_jit, jacobian=False)


def test_default_path_is_nvrtc_tiled_warp_specialised(lib):
    r = U.check_case(problems.config_c2(6), {}, lib, on_gpu=True)
    assert r["path"] in ("nvrtc-specialised+tiled-ws", "nvrtc-specialised+tiled-ws8", "nvrtc-specialised+tiled-wsk"), r


@pytest.mark.parametrize("n", [(5, 3, 7), (8, 4, 4), (17, 9, 6)])
def test_tiled_scatter_on_ragged_meshes(lib, n):
    """mesh sizes that do not divide the 8x4x4 tile, and one that is exactly one tile"""
    p = problems.config_c2(4)
    p["mesh"]["NX"], p["mesh"]["NY"], p["mesh"]["NZ"] = n
    U.check_case(p, {}, lib, on_gpu=True)
    U.check_case(p, {}, lib, on_gpu=True, no_ws=True)


@pytest.mark.parametrize("n", [(5, 3), (16, 8), (37, 19), (64, 40)])
@pytest.mark.parametrize("no_jit", [False, True])
def test_tiled_scatter_2d_on_ragged_meshes(lib, n, no_jit):
    """Q1 quadrilateral kernel with the tile-staged scatter: meshes that do not divide the 16x8 tile, exactly one tile, several tiles;
    NVRTC-printed and ahead-of-time kernels; distorted nodes and a caller-permuted numbering go through test_foreign_numbering"""
    p = problems.config_c1(4)
    p["mesh"]["NX"], p["mesh"]["NY"] = n
    r = U.check_case(p, {}, lib, on_gpu=True, no_jit=no_jit, tiled2d=True)      # flag bit 8: off by default (measured slower than the direct scatter)
    assert r["path"] == ("specialised-2d+tiled" if no_jit else "nvrtc-specialised-2d+tiled"), r
    r = U.check_case(p, {}, lib, on_gpu=True, no_jit=no_jit)
    assert r["path"] == ("specialised-2d" if no_jit else "nvrtc-specialised-2d"), r


def test_mid_size_c2_against_oracle(lib):
    """24^3 thermal Q1 (SURVEY.md 8d parity size) through the specialised kernel."""
    r = U.check_case(problems.config_c2(24), {}, lib, on_gpu=True)
    assert r["launches"] == 2          # one volume kernel + the DBC diagonal kernel


@pytest.mark.parametrize("n", [96, 256])
def test_large_mesh_properties(lib, n):
    """Size-independent properties at sizes the oracle does not reach in seconds (thermal Q1 at 96^3 and at
    BASELINE.json's full 256^3):
    Jacobian rows of non-fixed interior DOFs sum to ~0 (constants are in the Laplacian's kernel), J is symmetric in
    its action (x.Jy == y.Jx on free DOFs), res(u) = -(J u - b) is affine in u (linearity), and repeated assembly
    accumulates exactly twice the values up to atomics ordering; the stencil of an interior row equals the stencil
    of the oracle-checked 24^3 mesh scaled by the mesh-width ratio (uniform bricks, constant diffusion)."""
    import torch
    from mrhyde_b200 import Engine
    p = problems.config_c2(n)
    e = Engine.from_problem(p, lib=lib)
    s = e.sizes()
    rp, ci = e.crs()
    xyz, var = e.dof_coords()
    _, _, fixed, _, _ = e.dofmap()
    dev = torch.device("cuda:0")
    u = torch.from_numpy(problems.synthetic_state(xyz, var)).to(dev)
    res = torch.zeros(s["nlocal"], dtype=torch.float64, device=dev)
    jac = torch.zeros(s["nnz"], dtype=torch.float64, device=dev)
    e.assemble_jacres(u, res, jac)
    torch.cuda.synchronize()
    J = torch.sparse_csr_tensor(torch.from_numpy(rp).to(dev), torch.from_numpy(ci.astype(np.int64)).to(dev), jac,
                                size=(s["nlocal"], s["nlocal"]))
    free = torch.from_numpy(fixed == 0).to(dev)
    ones = torch.ones(s["nlocal"], dtype=torch.float64, device=dev)
    rowsum = (J @ ones)
    # rows whose whole stencil is free
    touches_fixed = (J.abs() @ (~free).double()) > 0
    interior = free & ~touches_fixed
    assert rowsum[interior].abs().max().item() < 1e-12 * jac.abs().max().item() * 27
    # fixed rows are identity rows
    assert torch.allclose(rowsum[~free], torch.ones_like(rowsum[~free]))
    # affine residual: res(u) - res(0) == -J u on free rows
    res0 = torch.zeros_like(res)
    e.assemble_res(torch.zeros_like(u), res0)
    lin = -(J @ u)
    d = (res - res0 - lin)[free].abs().max().item()
    assert d < 1e-11 * max(res.abs().max().item(), 1.0)
    # accumulate semantics
    e.assemble_jacres(u, res, jac)
    torch.cuda.synchronize()
    rowsum2 = (torch.sparse_csr_tensor(torch.from_numpy(rp).to(dev), torch.from_numpy(ci.astype(np.int64)).to(dev), jac,
                                       size=(s["nlocal"], s["nlocal"])) @ ones)
    assert torch.allclose(rowsum2[interior], 2 * rowsum[interior], atol=1e-9)
    assert s["npatterns"] >= 1 and s["npatterns"] < 4096
    # cross-size check against the mesh size that is compared with the oracle entry by entry
    def centre_row(eng, m, values):
        sz = eng.sizes(); rp_, ci_ = eng.crs(); xyz_, _ = eng.dof_coords()
        c = int(np.argmin(np.abs(xyz_ - 0.5).sum(axis=1)))          # the DOF at the centre of the unit cube
        seg = slice(int(rp_[c]), int(rp_[c + 1]))
        cols = ci_[seg]
        order = np.lexsort((xyz_[cols, 0], xyz_[cols, 1], xyz_[cols, 2]))      # geometric order of the 27 neighbours
        return values[seg].cpu().numpy()[order]
    small = Engine.from_problem(problems.config_c2(24), lib=lib)
    ss = small.sizes(); sxyz, svar = small.dof_coords()
    su = torch.from_numpy(problems.synthetic_state(sxyz, svar)).to(dev)
    sres = torch.zeros(ss["nlocal"], dtype=torch.float64, device=dev); sjac = torch.zeros(ss["nnz"], dtype=torch.float64, device=dev)
    small.assemble_jacres(su, sres, sjac)
    torch.cuda.synchronize()
    big_row = 0.5 * centre_row(e, n, jac)                              # jac holds two accumulated assemblies
    small_row = centre_row(small, 24, sjac)
    assert big_row.shape == small_row.shape == (27,)
    assert np.abs(big_row - small_row * (24.0 / n)).max() < 1e-12 * np.abs(big_row).max()


def test_host_buffer_entry_point(lib):
    from mrhyde_b200 import Engine
    p = problems.config_c2(12)
    e = Engine.from_problem(p, lib=lib)
    s = e.sizes()
    xyz, var = e.dof_coords()
    u = problems.synthetic_state(xyz, var)
    res = np.zeros(s["nlocal"]); jac = np.zeros(s["nnz"])
    e.assemble_jacres_host(u, res, jac)
    from oracle.pyoracle import Oracle
    o = Oracle(U.oracle_problem(p))
    ores, ojac = o.assemble(u)
    assert U.rel_errors(res, ores)[0] < U.TOL and U.rel_errors(jac, ojac)[0] < U.TOL


def test_host_buffer_entry_point_chunked_copy_out(lib):
    """96^3 is large enough for the host-buffer entry point to assemble in chunks of tiles and copy completed CRS
    rows out while later chunks run: the result must equal the device-buffer entry point's."""
    import torch
    from mrhyde_b200 import Engine
    e = Engine.from_problem(problems.config_c2(96), lib=lib)
    s = e.sizes()
    xyz, var = e.dof_coords()
    u = problems.synthetic_state(xyz, var)
    hu = torch.from_numpy(u).pin_memory()
    hres = torch.full((s["nlocal"],), 7.0, dtype=torch.float64).pin_memory()
    hjac = torch.full((s["nnz"],), 7.0, dtype=torch.float64).pin_memory()
    e.assemble_jacres_host(hu, hres, hjac)
    n_host_launches = e.launch_count()
    dev = torch.device("cuda:0")
    res = torch.zeros(s["nlocal"], dtype=torch.float64, device=dev); jac = torch.zeros(s["nnz"], dtype=torch.float64, device=dev)
    e.assemble_jacres(torch.from_numpy(u).to(dev), res, jac)
    torch.cuda.synchronize()
    assert n_host_launches > e.launch_count()        # several volume launches (chunks) against one
    assert U.rel_errors(hres.numpy(), res.cpu().numpy())[0] < U.TOL
    assert U.rel_errors(hjac.numpy(), jac.cpu().numpy())[0] < U.TOL


def _assemble(lib, p, **flags):
    import torch
    from mrhyde_b200 import Engine
    e = Engine.from_problem(p, lib=lib, **flags)
    s = e.sizes()
    xyz, var = e.dof_coords()
    dev = torch.device("cuda:0")
    u = torch.from_numpy(problems.synthetic_state(xyz, var)).to(dev)
    res = torch.zeros(s["nlocal"], dtype=torch.float64, device=dev)
    jac = torch.zeros(s["nnz"], dtype=torch.float64, device=dev)
    e.assemble_jacres(u, res, jac)
    torch.cuda.synchronize()
    return e, u, res, jac


@pytest.mark.parametrize("mk,n,jit", [(problems.config_c4, 24, False), (problems.config_c4, 24, True), (problems.config_c5, 24, False),
                                      (problems.config_c3, 12, False), (problems.config_c3, 12, True)],
                         ids=["C4-24", "C4-24-nvrtc", "C5-24", "C3-12", "C3-12-nvrtc"])
def test_batched_kernel_matches_generic_kernel_at_mid_size(lib, mk, n, jit):
    """meshes large enough that every warp of the persistent batched kernel walks several elements (software-pipelined
    gather, element block reuse): batched (and NVRTC-batched) against the oracle-checked generic kernel"""
    p = mk(n)
    e0, u0, res0, jac0 = _assemble(lib, p, force_generic=True, no_batched=True)
    assert e0.kernel_path() == "generic"
    e1, _, res1, jac1 = _assemble(lib, p, jit_batched=jit)
    assert e1.kernel_path() == ("nvrtc-batched" if jit else "batched")
    import torch
    for a, b in ((res1, res0), (jac1, jac0)):
        scale = b.abs().max().item()
        assert (a - b).abs().max().item() <= 1e-12 * scale
    # residual-only passes of the two kernels
    r0, r1 = torch.zeros_like(res0), torch.zeros_like(res0)
    e0.assemble_res(u0, r0); e1.assemble_res(u0, r1)
    torch.cuda.synchronize()
    assert e0.kernel_path() == "generic" and e1.kernel_path() == ("nvrtc-batched" if jit else "batched")
    assert (r1 - r0).abs().max().item() <= 1e-12 * r0.abs().max().item()
    assert (r1 - res1).abs().max().item() <= 1e-12 * r0.abs().max().item()
    e0.close(); e1.close()


@pytest.mark.parametrize("mk,n", [(problems.config_c5, 32), (problems.config_c3, 16), (problems.config_c4, 48)], ids=["C5-32", "C3-16", "C4-48"])
def test_batched_jacobian_is_the_derivative_of_the_residual(lib, mk, n):
    """size-independent property beyond the oracle's reach: for a random direction v on the free DOFs,
    (res(u + eps v) - res(u - eps v)) / (2 eps) == -J(u) v  (the residual vector holds -r, the matrix dr/du);
    exact up to round-off for linear elasticity, O(eps^2) for the nonlinear configurations"""
    import torch
    e, u, res, jac = _assemble(lib, mk(n), jit_batched=True)
    s = e.sizes()
    rp, ci = e.crs()
    _, _, fixed, _, _ = e.dofmap()
    dev = u.device
    J = torch.sparse_csr_tensor(torch.from_numpy(rp).to(dev), torch.from_numpy(ci.astype(np.int64)).to(dev), jac, size=(s["nlocal"], s["nlocal"]))
    free = torch.from_numpy(fixed == 0).to(dev)
    g = torch.Generator(device="cpu").manual_seed(7)
    v = torch.rand(s["nlocal"], dtype=torch.float64, generator=g).to(dev) - 0.5
    eps = 1e-5
    rp_, rm_ = torch.zeros_like(res), torch.zeros_like(res)
    e.assemble_res(u + eps * v, rp_)
    e.assemble_res(u - eps * v, rm_)
    torch.cuda.synchronize()
    fd = (rp_ - rm_) / (2 * eps)
    jv = -(J @ v)
    scale = jv[free].abs().max().item()
    assert (fd - jv)[free].abs().max().item() <= 2e-7 * scale
    e.close()


# ================================================================================================================
# round 2: parity holes named by VERDICT r1 - non-affine meshes, SURVEY 8d parity sizes against the ORACLE, every
# coefficient operator inside a kernel, caller-supplied numbering through the drop-in entry points.
# ================================================================================================================
DISTORTED = U.distorted_cases()
OPCODES = U.opcode_cases()
CONFIGS = U.config_parity_cases()


@pytest.mark.parametrize("name,p,kw", DISTORTED, ids=[c[0] for c in DISTORTED])
def test_non_affine_mesh_default_paths(lib, name, p, kw):
    U.check_case(p, kw, lib, on_gpu=True)
    U.check_case(p, kw, lib, on_gpu=True, jacobian=False)


@pytest.mark.parametrize("name,p,kw", DISTORTED, ids=[c[0] for c in DISTORTED])
def test_non_affine_mesh_generic_and_batched_kernels(lib, name, p, kw):
    r = U.check_case(p, kw, lib, on_gpu=True, force_generic=True, no_batched=True)
    assert r["path"] == "generic", r
    r = U.check_case(p, kw, lib, on_gpu=True, force_generic=True)
    assert r["path"] == "batched", r
    r = U.check_case(p, kw, lib, on_gpu=True, force_generic=True, jacobian=False)


@pytest.mark.parametrize("name,p,kw", [c for c in DISTORTED if c[1]["physics"] == "thermal"], ids=lambda v: v if isinstance(v, str) else "")
@pytest.mark.parametrize("no_jit,no_tiles,no_ws", [(True, False, False), (False, True, False), (True, True, False), (False, False, True)])
def test_non_affine_mesh_specialised_kernel_variants(lib, name, p, kw, no_jit, no_tiles, no_ws):
    """2-D and 3-D specialised thermal kernels: bytecode interpreter / NVRTC, direct / tile-staged / warp-specialised scatter"""
    U.check_case(p, kw, lib, on_gpu=True, no_jit=no_jit, no_tiles=no_tiles, no_ws=no_ws)


@pytest.mark.parametrize("name,p,kw", [c for c in DISTORTED if c[0].split(" [")[0] in JIT_BATCHED], ids=lambda v: v if isinstance(v, str) else "")
def test_non_affine_mesh_nvrtc_batched(lib, name, p, kw):
    r = U.check_case(p, kw, lib, on_gpu=True, force_generic=True, jit_batched=True)
    assert r["path"] == "nvrtc-batched", r


@pytest.mark.parametrize("name,p,kw", CONFIGS, ids=[c[0] for c in CONFIGS])
def test_baseline_configs_against_oracle_at_parity_sizes(lib, name, p, kw):
    """C2..C5 at 8^3 and 24^3, C1 at 40x40: entry by entry against the oracle, default kernels"""
    r = U.check_case(p, kw, lib, on_gpu=True)
    assert r["path"] in ("nvrtc-specialised+tiled-ws", "nvrtc-specialised+tiled-ws8", "nvrtc-specialised+tiled-wsk", "batched", "nvrtc-specialised-2d", "nvrtc-specialised-2d+tiled"), r
    U.check_case(p, kw, lib, on_gpu=True, jacobian=False)


@pytest.mark.parametrize("name,p,kw", [c for c in CONFIGS if c[0] in ("C3 24^3", "C4 24^3", "C5 24^3", "C3 8^3")], ids=lambda v: v if isinstance(v, str) else "")
def test_baseline_configs_nvrtc_batched_against_oracle(lib, name, p, kw):
    want = "batched" if name.startswith("C5") else "nvrtc-batched"     # C5's coefficients are constants: nothing to specialise
    r = U.check_case(p, kw, lib, on_gpu=True, jit_batched=True)
    assert r["path"] == want, r
    kw2 = dict(kw); kw2["distort"] = 0.2
    r = U.check_case(p, kw2, lib, on_gpu=True, jit_batched=True)
    assert r["path"] == want, r


@pytest.mark.parametrize("name,p,kw", OPCODES, ids=[c[0] for c in OPCODES])
def test_every_coefficient_operator_on_the_device(lib, name, p, kw):
    """each operator of bytecode.h inside a kernel: NVRTC-printed, in-kernel vector interpreter, PAD interpreter of the
    batched kernel, NVRTC-printed batched, generic kernel"""
    state = "(state)" in name
    r = U.check_case(p, kw, lib, on_gpu=True)
    assert r["path"].replace("-ws8", "-ws").replace("-wsk", "-ws").replace("-2d+tiled", "-2d") == ("batched" if state else ("nvrtc-specialised-2d" if p["mesh"]["dim"] == 2 else "nvrtc-specialised+tiled-ws")), r
    U.check_case(p, kw, lib, on_gpu=True, no_jit=True)
    r = U.check_case(p, kw, lib, on_gpu=True, force_generic=True, no_batched=True)
    assert r["path"] == "generic", r
    if not state:
        r = U.check_case(p, kw, lib, on_gpu=True, force_generic=True, jit_batched=True)
        assert r["path"] == "nvrtc-batched", r


FOREIGN = [("C1", "config_c1", 33), ("C2", "config_c2", 12), ("C3", "config_c3", 5), ("C4", "config_c4", 8), ("C5", "config_c5", 8)]


@pytest.mark.parametrize("tag,mk,n", FOREIGN, ids=[f[0] for f in FOREIGN])
def test_caller_supplied_numbering_on_the_device(lib, tag, mk, n):
    """Random element order, node ids and LIDs handed over through mrh_set_mesh / set_dofmap / set_crs / add_side on a
    distorted mesh (what a Panzer / Tpetra caller does): result equals the oracle's mapped through the permutations.
    The slot-pattern dictionary degenerates (about one pattern per element) and the tile plan falls back - both must
    still be correct."""
    p = getattr(problems, mk)(n)
    r = U.check_foreign_numbering(p, lib, True, distort=0.2)
    assert r["npatterns"] >= 1
    want = {"C1": ("nvrtc-specialised-2d", "nvrtc-specialised-2d+tiled"), "C2": ("nvrtc-specialised+tiled-ws", "nvrtc-specialised+tiled-ws8", "nvrtc-specialised+tiled-wsk", "nvrtc-specialised+tiled", "nvrtc-specialised")}.get(tag, ("batched",))
    assert r["path"] in want, r
    U.check_foreign_numbering(p, lib, True, distort=0.2, jacobian=False)
    r = U.check_foreign_numbering(p, lib, True, distort=0.2, force_generic=True, no_batched=True)
    assert r["path"] == "generic", r
    if tag == "C2":
        for fl in (dict(no_ws=True), dict(no_tiles=True), dict(no_jit=True)):
            U.check_foreign_numbering(p, lib, True, distort=0.2, **fl)


def test_boundary_cases_with_caller_supplied_numbering_on_the_device(lib):
    for name, p, kw in CASES:
        if "transient" in kw or not (p.get("neumann") or p.get("flux") or p.get("extra", {}).get("use weak Dirichlet")):
            continue
        U.check_foreign_numbering(p, lib, True, distort=0.2)


def test_set_nodes_after_finalize_moves_the_mesh(lib):
    """mrh_set_nodes on a finalized engine refreshes the device coordinates only (no re-plan): assemble, move, assemble"""
    import torch
    from mrhyde_b200 import Engine
    from oracle.pyoracle import Oracle
    p = problems.config_c2(9)
    e = Engine.from_problem(p, lib=lib)
    o = Oracle(U.oracle_problem(p))
    xyz, var = e.dof_coords()
    u = problems.synthetic_state(xyz, var)
    dev = torch.device("cuda:0")
    for amp in (0.0, 0.2, 0.1):
        nd = U.distorted_nodes(o.mesh()[0], p["mesh"], amp) if amp else o.mesh()[0]
        o.set_nodes(nd); e.set_nodes(nd)
        ores, ojac = o.assemble(u)
        res = torch.zeros(o.nlocal, dtype=torch.float64, device=dev); jac = torch.zeros(o.nnz, dtype=torch.float64, device=dev)
        e.assemble_jacres(torch.from_numpy(u).to(dev), res, jac)
        torch.cuda.synchronize()
        assert U.rel_errors(res.cpu().numpy(), ores)[0] <= U.TOL and U.rel_errors(jac.cpu().numpy(), ojac)[0] <= U.TOL
