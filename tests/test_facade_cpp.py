"""The C++ facade include/mrhyde_b200.hpp: a program written against MrHyDE-shaped classes (AssemblyManager, PhysicsBase
modules, FunctionManager, ParameterList laid out like input.yaml), compiled with plain g++ against include/ and linked to
libmrhyde_b200.so, compared with the oracle's C API.  Also pins the C-ABI struct layout against the ctypes mirror."""
import ctypes
import json
import os
import subprocess

import pytest

import parity_util as U

ROOT = U.ROOT
SRC = os.path.join(ROOT, "tests", "cpp", "facade_test.cpp")
EXE = os.path.join(ROOT, "tests", "cpp", "facade_test")


def build_facade_test():
    from mrhyde_b200 import build
    from oracle import pyoracle
    lib = build.build()
    orc = pyoracle.build()
    deps = [SRC, os.path.join(ROOT, "include", "mrhyde_b200.hpp"), os.path.join(ROOT, "include", "mrhyde_b200.h"), lib, orc]
    if (not os.path.exists(EXE)) or any(os.path.getmtime(d) > os.path.getmtime(EXE) for d in deps):
        cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
        subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), "-I" + cuda + "/include", SRC, "-o", EXE,
                               "-L" + os.path.dirname(lib), "-lmrhyde_b200", "-L" + os.path.dirname(orc), "-loracle", "-L" + cuda + "/lib64", "-lcudart",
                               "-Wl,-rpath,$ORIGIN/../../mrhyde_b200", "-Wl,-rpath,$ORIGIN/../../oracle", "-Wl,-rpath," + cuda + "/lib64"])
    return EXE


def test_facade_compiles_and_struct_layout_matches_the_ctypes_mirror():
    exe = build_facade_test()
    lay = json.loads(subprocess.check_output([exe, "--layout"]))
    from mrhyde_b200.engine import MrhConfig, MrhState
    assert lay["sizeof_config"] == ctypes.sizeof(MrhConfig) and lay["sizeof_state"] == ctypes.sizeof(MrhState)
    for name, off in lay["config"].items():
        assert getattr(MrhConfig, name).offset == off, name
    for name, off in lay["state"].items():
        assert getattr(MrhState, name).offset == off, name
    assert set(lay["config"]) == {f[0] for f in MrhConfig._fields_} and set(lay["state"]) == {f[0] for f in MrhState._fields_}


@pytest.mark.gpu
@pytest.mark.parametrize("case,n", [("C2", 12), ("C2T", 7), ("C4", 6), ("C5", 5)])
def test_facade_program_matches_oracle(case, n):
    exe = build_facade_test()
    r = subprocess.run([exe, case, str(n)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["ok"] and out["indexing_bit_exact"], out
    if case.startswith("C2"):
        assert out["bc_top"] == "Neumann", out


@pytest.mark.gpu
def test_yaml_deck_in_matrix_market_out(tmp_path):
    """a MrHyDE-format input deck read by mrh::readInput, J / res dumped as Matrix Market (the files
    linalg->writeToFile leaves behind, solverManager.cpp:1653) and compared entry by entry with the oracle on the same deck"""
    import numpy as np
    import scipy.io
    from oracle.pyoracle import Oracle, config_text
    exe = build_facade_test()
    deck = os.path.join(ROOT, "tests", "golden", "thermal_3d_input.yaml")
    pre = str(tmp_path / "dump")
    r = subprocess.run([exe, "--yaml", deck, pre], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    J = scipy.io.mmread(pre + "_J.mtx").tocsr()
    res = np.asarray(scipy.io.mmread(pre + "_res.mtx")).ravel()
    u = np.asarray(scipy.io.mmread(pre + "_u.mtx")).ravel()
    o = Oracle(config_text(mesh=dict(dim=3, NX=6, NY=5, NZ=7, xmin=0.0, xmax=1.0, ymin=0.0, ymax=2.0, zmin=0.0, zmax=1.0), physics="thermal", order=1,
                           functions={"thermal source": "12*(pi*pi)*sin(2*pi*x)*sin(2*pi*y)*sin(2*pi*z)", "thermal diffusion": "1.0+0.5*x*y"},
                           dirichlet={"e": {"all boundaries": "0.0"}}, neumann={"e": {"top": "1.0+x"}}))
    assert J.shape == (o.nlocal, o.nlocal) and J.nnz <= o.nnz
    ores, ojac = o.assemble(u)
    rp, ci = o.crs()
    import scipy.sparse as sp
    OJ = sp.csr_matrix((ojac, ci, rp), shape=J.shape)
    assert abs(J - OJ).max() <= 1e-12 * np.abs(ojac).max()
    assert np.abs(res - ores).max() <= 1e-12 * np.abs(ores).max()
