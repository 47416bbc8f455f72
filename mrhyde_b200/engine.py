"""ctypes binding of include/mrhyde_b200.h (test / bench harness; not on the hot path)."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBS = {}


class MrhError(RuntimeError):
    pass


class MrhConfig(C.Structure):
    _fields_ = [("dim", C.c_int32), ("physics", C.c_char_p), ("order", C.c_int32), ("quadrature", C.c_int32),
                ("side_quadrature", C.c_int32), ("use_strong_dbc", C.c_int32), ("fix_zero_rows", C.c_int32),
                ("lump_mass", C.c_int32), ("include_advection", C.c_int32), ("use_supg", C.c_int32),
                ("use_pspg", C.c_int32), ("form_param", C.c_double), ("le_penalty", C.c_double),
                ("device", C.c_int32), ("flags", C.c_int32), ("use_lsic", C.c_int32), ("use_evisc", C.c_int32)]


class MrhState(C.Structure):
    _fields_ = [("u", C.c_void_p), ("u_prev", C.c_void_p), ("u_stage", C.c_void_p), ("is_transient", C.c_int32),
                ("nstages", C.c_int32), ("nsteps", C.c_int32), ("stage", C.c_int32), ("butcher_A", C.c_void_p),
                ("butcher_b", C.c_void_p), ("bdf_wts", C.c_void_p), ("time", C.c_double), ("deltat", C.c_double),
                ("stream", C.c_void_p)]


class MrhSolverOpts(C.Structure):
    _fields_ = [("max_nonlinear_iters", C.c_int32), ("use_relative_tol", C.c_int32), ("use_absolute_tol", C.c_int32),
                ("allow_backtracking", C.c_int32), ("nonlinear_tol", C.c_double), ("nonlinear_abs_tol", C.c_double),
                ("linear_method", C.c_int32), ("max_linear_iters", C.c_int32), ("krylov_vectors", C.c_int32), ("verbosity", C.c_int32),
                ("linear_tol", C.c_double)]


class MrhSolverInfo(C.Structure):
    _fields_ = [("nonlinear_iters", C.c_int32), ("linear_iters", C.c_int32), ("converged", C.c_int32), ("time_steps", C.c_int32),
                ("resnorm_first", C.c_double), ("resnorm_last", C.c_double), ("last_linear_relres", C.c_double), ("final_time", C.c_double)]


SYMBOLS = ["mrh_default_config", "mrh_create", "mrh_destroy", "mrh_last_error", "mrh_build_brick", "mrh_set_mesh", "mrh_set_nodes",
           "mrh_set_dofmap", "mrh_set_crs", "mrh_add_side", "mrh_set_bc", "mrh_add_function", "mrh_set_parameter", "mrh_finalize",
           "mrh_finalize_host", "mrh_assemble_res", "mrh_assemble_jacres", "mrh_assemble_jacres_host", "mrh_assemble_res_host_static",
           "mrh_get_sizes", "mrh_get_mesh", "mrh_get_dofmap", "mrh_get_crs", "mrh_get_side", "mrh_get_bc",
           "mrh_get_varname", "mrh_eval_function_host", "mrh_last_launch_count", "mrh_last_volume_kernel_ms",
           "mrh_microbench", "mrh_last_kernel_path", "mrh_build_brick_part", "mrh_set_partition", "mrh_get_partition", "mrh_comm_unique_id",
           "mrh_comm_init", "mrh_exchange", "mrh_get_ghost_block", "mrh_exchange_meta_size", "mrh_exchange_meta_get", "mrh_exchange_meta_put",
           "mrh_get_exchange_table", "mrh_assemble_projection", "mrh_set_dirichlet", "mrh_set_initial", "mrh_default_solver_opts", "mrh_linear_solve", "mrh_nonlinear_solve",
           "mrh_transient_solve"]


def lib_path():
    return os.path.join(_HERE, "libmrhyde_b200.so")


def _bind(L, name, what, value):
    """argtypes / restype of one symbol; a library older than this file may lack newer entry points - the strict check that the
    library exports everything the header declares is tests/test_host.py::test_library_exports_every_declared_symbol"""
    if hasattr(L, name):
        setattr(getattr(L, name), what, value)


def load_library(path=None):
    path = path or lib_path()
    if path in _LIBS:
        return _LIBS[path]
    if not os.path.exists(path):
        raise MrhError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                       "(there is no CPU fallback)")
    L = C.CDLL(path)
    L.mrh_last_error.restype = C.c_char_p
    L.mrh_last_error.argtypes = [C.c_void_p]
    L.mrh_get_varname.restype = C.c_char_p
    L.mrh_get_varname.argtypes = [C.c_void_p, C.c_int32]
    L.mrh_last_kernel_path.restype = C.c_char_p
    L.mrh_last_kernel_path.argtypes = [C.c_void_p]
    L.mrh_last_launch_count.restype = C.c_int64
    L.mrh_last_launch_count.argtypes = [C.c_void_p]
    L.mrh_last_volume_kernel_ms.restype = C.c_double
    L.mrh_last_volume_kernel_ms.argtypes = [C.c_void_p]
    L.mrh_create.argtypes = [C.POINTER(MrhConfig), C.POINTER(C.c_void_p)]
    for n in ("mrh_destroy", "mrh_finalize", "mrh_finalize_host"):
        getattr(L, n).argtypes = [C.c_void_p]
    _bind(L, "mrh_build_brick", "argtypes", [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p])
    _bind(L, "mrh_set_mesh", "argtypes", [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64])
    _bind(L, "mrh_set_nodes", "argtypes", [C.c_void_p, C.c_void_p, C.c_int64])
    _bind(L, "mrh_set_dofmap", "argtypes", [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64])
    _bind(L, "mrh_set_crs", "argtypes", [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64])
    _bind(L, "mrh_add_side", "argtypes", [C.c_void_p, C.c_char_p, C.c_int32, C.c_void_p, C.c_int64])
    _bind(L, "mrh_set_bc", "argtypes", [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int32, C.c_char_p])
    _bind(L, "mrh_add_function", "argtypes", [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p])
    _bind(L, "mrh_set_parameter", "argtypes", [C.c_void_p, C.c_char_p, C.POINTER(C.c_double), C.c_int32])
    _bind(L, "mrh_assemble_res", "argtypes", [C.c_void_p, C.POINTER(MrhState), C.c_void_p])
    _bind(L, "mrh_assemble_jacres", "argtypes", [C.c_void_p, C.POINTER(MrhState), C.c_void_p, C.c_void_p])
    _bind(L, "mrh_assemble_jacres_host", "argtypes", [C.c_void_p, C.POINTER(MrhState), C.c_void_p, C.c_void_p])
    _bind(L, "mrh_assemble_res_host_static", "argtypes", [C.c_void_p, C.POINTER(MrhState), C.c_void_p])
    _bind(L, "mrh_get_sizes", "argtypes", [C.c_void_p, C.c_void_p])
    _bind(L, "mrh_get_mesh", "argtypes", [C.c_void_p, C.c_void_p, C.c_void_p])
    _bind(L, "mrh_get_dofmap", "argtypes", [C.c_void_p] + [C.c_void_p] * 5)
    _bind(L, "mrh_get_crs", "argtypes", [C.c_void_p, C.c_void_p, C.c_void_p])
    _bind(L, "mrh_get_side", "argtypes", [C.c_void_p, C.c_int32, C.c_char_p, C.c_int32, C.c_void_p, C.c_void_p])
    _bind(L, "mrh_get_bc", "argtypes", [C.c_void_p, C.c_int32, C.c_int32])
    _bind(L, "mrh_eval_function_host", "argtypes", [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int64, C.c_void_p, C.c_double, C.c_void_p])
    _bind(L, "mrh_microbench", "argtypes", [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p])
    _bind(L, "mrh_build_brick_part", "argtypes", [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32])
    _bind(L, "mrh_set_partition", "argtypes", [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32])
    _bind(L, "mrh_get_partition", "argtypes", [C.c_void_p, C.c_void_p, C.c_void_p])
    _bind(L, "mrh_comm_unique_id", "argtypes", [C.c_void_p])
    _bind(L, "mrh_comm_init", "argtypes", [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32])
    _bind(L, "mrh_exchange", "argtypes", [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p])
    _bind(L, "mrh_get_ghost_block", "argtypes", [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p])
    _bind(L, "mrh_get_ghost_block", "restype", C.c_int64)
    _bind(L, "mrh_exchange_meta_size", "argtypes", [C.c_void_p, C.c_int32])
    _bind(L, "mrh_exchange_meta_size", "restype", C.c_int64)
    _bind(L, "mrh_exchange_meta_get", "argtypes", [C.c_void_p, C.c_int32, C.c_void_p])
    _bind(L, "mrh_exchange_meta_put", "argtypes", [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64])
    _bind(L, "mrh_get_exchange_table", "argtypes", [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p])
    _bind(L, "mrh_get_exchange_table", "restype", C.c_int64)
    _bind(L, "mrh_assemble_projection", "argtypes", [C.c_void_p, C.c_int32, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p])
    if hasattr(L, "mrh_nonlinear_solve"):            # device-only entry points (absent from the CPU stepping harness)
        _bind(L, "mrh_default_solver_opts", "argtypes", [C.POINTER(MrhSolverOpts)])
        _bind(L, "mrh_default_solver_opts", "restype", None)
        _bind(L, "mrh_linear_solve", "argtypes", [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(MrhSolverOpts), C.POINTER(MrhSolverInfo), C.c_void_p])
        _bind(L, "mrh_nonlinear_solve", "argtypes", [C.c_void_p, C.POINTER(MrhState), C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(MrhSolverOpts), C.POINTER(MrhSolverInfo)])
        _bind(L, "mrh_set_dirichlet", "argtypes", [C.c_void_p, C.c_void_p, C.c_double, C.c_void_p, C.c_void_p, C.POINTER(MrhSolverOpts), C.POINTER(MrhSolverInfo), C.c_void_p])
        _bind(L, "mrh_set_initial", "argtypes", [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(MrhSolverOpts), C.POINTER(MrhSolverInfo), C.c_void_p])
        _bind(L, "mrh_transient_solve", "argtypes", [C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_char_p, C.c_int32, C.c_void_p, C.c_void_p,
                                                     C.POINTER(MrhSolverOpts), C.POINTER(MrhSolverInfo), C.c_void_p])
    _LIBS[path] = L
    return L


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    if hasattr(a, "data_ptr"):      # torch tensor
        return a.data_ptr()
    return int(a)


BC_CODES = {"dirichlet": 1, "weak_dirichlet": 2, "neumann": 3, "flux": 4}
SIZE_NAMES = ["dim", "nelem", "nnode", "nvert", "nvar", "nb", "ndof_elem", "nlocal", "nnz", "nq", "nqs", "nsides",
              "npatterns", "nfixed"]


class Engine:
    """One assembly engine handle (== one AssemblyManager on one GPU)."""

    def __init__(self, physics, dim, order=1, lib=None, device=0, quadrature=-1, side_quadrature=-1,
                 use_strong_dbc=True, fix_zero_rows=False, lump_mass=False, settings=None, force_generic=False, no_jit=False, no_tiles=False, no_ws=False, no_batched=False, jit_batched=False, tiled2d=False):
        self.L = load_library(lib)
        settings = settings or {}
        cfg = MrhConfig()
        self.L.mrh_default_config(C.byref(cfg))
        self._physics = physics.encode()
        cfg.dim, cfg.physics, cfg.order = dim, self._physics, order
        cfg.quadrature, cfg.side_quadrature = quadrature, side_quadrature
        cfg.use_strong_dbc, cfg.fix_zero_rows, cfg.lump_mass = int(use_strong_dbc), int(fix_zero_rows), int(lump_mass)
        cfg.include_advection = int(bool(settings.get("include advection", False)))
        cfg.use_supg = int(bool(settings.get("useSUPG", False) or settings.get("use SUPG", False)))     # "use SUPG": the Burgers spelling
        cfg.use_pspg = int(bool(settings.get("usePSPG", False)))
        cfg.use_lsic = int(bool(settings.get("useLSIC", False)))
        cfg.use_evisc = int(bool(settings.get("use entropy viscosity", False)))
        cfg.form_param = float(settings.get("form_param", 1.0))
        cfg.le_penalty = float(settings.get("penalty", 10.0))
        cfg.device = device
        cfg.flags = (1 if force_generic else 0) | (2 if no_jit else 0) | (4 if no_tiles else 0) | (8 if no_ws else 0) | (16 if no_batched else 0) | (32 if jit_batched else 0) | (256 if tiled2d else 0)
        self.h = C.c_void_p()
        rc = self.L.mrh_create(C.byref(cfg), C.byref(self.h))
        if rc != 0:
            raise MrhError(f"mrh_create failed ({rc}): {self.L.mrh_last_error(None).decode()}")
        self._keep = []

    # -------------------------------------------------------------------------------------------------
    def _ck(self, rc, what):
        if rc < 0:
            raise MrhError(f"{what} failed ({rc}): {self.L.mrh_last_error(self.h).decode()}")
        return rc

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            self.L.mrh_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @classmethod
    def from_problem(cls, p, lib=None, device=0, finalize=True, host_only=False, force_generic=False, no_jit=False, no_tiles=False, no_ws=False,
                     no_batched=False, jit_batched=False, part=None, comm_id=None, tiled2d=False):
        """Build an engine from a tests/problems.py style dict (same dict drives the oracle)."""
        extra = p.get("extra", {})
        weak = bool(extra.get("use weak Dirichlet", False))
        m = p["mesh"]
        e = cls(p["physics"], m["dim"], p.get("order", 1), lib=lib, device=device,
                quadrature=extra.get("quadrature", -1), side_quadrature=extra.get("side quadrature", -1),
                use_strong_dbc=extra.get("use strong DBCs", True), fix_zero_rows=extra.get("fix zero rows", False),
                lump_mass=extra.get("lump mass", False), settings=p.get("settings", {}),
                force_generic=force_generic, no_jit=no_jit, no_tiles=no_tiles, no_ws=no_ws, no_batched=no_batched, jit_batched=jit_batched, tiled2d=tiled2d)
        n = [m["NX"], m["NY"], m.get("NZ", 1)]
        lo = [m.get("xmin", 0.0), m.get("ymin", 0.0), m.get("zmin", 0.0)]
        hi = [m.get("xmax", 1.0), m.get("ymax", 1.0), m.get("zmax", 1.0)]
        if part is None:
            e.build_brick(n, lo, hi)
        else:                                    # (procs, rank): this rank's block of the global brick
            e.build_brick_part(n, lo, hi, part[0], part[1])
        for name, vals in p.get("parameters", {}).items():
            e.set_parameter(name, vals)
        for name, expr in p.get("functions", {}).items():       # physicsInterface.cpp:466-483: both locations
            e.add_function(name, expr, "ip")
            e.add_function(name, expr, "side ip")
        for var, expr in p.get("initial", {}).items():              # "Initial conditions" of the input: function "initial <var>"
            e.add_function("initial " + var, expr, "ip")
        for var, sides in p.get("dirichlet", {}).items():
            for side, expr in sides.items():
                e.set_bc(var, side, "weak_dirichlet" if weak else "dirichlet", expr)
        for kind in ("neumann", "flux"):
            for var, sides in p.get(kind, {}).items():
                for side, expr in sides.items():
                    e.set_bc(var, side, kind, expr)
        if comm_id is not None:                  # collective: every rank of the processor grid
            e.comm_init(comm_id, part[1], int(np.prod(part[0])))
        if host_only:
            e.finalize_host()
        elif finalize:
            e.finalize()
        return e

    # ---- set-up --------------------------------------------------------------------------------------
    def build_brick(self, nxyz, lo=(0, 0, 0), hi=(1, 1, 1)):
        n = np.asarray(list(nxyz) + [1] * (3 - len(nxyz)), dtype=np.int32)
        lo = np.asarray(list(lo) + [0.0] * (3 - len(lo)), dtype=np.float64)
        hi = np.asarray(list(hi) + [1.0] * (3 - len(hi)), dtype=np.float64)
        self._ck(self.L.mrh_build_brick(self.h, _ptr(n), _ptr(lo), _ptr(hi)), "mrh_build_brick")

    def build_brick_part(self, nxyz, lo, hi, procs, rank):
        n = np.asarray(list(nxyz) + [1] * (3 - len(nxyz)), dtype=np.int32)
        lo = np.asarray(list(lo) + [0.0] * (3 - len(lo)), dtype=np.float64)
        hi = np.asarray(list(hi) + [1.0] * (3 - len(hi)), dtype=np.float64)
        pr = np.asarray(list(procs) + [1] * (3 - len(procs)), dtype=np.int32)
        self._ck(self.L.mrh_build_brick_part(self.h, _ptr(n), _ptr(lo), _ptr(hi), _ptr(pr), rank), "mrh_build_brick_part")

    def set_partition(self, gids, owner, rank, world):
        gids = np.ascontiguousarray(gids, dtype=np.int64)
        owner = np.ascontiguousarray(owner, dtype=np.int32)
        self._ck(self.L.mrh_set_partition(self.h, _ptr(gids), _ptr(owner), rank, world), "mrh_set_partition")

    def partition(self):
        n = self.sizes()["nlocal"]
        gids, owner = np.zeros(n, dtype=np.int64), np.zeros(n, dtype=np.int32)
        self._ck(self.L.mrh_get_partition(self.h, _ptr(gids), _ptr(owner)), "mrh_get_partition")
        return gids, owner

    def comm_unique_id(self):
        buf = np.zeros(128, dtype=np.uint8)
        self._ck(self.L.mrh_comm_unique_id(_ptr(buf)), "mrh_comm_unique_id")
        return buf

    def comm_init(self, id128, rank, world):
        id128 = np.ascontiguousarray(id128, dtype=np.uint8)
        self._ck(self.L.mrh_comm_init(self.h, _ptr(id128), rank, world), "mrh_comm_init")

    def exchange(self, res, jac=None, stream=None):
        self._ck(self.L.mrh_exchange(self.h, _ptr(res), _ptr(jac), stream), "mrh_exchange")

    def exchange_meta(self, peer):
        n = int(self.L.mrh_exchange_meta_size(self.h, peer))
        out = np.zeros(n, dtype=np.int64)
        if n:
            self._ck(self.L.mrh_exchange_meta_get(self.h, peer, _ptr(out)), "mrh_exchange_meta_get")
        return out

    def exchange_meta_put(self, peer, meta):
        meta = np.ascontiguousarray(meta, dtype=np.int64)
        self._ck(self.L.mrh_exchange_meta_put(self.h, peer, _ptr(meta), len(meta)), "mrh_exchange_meta_put")

    def exchange_table(self, peer, which):
        n = int(self.L.mrh_get_exchange_table(self.h, peer, which, None))
        out = np.zeros(max(n, 0), dtype=np.int64)
        if n > 0:
            self.L.mrh_get_exchange_table(self.h, peer, which, _ptr(out))
        return out

    def ghost_block(self, values=False):
        n = int(self.L.mrh_get_ghost_block(self.h, None, None, None))
        rows, cols = np.zeros(n, dtype=np.int32), np.zeros(n, dtype=np.int64)
        vals = np.zeros(n) if values else None
        self._ck(self.L.mrh_get_ghost_block(self.h, _ptr(rows), _ptr(cols), _ptr(vals)), "mrh_get_ghost_block")
        return rows, cols, vals

    def set_mesh(self, nodes, elem2node):
        nodes = np.ascontiguousarray(nodes, dtype=np.float64)
        e2n = np.ascontiguousarray(elem2node, dtype=np.int32)
        self._ck(self.L.mrh_set_mesh(self.h, _ptr(nodes), nodes.shape[0], _ptr(e2n), e2n.shape[0]), "mrh_set_mesh")

    def set_nodes(self, nodes):
        nodes = np.ascontiguousarray(nodes, dtype=np.float64)
        self._ck(self.L.mrh_set_nodes(self.h, _ptr(nodes), nodes.shape[0]), "mrh_set_nodes")

    def set_dofmap(self, lids, offsets, nlocal):
        lids = np.ascontiguousarray(lids, dtype=np.int32)
        offsets = np.ascontiguousarray(offsets, dtype=np.int32)
        self._ck(self.L.mrh_set_dofmap(self.h, _ptr(lids), _ptr(offsets), nlocal), "mrh_set_dofmap")

    def set_crs(self, rowptr, colind):
        rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
        colind = np.ascontiguousarray(colind, dtype=np.int32)
        self._ck(self.L.mrh_set_crs(self.h, _ptr(rowptr), _ptr(colind), len(rowptr) - 1), "mrh_set_crs")

    def add_side(self, name, local_side, elems):
        elems = np.ascontiguousarray(elems, dtype=np.int32)
        self._ck(self.L.mrh_add_side(self.h, name.encode(), local_side, _ptr(elems), len(elems)), "mrh_add_side")

    def set_bc(self, var, side, kind, expr):
        self._ck(self.L.mrh_set_bc(self.h, var.encode(), side.encode(), BC_CODES[kind], str(expr).encode()),
                 "mrh_set_bc")

    def add_function(self, name, expr, location="ip"):
        self._ck(self.L.mrh_add_function(self.h, name.encode(), str(expr).encode(), location.encode()),
                 "mrh_add_function")

    def set_parameter(self, name, values):
        """scalar / vector parameter usable as a leaf `name` / `name(k)` of coefficient expressions (constant of the pass)"""
        v = np.ascontiguousarray(np.atleast_1d(values), dtype=np.float64)
        self._ck(self.L.mrh_set_parameter(self.h, name.encode(), v.ctypes.data_as(C.POINTER(C.c_double)), len(v)), "mrh_set_parameter")

    def finalize(self):
        self._ck(self.L.mrh_finalize(self.h), "mrh_finalize")

    def finalize_host(self):
        self._ck(self.L.mrh_finalize_host(self.h), "mrh_finalize_host")

    # ---- introspection -------------------------------------------------------------------------------
    def sizes(self):
        s = np.zeros(16, dtype=np.int64)
        self._ck(self.L.mrh_get_sizes(self.h, _ptr(s)), "mrh_get_sizes")
        return dict(zip(SIZE_NAMES, (int(v) for v in s)))

    def mesh(self):
        s = self.sizes()
        nodes = np.zeros((s["nnode"], s["dim"]))
        e2n = np.zeros((s["nelem"], s["nvert"]), dtype=np.int32)
        self._ck(self.L.mrh_get_mesh(self.h, _ptr(nodes), _ptr(e2n)), "mrh_get_mesh")
        return nodes, e2n

    def dofmap(self, fixed=True, coords=True):
        s = self.sizes()
        lids = np.zeros((s["nelem"], s["ndof_elem"]), dtype=np.int32)
        offs = np.zeros((s["nvar"], s["nb"]), dtype=np.int32)
        fx = np.zeros(s["nlocal"], dtype=np.uint8) if fixed else None
        xyz = np.zeros((s["nlocal"], s["dim"])) if coords else None
        var = np.zeros(s["nlocal"], dtype=np.int32) if coords else None
        self._ck(self.L.mrh_get_dofmap(self.h, _ptr(lids), _ptr(offs), _ptr(fx), _ptr(xyz), _ptr(var)),
                 "mrh_get_dofmap")
        return lids, offs, fx, xyz, var

    def dof_coords(self):
        s = self.sizes()
        xyz = np.zeros((s["nlocal"], s["dim"]))
        var = np.zeros(s["nlocal"], dtype=np.int32)
        self._ck(self.L.mrh_get_dofmap(self.h, None, None, None, _ptr(xyz), _ptr(var)), "mrh_get_dofmap")
        return xyz, var

    def crs(self):
        s = self.sizes()
        rowptr = np.zeros(s["nlocal"] + 1, dtype=np.int64)
        colind = np.zeros(s["nnz"], dtype=np.int32)
        self._ck(self.L.mrh_get_crs(self.h, _ptr(rowptr), _ptr(colind)), "mrh_get_crs")
        return rowptr, colind

    def sides(self):
        out = []
        for s in range(self.sizes()["nsides"]):
            buf = C.create_string_buffer(64)
            ls = C.c_int32(0)
            n = self._ck(self.L.mrh_get_side(self.h, s, buf, 64, C.byref(ls), None), "mrh_get_side")
            el = np.zeros(n, dtype=np.int32)
            self.L.mrh_get_side(self.h, s, None, 0, None, _ptr(el))
            out.append((buf.value.decode(), ls.value, el))
        return out

    def bc_codes(self):
        s = self.sizes()
        return np.array([[self.L.mrh_get_bc(self.h, v, k) for k in range(s["nsides"])] for v in range(s["nvar"])],
                        dtype=np.int32)

    def varnames(self):
        return [self.L.mrh_get_varname(self.h, v).decode() for v in range(self.sizes()["nvar"])]

    def eval_function(self, name, xyz, location="ip", time=0.0):
        xyz = np.ascontiguousarray(xyz, dtype=np.float64)
        out = np.zeros(xyz.shape[0])
        self._ck(self.L.mrh_eval_function_host(self.h, name.encode(), location.encode(), xyz.shape[0], _ptr(xyz),
                                               time, _ptr(out)), "mrh_eval_function_host")
        return out

    # ---- hot path ------------------------------------------------------------------------------------
    def _state(self, u, u_prev=None, u_stage=None, time_int=None, stream=None):
        st = MrhState()
        st.u, st.u_prev, st.u_stage = _ptr(u), _ptr(u_prev), _ptr(u_stage)
        st.time, st.deltat = 0.0, 1.0
        if time_int is not None:
            A = np.ascontiguousarray(time_int["A"], dtype=np.float64)
            b = np.ascontiguousarray(time_int["b"], dtype=np.float64)
            bdf = np.ascontiguousarray(time_int["bdf"], dtype=np.float64)
            self._keep = [A, b, bdf]
            st.is_transient = 1
            st.nstages, st.nsteps, st.stage = len(b), len(bdf) - 1, int(time_int.get("stage", 0))
            st.butcher_A, st.butcher_b, st.bdf_wts = _ptr(A), _ptr(b), _ptr(bdf)
            st.time, st.deltat = float(time_int.get("time", 0.0)), float(time_int["dt"])
        st.stream = stream
        return st

    def assemble_res(self, u, res, **kw):
        st = self._state(u, **kw)
        self._ck(self.L.mrh_assemble_res(self.h, C.byref(st), _ptr(res)), "mrh_assemble_res")

    def assemble_jacres(self, u, res, jac, **kw):
        st = self._state(u, **kw)
        self._ck(self.L.mrh_assemble_jacres(self.h, C.byref(st), _ptr(res), _ptr(jac)), "mrh_assemble_jacres")

    def assemble_jacres_host(self, u, res, jac, **kw):
        st = self._state(u, **kw)
        self._ck(self.L.mrh_assemble_jacres_host(self.h, C.byref(st), _ptr(res), _ptr(jac)),
                 "mrh_assemble_jacres_host")

    def assemble_res_host_static(self, u, res, **kw):
        """allocation-free host-buffer residual assembly (static device scratch; at most MRH_STATIC_SCRATCH_MAX_DOFS DOFs)"""
        st = self._state(u, **kw)
        self._ck(self.L.mrh_assemble_res_host_static(self.h, C.byref(st), _ptr(res)), "mrh_assemble_res_host_static")

    # ---- the caller of the hot path (Newton / time stepping / Krylov on the device) ------------------------------
    def solver_opts(self, **kw):
        o = MrhSolverOpts()
        self.L.mrh_default_solver_opts(C.byref(o))
        for k, v in kw.items():
            setattr(o, k, v)
        return o

    @staticmethod
    def _info(i):
        return {f[0]: getattr(i, f[0]) for f in MrhSolverInfo._fields_}

    def linear_solve(self, jac, rhs, x, stream=None, **opts):
        o, i = self.solver_opts(**opts), MrhSolverInfo()
        self._ck(self.L.mrh_linear_solve(self.h, _ptr(jac), _ptr(rhs), _ptr(x), C.byref(o), C.byref(i), stream), "mrh_linear_solve")
        return self._info(i)

    def nonlinear_solve(self, u, res, jac, u_prev=None, u_stage=None, time_int=None, stream=None, **opts):
        st = self._state(u, u_prev=u_prev, u_stage=u_stage, time_int=time_int, stream=stream)
        o, i = self.solver_opts(**opts), MrhSolverInfo()
        rc = self._ck(self.L.mrh_nonlinear_solve(self.h, C.byref(st), _ptr(u), _ptr(res), _ptr(jac), C.byref(o), C.byref(i)), "mrh_nonlinear_solve")
        out = self._info(i)
        out["status"] = rc
        return out

    def assemble_projection(self, kind, rhs, mass, time=0.0, stream=None):
        """kind 0: strong Dirichlet data (setDirichlet), 1: initial data (setInitial); rhs / mass are accumulated into"""
        self._ck(self.L.mrh_assemble_projection(self.h, kind, time, _ptr(rhs), _ptr(mass), stream), "mrh_assemble_projection")

    def set_dirichlet(self, u, rhs_work, mass_work, time=0.0, stream=None, **opts):
        o, i = self.solver_opts(linear_tol=1e-12, max_linear_iters=1000, **opts), MrhSolverInfo()
        rc = self._ck(self.L.mrh_set_dirichlet(self.h, _ptr(u), time, _ptr(rhs_work), _ptr(mass_work), C.byref(o), C.byref(i), stream), "mrh_set_dirichlet")
        out = self._info(i); out["status"] = rc
        return out

    def set_initial(self, u, rhs_work, mass_work, stream=None, **opts):
        o, i = self.solver_opts(linear_tol=1e-12, max_linear_iters=1000, **opts), MrhSolverInfo()
        rc = self._ck(self.L.mrh_set_initial(self.h, _ptr(u), _ptr(rhs_work), _ptr(mass_work), C.byref(o), C.byref(i), stream), "mrh_set_initial")
        out = self._info(i); out["status"] = rc
        return out

    def transient_solve(self, u, res, jac, t0, t1, dt, tableau="BWE", bdf_order=1, stream=None, **opts):
        o, i = self.solver_opts(**opts), MrhSolverInfo()
        rc = self._ck(self.L.mrh_transient_solve(self.h, _ptr(u), t0, t1, dt, tableau.encode(), bdf_order, _ptr(res), _ptr(jac), C.byref(o),
                                                 C.byref(i), stream), "mrh_transient_solve")
        out = self._info(i)
        out["status"] = rc
        return out

    def kernel_path(self):
        return self.L.mrh_last_kernel_path(self.h).decode()

    def launch_count(self):
        return int(self.L.mrh_last_launch_count(self.h))

    def volume_kernel_ms(self):
        return float(self.L.mrh_last_volume_kernel_ms(self.h))
