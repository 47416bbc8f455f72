"""ORACLE (test infrastructure, NOT product code): ctypes wrapper over oracle/liboracle.so.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, "src", f) for f in os.listdir(os.path.join(_HERE, "src"))]
    stale = (not os.path.exists(so)) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(so):
            build()
        L = C.CDLL(so)
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.c_char_p]
        L.orc_last_error.restype = C.c_char_p
        for name in ("orc_destroy", "orc_sizes", "orc_get_mesh", "orc_get_dofmap", "orc_get_crs",
                     "orc_get_group_basis", "orc_get_bgroup", "orc_set_time", "orc_last_timing", "orc_set_nodes"):
            getattr(L, name).restype = None
        _LIB = L
    return _LIB


def _p(a, t=C.c_double):
    return None if a is None else a.ctypes.data_as(C.POINTER(t))


def config_text(**kw):
    """Build the oracle's config text from keyword sections.

    mesh=dict(dim, NX, NY, NZ, xmin..), physics='thermal', order=1, functions={}, settings={},
    dirichlet={var:{side:expr}}, neumann=..., flux=..., extra={'workset size':100,...}
    """
    lines = []
    for k, v in kw.get("mesh", {}).items():
        lines.append(f"{k}: {v}")
    lines.append(f"physics: {kw.get('physics', 'thermal')}")
    lines.append(f"order: {kw.get('order', 1)}")
    for k, v in kw.get("extra", {}).items():
        lines.append(f"{k}: {str(v).lower() if isinstance(v, bool) else v}")
    for k, v in kw.get("functions", {}).items():
        lines.append(f"function.{k}: {v}")
    for k, v in kw.get("parameters", {}).items():
        lines.append(f"parameter.{k}: " + " ".join(repr(float(x)) for x in np.atleast_1d(v)))
    for k, v in kw.get("settings", {}).items():
        lines.append(f"setting.{k}: {str(v).lower() if isinstance(v, bool) else v}")
    for sec in ("dirichlet", "neumann", "flux"):
        for var, sides in kw.get(sec, {}).items():
            for side, expr in sides.items():
                lines.append(f"{sec}.{var}.{side}: {expr}")
    return "\n".join(lines) + "\n"


class Oracle:
    NAMES = ["dim", "nelem", "nnode", "nvert", "nvar", "nb", "ndof_elem", "nlocal", "nnz", "nq", "nqs",
             "nsides", "type_AD", "ngroups", "nbgroups"]

    def __init__(self, text):
        self.L = lib()
        self.text = text
        self.h = self.L.orc_create(text.encode())
        if not self.h:
            raise RuntimeError("oracle: " + self.L.orc_last_error().decode())
        s = np.zeros(15, dtype=np.int64)
        self.L.orc_sizes(C.c_void_p(self.h), _p(s, C.c_int64))
        for n, v in zip(self.NAMES, s):
            setattr(self, n, int(v))

    def __del__(self):
        try:
            if self.h:
                self.L.orc_destroy(C.c_void_p(self.h))
                self.h = None
        except Exception:
            pass

    def mesh(self):
        nodes = np.zeros((self.nnode, self.dim))
        e2n = np.zeros((self.nelem, self.nvert), dtype=np.int32)
        self.L.orc_get_mesh(C.c_void_p(self.h), _p(nodes), _p(e2n, C.c_int))
        return nodes, e2n

    def set_nodes(self, nodes):
        """Replace the vertex coordinates (same topology): distorted, non-affine test meshes."""
        nodes = np.ascontiguousarray(nodes, dtype=np.float64)
        assert nodes.shape == (self.nnode, self.dim)
        self.L.orc_set_nodes(C.c_void_p(self.h), _p(nodes))

    def dofmap(self):
        lids = np.zeros((self.nelem, self.ndof_elem), dtype=np.int32)
        offs = np.zeros((self.nvar, self.nb), dtype=np.int32)
        fixed = np.zeros(self.nlocal, dtype=np.uint8)
        xyz = np.zeros((self.nlocal, self.dim))
        var = np.zeros(self.nlocal, dtype=np.int32)
        self.L.orc_get_dofmap(C.c_void_p(self.h), _p(lids, C.c_int), _p(offs, C.c_int), _p(fixed, C.c_uint8),
                              _p(xyz), _p(var, C.c_int))
        return lids, offs, fixed, xyz, var

    def crs(self):
        rowptr = np.zeros(self.nlocal + 1, dtype=np.int64)
        colind = np.zeros(self.nnz, dtype=np.int32)
        self.L.orc_get_crs(C.c_void_p(self.h), _p(rowptr, C.c_int64), _p(colind, C.c_int))
        return rowptr, colind

    def sides(self):
        out = []
        for s in range(self.nsides):
            buf = C.create_string_buffer(64)
            ls = C.c_int(0)
            n = self.L.orc_get_side(C.c_void_p(self.h), s, buf, 64, C.byref(ls), None)
            el = np.zeros(n, dtype=np.int32)
            self.L.orc_get_side(C.c_void_p(self.h), s, None, 0, None, _p(el, C.c_int))
            out.append((buf.value.decode(), ls.value, el))
        return out

    def bc_codes(self):
        return np.array([[self.L.orc_get_bc(C.c_void_p(self.h), v, s) for s in range(self.nsides)]
                         for v in range(self.nvar)], dtype=np.int32)

    def group_basis(self, g, numElem):
        b = np.zeros((numElem, self.nb, self.nq))
        bg = np.zeros((numElem, self.nb, self.nq, self.dim))
        w = np.zeros((numElem, self.nq))
        ip = np.zeros((self.dim, numElem, self.nq))
        self.L.orc_get_group_basis(C.c_void_p(self.h), g, _p(b), _p(bg), _p(w), _p(ip))
        return b, bg, w, ip

    def bgroup(self, g):
        side = C.c_int(0)
        ne = C.c_int(0)
        self.L.orc_get_bgroup(C.c_void_p(self.h), g, C.byref(side), C.byref(ne), None, None, None, None)
        n = ne.value
        w = np.zeros((n, self.nqs))
        nrm = np.zeros((self.dim, n, self.nqs))
        b = np.zeros((n, self.nb, self.nqs))
        bg = np.zeros((n, self.nb, self.nqs, self.dim))
        self.L.orc_get_bgroup(C.c_void_p(self.h), g, None, None, _p(w), _p(nrm), _p(b), _p(bg))
        return side.value, w, nrm, b, bg

    def set_time(self, A, b, bdf, dt, time, stage=0):
        A = np.ascontiguousarray(A, dtype=np.float64)
        b = np.ascontiguousarray(b, dtype=np.float64)
        bdf = np.ascontiguousarray(bdf, dtype=np.float64)
        self.L.orc_set_time(C.c_void_p(self.h), len(b), _p(A), _p(b), len(bdf), _p(bdf), C.c_double(dt),
                            C.c_double(time), stage)

    def assemble(self, u, jacobian=True, u_prev=None, u_stage=None, nthreads=1):
        u = np.ascontiguousarray(u, dtype=np.float64)
        res = np.zeros(self.nlocal)
        jac = np.zeros(self.nnz) if jacobian else None
        rc = self.L.orc_assemble(C.c_void_p(self.h), _p(u), _p(u_prev), _p(u_stage), int(jacobian), _p(res),
                                 _p(jac), nthreads)
        if rc != 0:
            raise RuntimeError("oracle: " + self.L.orc_last_error().decode())
        return res, jac

    def assemble_dirichlet(self, time=0.0):
        """setDirichlet: (rhs, mass values) of the L2 projection of the strong Dirichlet data onto the fixed DOFs"""
        rhs, mass = np.zeros(self.nlocal), np.zeros(self.nnz)
        rc = self.L.orc_assemble_dirichlet(C.c_void_p(self.h), C.c_double(time), _p(rhs), _p(mass))
        if rc != 0:
            raise RuntimeError("oracle: " + self.L.orc_last_error().decode())
        return rhs, mass

    def timing(self):
        t = np.zeros(4)
        self.L.orc_last_timing(C.c_void_p(self.h), _p(t))
        return dict(gather=t[0], physics=t[1], scatter=t[2], total=t[3])

    def eval_expression(self, expr, g, numElem, time=0.0):
        out = np.zeros((numElem, self.nq))
        rc = self.L.orc_eval_expression(C.c_void_p(self.h), expr.encode(), g, C.c_double(time), _p(out))
        if rc != 0:
            raise RuntimeError("oracle: " + self.L.orc_last_error().decode())
        return out

    def l2_error(self, u, var, true_expr, time=0.0):
        u = np.ascontiguousarray(u, dtype=np.float64)
        out = C.c_double(0.0)
        rc = self.L.orc_l2_error(C.c_void_p(self.h), _p(u), var, true_expr.encode(), C.c_double(time), C.byref(out))
        if rc != 0:
            raise RuntimeError("oracle: " + self.L.orc_last_error().decode())
        return out.value
