"""Multi-GPU plumbing: z-slab partition of the brick and the additive exchange of interface rows.

Mirrors what the reference does with Tpetra after each rank assembled its own elements into the OVERLAPPED
vector/matrix (/root/reference/src/interfaces/linearAlgebraInterface.hpp:296-337, doExport(ADD)): rows of DOFs on
a partition interface are summed onto their owner.  Here: one process per GPU, rank r owns slab r of the brick;
a DOF on the plane shared by slabs r and r+1 is owned by the LOWER rank.  Every step rank r+1 sends,
in ONE message per neighbour, [residual of its bottom plane | Jacobian entries of those rows whose column is also
on the plane | the remaining entries (columns in its own slab)] to rank r; the owner adds the first two into its
overlapped arrays and keeps the third as the ghost-column block of its owned rows.
torch.distributed (NCCL on GPUs, gloo in the CPU tests) carries the message; nothing else is collective.
"""
import copy

import numpy as np

SIDES3 = ("back", "bottom", "front", "left", "right", "top")


def slab_problem(p, rank, world):
    """Problem dict of slab `rank`: same NX,NY,NZ per rank (weak scaling), z in [zmin + rank*Lz, zmin + (rank+1)*Lz];
    "all boundaries" conditions are expanded to the sides that are still physical boundaries of this slab."""
    q = copy.deepcopy(p)
    if world == 1:
        return q
    m = q["mesh"]
    assert m["dim"] == 3, "slab partition is along z"
    z0, L = m.get("zmin", 0.0), m.get("zmax", 1.0) - m.get("zmin", 0.0)
    m["zmin"], m["zmax"] = z0 + rank * L, z0 + (rank + 1) * L
    phys = [s for s in SIDES3 if not ((s == "back" and rank > 0) or (s == "front" and rank < world - 1))]   # z planes: back (z-) / front (z+)
    for kind in ("dirichlet", "neumann", "flux"):
        for var, sides in list(q.get(kind, {}).items()):
            new = {}
            for side, expr in sides.items():
                if side == "all boundaries":
                    for s in phys:
                        new.setdefault(s, expr)
                elif side in phys:
                    new[side] = expr
            q[kind][var] = new
    return q


class InterfaceExchange:
    """Index tables + per-step exchange for one rank.  Works on any torch device (cuda+nccl, cpu+gloo)."""

    launches_per_step = 0       # pack/unpack are torch index ops, not kernels of this library

    def __init__(self, engine, rank, world, device, xyz=None, var=None, crs=None, zlo=None, zhi=None):
        import torch
        self.torch, self.rank, self.world, self.device = torch, rank, world, device
        if xyz is None:
            xyz, var = engine.dof_coords()
        rp, ci = crs if crs is not None else engine.crs()
        z = xyz[:, 2]
        zlo = z.min() if zlo is None else zlo
        zhi = z.max() if zhi is None else zhi
        self.send = self._tables(xyz, var, rp, ci, z == zlo, sender=True) if rank > 0 else None
        self.recv = self._tables(xyz, var, rp, ci, z == zhi, sender=False) if rank < world - 1 else None
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.int64)).to(device)
        if self.send:
            self.s_rows, self.s_in, self.s_out = t(self.send["rows"]), t(self.send["in_slots"]), t(self.send["out_slots"])
            n = len(self.send["rows"]) + len(self.send["in_slots"]) + len(self.send["out_slots"])
            self.sbuf = torch.empty(n, dtype=torch.float64, device=device)
        if self.recv:
            self.r_rows, self.r_in = t(self.recv["rows"]), t(self.recv["in_slots"])
            self.n_rows, self.n_in = len(self.recv["rows"]), len(self.recv["in_slots"])
            self.rbuf = None     # sized on first exchange from the neighbour's announced length
            self.ghost_vals = None
        self._sized = False

    @staticmethod
    def _tables(xyz, var, rp, ci, mask, sender):
        """Plane rows ordered by (y, x, var); their CRS slots split into in-plane columns (ordered by the column's
        plane rank) and the rest (ordered by the column's (z, y, x, var))."""
        dofs = np.nonzero(mask)[0]
        order = np.lexsort((var[dofs], xyz[dofs, 0], xyz[dofs, 1]))
        rows = dofs[order]
        prank = -np.ones(len(mask), dtype=np.int64)
        prank[rows] = np.arange(len(rows))
        lens = (rp[rows + 1] - rp[rows]).astype(np.int64)
        rowrank = np.repeat(np.arange(len(rows)), lens)
        slots = np.concatenate([np.arange(rp[r], rp[r + 1]) for r in rows]) if len(rows) else np.zeros(0, dtype=np.int64)
        cols = ci[slots]
        cr = prank[cols]
        inp = cr >= 0
        o_in = np.lexsort((cr[inp], rowrank[inp]))
        out = {"rows": rows, "in_slots": slots[inp][o_in]}
        if sender:
            c = cols[~inp]
            o_out = np.lexsort((var[c], xyz[c, 0], xyz[c, 1], xyz[c, 2], rowrank[~inp]))
            out["out_slots"] = slots[~inp][o_out]
            out["out_rowrank"] = rowrank[~inp][o_out]
            out["out_cols"] = c[o_out]
        return out

    def exchange(self, res, jac):
        """res/jac: this rank's overlapped residual and CRS values (torch tensors on self.device)."""
        import torch.distributed as dist
        torch = self.torch
        ops = []
        if self.send:
            nr, ni = len(self.s_rows), len(self.s_in)
            torch.index_select(res, 0, self.s_rows, out=self.sbuf[:nr])
            torch.index_select(jac, 0, self.s_in, out=self.sbuf[nr:nr + ni])
            torch.index_select(jac, 0, self.s_out, out=self.sbuf[nr + ni:])
            ops.append(dist.P2POp(dist.isend, self.sbuf, self.rank - 1))
        # one-time size handshake (blocking, outside the steady state)
        if not self._sized:
            if self.send:
                n = torch.tensor([self.sbuf.numel()], dtype=torch.int64, device=self.device)
                dist.send(n, self.rank - 1)
            if self.recv:
                n = torch.zeros(1, dtype=torch.int64, device=self.device)
                dist.recv(n, self.rank + 1)
                self.rbuf = torch.empty(int(n.item()), dtype=torch.float64, device=self.device)
            self._sized = True
        if self.recv:
            ops.append(dist.P2POp(dist.irecv, self.rbuf, self.rank + 1))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        if self.recv:
            nr, ni = self.n_rows, self.n_in
            res.index_add_(0, self.r_rows, self.rbuf[:nr])
            jac.index_add_(0, self.r_in, self.rbuf[nr:nr + ni])
            self.ghost_vals = self.rbuf[nr + ni:]
