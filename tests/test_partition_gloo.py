"""N>1 host logic on CPU: world_size-2 gloo run of the slab partition + interface-row exchange.
Each rank's overlapped contributions come from the oracle on its slab (no GPU here); after the exchange every
OWNED row (residual, in-graph Jacobian entries and the ghost-column block) must equal the single-domain oracle."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import parity_util as U
import problems


def _worker(rank, world, port, n, q):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        dist.init_process_group("gloo", rank=rank, world_size=world)
        from mrhyde_b200 import Engine, build, partition
        from oracle.pyoracle import Oracle
        base = problems.config_c2(n)
        base["dirichlet"] = {"e": {"left": "0.0", "bottom": "0.0", "back": "0.0"}}      # x-, y-, z- (back / front are the z planes)
        base["neumann"] = {"e": {"front": "1.0+x"}}
        p = partition.slab_problem(base, rank, world)
        e = Engine.from_problem(p, lib=build.build(), host_only=True)
        o = Oracle(U.oracle_problem(p))
        xyz, var = e.dof_coords()
        rp, ci = e.crs()
        fixed = e.dofmap(coords=False)[2]
        u = np.prod(np.sin(1.3 * xyz + 0.2), axis=1)          # a function of position: consistent across ranks
        r, j = o.assemble(u)
        res, jac = torch.from_numpy(r.copy()), torch.from_numpy(j.copy())
        x = partition.InterfaceExchange(e, rank, world, torch.device("cpu"))
        x.exchange(res, jac)
        x.exchange(torch.from_numpy(r.copy()), torch.from_numpy(j.copy()))     # steady-state path (sizes known)
        # single-domain reference
        g = problems.config_c2(n)
        g["mesh"]["NZ"] = n * world; g["mesh"]["zmax"] = float(world)
        g["dirichlet"], g["neumann"] = base["dirichlet"], base["neumann"]
        og = Oracle(U.oracle_problem(g))
        _, _, _, gxyz, _ = og.dofmap()
        grp, gci = og.crs()
        gu = np.prod(np.sin(1.3 * gxyz + 0.2), axis=1)
        gr, gj = og.assemble(gu)
        key = lambda a: tuple(np.round(a * 4096).astype(np.int64))
        gid = {key(c): i for i, c in enumerate(gxyz)}
        owned = np.ones(len(xyz), dtype=bool)
        if rank > 0:
            owned[xyz[:, 2] == xyz[:, 2].min()] = False
        l2g = np.array([gid[key(c)] for c in xyz])
        res, jac = res.numpy(), jac.numpy()
        scale_r, scale_j = np.abs(gr).max(), np.abs(gj).max()
        ghost = {}
        if x.recv:
            gv = x.ghost_vals.numpy()
            ghost_n = len(gv)
        for row in np.nonzero(owned)[0]:
            G = l2g[row]
            assert abs(res[row] - gr[G]) <= 1e-12 * scale_r
            grow = dict(zip(gci[grp[G]:grp[G + 1]], gj[grp[G]:grp[G + 1]]))
            seen = 0.0
            shared = bool(x.recv) and xyz[row, 2] == xyz[:, 2].max()
            for s in range(rp[row], rp[row + 1]):
                want = grow[l2g[ci[s]]]
                # like the reference, updateJacDBC sets diag=1 on every rank's OVERLAPPED matrix before the ADD
                # export (assemblyManager.cpp:1167-1180, :3158), so a fixed DOF shared by two slabs ends with 2
                if shared and fixed[row] and ci[s] == row:
                    want = 2.0
                assert abs(jac[s] - want) <= 1e-12 * scale_j, (rank, row, s, jac[s], want)
                seen += 1
            if not (x.recv and xyz[row, 2] == xyz[:, 2].max()):
                assert seen == len(grow)
        # ghost-column block: the owner's interface rows get their remaining columns from the neighbour
        if x.recv:
            tot = sum(len(gci[grp[l2g[r_]]:grp[l2g[r_] + 1]]) - (rp[r_ + 1] - rp[r_]) for r_ in x.recv["rows"])
            assert tot == ghost_n
            gsum = sum(gj[grp[l2g[r_]]:grp[l2g[r_] + 1]].sum() + (1.0 if fixed[r_] else 0.0) for r_ in x.recv["rows"])
            lsum = sum(jac[rp[r_]:rp[r_ + 1]].sum() for r_ in x.recv["rows"]) + gv.sum()
            assert abs(gsum - lsum) <= 1e-10 * scale_j * len(x.recv["rows"])
        q.put((rank, "ok"))
    except Exception as ex:   # noqa: BLE001
        import traceback
        q.put((rank, "FAIL " + traceback.format_exc()))
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_slab_exchange_matches_single_domain(world):
    from mrhyde_b200 import build
    build.build()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + world + (os.getpid() % 200)
    procs = [ctx.Process(target=_worker, args=(r, world, port, 4, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(30)
    for rank, msg in out:
        assert msg == "ok", f"rank {rank}: {msg}"
