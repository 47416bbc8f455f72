"""Multi-GPU host logic on CPU (gloo, several processes): the library's element-block partition (mrh_build_brick_part), its
exchange plan (mrh_exchange_meta_* / mrh_get_exchange_table) and the additive export, emulated with numpy over gloo on the
arrays the CPU stepping harness assembles, against a single-domain assembly of the oracle.  The NCCL transport of the same
plan (mrh_comm_init / mrh_exchange) is covered by tests/test_gpu_multi.py on >= 2 GPUs."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import parity_util as U
import problems


def _problem(kind, n):
    if kind == "thermal":
        p = problems.config_c2(n)
        p["mesh"].update(NX=n, NY=n + 1, NZ=n + 2)
        p["dirichlet"] = {"e": {"left": "0.0", "bottom": "0.1*x", "back": "0.0"}}
        p["neumann"] = {"e": {"front": "1.0+x"}}
        p["functions"] = dict(p["functions"]); p["functions"]["thermal diffusion"] = "1.0+x*y"
    else:
        p = problems.config_c4(n)
        p["mesh"].update(NX=n + 1, NY=n, NZ=n)
    return p


def _worker(rank, world, port, procs, kind, n, q):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        dist.init_process_group("gloo", rank=rank, world_size=world)
        from mrhyde_b200 import Engine
        from oracle.pyoracle import Oracle
        lib = U.EMU_LIB
        p = _problem(kind, n)
        e = Engine.from_problem(p, lib=lib, part=(procs, rank))
        s = e.sizes()
        xyz, var = e.dof_coords()
        gids, owner = e.partition()
        rp, ci = e.crs()
        fixed = e.dofmap(coords=False)[2]
        u = np.prod(np.sin(1.3 * xyz + 0.2), axis=1) + 0.1 * var
        res, jac = np.zeros(s["nlocal"]), np.zeros(s["nnz"])
        e.assemble_jacres(u, res, jac)                       # this rank's elements only: the OVERLAPPED vector / matrix
        # ---- plan metadata over gloo ----
        sizes = torch.tensor([int(e.L.mrh_exchange_meta_size(e.h, pr)) if pr != rank else 0 for pr in range(world)], dtype=torch.int64)
        allsz = [torch.zeros(world, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(allsz, sizes)
        ops, inbox, keep = [], {}, []
        for pr in range(world):
            if pr == rank:
                continue
            if sizes[pr] > 0:
                t = torch.from_numpy(e.exchange_meta(pr)); keep.append(t)
                ops.append(dist.P2POp(dist.isend, t, pr))
            if allsz[pr][rank] > 0:
                inbox[pr] = torch.zeros(int(allsz[pr][rank]), dtype=torch.int64)
                ops.append(dist.P2POp(dist.irecv, inbox[pr], pr))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        for pr, t in inbox.items():
            e.exchange_meta_put(pr, t.numpy())
        # ---- the additive export, emulated: pack -> send / receive -> unpack ----
        grow_, gcol_, _ = e.ghost_block()
        ghost = np.zeros(len(grow_))
        ops, inbox, keep = [], {}, []
        for pr in range(world):
            if pr == rank:
                continue
            srows, sslots = e.exchange_table(pr, 0), e.exchange_table(pr, 1)
            rrows, rslots = e.exchange_table(pr, 2), e.exchange_table(pr, 3)
            if len(srows):
                assert np.all(owner[srows] == pr)
                t = torch.from_numpy(np.concatenate([res[srows], jac[sslots]])); keep.append(t)
                ops.append(dist.P2POp(dist.isend, t, pr))
            if len(rrows):
                assert np.all(owner[rrows] == rank)
                inbox[pr] = (torch.zeros(len(rrows) + len(rslots), dtype=torch.float64), rrows, rslots)
                ops.append(dist.P2POp(dist.irecv, inbox[pr][0], pr))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        for pr, (t, rrows, rslots) in inbox.items():
            v = t.numpy()
            np.add.at(res, rrows, v[:len(rrows)])
            jv = v[len(rrows):]
            loc = rslots >= 0
            np.add.at(jac, rslots[loc], jv[loc])
            np.add.at(ghost, -rslots[~loc] - 1, jv[~loc])
        # ---- single-domain reference ----
        eg = Engine.from_problem(p, lib=lib, part=((1, 1, 1), 0), host_only=True)
        ggid, _ = eg.partition()
        og = Oracle(U.oracle_problem(p))
        _, _, gfixed, gxyz, gvar = og.dofmap()
        grp, gci = og.crs()
        gu = np.prod(np.sin(1.3 * gxyz + 0.2), axis=1) + 0.1 * gvar
        gr, gj = og.assemble(gu)
        g2G = {int(g): i for i, g in enumerate(ggid)}
        # holders of every DOF (for the diagonal of shared fixed rows)
        allg = [None] * world
        dist.all_gather_object(allg, gids.tolist())
        holders = {}
        for lst in allg:
            for g in lst:
                holders[g] = holders.get(g, 0) + 1
        sr, sj = np.abs(gr).max(), np.abs(gj).max()
        ghost_of = {}
        for k in range(len(grow_)):
            ghost_of.setdefault(int(grow_[k]), {})[int(gcol_[k])] = ghost[k]
        nowned = 0
        for row in np.nonzero(owner == rank)[0]:
            G = g2G[int(gids[row])]
            nowned += 1
            assert bool(fixed[row]) == bool(gfixed[G])
            assert abs(res[row] - gr[G]) <= 1e-12 * sr, (rank, row, res[row], gr[G])
            want = {int(c): v for c, v in zip(gci[grp[G]:grp[G + 1]], gj[grp[G]:grp[G + 1]])}
            got = {g2G[int(gids[ci[t]])]: jac[t] for t in range(rp[row], rp[row + 1])}
            for cg, v in ghost_of.get(int(row), {}).items():
                assert g2G[cg] not in got
                got[g2G[cg]] = v
            assert set(got) == set(want), (rank, row)
            for c, v in got.items():
                w = want[c]
                if fixed[row] and c == G:
                    w = float(holders[int(gids[row])])     # unit diagonal set on every holder's overlapped matrix, then ADD
                assert abs(v - w) <= 1e-12 * sj, (rank, row, c, v, w)
        tot = torch.tensor([nowned], dtype=torch.int64)
        dist.all_reduce(tot)
        assert int(tot) == og.nlocal                         # every DOF has exactly one owner
        q.put((rank, "ok"))
    except Exception:   # noqa: BLE001
        import traceback
        q.put((rank, "FAIL " + traceback.format_exc()))
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


@pytest.mark.parametrize("procs,kind,n", [((1, 1, 2), "thermal", 4), ((2, 1, 1), "le", 3), ((2, 2, 1), "thermal", 4), ((2, 2, 2), "thermal", 4),
                                           ((1, 3, 1), "le", 3)], ids=["1x1x2", "2x1x1-le", "2x2x1", "2x2x2", "1x3x1-le"])
def test_block_partition_and_exchange_plan(procs, kind, n):
    U.build_emu()
    world = int(np.prod(procs))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29700 + world * 7 + (os.getpid() % 150)
    ps = [ctx.Process(target=_worker, args=(r, world, port, procs, kind, n, q)) for r in range(world)]
    for p in ps:
        p.start()
    out = [q.get(timeout=300) for _ in ps]
    for p in ps:
        p.join(30)
    for rank, msg in out:
        assert msg == "ok", f"rank {rank}: {msg}"
