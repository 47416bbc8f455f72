"""Engine + NCCL exchange on >= 2 GPUs against the single-GPU engine (run with gpurun --gpus 2; skipped on one GPU):
every rank assembles its block with mrh_assemble_jacres (interface tiles first, exchange overlapped), owned rows must equal the
single-domain rows - bit-exact in structure (global ids), <= 1e-12 in value - and shared fixed rows end with diagonal = holders."""
import os

import numpy as np
import pytest

import parity_util as U
import problems

pytestmark = pytest.mark.gpu


def _ngpu():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _worker(rank, world, port, procs, mk, n, q):
    try:
        import torch
        import torch.distributed as dist
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        torch.cuda.set_device(rank)
        dev = torch.device("cuda", rank)
        dist.init_process_group("gloo", rank=rank, world_size=world)            # plumbing for the 128-byte id only
        from mrhyde_b200 import Engine, lib_path
        p = getattr(problems, mk)(n)
        idt = torch.zeros(128, dtype=torch.uint8)
        e = Engine(p["physics"], 3, p.get("order", 1), device=rank, settings=p.get("settings", {}))
        if rank == 0:
            idt = torch.from_numpy(e.comm_unique_id().copy())
        dist.broadcast(idt, 0)
        e.close()
        e = Engine.from_problem(p, lib=lib_path(), device=rank, part=(procs, rank), comm_id=idt.numpy())
        s = e.sizes()
        xyz, var = e.dof_coords()
        gids, owner = e.partition()
        rp, ci = e.crs()
        fixed = e.dofmap(coords=False)[2]
        u = np.prod(np.sin(1.3 * xyz + 0.2), axis=1) + 0.1 * var
        tu = torch.from_numpy(u).to(dev)
        outs = []
        for jacobian in (True, False):
            res = torch.zeros(s["nlocal"], dtype=torch.float64, device=dev); jac = torch.zeros(s["nnz"], dtype=torch.float64, device=dev)
            for _ in range(2):                                   # twice: the steady-state path (buffers and plan reused)
                res.zero_(); jac.zero_()
                if jacobian:
                    e.assemble_jacres(tu, res, jac)
                else:
                    e.assemble_res(tu, res)
                torch.cuda.synchronize()
            outs.append((res.cpu().numpy(), jac.cpu().numpy()))
            if jacobian:
                path, launches = e.kernel_path(), e.launch_count()          # (residual-only passes scatter directly)
        # host-buffer entry point: with a communicator attached it exchanges before the copy-out (same rows as the device buffers)
        hres, hjac = np.zeros(s["nlocal"]), np.zeros(s["nnz"])
        e.assemble_jacres_host(u, hres, hjac)
        assert np.abs(hres - outs[0][0]).max() <= 1e-13 * max(np.abs(outs[0][0]).max(), 1e-300), "host-buffer residual differs after the exchange"
        assert np.abs(hjac - outs[0][1]).max() <= 1e-13 * np.abs(outs[0][1]).max(), "host-buffer Jacobian differs after the exchange"
        grow_, gcol_, gval = e.ghost_block(values=False)
        (res, jac), (res_only, _) = outs
        # ghost values of the LAST Jacobian exchange were overwritten by nothing (the residual-only pass does not touch them)
        _, _, gval = e.ghost_block(values=True)
        # ---- single-GPU engine on the whole mesh (rank-local GPU) ----
        eg = Engine.from_problem(p, lib=lib_path(), device=rank, part=((1, 1, 1), 0))
        sg = eg.sizes()
        gxyz, gvar = eg.dof_coords()
        ggid, _ = eg.partition()
        grp, gci = eg.crs()
        gu = torch.from_numpy(np.prod(np.sin(1.3 * gxyz + 0.2), axis=1) + 0.1 * gvar).to(dev)
        gres = torch.zeros(sg["nlocal"], dtype=torch.float64, device=dev); gjac = torch.zeros(sg["nnz"], dtype=torch.float64, device=dev)
        eg.assemble_jacres(gu, gres, gjac)
        torch.cuda.synchronize()
        gr, gj = gres.cpu().numpy(), gjac.cpu().numpy()
        g2G = {int(g): i for i, g in enumerate(ggid)}
        allg = [None] * world
        dist.all_gather_object(allg, gids.tolist())
        holders = {}
        for lst in allg:
            for g in lst:
                holders[g] = holders.get(g, 0) + 1
        sr, sj = np.abs(gr).max(), np.abs(gj).max()
        ghost_of = {}
        for k in range(len(grow_)):
            ghost_of.setdefault(int(grow_[k]), {})[int(gcol_[k])] = gval[k]
        own_rows = np.nonzero(owner == rank)[0]
        G = np.array([g2G[int(g)] for g in gids])
        assert np.abs(res[own_rows] - gr[G[own_rows]]).max() <= 1e-12 * sr
        assert np.abs(res_only[own_rows] - gr[G[own_rows]]).max() <= 1e-12 * sr
        for row in own_rows:
            want = {int(c): v for c, v in zip(gci[grp[G[row]]:grp[G[row] + 1]], gj[grp[G[row]]:grp[G[row] + 1]])}
            got = {int(G[ci[t]]): jac[t] for t in range(rp[row], rp[row + 1])}
            for cg, v in ghost_of.get(int(row), {}).items():
                got[g2G[cg]] = v
            assert set(got) == set(want), (rank, row)
            for c, v in got.items():
                w = float(holders[int(gids[row])]) if (fixed[row] and c == G[row]) else want[c]
                assert abs(v - w) <= 1e-12 * sj, (rank, row, c, v, w)
        q.put((rank, "ok", path, launches))
    except Exception:   # noqa: BLE001
        import traceback
        q.put((rank, "FAIL " + traceback.format_exc(), "", 0))


@pytest.mark.parametrize("procs,mk,n", [((1, 1, 2), "config_c2", 16), ((2, 1, 1), "config_c2", 12), ((1, 1, 2), "config_c4", 6), ((1, 2, 1), "config_c5", 5),
                                          ((2, 2, 1), "config_c2", 16), ((2, 2, 2), "config_c2", 16)],
                         ids=["C2-1x1x2", "C2-2x1x1", "C4-1x1x2", "C5-1x2x1", "C2-2x2x1", "C2-2x2x2"])
def test_engine_with_nccl_exchange_matches_single_gpu(procs, mk, n):
    import torch.multiprocessing as mp
    world = int(np.prod(procs))
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29900 + world * 3 + (os.getpid() % 90)
    ps = [ctx.Process(target=_worker, args=(r, world, port, procs, mk, n, q)) for r in range(world)]
    for p in ps:
        p.start()
    out = [q.get(timeout=600) for _ in ps]
    for p in ps:
        p.join(60)
    for rank, msg, path, launches in out:
        assert msg == "ok", f"rank {rank}: {msg}"
    if mk == "config_c2":
        assert all("tiled-ws" in o[2] for o in out), out
