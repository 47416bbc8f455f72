#!/usr/bin/env python3
"""bench.py — elements/s assembled (FP64 residual + Jacobian) on BASELINE.json configs[1]:
3-D thermal, HGRAD Q1 hexes, 256^3 synthetic structured brick, general trilinear geometry path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--n 256] [--impl reference]

One "step" = one mrh_assemble_jacres call (volume + boundary terms + DBC fix-up) accumulating into the
caller's residual vector and CRS values, all tables prebuilt — what the reference times as
"computeJacRes() - total assembly" (/root/reference/src/managers/assemblyManager.hpp:1218).
N > 1 (torchrun, one rank per GPU): weak scaling by default, each rank owns a 256x256x256 z-slab of a 256x256x(256 N)
brick (the line also carries a "strong" record: the 256^3 mesh itself cut into N blocks); --scaling strong makes that the
headline.  Interface rows go to their owner rank inside mrh_assemble_jacres (library pack / NCCL send-recv / unpack,
interface tiles first and overlapped with the interior; csrc/exchange.inl).
--impl reference: the CPU restatement of the reference path (oracle/, kind "port" — the reference itself needs
Trilinos and cannot be built here) on all host cores, same metric and config, bounded sample per step.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "elements/sec assembled (FP64 residual+Jacobian)"
UNIT = "elements/s"
# BASELINE.json configs: (problem factory, timed mesh size, algorithmic bytes / element B, flops / element F)  - SURVEY.md 8d
CONFIGS = {
    "C1": ("config_c1", 4096, 120.0, 800.0, "C1: 2D thermal, HGRAD Q1 quads, structured mesh, FP64 residual+Jacobian"),
    "C2": ("config_c2", 256, 289.0, 7208.0, "C2: 3D thermal, HGRAD Q1 hexes, structured brick, FP64 residual+Jacobian"),
    "C3": ("config_c3", 128, 4370.0, 213e3, "C3: 3D convection-diffusion-reaction, HGRAD Q2 hexes, FP64 residual+Jacobian"),
    "C4": ("config_c4", 192, 2123.0, 36.5e3, "C4: 3D linear elasticity, Q1 hexes, 3 DOFs/node, FP64 residual+Jacobian"),
    "C5": ("config_c5", 128, 3701.0, 150e3, "C5: 3D Navier-Stokes SUPG/PSPG, Q1 hexes, FP64 residual+Jacobian"),
}
BYTES_PER_ELEM, FLOPS_PER_ELEM = CONFIGS["C2"][2], CONFIGS["C2"][3]
CPU_SAMPLE_N = {"C1": 512, "C2": 64, "C3": 16, "C4": 32, "C5": 32}     # bounded CPU sample of each workload (elements per direction)


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.rows.append((time.time(), ln.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for ts, ln in self.rows:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                mx = float(f[1])
                if t0 - 0.05 <= ts <= t1 + 0.15:
                    sm.append(float(f[0]))
                    for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                        if v.lower().startswith("active"):
                            reasons.add(name)
            except ValueError:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def cpu_port_rate(config, n, threads, steps, warmup):
    """Oracle (CPU restatement of the reference path, reference-shaped groups of 100 elements) on ONE n^d mesh of the
    configuration, `threads` host threads working through the groups (atomic scatter, like the reference's OpenMP build).
    Returns (median elements/s, median s per pass, timing buckets of the last pass)."""
    import parity_util as U
    import problems
    from oracle.pyoracle import Oracle
    o = Oracle(U.oracle_problem(getattr(problems, CONFIGS[config][0])(n)))
    _, _, _, xyz, var = o.dofmap()
    u = problems.synthetic_state(xyz, var)
    for _ in range(warmup):
        o.assemble(u, jacobian=True, nthreads=threads)
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter()
        o.assemble(u, jacobian=True, nthreads=threads)
        ts.append(time.perf_counter() - t0)
    med = float(np.median(ts))
    return o.nelem / med, med, {k: float(v) for k, v in o.timing().items()}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    cfg = args.config
    n = CPU_SAMPLE_N[cfg]
    warm = 1
    steps = max(5, min(args.steps, 20))
    rate, sps, buckets = cpu_port_rate(cfg, n, cores, steps, warm)
    dim = 2 if cfg == "C1" else 3
    sample = (f"oracle port, one {n}^{dim} mesh of the workload per step, {cores} threads over the 100-element groups (atomic scatter), "
              f"1 warm-up + median of {steps}; buckets of the last pass (thread-averaged s): gather {buckets['gather']:.3f}, "
              f"physics {buckets['physics']:.3f}, scatter {buckets['scatter']:.3f}")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warm,
        "ms_per_step": sps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": CONFIGS[cfg][4], "mesh": f"{n}^{dim} (bounded sample of {CONFIGS[cfg][1]}^{dim}; per-element cost is size-independent)"},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def time_config(cfg, n, dev, local, K, W):
    """Device-buffer timing of one configuration on one GPU: returns a dict with ms per assembly, elements/s and the
    fraction of the FP64 roof (filled by the caller), plus the engine for further use (caller closes it)."""
    import torch
    import problems
    from mrhyde_b200 import Engine
    t0 = time.time()
    e = Engine.from_problem(getattr(problems, CONFIGS[cfg][0])(n), device=local)
    s = e.sizes()
    xyz, var = e.dof_coords()
    u = torch.from_numpy(problems.synthetic_state(xyz, var)).to(dev)
    res = torch.zeros(s["nlocal"], dtype=torch.float64, device=dev)
    jac = torch.zeros(s["nnz"], dtype=torch.float64, device=dev)
    setup = time.time() - t0
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(W):
        e.assemble_jacres(u, res, jac, stream=st)
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(K):
        e.assemble_jacres(u, res, jac, stream=st)
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1) / K
    ok = bool(torch.isfinite(res).all().item() and torch.isfinite(jac).all().item())
    out = {"mesh": n, "elements": s["nelem"], "ms": ms, "elements_per_s": s["nelem"] / (ms * 1e-3), "kernel_ms": e.volume_kernel_ms(),
           "path": e.kernel_path(), "launches": e.launch_count(), "setup_s": round(setup, 1), "finite": ok}
    e.close()
    del u, res, jac
    torch.cuda.empty_cache()
    return out


def run_gpu(args):
    import torch
    import torch.distributed as dist
    import problems
    from mrhyde_b200 import Engine, load_library

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a GPU: the engine has no CPU fallback"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cfg = args.config
    mk, n_full, B_ELEM, F_ELEM, workload = CONFIGS[cfg]
    n = args.n if args.n > 0 else n_full
    dim = 2 if cfg == "C1" else 3
    W, K = max(args.warmup, 3), args.steps
    if world > 1 and cfg != "C2":
        pass     # any configuration partitions; only C2's kernels overlap the exchange with the interior

    # ---- set-up (untimed): mesh, DOFs, CRS graph, slot patterns, synthetic state -------------------------
    STRONG_GRID = {1: (1, 1, 1), 2: (1, 1, 2), 4: (1, 2, 2), 8: (2, 2, 2)}

    def make_engine(strong):
        """this rank's block: weak = n^3 elements per GPU (z-slabs of an n x n x (n world) brick), strong = the n^3 mesh itself
        cut into a Px x Py x Pz grid; the NCCL id travels over torch.distributed (plumbing), the exchange is the library's"""
        p = getattr(problems, mk)(n)
        if world == 1:
            return Engine.from_problem(p, device=local), (1, 1, 1)
        if strong:
            procs = STRONG_GRID.get(world) or (1, 1, world)
        else:
            procs = (1, 1, world)
            p["mesh"]["NZ"] = n * world
            p["mesh"]["zmax"] = float(world)
        idt = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            idt.copy_(torch.from_numpy(Engine(p["physics"], dim, p.get("order", 1), device=local).comm_unique_id().copy()))
        dist.broadcast(idt, 0)
        return Engine.from_problem(p, device=local, part=(procs, rank), comm_id=idt.cpu().numpy()), procs

    t_setup = time.time()
    strong_mode = args.scaling == "strong"
    e, procs = make_engine(strong_mode)
    s = e.sizes()
    xyz, var = e.dof_coords()
    u_host = problems.synthetic_state(xyz, var)
    u = torch.from_numpy(u_host).to(dev)
    res = torch.zeros(s["nlocal"], dtype=torch.float64, device=dev)
    jac = torch.zeros(s["nnz"], dtype=torch.float64, device=dev)
    xch = None
    t_setup = time.time() - t_setup
    stream = torch.cuda.current_stream().cuda_stream
    nelem_total = torch.tensor([s["nelem"]], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(nelem_total)
    nelem_total = int(nelem_total.item())

    def step():
        e.assemble_jacres(u, res, jac, stream=stream)      # with a communicator attached this includes the interface exchange

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(W):
        step()
    sync()
    launches_per_step = e.launch_count()
    sampler = ClockSampler(local) if rank == 0 else None
    vol_ms = []
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync()
    t0 = time.time()
    ev0.record()
    for _ in range(K):
        step()
    ev1.record()
    sync()
    t1 = time.time()
    ms_total = ev0.elapsed_time(ev1)
    # dominant-kernel time: CUDA events recorded by the engine around the volume kernel on this stream
    for _ in range(3):
        e.assemble_jacres(u, res, jac, stream=stream)
        vol_ms.append(e.volume_kernel_ms())
    torch.cuda.synchronize()
    clocks = sampler.stop(t0, t1) if sampler else None
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = t.item() / K
    value = nelem_total / (ms_step * 1e-3)

    # ---- strong scaling of the SAME n^3 mesh (BASELINE.json north_star: "the mesh is partitioned ... across the 8 GPUs") ----
    strong_rec = None
    if world > 1 and not strong_mode and not args.no_other and not args.kernel_only:
        try:          # a failure of this extra leg must not lose the headline line
            e2, procs2 = make_engine(True)
            s2 = e2.sizes()
            xyz2, var2 = e2.dof_coords()
            u2 = torch.from_numpy(problems.synthetic_state(xyz2, var2)).to(dev)
            res2 = torch.zeros(s2["nlocal"], dtype=torch.float64, device=dev)
            jac2 = torch.zeros(s2["nnz"], dtype=torch.float64, device=dev)
            for _ in range(W):
                e2.assemble_jacres(u2, res2, jac2, stream=stream)
            sync()
            ev0.record()
            for _ in range(K):
                e2.assemble_jacres(u2, res2, jac2, stream=stream)
            ev1.record()
            sync()
            t2 = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
            dist.all_reduce(t2, op=dist.ReduceOp.MAX)
            ms2 = t2.item() / K
            strong_rec = {"scaling": "strong", "mesh": f"{n}^{dim} cut into {procs2[0]}x{procs2[1]}x{procs2[2]} blocks", "ms_per_step": ms2,
                          "value": n ** dim / (ms2 * 1e-3), "unit": UNIT, "elements_per_gpu": s2["nelem"], "gpu_launches_per_step": e2.launch_count(),
                          "volume_kernel": e2.kernel_path(), "note": "efficiency = value / (N x the N=1 value of this bench)"}
            e2.close()
            del u2, res2, jac2
            torch.cuda.empty_cache()
        except Exception as ex:
            strong_rec = {"scaling": "strong", "error": str(ex)[:300]}

    if args.kernel_only:
        if rank == 0:
            print(json.dumps({"kernel_only": True, "ms_per_step": ms_step, "value": value, "volume_kernel_ms": float(np.mean(vol_ms)), "path": e.kernel_path()}))
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- end to end through the host-buffer entry point: pinned u in, res + Jacobian values out, every step --
    Ke = max(1, min(K, 5))
    hu = torch.from_numpy(u_host).pin_memory()
    hres = torch.empty(s["nlocal"], dtype=torch.float64).pin_memory()
    hjac = torch.empty(s["nnz"], dtype=torch.float64).pin_memory()
    e.assemble_jacres_host(hu, hres, hjac, stream=stream)
    sync()
    te = time.perf_counter()
    for _ in range(Ke):
        e.assemble_jacres_host(hu, hres, hjac, stream=stream)     # synchronises its stream before returning
    torch.cuda.synchronize()
    te = time.perf_counter() - te
    t = torch.tensor([te], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = nelem_total * Ke / t.item()
    checksum = float(hres.abs().sum())

    kpath = e.kernel_path()
    if rank == 0:
        pk = peaks()
        vk = float(np.mean(vol_ms)) * 1e-3
        L = load_library()
        import ctypes as C
        dfma, red, cpy = C.c_double(0), C.c_double(0), C.c_double(0)
        L.mrh_microbench(local, C.byref(dfma), C.byref(red), C.byref(cpy))
        hbm_peak = pk.get("hbm_gbs", 6650.0)
        ach_gbs = B_ELEM * s["nelem"] / vk / 1e9
        ach_tf = F_ELEM * s["nelem"] / vk / 1e12
        cores = os.cpu_count() or 1
        cpu_n = CPU_SAMPLE_N[cfg]
        cpu_rate, cpu_sps, cpu_buckets = cpu_port_rate(cfg, cpu_n, cores, 5, 1)
        # the other BASELINE.json configurations at their full sizes, same process, same clocks (N = 1, default run only)
        other = {}
        if world == 1 and cfg == "C2" and args.n <= 0 and not args.no_other:
            del u, res, jac, hu, hres, hjac
            e.close()
            torch.cuda.empty_cache()
            for oc in ("C1", "C3", "C4", "C5"):
                try:
                    r = time_config(oc, CONFIGS[oc][1], dev, local, 5, 3)
                    r["fp64_tflops"] = CONFIGS[oc][3] * r["elements_per_s"] / 1e12
                    r["fp64_frac"] = r["fp64_tflops"] / dfma.value if dfma.value > 0 else None
                    r["hbm_gbs_algorithmic"] = CONFIGS[oc][2] * r["elements_per_s"] / 1e9
                    r["flops_per_elem"], r["bytes_per_elem"] = CONFIGS[oc][3], CONFIGS[oc][2]
                    other[oc] = r
                except Exception as ex:          # a failure here must not lose the headline line
                    other[oc] = {"error": str(ex)[:200]}
        traffic, traffic_src = None, None
        try:      # DRAM bytes of one launch of the dominant kernel from the committed ncu --set full capture of this command
            pname = next(f for f in ("r2_ws_thermal_n256_ncu_summary.json", "r1_ws_thermal_n256_ncu_summary.json") if os.path.exists(os.path.join(ROOT, "profiles", f)))
            prof = json.load(open(os.path.join(ROOT, "profiles", pname)))
            if cfg == "C2" and n == 256 and world == 1 and kpath == prof.get("kernel_path", "nvrtc-specialised+tiled-ws"):
                traffic = prof["derived"]["dram_bytes_per_launch"]
                traffic_src = f"profiles/{pname} (dram__bytes_read.sum + dram__bytes_write.sum, one launch)"
        except Exception:
            pass
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "strong" if strong_mode else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload,
                       "mesh": (f"{n}^{dim} cut into {procs[0]}x{procs[1]}x{procs[2]} blocks" if strong_mode and world > 1 else
                                f"{n}^{dim} per GPU" + (f", {world} z-slabs" if world > 1 else "")), "elements_per_gpu": s["nelem"],
                       "exchange": "library: pack kernel -> ncclSend/ncclRecv per neighbour -> unpack kernel, interface tiles first and overlapped" if world > 1 else None,
                       "dofs_per_gpu": s["nlocal"], "nnz_per_gpu": s["nnz"], "slot_patterns": s["npatterns"], "volume_kernel": kpath,
                       "l2": f"inputs+outputs ({8e-9 * (s['nnz'] + 2 * s['nlocal']):.1f} GB/step) exceed the 126 MB L2; no flush needed",
                       "timing": "CUDA events on the launch stream, K steps, max over ranks; zeroing of res / J is the caller's job "
                                 "(solverManager.cpp:1528-1533) and outside the timed step, which accumulates", "setup_s": round(t_setup, 1)},
            "roofline": {"bound": "hbm", "achieved": ach_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": ach_gbs / hbm_peak,
                         "traffic": traffic, "traffic_unit": "bytes per launch", "traffic_source": traffic_src, "peak_source": "MEASURED_PEAKS.json (of measured)" if pk else "fallback 6650 (of fallback)",
                         "kernel": "volume assembly kernel (" + kpath + ")", "kernel_ms": vk * 1e3,
                         "bytes_per_elem": B_ELEM, "binding_roof": "fp64 (see roofline_fp64)"},
            "roofline_fp64": {"bound": "fp64", "achieved": ach_tf, "peak": dfma.value, "unit": "TFLOP/s",
                              "frac": ach_tf / dfma.value if dfma.value > 0 else None, "flops_per_elem": F_ELEM,
                              "peak_source": "DFMA micro-benchmark run in this process (mrh_microbench)",
                              "red_f64_gops": red.value, "copy_gbs_here": cpy.value},
            "cpu_baseline": {"value": cpu_rate, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"oracle (CPU restatement, reference-shaped groups of 100, atomic scatter) on one {cpu_n}^{dim} mesh of the workload, "
                                       f"{cores} threads, 1 warm-up + median of 5", "ms_per_pass": cpu_sps * 1e3, "buckets_s": cpu_buckets},
            "other_configs": other,
            "strong": strong_rec,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 8 * s["nlocal"],
                    "d2h_bytes_per_step": 8 * (s["nlocal"] + s["nnz"]), "steps": Ke, "checksum_abs_res": checksum,
                    "api": "mrh_assemble_jacres_host (pinned host buffers; u in, residual + CRS values out)"},
            "gpu_launches": launches_per_step * K,
            "clocks": clocks,
        }
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--n", type=int, default=0, help="elements per direction per GPU (default: the configuration's BASELINE.json size)")
    ap.add_argument("--config", default="C2", choices=sorted(CONFIGS), help="BASELINE.json configuration (default C2, the headline)")
    ap.add_argument("--no-other", action="store_true", help="skip the other_configs leg of the default run")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"], help="N > 1: n^3 elements per GPU (weak, default) or the n^3 mesh cut into N blocks (strong)")
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--kernel-only", action="store_true", help="profiling aid: skip the e2e and cpu_baseline legs")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
