#!/usr/bin/env python
"""bench.py — FP64 RHS+Jacobian cell-updates/s of the FVMCC ComputeRHS path on N B200s of one node.

One "step" = one full ComputeRhsJacob evaluation of BASELINE.json configs[2] (3D Navier-Stokes, hexahedra, Roe +
viscous flux, LSQ + Venkatakrishnan, finite-difference block Jacobian) on a synthetic jittered hex box: applyBC,
LSQ gradients, nodal extrapolation, limiter, face fluxes, residual gather, interior + boundary FD Jacobian blocks,
diagonal gather, pseudo-steady time term.  `value` = updatable cells / step time with everything
resident in HBM; `e2e` = the same through the C ABI with HOST buffers (states H2D from pinned memory, rhs and
updateCoeff D2H inside the timed region; the block Jacobian stays on the device for the linear solver, which is how
the plugin hands it to PETSc-cusparse / Paralution).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--cells-per-gpu M] [--impl reference]

N>1 is launched by torchrun (one rank per GPU); every rank owns one partition of the same size (weak scaling).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "fp64_rhs_jacobian_cell_updates_per_s"
UNIT = "cells/s"
# algorithmic bytes per cell-update, 3D NS hex, 2nd order, viscous: A_rhs 744 + A_jac 1400
A_RHS, A_JAC = 744, 1400


# Compulsory bytes of the interior-face Jacobian kernel per face: per-face arrays once (normal 24, area 8,
# face centre 24, L/R ids 8, node ptr 8, node ids 16, diamond volume 8, two BSR slots 8 = 104), per-cell arrays once per
# cell = 1/3 per face (state 40, gradients 120, limiter 40, centre 24 -> 75), per-node arrays once per node = 1/3 per face
# (coordinates 24, nodal state 40 -> 21), writes: two 5x5 blocks 400, face flux 40, three update coefficients 24.
INNER_FACE_JAC_BYTES = 104 + 75 + 21 + 400 + 40 + 24


def measured_traffic(kernel, items):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch from the committed `ncu --set full` capture
    (profiles/traffic.json: bytes per item measured at the capture's size), scaled to this launch's item count."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(p):
        return None, None
    t = json.load(open(p)).get(kernel)
    if not t:
        return None, None
    return int(t["bytes_per_item"]*items), t["source"]


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def flagship(n, workload="ns3d"):
    """configs[2]: jittered hex box n^3, Pvt variables, one inlet / outlet / no-slip wall / mirrors;
    `mhd3d` = configs[3]: the same box with MHD3DProjection (q = 9), HLL, LSQ + Venkatakrishnan."""
    import cases
    from coolfluid_b200 import meshgen
    em = meshgen.hex_box(n, n, n) if not workload.startswith("euler2d") else None
    if workload == "mhd3d":
        return em, cases.mhd3d_cfg(), cases.bcs_mhd3d()
    if workload.startswith("euler2d"):
        from coolfluid_b200 import config as B
        em = meshgen.quad_channel(n, n)
        jac = "analytic" if workload == "euler2d_analytic" else True
        return em, B.make_config(2, flux=B.FS_AUSMPLUSUP, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, jacobian=jac,
                                 mach_inf=2.0, choice_a12=1), cases.bcs_euler2d()
    if workload == "ns3d_corrected":
        return em, cases.ns3d_cfg(derivative="corrected"), cases.bcs_ns3d()
    return em, cases.ns3d_cfg(), cases.bcs_ns3d()


WORKLOADS = {
    # name: (description, pvt flag of cases.init_state, A_rhs, A_jac)   [SURVEY.md §8d algorithmic bytes per cell-update]
    "ns3d": ("configs[2]: 3D Navier-Stokes, %d^3 = %d jittered hex cells per GPU, Pvt variables, Roe + viscous "
             "(DiamondVolume3D) flux, LinearLS3D + Venktn3D, finite-difference block Jacobian (NumJacob) into device "
             "BSR, pseudo-steady time term", True, 744, 1400),
    "ns3d_corrected": ("configs[2] with DerivativeStrategy = Corrected3D instead of DiamondVolume3D: 3D Navier-Stokes, "
                       "%d^3 = %d jittered hex cells per GPU, Pvt variables, Roe + viscous flux, LinearLS3D + Venktn3D, "
                       "NumJacob into device BSR, pseudo-steady time term", True, 744, 1400),
    "euler2d_analytic": ("configs[1]: 2D Euler, %d^2 = %d quadrilaterals per GPU, Cons variables, AUSM+up (machInf 2, choiceA12 1), "
                         "LinearLS2D + Venktn2D, ANALYTIC flux Jacobian (NumJacobAnalytic) into device BSR, pseudo-steady time "
                         "term", False, 368, 640),
    "euler2d_numjac": ("configs[1] with the finite-difference Jacobian instead: 2D Euler, %d^2 = %d quadrilaterals per GPU, "
                       "AUSM+up, LinearLS2D + Venktn2D, NumJacob, pseudo-steady time term", False, 368, 640),
    "mhd3d": ("configs[3]: 3D ideal MHD (MHD3DProjection, q = 9), %d^3 = %d jittered hex cells per GPU, Cons variables, "
              "HLL flux, LinearLS3D + Venktn3D, finite-difference 9x9 block Jacobian into device BSR, pseudo-steady "
              "time term", "mhd", 960, 4536),
}


def slab_partition(n, rank, world, workload):
    """N > 1: rank r owns a slab of n^3 cells of an n x n x (n*world) box cut along z and carries the two overlap
    layers of each interior side (LSQ needs 2: LeastSquareP1Setup.cxx:60-63).  The slab is generated locally; its cells
    are renumbered [own cells that see no overlap cell | own cells next to the overlap | overlap cells] so that the
    gradients of the first group overlap the state halo exchange; the halo lists pair this slab's two outermost own
    layers with the neighbour's overlap layers (same (i, j, k) order on both sides).  Returns (ElementMesh, Partition)."""
    from coolfluid_b200 import meshgen, partition
    lo = 2 if rank > 0 else 0
    hi = 2 if rank < world - 1 else 0
    K = n + lo + hi
    em = meshgen.hex_box(n, n, K, lengths=(1.0, 1.0, K/float(n)), seed=12345 + rank)
    nl = n*n
    k_of = np.arange(em.n_cells)//nl
    own = (k_of >= lo) & (k_of < lo + n)
    near = own & (((k_of == lo) & (lo > 0)) | ((k_of == lo + n - 1) & (hi > 0)))
    group = np.where(own & ~near, 0, np.where(near, 1, 2))
    perm = np.argsort(group, kind="stable").astype(np.int64)        # new -> old
    inv = np.empty_like(perm); inv[perm] = np.arange(em.n_cells)    # old -> new
    em.cell_nodes = np.ascontiguousarray(em.cell_nodes.reshape(-1, 8)[perm].reshape(-1))
    em.trs_face_cell = np.ascontiguousarray(inv[em.trs_face_cell]).astype(em.trs_face_cell.dtype)
    p = partition.Partition()
    p.rank, p.nranks, p.em = rank, world, em
    p.updatable = (group[perm] < 2).astype(np.uint8)
    p.n_interior = int((group == 0).sum())
    p.n_updatable = int(own.sum())
    layer = lambda k0: inv[np.arange(k0*nl, (k0 + 2)*nl)].astype(np.int32)
    if lo:
        p.send[rank - 1] = layer(lo); p.recv[rank - 1] = layer(0)
    if hi:
        p.send[rank + 1] = layer(lo + n - 2); p.recv[rank + 1] = layer(lo + n)
    return em, p


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.gpu = gpu
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=3)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def cpu_baseline(sample_n, evals, threads):
    """The oracle (CPU restatement of the reference path, `kind: port`) on `threads` host threads, one independent
    partition of sample_n^3 hex cells per thread (the MPI path runs one rank per core on its own partition)."""
    import cases
    from oracle import binding as ob
    em, cfg, bcs = flagship(sample_n)
    fm = ob.build_mesh(em)
    q = 5
    U0 = cases.init_state(fm, q, True)
    ob.lib(True)
    outs = [ob.alloc_outputs(cfg, fm, True) for _ in range(threads)]
    Us = [U0.copy() for _ in range(threads)]
    past = np.concatenate([U0[:fm.nCells*q]*(1 + 1e-3), np.zeros(fm.nGhosts*q)])

    def work(i):
        for _ in range(evals):
            outs[i].updateCoeff[:] = 0.0
            ob.compute(cfg, fm, bcs, Us[i], jacobian=True, omp=True, out=outs[i])
            ob.time_rhs(cfg, fm, bcs, Us[i], past, outs[i], omp=True)
    th = [threading.Thread(target=work, args=(i,)) for i in range(threads)]
    t0 = time.perf_counter()
    for t in th:
        t.start()
    for t in th:
        t.join()
    dt = time.perf_counter() - t0
    cells = fm.nCells*threads*evals
    return cells/dt, dt, fm.nCells


def run_reference(args, rank):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    n = args.cpu_sample
    vals = []
    for _ in range(args.warmup):
        cpu_baseline(n, 1, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        v, dt, nc = cpu_baseline(n, 1, threads)
        vals.append(v)
    total = time.perf_counter() - t0
    v = statistics.median(vals)
    sample = "%d independent partitions of %d^3 hex cells (one per host thread), 1 evaluation per step" % (threads, n)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3*total/max(args.steps, 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "3D Navier-Stokes hex box, Roe + viscous (DiamondVolume3D), LSQ + Venktn3D, FD Jacobian, "
                               "pseudo-steady time term; CPU sample %s" % sample},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "the reference itself (Boost/MPI/PETSc) cannot be built on this image; this is the oracle port "
                "(-O3, no FP contraction), one instance per host thread"}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cfb200")
    ap.add_argument("--n", "--edge", dest="n", type=int, default=220, help="hex box edge per GPU (220^3 = 10.6 M cells, configs[2])")
    ap.add_argument("--workload", default="ns3d", choices=sorted(WORKLOADS), help="ns3d = configs[2] (the metric), euler2d_* = configs[1], mhd3d = configs[3]")
    ap.add_argument("--cpu-sample", type=int, default=36, help="edge of the per-thread hex box of the CPU baseline")
    ap.add_argument("--cpu-evals", type=int, default=24, help="evaluations per thread of the cpu_baseline leg (~10-20 s)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import cases
    from coolfluid_b200 import lib as cl
    L = cl.load()
    if L.device_count() < 1:
        raise SystemExit("bench.py needs a CUDA device: cfb200 has no CPU path")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_
        dist = dist_
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    wl_desc, wl_pvt, a_rhs, a_jac = WORKLOADS[args.workload]
    t0 = time.perf_counter()
    em, cfg, bcs = flagship(args.n if world == 1 else 2, args.workload)
    q = cfg.nEqs
    part = None
    if world > 1:
        em, part = slab_partition(args.n, rank, world, args.workload)
    fm = cl.build_mesh(em, L)
    if part is not None:
        fm.updatable = np.ascontiguousarray(part.updatable)
    t_mesh = time.perf_counter() - t0
    U0 = cases.init_state(fm, q, wl_pvt)[:fm.nCells*q]
    eng = cl.Engine(cfg, local_rank, lib=L)
    t0 = time.perf_counter()
    eng.set_mesh(fm, bcs)
    t_setup = time.perf_counter() - t0
    nc = fm.nCells
    n_upd = nc if part is None else part.n_updatable
    if part is not None:
        # NCCL communicator of the engine (cfb200_comm_init) + halo lists; the unique id travels over torch.distributed
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(cl.nccl_unique_id(L)), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        eng.comm_init(rank, world, bytes(idt.cpu().numpy().tobytes()))
        eng.set_halo(part)
    # pinned host buffers of the reference-facing call (states in, rhs / updateCoeff out)
    hU = torch.empty(nc*q, dtype=torch.float64, pin_memory=True)
    hU.numpy()[:] = U0
    hR = torch.empty(nc*q, dtype=torch.float64, pin_memory=True)
    hC = torch.empty(nc, dtype=torch.float64, pin_memory=True)
    eng.set_states(hU.numpy())
    eng.set_past_states(hU.numpy()*(1 + 1e-3))
    stream = torch.cuda.ExternalStream(eng.stream(), device=torch.device("cuda", local_rank))

    def step_resident():
        eng.set_update_coeff(None)           # the updater zeroes updateCoeff between evaluations (UpdateSol.cxx:192)
        if part is not None:
            eng.halo_begin()                 # state halo over NCCL; compute_rhs overlaps it with the interior gradients
        eng.compute_rhs(apply_bc=True, download=False)
        eng.time_rhs(download=False)

    def step_e2e():
        eng.set_states(hU.numpy())
        eng.set_update_coeff(None)
        if part is not None:
            eng.halo_begin()
        eng.compute_rhs(apply_bc=True, download=True, out_rhs=hR.numpy(), out_uc=hC.numpy())
        eng.time_rhs(download=False)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(steps):
            fn()
        b.record(stream)
        barrier()
        ms = a.elapsed_time(b)
        if dist is not None:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    for _ in range(args.warmup):
        step_resident()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    eng.enable_timing(True)
    kern = {}
    launches = [0]

    def step_resident_timed():
        step_resident()
        tl = eng.last_timings()
        launches[0] += len(tl) + 1          # + the time-term kernel is in the list; +1 = updateCoeff memset excluded
        per = {}
        for name, ms in tl:                   # chunked kernels appear once per chunk: sum them per step
            per[name] = per.get(name, 0.0) + ms
        for name, ms in per.items():
            kern.setdefault(name, []).append(ms)
    ms_total = timed(step_resident_timed, args.steps)
    clocks = sampler.stop() if rank == 0 else None
    eng.enable_timing(False)
    ms_step = ms_total/args.steps
    total_cells = n_upd*world            # updatable cells only (overlap cells are recomputed copies)
    value = total_cells/(ms_step*1e-3)

    for _ in range(max(1, args.warmup//2)):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps)/args.steps
    e2e_value = total_cells/(ms_e2e*1e-3)

    status = eng.L.status(eng.h)
    if rank == 0:
        peak, peak_src = measured_peaks()
        kavg = {k: sum(v)/len(v) for k, v in kern.items()}
        top = max(kavg, key=kavg.get)
        # algorithmic bytes of the dominant kernel per launch (DESIGN.md "kernels"): the interior-face FD Jacobian
        # writes two q x q blocks per interior face and reads both cell states, gradients, limiters, the face flux,
        # normal, area and the 4 nodal states: see DESIGN.md for the per-face count
        per_launch = {"innerFaceJac": fm.nInner*INNER_FACE_JAC_BYTES} if args.workload == "ns3d" else {}
        algo = per_launch.get(top, nc*(a_rhs + a_jac))
        achieved = algo/(kavg[top]*1e-3)/1e9
        traffic, traffic_src = (measured_traffic(top, fm.nInner if top == "innerFaceJac" else nc)
                                if args.workload == "ns3d" else (None, None))
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": wl_desc % (args.n, n_upd),
                       "cells_per_gpu": nc, "faces_per_gpu": int(fm.nFaces), "bsr_blocks_per_gpu": int(eng.nnzb),
                       "l2": "inputs (%.1f GB BSR + %.2f GB states/gradients) far larger than the 126 MB L2; no flush needed"
                             % (eng.nnzb*q*q*8/1e9, nc*q*8*6/1e9),
                       "partitioning": ("single partition" if part is None else
                                        "slab per GPU (%d updatable + %d overlap cells on rank 0), state halo exchange over "
                                        "NCCL every step (%d cells x %d doubles per neighbour), overlapped with the "
                                        "interior gradients" % (n_upd, nc - n_upd, 2*args.n*args.n, q)),
                       "mesh_build_s": round(t_mesh, 2), "device_setup_s": round(t_setup, 2)},
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": int(nc*q*8*world), "d2h_bytes_per_step": int((nc*q*8 + nc*8)*world)},
            "gpu_launches": int(launches[0]),
            "kernels_ms": {k: round(v, 4) for k, v in sorted(kavg.items(), key=lambda kv: -kv[1])},
            "roofline": {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved/peak, "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": int(algo), "avg_launch_ms": kavg[top]},
            "step_roofline": {"bytes_per_cell": a_rhs + a_jac, "achieved_gbs": (a_rhs + a_jac)*value/world/1e9,
                              "frac_of_hbm_peak": (a_rhs + a_jac)*value/world/1e9/peak},
            "clocks": clocks, "status": int(status),
        }
        if world == 1 and not args.no_cpu_baseline and args.workload == "ns3d":
            threads = os.cpu_count() or 1
            v, dt, ncs = cpu_baseline(args.cpu_sample, args.cpu_evals, threads)
            out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                   "sample": "%d independent partitions of %d^3 hex cells (one oracle instance per host "
                                             "thread), %d evaluations each, %.1f s" % (threads, args.cpu_sample,
                                                                                       args.cpu_evals, dt)}
        print(json.dumps(out))
    eng.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
