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


# FP64 roofline of the dominant kernel: thread-level FP64 instructions per interior face of the
# finite-difference Jacobian kernel from the committed ncu source-page capture
# (profiles/r02_ncu_innerface_pipe1_source_lines.md: DMUL 966 M + DADD 739 M + DFMA 282 M warp instructions per launch
# at 6.24 M faces), and the measured no-FMA FP64 issue peak of tools/microbench/fp64_lat.cu
# (profiles/r02_fp64_microbench.txt: DADD 567.7 G warp-instr/s over 148 SMs)
INNER_FACE_JAC_FP64_PER_FACE = 10190.0
FP64_PEAK_WARP_INSTR = 567.7e9


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
    em = meshgen.hex_box(n, n, n, nz_total=n) if not workload.startswith("euler2d") else None
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
    if workload == "ns3d_tet":
        return meshgen.tet_box(n, n, n), cases.ns3d_cfg(), cases.bcs_ns3d()
    if workload == "ns3d_roesa":
        from coolfluid_b200 import config as B
        return em, cases.ns3d_cfg(flux=B.FS_ROESA), cases.bcs_ns3d()
    if workload == "ns3d_ausm_analytic":
        from coolfluid_b200 import config as B
        return em, cases.ns3d_cfg(flux=B.FS_AUSMPLUSUP, jacobian="analytic", mach_inf=0.35), cases.bcs_ns3d()
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
    "ns3d_ausm_analytic": ("configs[2] with the analytic Jacobians of a19 (NumJacobAnalytic: AUSM+up ComputedFluxdPvt + "
                           "NSFluxAndJacob): 3D Navier-Stokes, %d^3 = %d jittered hex cells per GPU, Pvt variables, AUSM+up "
                           "(machInf 0.35) + viscous (DiamondVolume3D) flux, LinearLS3D + Venktn3D, block Jacobian into device "
                           "BSR, pseudo-steady time term", True, 744, 1400),
    "ns3d_roesa": ("configs[2] with FluxSplitter = RoeSA (Sanders' H-correction over the faces of the two cells, the "
                   "splitter of the reference's cylinder case): 3D Navier-Stokes, %d^3 = %d jittered hex cells per GPU, Pvt "
                   "variables, RoeSA + viscous (DiamondVolume3D) flux, LinearLS3D + Venktn3D, NumJacob into device BSR, "
                   "pseudo-steady time term", True, 744, 1400),
    "ns3d_tet": ("configs[2] on tetrahedra: 3D Navier-Stokes, %d^3 boxes split into 6 tetrahedra each (Kuhn), %d cells "
                 "per GPU, Pvt variables, Roe + viscous (DiamondVolume3D) flux, LinearLS3D + Venktn3D (stencils of ~60 vertex "
                 "neighbours), NumJacob into device BSR, pseudo-steady time term", True, 763, 1000),
    "mhd3d": ("configs[3]: 3D ideal MHD (MHD3DProjection, q = 9), %d^3 = %d jittered hex cells per GPU, Cons variables, "
              "HLL flux, LinearLS3D + Venktn3D, finite-difference 9x9 block Jacobian into device BSR, pseudo-steady "
              "time term", "mhd", 960, 4536),
}


def slab_partition(n, rank, world, workload, box=None):
    """N > 1: ONE global box of n x n x (n*world) cells (jitter = hash of the global node index: every rank generates
    only its own slab, node for node a piece of that mesh), cut along z.  Rank r owns cell layers [r*n, (r+1)*n) and
    carries the two overlap layers of each interior side (LSQ needs 2: LeastSquareP1Setup.cxx:60-63).  Local cells keep
    the global order (so every gather of an updatable cell runs in the single-domain order); the cells whose gradients
    may overlap the halo exchange are given as a list.  The cut planes appear as the slab's zmin / zmax TRSs: their inner
    cells are not updatable, so they are never evaluated.  Returns (ElementMesh, Partition).
    box = (nx, ny, layers per rank) overrides the cube (tests)."""
    from coolfluid_b200 import meshgen, partition
    nx, ny, nzr = box if box is not None else (n, n, n)
    lo = 2 if rank > 0 else 0
    hi = 2 if rank < world - 1 else 0
    K = nzr + lo + hi
    nzt = nzr*world
    em = meshgen.hex_box(nx, ny, K, lengths=(1.0, 1.0*ny/nx, 1.0*nzt/nx), k0=rank*nzr - lo, nz_total=nzt)
    nl = nx*ny
    k_of = np.arange(em.n_cells)//nl
    own = (k_of >= lo) & (k_of < lo + nzr)
    near = own & (((k_of == lo) & (lo > 0)) | ((k_of == lo + nzr - 1) & (hi > 0)))
    p = partition.Partition()
    p.rank, p.nranks, p.em = rank, world, em
    p.updatable = own.astype(np.uint8)
    p.indep = np.nonzero(own & ~near)[0].astype(np.int32)
    p.n_interior = len(p.indep)
    p.n_updatable = int(own.sum())
    p.cell_l2g = (np.arange(em.n_cells) + (rank*nzr - lo)*nl).astype(np.int64)
    layer = lambda k0: np.arange(k0*nl, (k0 + 2)*nl).astype(np.int32)
    if lo:
        p.send[rank - 1] = layer(lo); p.recv[rank - 1] = layer(0)
    if hi:
        p.send[rank + 1] = layer(lo + nzr - 2); p.recv[rank + 1] = layer(lo + nzr)
    return em, p


def multi_gpu_parity(rank, world, local_rank, L, dist, torch):
    """Cross-rank correctness statement carried by every N > 1 bench line (untimed): a global 32 x 32 x (16*N) jittered
    hex box is partitioned with the general partitioner (recursive coordinate bisection, node-adjacency overlap, order
    preserving local numbering), each rank evaluates its partition on its GPU (NCCL state halo in flight while the
    independent gradients run; residual + FD block Jacobian + fused pseudo-steady term), and rank 0 compares every
    updatable cell's residual, update coefficient and Jacobian block row, bit for bit, with a single-domain evaluation of
    the same mesh on its own GPU."""
    import cases
    from coolfluid_b200 import meshgen, partition, lib as cl, config as B
    em = meshgen.hex_box(32, 32, 16*world, lengths=(1.0, 1.0, 0.5*world), nz_total=16*world)
    part = partition.rcb_partition(em, world)
    p = partition.extract(em, part, rank, world)
    cfg = cases.ns3d_cfg(fuse_time_rhs=1, cfl=3.0)
    q = cfg.nEqs
    bcs = cases.bcs_ns3d()
    fm_g = cl.build_mesh(em, L)
    Ug = cases.init_state(fm_g, q, True)[:fm_g.nCells*q].reshape(-1, q)
    fm = cl.build_mesh(p.em, L)
    fm.updatable = np.ascontiguousarray(p.updatable)
    lb = bcs + ([B.make_bc(B.BC_SUPER_OUTLET)] if p.em.trs_names[-1] == "PartitionFaces" else [])
    e = cl.Engine(cfg, local_rank, lib=L)
    e.set_mesh(fm, lb)
    # overlap cells start from garbage: only the halo exchange can make the result right
    Ul = Ug[p.cell_l2g].copy()
    Ul[p.updatable == 0] = 1e30
    e.set_states(Ul.reshape(-1))
    idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        idt.copy_(torch.frombuffer(bytearray(cl.nccl_unique_id(L)), dtype=torch.uint8))
    dist.broadcast(idt, 0)
    e.comm_init(rank, world, bytes(idt.cpu().numpy().tobytes()))
    e.set_halo(p)
    e.set_past_states(Ug[p.cell_l2g].reshape(-1)*(1 + 1e-3))
    e.halo_begin()
    rhs, uc = e.compute_rhs()
    jac = e.bsr_values().reshape(-1, q*q)
    rp, ci = e.bsr_pattern()
    e.close()
    upd = p.updatable == 1
    brow = np.repeat(np.arange(fm.nCells), np.diff(rp))
    keep = upd[brow]
    mine = dict(gid=p.cell_l2g[upd], rhs=rhs.reshape(-1, q)[upd], uc=uc[upd], jac=jac[keep],
                jrow=p.cell_l2g[brow[keep]], jcol=p.cell_l2g[ci[keep]])
    allr = [None]*world
    dist.gather_object(mine, allr if rank == 0 else None, dst=0)
    if rank != 0:
        return None
    e = cl.Engine(cfg, local_rank, lib=L)
    e.set_mesh(fm_g, bcs)
    e.set_states(Ug.reshape(-1))
    e.set_past_states(Ug.reshape(-1)*(1 + 1e-3))
    rhs_g, uc_g = e.compute_rhs()
    jac_g = e.bsr_values().reshape(-1, q*q)
    rp_g, ci_g = e.bsr_pattern()
    e.close()
    rhs_g = rhs_g.reshape(-1, q)
    key_g = np.repeat(np.arange(fm_g.nCells, dtype=np.int64), np.diff(rp_g))*fm_g.nCells + ci_g
    order = np.argsort(key_g)
    ok_r = ok_u = ok_j = True
    ncell = nblk = 0
    for z in allr:
        ok_r &= bool(np.array_equal(z["rhs"], rhs_g[z["gid"]]))
        ok_u &= bool(np.array_equal(z["uc"], uc_g[z["gid"]]))
        k = order[np.searchsorted(key_g[order], z["jrow"]*fm_g.nCells + z["jcol"])]
        ok_j &= bool(np.array_equal(z["jac"], jac_g[k]))
        ncell += len(z["gid"]); nblk += len(k)
    complete = (ncell == fm_g.nCells) and (nblk == len(ci_g))
    return {"ok": bool(ok_r and ok_u and ok_j and complete), "mesh": "32x32x%d jittered hexahedra, RCB into %d partitions, "
            "2 overlap layers" % (16*world, world), "cells": int(ncell), "jacobian_blocks": int(nblk),
            "rhs_bitwise_equal_to_single_domain": ok_r, "update_coeff_bitwise_equal": ok_u,
            "jacobian_rows_bitwise_equal": ok_j, "every_cell_and_block_covered_once": bool(complete),
            "sum_rhs_sq": float((rhs_g*rhs_g).sum())}


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


def bind_to_gpu_numa_node(local_rank):
    """One rank per GPU: run (and first-touch the pinned staging buffers) on the host cores next to that GPU.  torchrun does
    not bind its workers; with 8 ranks moving 0.94 GB per step each, remote-socket pinned memory costs the e2e figure."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        words = (os.cpu_count() + 63)//64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = [64*w + b for w in range(words) for b in range(64) if (mask[w] >> b) & 1]
        cpus = [c for c in cpus if c in os.sched_getaffinity(0)]
        if cpus:
            os.sched_setaffinity(0, cpus)
    except Exception:
        pass


class CpuArm:
    """The oracle (CPU restatement of the reference path, `kind: port`) on `threads` host threads, the way the MPI path
    runs: one rank per core, each on its own partition of THE SAME n^3 mesh the GPU arm evaluates (a z-slab of
    n x n x layers cells plus the two overlap layers per interior side, cut from the global box node for node), full
    RHS + FD Jacobian + pseudo-steady term per evaluation.  `layers` bounds the sample: every thread gets the first
    `layers` cell layers of its share n/threads of the box (all of it when the share is not larger)."""

    def __init__(self, n, threads, layers):
        import cases
        from coolfluid_b200 import lib as cl, meshgen
        from oracle import binding as ob
        self.ob, self.cfg, self.bcs = ob, cases.ns3d_cfg(), cases.bcs_ns3d()
        self.n, self.threads = n, threads
        share = max(1, n//threads)
        self.layers = layers = share if not layers else max(1, min(layers, share))
        L = cl.load()
        ob.lib(True)
        q = 5
        self.items = []
        for t in range(threads):
            k_first = t*share
            lo = 2 if k_first > 0 else 0
            hi = 2 if k_first + layers < n else 0
            em = meshgen.hex_box(n, n, layers + lo + hi, k0=k_first - lo, nz_total=n)
            fm = cl.build_mesh(em, L)                            # host mesh-data builder (setup, not timed)
            k_of = np.arange(fm.nCells)//(n*n)
            fm.updatable = np.ascontiguousarray(((k_of >= lo) & (k_of < lo + layers)).astype(np.uint8))
            U = cases.init_state(fm, q, True)
            past = np.concatenate([U[:fm.nCells*q]*(1 + 1e-3), np.zeros(fm.nGhosts*q)])
            self.items.append((fm, U, past, ob.alloc_outputs(self.cfg, fm, True)))
        self.cells = n*n*layers*threads
        self.sample = ("%d host threads, one oracle instance each on its own z-slab (%d of the %d cell layers of its share, "
                       "+ 2 overlap layers per interior side) of the same %d^3 mesh: %d updatable cells per evaluation"
                       % (threads, layers, share, n, self.cells))

    def run(self, evals):
        """-> (updatable cells per second over all threads, seconds)"""
        ob, cfg, bcs = self.ob, self.cfg, self.bcs

        def work(i):
            fm, U, past, out = self.items[i]
            for _ in range(evals):
                out.updateCoeff[:] = 0.0
                ob.compute(cfg, fm, bcs, U, jacobian=True, omp=True, out=out)
                ob.time_rhs(cfg, fm, bcs, U, past, out, omp=True)
        th = [threading.Thread(target=work, args=(i,)) for i in range(self.threads)]
        t0 = time.perf_counter()
        for t in th:
            t.start()
        for t in th:
            t.join()
        dt = time.perf_counter() - t0
        return self.cells*evals/dt, dt


def workload_text(workload, n, n_upd):
    return WORKLOADS[workload][0] % (n, n_upd)


def run_reference(args, rank):
    """`--impl reference`: the reference's CPU path for the same workload on the box's host cores.  COOLFluiD itself cannot
    be built on this image (no Boost / MPI / PETSc: DESIGN.md), so this is the oracle port, one instance per host thread
    like one MPI rank per core; each step is one evaluation of a bounded sample of the workload (CpuArm)."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    arm = CpuArm(args.n, threads, args.cpu_layers)
    vals = []
    for _ in range(args.warmup):
        arm.run(1)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        v, dt = arm.run(1)
        vals.append(v)
    total = time.perf_counter() - t0
    v = statistics.median(vals) if vals else 0.0
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3*total/max(args.steps, 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_text("ns3d", args.n, args.n**3)},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": arm.sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "the reference itself (Boost/MPI/PETSc) cannot be built on this image; this is the oracle port "
                "(-O3, no FP contraction), one instance per host thread"}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cfb200")
    # (--edge is the spelling to use under torchrun, whose own parser takes a bare --n for an abbreviation of --nnodes)
    ap.add_argument("--n", "--edge", dest="n", type=int, default=None, help="hex box edge per GPU (default 220: 220^3 = 10.6 M cells, "
                    "configs[2]; ns3d_tet: 100, i.e. 6 M tetrahedra, whose ~60-entry stencils keep the 32-bit tables in range)")
    ap.add_argument("--workload", default="ns3d", choices=sorted(WORKLOADS), help="ns3d = configs[2] (the metric), euler2d_* = configs[1], mhd3d = configs[3]")
    ap.add_argument("--cpu-layers", type=int, default=8, help="cell layers (of n x n cells) per host thread of the CPU arm: the "
                    "bounded sample of each thread's share of the n^3 mesh (0 = the whole share)")
    ap.add_argument("--cpu-evals", type=int, default=2, help="timed evaluations of the cpu_baseline leg (after one warm-up)")
    ap.add_argument("--no-fuse", action="store_true", help="pseudo-steady time term as a separate kernel (cfb200_time_rhs) "
                    "instead of inside the residual / diagonal gathers (measured equal on B200: 6.31 vs 6.33 ms at 220^3)")
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the cross-rank parity mini-problem")
    ap.add_argument("--strong", action="store_true", help="N > 1: strong scaling, ONE n^3 box cut into N slabs of n/N layers "
                    "(n must be a multiple of N) instead of one n^3 slab per GPU")
    ap.add_argument("--rhs-only", action="store_true", help="explicit configuration: ComputeRHS without the Jacobian (C5 sweep)")
    ap.add_argument("--shuffle", action="store_true", help="renumber the cells of the mesh at random first (no locality)")
    ap.add_argument("--renumber", action="store_true", help="cfg.renumber: space-filling-curve storage order inside the engine")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.n is None:
        args.n = 100 if args.workload == "ns3d_tet" else 220
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
    bind_to_gpu_numa_node(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_
        dist = dist_
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    wl_desc, wl_pvt, a_rhs, a_jac = WORKLOADS[args.workload]
    t0 = time.perf_counter()
    parity = None
    if world > 1 and not args.no_parity:
        parity = multi_gpu_parity(rank, world, local_rank, L, dist, torch)
    em, cfg, bcs = flagship(args.n if world == 1 else 2, args.workload)
    cfg.fuseTimeRhs = 0 if args.no_fuse else 1
    cfg.renumber = 1 if args.renumber else 0
    if args.rhs_only:
        cfg.computeJacobian = 0
        cfg.fuseTimeRhs = 0
        a_jac = 0
    if args.shuffle and world == 1:
        from coolfluid_b200 import meshgen as _mg
        em, _ = _mg.shuffle_cells(em)
    q = cfg.nEqs
    part = None
    if world > 1:
        if args.strong:
            if args.n % world:
                raise SystemExit("--strong needs an edge that is a multiple of the number of GPUs")
            em, part = slab_partition(args.n, rank, world, args.workload, box=(args.n, args.n, args.n//world))
        else:
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
        if args.no_fuse and not args.rhs_only:
            eng.time_rhs(download=False)

    def step_e2e():
        eng.set_states(hU.numpy())
        eng.set_update_coeff(None)
        if part is not None:
            eng.halo_begin()
        eng.compute_rhs(apply_bc=True, download=True, out_rhs=hR.numpy(), out_uc=hC.numpy())
        if args.no_fuse and not args.rhs_only:
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
        launches[0] += len(tl)              # every kernel of the step is in the list (the updateCoeff memset is not a kernel)
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
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if (args.strong and world > 1) else "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": workload_text(args.workload, args.n, n_upd) + (" -- RHS ONLY (explicit: no Jacobian, no time "
                                   "term)" if args.rhs_only else ""),
                       "cells_per_gpu": nc, "faces_per_gpu": int(fm.nFaces), "bsr_blocks_per_gpu": int(eng.nnzb),
                       "l2": "inputs (%.1f GB BSR + %.2f GB states/gradients) far larger than the 126 MB L2; no flush needed"
                             % (eng.nnzb*q*q*8/1e9, nc*q*8*6/1e9),
                       "partitioning": ("single partition" if part is None else
                                        ("STRONG scaling: " if args.strong else "") +
                                        "one global %d x %d x %d box cut along z, one slab per GPU (%d updatable + %d overlap "
                                        "cells on rank 0, order-preserving local numbering), state halo exchange over NCCL "
                                        "every step (%d cells x %d doubles per neighbour), overlapped with the independent "
                                        "gradients" % (args.n, args.n, args.n if args.strong else args.n*world, n_upd, nc - n_upd, 2*args.n*args.n, q)),
                       "time_term": "separate kernel" if args.no_fuse else "fused into the residual / diagonal gathers",
                       "numbering": ("cells renumbered at random" if args.shuffle else "generator order (x fastest)") +
                                    (", Morton storage order inside the engine (cfg.renumber)" if args.renumber else ""),
                       "mesh_build_s": round(t_mesh, 2), "device_setup_s": round(t_setup, 2)},
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": int(nc*q*8*world), "d2h_bytes_per_step": int((nc*q*8 + nc*8)*world)},
            "gpu_launches": int(launches[0]),
            "kernels_ms": {k: round(v, 4) for k, v in sorted(kavg.items(), key=lambda kv: -kv[1])},
            "roofline": {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved/peak, "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": int(algo), "avg_launch_ms": kavg[top]},
            # the same kernel against the pipe that actually bounds it: FP64 instructions per launch / time / measured peak
            **({"roofline_fp64": {
                "kernel": top, "bound": "fp64 issue (no FMA: reference operation order)",
                "achieved": fm.nInner*INNER_FACE_JAC_FP64_PER_FACE/32./(kavg[top]*1e-3)/1e9,
                "peak": FP64_PEAK_WARP_INSTR/1e9, "unit": "G warp-instr/s",
                "frac": fm.nInner*INNER_FACE_JAC_FP64_PER_FACE/32./(kavg[top]*1e-3)/FP64_PEAK_WARP_INSTR,
                "instr_source": "profiles/r02_ncu_innerface_pipe1_source_lines.md",
                "peak_source": "profiles/r02_fp64_microbench.txt"}}
               if (top == "innerFaceJac" and args.workload == "ns3d") else {}),
            "step_roofline": {"bytes_per_cell": a_rhs + a_jac, "achieved_gbs": (a_rhs + a_jac)*value/world/1e9,
                              "frac_of_hbm_peak": (a_rhs + a_jac)*value/world/1e9/peak},
            "clocks": clocks, "status": int(status),
        }
        if parity is not None:
            out["parity"] = parity
        if world == 1 and not args.no_cpu_baseline and args.workload == "ns3d":
            threads = os.cpu_count() or 1
            arm = CpuArm(args.n, threads, args.cpu_layers)
            arm.run(1)
            v, dt = arm.run(args.cpu_evals)
            out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                   "sample": arm.sample + "; %d evaluations, %.1f s" % (args.cpu_evals, dt)}
        print(json.dumps(out))
    eng.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
