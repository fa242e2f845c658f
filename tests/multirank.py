"""Worker used by the N>1 tests: one process per rank (spawned by the tests; `gloo` on the CPU tier, NCCL on the GPU
box).  Each rank extracts its partition of a shared case, runs K ForwardEuler iterations
(applyBC -> ComputeRHS -> UpdateSol -> state halo exchange) and writes its updatable states to <out>/rank<r>.npz.

  CPU tier  : engine = tests/hostsim (TEST ONLY host build of the kernel functors); the halo travels through
              torch.distributed (gloo) with the same send / receive lists the NCCL path uses.
  GPU tier  : engine = libcfb200.so on cuda:<rank>; the halo travels through cfb200_halo_begin/exchange (NCCL) and the
              residual norm through cfb200_allreduce_sum.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def run(rank, world, mode, name, steps, out_dir, cfl=0.5, overlap_mode=True):
    import torch.distributed as dist
    import cases
    from coolfluid_b200 import partition, config as B, lib as cl
    from oracle import binding as oracle

    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mesh_f, cfg_f, bcs_f, pvt = cases.CASES[name]
        em = mesh_f()
        part = partition.rcb_partition(em, world)
        p = partition.extract(em, part, rank, world)
        cfg = cfg_f()
        cfg.computeJacobian = 0
        renumber = int(os.environ.get("CFB_TEST_RENUMBER", "0"))     # storage renumbering inside every rank's engine
        cfg.renumber = renumber
        q = cfg.nEqs
        if mode == "hostsim":
            L = cl.load(os.path.join(ROOT, "tests", "hostsim", "libhostsim.so"), "hostsim_")
            build = oracle.build_mesh
            dev = 0
        else:
            L = cl.load()
            build = lambda m: cl.build_mesh(m, L)
            dev = rank
        fm = build(p.em)
        fm.updatable = np.ascontiguousarray(p.updatable)
        bcs = bcs_f()
        if p.em.trs_names[-1] == "PartitionFaces":
            bcs = bcs + [B.make_bc(B.BC_SUPER_OUTLET)]
        # the global initial state, restricted (every rank builds the same global field)
        fm_g = oracle.build_mesh(em)
        U_g = cases.init_state(fm_g, q, pvt)[:fm_g.nCells*q].reshape(-1, q)
        e = cl.Engine(cfg, dev, lib=L)
        e.set_mesh(fm, bcs)
        e.set_states(U_g[p.cell_l2g].reshape(-1))
        if mode != "hostsim":
            ids = [cl.nccl_unique_id(L) if rank == 0 else None]
            dist.broadcast_object_list(ids, src=0)
            e.comm_init(rank, world, ids[0])
            e.set_halo(p)

        def exchange_gloo():
            import torch
            U = e.get_states().reshape(-1, q)
            reqs, bufs = [], {}
            for peer, ids_ in sorted(p.send.items()):
                reqs.append(dist.isend(torch.from_numpy(np.ascontiguousarray(U[ids_])), peer))
            for peer, ids_ in sorted(p.recv.items()):
                bufs[peer] = torch.empty((len(ids_), q), dtype=torch.float64)
                reqs.append(dist.irecv(bufs[peer], peer))
            for r in reqs:
                r.wait()
            for peer, ids_ in p.recv.items():
                U[ids_] = bufs[peer].numpy()
            e.set_states(U.reshape(-1))

        hist = []
        for it in range(steps):
            e.compute_rhs(apply_bc=True, download=False)
            res = e.update_sol_fwd_euler(cfl, 0)
            local = 10.0**(2.0*res) if res > -1e300 else 0.0
            if mode == "hostsim":
                import torch
                t = torch.tensor([local], dtype=torch.float64)
                dist.all_reduce(t)
                tot = float(t.item())
                exchange_gloo()
            else:
                tot = float(e.allreduce_sum([local])[0])
                if overlap_mode and it % 2 == 0:
                    e.halo_begin()          # the next compute_rhs overlaps it with the interior gradients
                else:
                    e.halo_exchange()
            hist.append(0.5*np.log10(tot) if tot > 0 else -np.inf)
        U = e.get_states().reshape(-1, q)
        # one implicit evaluation on top (overlap states in place): residual + block Jacobian + fused pseudo-steady term
        jc = cfg_f()
        jc.fuseTimeRhs = 1
        jc.cfl = 3.0
        jc.renumber = renumber
        e.set_step(jc)
        e.backup_past_states()
        e.set_update_coeff(None)
        if mode != "hostsim":
            e.halo_begin()              # joined inside compute_rhs (overlapped with the independent gradients)
        rhs, uc = e.compute_rhs()
        jac = e.bsr_values().reshape(-1, q*q)
        rp, ci = e.bsr_pattern()
        upd = p.updatable == 1
        rows = np.nonzero(upd)[0]
        blk_row = np.repeat(np.arange(fm.nCells), np.diff(rp))
        keep = upd[blk_row]
        np.savez(os.path.join(out_dir, "rank%d.npz" % rank), U=U[upd], gid=p.cell_l2g[upd],
                 hist=np.array(hist), Uall=U, gall=p.cell_l2g, rhs=rhs.reshape(-1, q)[upd], uc=uc[upd],
                 jac=jac[keep], jrow=p.cell_l2g[blk_row[keep]], jcol=p.cell_l2g[ci[keep]])
        e.close()
    finally:
        dist.barrier()
        dist.destroy_process_group()


def single_domain(name, steps, lib, cfl=0.5, device=0):
    """The same iterations on the whole mesh with one engine: (states [nCells, q], residual history)."""
    import cases
    from coolfluid_b200 import lib as cl
    from oracle import binding as oracle
    mesh_f, cfg_f, bcs_f, pvt = cases.CASES[name]
    em = mesh_f()
    cfg = cfg_f()
    cfg.computeJacobian = 0
    q = cfg.nEqs
    fm = oracle.build_mesh(em)
    U0 = cases.init_state(fm, q, pvt)[:fm.nCells*q]
    e = cl.Engine(cfg, device, lib=lib)
    e.set_mesh(fm, bcs_f())
    e.set_states(U0)
    hist = []
    for _ in range(steps):
        e.compute_rhs(apply_bc=True, download=False)
        hist.append(e.update_sol_fwd_euler(cfl, 0))
    U = e.get_states().reshape(-1, q)
    jc = cfg_f()
    jc.fuseTimeRhs = 1
    jc.cfl = 3.0
    e.set_step(jc)
    e.backup_past_states()
    e.set_update_coeff(None)
    rhs, uc = e.compute_rhs()
    jac = e.bsr_values().reshape(-1, q*q)
    rp, ci = e.bsr_pattern()
    e.close()
    extra = dict(rhs=rhs.reshape(-1, q), uc=uc, jac=jac, rowPtr=rp, colInd=ci)
    return U, np.array(hist), extra


if __name__ == "__main__":
    # python multirank.py <mode> <case> <steps> <out_dir>   (RANK / WORLD_SIZE / MASTER_* from the environment)
    run(int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), sys.argv[1], sys.argv[2], int(sys.argv[3]), sys.argv[4])
