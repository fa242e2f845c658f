"""CPU tier: the harness partitioner (coolfluid_b200/partition.py) and the partition semantics of the engine
(SURVEY.md §8e): every rank holds its updatable cells + 2 overlap layers, skips faces with no updatable side and fills
only updatable rows; with the states of the overlap cells in place, the residual and the Jacobian rows of the updatable
cells equal the single-domain ones (to summation order: the local face numbering differs), and the engine is bit-exact
against the oracle run on the same local mesh.  The N>1 exchange itself is covered by test_halo_gloo.py (world_size 2,
gloo) here and by the NCCL path in test_gpu_multi.py on the GPU box."""
import numpy as np
import pytest

import cases
from coolfluid_b200 import partition, config as B, lib as cl


def local_problem(name, nparts, rank, oracle):
    mesh_f, cfg_f, bcs_f, pvt = cases.CASES[name]
    em = mesh_f()
    part = partition.rcb_partition(em, nparts)
    p = partition.extract(em, part, rank, nparts)
    fm = oracle.build_mesh(p.em)
    fm.updatable = np.ascontiguousarray(p.updatable)
    bcs = bcs_f()
    if p.em.trs_names[-1] == "PartitionFaces":
        bcs = bcs + [B.make_bc(B.BC_SUPER_OUTLET)]       # never evaluated: the inner cell is not updatable
    return em, part, p, fm, cfg_f(), bcs, pvt


def test_partition_structure():
    from coolfluid_b200 import meshgen
    for em, n in ((meshgen.hex_box(7, 6, 5), 3), (meshgen.quad_channel(11, 8, triangles=True), 4), (meshgen.tet_box(3, 3, 3), 2)):
        part = partition.rcb_partition(em, n)
        assert sorted(np.unique(part)) == list(range(n))
        ps = [partition.extract(em, part, r, n) for r in range(n)]
        owned = np.concatenate([p.cell_l2g[:p.n_updatable] for p in ps])
        assert sorted(owned) == list(range(em.n_cells))                      # every cell updatable on exactly one rank
        for p in ps:
            assert (part[p.cell_l2g[:p.n_updatable]] == p.rank).all()
            assert (part[p.cell_l2g[p.n_updatable:]] != p.rank).all()
            assert (np.diff(p.cell_l2g[:p.n_interior]) > 0).all()
            for peer, ids in p.recv.items():
                q = ps[peer]
                assert np.array_equal(p.cell_l2g[ids], q.cell_l2g[q.send[p.rank]])    # packed orders agree
                assert (ids >= p.n_updatable).all() and (q.send[p.rank] < q.n_updatable).all()
            recv_all = np.sort(np.concatenate(list(p.recv.values()))) if p.recv else np.zeros(0)
            assert np.array_equal(recv_all, np.arange(p.n_updatable, p.em.n_cells))   # every overlap cell is received
            sp, sc, rp, rc = p.halo_csr()
            assert sp[-1] == len(sc) and rp[-1] == len(rc) and len(sp) == n + 1


@pytest.mark.parametrize("name,nparts", [("ns3d_hex", 2), ("euler2d_2nd_tri_fullstencil", 3), ("ns3d_tet", 2),
                                         ("ns2d_quad", 2)])
def test_partition_matches_single_domain(name, nparts, oracle, hostsim_lib):
    mesh_f, cfg_f, bcs_f, pvt = cases.CASES[name]
    em = mesh_f()
    fm_g = oracle.build_mesh(em)
    cfg = cfg_f()
    q = cfg.nEqs
    U_g = cases.init_state(fm_g, q, pvt)
    r_g = oracle.compute(cfg, fm_g, bcs_f(), U_g.copy(), jacobian=True)
    assert r_g.status == 0
    Ug = U_g[:fm_g.nCells*q].reshape(-1, q)
    rhs_g = r_g.rhs.reshape(-1, q)
    jac_g = r_g.bsrVal.reshape(-1, q*q)
    gslot = {}
    for c in range(fm_g.nCells):
        for k in range(fm_g.bsrRowPtr[c], fm_g.bsrRowPtr[c+1]):
            gslot[(c, int(fm_g.bsrColInd[k]))] = k
    scale_r = np.abs(rhs_g).max(axis=0)
    for rank in range(nparts):
        _, _, p, fm, cfg_l, bcs, _ = local_problem(name, nparts, rank, oracle)
        U = np.zeros((fm.nCells + fm.nGhosts, q))
        U[:fm.nCells] = Ug[p.cell_l2g]
        U = U.reshape(-1)
        r = oracle.compute(cfg_l, fm, bcs, U.copy(), jacobian=True)
        assert r.status == 0
        e = cl.Engine(cfg_l, 0, lib=hostsim_lib)
        try:
            e.set_mesh(fm, bcs)
            e.set_states(U[:fm.nCells*q])
            rhs, uc = e.compute_rhs()
            jac = e.bsr_values()
        finally:
            e.close()
        # engine == oracle on the same local mesh, bit for bit
        assert np.array_equal(rhs, r.rhs) and np.array_equal(uc, r.updateCoeff) and np.array_equal(jac, r.bsrVal)
        # updatable cells == single domain (summation order differs)
        nu = p.n_updatable
        rl = rhs.reshape(-1, q)
        assert (np.abs(rl[:nu] - rhs_g[p.cell_l2g[:nu]]) <= 1e-12*scale_r).all()
        assert np.allclose(uc[:nu], r_g.updateCoeff[p.cell_l2g[:nu]], rtol=1e-12, atol=0)
        jl = jac.reshape(-1, q*q)
        for c in range(fm.nCells):
            for k in range(fm.bsrRowPtr[c], fm.bsrRowPtr[c+1]):
                if c >= nu:
                    assert not jl[k].any()                                   # rows of non-updatable cells stay empty
                    continue
                gk = gslot[(int(p.cell_l2g[c]), int(p.cell_l2g[fm.bsrColInd[k]]))]
                s = np.abs(jac_g[gk]).max()
                # FD entries amplify the 1e-16 summation-order differences of gradients / nodal states by 1/eps ~ 1e6
                assert np.abs(jl[k] - jac_g[gk]).max() <= 1e-8*s + 1e-300, (c, k)


def test_bench_slab_partition_is_consistent(product_lib):
    """bench.py --gpus N: every rank's send list pairs, cell for cell (same (i, j, k) order), with the neighbour's
    overlap list; the leading n_interior cells see no overlap cell; own cells are updatable, overlap cells are not."""
    import bench
    from coolfluid_b200 import lib as cl
    n, world = 5, 3
    parts = []
    for r in range(world):
        em, p = bench.slab_partition(n, r, world, "ns3d")
        fm = cl.build_mesh(em, product_lib)
        x = fm.centres.reshape(-1, 3)[:fm.nCells]
        lo = 2 if r > 0 else 0
        parts.append((p, x, lo))
        assert p.n_updatable == n**3 and int(p.updatable.sum()) == n**3
        assert p.updatable[:p.n_updatable].all() and not p.updatable[p.n_updatable:].any()
        # interior cells: no stencil entry is an overlap cell
        sp, st = fm.stencilPtr, fm.stencil
        for c in range(p.n_interior):
            nb = st[sp[c]:sp[c+1]]
            nb = nb[nb < fm.nCells]
            assert p.updatable[nb].all()
    h = 1.0/n
    for r in range(world):
        p, x, lo = parts[r]
        for peer, ids in p.send.items():
            q, xq, loq = parts[peer]
            rid = q.recv[r]
            assert len(ids) == len(rid) == 2*n*n
            # same cell of the global box: x, y match to the jitter, z matches after removing each slab's own offset
            zs = x[ids, 2] - lo*h + r*n*h
            zr = xq[rid, 2] - loq*h + peer*n*h
            assert np.abs(zs - zr).max() < 0.5*h
            assert np.abs(x[ids, :2] - xq[rid, :2]).max() < 0.5*h
