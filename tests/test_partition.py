"""CPU tier: the harness partitioner (coolfluid_b200/partition.py) and the partition semantics of the engine
(SURVEY.md §8e): every rank holds its updatable cells + 2 overlap layers, skips faces with no updatable side and fills
only updatable rows; with the states of the overlap cells in place, the residual and the Jacobian rows of the updatable
cells equal the single-domain ones -- BIT FOR BIT with the default order-preserving local numbering (ascending global
ID), to summation order with the grouped numbering -- and the engine is bit-exact against the oracle run on the same
local mesh.  The N>1 exchange itself is covered by test_halo_gloo.py (world_size 2,
gloo) here and by the NCCL path in test_gpu_multi.py on the GPU box."""
import numpy as np
import pytest

import cases
from coolfluid_b200 import partition, config as B, lib as cl


def local_problem(name, nparts, rank, oracle, order="global"):
    mesh_f, cfg_f, bcs_f, pvt = cases.CASES[name]
    em = mesh_f()
    part = partition.rcb_partition(em, nparts)
    p = partition.extract(em, part, rank, nparts, order=order)
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
      for order in ("global", "grouped"):
        ps = [partition.extract(em, part, r, n, order=order) for r in range(n)]
        owned = np.concatenate([p.cell_l2g[p.updatable == 1] for p in ps])
        assert sorted(owned) == list(range(em.n_cells))                      # every cell updatable on exactly one rank
        for p in ps:
            upd = p.updatable == 1
            assert int(upd.sum()) == p.n_updatable
            assert (part[p.cell_l2g[upd]] == p.rank).all()
            assert (part[p.cell_l2g[~upd]] != p.rank).all()
            if order == "global":
                assert (np.diff(p.cell_l2g) > 0).all()                        # local numbering preserves the global order
                assert len(p.indep) == p.n_interior and upd[p.indep].all()
            else:
                assert upd[:p.n_updatable].all() and (np.diff(p.cell_l2g[:p.n_interior]) > 0).all()
            for peer, ids in p.recv.items():
                q = ps[peer]
                assert np.array_equal(p.cell_l2g[ids], q.cell_l2g[q.send[p.rank]])    # packed orders agree
                assert (~upd[ids]).all() and (q.updatable[q.send[p.rank]] == 1).all()
            recv_all = np.sort(np.concatenate(list(p.recv.values()))) if p.recv else np.zeros(0)
            assert np.array_equal(recv_all, np.nonzero(~upd)[0])              # every overlap cell is received
            sp, sc, rp, rc = p.halo_csr()
            assert sp[-1] == len(sc) and rp[-1] == len(rc) and len(sp) == n + 1


@pytest.mark.parametrize("name,nparts,order", [("ns3d_hex", 2, "global"), ("euler2d_2nd_tri_fullstencil", 3, "global"),
                                               ("ns3d_tet", 2, "global"), ("ns2d_quad", 2, "global"),
                                               ("mhd3d_hex_hll", 2, "global"), ("ns3d_hex_corrected", 3, "global"),
                                               ("ns3d_hex_roesa_fix2", 2, "global"), ("ns3d_tet_roesa", 3, "global"),
                                               ("ns3d_hex_correctedgg", 2, "global"), ("ns2d_quad_statediff", 2, "global"),
                                               ("ns3d_hex_ausmplusup_analytic", 2, "global"),
                                               ("mhd3d_hex_hll_dipole", 2, "global"),
                                               ("mhd3d_hex_hll_outlet_coronal", 2, "global"),      # + MHDConsACAST source term
                                               ("ns3d_farfield_char_sub", 2, "global"),
                                               ("ns3d_hex_facevertex", 2, "global"),               # Setup1.stencil = FaceVertex
                                               ("euler2d_2nd_quad_facevertex", 3, "global"),
                                               ("ns3d_tet_faceedge", 2, "global"), ("euler2d_2nd_quad_face", 2, "global"),
                                               ("ns3d_hex", 2, "grouped"), ("ns3d_tet", 2, "grouped")])
def test_partition_matches_single_domain(name, nparts, order, oracle, hostsim_lib):
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
        _, _, p, fm, cfg_l, bcs, _ = local_problem(name, nparts, rank, oracle, order)
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
        # updatable cells == single domain: bit for bit with the order-preserving numbering, to summation order otherwise
        upd = p.updatable == 1
        rl = rhs.reshape(-1, q)
        if order == "global":
            assert np.array_equal(rl[upd], rhs_g[p.cell_l2g[upd]])
            assert np.array_equal(uc[upd], r_g.updateCoeff[p.cell_l2g[upd]])
        else:
            assert (np.abs(rl[upd] - rhs_g[p.cell_l2g[upd]]) <= 1e-12*scale_r).all()
            assert np.allclose(uc[upd], r_g.updateCoeff[p.cell_l2g[upd]], rtol=1e-12, atol=0)
        jl = jac.reshape(-1, q*q)
        for c in range(fm.nCells):
            for k in range(fm.bsrRowPtr[c], fm.bsrRowPtr[c+1]):
                if not upd[c]:
                    assert not jl[k].any()                                   # rows of non-updatable cells stay empty
                    continue
                gk = gslot[(int(p.cell_l2g[c]), int(p.cell_l2g[fm.bsrColInd[k]]))]
                if order == "global":
                    assert np.array_equal(jl[k], jac_g[gk]), (c, k)
                else:
                    s = np.abs(jac_g[gk]).max()
                    # FD entries amplify the 1e-16 summation-order differences of gradients / nodal states by 1/eps ~ 1e6
                    assert np.abs(jl[k] - jac_g[gk]).max() <= 1e-8*s + 1e-300, (c, k)


def test_bench_slabs_are_pieces_of_one_global_mesh(oracle, hostsim_lib):
    """bench.py --gpus N: the slabs the ranks generate independently are, node for node, pieces of ONE global box (hash
    jitter of the global node index); send lists pair cell for cell with the neighbour's overlap lists; the independent
    cells see no overlap cell; and every rank's residual / update coefficient / Jacobian rows of its updatable cells are
    bit-identical to the single-domain evaluation of the global mesh (checked with the host build of the kernels)."""
    import bench
    from coolfluid_b200 import meshgen
    world, box = 3, (5, 4, 4)
    nx, ny, nzr = box
    em_g = meshgen.hex_box(nx, ny, nzr*world, lengths=(1.0, 1.0*ny/nx, 1.0*nzr*world/nx), nz_total=nzr*world)
    fm_g = oracle.build_mesh(em_g)
    cfg = cases.ns3d_cfg(fuse_time_rhs=1, cfl=2.0)
    q = 5
    bcs = cases.bcs_ns3d()
    Ug = cases.init_state(fm_g, q, True)[:fm_g.nCells*q].reshape(-1, q)

    def evaluate(fm, U):
        e = cl.Engine(cfg, 0, lib=hostsim_lib)
        try:
            e.set_mesh(fm, bcs)
            e.set_states(U.reshape(-1))
            e.set_past_states(U.reshape(-1)*(1 + 1e-3))
            rhs, uc = e.compute_rhs()
            return rhs.reshape(-1, q), uc, e.bsr_values().reshape(-1, q*q), e.bsr_pattern()
        finally:
            e.close()
    rhs_g, uc_g, jac_g, (rp_g, ci_g) = evaluate(fm_g, Ug)
    gslot = {(c, int(ci_g[k])): k for c in range(fm_g.nCells) for k in range(rp_g[c], rp_g[c+1])}
    parts = []
    for r in range(world):
        em, p = bench.slab_partition(0, r, world, "ns3d", box=box)
        fm = oracle.build_mesh(em)
        parts.append(p)
        assert np.array_equal(fm.centres.reshape(-1, 3)[:fm.nCells], fm_g.centres.reshape(-1, 3)[p.cell_l2g])
        upd = p.updatable == 1
        assert p.n_updatable == nx*ny*nzr == int(upd.sum())
        sp, st = fm.stencilPtr, fm.stencil
        for c in p.indep:                                   # no stencil entry of an independent cell is an overlap cell
            nb = st[sp[c]:sp[c+1]]
            assert upd[nb[nb < fm.nCells]].all()
        fm.updatable = np.ascontiguousarray(p.updatable)
        rhs, uc, jac, (rp, ci) = evaluate(fm, Ug[p.cell_l2g])
        assert np.array_equal(rhs[upd], rhs_g[p.cell_l2g[upd]]) and np.array_equal(uc[upd], uc_g[p.cell_l2g[upd]])
        for c in np.nonzero(upd)[0]:
            for k in range(rp[c], rp[c+1]):
                assert np.array_equal(jac[k], jac_g[gslot[(int(p.cell_l2g[c]), int(p.cell_l2g[ci[k]]))]])
    for r in range(world):
        p = parts[r]
        for peer, ids in p.send.items():
            qq = parts[peer]
            assert np.array_equal(p.cell_l2g[ids], qq.cell_l2g[qq.recv[r]])          # same global cells, same order
            assert (p.updatable[ids] == 1).all() and (qq.updatable[qq.recv[r]] == 0).all()
        recv_all = np.sort(np.concatenate(list(p.recv.values())))
        assert np.array_equal(recv_all, np.nonzero(p.updatable == 0)[0])              # every overlap cell is received
