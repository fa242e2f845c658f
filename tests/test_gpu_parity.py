"""GPU tier (-m gpu): the CUDA path through the C ABI (libcfb200.so on cuda:0) against the oracle on the same
seeded inputs.  Tolerance: BASELINE.json asks for 1e-12 relative per-cell residual and Jacobian entry; the kernels
are built without FMA contraction and with the reference's operation order, so the bar here is BIT-EXACT
(np.array_equal) for every field of every case, from the 1-cell meshes up to 2.1 M hexahedra / 1.05 M tetrahedra."""
import numpy as np
import pytest

import cases
from coolfluid_b200 import lib as cl, meshgen, config as B

pytestmark = pytest.mark.gpu
TOL = 1e-13


def _factory(product_lib):
    return lambda cfg: cl.Engine(cfg, 0, lib=product_lib)


def _check(out):
    """Every quantity bit-identical to the oracle's (the message reports the per-row relative error when it is not)."""
    for k, ab in out.items():
        if k.startswith("_") or ab[1] is None:
            continue
        a, b = ab
        assert a.shape == b.shape, k
        if not np.array_equal(a, b):
            width = out["_fm"].dim + 2 if k in ("rhs", "time_rhs") else 1
            raise AssertionError("%s differs from the oracle: per-row rel err %.3e, %d of %d entries"
                                 % (k, cases.relerr_rows(a, b, width), int((a != b).sum()), a.size))
    assert np.abs(out["rhs"][1]).max() > 0.0


@pytest.mark.parametrize("name", sorted(cases.CASES))
def test_cuda_matches_oracle(name, oracle, product_lib):
    _check(cases.run_case(name, oracle, _factory(product_lib)))


def test_cuda_bit_exact_on_flagship_case(oracle, product_lib):
    """3D NS hex (BASELINE config 3) is expected to match the oracle bit for bit (no FMA, IEEE div/sqrt)."""
    out = cases.run_case("ns3d_hex", oracle, _factory(product_lib))
    for k in ("rhs", "updateCoeff", "jacobian", "time_jacobian"):
        assert np.array_equal(out[k][0], out[k][1]), k


def test_partition_overlap_rows(oracle, product_lib):
    def upd(fm):
        x = fm.centres.reshape(-1, fm.dim)[:fm.nCells, 0]
        return (x < 0.6).astype(np.uint8)
    for name in ("ns3d_hex", "euler2d_2nd_tri_fullstencil"):
        out = cases.run_case(name, oracle, _factory(product_lib), updatable=upd)
        _check(out)
        fm = out["_fm"]
        jac = out["jacobian"][0].reshape(len(fm.bsrColInd), -1)
        for c in np.nonzero(fm.updatable == 0)[0]:
            assert not jac[fm.bsrRowPtr[c]:fm.bsrRowPtr[c+1]].any()


def test_historical_limiter(oracle, product_lib):
    rng = np.random.RandomState(3)

    def mut(cfg):
        cfg.residual = -3.0
    n = cases.CASES["ns3d_hex"][0]().n_cells*5
    lim0 = 0.3 + 0.7*rng.random_sample(n)
    _check(cases.run_case("ns3d_hex", oracle, _factory(product_lib), limiter_init=lim0, cfg_mut=mut))


def _medium(oracle, product_lib, em, cfg, bcs, pvt, fields=True):
    """All fields and the full block Jacobian (pseudo-steady term included) against the oracle, bit for bit; the
    per-cell / per-block relative errors BASELINE.json states (1e-12) are evaluated as well, for the failure message."""
    fm = cl.build_mesh(em, product_lib)
    q = cfg.nEqs
    U = cases.init_state(fm, q, pvt)
    Ug = U.copy()
    r = oracle.compute(cfg, fm, bcs, Ug, jacobian=True, omp=True)
    assert r.status == 0
    e = cl.Engine(cfg, 0, lib=product_lib)
    e.set_mesh(fm, bcs)
    e.set_states(U[:fm.nCells*q])
    rhs, uc = e.compute_rhs()
    jac = e.bsr_values()
    got = {}
    if fields:
        for f in cases.FIELDS:
            if not (f == "uZ" and fm.dim == 2):
                got[f] = (e.field(f), getattr(r, f))
        got["gstates"] = (e.field("gstates"), Ug[fm.nCells*q:])
    # determinism: a second evaluation is bitwise identical (no atomics, fixed gather order)
    e.set_update_coeff(None)
    rhs2, uc2 = e.compute_rhs()
    assert np.array_equal(rhs, rhs2) and np.array_equal(uc, uc2)
    assert np.array_equal(jac, e.bsr_values())
    assert cases.relerr_rows(rhs, r.rhs, q) <= 1e-12, "per-cell residual"
    assert cases.relerr_rows(jac, r.bsrVal, q*q) <= 1e-12, "per-block Jacobian"
    assert np.array_equal(rhs, r.rhs) and np.array_equal(uc, r.updateCoeff) and np.array_equal(jac, r.bsrVal)
    for k, (a, b) in got.items():
        assert np.array_equal(a, b), k
    # pseudo-steady time term on top (FVMCC_PseudoSteadyTimeRhs)
    past = U[:fm.nCells*q]*(1 + 1e-3)
    oracle.time_rhs(cfg, fm, bcs, Ug, np.concatenate([past, np.zeros(fm.nGhosts*q)]), r)
    e.set_past_states(past)
    trhs = e.time_rhs()
    assert np.array_equal(trhs, r.rhs), "time_rhs"
    assert np.array_equal(e.bsr_values(), r.bsrVal), "time_jacobian"
    e.close()


def test_medium_hex_ns(oracle, product_lib):
    _medium(oracle, product_lib, meshgen.hex_box(24, 22, 20), cases.ns3d_cfg(), cases.bcs_ns3d(), True)


def test_medium_tet_ns(oracle, product_lib):
    _medium(oracle, product_lib, meshgen.tet_box(10, 9, 8), cases.ns3d_cfg(), cases.bcs_ns3d(), True)


def test_large_hex_ns_2M_cells(oracle, product_lib):
    """128^3 = 2.1 M jittered hexahedra (above the 1 M-cell threshold of the two-level ordered sum, L2-missing
    gathers): every field, the residual and all 14.6 M Jacobian blocks bit-identical to the oracle."""
    _medium(oracle, product_lib, meshgen.hex_box(128, 128, 128), cases.ns3d_cfg(), cases.bcs_ns3d(), True)


def test_large_tet_ns_1M_cells(oracle, product_lib):
    """56^3 boxes split into 6 tetrahedra each (Kuhn) = 1.05 M cells, stencils of ~60 vertex neighbours."""
    _medium(oracle, product_lib, meshgen.tet_box(56, 56, 56), cases.ns3d_cfg(), cases.bcs_ns3d(), True)


def test_large_update_norm_two_level_sum(product_lib):
    """ForwardEuler update above 1 M cells: the two-level ordered block sum of the residual norm against numpy
    (the same squares; the sum itself is compared to 1e-13 relative) and a second identical run bit for bit."""
    fm = cl.build_mesh(meshgen.hex_box(110, 100, 100), product_lib)
    cfg = B.make_config(3, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, cfl=0.5)
    U = cases.init_state(fm, 5, False)[:fm.nCells*5]
    res = []
    for _ in range(2):
        e = cl.Engine(cfg, 0, lib=product_lib)
        e.set_mesh(fm, cases.bcs_euler3d())
        e.set_states(U)
        rhs, uc = e.compute_rhs()
        res.append(e.update_sol_fwd_euler(0.5, 0))
        e.close()
    assert res[0] == res[1]
    ref = np.log10(np.sqrt((rhs.reshape(-1, 5)[:, 0]**2).sum()))
    assert abs(res[0] - ref) <= 1e-12


def test_medium_tri_euler(oracle, product_lib):
    cfg = B.make_config(2, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, jacobian=True)
    _medium(oracle, product_lib, meshgen.quad_channel(90, 70, triangles=True), cfg, cases.bcs_euler2d(), False)


def test_forward_euler_history(oracle, product_lib):
    """BASELINE config 1/2 explicit loop, device resident (applyBC -> ComputeRHS -> UpdateSol + L2 norm) against the
    oracle's ForwardEuler driver: residual history to 1e-10 (BASELINE.json north_star)."""
    em = meshgen.quad_channel(40, 30, triangles=True)
    fm = cl.build_mesh(em, product_lib)
    cfg = B.make_config(2, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, limit_res=-1.42, cfl=0.7)
    bcs = cases.bcs_euler2d()
    U0 = cases.init_state(fm, 4, False)
    nit = 60
    ref = oracle.run_fwd_euler(B.make_config(2, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, limit_res=-1.42, cfl=0.7),
                               fm, bcs, U0.copy(), 0.7, -99.0, nit, omp=True)
    e = cl.Engine(cfg, 0, lib=product_lib)
    e.set_mesh(fm, bcs)
    e.set_states(U0[:fm.nCells*4])
    hist = []
    cfg.residual = 0.0
    for it in range(nit):
        cfg.iter = it + 1
        e.set_step(cfg)
        e.compute_rhs(apply_bc=True, download=False)
        res = e.update_sol_fwd_euler(0.7, 0)
        hist.append(res)
        cfg.residual = res
    e.close()
    assert np.abs(np.array(hist) - ref.history).max() <= 1e-10


def test_wedge_fvm_reference_case_on_gpu(product_lib):
    """The reference's own regression case (plugins/NavierStokes/testcases/Wedge/wedgeFVM.CFcase: Euler2D, Roe,
    LinearLS2D + Venktn2D, ForwardEuler CFL 0.7, stop at Norm -4.0) on its own mesh (tests/golden/wedge.thor),
    device resident: the residual history equals the oracle's committed history to 1e-10 at every iteration, the run
    stops at the same iteration, and the final value is within the reference's 0.1 % of `### Residual = -4.0017502`
    (wedgeFVM.CFcase:14)."""
    import os
    import sys
    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    sys.path.insert(0, gold)
    import make_golden
    g = np.load(os.path.join(gold, "wedge_history.npz"))
    em, cfg, bcmap = make_golden.wedge_setup()
    fm = cl.build_mesh(em, product_lib)
    assert fm.nCells == 6871 and fm.nNodes == 3568
    bcs = [bcmap[n] for n in fm.trs_names]
    U = np.zeros((fm.nCells, 4))
    U[:] = make_golden.WEDGE_INLET
    e = cl.Engine(cfg, 0, lib=product_lib)
    e.set_mesh(fm, bcs)
    e.set_states(U.reshape(-1))
    hist = []
    cfg.residual = 0.0
    for it in range(5000):
        cfg.iter = it + 1
        e.set_step(cfg)
        e.compute_rhs(apply_bc=True, download=False)
        res = e.update_sol_fwd_euler(0.7, 0)
        hist.append(res)
        cfg.residual = res
        if res < -4.0:
            break
    e.close()
    hist = np.array(hist)
    assert len(hist) == len(g["history"]) == 1984
    assert np.abs(hist - g["history"]).max() <= 1e-10
    pinned = float(g["reference_residual"])
    assert abs(hist[-1] - pinned)/abs(pinned)*100.0 < 0.1


def test_frozen_limiter_and_stop_limiting(oracle, product_lib):
    """GPU twins of the host-simulation cases: freezeLimiter keeps the stored limiter (FVMCC_PolyRec.cxx:131-165),
    StopLimiting reconstructs with phi = 1."""
    rng = np.random.RandomState(3)
    n = cases.CASES["ns3d_hex"][0]().n_cells*5
    for freeze in (0, 1):
        def mut(cfg):
            cfg.residual = -3.0
            cfg.freezeLimiter = freeze
        lim0 = 0.3 + 0.7*rng.random_sample(n)
        out = cases.run_case("ns3d_hex", oracle, _factory(product_lib), limiter_init=lim0, cfg_mut=mut)
        _check(out)
        if freeze:
            assert np.array_equal(out["limiter"][0], lim0)
        else:
            assert (out["limiter"][0] <= lim0).all() and (out["limiter"][0] < lim0).any()

    def stop(cfg):
        cfg.stopLimiting = 1
    _check(cases.run_case("euler2d_2nd_quad", oracle, _factory(product_lib), cfg_mut=stop))


def test_bad_value_is_reported(product_lib):
    """a^2 < 0 in the Roe linearisation -> CFB200_ERR_BAD_VALUE (BadValueException in the reference), with and without
    the Jacobian (the pipelined finite-difference loop carries the flag instead of branching)."""
    for jac in (False, True):
        fm = cl.build_mesh(meshgen.quad_channel(4, 4), product_lib)
        cfg = B.make_config(2, jacobian=jac)
        U = cases.init_state(fm, 4, False)
        U[4*5 + 3] = -50.0          # negative total energy in cell 5
        e = cl.Engine(cfg, 0, lib=product_lib)
        e.set_mesh(fm, cases.bcs_euler2d())
        e.set_states(U[:fm.nCells*4])
        with pytest.raises(cl.Cfb200Error) as ei:
            e.compute_rhs()
        assert ei.value.code == 1
        e.close()


def test_time_term_variants_separate_and_fused(oracle, product_lib):
    cases.check_time_term_variants(oracle, _factory(product_lib))


def test_renumbering_layer_is_a_pure_permutation(oracle, product_lib):
    cases.check_renumbering_layer(oracle, _factory(product_lib))


def test_renumbering_layer_at_size(product_lib):
    """96^3 hexahedra numbered at random: the renumbering engine returns, in the caller's numbering, exactly what the plain
    engine returns on the same mesh (which is bit-exact against the oracle: test_large_hex_ns_2M_cells)."""
    em, _ = meshgen.shuffle_cells(meshgen.hex_box(96, 96, 96))
    fm = cl.build_mesh(em, product_lib)
    U = cases.init_state(fm, 5, True)[:fm.nCells*5]
    got = []
    for ren in (0, 1):
        e = cl.Engine(cases.ns3d_cfg(renumber=ren, fuse_time_rhs=1, cfl=5.0), 0, lib=product_lib)
        e.set_mesh(fm, cases.bcs_ns3d())
        e.set_states(U)
        e.set_past_states(U*(1 + 1e-3))
        rhs, uc = e.compute_rhs()
        got.append((rhs, uc, e.bsr_values(), e.field("limiter"), e.bsr_pattern()))
        e.close()
    for a, b in zip(got[0][:4], got[1][:4]):
        assert np.array_equal(a, b)
    assert np.array_equal(got[0][4][0], got[1][4][0]) and np.array_equal(got[0][4][1], got[1][4][1])


def test_isothermal_wall_reports_ghost_reposition(oracle, product_lib):
    cases.check_ghost_reposition_is_reported(oracle, _factory(product_lib))


def test_position_dependent_inlet(oracle, product_lib):
    cases.check_bc_face_values(oracle, _factory(product_lib))


@pytest.mark.gpu
def test_nodal_values_prescribed_on_wall_nodes(oracle, product_lib):
    cases.check_nodal_prescribed(oracle, _factory(product_lib))


def test_newton_update(product_lib):
    cases.check_newton_update(_factory(product_lib), lambda em: cl.build_mesh(em, product_lib))


def test_large_size_properties(product_lib):
    """At a size the oracle does not finish quickly: size-independent properties.  (i) determinism; (ii) block-row
    sums: every flux is added to L and subtracted from R, so over a mesh whose cells are all updatable the column
    sums of the Jacobian restricted to interior-face terms vanish: sum_rows J[:, c] == boundary blocks only, i.e.
    for interior cells (no boundary face) diag(c) + sum_n block(n, c) == 0 exactly."""
    em = meshgen.hex_box(64, 64, 48)
    fm = cl.build_mesh(em, product_lib)
    cfg = cases.ns3d_cfg()
    e = cl.Engine(cfg, 0, lib=product_lib)
    e.set_mesh(fm, cases.bcs_ns3d())
    U = cases.init_state(fm, 5, True)
    e.set_states(U[:fm.nCells*5])
    rhs, uc = e.compute_rhs()
    jac = e.bsr_values().reshape(-1, 25)
    assert np.isfinite(rhs).all() and np.isfinite(jac).all() and (uc > 0).all()
    rp, ci = e.bsr_pattern()
    rows = np.repeat(np.arange(fm.nCells), np.diff(rp))
    colsum = np.zeros((fm.nCells, 25))
    np.add.at(colsum, ci, jac)
    interior = np.ones(fm.nCells, dtype=bool)
    interior[fm.faceL[fm.nInner:]] = False
    scale = np.abs(jac).max()
    assert np.abs(colsum[interior]).max() <= 1e-12*scale
    # (i) a second evaluation of the same state is bit-identical (no atomics, fixed gather order)
    e.set_update_coeff(None)          # the updateCoeff socket accumulates until the updater zeroes it
    rhs2, uc2 = e.compute_rhs()
    assert np.array_equal(rhs, rhs2) and np.array_equal(uc, uc2)
    assert np.array_equal(jac, e.bsr_values().reshape(-1, 25))
    e.close()


def test_full_size_freestream_and_determinism(product_lib):
    """BASELINE config 3 at its full size (220^3 = 10.6 M hexahedra): a uniform state is preserved by every cell
    without a boundary face (closed control volumes, zero gradients: the residual is rounding noise of the flux sums),
    and two evaluations of a non-uniform state give bit-identical residuals."""
    n = 220
    fm = cl.build_mesh(meshgen.hex_box(n, n, n), product_lib)
    cfg = cases.ns3d_cfg()
    e = cl.Engine(cfg, 0, lib=product_lib)
    e.set_mesh(fm, cases.bcs_ns3d())
    U = np.empty((fm.nCells, 5))
    U[:] = [1.0, 0.4, 0.03, 0.02, 1.0]                    # p u v w T in the units of cases.ns3d_cfg()
    e.set_states(U.reshape(-1))
    rhs, uc = e.compute_rhs()
    # two cell layers next to the boundary see the ghost states through their gradients
    x = fm.centres.reshape(-1, 3)[:fm.nCells]
    interior = ((x > 2.0/n) & (x < 1.0 - 2.0/n)).all(axis=1)
    assert interior.sum() == (n - 4)**3
    flux_scale = 3.5*1.0*0.4*(1.0/n)**2                   # ~ rho u H * face area
    assert np.abs(rhs.reshape(fm.nCells, 5)[interior]).max() <= 1e-11*flux_scale
    assert (uc > 0).all()
    U2 = cases.init_state(fm, 5, True)
    e.set_states(U2[:fm.nCells*5])
    e.set_update_coeff(None)
    a, ua = e.compute_rhs()
    e.set_update_coeff(None)
    b, ub = e.compute_rhs()
    assert np.isfinite(a).all() and np.array_equal(a, b) and np.array_equal(ua, ub)
    e.close()


def test_cuda_matches_committed_golden_vectors(product_lib):
    """The CUDA path against tests/golden/cases.npz (vectors generated by tests/golden/make_golden.py): no oracle
    run involved on the GPU box."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cases.npz"))
    for name in sorted(cases.CASES):
        mesh_f, cfg_f, bcs_f, pvt = cases.CASES[name]
        fm = cl.build_mesh(mesh_f(), product_lib)
        c = cfg_f()
        U = cases.init_state(fm, c.nEqs, pvt)
        uc0 = 0.1*np.arange(fm.nCells, dtype=np.float64)/max(fm.nCells, 1)
        e = cl.Engine(c, 0, lib=product_lib)
        e.set_mesh(fm, bcs_f())
        e.set_states(U[:fm.nCells*c.nEqs])
        e.set_update_coeff(uc0)
        rhs, uc = e.compute_rhs()
        got = {"rhs": rhs, "updateCoeff": uc, "limiter": e.field("limiter"), "jacobian": e.bsr_values()}
        e.close()
        for k, a in got.items():
            assert np.array_equal(a, g[name + "/" + k]), (name, k)


def test_div_sqrt_bitexact(product_lib):
    """The kernels' branch-free FP64 divide / square root (csrc/cfb_math.cuh) against the compiler's IEEE `/` and
    sqrt() on the device: 2^28 operand pairs per exponent span, zeros, exact quotients and shared reciprocals included.
    Bar: zero mismatching bit patterns."""
    import ctypes as C
    for span, seed in ((4, 1), (60, 2), (300, 3), (480, 4)):
        bad = C.c_int64(-1)
        rc = product_lib.selftest_divsqrt(0, 1 << 28, seed, span, C.byref(bad))
        assert rc == 0
        assert bad.value == 0, "span 2^+-%d: %d mismatches" % (span, bad.value)
